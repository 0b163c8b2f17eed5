"""Smallest run that touches every kernel once (for compute-sanitizer --tool memcheck)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from absolutetrack_b200 import synthetic
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
from absolutetrack_b200.lib.tracker.batch_tracker import BatchedHandTracker
for prec in ("fp16", "fp32"):
    model = model_from_state_dict(make_random_state_dict(0), precision=prec)
    hm = synthetic.hand_model()
    S = 3
    tr = BatchedHandTracker(model, hm, max_slots=2 * S)
    seq = np.arange(S)
    frames = torch.randint(0, 256, (S, synthetic.H, synthetic.V, synthetic.W), dtype=torch.uint8)
    for step in range(2):
        hb = synthetic.make_batch(seq, step, frames=frames, first_step=(step == 0)).pin()
        out = tr.step(hb)                       # zero-copy staged crop + encoder + tail + FK
        r = tr.step_device(hb.to("cuda"))       # device-resident crop path
    torch.cuda.synchronize(); tr.check()
    print(prec, "ok", float(r.landmarks.abs().mean()))
