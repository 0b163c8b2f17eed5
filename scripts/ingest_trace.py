"""Timeline of BatchedHandTracker.step: when does the ingest (side stream) run relative to the compute?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from absolutetrack_b200 import synthetic
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
from absolutetrack_b200.lib.tracker.batch_tracker import BatchedHandTracker
S = 512
ingest = sys.argv[1] if len(sys.argv) > 1 else "pull"
model = model_from_state_dict(make_random_state_dict(0), precision="fp16")
tr = BatchedHandTracker(model, synthetic.hand_model(), max_slots=2 * S, map_mode="fp32", ingest=ingest)
hbs = [synthetic.make_batch(np.arange(S), i, frames=torch.randint(0, 256, (S, synthetic.H, synthetic.V, synthetic.W), dtype=torch.uint8)).pin() for i in range(2)]
out = None
for i in range(4): out = tr.step(hbs[i % 2], out)
torch.cuda.synchronize()
tr.trace = []
for i in range(8): out = tr.step(hbs[i % 2], out)
torch.cuda.synchronize()
t0 = tr.trace[0][0]
for i, ev in enumerate(tr.trace):
    a, b, c, d = [t0.elapsed_time(e) for e in ev]
    print(f"step {i}: ingest {a:7.3f} -> {b:7.3f} ({b - a:5.3f} ms)   compute {c:7.3f} -> {d:7.3f} ({d - c:5.3f} ms)")
