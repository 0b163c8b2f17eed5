"""Smallest run that touches every kernel family once (for compute-sanitizer --tool memcheck / racecheck)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from absolutetrack_b200 import synthetic
from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
from absolutetrack_b200.lib.tracker.batch_tracker import BatchedHandTracker
from absolutetrack_b200.lib.tracker.tracker import crop_resample_cuda

sd = make_random_state_dict(0)
hm = synthetic.hand_model()
S = 3
seq = np.arange(S)
frames = torch.randint(0, 256, (S, synthetic.H, synthetic.V, synthetic.W), dtype=torch.uint8)
for prec in ("fp16", "bf16", "fp32"):
    model = model_from_state_dict(sd, precision=prec)
    tr = BatchedHandTracker(model, hm, max_slots=2 * S)
    for step in range(2):
        hb = synthetic.make_batch(seq, step, frames=frames, first_step=(step == 0)).pin()
        out = tr.step(hb)                       # source-box pull + lattice crop + encoder + tail + FK
        r = tr.step_device(hb.to("cuda"))       # device-resident crop path
    torch.cuda.synchronize(); tr.check()
    print(prec, "ok", float(r.landmarks.abs().mean()), flush=True)
# crop-camera generator (step_poses) + evaluation sums
tr.attach_rig(*synthetic.rig_cameras())
pb = synthetic.make_pose_batch(seq, 0, frames=frames.pin_memory()).pin()
o = tr.step_poses(pb); torch.cuda.synchronize(); print("step_poses ok", o["n"], flush=True)
# exact fp64 crop map, float outputs, staged / fast variants
db = hb.to("cuda")
for mode, dt in (("fp64", torch.uint8), ("fp32", torch.float32), ("fp64", torch.bfloat16)):
    crop_resample_cuda(db.frames, synthetic.H * synthetic.V * synthetic.W, synthetic.W, synthetic.V * synthetic.W, db.crop_params, (96, 96), dt, mode)
crop_resample_cuda(db.frames, synthetic.H * synthetic.V * synthetic.W, synthetic.W, synthetic.V * synthetic.W, db.crop_params, (64, 64), torch.uint8, "fp32")
# FK with per-sample models + evaluation table kernels
from absolutetrack_b200.lib.common.hand import HandModel
n = 70
rep = HandModel(*[(v.unsqueeze(0).expand(n, *v.shape).contiguous() if isinstance(v, torch.Tensor) and v.dim() >= 2 else v) for v in hm])
lm = skin_landmarks_cuda(rep, torch.rand(n, 22) * 0.3, torch.eye(4).repeat(n, 1, 1))
from absolutetrack_b200 import load_eval
es = load_eval.EvalSums()
kp = torch.rand(2, 9, 21, 3)
load_eval.compute_metric_sums_cuda(kp.numpy().astype(np.float64), (kp + 1).numpy().astype(np.float64), np.ones((2, 9), bool), es)
torch.cuda.synchronize(); print("fk per-sample + eval ok", flush=True)
# config-3 preprocessing: crop matrices + homography resampler on the committed fixture files
from absolutetrack_b200.lib.batched_dataset import data_transform as DT
d = os.path.join(ROOT, "tests", "golden", "torchdata")
for i, mi, mt in DT.iter_preprocessed_sequences(os.path.join(d, "seq.mono.torch.idx"), os.path.join(d, "seq.labels.torch.idx")):
    pass
torch.cuda.synchronize(); print("config 3 preprocess ok", flush=True)
