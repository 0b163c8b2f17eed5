"""Per-launch role counters of one encoder pass (debug instrumentation of conv_tc_kernel, AT_TC_DBG=16):
for every conv launch the MMA warp's cycles and the share it spent waiting for an accumulator stage / the
activation ring / the weight ring, and the epilogue groups' wait share.

  AT_TC_DBG=16 python scripts/encoder_prof.py [--crops 2048]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("AT_TC_DBG", "16")
import torch  # noqa: E402

from absolutetrack_b200 import _cabi  # noqa: E402
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--crops", type=int, default=2048)
args = ap.parse_args()
lib = _cabi.lib()
dev = torch.device("cuda", 0)
model = model_from_state_dict(make_random_state_dict(0), precision="fp16").cuda()
crops = torch.randint(0, 256, (args.crops, 96, 96), dtype=torch.uint8, device=dev)
for _ in range(3):
    model.encode(crops)
torch.cuda.synchronize()
prof = torch.zeros(32, 148, 16, dtype=torch.int64, device=dev)
lib.at_debug_set_prof_seq(prof.data_ptr())
model.encode(crops)
torch.cuda.synchronize()
lib.at_debug_set_prof_seq(None)
P = prof.double().cpu()
print("launch  span_us  tiles/CTA  mma_cyc/tile  wait_acc%  wait_a%  wait_w%  epi0_wait%  epi1_wait%")
for k in range(32):
    q = P[k]
    if q[:, 4].sum() == 0:
        continue
    span = (q[:, 6].max() - q[:, 5][q[:, 5] > 0].min()).item() / 1e3 if (q[:, 6] > 0).any() else float("nan")
    m = q.mean(0)
    print(f"{k:4d}  {span:8.1f}  {m[4]:8.1f}  {m[0] / max(m[4], 1):10.0f}  {100 * m[1] / m[0]:8.1f}  {100 * m[2] / m[0]:7.1f}  {100 * m[3] / m[0]:7.1f}"
          f"  {100 * m[9] / max(m[8], 1):9.1f}  {100 * m[11] / max(m[10], 1):9.1f}")
