"""Stand-alone duration of the source-box pull kernel (pinned host frames -> device) vs CTA count."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from absolutetrack_b200 import synthetic
from absolutetrack_b200.lib.tracker.tracker import crop_pull_source_boxes_cuda
S = int(sys.argv[1]) if len(sys.argv) > 1 else 512
fr = torch.randint(0, 256, (S, synthetic.H, synthetic.V, synthetic.W), dtype=torch.uint8)
hb = synthetic.make_batch(np.arange(S), 1, frames=fr).pin()
params = hb.crop_params.cuda()
dst = torch.empty_like(hb.frames, device="cuda")
cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
for ctas in [int(a) for a in sys.argv[2:]] or [37, 74, 148, 296, 592, 1184, 2048]:
    for _ in range(2):
        crop_pull_source_boxes_cuda(hb.frames, dst, synthetic.H * synthetic.V * synthetic.W, synthetic.W, synthetic.V * synthetic.W, params, (96, 96), ctas, cnt)
    torch.cuda.synchronize(); cnt.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        crop_pull_source_boxes_cuda(hb.frames, dst, synthetic.H * synthetic.V * synthetic.W, synthetic.W, synthetic.V * synthetic.W, params, (96, 96), ctas, cnt)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    mb = cnt.item() / 5 / 1e6
    print(f"ctas {ctas:5d}: {ms:7.3f} ms  {mb:6.1f} MB pulled  {mb / ms:6.1f} GB/s")
