"""Roofline of the two HBM-bound kernels at BASELINE config-4 sizes:
FK + skinning (404 B/hand shared model, 1,688 B/hand per-sample models) and the crop resampler
(source bounding box + output + 256 B params per crop).  CUDA-event timed on the launching stream,
>= 3 warm-ups, working sets larger than L2.  Prints one JSON line per kernel."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from absolutetrack_b200 import synthetic
from absolutetrack_b200.lib.common.hand import HandModel
from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda
from absolutetrack_b200.lib.tracker.tracker import crop_resample_cuda

peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
HBM = peaks["hbm_gbs"]
dev = torch.device("cuda")

def timed(fn, reps=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

hm = synthetic.hand_model()
hm_dev = HandModel(*[(v.to(dev) if isinstance(v, torch.Tensor) else v) for v in hm])
g = torch.Generator(device=dev).manual_seed(0)
for n, per_sample in ((1 << 20, False), (1 << 22, False), (1 << 18, True)):
    ja = (torch.rand(n, 22, device=dev, generator=g) - 0.5) * 2.0
    wx = torch.eye(4, device=dev).repeat(n, 1, 1)
    wx[:, :3, 3] = torch.randn(n, 3, device=dev, generator=g) * 100
    hidx = torch.randint(0, 2, (n,), device=dev, generator=g)
    model = hm_dev
    if per_sample:
        model = HandModel(*[(v.unsqueeze(0).expand(n, *v.shape).contiguous() if isinstance(v, torch.Tensor) and v.dim() >= 2 else v) for v in hm_dev])
    out = torch.empty(n, 21, 3, device=dev)
    ms = timed(lambda: skin_landmarks_cuda(model, ja, wx, hand_idx=hidx, out=out))
    bytes_per = 404 + 8 + (1284 if per_sample else 0)   # + hand_idx int64
    gbs = n * bytes_per / (ms * 1e-3) / 1e9
    print(json.dumps({"kernel": "fk_skin", "n_hands": n, "per_sample_model": per_sample, "ms": ms, "bytes_per_hand": bytes_per,
                      "achieved_gbs": gbs, "peak_gbs": HBM, "frac": gbs / HBM, "hands_per_s": n / (ms * 1e-3)}), flush=True)
    del ja, wx, out

for S, map_mode, od in ((1024, "fp32", torch.uint8), (1024, "fp32", torch.float32), (256, "fp64", torch.uint8)):
    seq = np.arange(S)
    frames = torch.randint(0, 256, (S, synthetic.H, synthetic.V, synthetic.W), dtype=torch.uint8, device=dev, generator=g)
    hb = synthetic.make_batch(seq, 0, frames=None, first_step=False)
    params = hb.crop_params.to(dev)
    c = params.shape[0]
    out = torch.empty(c, 96, 96, dtype=od, device=dev)
    ms = timed(lambda: crop_resample_cuda(frames, synthetic.H * synthetic.V * synthetic.W, synthetic.W, synthetic.V * synthetic.W,
                                          params, (96, 96), od, map_mode, out=out))
    bytes_per = 32160 + 96 * 96 * out.element_size() + 256
    gbs = c * bytes_per / (ms * 1e-3) / 1e9
    print(json.dumps({"kernel": "crop_resample", "crops": c, "map_mode": map_mode, "out": str(od), "ms": ms, "bytes_per_crop": bytes_per,
                      "achieved_gbs": gbs, "peak_gbs": HBM, "frac": gbs / HBM, "crops_per_s": c / (ms * 1e-3)}), flush=True)
    del frames, out
