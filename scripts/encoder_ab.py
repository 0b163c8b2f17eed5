"""Encoder-only A/B on the GPU box: ms per pass at C crops and a hash of the features (env knobs are read once per
process, so every configuration is its own run).

  AT_TC_SUB=256 python scripts/encoder_ab.py [--crops 2048] [--iters 30]
prints one JSON line: {"env": {...AT_*...}, "crops": C, "ms": median, "ms_min": .., "tflops": .., "features_sha1": ".."}
"""
import argparse
import hashlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--crops", type=int, default=2048)
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--precision", default="fp16")
    args = ap.parse_args()
    import torch

    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict

    dev = torch.device("cuda", 0)
    model = model_from_state_dict(make_random_state_dict(0), precision=args.precision).cuda()
    g = torch.Generator(device=dev).manual_seed(5)
    crops = [torch.randint(0, 256, (args.crops, 96, 96), dtype=torch.uint8, device=dev, generator=g) for _ in range(2)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for i in range(3):
        f = model.encode(crops[i & 1])
    torch.cuda.synchronize()
    sha = hashlib.sha1(model.encode(crops[0]).float().cpu().numpy().tobytes()).hexdigest()
    # accuracy of the 16-bit features against the fp32 CUDA-core encoder on the first 64 crops
    ref_model = model_from_state_dict(make_random_state_dict(0), precision="fp32").cuda()
    fr = ref_model.encode(crops[0][:64]).float()
    ft = model.encode(crops[0][:64]).float()
    err = (ft - fr).abs()
    err_vs_fp32 = {"max": float(err.max()), "mean": float(err.mean()), "ref_abs_mean": float(fr.abs().mean())}
    ms = []
    for i in range(args.iters):
        flush.zero_()          # L2 holds nothing of the previous pass
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        model.encode(crops[i & 1])
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    med = ms[len(ms) // 2]
    print(json.dumps({"env": {k: v for k, v in os.environ.items() if k.startswith("AT_")}, "crops": args.crops, "ms": med,
                      "ms_min": ms[0], "tflops": 969228288 * args.crops / (med * 1e-3) / 1e12, "features_sha1": sha,
                      "err_vs_fp32_encoder": err_vs_fp32}))


if __name__ == "__main__":
    main()
