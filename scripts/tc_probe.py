"""GPU probe of the tcgen05 conv path: every backbone conv layer through at_debug_conv_tc vs a
torch CPU conv on identically rounded operands, then the whole encoder vs the fp32 CUDA path.
Prints diagnostics instead of asserting (used while bringing the kernel up)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch, torch.nn.functional as F
from absolutetrack_b200 import _cabi
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict

lib = _cabi.lib()
dev = torch.device("cuda")
prec_name = sys.argv[1] if len(sys.argv) > 1 else "bf16"
mt = int(sys.argv[2]) if len(sys.argv) > 2 else 0
PREC = {"bf16": _cabi.AT_PREC_BF16, "fp16": _cabi.AT_PREC_FP16}[prec_name]
tdt = {"bf16": torch.bfloat16, "fp16": torch.float16}[prec_name]
model = model_from_state_dict(make_random_state_dict(0), precision=prec_name).cuda()
blob = model._blob
tcb = model._bf16_weights(dev)
torch.cuda.synchronize()
table = _cabi.layer_table()
q = lambda t: t.to(tdt).float()

def run_layer(lid, n_img, H, W, use_res, relu, seed=0):
    L = table[lid]
    g = torch.Generator().manual_seed(seed)
    cin, cout, k, st = L["cin"], L["cout"], L["ksize"], L["stride"]
    x = q(torch.randn(n_img, H, W, cin, generator=g))
    Ho, Wo = H // st, W // st
    res = q(torch.randn(n_img, Ho, Wo, cout, generator=g)) if use_res else None
    coutp = (cout + 3) // 4 * 4
    w = blob[L["weight_offset"]: L["weight_offset"] + k * k * cin * coutp].cpu().reshape(k, k, cin, coutp)[..., :cout]
    b = blob[L["bias_offset"]: L["bias_offset"] + cout].cpu()
    ref = F.conv2d(x.permute(0, 3, 1, 2), q(w).permute(3, 2, 0, 1), b, st, k // 2).permute(0, 2, 3, 1)
    if use_res: ref = ref + res
    if relu: ref = F.relu(ref)
    out = torch.full((n_img, Ho, Wo, cout), float("nan"), device=dev)
    nbytes = lib.at_debug_conv_tc_workspace_bytes(n_img, H, W)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    xd = x.to(dev).contiguous(); rd = res.to(dev).contiguous() if use_res else None
    rc = lib.at_debug_conv_tc(blob.data_ptr(), tcb.data_ptr(), lid, xd.data_ptr(), n_img, H, W, _cabi.ptr(rd), int(relu), PREC, mt,
                              out.data_ptr(), ws.data_ptr(), nbytes, _cabi.stream_ptr(dev))
    if rc != 0:
        print(f"  layer {lid} {L['name']}: rc={rc} {lib.at_last_error_string().decode()}"); return None
    torch.cuda.synchronize()
    o = out.cpu()
    err = (o - ref).abs()
    scale = ref.abs().max().item()
    nan = int(torch.isnan(o).sum())
    tol = 1e-4 * scale + 1e-5 if lid == 28 else (2.0 ** -7) * scale
    bad = err > tol
    msg = f"  layer {lid:2d} {L['name']:28s} cin={cin:3d} cout={cout:3d} k={k} s={st} n_img={n_img} {H}x{W} res={int(use_res)}: max_err={np.nanmax(err.numpy()):.4g} scale={scale:.3g} nan={nan} bad={int(bad.sum())}/{bad.numel()}"
    if bad.any() or nan:
        idx = torch.nonzero(bad | torch.isnan(o))
        msg += f"  first_bad={idx[0].tolist()} last_bad={idx[-1].tolist()} out={o[tuple(idx[0])].item():.4g} ref={ref[tuple(idx[0])].item():.4g}"
        # per-channel / per-row pattern
        msg += f"  bad_by_channel_mod8={[int(bad[..., c::8].sum()) for c in range(8)]}"
    print(msg, flush=True)
    return bool(not bad.any() and not nan)

cases = [  # (layer id, n_img, H, W, res, relu)
    (1, 3, 48, 48, False, True), (2, 3, 48, 48, True, True),          # 32->32
    (5, 3, 48, 48, False, True), (7, 3, 48, 48, False, False),        # stride-2 3x3 32->64, 1x1 s2 down
    (6, 5, 24, 24, True, True),                                       # 64->64
    (12, 5, 24, 24, False, True), (14, 5, 24, 24, False, False), (13, 9, 12, 12, True, True),   # level 3
    (23, 9, 12, 12, False, True), (25, 9, 12, 12, False, False), (24, 33, 6, 6, True, True),    # level 4
    (28, 33, 6, 6, False, False),                                     # proj -> fp32
]
ok = True
for c in cases:
    r = run_layer(*c)
    ok = ok and bool(r)
print("layers ok:", ok, flush=True)

# whole encoder vs the fp32 CUDA-core path
m32 = model_from_state_dict(make_random_state_dict(0), precision="fp32").cuda()
g = torch.Generator().manual_seed(1)
for c in (1, 5, 64):
    crops = torch.rand(c, 96, 96, generator=g).cuda()
    f32 = m32.encode(crops)
    ftc = model.encode(crops)
    torch.cuda.synchronize()
    d = (f32 - ftc).abs()
    print(f"encoder c={c}: max|dfeat|={d.max().item():.4g} mean={d.mean().item():.4g} scale={f32.abs().max().item():.3g} nan={int(torch.isnan(ftc).sum())}", flush=True)
# timing
for c in (256, 1024, 2048):
    crops = (torch.rand(c, 96, 96, generator=g) * 255).to(torch.uint8).cuda()
    for _ in range(3): model.encode(crops)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): model.encode(crops)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"encoder {prec_name} c={c}: {ms:.3f} ms  {969.228288e6 * c / ms / 1e9:.1f} TFLOP/s", flush=True)
