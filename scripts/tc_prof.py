"""Per-role cycle counters of single conv_tc layers at full batch size (debug instrumentation)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from absolutetrack_b200 import _cabi
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
lib = _cabi.lib(); dev = torch.device("cuda")
model = model_from_state_dict(make_random_state_dict(0), precision="bf16").cuda()
blob, tcb = model._blob, model._bf16_weights(dev)
table = _cabi.layer_table()
prof = torch.zeros(148, 16, dtype=torch.int64, device=dev)
cases = [(1, 1024, 48, False), (2, 1024, 48, True), (5, 1024, 48, False), (7, 1024, 48, False), (8, 2048, 24, False), (6, 2048, 24, True),
         (15, 2048, 12, False), (13, 2048, 12, True), (26, 2048, 6, False), (24, 2048, 6, True), (28, 2048, 6, False)]
mt = int(sys.argv[1]) if len(sys.argv) > 1 else 0
for lid, n_img, H, use_res in cases:
    L = table[lid]
    Ho = H // L["stride"]
    x = torch.randn(n_img, H, H, L["cin"], device=dev)
    res = torch.randn(n_img, Ho, Ho, L["cout"], device=dev) if use_res else None
    out = torch.empty(n_img, Ho, Ho, L["cout"], device=dev)
    nbytes = lib.at_debug_conv_tc_workspace_bytes(n_img, H, H)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    for rep in range(2):
        prof.zero_()
        lib.at_debug_set_prof(prof.data_ptr())
        rc = lib.at_debug_conv_tc(blob.data_ptr(), tcb.data_ptr(), lid, x.data_ptr(), n_img, H, H, _cabi.ptr(res), 1, _cabi.AT_PREC_BF16, mt,
                                  out.data_ptr(), ws.data_ptr(), nbytes, _cabi.stream_ptr(dev))
        lib.at_debug_set_prof(None)
        torch.cuda.synchronize()
    assert rc == 0, lib.at_last_error_string()
    P = prof.double()
    span = (P[:, 6].max() - P[:, 5].min()).item() / 1e3
    skew = (P[:, 5].max() - P[:, 5].min()).item() / 1e3
    pro = (P[:, 7] - P[:, 5]).mean().item() / 1e3
    endskew = (P[:, 6].max() - P[:, 6].min()).item() / 1e3
    mm = P[:, 0]
    print(f"   span={span:.1f}us start_skew={skew:.1f}us prologue={pro:.1f}us end_skew={endskew:.1f}us mma_total min/mean/max={mm.min().item()/1.9e3:.1f}/{mm.mean().item()/1.9e3:.1f}/{mm.max().item()/1.9e3:.1f}us")
    p = prof.double().mean(0).tolist()
    tiles = p[4]
    flops = 2.0 * n_img * Ho * Ho * L["cout"] * L["cin"] * L["ksize"] ** 2
    us = p[0] / 1.9e3
    print(f"layer {lid:2d} {L['name']:26s} res={int(use_res)} tiles/CTA={tiles:6.1f} mma_total={p[0]:9.0f}cyc ({us:6.1f}us, {flops/us/1e6:6.0f} TF) per-tile={p[0]/max(tiles,1):7.0f} "
          f"wait acc={100*p[1]/p[0]:4.1f}% a={100*p[2]/p[0]:4.1f}% w={100*p[3]/p[0]:4.1f}% | epi0 total={p[8]:9.0f} wait={100*p[9]/max(p[8],1):4.1f}% epi1 wait={100*p[11]/max(p[10],1):4.1f}%", flush=True)
    del ws, x, out
