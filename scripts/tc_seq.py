"""Per-launch timeline of one encoder pass: globaltimer span of every conv_tc launch, the gap to
the next launch and the effective SM clock (MMA-warp cycles / span)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from absolutetrack_b200 import _cabi
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
c = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
lib = _cabi.lib()
model = model_from_state_dict(make_random_state_dict(0), precision=(sys.argv[2] if len(sys.argv) > 2 else "fp16")).cuda()
crops = (torch.rand(c, 96, 96) * 255).to(torch.uint8).cuda()
for _ in range(20): model.encode(crops)
torch.cuda.synchronize()
prof = torch.zeros(64, 148, 16, dtype=torch.int64, device="cuda")
lib.at_debug_set_prof_seq(prof.data_ptr())
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); model.encode(crops); e1.record()
lib.at_debug_set_prof(None)
torch.cuda.synchronize()
print(f"encode (instrumented) {e0.elapsed_time(e1)*1e3:.0f} us")
P = prof.double().cpu()
prev_end = None; t_first = None; tot_span = 0; tot_gap = 0
for k in range(64):
    if P[k, :, 5].max() == 0: break
    act = P[k, :, 5] > 0
    st, en = P[k, act, 5].min().item(), P[k, act, 6].max().item()
    if t_first is None: t_first = st
    gap = (st - prev_end) / 1e3 if prev_end is not None else 0.0
    span = (en - st) / 1e3
    cyc = P[k, act, 0].max().item()
    tiles = P[k, act, 4].max().item()
    wa, wA, ww = [100 * P[k, act, i].mean().item() / max(P[k, act, 0].mean().item(), 1) for i in (1, 2, 3)]
    print(f"launch {k:2d}: start {(st - t_first)/1e3:8.1f} span {span:7.1f} us  gap_before {gap:5.1f}  mma_cyc_max {cyc:9.0f}  clk {cyc/span/1e3:5.2f} GHz  tiles/CTA {tiles:5.0f} cyc/tile {cyc/max(tiles,1):7.0f}  wait acc {wa:4.1f}% a {wA:4.1f}% w {ww:4.1f}%")
    prev_end = en; tot_span += span; tot_gap += gap
print(f"sum span {tot_span:.0f} us  sum gaps {tot_gap:.0f} us  first->last {(prev_end - t_first)/1e3:.0f} us")
