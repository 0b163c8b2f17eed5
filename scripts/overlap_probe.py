"""Encoder time with a tiny side kernel co-resident on every SM (scripts/overlap_probe.cu)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
P = C.CDLL(os.path.join(ROOT, "scripts", "liboverlap_probe.so"))
P.probe_launch.argtypes = [C.c_int, C.c_int, C.c_ulonglong, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
model = model_from_state_dict(make_random_state_dict(0), precision="fp16").cuda()
crops = (torch.rand(2048, 96, 96) * 255).to(torch.uint8).cuda()
host = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
devsrc = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
scratch = torch.zeros(4, dtype=torch.int32, device="cuda")
side = torch.cuda.Stream(priority=-1)
for _ in range(5): model.encode(crops)
torch.cuda.synchronize()
def run(kind, ctas, src):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if kind >= 0:
        with torch.cuda.stream(side):
            P.probe_launch(kind, ctas, 12_000_000, src.data_ptr() if src is not None else None, src.numel() if src is not None else 0, scratch.data_ptr(), side.cuda_stream)
        torch.cuda._sleep(200_000)  # let the side kernel become resident first
    e0.record()
    for _ in range(3): model.encode(crops)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 3
print(f"encoder alone                       {run(-1, 0, None):.3f} ms")
for carve in (100,):
    for name, kind, src in (("alu spin", 1, None), ("cp.async.ca pcie linear", 3, host), ("cp.async.ca pcie scatter", 4, host),
                            ("cp.async.ca hbm linear", 3, devsrc), ("cp.async.ca hbm scatter", 4, devsrc)):
        for ctas in (148, 37, 16):
            print(f"carveout {carve:4d} side {name:26s} ctas {ctas:4d}:   {run(kind | ((carve + 1) << 8), ctas, src):.3f} ms")
print(f"encoder alone                       {run(-1, 0, None):.3f} ms")
