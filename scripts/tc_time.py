"""Times every conv launch of the tcgen05 encoder with CUDA events (per-layer TFLOP/s)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
c = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
model = model_from_state_dict(make_random_state_dict(0), precision=prec).cuda()
crops = (torch.rand(c, 96, 96) * 255).to(torch.uint8).cuda()
for _ in range(3): model.encode(crops)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): model.encode(crops)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"encoder {prec} c={c}: {ms:.3f} ms  {969.228288e6 * c / ms / 1e9:.1f} TFLOP/s", flush=True)
