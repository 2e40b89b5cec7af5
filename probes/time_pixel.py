"""Quick timing of the fused pixel kernel alone (probe; the judged numbers come from bench.py)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops
B, h, w = int(sys.argv[1]) if len(sys.argv) > 1 else 16, 240, 320
dev = torch.device("cuda:0")
x = torch.randn(B, h, w, 128, device=dev).bfloat16()
w3p = (torch.randn(9, 128, 128, device=dev) * 0.03).bfloat16()
Mb = (torch.randn(B, 256, 128, device=dev) * 0.1).bfloat16()
blog = torch.randn(B, 256, device=dev) * 0.1
cen = torch.rand(B, 256, device=dev).cumsum(1)
for _ in range(3):
    d = ops.pixel_fused_bf16(x, w3p, Mb, blog, cen)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 20
e0.record()
for _ in range(n):
    d = ops.pixel_fused_bf16(x, w3p, Mb, blog, cen)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
gf = (22.649 + 2.517 + 5.033) * B * (h * w / 76800.0)
print(f"pixel_fused B={B}: {ms:.3f} ms/launch, {B / ms * 1e3:.0f} maps/s, algorithmic {gf / ms:.1f} TFLOP/s, "
      f"executed {(22.649 + 5.033) * B / ms:.1f} TFLOP/s, finite={bool(torch.isfinite(d).all())}")
