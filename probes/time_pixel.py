"""Timing of the fused pixel kernel alone, at several grid sizes (probe; the judged numbers come from bench.py).
Per-pair cycles at fewer CTAs tell a shared-resource (L2) limit from a per-SM latency limit."""
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
pairs = B * ((h + 15) // 16) * ((w + 15) // 16)
for ctas in ([int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else (0, 111, 74, 37, 8)):
    for _ in range(3):
        d = ops.pixel_fused_bf16(x, w3p, Mb, blog, cen, max_ctas=ctas)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for _ in range(n):
        d = ops.pixel_fused_bf16(x, w3p, Mb, blog, cen, max_ctas=ctas)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    g = ctas if ctas else 148
    per_cta = -(-pairs // g)
    gf = (22.649 + 2.517 + 5.033) * B * (h * w / 76800.0)
    print(f"pixel_fused B={B} ctas={g}: {ms:.3f} ms/launch, {B / ms * 1e3:.0f} maps/s, algorithmic {gf / ms:.1f} TFLOP/s, "
          f"~{ms * 1e-3 * 1.965e9 / per_cta:.0f} clk/pair-block, finite={bool(torch.isfinite(d).all())}")
