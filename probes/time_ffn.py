"""Fused FFN kernel vs the two-GEMM path at the head's token count (M = 4800): CUDA-event time per call, L2-warm (back to back)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops
dev = torch.device("cuda:0")
for M in (4800, 512 * 16 // 16):
    g = torch.Generator().manual_seed(0)
    x32 = torch.randn(M, 128, generator=g).to(dev); x16 = x32.bfloat16()
    w1 = (torch.randn(1024, 128, generator=g) * 0.05).to(dev).bfloat16(); w2 = (torch.randn(128, 1024, generator=g) * 0.03).to(dev).bfloat16()
    b1 = torch.zeros(1024, device=dev); b2 = torch.zeros(128, device=dev); ga = torch.ones(128, device=dev); be = torch.zeros(128, device=dev)
    def fused():
        ops.ffn_fused_bf16(x16, x32, w1, b1, w2, b2, ga, be)
    def two():
        _, h = ops.gemm_bf16_ex(x16, w1, b1, act=ops.ACT_RELU)
        ops.gemm_bf16_ex(h, w2, b2, residual=x32, ln=(ga, be), want32=True)
    for name, fn in (("fused", fused), ("two-gemm", two)):
        for _ in range(10): fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(200): fn()
        e1.record(); torch.cuda.synchronize()
        print(f"M={M} {name}: {e0.elapsed_time(e1) / 200 * 1e3:.1f} us/call")
