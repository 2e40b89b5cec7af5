"""Probe: one launch each of the round-2 kernels at the config-5 / config-2 sizes, for `ncu --set full -k regex:...`:
tail_fwd / tail_bwd (training pixel path without logits), gemm_bf16_kernel in its training forms, the tcgen05 attention core
forward / backward, and a cuBLAS bf16 8192^3 GEMM (torch.matmul) as the tensor-pipe reference the roofline peak comes from."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
B, Pn = 8, 208 * 272
x3 = (torch.randn(B, Pn, 128, generator=g) * 0.7).to(dev).bfloat16()
Mb2 = (torch.randn(B, 256, 128, generator=g) * 0.25).to(dev).bfloat16()
blog2 = (torch.randn(B, 256, generator=g) * 0.5).to(dev)
centres = torch.rand(B, 256, generator=g).cumsum(1).to(dev).contiguous()
gd = torch.randn(B, Pn, generator=g).to(dev)
for _ in range(2):
    depth, smax, ssum = ops.tail_fused_fwd_bf16(x3, Mb2, blog2, centres)
    ops.tail_fused_bwd_bf16(x3, Mb2, blog2, centres, depth, smax, ssum, gd)
# attention: training shape (B=8, S=221) and inference shape (B=16, S=300)
for Bq, S in ((8, 221), (16, 300)):
    qkv = torch.randn(Bq, S, 384, generator=g).to(dev).bfloat16()
    q, k, v = qkv[:, :, 0:128], qkv[:, :, 128:256], qkv[:, :, 256:]
    for _ in range(2):
        o16, _, st = ops.mha_core_tc_bf16(q, k, v, None, want_stats=True)
        if S <= 256:
            ops.mha_core_bwd_tc_bf16(q, k, v, o16, o16, st)
# encoder GEMMs in their training forms (M = 1768 token rows)
M = 8 * 221
a = torch.randn(M, 128, generator=g).to(dev).bfloat16()
w1 = torch.randn(1024, 128, generator=g).to(dev).bfloat16()
h = torch.randn(M, 1024, generator=g).to(dev).bfloat16()
for _ in range(2):
    ops.gemm_bf16_v2(a, w1, act=ops.ACT_RELU, want16=True)                 # linear1 forward
    ops.gemm_bf16_v2(h, w1, b_mn=True)                                     # dX = dY W (B MN-major)
    ops.gemm_bf16_v2(h, a, a_mn=True, b_mn=True)                           # dW = dY^T X (both MN-major)
# cuBLAS reference
A = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
Bm = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
for _ in range(3):
    C = A @ Bm
torch.cuda.synchronize()
print("ok")
