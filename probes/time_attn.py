"""Probe: the attention core on tcgen05 (csrc/attn_tc.cu) against the mma.sync kernel (csrc/attn_bf16.cu) at the reference shapes,
CUDA events over 200 back-to-back launches (warm), us per launch."""
import os, sys, subprocess, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)


def t(fn, n=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


which = "tcgen05" if os.environ.get("OCV_ATTN_TC", "1") != "0" else "mma.sync"
for B, S, label in ((16, 300, "config 2 image encoder"), (8, 418, "config 3 KITTI"), (32, 300, "config 4"), (16, 32, "object encoder"), (8, 221, "config 5")):
    qkv = torch.randn(B, S, 384, generator=g).to(dev).bfloat16()
    q, k, v = qkv[:, :, 0:128], qkv[:, :, 128:256], qkv[:, :, 256:]
    print(f"{which:9s} B={B:2d} S={S:3d} ({label}): {t(lambda: ops.mha_core_bf16(q, k, v, None)):6.1f} us")
