"""Probe: conv3x3 weight-gradient kernel (csrc/conv_wgrad_bf16.cu) at the config-5 shape for each strip width.

Needs a measurement build: ``OCV_NVCC_EXTRA=-DOCV_PROBES python -m objcavit_b200.build -f`` (the shipped library ignores
OCV_WGRAD_SW).  Prints the mean time per launch over 100 back-to-back launches (operands 2 x 116 MB > L2) and checks that every
strip width gives the same gradient as the default."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops

dev = torch.device("cuda")
shapes = [(8, 208, 272), (16, 240, 320), (8, 176, 608)]
for B, h, w in shapes:
    g = torch.Generator().manual_seed(0)
    xh = (torch.randn(B, h, w, 128, generator=g) * 0.5).bfloat16().to(dev)
    dyT = (torch.randn(128, B * h * w, generator=g) * 0.1).bfloat16().to(dev)
    base = None
    for sw in ("0", "16", "32", "48", "64"):
        os.environ["OCV_WGRAD_SW"] = sw
        gw = torch.zeros(128, 128, 3, 3, device=dev); gb = torch.zeros(128, device=dev)
        for _ in range(10):
            ops.conv3x3_wgrad_bf16(dyT, xh, gw, gb, beta=0.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(100):
            ops.conv3x3_wgrad_bf16(dyT, xh, gw, gb, beta=0.0)
        e1.record(); torch.cuda.synchronize()
        if base is None:
            base = gw.clone()
        err = float((gw - base).abs().max() / base.abs().max())
        print(f"B{B} {h}x{w} sw={sw:>2}: {e0.elapsed_time(e1) * 10:.1f} us/launch, max diff vs default {err:.2e}", flush=True)
