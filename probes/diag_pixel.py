import sys, os, math, torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from objcavit_b200 import ops
from oracle import head as oh
from test_gpu_stages import _pixel_reference, g
cuda = torch.device("cuda:0")
for (B, h, w, mc) in [(1, 16, 8, 0), (2, 48, 40, 0), (3, 37, 29, 2)]:
    x = torch.randn(B, 128, h, w, generator=g(1))
    w3 = torch.randn(128, 128, 3, 3, generator=g(2)) * 0.03; b3 = torch.randn(128, generator=g(3)) * 0.1
    wout = torch.randn(256, 128, 1, 1, generator=g(4)) * 0.1; bout = torch.randn(256, generator=g(5)) * 0.1
    tok = torch.randn(B, 140, 128, generator=g(6))
    wn = torch.rand(B, 256, generator=g(7)) + 0.1
    wn = wn / wn.sum(1, keepdim=True)
    _, centres = oh.bin_edges_and_centres(wn, 0.001, 10.0)
    want16 = _pixel_reference(x, w3, b3, wout, bout, tok[:, 1:129, :], centres)
    xd = ops.nchw_to_nhwc_bf16(x.to(cuda)); w3p = ops.pack_conv3x3_bf16(w3.to(cuda))
    Mb, blog = ops.fold_queries_bf16(wout.to(cuda), bout.to(cuda), tok.to(cuda)[:, 1:129, :], b3.to(cuda))
    got = ops.pixel_fused_bf16(xd, w3p, Mb, blog, centres.to(cuda), max_ctas=mc).cpu()
    err = (got - want16).abs()
    print(B, h, w, "max", float(err.max()), "mean", float(err.mean()), "want mean", float(want16.abs().mean()),
          "n>1e-2:", int((err > 1e-2).sum()), "of", err.numel())
    idx = torch.nonzero(err > 0.5 * err.max())
    print("  worst at", idx[:8].tolist())
