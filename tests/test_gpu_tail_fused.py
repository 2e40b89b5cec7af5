"""GPU: the logits-free pixel path of the training step (csrc/train_tail_fused_bf16.cu) against fp64 torch math on the SAME bf16
operands: forward (depth + softmax statistics) and the flash-style backward (d x3, d Mb, d centres, d conv_out bias) whose logits
tile is recomputed on the tensor cores.  Shapes make CTAs walk several tiles, cross an image boundary inside one CTA (363 tiles on
148 CTAs, 121 per image) and leave CTAs without work (5 tiles).
reference sites: modules/GraphBins.py:109-119 + modules/layers.py:31-36 and their autograd.
Tolerances: forward 2e-5 relative (fp32 accumulation of exact bf16 products, ex2.approx); backward 1e-2 of the largest |value| per
tensor — d logits is rounded to bf16 (2^-9 relative per element) before the two gradient GEMMs."""
import math

import pytest
import torch

from objcavit_b200 import ops

pytestmark = pytest.mark.gpu
LOG2E = 1.4426950408889634


def _case(cuda, B, Pn, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x3 = (torch.randn(B, Pn, 128, generator=g) * 0.7).to(cuda).bfloat16()
    Mb2 = (torch.randn(B, 256, 128, generator=g) * 0.25).to(cuda).bfloat16()
    blog2 = (torch.randn(B, 256, generator=g) * 0.5).to(cuda)
    widths = torch.rand(B, 256, generator=g) + 0.1
    edges = torch.cumsum(torch.nn.functional.pad(widths / widths.sum(1, keepdim=True) * 9.999, (1, 0), value=0.001), 1)
    centres = (0.5 * (edges[:, :-1] + edges[:, 1:])).to(cuda).contiguous()
    gd = torch.randn(B, Pn, generator=g).to(cuda)
    return x3, Mb2, blog2, centres, gd


def _reference(x3, Mb2, blog2, centres, gd):
    x = x3.double().requires_grad_(True)
    M = Mb2.double().requires_grad_(True)
    bl = blog2.double().requires_grad_(True)
    c = centres.double().requires_grad_(True)
    l2 = torch.einsum("bpc,bkc->bpk", x, M) + bl[:, None, :]
    mx = l2.max(dim=2, keepdim=True).values
    e = torch.exp2(l2 - mx)
    s = e.sum(2, keepdim=True)
    depth = ((e / s) * c[:, None, :]).sum(2)
    (depth * gd.double()).sum().backward()
    ln2 = math.log(2.0)
    return dict(depth=depth.detach(), smax=mx.detach()[..., 0], ssum=s.detach()[..., 0], dx3=x.grad, dMb=M.grad / ln2, dcent=c.grad,
                dbias=(bl.grad * LOG2E).sum(0))


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("B,Pn", [(3, 128 * 121), (1, 128 * 5), (2, 128 * 300)])
def test_tail_fused_forward_and_backward(cuda, B, Pn):
    x3, Mb2, blog2, centres, gd = _case(cuda, B, Pn, seed=B * 1000 + Pn)
    ref = _reference(x3, Mb2, blog2, centres, gd)
    depth, smax, ssum = ops.tail_fused_fwd_bf16(x3, Mb2, blog2, centres)
    torch.cuda.synchronize()
    assert _rel(depth, ref["depth"]) < 2e-5
    assert float((smax.double() - ref["smax"]).abs().max()) < 1e-4
    assert _rel(ssum, ref["ssum"]) < 1e-4
    gbias = torch.full((256,), 3.0, device=cuda)
    dx3T, dMb, dcent = ops.tail_fused_bwd_bf16(x3, Mb2, blog2, centres, depth, smax, ssum, gd, grad_bias=gbias, beta=1.0)
    torch.cuda.synchronize()
    assert tuple(dx3T.shape) == (128, B * Pn)
    got_dx3 = dx3T.view(128, B, Pn).permute(1, 2, 0)
    e = dict(dx3=_rel(got_dx3, ref["dx3"]), dMb=_rel(dMb, ref["dMb"]), dcent=_rel(dcent, ref["dcent"]),
             dbias=_rel(gbias - 3.0, ref["dbias"]))
    print(f"B={B} P={Pn}: {e}")
    assert e["dx3"] < 1e-2 and e["dMb"] < 1e-2 and e["dcent"] < 1e-3 and e["dbias"] < 1e-2, e
    # deterministic: same bits on a second run (fixed-order reduction of the per-CTA partial accumulators)
    gbias2 = torch.full((256,), 3.0, device=cuda)
    dx3T2, dMb2, dcent2 = ops.tail_fused_bwd_bf16(x3, Mb2, blog2, centres, depth, smax, ssum, gd, grad_bias=gbias2, beta=1.0)
    torch.cuda.synchronize()
    assert torch.equal(dx3T, dx3T2) and torch.equal(dMb, dMb2) and torch.equal(dcent, dcent2) and torch.equal(gbias, gbias2)
