"""GPU: SILog and BinsChamfer kernels (forward + gradients w.r.t. the head outputs) against the CPU oracle
(``oracle/head.py``; gradients by autograd through the oracle) and the committed loss fixture."""
import os
from collections import namedtuple

import pytest
import torch

from helpers import GOLD, rel_err
from objcavit_b200 import ops
from objcavit_b200.config import AttrDict
from objcavit_b200.losses import BinsChamferLoss, LossWrapper, SILogLoss
from oracle import head as oh
from oracle import synth

pytestmark = pytest.mark.gpu


def _case(B, h, w, H, W, seed, max_depth=10.0, sparse=0.2, K=256):
    g = torch.Generator().manual_seed(seed)
    pred = 0.5 + torch.rand(B, 1, h, w, generator=g) * (max_depth - 1.0)
    gt = 0.5 + torch.rand(B, 1, H, W, generator=g) * (max_depth - 0.5)
    gt[torch.rand(B, 1, H, W, generator=g) < sparse] = 0.0
    mask = gt > 0.001
    widths = torch.rand(B, K, generator=g) + 0.1
    widths = widths / widths.sum(1, keepdim=True) * (max_depth - 0.001)
    edges = torch.cumsum(torch.cat([torch.full((B, 1), 0.001), widths], 1), 1)
    return pred, gt, mask, edges


@pytest.mark.parametrize("B,h,w,H,W,sparse", [(2, 24, 32, 48, 64, 0.2), (3, 13, 17, 26, 34, 0.7), (2, 20, 20, 20, 20, 0.0),
                                              (1, 9, 11, 31, 23, 0.3)])
def test_silog_forward_backward(cuda, B, h, w, H, W, sparse):
    pred, gt, mask, _ = _case(B, h, w, H, W, seed=B * 100 + h, sparse=sparse)
    p = pred.clone().requires_grad_(True)
    want = oh.silog_loss(p, gt, mask)
    want.backward()
    loss, grad = ops.silog(pred.to(cuda), gt.to(cuda), mask.to(cuda), want_grad=True)
    assert rel_err(loss.view(()), want.detach()) < 1e-5
    assert rel_err(grad, p.grad) < 1e-4
    # no mask
    want2 = oh.silog_loss(pred, gt.clamp_min(0.3), None)
    assert rel_err(ops.silog(pred.to(cuda), gt.clamp_min(0.3).to(cuda), None).view(()), want2) < 1e-5


@pytest.mark.parametrize("B,H,W,sparse,maxd", [(2, 48, 64, 0.2, 10.0), (3, 40, 56, 0.9, 80.0), (1, 16, 16, 0.0, 10.0)])
def test_bins_chamfer_forward_backward(cuda, B, H, W, sparse, maxd):
    _, gt, mask, edges = _case(B, 8, 8, H, W, seed=7 + B, max_depth=maxd, sparse=sparse)
    e = edges.clone().requires_grad_(True)
    want = oh.bins_chamfer_loss(e, gt, mask)
    want.backward()
    brute = oh.bins_chamfer_loss_bruteforce(edges, gt, mask)
    loss, grad = ops.bins_chamfer(edges.to(cuda), gt.to(cuda), mask.to(cuda), want_grad=True)
    assert rel_err(loss.view(()), want.detach()) < 1e-5 and rel_err(loss.view(()), brute) < 1e-5
    assert rel_err(grad, e.grad) < 1e-4
    # deterministic: integer atomics + fixed-order partials
    loss2, grad2 = ops.bins_chamfer(edges.to(cuda), gt.to(cuda), mask.to(cuda), want_grad=True)
    assert torch.equal(loss, loss2) and torch.equal(grad, grad2)


def test_loss_fixture_and_wrapper(cuda):
    """The committed fixture (SILog from the REAL reference module; chamfer from the restated formula) through the
    reference-shaped modules, and LossWrapper's 1*silog + 0.1*chamfer with autograd into both head outputs."""
    g = torch.load(os.path.join(GOLD, "losses.pt"))
    from oracle import synth
    depth, edges = g["depth"], g["edges"]
    gt, mask = synth.make_gt(2, 24, 32, 10.0)     # the seeded GT oracle/make_golden.py used for the fixture
    RT = namedtuple("ReturnType", ["depth_pred", "bin_edges"])
    out = RT(depth.to(cuda).requires_grad_(True), edges.to(cuda).requires_grad_(True))
    s = SILogLoss(None)(out.depth_pred, gt.to(cuda), mask.to(cuda))
    c = BinsChamferLoss(None)(out.depth_pred, gt.to(cuda), mask.to(cuda), out)
    assert rel_err(s.detach(), g["silog"]) < 1e-5 and rel_err(c.detach(), g["chamfer"]) < 1e-5
    args = AttrDict({"loss": {"names": ["silog", "bins_chamfer"], "coeffs": [1, 0.1]}})
    total = LossWrapper(args)(out.depth_pred, gt.to(cuda), mask.to(cuda), out)
    total.backward()
    dp = depth.clone().requires_grad_(True)
    de = edges.clone().requires_grad_(True)
    want = oh.silog_loss(dp, gt, mask) + 0.1 * oh.bins_chamfer_loss(de, gt, mask)
    want.backward()
    assert rel_err(total.detach(), want.detach()) < 1e-5
    assert rel_err(out.depth_pred.grad, dp.grad) < 1e-4 and rel_err(out.bin_edges.grad, de.grad) < 1e-4


def test_bins_chamfer_edge_cases(cuda):
    """An image without valid ground truth (contributes 0; the batch mean still divides by B), constant ground truth, one valid pixel
    beyond the last bin: kernel == the three CPU statements of the loss (tests/test_oracle_golden.py), gradient == autograd."""
    g = torch.Generator().manual_seed(11)
    widths = torch.rand(4, 256, generator=g) + 0.05
    edges, _ = oh.bin_edges_and_centres(widths / widths.sum(1, keepdim=True), 0.001, 10.0)
    gt, _ = synth.make_gt(4, 24, 32, 10.0, seed=5)
    gt[1] = 0.0
    gt[2] = torch.where(gt[2] > 0, torch.full_like(gt[2], 3.25), gt[2])
    gt[3] = 0.0
    gt[3, 0, 5, 7] = 25.0
    mask = gt > 0.001
    e = edges.clone().requires_grad_(True)
    want = oh.bins_chamfer_loss(e, gt, mask)
    want.backward()
    loss, grad = ops.bins_chamfer(edges.to(cuda), gt.to(cuda), mask.to(cuda), want_grad=True)
    assert rel_err(loss.view(()), want.detach()) < 1e-5
    assert abs(float(loss) - oh.bins_chamfer_loss_np(edges, gt, mask)) < 1e-5 * abs(float(loss))
    assert rel_err(grad, e.grad) < 1e-4
    assert float(grad[1].abs().max()) == 0.0          # the empty image has no gradient
