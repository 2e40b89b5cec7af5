"""Loss modules with the reference's names and call signatures (``losses/SILogLoss.py``, ``losses/BinsChamferLoss.py``,
``losses/LossWrapper.py``), running on the objcavit_b200 loss kernels.  Each returns a 0-d tensor; under autograd the
backward pass is the matching gradient kernel (w.r.t. ``depth_pred`` for SILog, ``bin_edges`` for the chamfer term)."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import ops


class _SILogFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, depth_pred, depth_gt, depth_mask, alph, lam):
        pred = depth_pred.detach().contiguous().float()
        gt = depth_gt.detach().contiguous().float()
        if depth_pred.requires_grad:
            loss, grad = ops.silog(pred, gt, depth_mask, alph, lam, want_grad=True)
            ctx.save_for_backward(grad)
        else:
            loss = ops.silog(pred, gt, depth_mask, alph, lam)
        return loss.view(())

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g, None, None, None, None   # scalar chain-rule factor of the caller (1 for a plain .backward())


class _ChamferFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, bin_edges, depth_gt, depth_mask):
        e = bin_edges.detach().contiguous().float()
        gt = depth_gt.detach().contiguous().float()
        if bin_edges.requires_grad:
            loss, grad = ops.bins_chamfer(e, gt, depth_mask, want_grad=True)
            ctx.save_for_backward(grad)
        else:
            loss = ops.bins_chamfer(e, gt, depth_mask)
        return loss.view(())

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g, None, None


class SILogLoss(nn.Module):
    """``losses/SILogLoss.py:10-56`` (alpha 10, lambda 0.85; the prediction is up-sampled to the GT size inside the kernel)."""

    def __init__(self, args=None):
        super().__init__()
        self.name = 'SILog'
        self.args = args
        self.alph = 10
        self.lam = 0.85

    def forward(self, depth_pred, depth_gt, depth_mask=None, interpolate=True, **kwargs):
        if not interpolate:
            assert depth_pred.shape[-2:] == depth_gt.shape[-2:], "interpolate=False needs a prediction at the GT size"
        return _SILogFn.apply(depth_pred, depth_gt, depth_mask, float(self.alph), float(self.lam))


class BinsChamferLoss(nn.Module):
    """``losses/BinsChamferLoss.py:14-37``: indexes the output tuple positionally (``output_named_tuple[1]`` = bin edges)."""

    def __init__(self, args=None):
        super().__init__()
        self.name = 'ChamferLoss'
        self.args = args

    def forward(self, depth_pred, depth_gt, depth_mask, output_named_tuple):
        return _ChamferFn.apply(output_named_tuple[1], depth_gt, depth_mask)


class LossWrapper(nn.Module):
    """``losses/LossWrapper.py:14-67``: ``args.loss.names`` / ``args.loss.coeffs`` driven sum of components."""

    def __init__(self, args):
        super().__init__()
        self.args = args
        possible_losses = ['silog', 'bins_chamfer']   # 'mse' of the reference is not on the hot path
        assert 'loss' in self.args, "Error: loss section missing from args file/dict."
        assert 'names' in self.args.loss, "Error: Loss names key not found in args.loss."
        assert len(self.args.loss.names) >= 1, "Error: no loss names given"
        assert all([name in possible_losses for name in self.args.loss.names]), "Error: unrecognised loss function"
        assert 'coeffs' in self.args.loss, "Error: Loss component coefficients array not found."
        assert len(self.args.loss.coeffs) == len(self.args.loss.names), \
            "Error: mismatched number of loss components and coefficients"
        self.loss_components = nn.ModuleList()
        for name in self.args.loss.names:
            self.loss_components.append(SILogLoss(self.args) if name == 'silog' else BinsChamferLoss(self.args))

    def forward(self, depth_pred, depth_gt, depth_mask, output_named_tuple):
        loss = 0
        for i, component in enumerate(self.loss_components):
            if isinstance(component, SILogLoss):   # the reference passes the tuple positionally as `interpolate` (truthy)
                term = component(depth_pred, depth_gt, depth_mask, True)
            else:
                term = component(depth_pred, depth_gt, depth_mask, output_named_tuple)
            loss = loss + self.args.loss.coeffs[i] * term
        return loss
