"""GPU: the training step of the head (config 5) — forward in grad mode, SILog + 0.1 chamfer, backward — against
(a) the committed fixtures generated from the UNMODIFIED reference modules in train() mode with dropout disabled
    (``oracle/make_golden.py:run_train_case``: loss, per-parameter gradient norm / sum / 8 probes), and
(b) full per-parameter gradients from autograd through the CPU oracle (``oracle/head.py``, ``train=True``).

Tolerance (written here because north_star states none for gradients): per parameter, max |g - g_ref| <= 2e-3 * max(|g_ref|_max,
1e-3 * largest |g| of the head) — the floor covers parameters whose true gradient is a rounding-noise cancellation (the block-2
object path of the 2-SACA head, |g| ~ 1e-9); typical measured errors are 1e-5.  Parameters the reference leaves without
gradient must come out exactly zero.
"""
import copy
import json
import os

import pytest
import torch

from helpers import GOLD, rel_err
from objcavit_b200.config import AttrDict, builtin_args
from objcavit_b200.losses import LossWrapper
from objcavit_b200.modules import build_head
from oracle import head as oh
from oracle import synth
from oracle.cases import TRAIN_CASES, case_cfg, case_gt, case_inputs, grad_probe_indices

pytestmark = pytest.mark.gpu

GRAD_TOL, GRAD_FLOOR = 2e-3, 1e-3


def _load(name):
    case = TRAIN_CASES[name]
    args = builtin_args(case["args"])
    with open(os.path.join(GOLD, f"manifest_{case['args']}.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=case["seed"])
    feats, objs, boxes = case_inputs(case)
    gt, mask = case_gt(case, args)
    return case, args, sd, feats, objs, boxes, gt, mask, torch.load(os.path.join(GOLD, f"{name}.pt"))


def _run_cuda(args, sd, feats, objs, boxes, gt, mask, cuda):
    head = build_head(args, precision="fp32").to(cuda).train()
    head.dropout = 0.0          # the fixtures were generated with the reference's dropout modules switched off (p = 0)
    head.load_state_dict(sd, strict=True)
    largs = AttrDict({"loss": {"names": ["silog", "bins_chamfer"], "coeffs": [1, 0.1]}})
    crit = LossWrapper(largs)
    if objs is None:
        out = head(feats.to(cuda))
    else:
        out = head(feats.to(cuda), [o.to(cuda) for o in objs], [None if b is None else b.to(cuda) for b in boxes])
    loss = crit(out.depth_pred, gt.to(cuda), mask.to(cuda), out)
    loss.backward()
    torch.cuda.synchronize()
    return head, out, loss


@pytest.mark.parametrize("name", sorted(TRAIN_CASES))
def test_training_gradients_match_reference_and_oracle(cuda, name):
    case, args, sd, feats, objs, boxes, gt, mask, gold = _load(name)
    head, out, loss = _run_cuda(args, sd, feats, objs, boxes, gt, mask, cuda)
    # (a) reference fixture
    assert abs(float(loss) - float(gold["loss"])) < 2e-5 * abs(float(gold["loss"]))
    assert rel_err(out.depth_pred.detach()[:, :, ::4, ::4], gold["depth"]) < 1e-4
    assert rel_err(out.bin_edges.detach(), gold["bin_edges"]) < 1e-4
    gmax = float(gold["gmax"])
    grads = {k: p.grad for k, p in head.named_parameters()}
    for k, fx in gold["grads"].items():
        g = grads[k]
        if fx is None:
            assert g is None or float(g.abs().max()) == 0.0, f"{k}: the reference gives no gradient"
            continue
        assert g is not None, f"{k}: no gradient produced"
        scale = max(float(fx["probes"].abs().max()), GRAD_FLOOR * gmax)
        idx = torch.tensor(grad_probe_indices(g.numel()))
        err = float((g.flatten().cpu()[idx] - fx["probes"]).abs().max()) / scale
        assert err < GRAD_TOL, f"{k}: probe error {err:.2e} vs the reference fixture"
        nscale = max(float(fx["norm"]), GRAD_FLOOR * gmax * g.numel() ** 0.5)
        assert abs(float(g.double().norm()) - float(fx["norm"])) < GRAD_TOL * nscale, f"{k}: gradient norm differs"
    # (b) full gradients through the oracle
    sd_g = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    o_depth, o_edges, _, _ = oh.head_forward(sd_g, dict(case_cfg(case, args), train=True), feats, objs, boxes)
    (oh.silog_loss(o_depth, gt, mask) + 0.1 * oh.bins_chamfer_loss(o_edges, gt, mask)).backward()
    worst = 0.0
    for k, g in grads.items():
        og = sd_g[k].grad
        if og is None:
            continue
        scale = max(float(og.abs().max()), GRAD_FLOOR * gmax)
        err = float((g.cpu() - og).abs().max()) / scale
        worst = max(worst, err)
        assert err < GRAD_TOL, f"{k}: gradient error {err:.2e} vs the oracle"
    print(f"{name}: worst per-parameter gradient error {worst:.2e}")


def test_training_forward_equals_inference_forward_without_padding(cuda):
    """AdaBins head: no padded rows exist, so the grad-mode forward must equal the no_grad forward bit for bit (same kernels)."""
    case, args, sd, feats, objs, boxes, gt, mask, gold = _load("train_adabins_small")
    head = build_head(args, precision="fp32").to(cuda)
    head.dropout = 0.0
    head.load_state_dict(sd, strict=True)
    a = head(feats.to(cuda))
    with torch.no_grad():
        b = head.eval()(feats.to(cuda))
    assert rel_err(a.depth_pred.detach(), b.depth_pred) < 1e-6 and rel_err(a.bin_edges.detach(), b.bin_edges) < 1e-6


def test_gradient_accumulation_and_flat_buffer(cuda):
    """Two backward passes accumulate; with a FlatGradBuffer attached the gradients land in the flat buffer."""
    from objcavit_b200.grad_sync import FlatGradBuffer
    case, args, sd, feats, objs, boxes, gt, mask, gold = _load("train_no_obj_sa_small")
    head, out, loss = _run_cuda(args, sd, feats, objs, boxes, gt, mask, cuda)
    once = {k: p.grad.clone() for k, p in head.named_parameters() if p.grad is not None}
    buf = FlatGradBuffer(head.parameters())
    for _ in range(2):
        out = head(feats.to(cuda), [o.to(cuda) for o in objs], [None if b is None else b.to(cuda) for b in boxes])
        LossWrapper(AttrDict({"loss": {"names": ["silog", "bins_chamfer"], "coeffs": [1, 0.1]}}))(
            out.depth_pred, gt.to(cuda), mask.to(cuda), out).backward()
    for (k, p), off in zip(head.named_parameters(), buf.offsets):
        assert p.grad.data_ptr() == buf.flat[off:].data_ptr(), f"{k}: .grad no longer a view of the flat buffer"
        if k in once:
            assert rel_err(p.grad, 2 * once[k]) < 1e-5 or float(once[k].abs().max()) < 1e-7


GRID_CASES = {
    "grid_random": dict(args="nyu_ocv_grid_random", B=2, h=176, w=192, n_obj=6, ragged=True, none_at=1, seed=21),
    "grid_random_roi_align": dict(args="kitti_ocv_grid_random_roi_align", B=2, h=128, w=320, n_obj=7, ragged=True, seed=22),
}


@pytest.mark.parametrize("name", sorted(GRID_CASES))
def test_positional_table_gradients_of_the_grid_strategies(cuda, name):
    """Training with the grid_random / grid_random_roi_align strategies: every parameter gradient, including the positional
    TABLE (adjoint of F.grid_sample / ps_roi_align incl. the quirks F13), against autograd through the fp64 oracle."""
    case = GRID_CASES[name]
    args = builtin_args(case["args"])
    with open(os.path.join(GOLD, f"manifest_{case['args']}.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=case["seed"])
    feats, objs, boxes = case_inputs(case)
    gt, mask = case_gt(case, args)
    head, out, loss = _run_cuda(args, sd, feats, objs, boxes, gt, mask, cuda)
    sd_g = {k: v.double().requires_grad_(True) for k, v in sd.items()}
    f64 = lambda lst: [None if t is None else t.double() for t in lst]
    o_depth, o_edges, _, _ = oh.head_forward(sd_g, dict(case_cfg(case, args), train=True), feats.double(), f64(objs), f64(boxes))
    o_loss = oh.silog_loss(o_depth, gt.double(), mask) + 0.1 * oh.bins_chamfer_loss(o_edges, gt.double(), mask)
    o_loss.backward()
    assert abs(float(loss.detach()) - float(o_loss.detach())) < 2e-5 * abs(float(o_loss.detach()))
    grads = {k: p.grad for k, p in head.named_parameters()}
    gmax = max(float(v.grad.abs().max()) for v in sd_g.values() if v.grad is not None)
    tkey = "objcavit.positional_encoder.positional_encodings"
    assert sd_g[tkey].grad is not None and float(sd_g[tkey].grad.abs().max()) > 0 and grads[tkey] is not None
    worst, bad = 0.0, []
    for k, g in grads.items():
        og = sd_g[k].grad
        if og is None:
            assert g is None or float(g.abs().max()) == 0.0, k
            continue
        assert g is not None, k
        err = float((g.cpu().double() - og).abs().max()) / max(float(og.abs().max()), GRAD_FLOOR * gmax)
        worst = max(worst, err)
        if err >= GRAD_TOL:
            bad.append(f"{k}: {err:.2e}")
    print(f"{name}: worst per-parameter gradient error {worst:.2e}; table gradient max {float(grads[tkey].abs().max()):.3e}")
    assert not bad, "; ".join(bad)
    # deterministic scatter (integer fixed-point atomics)
    head2, _, _ = _run_cuda(args, sd, feats, objs, boxes, gt, mask, cuda)
    assert torch.equal(dict(head2.named_parameters())[tkey].grad, grads[tkey])
