"""GPU: the INNER modules train under autograd — the reference trains through ``self.objcavit(...)`` (``modules/GraphBins.py:107``)
and ``self.adaptive_bins_layer(...)`` (``modules/AdaBins.py:76``), so INTEGRATION.md §1's import swap (replace the two classes,
keep the reference's own ``conv_out`` / tail / loss in torch) must produce the reference's gradients.

Checked against autograd through the CPU oracle in fp64 (``oracle/head.py``, slow-path encoder semantics) with dropout off: outputs ``(bin_widths_normed, range_attention_maps)`` and the gradient of every parameter for a loss that uses BOTH outputs
through the reference's own tail restated in torch.  Tolerance: 5e-3 of max(|g|, 1e-3 largest |g|) per parameter — the test loss
((log depth)^2, mixed-sign pixel gradients) makes the conv3x3 bias gradient a cancelling fp32 sum over 34k pixels (measured 2.0e-3
against fp64; every other parameter is at 1e-5 .. 4e-4 as in tests/test_gpu_train.py, which keeps its 2e-3)."""
import json
import os

import pytest
import torch
import torch.nn.functional as F

from helpers import GOLD, rel_err
from objcavit_b200.config import builtin_args
from objcavit_b200.modules import ObjCAViT, mViT
from oracle import head as oh
from oracle import synth
from oracle.cases import TRAIN_CASES, case_cfg, case_inputs

pytestmark = pytest.mark.gpu
GRAD_TOL, GRAD_FLOOR = 5e-3, 1e-3


def _tail_loss(widths, ram, wout, bout):
    """The caller's side of the boundary, in plain torch on whatever device the tensors live on (AdaBins.py:77-87 + a scalar)."""
    out = torch.softmax(F.conv2d(ram, wout, bout), dim=1)
    bw = F.pad((10.0 - 0.001) * widths, (1, 0), value=0.001)
    edges = torch.cumsum(bw, dim=1)
    centres = 0.5 * (edges[:, :-1] + edges[:, 1:])
    depth = (out * centres.view(*centres.shape, 1, 1)).sum(1)
    return (depth.log() ** 2).mean() + 0.3 * (edges[:, 1:-1] ** 2).mean()


@pytest.mark.parametrize("name", ["train_adabins_small", "train_learned_bbox_wh_small", "train_2saca_small"])
def test_inner_modules_train_under_autograd(cuda, name):
    case = TRAIN_CASES[name]
    args = builtin_args(case["args"])
    with open(os.path.join(GOLD, f"manifest_{case['args']}.json")) as f:
        sd_head = synth.make_state_dict(json.load(f), seed=case["seed"])
    feats, objs, boxes = case_inputs(case)
    ada = args.model.name == "adabins"
    pre = "adaptive_bins_layer." if ada else "objcavit."
    sd = {k[len(pre):]: v for k, v in sd_head.items() if k.startswith(pre)}
    wout, bout = sd_head["conv_out.0.weight"], sd_head["conv_out.0.bias"]
    mod = (mViT(128, dim_out=256) if ada else ObjCAViT(args)).to(cuda).train()
    mod.dropout = 0.0
    mod.load_state_dict(sd, strict=True)
    wg, bg = wout.to(cuda).requires_grad_(True), bout.to(cuda).requires_grad_(True)
    if ada:
        widths, ram = mod(feats.to(cuda))
    else:
        lst = [o.to(cuda) for o in objs]
        widths, ram = mod(feats.to(cuda), lst, [None if b is None else b.to(cuda) for b in boxes])
        assert lst[0].shape[1] == 128          # F11: the caller's list now holds the embedded object tokens
    assert widths.requires_grad and ram.requires_grad
    loss = _tail_loss(widths, ram, wg, bg)
    loss.backward()
    torch.cuda.synchronize()
    # oracle
    sd_g = {k: v.double().requires_grad_(True) for k, v in sd.items()}       # fp64: the checker's own rounding stays out of it
    wo, bo = wout.double().requires_grad_(True), bout.double().requires_grad_(True)
    if ada:
        o_w, o_ram = oh.mvit_forward(sd_g, feats.double(), pre="")
    else:
        cfg = case_cfg(case, args)
        f64 = lambda lst: [None if t is None else t.double() for t in lst]
        o_w, o_ram = oh.objcavit_forward(sd_g, feats.double(), f64(objs), f64(boxes), cfg["strategy"], pre="",
                                         no_obj_sa=cfg["no_obj_sa"], use_2_saca=cfg["use_2_saca"], fast_path=False)
    assert rel_err(widths.detach(), o_w.detach()) < 1e-4 and rel_err(ram.detach(), o_ram.detach()) < 1e-4
    o_loss = _tail_loss(o_w, o_ram, wo, bo)
    o_loss.backward()
    assert abs(float(loss.detach()) - float(o_loss.detach())) < 2e-5 * abs(float(o_loss.detach()))
    assert rel_err(wg.grad, wo.grad.float()) < GRAD_TOL          # the caller's own parameters: torch autograd through our outputs
    gmax = max(float(v.grad.abs().max()) for v in sd_g.values() if v.grad is not None)
    worst = 0.0
    for k, p in mod.named_parameters():
        og = sd_g[k].grad
        if og is None or float(og.abs().max()) == 0.0:
            assert p.grad is None or float(p.grad.abs().max()) <= 1e-6 * gmax, f"{k}: the reference gives no gradient"
            continue
        assert p.grad is not None, f"{k}: no gradient produced"
        err = float((p.grad.cpu().double() - og).abs().max()) / max(float(og.abs().max()), GRAD_FLOOR * gmax)
        worst = max(worst, err)
        assert err < GRAD_TOL, f"{k}: gradient error {err:.2e} vs the oracle"
    print(f"{name}: module-surface worst per-parameter gradient error {worst:.2e}")


def test_only_one_output_used(cuda):
    """A loss on the range-attention maps alone (bin widths unused -> their upstream gradient is None) still trains."""
    args = builtin_args("nyu_adabins")
    mod = mViT(128, dim_out=256).to(cuda).train()
    mod.dropout = 0.0
    x = synth.make_features(1, 176, 192, seed=5).to(cuda)
    _, ram = mod(x)
    ram.square().mean().backward()
    g = mod.conv3x3.weight.grad
    assert g is not None and bool(torch.isfinite(g).all()) and float(g.abs().max()) > 0
    assert mod.regressor[0].weight.grad is None or float(mod.regressor[0].weight.grad.abs().max()) == 0.0
