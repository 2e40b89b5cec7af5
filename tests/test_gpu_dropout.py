"""GPU: dropout of the training step (the reference trains with nn.TransformerEncoderLayer's default dropout 0.1 at four sites
per encoder layer).

torch's random stream cannot be reproduced by another kernel, so parity is stated the only way the domain allows:
(1) the masks themselves: counter-based, deterministic per (seed, element), keep-rate 1-p, scale 1/(1-p), independent across sites;
(2) forward AND every parameter gradient against autograd through the CPU oracle with the SAME masks applied at the same four
    sites of every layer (the masks are read back from the dropout kernel, the arithmetic around them is the oracle's);
(3) p = 0 and eval() take the dropout-free path (bit-equal to it).
Tolerance for (2): 5e-3 of the per-parameter gradient scale, floor 1e-3 of the head's largest (tests/test_gpu_train.py uses 2e-3
without dropout; with dropout the block-2 object path of the 2-SACA head — a rounding-noise cancellation, |g| ~ 1e-9 — has been
measured at 3.7e-3 of the floor after a change of summation order in one GEMM; every other parameter stays below 3e-5).
"""
import json
import os

import pytest
import torch

from helpers import GOLD, rel_err
from objcavit_b200 import engine_train, ops
from objcavit_b200.config import AttrDict, builtin_args
from objcavit_b200.losses import LossWrapper
from objcavit_b200.modules import build_head
from objcavit_b200.modules.heads import draw_dropout_seed
from oracle import head as oh
from oracle import synth
from oracle.cases import TRAIN_CASES, case_cfg, case_gt, case_inputs

pytestmark = pytest.mark.gpu

GRAD_TOL, GRAD_FLOOR = 5e-3, 1e-3


def test_mask_statistics_and_determinism(cuda):
    n, p = 1 << 22, 0.1
    ones = torch.ones(n, device=cuda)
    a = ops.dropout(ones, p, 1234)
    assert torch.equal(a, ops.dropout(ones, p, 1234))                               # same (seed, index) -> same decision
    vals = torch.unique(a)
    assert vals.numel() == 2 and float(vals[0]) == 0.0 and abs(float(vals[1]) - 1.0 / (1.0 - p)) < 1e-6
    keep = float((a > 0).float().mean())
    assert abs(keep - (1 - p)) < 5 * (p * (1 - p) / n) ** 0.5                     # 5 sigma of a Bernoulli(1-p) mean
    b = ops.dropout(ones, p, 1235)                                                  # another site: independent decisions
    both = float(((a > 0) & (b > 0)).float().mean())
    assert abs(both - (1 - p) ** 2) < 1e-3
    # neighbouring elements are uncorrelated (a counter hash, not a stream)
    k = (a > 0).float() - (1 - p)
    assert abs(float((k[1:] * k[:-1]).mean())) < 1e-3
    # residual form and the in-place form
    x = torch.randn(n, device=cuda)
    r = torch.randn(n, device=cuda)
    y = ops.dropout(x, p, 99, residual=r)
    assert torch.equal(y, r + x * ops.dropout(ones, p, 99))
    x2 = x.clone()
    ops.dropout(x2, p, 99, residual=r, out=x2)
    assert torch.equal(x2, y)
    assert torch.equal(ops.dropout(x, 0.0, 5), x)


def test_attention_dropout_forward_and_backward(cuda):
    """The attention kernel drops probabilities AFTER the normaliser (softmax, then dropout, then @V) with the mask the
    elementwise kernel produces for index ((b*heads+h)*Sq+i)*Sk+j; its backward regenerates the same mask."""
    torch.manual_seed(3)
    B, Sq, Sk, H, E, p, seed = 2, 37, 21, 4, 128, 0.25, 777
    q, k, v = (torch.randn(B, s, E, device=cuda) for s in (Sq, Sk, Sk))
    mask = torch.zeros(B, Sk, dtype=torch.bool, device=cuda)
    mask[1, -5:] = True
    m8 = mask.to(torch.uint8)
    out = ops.mha_core(q, k, v, m8, H, drop=(p, seed))
    m = ops.dropout(torch.ones(B, H, Sq, Sk, device=cuda), p, seed)
    qd, kd, vd = (t.double().requires_grad_(True) for t in (q, k, v))
    d = E // H
    split = lambda t, s: t.view(B, s, H, d).transpose(1, 2)
    sc = (split(qd, Sq) / d ** 0.5) @ split(kd, Sk).transpose(-1, -2)
    sc = sc.masked_fill(mask[:, None, None, :], float("-inf"))
    ref = ((torch.softmax(sc, -1) * m.double()) @ split(vd, Sk)).transpose(1, 2).reshape(B, Sq, E)
    assert rel_err(out, ref.detach().float()) < 1e-5
    go = torch.randn(B, Sq, E, device=cuda)
    ref.backward(go.double())
    dq, dk, dv = ops.mha_core_bwd(q, k, v, out, go, m8, H, drop=(p, seed))[:3]
    for mine, r, nm in ((dq, qd.grad, "dq"), (dk, kd.grad, "dk"), (dv, vd.grad, "dv")):
        assert rel_err(mine, r.float()) < 2e-5, nm


def _setup(name):
    case = TRAIN_CASES[name]
    args = builtin_args(case["args"])
    with open(os.path.join(GOLD, f"manifest_{case['args']}.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=case["seed"])
    feats, objs, boxes = case_inputs(case)
    gt, mask = case_gt(case, args)
    return case, args, sd, feats, objs, boxes, gt, mask


class _SharedMasks:
    """The oracle's dropout hook: site k of the forward gets the mask the CUDA kernel generates for site k (same DropCtx walk)."""

    def __init__(self, p, seed, cuda):
        self.ctx, self.p, self.cuda, self.sites = engine_train.DropCtx(p, seed), p, cuda, 0

    def __call__(self, t):
        self.sites += 1
        m = ops.dropout(torch.ones(t.shape, device=self.cuda), self.p, self.ctx.site())
        return t * m.cpu().to(t.dtype)


@pytest.mark.parametrize("name", ["train_adabins_small", "train_learned_bbox_wh_small", "train_2saca_small"])
def test_training_with_dropout_matches_oracle_with_shared_masks(cuda, name):
    if name not in TRAIN_CASES:
        pytest.skip(f"{name}: not a committed training case")
    case, args, sd, feats, objs, boxes, gt, mask = _setup(name)
    p = 0.1
    head = build_head(args, precision="fp32").to(cuda).train()
    assert head.dropout == p                                                        # the reference's default
    head.load_state_dict(sd, strict=True)
    torch.manual_seed(11)
    seed = draw_dropout_seed()
    torch.manual_seed(11)                                                           # the head draws the same seed
    dev = lambda lst: [None if t is None else t.to(cuda) for t in lst]
    out = head(feats.to(cuda)) if objs is None else head(feats.to(cuda), dev(objs), dev(boxes))
    crit = LossWrapper(AttrDict({"loss": {"names": ["silog", "bins_chamfer"], "coeffs": [1, 0.1]}}))
    loss = crit(out.depth_pred, gt.to(cuda), mask.to(cuda), out)
    loss.backward()
    torch.cuda.synchronize()

    hook = _SharedMasks(p, seed, cuda)
    # the oracle runs in fp64: with only 2 x 129 gradient-carrying rows in the last AdaBins layer, one ReLU gate that flips
    # between two fp32 implementations (pre-activation ~ 0) moves a linear1 gradient column by ~1e-2 — measured with an fp32 oracle
    sd_g = {k: v.double().requires_grad_(True) for k, v in sd.items()}
    f64 = lambda lst: None if lst is None else [None if t is None else t.double() for t in lst]
    oh.DROPOUT_HOOK = hook
    try:
        o_depth, o_edges, _, _ = oh.head_forward(sd_g, dict(case_cfg(case, args), train=True), feats.double(), f64(objs), f64(boxes))
    finally:
        oh.DROPOUT_HOOK = None
    o_loss = oh.silog_loss(o_depth, gt.double(), mask) + 0.1 * oh.bins_chamfer_loss(o_edges, gt.double(), mask)
    o_loss.backward()
    assert hook.sites > 0 and hook.sites % 16 == 0                                 # 4 sites x 4 layers per encoder
    assert rel_err(out.depth_pred.detach(), o_depth.detach().float()) < 1e-4
    assert rel_err(out.bin_edges.detach(), o_edges.detach().float()) < 1e-4
    assert abs(float(loss.detach()) - float(o_loss.detach())) < 2e-5 * abs(float(o_loss))
    grads = {k: q.grad for k, q in head.named_parameters()}
    gmax = max(float(v.grad.abs().max()) for v in sd_g.values() if v.grad is not None)
    worst, bad = 0.0, []
    for k, g in grads.items():
        og = sd_g[k].grad
        if og is None:
            assert g is None or float(g.abs().max()) == 0.0, k
            continue
        err = float((g.cpu().double() - og).abs().max()) / max(float(og.abs().max()), GRAD_FLOOR * gmax)
        worst = max(worst, err)
        if err >= GRAD_TOL:
            bad.append(f"{k}: {err:.2e}")
    assert not bad, "gradient error vs the oracle under shared dropout masks: " + "; ".join(bad)
    print(f"{name}: dropout {p}, {hook.sites} sites, worst per-parameter gradient error {worst:.2e}")
    # dropout really happened: the dropout-free forward differs
    head.dropout = 0.0
    out0 = head(feats.to(cuda)) if objs is None else head(feats.to(cuda), dev(objs), dev(boxes))
    assert rel_err(out0.depth_pred.detach(), out.depth_pred.detach()) > 1e-4


def test_eval_mode_and_p0_skip_dropout(cuda):
    case, args, sd, feats, objs, boxes, gt, mask = _setup("train_adabins_small")
    head = build_head(args, precision="fp32").to(cuda)
    head.load_state_dict(sd, strict=True)
    x = feats.to(cuda)
    a = head.eval()(x).depth_pred.detach()                                          # grad mode, eval(): no dropout (as nn.Dropout)
    head.train()
    head.dropout = 0.0
    b = head(x).depth_pred.detach()
    assert torch.equal(a, b)
    head.dropout = 0.1
    torch.manual_seed(1)
    c = head(x).depth_pred.detach()
    torch.manual_seed(1)
    d = head(x).depth_pred.detach()
    e = head(x).depth_pred.detach()                                                 # next draw of the generator: other masks
    assert torch.equal(c, d) and not torch.equal(c, e) and not torch.equal(a, c)


def test_trainer_step_with_dropout_is_reproducible(cuda):
    from objcavit_b200.trainer import HeadTrainer
    case, args, sd, feats, objs, boxes, gt, mask = _setup("train_learned_bbox_wh_small")
    dev = lambda lst: [None if t is None else t.to(cuda) for t in lst]
    res = []
    for rep in range(2):
        h = build_head(args).to(cuda).train()
        h.load_state_dict(sd)
        tr = HeadTrainer(h, lr=3.57e-4, weight_decay=0.1, clip_grad_norm=0.1)
        torch.manual_seed(5)
        l1, l2 = tr.step(feats.to(cuda), dev(objs), dev(boxes), gt.to(cuda), mask.to(cuda))
        res.append((tr.loss_value(l1, l2), tr.flat_params.clone()))
    assert res[0][0] == res[1][0] and torch.equal(res[0][1], res[1][1])
    h0 = build_head(args).to(cuda).train()
    h0.load_state_dict(sd)
    h0.dropout = 0.0
    tr0 = HeadTrainer(h0, lr=3.57e-4, weight_decay=0.1, clip_grad_norm=0.1)
    l1, l2 = tr0.step(feats.to(cuda), dev(objs), dev(boxes), gt.to(cuda), mask.to(cuda))
    assert tr0.loss_value(l1, l2) != res[0][0]


def test_fused_dropout_kernels_of_the_layer_sequencer_equal_the_two_call_sequences(cuda):
    """``ocv_dropout_add_layernorm_f32`` == ``ocv_dropout_f32`` (+ residual) then ``ocv_add_layernorm_f32``, and
    ``ocv_dropout_act_bwd_f32`` == ``ocv_dropout_f32`` then ``ocv_act_bwd_f32``: same mask, same arithmetic, bit for bit."""
    from objcavit_b200 import ops
    g = torch.Generator().manual_seed(5)
    rows, E, F = 8 * 221, 128, 1024
    x = torch.randn(rows, E, generator=g).to(cuda)
    r = torch.randn(rows, E, generator=g).to(cuda)
    gamma, beta = torch.randn(E, generator=g).to(cuda), torch.randn(E, generator=g).to(cuda)
    for p, seed in ((0.1, 0x1234567890ABCDEF), (0.5, 7)):
        want_sum = ops.dropout(x, p, seed, residual=r)
        want_y = ops.add_layernorm(want_sum, None, gamma, beta)
        got_sum, got_y = ops.dropout_add_layernorm(x, r, gamma, beta, p, seed)
        assert torch.equal(got_sum, want_sum) and torch.equal(got_y, want_y)
        dy = torch.randn(rows, F, generator=g).to(cuda)
        h = torch.relu(torch.randn(rows, F, generator=g)).to(cuda)
        want = ops.act_bwd(ops.dropout(dy, p, seed), h, 1)
        assert torch.equal(ops.dropout_act_bwd(dy, h, 1, p, seed), want)
        # LayerNorm backward: data half + dropout in one kernel, parameter half as its own call
        dg0, db0 = torch.randn(E, generator=g).to(cuda), torch.randn(E, generator=g).to(cuda)
        dg1, db1 = dg0.clone(), db0.clone()
        want_dz = ops.layernorm_bwd(x, r, gamma, dg0, db0, beta=1.0)
        got_dz, got_dzd = ops.layernorm_bwd_dropout(x, r, gamma, p, seed, dg1, db1, beta=1.0)
        assert torch.equal(got_dz, want_dz) and torch.equal(got_dzd, ops.dropout(want_dz, p, seed))
        assert torch.equal(dg1, dg0) and torch.equal(db1, db0)
