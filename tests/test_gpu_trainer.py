"""GPU: the native training step (objcavit_b200/trainer.py) against torch: flat-buffer gradients equal the autograd-boundary
gradients, and clip + AdamW (+ OneCycle momentum cycling) equals ``clip_grad_norm_`` + ``torch.optim.AdamW`` + ``OneCycleLR``."""
import json
import os

import pytest
import torch

from helpers import GOLD, rel_err
from objcavit_b200 import ops
from objcavit_b200.config import builtin_args
from objcavit_b200.modules import build_head
from objcavit_b200.trainer import HeadTrainer, OneCycle
from oracle import synth
from oracle.cases import TRAIN_CASES, case_gt, case_inputs

pytestmark = pytest.mark.gpu


def test_onecycle_matches_torch():
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.AdamW([p], lr=0.000357)
    sch = torch.optim.lr_scheduler.OneCycleLR(opt, max_lr=0.000357, total_steps=50, cycle_momentum=True, base_momentum=0.85,
                                              max_momentum=0.95, div_factor=25, final_div_factor=100)
    mine = OneCycle(0.000357, 50, 25, 100)
    for i in range(50):
        lr, b1 = mine.at(i)
        assert abs(lr - opt.param_groups[0]["lr"]) < 1e-12 and abs(b1 - opt.param_groups[0]["betas"][0]) < 1e-12
        opt.step()
        sch.step()


def test_sumsq_and_adamw_against_torch(cuda):
    g = torch.Generator().manual_seed(0)
    n = 100003 * 4
    p0 = torch.randn(n, generator=g)
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref], lr=3.57e-4, weight_decay=0.1)
    sch = torch.optim.lr_scheduler.OneCycleLR(opt, max_lr=3.57e-4, total_steps=10, cycle_momentum=True, base_momentum=0.85,
                                              max_momentum=0.95, div_factor=25, final_div_factor=100)
    mine = OneCycle(3.57e-4, 10, 25, 100)
    p = p0.clone().to(cuda)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for step in range(1, 6):
        grad = torch.randn(n, generator=g) * (0.01 if step % 2 else 10.0)      # below / above the clipping threshold
        ref.grad = grad.clone()
        norm = torch.nn.utils.clip_grad_norm_([ref], 0.1 if step > 1 else 1e9)
        opt.step(); sch.step()
        gd = (grad * 2.0).to(cuda)                                              # "sum over 2 ranks", scaled back by 0.5
        ss = ops.sumsq(gd, 0.5)
        assert rel_err(ss.sqrt(), norm.view(1)) < 1e-5
        lr, b1 = mine.at(step - 1)
        ops.adamw_step(p, gd, m, v, lr, b1, 0.999, 1e-8, 0.1, step, grad_scale=0.5, grad_sumsq=ss if step > 1 else None,
                       max_norm=0.1)
        assert rel_err(p, ref.data) < 2e-6, step
    assert torch.equal(ops.sumsq(gd, 0.5), ops.sumsq(gd, 0.5))                  # deterministic


def test_trainer_step_equals_autograd_boundary_plus_torch_adamw(cuda):
    name = "train_learned_bbox_wh_small"
    case = TRAIN_CASES[name]
    args = builtin_args(case["args"])
    with open(os.path.join(GOLD, f"manifest_{case['args']}.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=case["seed"])
    feats, objs, boxes = case_inputs(case)
    gt, mask = case_gt(case, args)
    dev = lambda lst: [None if t is None else t.to(cuda) for t in lst]
    # torch side: the autograd boundary + torch's clip + AdamW
    from objcavit_b200.config import AttrDict
    from objcavit_b200.losses import LossWrapper
    a = build_head(args).to(cuda).train(); a.load_state_dict(sd); a.dropout = 0.0   # the two arms must see the same (no) masks
    opt = torch.optim.AdamW(a.parameters(), lr=3.57e-4, weight_decay=0.1)
    out = a(feats.to(cuda), dev(objs), dev(boxes))
    loss = LossWrapper(AttrDict({"loss": {"names": ["silog", "bins_chamfer"], "coeffs": [1, 0.1]}}))(
        out.depth_pred, gt.to(cuda), mask.to(cuda), out)
    loss.backward()
    torch.nn.utils.clip_grad_norm_([p for p in a.parameters() if p.grad is not None], 0.1)
    opt.step()
    # native step
    b = build_head(args).to(cuda).train(); b.load_state_dict(sd); b.dropout = 0.0
    tr = HeadTrainer(b, lr=3.57e-4, weight_decay=0.1, clip_grad_norm=0.1)
    l1, l2 = tr.step(feats.to(cuda), dev(objs), dev(boxes), gt.to(cuda), mask.to(cuda))
    torch.cuda.synchronize()
    assert abs(tr.loss_value(l1, l2) - float(loss)) < 1e-5 * abs(float(loss))
    pa = dict(a.named_parameters())
    for k, p in b.named_parameters():
        if pa[k].grad is None:       # torch skips parameters without gradient entirely (no weight decay either): F10
            assert torch.equal(p.data.cpu(), sd[k]), f"{k}: a parameter without gradient was modified"
            continue
        assert rel_err(p.data, pa[k].data) < 1e-5, k
    assert set(b.state_dict()) == set(sd)                                       # checkpoint surface unchanged by the flat buffers
    with torch.no_grad():                                                       # and inference still runs on the updated weights
        o2 = b.eval()(feats.to(cuda), dev(objs), dev(boxes))
    assert torch.isfinite(o2.depth_pred).all()


def _fullsize_inputs(cuda, seed=3):
    """BASELINE.json config 5 at its full size: batch 8, features 128x208x272, GT 416x544, 32 boxes per image."""
    B, H, W, NOBJ = 8, 208, 272, 32
    x = synth.make_features(B, H, W, seed=seed).to(cuda)
    objs, boxes = synth.make_objects(B, H, W, NOBJ, seed=seed)
    gt, mask = synth.make_gt(B, H, W, 10.0, seed=seed)
    return x, [o.to(cuda) for o in objs], [b.to(cuda) for b in boxes], gt.to(cuda), mask.to(cuda)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_fullsize_training_steps_are_deterministic_and_reduce_the_loss(cuda, precision):
    """Size-independent properties at the full config-5 size: two runs from the same seed give bit-identical parameters (every
    reduction is fixed-order, the weight-gradient stream writes disjoint buffers, dropout masks are counter-based), all values
    stay finite, and six optimizer steps on one batch lower the loss."""
    args = builtin_args("nyu_ocv_learned_bbox_wh")
    with open(os.path.join(GOLD, "manifest_nyu_ocv_learned_bbox_wh.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=5)
    x, objs, boxes, gt, mask = _fullsize_inputs(cuda)
    runs = []
    for rep in range(2):
        head = build_head(args, precision=precision).to(cuda).train()
        head.load_state_dict(sd)
        tr = HeadTrainer(head, lr=3.57e-4, weight_decay=0.1, clip_grad_norm=0.1)
        torch.manual_seed(123)
        losses = []
        for _ in range(6):
            l1, l2 = tr.step(x, list(objs), list(boxes), gt, mask)
            losses.append(tr.loss_value(l1, l2))
        runs.append((losses, tr.flat_params.clone()))
    assert runs[0][0] == runs[1][0] and torch.equal(runs[0][1], runs[1][1])
    assert all(l == l and abs(l) < 1e6 for l in runs[0][0]) and bool(torch.isfinite(runs[0][1]).all())
    assert runs[0][0][-1] < runs[0][0][0], runs[0][0]


def test_fullsize_bf16_pixel_path_loss_matches_fp32_path(cuda):
    args = builtin_args("nyu_ocv_learned_bbox_wh")
    with open(os.path.join(GOLD, "manifest_nyu_ocv_learned_bbox_wh.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=5)
    x, objs, boxes, gt, mask = _fullsize_inputs(cuda)
    vals = {}
    for precision in ("fp32", "bf16"):
        head = build_head(args, precision=precision).to(cuda).train()
        head.dropout = 0.0
        head.load_state_dict(sd)
        tr = HeadTrainer(head, lr=3.57e-4, weight_decay=0.1, clip_grad_norm=0.1)
        l1, l2 = tr.forward_backward(x, list(objs), list(boxes), gt, mask)
        torch.cuda.synchronize()
        vals[precision] = (float(l1), float(l2), tr.last_outputs[0].clone(), float(tr.grads.flat.double().norm()))
    a, b = vals["fp32"], vals["bf16"]
    assert abs(a[0] - b[0]) < 2e-2 * abs(a[0]) and abs(a[1] - b[1]) < 2e-2 * abs(a[1])
    assert rel_err(b[2], a[2]) < 2e-2                                   # north_star: bf16 path within 2e-2 relative
    assert abs(a[3] - b[3]) < 5e-2 * a[3]                               # size of the whole gradient


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_graph_replay_of_the_step_is_bit_identical_to_eager_steps(cuda, precision):
    """``HeadTrainer(graph=True)`` captures zero-grads + forward + losses + backward into a CUDA graph after two eager steps and
    replays it: same kernels in the same order on the same streams, so with dropout off the parameters after six steps (two
    eager, four replayed, fresh inputs copied into the static buffers each step) equal the eager trainer's bit for bit."""
    args = builtin_args("nyu_ocv_learned_bbox_wh")
    with open(os.path.join(GOLD, "manifest_nyu_ocv_learned_bbox_wh.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=5)
    batches = [_fullsize_inputs(cuda, seed=3), _fullsize_inputs(cuda, seed=4)]
    runs = []
    for graph in (False, True):
        head = build_head(args, precision=precision).to(cuda).train()
        head.dropout = 0.0
        head.load_state_dict(sd)
        tr = HeadTrainer(head, lr=3.57e-4, weight_decay=0.1, clip_grad_norm=0.1, graph=graph)
        losses = []
        for i in range(6):
            x, objs, boxes, gt, mask = batches[i % 2]
            l1, l2 = tr.step(x, list(objs), list(boxes), gt, mask)
            losses.append(tr.loss_value(l1, l2))
        runs.append((losses, tr.flat_params.clone(), tr.last_outputs[0].clone()))
        if graph:
            assert len(tr._graphs) == 1
            tr.enable_graph(False)
    assert runs[0][0] == runs[1][0], (runs[0][0], runs[1][0])
    assert torch.equal(runs[0][1], runs[1][1]) and torch.equal(runs[0][2], runs[1][2])


def test_graph_replay_draws_new_dropout_masks_every_step_and_is_reproducible(cuda):
    """Dropout on: the site seeds are frozen into the graph, the device-side salt is what changes — replaying the SAME batch with
    the learning rate at 0 must give different losses from step to step (new masks), two runs from the same torch seed must agree
    bit for bit, and the loss must stay close to the eager trainer's (same distribution of masks)."""
    args = builtin_args("nyu_ocv_learned_bbox_wh")
    with open(os.path.join(GOLD, "manifest_nyu_ocv_learned_bbox_wh.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=5)
    x, objs, boxes, gt, mask = _fullsize_inputs(cuda)
    runs = []
    for graph in (True, True, False):
        head = build_head(args, precision="bf16").to(cuda).train()
        head.load_state_dict(sd)
        torch.manual_seed(77)
        tr = HeadTrainer(head, lr=0.0, weight_decay=0.0, clip_grad_norm=0.1, graph=graph)
        losses = []
        for _ in range(6):
            l1, l2 = tr.step(x, list(objs), list(boxes), gt, mask)
            losses.append(tr.loss_value(l1, l2))
        runs.append(losses)
        if graph:
            tr.enable_graph(False)
    assert runs[0] == runs[1]                                             # reproducible
    assert len(set(runs[0][2:])) == 4, runs[0]                            # replays: a new mask stream every step
    mean = lambda v: sum(v) / len(v)
    assert abs(mean(runs[0]) - mean(runs[2])) < 0.05 * abs(mean(runs[2])), (runs[0], runs[2])


def test_branch_streams_do_not_change_a_single_bit_of_the_step(cuda):
    """Every scheduling change of the step (object / pixel / loss / regressor branches, the per-image GEMM queues, the forward
    hoists) moves launches between streams without touching a kernel: three optimizer steps with dropout on, from the same
    torch seed, must leave exactly the parameters of the single-stream step (``BRANCH_OBJECTS = 0``)."""
    import objcavit_b200.engine_train as et
    args = builtin_args("nyu_ocv_learned_bbox_wh")
    with open(os.path.join(GOLD, "manifest_nyu_ocv_learned_bbox_wh.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=5)
    x, objs, boxes, gt, mask = _fullsize_inputs(cuda)
    saved = (et.BRANCH_OBJECTS, et.UNFOLD_IN_FORWARD, et.PIXEL_QUEUES)
    runs = []
    try:
        for branches, hoist, queues in ((0, False, 1), saved):
            et.BRANCH_OBJECTS, et.UNFOLD_IN_FORWARD, et.PIXEL_QUEUES = branches, hoist, queues
            head = build_head(args, precision="bf16").to(cuda).train()
            head.load_state_dict(sd)
            torch.manual_seed(11)
            tr = HeadTrainer(head, lr=3.57e-4, weight_decay=0.1, clip_grad_norm=0.1)
            losses = [tr.loss_value(*tr.step(x, list(objs), list(boxes), gt, mask)) for _ in range(3)]
            runs.append((losses, tr.flat_params.clone()))
    finally:
        et.BRANCH_OBJECTS, et.UNFOLD_IN_FORWARD, et.PIXEL_QUEUES = saved
    assert saved[0] != 0, "the default configuration is the branched one"
    assert runs[0][0] == runs[1][0] and torch.equal(runs[0][1], runs[1][1])
