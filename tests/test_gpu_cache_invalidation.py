"""GPU: the derived bf16 / packed weight copies follow the parameters (the caches were keyed on ``p.data._version``,
which is always 0, so ``load_state_dict`` / optimizer steps after the first bf16 forward kept serving the old weights)."""
import json
import os

import pytest
import torch

from helpers import GOLD, rel_err
from objcavit_b200 import engine
from objcavit_b200.config import builtin_args
from objcavit_b200.modules import build_head
from oracle import synth

pytestmark = pytest.mark.gpu
NAME = "nyu_ocv_learned_bbox_wh"


def _sd(seed):
    with open(os.path.join(GOLD, f"manifest_{NAME}.json")) as f:
        return synth.make_state_dict(json.load(f), seed=seed)


def _inputs(cuda):
    feats = synth.make_features(2, 176, 192, seed=41).to(cuda)
    objs, boxes = synth.make_objects(2, 176, 192, 5, seed=41)
    return feats, [o.to(cuda) for o in objs], [b.to(cuda) for b in boxes]


def _run(head, feats, objs, boxes):
    with torch.no_grad():
        out = head(feats, list(objs), boxes)
    torch.cuda.synchronize()
    return out.depth_pred.clone(), out.bin_edges.clone()


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_load_state_dict_after_first_forward_is_seen(cuda, precision):
    feats, objs, boxes = _inputs(cuda)
    head = build_head(builtin_args(NAME), precision=precision).to(cuda).eval()
    head.load_state_dict(_sd(0), strict=True)
    d0, _ = _run(head, feats, objs, boxes)
    head.load_state_dict(_sd(1), strict=True)              # same storages, new values: only the version counters move
    d1, e1 = _run(head, feats, objs, boxes)
    fresh = build_head(builtin_args(NAME), precision=precision).to(cuda).eval()
    fresh.load_state_dict(_sd(1), strict=True)
    dw, ew = _run(fresh, feats, objs, boxes)
    assert not torch.equal(d0, d1), "the second checkpoint was ignored (stale packed weights)"
    assert torch.equal(d1, dw) and torch.equal(e1, ew)


def test_in_place_updates_and_optimizer_steps_are_seen(cuda):
    feats, objs, boxes = _inputs(cuda)
    head = build_head(builtin_args(NAME), precision="bf16").to(cuda).eval()
    head.load_state_dict(_sd(0), strict=True)
    d0, _ = _run(head, feats, objs, boxes)
    with torch.no_grad():
        head.conv_out[0].bias.add_(torch.linspace(-1, 1, 256, device=cuda))       # in-place through the Parameter
    d1, _ = _run(head, feats, objs, boxes)
    assert not torch.equal(d0, d1)
    # a torch optimizer step through the autograd boundary (periodic validation during torch-optimizer training)
    opt = torch.optim.SGD(head.parameters(), lr=0.05)
    head.train()
    head.dropout = 0.0
    out = head(feats, list(objs), boxes)
    out.depth_pred.mean().backward()
    opt.step()
    head.eval()
    d2, e2 = _run(head, feats, objs, boxes)
    fresh = build_head(builtin_args(NAME), precision="bf16").to(cuda).eval()
    fresh.load_state_dict({k: v.detach().clone() for k, v in head.state_dict().items()}, strict=True)
    dw, ew = _run(fresh, feats, objs, boxes)
    assert not torch.equal(d1, d2)
    assert torch.equal(d2, dw) and torch.equal(e2, ew)
    # writes torch cannot see (p.data) need the explicit hook
    head.conv_out[0].bias.data.zero_()
    head.invalidate_caches()
    d3, _ = _run(head, feats, objs, boxes)
    assert not torch.equal(d2, d3)


def test_offsets_cache_generations_keep_live_tensors(cuda):
    """The CSR offsets cache rotates generations instead of clearing: a tensor that is still referenced (a captured graph owns
    its offsets) keeps its contents after many other signatures went through."""
    rows = torch.zeros(8, 4, device=cuda)
    first = engine.ObjectBatch(rows, [3, 5])
    keep = first.offsets
    want = keep.cpu().clone()
    engine.offsets_capture_refs = refs = []
    try:
        owned = engine.ObjectBatch(rows, [1, 7]).offsets
    finally:
        engine.offsets_capture_refs = None
    assert any(t is owned for t in refs)
    for i in range(3 * engine._OFFSETS_GENERATION):
        engine.ObjectBatch(torch.zeros(i % 7 + 2, 4, device=cuda), [1, i % 7 + 1] if i % 7 else [2])
        engine.ObjectBatch(torch.zeros(i + 2, 4, device=cuda), [i + 2])
    torch.cuda.synchronize()
    assert len(engine._OFFSETS_CACHE) <= engine._OFFSETS_GENERATION and len(engine._OFFSETS_OLD) <= engine._OFFSETS_GENERATION
    assert torch.equal(keep.cpu(), want) and owned.cpu().tolist() == [0, 1, 8]
    again = engine.ObjectBatch(rows, [3, 5]).offsets
    assert again.cpu().tolist() == [0, 3, 8]
