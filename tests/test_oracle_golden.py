"""CPU: the oracle restatement against the fixtures produced by the REAL reference (oracle/make_golden.py)."""
import os

import pytest
import torch

from helpers import GOLD, load_case, rel_err
from oracle import head as oh
from oracle import synth
from oracle.cases import CASES


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_matches_reference_fixture(name):
    case, args, sd, feats, objs, boxes, gold, cfg = load_case(name)
    depth, edges, widths, ram = oh.head_forward(sd, cfg, feats, objs, boxes)
    assert rel_err(depth, gold["depth"]) < 2e-5
    assert rel_err(edges, gold["bin_edges"]) < 2e-5
    assert rel_err(widths, gold["bin_widths_normed"]) < 2e-5
    assert rel_err(ram[:, ::8, ::8, ::8], gold["ram_sample"]) < 5e-5


def test_grid_ops_fixture():
    g = torch.load(os.path.join(GOLD, "grid_ops.pt"))
    table = g["table"]
    for key, c in g["cases"].items():
        h, w = (int(v) for v in key.split("_")[-1].split("x"))
        grid, gh, gw = oh._grid_for(table, h, w)
        if key.startswith("roi_"):
            scale = 1 / 32 if "_obj_" in key else 1 / 16
            got, _ = oh.ps_roi_align_1x1(grid, c["xyxy"], scale)
            want = c["want"]
            assert torch.equal(torch.isnan(got), torch.isnan(want))
            m = ~torch.isnan(want)
            assert torch.equal(got[m], want[m])          # bit-exact restatement of torchvision's kernel
        else:
            got, idx, _ = oh.grid_sample_points(grid, c["pts"][:, 0], c["pts"][:, 1])
            assert float((got - c["want"]).abs().max()) < 5e-6


def test_f5_quirk_object_content_invariance():
    """with one SACA block and N_max <= S/2 the head output does not depend on the
    object embeddings/boxes, only on the object counts."""
    case, args, sd, feats, objs, boxes, gold, cfg = load_case("ocv_learned_small")
    d0, *_ = oh.head_forward(sd, cfg, feats, objs, boxes)
    objs2 = [torch.randn_like(o) for o in objs]
    boxes2 = [b + 3.0 for b in boxes]
    d1, *_ = oh.head_forward(sd, cfg, feats, objs2, boxes2)
    assert torch.equal(d0, d1)


def test_losses_fixture():
    g = torch.load(os.path.join(GOLD, "losses.pt"))
    gt, mask = synth.make_gt(2, 24, 32, 10.0)
    assert abs(float(oh.silog_loss(g["depth"], gt, mask)) - float(g["silog"])) < 1e-5
    assert abs(float(oh.bins_chamfer_loss(g["edges"], gt, mask)) - float(g["chamfer"])) < 1e-6
