"""Rows N2-N4 (evaluation epilogue, metrics, object glue): the oracle against the committed reference fixture (CPU) and the
kernels against the oracle + fixture (GPU).  Tolerances: index / bucket / mask work bit-exact; up-sampled depth 1e-6 relative
(same arithmetic order as torch's upsample_bilinear2d); metric sums 1e-5 relative (fp32 per-pixel terms, fp64 accumulation, a
different summation order than torch.sum)."""
import os

import pytest
import torch

from helpers import GOLD, rel_err
from oracle import eval as oe

FIX = torch.load(os.path.join(GOLD, "eval_rows.pt"))
KEYS = ("abs_rel", "sq_rel", "rmse", "rmse_log", "log10", "acc_1", "acc_2", "acc_3")


@pytest.mark.parametrize("tag", sorted(oe.EVAL_CASES))
def test_oracle_eval_rows_match_reference_fixture(tag):
    c, fx = oe.EVAL_CASES[tag], FIX[tag]
    a, b, gt = oe.eval_case_inputs(tag)
    final = oe.tta_merge(a, b, c["lo"], c["hi"])
    assert torch.equal(final[:, :, ::3, ::3].nan_to_num(-1.0), fx["final"].nan_to_num(-1.0))
    crop = oe.crop_rect(c["dataset"], c["garg"], c["eigen"], *gt.shape[-2:])
    assert crop == fx["crop"]
    up, mask = oe.metrics_preprocess(final, gt, c["lo"], c["hi"], crop)
    assert torch.equal(up[:, :, ::5, ::5], fx["upsampled"])
    sums = oe.metric_sums(up[mask], gt[mask])
    assert rel_err(sums, fx["sums"]) < 1e-12
    m = oe.metrics_from_sums(sums)
    assert rel_err(torch.tensor([m[k] for k in KEYS]), fx["metrics"]) < 2e-5


def test_oracle_object_glue_matches_reference_fixture():
    xywh, bucket = oe.object_glue(oe.glue_case_inputs())
    for x, fx in zip(xywh, FIX["glue"]["xywh"]):
        assert torch.equal(x, fx)
    assert bucket == FIX["glue"]["bucket"]


@pytest.mark.gpu
@pytest.mark.parametrize("tag", sorted(oe.EVAL_CASES))
def test_gpu_eval_epilogue_and_metrics(cuda, tag):
    from objcavit_b200 import evaluation as ev
    from objcavit_b200 import ops
    from objcavit_b200.config import AttrDict
    c, fx = oe.EVAL_CASES[tag], FIX[tag]
    a, b, gt = oe.eval_case_inputs(tag)
    lo, hi = c["lo"], c["hi"]
    final = ev.tta_merge(a.to(cuda), b.to(cuda), lo, hi)
    want_final = oe.tta_merge(a, b, lo, hi)
    assert torch.equal(final.cpu().nan_to_num(-1.0), want_final.nan_to_num(-1.0))              # clamp / flip / average: bit-exact
    assert torch.equal(ev.tta_merge(a.to(cuda), None, lo, hi).cpu().nan_to_num(-1.0), torch.clamp(a, lo, hi).nan_to_num(-1.0))
    args = AttrDict({"basic": {"dataset": c["dataset"]},
                     c["dataset"]: {"min_depth": lo, "max_depth": hi, "garg_crop": c["garg"], "eigen_crop": c["eigen"]}})
    dm = ev.DepthMetrics(args, cuda)
    up = dm.update(final, gt.to(cuda), want_upsampled=True)
    crop = oe.crop_rect(c["dataset"], c["garg"], c["eigen"], *gt.shape[-2:])
    want_up, mask = oe.metrics_preprocess(want_final, gt, lo, hi, crop)
    assert rel_err(up, want_up) < 1e-6
    assert rel_err(up.cpu()[:, :, ::5, ::5], fx["upsampled"]) < 1e-6
    want = oe.metric_sums(want_up[mask], gt[mask])
    got = dm.acc.cpu()
    assert float(got[8]) == float(want[8]) == float(fx["sums"][8])                                # valid-pixel count: the mask is bit-exact
    assert rel_err(got[:5], want[:5]) < 1e-5
    assert (got[5:8] - want[5:8]).abs().max() <= 2                                               # threshold counts: fp32 ratio vs fp64 at the boundary
    vals = dm.compute()
    assert rel_err(torch.tensor([vals[k] for k in KEYS]), fx["metrics"]) < 2e-5                  # vs the reference's metric classes
    # second batch: accumulators add up, running averages average; deterministic
    dm.update(final, gt.to(cuda))
    assert rel_err(dm.acc.cpu(), 2 * got) < 1e-12
    v2 = dm.compute()
    assert abs(v2["abs_rel_ra"] - vals["abs_rel"]) < 1e-12 and abs(v2["rmse_log_ra"] - vals["rmse_log"] ** 2) < 1e-12
    s1, _ = ops.depth_metrics(final, gt.to(cuda), lo, hi, crop)
    s2, _ = ops.depth_metrics(final, gt.to(cuda), lo, hi, crop)
    assert torch.equal(s1, s2)


@pytest.mark.gpu
def test_gpu_object_glue_bit_exact(cuda):
    from objcavit_b200 import evaluation as ev
    xyxy = oe.glue_case_inputs()
    want_xywh, want_bucket = oe.object_glue(xyxy)
    xs, bs = ev.object_frontend([t.to(cuda) for t in xyxy] + [None])
    assert xs[-1] is None and bs[-1] is None
    for x, b, wx, wb, fx in zip(xs, bs, want_xywh, want_bucket, FIX["glue"]["xywh"]):
        assert torch.equal(x.cpu(), wx) and torch.equal(x.cpu(), fx)
        assert b.cpu().tolist() == wb
    # the reference raises on a degenerate (zero-area) box: flagged, never a silent bucket
    bad = torch.tensor([[0., 0., 10., 10.], [5., 5., 5., 9.]], device=cuda)
    _, bb = ev.object_frontend([bad])
    # reversed order: [zero-area box, 10x10 box] -> ratio 0 (math.log raises: flagged) and ratio inf (log = inf -> top bucket)
    assert bb[0].cpu().tolist() == [-2 ** 31, 6]
