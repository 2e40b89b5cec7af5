"""TEST INFRASTRUCTURE ONLY — CPU restatement of the rows either side of the head: the evaluation
epilogue, the metrics and the object front-end glue.  Plain torch / python, each function citing the reference lines it follows.
Pinned against the reference's own code by ``oracle/make_golden.py:pin_eval`` (``MetricsPreprocess`` and the metric classes are
imported unmodified through a ``torchmetrics`` stub; ``get_single_relative_size_clause`` through an ``nltk`` stub).
"""
from __future__ import annotations

import math
from typing import List, Optional

import numpy as np
import torch
import torch.nn.functional as F

SUM_NAMES = ("abs_rel", "sq_rel", "sq_diff", "sq_log_diff", "abs_log10_diff", "d1", "d2", "d3", "count")
REL_SIZE_SCALE = ("much smaller than", "smaller than", "a bit smaller than", "about the same size as", "a bit bigger than",
                  "bigger than", "much bigger than")   # modules/ObjectLanguageStrategy.py:23-31


def tta_merge(depth, depth_mirror, min_depth: float, max_depth: float):
    """``modules/GraphBinsLM.py:159-183``: clamp both passes, un-mirror the second, average."""
    a = torch.clamp(depth, min=min_depth, max=max_depth)
    if depth_mirror is None:
        return a
    b = torch.clamp(depth_mirror.flip(dims=[3]), min=min_depth, max=max_depth)
    return 0.5 * (a + b)


def crop_rect(dataset: str, garg_crop: bool, eigen_crop: bool, H: int, W: int):
    """``metrics/MetricsPreprocess.py:31-43`` -> (y0, y1, x0, x1) or None."""
    if garg_crop:
        return int(0.40810811 * H), int(0.99189189 * H), int(0.03594771 * W), int(0.96405229 * W)
    if eigen_crop:
        if dataset == "kitti":
            return int(0.3324324 * H), int(0.91351351 * H), int(0.0359477 * W), int(0.96405229 * W)
        return 45, 471, 41, 601
    return None


def metrics_preprocess(depth_pred, depth_gt, min_depth: float, max_depth: float, crop=None):
    """``metrics/MetricsPreprocess.py:17-45`` -> (up-sampled + de-nan'd prediction, validity mask)."""
    p = F.interpolate(depth_pred, depth_gt.shape[-2:], mode="bilinear", align_corners=True)
    p = p.nan_to_num(nan=min_depth, posinf=max_depth, neginf=max_depth)
    mask = (depth_gt > min_depth) & (depth_gt <= max_depth)
    if crop is not None:
        ev = torch.zeros(depth_gt.shape[-2:], dtype=torch.bool)
        ev[crop[0]:crop[1], crop[2]:crop[3]] = True
        mask = mask & ev
    return p, mask


def metric_sums(pred_m, gt_m):
    """The accumulator updates of ``metrics/{AbsRel,SqRel,RMSE,RMSELog,Log10,AccThresh}.py`` (non-running-average classes) on
    pre-masked values -> float64 [9] in ``SUM_NAMES`` order."""
    p, t = pred_m.double(), gt_m.double()
    thr = torch.maximum(t / p, p / t)
    return torch.stack([
        (torch.abs(t - p) / t).sum(), (((t - p) ** 2) / t).sum(), ((t - p) ** 2).sum(),
        ((torch.log(t) - torch.log(p)) ** 2).sum(), torch.abs(torch.log10(t) - torch.log10(p)).sum(),
        (thr < 1.25).sum().double(), (thr < 1.25 ** 2).sum().double(), (thr < 1.25 ** 3).sum().double(),
        torch.tensor(float(t.numel()), dtype=torch.float64)])


def metrics_from_sums(s):
    """``compute()`` of the same classes -> dict."""
    n = float(s[8])
    return dict(abs_rel=float(s[0]) / n, sq_rel=float(s[1]) / n, rmse=math.sqrt(float(s[2]) / n), rmse_log=math.sqrt(float(s[3]) / n),
                log10=float(s[4]) / n, acc_1=float(s[5]) / n, acc_2=float(s[6]) / n, acc_3=float(s[7]) / n)


def object_glue(xyxy_list: List[Optional[torch.Tensor]], n_scale: int = 7, reverse: bool = True):
    """``modules/Yolov7Wrapper.py:118-123`` (reversed(det), xyxy2xywh) + the bucket arithmetic of
    ``modules/ObjectLanguageStrategy.py:61-81`` -> (list of xywh tensors, list of int bucket lists; -1 = lone object)."""
    xywhs, buckets = [], []
    for det in xyxy_list:
        x = torch.flip(det, dims=[0]) if reverse else det
        y = x.clone()
        y[:, 0] = (x[:, 0] + x[:, 2]) / 2
        y[:, 1] = (x[:, 1] + x[:, 3]) / 2
        y[:, 2] = x[:, 2] - x[:, 0]
        y[:, 3] = x[:, 3] - x[:, 1]
        xywhs.append(y)
        bl = []
        for j, o in enumerate(y):
            if len(y) <= 1:
                bl.append(-1)
                continue
            nxt = y[(j + 1) % len(y)]
            rel = math.log(o[2] * o[3] / (nxt[2] * nxt[3])) + 1
            rel /= 2
            rel *= (n_scale - 3)
            rel = np.round(rel) + 1
            bl.append(int(np.clip(rel, 0, n_scale - 1)))
        buckets.append(bl)
    return xywhs, buckets


# seeded inputs of the evaluation-row parity cases (shared by the golden generator and the tests)
EVAL_CASES = {
    "nyu_eigen": dict(dataset="nyu", hw=(240, 320), lo=0.001, hi=10.0, garg=False, eigen=True, seed=21),
    "kitti_garg": dict(dataset="kitti", hw=(88, 304), lo=0.001, hi=80.0, garg=True, eigen=False, seed=22),
    "nyu_nocrop": dict(dataset="nyu", hw=(30, 40), lo=0.001, hi=10.0, garg=False, eigen=False, seed=23),
}


def eval_case_inputs(tag: str):
    """-> (pass-1 depth, mirrored-pass depth, ground truth): values above max_depth, NaN and +-inf included on purpose."""
    c = EVAL_CASES[tag]
    g = torch.Generator().manual_seed(c["seed"])
    (h, w), hi = c["hw"], c["hi"]
    B, H, W = 2, 2 * h, 2 * w
    a = 0.2 + torch.rand(B, 1, h, w, generator=g) * hi * 1.1
    b = 0.2 + torch.rand(B, 1, h, w, generator=g) * hi * 1.1
    a[0, 0, 3, 5] = float("nan")
    b[1, 0, 7, 2] = float("inf")
    a[1, 0, 1, 1] = float("-inf")
    gt = torch.rand(B, 1, H, W, generator=g) * hi * 1.05
    gt[torch.rand(B, 1, H, W, generator=g) < 0.3] = 0.0
    return a, b, gt


def glue_case_inputs():
    """Corner boxes of 3 ragged images + one lone detection."""
    from oracle import synth
    _, boxes = synth.make_objects(3, 240, 320, 9, seed=77, ragged=True)
    xyxy = [torch.stack([c[:, 0] - c[:, 2] / 2, c[:, 1] - c[:, 3] / 2, c[:, 0] + c[:, 2] / 2, c[:, 1] + c[:, 3] / 2], 1) for c in boxes]
    xyxy.append(torch.tensor([[10., 20., 50., 80.]]))
    return xyxy
