"""The rows either side of the head behind reference-shaped host code:

* :func:`tta_merge` — the clamp / un-mirror / average of the two validation passes (``modules/GraphBinsLM.py:159-183``);
* :class:`DepthMetrics` — ``MetricsPreprocess`` + the 16 torchmetrics objects of ``GraphBinsLM`` (8 metrics, each as a global
  ratio of sums and as a running average over batches; ``modules/GraphBinsLM.py:198-222``, ``metrics/*.py``) as ONE fused pass
  over the ground truth per batch, state = nine fp64 accumulators on the device (sum-reduced across ranks with one all-reduce);
* :func:`object_frontend` — per-image reversal + ``xyxy2xywh`` + the relative-size bucket of every detection
  (``modules/Yolov7Wrapper.py:118-123``, ``modules/ObjectLanguageStrategy.py:61-81``) on device, CSR in / lists out.

Everything numeric is a kernel of ``libobjcavit_b200.so`` (``csrc/eval_ops.cu``); there is no CPU path.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

from . import ops

REL_SIZE_SCALE = ("much smaller than", "smaller than", "a bit smaller than", "about the same size as", "a bit bigger than",
                  "bigger than", "much bigger than")   # modules/ObjectLanguageStrategy.py:23-31
METRIC_NAMES = ("abs_rel", "sq_rel", "rmse", "rmse_log", "log10", "acc_1", "acc_2", "acc_3")


def tta_merge(depth_pred: torch.Tensor, depth_pred_mirror: Optional[torch.Tensor], min_depth: float, max_depth: float):
    """``depth_pred_mirror`` is the head's output on ``image.flip(dims=[3])`` (NOT flipped back: the kernel reads it mirrored)."""
    return ops.tta_merge(depth_pred.contiguous().float(),
                         None if depth_pred_mirror is None else depth_pred_mirror.contiguous().float(), min_depth, max_depth)


def crop_rect(dataset: str, garg_crop: bool, eigen_crop: bool, H: int, W: int):
    """The evaluation rectangle of ``metrics/MetricsPreprocess.py:31-43`` as (y0, y1, x0, x1), or None."""
    if garg_crop:
        return int(0.40810811 * H), int(0.99189189 * H), int(0.03594771 * W), int(0.96405229 * W)
    if eigen_crop:
        if dataset == "kitti":
            return int(0.3324324 * H), int(0.91351351 * H), int(0.0359477 * W), int(0.96405229 * W)
        return 45, 471, 41, 601
    return None


def _values(s: Sequence[float]) -> Dict[str, float]:
    n = s[8]
    if n <= 0:
        return {k: float("nan") for k in METRIC_NAMES}
    return dict(abs_rel=s[0] / n, sq_rel=s[1] / n, rmse=math.sqrt(s[2] / n), rmse_log=math.sqrt(s[3] / n), log10=s[4] / n,
                acc_1=s[5] / n, acc_2=s[6] / n, acc_3=s[7] / n)


class DepthMetrics:
    """``args``-driven like the reference's metric objects: depth range and crop flags come from ``args[args.basic.dataset]``."""

    def __init__(self, args, device):
        ds = args[args.basic.dataset]
        self.dataset = args.basic.dataset
        self.lo, self.hi = float(ds.min_depth), float(ds.max_depth)
        self.garg, self.eigen = bool(ds.get("garg_crop")), bool(ds.get("eigen_crop"))
        self.acc = torch.zeros(9, device=device, dtype=torch.float64)
        self.batch_sums: List[torch.Tensor] = []

    def reset(self):
        self.acc.zero_()
        self.batch_sums = []

    def update(self, depth_pred: torch.Tensor, depth_gt: torch.Tensor, want_upsampled: bool = False):
        """``depth_pred`` [B,1,h,w] (e.g. :func:`tta_merge`'s result), ``depth_gt`` [B,1,H,W].  No host synchronisation."""
        H, W = depth_gt.shape[-2:]
        sums, up = ops.depth_metrics(depth_pred.contiguous().float(), depth_gt.contiguous().float(), self.lo, self.hi,
                                     crop_rect(self.dataset, self.garg, self.eigen, H, W), self.acc, want_upsampled)
        self.batch_sums.append(sums)
        return up

    def sync(self, group=None):
        """Sum the accumulators over the ranks (``dist_reduce_fx="sum"`` of the reference's metric states)."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.acc, op=dist.ReduceOp.SUM, group=group)

    def compute(self) -> Dict[str, float]:
        """-> the 8 metrics (global ratios of sums) and their ``*_ra`` running averages over this rank's batches."""
        out = _values(self.acc.cpu().tolist())
        if self.batch_sums:
            per = [_values(s) for s in torch.stack(self.batch_sums).cpu().tolist()]
            for k in METRIC_NAMES:
                # RMSELogRunningAvg averages the MEAN squared log difference (no square root): metrics/RMSELog.py
                vals = [p[k] ** 2 if k == "rmse_log" else p[k] for p in per]
                out[k + "_ra"] = sum(vals) / len(vals)
        return out


def object_frontend(xyxy_list: List[Optional[torch.Tensor]], reverse: bool = True):
    """Detections per image as corner boxes (``None`` / empty = no detection) -> (list of centre/size boxes or None, list of
    relative-size bucket tensors (index into ``REL_SIZE_SCALE``; -1 = lone object) or None)."""
    counts = [0 if t is None else int(t.shape[0]) for t in xyxy_list]
    if sum(counts) == 0:        # no detection in the whole batch (all None: what Yolov7Wrapper returns, common at batch size 1)
        return [None] * len(xyxy_list), [None] * len(xyxy_list)
    dev = next(t.device for t in xyxy_list if t is not None and t.shape[0] > 0)
    rows = torch.cat([t.to(device=dev, dtype=torch.float32) for t in xyxy_list if t is not None and t.shape[0] > 0], 0).contiguous()
    offs = [0]
    for c in counts:
        offs.append(offs[-1] + c)
    offsets = torch.tensor(offs, dtype=torch.int32).to(dev)
    xywh, bucket = ops.object_glue(rows, offsets, max(counts), reverse, len(REL_SIZE_SCALE))
    xs = [xywh[offs[i]:offs[i + 1]] if counts[i] else None for i in range(len(counts))]   # Yolov7Wrapper returns None when empty
    bs = [bucket[offs[i]:offs[i + 1]] if counts[i] else None for i in range(len(counts))]
    return xs, bs
