"""TEST INFRASTRUCTURE ONLY — seeded synthetic inputs and weights.

Weights are generated from a ``{state_dict key: shape}`` manifest (committed under
``tests/golden/manifest_*.json``, dumped from the real reference modules by
``oracle/make_golden.py``) so that the GPU box — which has no reference tree — regenerates exactly
the tensors the golden outputs were produced with.  Scales mimic torch's default initialisers
(U(±1/sqrt(fan_in)) weights, U(0,1) positional tables, LayerNorm gains around 1).
"""
from __future__ import annotations

import math
import zlib
from typing import Dict, List, Optional, Tuple

import torch


def _gen(seed: int, key: str) -> torch.Generator:
    return torch.Generator().manual_seed((seed * 1000003 + zlib.crc32(key.encode())) % (2 ** 63 - 1))


def make_state_dict(manifest: Dict[str, List[int]], seed: int = 0) -> Dict[str, torch.Tensor]:
    sd = {}
    for key in sorted(manifest):
        shape = tuple(manifest[key])
        g = _gen(seed, key)
        if key.endswith("positional_encodings"):
            t = torch.rand(shape, generator=g)
        elif len(shape) >= 2:
            fan_in = 1
            for s in shape[1:]:
                fan_in *= s
            a = 1.0 / math.sqrt(fan_in)
            t = (torch.rand(shape, generator=g) * 2 - 1) * a
        elif ".norm" in key and key.endswith("weight"):
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        else:
            t = 0.05 * torch.randn(shape, generator=g)
        sd[key] = t.float().contiguous()
    return sd


def make_features(B: int, h: int, w: int, seed: int = 1234, C: int = 128) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, C, h, w, generator=g)


def make_objects(B: int, h: int, w: int, n_obj: int, seed: int = 1234, ragged: bool = False,
                 none_at: Optional[int] = None) -> Tuple[List[torch.Tensor], List[Optional[torch.Tensor]]]:
    """CLIP-like embeddings (``N(0,1).half().float()``, ``GraphBins.py:106``) and YOLO-like rounded
    xywh boxes in FULL-resolution pixels (``Yolov7Wrapper.py:110,118-123``).  ``none_at``: image with
    no detections -> box entry ``None`` and a single <UNK> embedding row (``ObjCAViT.py:312-315``)."""
    g = torch.Generator().manual_seed(seed + 7919)
    H, W = 2 * h, 2 * w
    feats, boxes = [], []
    for b in range(B):
        n = n_obj
        if ragged:
            n = int(torch.randint(1, n_obj + 1, (1,), generator=g))
        if none_at is not None and b == none_at:
            feats.append(torch.randn(1, 512, generator=g).half().float())
            boxes.append(None)
            continue
        feats.append(torch.randn(n, 512, generator=g).half().float())
        x1 = torch.rand(n, generator=g) * (W - 8)
        y1 = torch.rand(n, generator=g) * (H - 8)
        bw = 4 + torch.rand(n, generator=g) * (W / 4 - 4)
        bh = 4 + torch.rand(n, generator=g) * (H / 4 - 4)
        x2 = torch.minimum(x1 + bw, torch.tensor(float(W)))
        y2 = torch.minimum(y1 + bh, torch.tensor(float(H)))
        xyxy = torch.stack([x1, y1, x2, y2], 1).round()
        xywh = torch.stack([(xyxy[:, 0] + xyxy[:, 2]) / 2, (xyxy[:, 1] + xyxy[:, 3]) / 2,
                            xyxy[:, 2] - xyxy[:, 0], xyxy[:, 3] - xyxy[:, 1]], 1)
        boxes.append(xywh)
    return feats, boxes


def make_gt(B: int, h: int, w: int, max_depth: float, min_depth: float = 0.001, zero_frac: float = 0.2,
            seed: int = 1234):
    """Config-5 ground truth ``[B,1,2h,2w]`` with invalid (zero) pixels; mask = gt > min_depth
    (``GraphBinsLM.py:134``)."""
    g = torch.Generator().manual_seed(seed + 104729)
    gt = 0.5 + torch.rand(B, 1, 2 * h, 2 * w, generator=g) * (max_depth - 0.5)
    drop = torch.rand(B, 1, 2 * h, 2 * w, generator=g) < zero_frac
    gt = gt.masked_fill(drop, 0.0)
    return gt, gt > min_depth
