"""Head classes: the part of ``AdaBins`` / ``GraphBins`` between the decoder's feature map and
``depth_pred`` (``modules/AdaBins.py:42-54,76-89``; ``modules/GraphBins.py:45-62,105-121``).

The encoder/decoder, YOLOv7 and CLIP stay outside (SURVEY.md §2 "out of scope"): these classes take
the dense features (and the object embeddings / boxes) the reference models compute just before the
head, expose the same attribute names (``adaptive_bins_layer`` / ``objcavit``, ``conv_out``) so
checkpoints map with a prefix strip, and return the reference's ``ReturnType`` namedtuples.
"""
from __future__ import annotations

from collections import namedtuple

import torch
import torch.nn as nn

from .. import engine
from ._base import KernelModule
from .miniViT import mViT
from .ObjCAViT import ObjCAViT


class _HeadBase(KernelModule):
    precision = "fp32"   # "fp32": parity path (SIMT fp32).  "bf16": tensor-core path (tcgen05).

    def _depth_range(self):
        ds = self.args[self.args.basic.dataset]
        return float(ds.min_depth), float(ds.max_depth)

    def get_non_encoder_params(self):  # AdaBins.py:63-65 — the head is all "non encoder"
        yield from self.parameters()

    def get_encoder_params(self):
        return iter(())

    def get_frozen_params(self):
        return iter(())


class AdaBinsHead(_HeadBase):
    def __init__(self, args, precision: str = "fp32"):
        super().__init__()
        self.args = args
        self.precision = precision
        self.n_bins = self.args.adabins.n_bins
        self.ReturnType = namedtuple('ReturnType', ['depth_pred', 'bin_edges'])
        max_seq_len = 1200 if self.args[self.args.model.name].get('do_final_upscale') else 500
        self.adaptive_bins_layer = mViT(128, n_query_channels=128, patch_size=16, dim_out=self.n_bins,
                                        embedding_dim=128, norm='linear', max_seq_len=max_seq_len)
        self.conv_out = nn.Sequential(nn.Conv2d(128, self.n_bins, kernel_size=1, stride=1, padding=0),
                                      nn.Softmax(dim=1))

    def forward(self, unet_out):
        """``unet_out``: decoder features ``[B,128,h,w]`` -> ``ReturnType(depth_pred [B,1,h,w], bin_edges [B,K+1])``."""
        self._guard_inference(unet_out)
        lo, hi = self._depth_range()
        P = self._params()
        if self.precision == "bf16":
            from .. import engine_bf16
            depth, edges = engine_bf16.head_forward(self, P, "adaptive_bins_layer.", unet_out, None, None, lo, hi)
        else:
            x = unet_out.contiguous().float()
            widths, ram = engine.mvit_forward(P, "adaptive_bins_layer.", x, 128, 'linear')
            depth, edges = engine.head_tail(P, "conv_out.", ram, widths, lo, hi)
        return self.ReturnType(depth_pred=depth, bin_edges=edges)


class GraphBinsHead(_HeadBase):
    def __init__(self, args, precision: str = "fp32", obj_feature_dim: int = 512):
        super().__init__()
        self.args = args
        self.precision = precision
        self.ReturnType = namedtuple('ReturnType', ['depth_pred', 'bin_edges', 'detections'])
        emb = self.args[self.args.model.name].objcavit.embedding_dim
        max_seq_len = 1200 if self.args[self.args.model.name].get('do_final_upscale') else 500
        self.objcavit = ObjCAViT(self.args, n_query_channels=128, patch_size=16, im_feature_dim=128,
                                 obj_feature_dim=obj_feature_dim, embedding_dim=emb,
                                 dim_out=self.args.graphbins.n_bins, norm='linear', max_seq_len=max_seq_len)
        self.conv_out = nn.Sequential(nn.Conv2d(emb, self.args.graphbins.n_bins, kernel_size=1, stride=1, padding=0),
                                      nn.Softmax(dim=1))

    def forward(self, dense_features, object_features, object_xywh_list, detections=None):
        """-> ``ReturnType(depth_pred, bin_edges, detections)`` (``GraphBins.py:105-121``)."""
        self._guard_inference(dense_features)
        lo, hi = self._depth_range()
        P = self._params()
        oc = self.objcavit
        if self.precision == "bf16":
            from .. import engine_bf16
            depth, edges = engine_bf16.head_forward(self, P, "objcavit.", dense_features, object_features,
                                                    object_xywh_list, lo, hi)
        else:
            x = dense_features.contiguous().float()
            widths, ram = engine.objcavit_forward(P, "objcavit.", x, object_features, object_xywh_list, oc.strategy,
                                                  oc.no_obj_sa, oc.use_2_saca, 128, 'linear')
            depth, edges = engine.head_tail(P, "conv_out.", ram, widths, lo, hi)
        return self.ReturnType(depth_pred=depth, bin_edges=edges, detections=detections)


def build_head(args, precision: str = "fp32"):
    """Head for a params file: ``model.name == 'adabins'`` -> :class:`AdaBinsHead`, else :class:`GraphBinsHead`."""
    return AdaBinsHead(args, precision) if args.model.name == "adabins" else GraphBinsHead(args, precision)
