"""TEST INFRASTRUCTURE ONLY. Import the UNMODIFIED reference head modules from /root/reference.

In the build container the files come from ``/root/reference``; on the GPU box from ``oracle/_ref`` (unmodified copies made by
``oracle/vendor_ref.py``, git-ignored).  Follows
SURVEY.md Appendix A: the reference's ``modules/ObjCAViT.py:10`` imports ``pytorch_lightning`` only
for the ``pl.LightningModule`` base class (``ObjCAViT.py:150,216``), so a stub whose
``LightningModule`` is ``nn.Module`` (+ a ``device`` property) lets the file import as it is.
"""
from __future__ import annotations

import os
import sys
import types

import torch
import torch.nn as nn

def _find_root() -> str:
    """``$OBJCAVIT_REFERENCE_ROOT`` > ``/root/reference`` (build container) > ``oracle/_ref`` (the byte-identical copies made by
    ``oracle/vendor_ref.py``; what the GPU box has)."""
    env = os.environ.get("OBJCAVIT_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isfile("/root/reference/modules/ObjCAViT.py"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


REFERENCE_ROOT = _find_root()


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "modules", "ObjCAViT.py"))


def _install_lightning_stub() -> None:
    if "pytorch_lightning" in sys.modules:
        return

    class LightningModule(nn.Module):
        @property
        def device(self):
            try:
                return next(self.parameters()).device
            except StopIteration:
                return torch.device("cpu")

    stub = types.ModuleType("pytorch_lightning")
    stub.LightningModule = LightningModule
    stub.__graft_stub__ = True
    sys.modules["pytorch_lightning"] = stub


def import_reference():
    """Returns a namespace with the reference classes (mViT, ObjCAViT, ..., SILogLoss)."""
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    _install_lightning_stub()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from modules.layers import PatchTransformerEncoder, PixelWiseDotProduct  # type: ignore
    from modules.miniViT import mViT  # type: ignore
    from modules.ObjCAViT import (GridRandomPositionalEmbeddings, ObjCAViT,  # type: ignore
                                  SelfAttnCrossAttn)
    from losses.SILogLoss import SILogLoss  # type: ignore

    return types.SimpleNamespace(
        PatchTransformerEncoder=PatchTransformerEncoder,
        PixelWiseDotProduct=PixelWiseDotProduct,
        mViT=mViT,
        GridRandomPositionalEmbeddings=GridRandomPositionalEmbeddings,
        SelfAttnCrossAttn=SelfAttnCrossAttn,
        ObjCAViT=ObjCAViT,
        SILogLoss=SILogLoss,
    )


class ReferenceHead(nn.Module):
    """The reference head + the 8-line tail re-stated from ``modules/AdaBins.py:76-87`` /
    ``modules/GraphBins.py:107-119`` around the imported reference module (the AdaBins/GraphBins
    classes themselves cannot be constructed offline: ``DenseFeatureExtractor.py:149`` downloads
    weights, ``GraphBins.py:11-17`` imports nltk/yolo/clip).  ``state_dict`` keys follow the
    reference: ``adaptive_bins_layer.*`` / ``objcavit.*`` and ``conv_out.0.*``.
    """

    def __init__(self, args):
        super().__init__()
        ref = import_reference()
        self.args = args
        self.kind = args.model.name
        n_bins = args[args.model.name].n_bins
        max_seq_len = 1200 if args[args.model.name].get("do_final_upscale") else 500
        if self.kind == "adabins":
            self.adaptive_bins_layer = ref.mViT(128, n_query_channels=128, patch_size=16, dim_out=n_bins,
                                                embedding_dim=128, norm="linear", max_seq_len=max_seq_len)
            emb = 128
        else:
            emb = args[args.model.name].objcavit.embedding_dim
            self.objcavit = ref.ObjCAViT(args, n_query_channels=128, patch_size=16, im_feature_dim=128,
                                         obj_feature_dim=512, embedding_dim=emb, dim_out=n_bins,
                                         norm="linear", max_seq_len=max_seq_len)
        self.conv_out = nn.Sequential(nn.Conv2d(emb, n_bins, kernel_size=1, stride=1, padding=0),
                                      nn.Softmax(dim=1))

    def forward(self, features, object_features=None, object_xywh_list=None):
        ds = self.args[self.args.basic.dataset]
        if self.kind == "adabins":
            bin_widths_normed, ram = self.adaptive_bins_layer(features)
        else:
            bin_widths_normed, ram = self.objcavit(features, object_features, object_xywh_list)
        out = self.conv_out(ram)
        bin_widths = (ds.max_depth - ds.min_depth) * bin_widths_normed
        bin_widths = nn.functional.pad(bin_widths, (1, 0), mode="constant", value=ds.min_depth)
        bin_edges = torch.cumsum(bin_widths, dim=1)
        centers = 0.5 * (bin_edges[:, :-1] + bin_edges[:, 1:])
        n, dout = centers.size()
        centers = centers.view(n, dout, 1, 1)
        depth_pred = torch.sum(out * centers, dim=1, keepdim=True)
        return depth_pred, bin_edges, bin_widths_normed, ram
