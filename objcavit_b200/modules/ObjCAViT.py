"""``GridRandomPositionalEmbeddings``, ``SelfAttnCrossAttn`` and ``ObjCAViT`` — the reference's
``modules/ObjCAViT.py`` surface (``ObjCAViT.py:18-390``) on the objcavit_b200 kernels."""
from __future__ import annotations

import math
import sys

import torch
import torch.nn as nn

from .. import engine, ops
from ._base import KernelModule
from .layers import PixelWiseDotProduct


class GridRandomPositionalEmbeddings(KernelModule):
    """One learnable vector per image patch (``ObjCAViT.py:18-47``); table length from the larger of the
    train/test image grids, exactly as the reference computes it (full-resolution dims / patch size)."""

    def __init__(self, args, embedding_dim, patch_size, mode="centre"):
        super().__init__()
        self.args = args
        self.embedding_dim = embedding_dim
        self.patch_size = patch_size
        self.mode = mode
        assert self.mode in ["centre", "roi_align"], "Error: unrecognised GridRandomPositionalEmbeddings mode."
        ds = self.args[self.args.basic.dataset]
        tot = []
        for dims in (ds.dimensions_train, ds.dimensions_test):
            tot.append(math.ceil(dims[0] / self.patch_size) * math.ceil(dims[1] / self.patch_size))
        self.sequence_length = max(tot)
        self.positional_encodings = nn.Parameter(torch.rand(self.sequence_length, self.embedding_dim), requires_grad=True)

    def forward(self, coords, image_features, input_coord_space="img", factor=2.0):
        """``ObjCAViT.py:50-147`` incl. quirks F13.  ``img``: ``[B,S,2|4] -> [B,S,E]``; ``obj``: ``[N,2|4] -> [N,E]``."""
        table = self.positional_encodings.data
        h, w = image_features.shape[2], image_features.shape[3]
        gh, gw = math.ceil(h / self.patch_size), math.ceil(w / self.patch_size)
        c = coords.float()
        if input_coord_space == "img":
            B, S = c.shape[0], c.shape[1]
            c2 = c.reshape(B * S, -1)
        else:
            c2 = c
        if self.mode == "centre":
            pts = c2[:, 0:2].contiguous()
            if input_coord_space == "img":
                # F13a: only tokens 0 and 1 of EVERY image are normalised; the kernel keys on the point index,
                # so feed one image's S points at a time semantics by tiling the index (B identical rows).
                out = ops.grid_sample_pos(table, c[0, :, 0:2].contiguous(), gh, gw, 2, float(gh), float(gw))
                return out.unsqueeze(0).expand(B, -1, -1).contiguous() if _all_rows_equal(c) else \
                    torch.stack([ops.grid_sample_pos(table, c[b, :, 0:2].contiguous(), gh, gw, 2, float(gh), float(gw))
                                 for b in range(B)], 0)
            return ops.grid_sample_pos(table, pts, gh, gw, 1, h * factor, w * factor)
        scale = 1.0 / self.patch_size if input_coord_space == "img" else 1.0 / (self.patch_size * factor)
        out = ops.ps_roi_align_pos(table, c2[:, 0:4].contiguous(), gh, gw, scale, True)
        return out.view(B, S, -1) if input_coord_space == "img" else out


def _all_rows_equal(c):
    return c.shape[0] == 1 or bool((c == c[0:1]).all())


class SelfAttnCrossAttn(KernelModule):
    """``ObjCAViT.py:150-213``.  Registers the same sub-modules as the reference — including the
    prototype layers ``image_encoder_layers`` / ``obj_encoder_layers`` that are never executed but are
    part of every checkpoint."""

    def __init__(self, args, embedding_dim=128, num_heads=4, dim_feedforward=1024):
        super().__init__()
        self.args = args
        assert embedding_dim == 128 and num_heads == 4
        self.image_encoder_layers = nn.TransformerEncoderLayer(embedding_dim, num_heads, dim_feedforward=dim_feedforward, batch_first=True)
        self.image_transformer_encoder = nn.TransformerEncoder(self.image_encoder_layers, num_layers=4, enable_nested_tensor=False)
        if self.args.graphbins.objcavit.get("no_obj_sa") != True:  # noqa: E712 (reference idiom)
            self.obj_encoder_layers = nn.TransformerEncoderLayer(embedding_dim, num_heads, dim_feedforward=dim_feedforward, batch_first=True)
            self.obj_transformer_encoder = nn.TransformerEncoder(self.obj_encoder_layers, num_layers=4, enable_nested_tensor=False)
        self.cross_attn_obj_im = nn.MultiheadAttention(embed_dim=embedding_dim, num_heads=4, batch_first=True)
        self.cross_attn_im_obj = nn.MultiheadAttention(embed_dim=embedding_dim, num_heads=4, batch_first=True)

    def forward(self, image_patch_embeddings, object_features):
        """-> ``(final_image_features [B,S,E], final_object_features [B,S,E])``."""
        self._guard_inference(image_patch_embeddings)
        no_sa = self.args.graphbins.objcavit.get("no_obj_sa") == True  # noqa: E712
        x = image_patch_embeddings.contiguous().float()
        if torch.is_tensor(object_features):
            objs = object_features.contiguous().float()
        else:
            counts = [int(o.shape[0]) for o in object_features]
            objs = engine.ObjectBatch(torch.cat([o.float() for o in object_features], 0).contiguous(), counts)
        return engine.saca_forward(self._params(), "", x, objs, no_sa, need_object_output=True)


class ObjCAViT(KernelModule):
    """Object Cross-Attention ViT block (``ObjCAViT.py:216-390``): same constructor, parameters and
    ``forward(image_features, object_features, object_xywh_list)`` contract."""

    def __init__(self, args, im_feature_dim=128, obj_feature_dim=512, n_query_channels=128, patch_size=16, dim_out=256,
                 embedding_dim=128, num_heads=4, norm='linear', max_seq_len=50):
        super().__init__()
        self.args = args
        self.norm = norm
        self.n_query_channels = n_query_channels
        self.patch_size = patch_size
        self.half_patch_size = self.patch_size // 2
        self.obj_feature_dim = obj_feature_dim
        self.strategy = self.args[self.args.model.name].objcavit.positional_embedding_strategy

        def mlp(nin):
            return nn.Sequential(nn.Linear(nin, 32, bias=True), nn.LeakyReLU(),
                                 nn.Linear(32, 64, bias=True), nn.LeakyReLU(),
                                 nn.Linear(64, 128, bias=True), nn.LeakyReLU(),
                                 nn.Linear(128, 256, bias=True), nn.LeakyReLU(),
                                 nn.Linear(256, embedding_dim, bias=True))

        if self.strategy == "grid_random":
            self.positional_encoder = GridRandomPositionalEmbeddings(args, embedding_dim=embedding_dim, patch_size=patch_size, mode="centre")
        elif self.strategy == "grid_random_roi_align":
            self.positional_encoder = GridRandomPositionalEmbeddings(args, embedding_dim=embedding_dim, patch_size=patch_size, mode="roi_align")
        elif self.strategy == "learned":
            self.positional_encoder = mlp(2)
        elif self.strategy == "learned_bbox_wh":
            self.positional_encoder = mlp(4)
        else:
            sys.exit("Error: ObjCAViT positional embedding strategy not recognised.")  # ObjCAViT.py:284

        self.image_embedding_convPxP = nn.Conv2d(im_feature_dim, embedding_dim, kernel_size=self.patch_size,
                                                 stride=patch_size, padding=0)
        self.obj_embedding_layer = nn.Linear(self.obj_feature_dim, embedding_dim)
        self.saca_1 = SelfAttnCrossAttn(self.args, embedding_dim, num_heads, dim_feedforward=1024)
        if self.args.graphbins.objcavit.get("use_2_saca") == True:  # noqa: E712
            self.saca_2 = SelfAttnCrossAttn(self.args, embedding_dim, num_heads, dim_feedforward=1024)
        self.dot_product_layer = PixelWiseDotProduct()
        self.conv3x3 = nn.Conv2d(im_feature_dim, embedding_dim, kernel_size=3, stride=1, padding=1)
        self.regressor = nn.Sequential(nn.Linear(embedding_dim, 256),
                                       nn.LeakyReLU(),
                                       nn.Linear(256, 256),
                                       nn.LeakyReLU(),
                                       nn.Linear(256, dim_out))

    @property
    def no_obj_sa(self):
        return self.args.graphbins.objcavit.get("no_obj_sa") == True  # noqa: E712

    @property
    def use_2_saca(self):
        return self.args.graphbins.objcavit.get("use_2_saca") == True  # noqa: E712

    def forward(self, image_features, object_features, object_xywh_list):
        """-> ``(bin_widths_normed, range_attention_maps)``; overwrites ``object_features[i]`` with the
        embedded object tokens like the reference does (``ObjCAViT.py:330``, F11)."""
        if self._wants_grad(image_features):     # training through ``self.objcavit(...)`` (GraphBins.py:107)
            if self.norm != 'linear':
                raise NotImplementedError("objcavit_b200: the training kernels implement norm='linear', the only value the "
                                          "reference's callers pass (AdaBins.py:46, GraphBins.py:54)")
            from ._autograd import train_forward
            cfg = dict(kind="graphbins", prefix="", surface="module", strategy=self.strategy, no_obj_sa=self.no_obj_sa,
                       use_2_saca=self.use_2_saca)
            return train_forward(self, cfg, image_features, object_features, object_xywh_list)
        return engine.objcavit_forward(self._params(), "", image_features.contiguous().float(), object_features,
                                       object_xywh_list, self.strategy, self.no_obj_sa, self.use_2_saca,
                                       self.n_query_channels, self.norm)
