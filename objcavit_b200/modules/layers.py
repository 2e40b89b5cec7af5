"""``PatchTransformerEncoder`` and ``PixelWiseDotProduct`` — the reference's ``modules/layers.py``
surface (``layers.py:5-36``) on the objcavit_b200 kernels."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import engine, ops
from ._base import KernelModule


class PatchTransformerEncoder(KernelModule):
    """Same constructor, parameters and ``state_dict`` keys as ``modules/layers.py:5-14``; ``forward``
    returns the encoded tokens in the reference's ``[S, N, E]`` layout (``layers.py:22-24``)."""

    def __init__(self, in_channels, patch_size=10, embedding_dim=128, num_heads=4, max_seq_len=500):
        super().__init__()
        assert embedding_dim == 128 and num_heads == 4, "kernels are built for E=128, 4 heads (all reference configs)"
        encoder_layers = nn.TransformerEncoderLayer(embedding_dim, num_heads, dim_feedforward=1024)
        self.transformer_encoder = nn.TransformerEncoder(encoder_layers, num_layers=4, enable_nested_tensor=False)
        self.embedding_convPxP = nn.Conv2d(in_channels, embedding_dim, kernel_size=patch_size, stride=patch_size,
                                           padding=0)
        self.positional_encodings = nn.Parameter(torch.rand(max_seq_len, embedding_dim), requires_grad=True)
        self.patch_size = patch_size

    def forward(self, x):
        self._guard_inference(x)
        tok = engine.mvit_tokens(self._params(), "", x.contiguous().float(), self.patch_size)  # [N,S,E]
        return tok.permute(1, 0, 2)


class PixelWiseDotProduct(KernelModule):
    """``modules/layers.py:27-36``: ``[n,c,h,w] x [n,cout,c] -> [n,cout,h,w]`` in one batched kernel
    (no permute/contiguous copies)."""

    def forward(self, x, K):
        n, c, h, w = x.size()
        _, cout, ck = K.size()
        assert c == ck, "Number of channels in x and Embedding dimension (at dim 2) of K matrix must match"
        y = ops.batched_matmul_nn(K.float(), x.contiguous().float().view(n, c, h * w))
        return y.view(n, cout, h, w)
