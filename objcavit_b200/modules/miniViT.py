"""``mViT`` — the AdaBins adaptive-bins block (reference ``modules/miniViT.py:6-44``)."""
from __future__ import annotations

import torch.nn as nn

from .. import engine
from ._base import KernelModule
from .layers import PatchTransformerEncoder, PixelWiseDotProduct


class mViT(KernelModule):
    def __init__(self, in_channels, n_query_channels=128, patch_size=16, dim_out=256,
                 embedding_dim=128, num_heads=4, norm='linear', max_seq_len=500):
        super().__init__()
        self.norm = norm
        self.n_query_channels = n_query_channels
        self.patch_size = patch_size
        self.patch_transformer = PatchTransformerEncoder(in_channels, patch_size, embedding_dim, num_heads, max_seq_len)
        self.dot_product_layer = PixelWiseDotProduct()
        self.conv3x3 = nn.Conv2d(in_channels, embedding_dim, kernel_size=3, stride=1, padding=1)
        self.regressor = nn.Sequential(nn.Linear(embedding_dim, 256),
                                       nn.LeakyReLU(),
                                       nn.Linear(256, 256),
                                       nn.LeakyReLU(),
                                       nn.Linear(256, dim_out))

    def forward(self, x):
        """-> ``(bin_widths_normed [N,dim_out], range_attention_maps [N,n_query_channels,h,w])``."""
        if self._wants_grad(x):        # training through ``self.adaptive_bins_layer(unet_out)`` (AdaBins.py:76)
            if self.norm != 'linear':
                raise NotImplementedError("objcavit_b200: the training kernels implement norm='linear', the only value the "
                                          "reference's callers pass (AdaBins.py:46, GraphBins.py:54)")
            from ._autograd import train_forward
            return train_forward(self, dict(kind="adabins", prefix="", surface="module"), x, None, None)
        return engine.mvit_forward(self._params(), "", x.contiguous().float(), self.n_query_channels, self.norm)
