"""Decoder hand-off: the decoder's last layer — ``DenseFeatureExtractor.conv3``, a 3x3 / pad 1 convolution
128 -> 128 with bias and no activation (``modules/DenseFeatureExtractor.py:97,116``) — on the tcgen05 implicit-GEMM pipeline of
the fused pixel kernel, writing the feature map directly as bf16 NHWC, the layout / dtype the bf16 head consumes in place.
Without it the reference path writes the map as fp32 NCHW (39 MB / NYU image) and the head re-reads and converts it.

``state_dict`` keys are those of the ``nn.Conv2d`` it replaces (``weight`` [128,128,3,3], ``bias`` [128]): load the decoder
checkpoint's ``conv3.*`` entries into it.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import ops
from ._base import KernelModule


class FinalConvNHWC(KernelModule):
    def __init__(self, in_channels: int = 128, out_channels: int = 128):
        super().__init__()
        assert in_channels == 128 and out_channels == 128, "the tcgen05 conv pipeline is built for 128 -> 128 channels"
        ref = nn.Conv2d(in_channels, out_channels, kernel_size=3, stride=1, padding=1)   # same default init as the reference
        self.weight = nn.Parameter(ref.weight.detach().clone())
        self.bias = nn.Parameter(ref.bias.detach().clone())

    @torch.no_grad()
    def forward(self, x_d4: torch.Tensor) -> torch.Tensor:
        """``x_d4`` [B,128,h,w] (any float dtype / memory format) -> logical [B,128,h,w] view of bf16 NHWC memory."""
        self._guard_inference(x_d4)
        cache = self.__dict__.get("_ocv_packed")
        sig = (self.weight.data_ptr(), self.weight._version)
        if cache is None or cache[0] != sig:
            cache = (sig, ops.pack_conv3x3_bf16(self.weight.data.contiguous()))
            self.__dict__["_ocv_packed"] = cache
        if x_d4.dtype == torch.bfloat16 and x_d4.permute(0, 2, 3, 1).is_contiguous():
            x_nhwc = x_d4.permute(0, 2, 3, 1)                       # channels_last bf16: consumed in place
        else:
            x_nhwc = ops.nchw_to_nhwc_bf16(x_d4.contiguous())
        y = ops.conv3x3_nhwc_bf16(x_nhwc, cache[1], self.bias.data)
        return y.permute(0, 3, 1, 2)
