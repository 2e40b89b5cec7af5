"""bf16 tensor-core sequencing of the head (throughput path).

The pixel path (conv3x3 -> range-attention dot product -> conv_out -> softmax -> bin expectation) is ONE
persistent tcgen05 kernel (``csrc/pixel_fused_bf16.cu``); the 128 range-attention channels are folded into
the 1x1 conv weights per image (``ocv_fold_queries_bf16``), so neither ``range_attention_maps`` nor the
[B,256,h,w] probability volume is ever written.  Features are consumed as NHWC bf16: a ``channels_last``
bf16 tensor is used in place (zero copy), anything else goes through the layout kernel once.
"""
from __future__ import annotations

import torch

from . import engine, ops


def features_nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    """[B,128,h,w] (any of fp32/bf16, NCHW/channels_last) -> contiguous [B,h,w,128] bf16 without torch math."""
    if x.dtype == torch.bfloat16 and x.permute(0, 2, 3, 1).is_contiguous():
        return x.permute(0, 2, 3, 1)
    if x.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError(f"features must be fp32 or bf16, got {x.dtype}")
    return ops.nchw_to_nhwc_bf16(x.contiguous())


def packed_weights(head, P, pre):
    """Derived, kernel-friendly copies of parameters (dropped by ``KernelModule._apply`` / ``load_state_dict``)."""
    cache = head.__dict__.get("_ocv_packed")
    key = tuple(p.data_ptr() for p in (P[pre + "conv3x3.weight"],)) + (P[pre + "conv3x3.weight"]._version,)
    if cache is None or cache.get("key") != key:
        cache = {"key": key, "w3p": ops.pack_conv3x3_bf16(P[pre + "conv3x3.weight"].contiguous())}
        head.__dict__["_ocv_packed"] = cache
    return cache


def head_forward(head, P, pre, x, object_features, object_xywh_list, lo, hi):
    """-> (depth [B,1,h,w] fp32, bin_edges [B,K+1] fp32)."""
    x_nhwc = features_nhwc_bf16(x)
    if object_features is None:
        tok = engine.mvit_tokens_any(P, pre + "patch_transformer.", x, x_nhwc)
    else:
        oc = head.objcavit
        tok = engine.objcavit_tokens(P, pre, x, object_features, object_xywh_list, oc.strategy, oc.no_obj_sa,
                                     oc.use_2_saca, x_nhwc=x_nhwc)
    _, edges, centres = ops.bins(engine.regress(P, pre + "regressor.", tok), "linear", lo, hi)
    pk = packed_weights(head, P, pre)
    Mb, blog = ops.fold_queries_bf16(P["conv_out.0.weight"], P["conv_out.0.bias"], tok[:, 1:129, :],
                                     P[pre + "conv3x3.bias"])
    depth = ops.pixel_fused_bf16(x_nhwc, pk["w3p"], Mb, blog, centres)
    return depth, edges
