"""bf16 tensor-core sequencing of the head (throughput path).

Pixel path: ONE persistent tcgen05 kernel (``csrc/pixel_fused_bf16.cu``) does conv3x3 -> range-attention dot
product -> conv_out -> softmax -> bin expectation; the 128 range-attention channels are folded into the 1x1
conv weights per image (``ocv_fold_queries_bf16``), so neither ``range_attention_maps`` nor the [B,256,h,w]
probability volume is ever written.  Token path: patch embedding as a TMA-fed tcgen05 GEMM straight from the NHWC
feature map, encoder layers as 5 launches each (QKV GEMM, attention core, out-proj+residual+LayerNorm GEMM,
FFN1 GEMM+ReLU, FFN2+residual+LayerNorm GEMM), bf16 operands with fp32 accumulation and an fp32 residual stream.

Features are consumed as NHWC bf16: a ``channels_last`` bf16 tensor is used in place (zero copy), anything else
goes through the layout kernel once.  Reference semantics incl. quirks F5/F6/F7/F11/F13 are those of
:mod:`objcavit_b200.engine` (same control flow, bf16 kernels).
"""
from __future__ import annotations

from typing import Dict

import torch

from . import engine, ops

E = engine.E
HEADS = engine.HEADS


def features_nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    """[B,128,h,w] (fp32/bf16, NCHW/channels_last) -> contiguous [B,h,w,128] bf16 without torch math."""
    if x.dtype == torch.bfloat16 and x.permute(0, 2, 3, 1).is_contiguous():
        return x.permute(0, 2, 3, 1)
    if x.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError(f"features must be fp32 or bf16, got {x.dtype}")
    return ops.nchw_to_nhwc_bf16(x.contiguous())


class Packed:
    """Derived, kernel-friendly copies of parameters (bf16 GEMM weights, packed conv weights, the batch-invariant
    image positional rows).  Rebuilt when any parameter's storage or version counter changes."""

    def __init__(self):
        self.key = None
        self.w: Dict[str, torch.Tensor] = {}

    @staticmethod
    def signature(P):
        return tuple((v.data_ptr(), v._version) for v in P.values())

    def get(self, P, name, maker):
        t = self.w.get(name)
        if t is None:
            t = maker(P)
            self.w[name] = t
        return t

    def bf16(self, P, name):
        return self.get(P, "bf16:" + name, lambda P: ops.cast_bf16(P[name].contiguous()))


def packed_for(head, P) -> Packed:
    pk = head.__dict__.get("_ocv_packed")
    sig = Packed.signature(P)
    if pk is None or pk.key != sig:
        pk = Packed()
        pk.key = sig
        head.__dict__["_ocv_packed"] = pk
    return pk


# ----------------------------------------------------------------------------------------------------
# transformer blocks
# ----------------------------------------------------------------------------------------------------

def encoder_layer(P, pk: Packed, pre, x32, x16, mask, zero_rows):
    B, S, _ = x32.shape
    M = B * S
    _, qkv = ops.gemm_bf16_ex(x16.view(M, E), pk.bf16(P, pre + "self_attn.in_proj_weight"),
                              P[pre + "self_attn.in_proj_bias"])
    qkv = qkv.view(B, S, 3 * E)
    att = ops.mha_core_bf16(qkv[:, :, 0:E], qkv[:, :, E:2 * E], qkv[:, :, 2 * E:], mask, HEADS)
    y32, y16 = ops.gemm_bf16_ex(att.view(M, E), pk.bf16(P, pre + "self_attn.out_proj.weight"),
                                P[pre + "self_attn.out_proj.bias"], residual=x32.view(M, E),
                                ln=(P[pre + "norm1.weight"], P[pre + "norm1.bias"]), want32=True)
    _, h16 = ops.gemm_bf16_ex(y16, pk.bf16(P, pre + "linear1.weight"), P[pre + "linear1.bias"], act=ops.ACT_RELU)
    z32, z16 = ops.gemm_bf16_ex(h16, pk.bf16(P, pre + "linear2.weight"), P[pre + "linear2.bias"], residual=y32,
                                ln=(P[pre + "norm2.weight"], P[pre + "norm2.bias"]), zero_rows=zero_rows, want32=True)
    return z32.view(B, S, E), z16.view(B, S, E)


def transformer_encoder(P, pk, pre, x32, x16, mask=None, num_layers=4):
    for i in range(num_layers):
        last = i == num_layers - 1
        x32, x16 = encoder_layer(P, pk, f"{pre}layers.{i}.", x32, x16, mask, mask if last else None)
    return x32, x16


def cross_attention(P, pk, pre, qv16, k16, mask):
    """nn.MultiheadAttention(query=qv, key=k, value=qv) — both reference call sites pass the same tensor as query
    and value (``ObjCAViT.py:195-207``), so q and v come out of one N=384 GEMM."""
    B, Sq, _ = qv16.shape
    Sk = k16.shape[1]
    w = pk.bf16(P, pre + "in_proj_weight")
    b = P[pre + "in_proj_bias"]
    _, qkv = ops.gemm_bf16_ex(qv16.reshape(B * Sq, E), w, b)
    qkv = qkv.view(B, Sq, 3 * E)
    _, kk = ops.gemm_bf16_ex(k16.reshape(B * Sk, E), w[E:2 * E], b[E:2 * E])
    att = ops.mha_core_bf16(qkv[:, :, 0:E], kk.view(B, Sk, E), qkv[:, :, 2 * E:], mask, HEADS)
    o32, o16 = ops.gemm_bf16_ex(att.view(B * Sq, E), pk.bf16(P, pre + "out_proj.weight"), P[pre + "out_proj.bias"],
                                want32=True)
    return o32.view(B, Sq, E), o16.view(B, Sq, E)


# ----------------------------------------------------------------------------------------------------
# tokens
# ----------------------------------------------------------------------------------------------------

def mvit_tokens(P, pk, pre, x_nhwc, patch_size=16):
    wp = pk.get(P, "patch:" + pre, lambda P: ops.pack_patch_weights_bf16(P[pre + "embedding_convPxP.weight"].contiguous()))
    t32, t16 = ops.patch_embed_bf16(x_nhwc, wp, P[pre + "embedding_convPxP.bias"], P[pre + "positional_encodings"],
                                    patch_size)
    t32, _ = transformer_encoder(P, pk, pre + "transformer_encoder.", t32, t16)
    return t32


def _mlp_keys(pre, idxs):
    return [f"{pre}{i}.weight" for i in idxs], [f"{pre}{i}.bias" for i in idxs]


def positional_embeddings(P, pk, pre, strategy, boxes, h, w, patch_size=16):
    """(object rows [total,E] fp32, image rows [S,E] fp32).  The image side depends only on (h, w) and the weights:
    it is produced once and kept with the packed weights."""
    if strategy in ("learned", "learned_bbox_wh"):
        nin = 4 if strategy == "learned_bbox_wh" else 2
        wk, bk = _mlp_keys(pre, (0, 2, 4, 6, 8))
        Ws, bs = [P[k] for k in wk], [P[k] for k in bk]
        obj, _ = ops.mlp(boxes[:, 0:nin], Ws, bs)

        def make_img(P):
            coords = ops.patch_coords(h // patch_size, w // patch_size, patch_size, boxes.device)
            return ops.mlp(coords[:, 0:nin], Ws, bs)[0]
        return obj, pk.get(P, f"imgpos:{h}x{w}", make_img)
    obj, img = engine.positional_embeddings(P, pre, strategy, boxes, h, w, patch_size)
    return obj, img


def saca(P, pk, pre, img32, img16, objs, no_obj_sa, need_object_output):
    """``SelfAttnCrossAttn.forward`` (``ObjCAViT.py:167-213``), bf16 kernels.  ``objs``: :class:`engine.ObjectBatch`
    (block 1) or the (fp32, bf16) dense [B,S,E] pair block 1 produced (block 2, F6)."""
    B, S, _ = img32.shape
    att32, att16 = transformer_encoder(P, pk, pre + "image_transformer_encoder.", img32, img16)
    if isinstance(objs, engine.ObjectBatch):
        n_max = objs.n_max
        if n_max > S:
            raise ValueError("more objects than image tokens: the reference's F.pad would crop; unsupported")
        padded, mask = ops.pad_objects(objs.rows, objs.offsets, B, n_max, n_max, 0.0001, False)          # :180-183
        if no_obj_sa:
            obj32 = padded                                                                            # :186
        else:
            obj32, _ = transformer_encoder(P, pk, pre + "obj_transformer_encoder.", padded, ops.cast_bf16(padded), mask)
        uni = objs.uniform_offsets(n_max)
        obj_s32, kpm = ops.pad_objects(obj32.reshape(B * n_max, E), uni, B, S, n_max, 0.0001, True,
                                       mask_offsets=objs.offsets)                                     # :192-194 (F5)
        obj_s16 = ops.cast_bf16(obj_s32)
    else:
        o32, o16 = objs
        if no_obj_sa:
            obj_s32, obj_s16 = o32, o16
        else:
            obj_s32, obj_s16 = transformer_encoder(P, pk, pre + "obj_transformer_encoder.", o32, o16)
        kpm = None
    fin32, fin16 = cross_attention(P, pk, pre + "cross_attn_obj_im.", att16, obj_s16, kpm)               # :195-201
    fobj = None
    if need_object_output:
        fobj = cross_attention(P, pk, pre + "cross_attn_im_obj.", obj_s16, att16, None)                  # :202-207
    return fin32, fin16, fobj


def objcavit_tokens(P, pk, pre, x_nhwc, object_features, object_xywh_list, strategy, no_obj_sa, use_2_saca,
                    patch_size=16, mutate_inputs=True):
    B, h, w, _ = x_nhwc.shape
    feats, boxes, counts = engine.marshal_objects(object_features, object_xywh_list, x_nhwc.device)
    obj_pos, img_pos = positional_embeddings(P, pk, pre + "positional_encoder.", strategy, boxes, h, w, patch_size)
    obj_tok = ops.linear(feats, P[pre + "obj_embedding_layer.weight"], P[pre + "obj_embedding_layer.bias"],
                         residual=obj_pos)                                                             # :330
    batch = engine.ObjectBatch(obj_tok, counts)
    if mutate_inputs:  # F11
        for i in range(B):
            object_features[i] = obj_tok[batch.offsets_host[i]:batch.offsets_host[i + 1]]
    wp = pk.get(P, "patch:" + pre, lambda P: ops.pack_patch_weights_bf16(P[pre + "image_embedding_convPxP.weight"].contiguous()))
    t32, t16 = ops.patch_embed_bf16(x_nhwc, wp, P[pre + "image_embedding_convPxP.bias"], img_pos.contiguous(), patch_size)
    t32, t16, fobj = saca(P, pk, pre + "saca_1.", t32, t16, batch, no_obj_sa, need_object_output=use_2_saca)
    if use_2_saca:
        t32, t16, _ = saca(P, pk, pre + "saca_2.", t32, t16, fobj, no_obj_sa, need_object_output=False)
    return t32


def head_forward(head, P, pre, x, object_features, object_xywh_list, lo, hi, mutate_inputs=True, fingerprint=None):
    """-> (depth [B,1,h,w] fp32, bin_edges [B,K+1] fp32): ONE native call (``ocv_head_forward``) sequences every kernel;
    python only marshals the ragged object lists into CSR form."""
    x_nhwc = features_nhwc_bf16(x)
    nh = head.__dict__.get("_ocv_native")
    if nh is None:
        if object_features is None:
            nh = ops.NativeHead(0, "learned", False, False, head.n_bins, 512, lo, hi)
        else:
            oc = head.objcavit
            nh = ops.NativeHead(1, oc.strategy, oc.no_obj_sa, oc.use_2_saca, head.args.graphbins.n_bins,
                                oc.obj_embedding_layer.in_features, lo, hi)
        head.__dict__["_ocv_native"] = nh
        pipe = head.__dict__.get("_ocv_pipeline")
        if pipe:
            nh.set_pipeline(*pipe)
    nh.bind(P, fingerprint)
    if object_features is None:
        return nh.forward(x_nhwc, None, None, None, 0, 0)
    feats, boxes, counts = engine.marshal_objects(object_features, object_xywh_list, x_nhwc.device)
    batch = engine.ObjectBatch(feats, counts)
    obj_tok = torch.empty(feats.shape[0], E, device=feats.device, dtype=torch.float32) if mutate_inputs else None
    depth, edges = nh.forward(x_nhwc, feats, boxes, batch.offsets, feats.shape[0], batch.n_max, obj_tok)
    if mutate_inputs:  # F11: the reference overwrites the caller's list with the embedded tokens (one split: B views)
        object_features[:] = torch.split(obj_tok, counts)
    return depth, edges


def head_forward_staged(head, P, pre, x, object_features, object_xywh_list, lo, hi, want_argmax=False):
    """Same computation sequenced stage by stage from python (one C-ABI call per kernel): kept for the stage-level
    parity tests and for debugging; the module forward uses :func:`head_forward`."""
    x_nhwc = features_nhwc_bf16(x)
    pk = packed_for(head, P)
    if object_features is None:
        tok = mvit_tokens(P, pk, pre + "patch_transformer.", x_nhwc)
    else:
        oc = head.objcavit
        tok = objcavit_tokens(P, pk, pre, x_nhwc, object_features, object_xywh_list, oc.strategy, oc.no_obj_sa,
                              oc.use_2_saca)
    wk, bk = _mlp_keys(pre + "regressor.", (0, 2, 4))
    y_raw, _ = ops.mlp(tok[:, 0, :], [P[k] for k in wk], [P[k] for k in bk])
    _, edges, centres = ops.bins(y_raw, "linear", lo, hi)
    w3p = pk.get(P, "w3p", lambda P: ops.pack_conv3x3_bf16(P[pre + "conv3x3.weight"].contiguous()))
    Mb, blog = ops.fold_queries_bf16(P["conv_out.0.weight"], P["conv_out.0.bias"], tok[:, 1:129, :],
                                     P[pre + "conv3x3.bias"])
    if want_argmax:  # parity tests: the bin with the largest logit per pixel (north star: "identical bin argmax")
        depth, amax = ops.pixel_fused_bf16(x_nhwc, w3p, Mb, blog, centres, want_argmax=True)
        return depth, edges, amax
    depth = ops.pixel_fused_bf16(x_nhwc, w3p, Mb, blog, centres)
    return depth, edges
