"""Host-side sequencing of the head: which C-ABI stage runs on which buffers.

Functions take a ``params`` mapping (reference ``state_dict`` key -> CUDA tensor, fp32) and a key prefix,
so the same code serves ``mViT`` (``modules/miniViT.py``), ``ObjCAViT`` (``modules/ObjCAViT.py``) and
the head classes.  Everything numeric happens inside ``libobjcavit_b200.so``; torch is used for
allocation, views and the ragged-list -> CSR marshalling of the object inputs.

Inference semantics (``eval()`` + ``no_grad``) of the reference are reproduced, including its
quirks F5/F6/F7/F11/F13.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple, Union

import torch

from . import ops

Params = Dict[str, torch.Tensor]
E = 128
HEADS = 4


# ----------------------------------------------------------------------------------------------------
# transformer blocks (nn.TransformerEncoderLayer post-norm / nn.MultiheadAttention), fp32
# ----------------------------------------------------------------------------------------------------

def _encoder_layer(P: Params, pre: str, x: torch.Tensor, mask: Optional[torch.Tensor], zero_rows) -> torch.Tensor:
    B, S, _ = x.shape
    x2 = x.view(B * S, E)
    qkv = ops.linear(x2, P[pre + "self_attn.in_proj_weight"], P[pre + "self_attn.in_proj_bias"]).view(B, S, 3 * E)
    att = ops.mha_core(qkv[:, :, 0:E], qkv[:, :, E:2 * E], qkv[:, :, 2 * E:], mask, HEADS)
    y = ops.linear(att.view(B * S, E), P[pre + "self_attn.out_proj.weight"], P[pre + "self_attn.out_proj.bias"],
                   residual=x2)
    x1 = ops.add_layernorm(y, None, P[pre + "norm1.weight"], P[pre + "norm1.bias"])
    h = ops.linear(x1, P[pre + "linear1.weight"], P[pre + "linear1.bias"], act=ops.ACT_RELU)
    y2 = ops.linear(h, P[pre + "linear2.weight"], P[pre + "linear2.bias"], residual=x1)
    out = ops.add_layernorm(y2, None, P[pre + "norm2.weight"], P[pre + "norm2.bias"], zero_rows=zero_rows)
    return out.view(B, S, E)


def transformer_encoder(P: Params, pre: str, x: torch.Tensor, mask: Optional[torch.Tensor] = None,
                        num_layers: int = 4) -> torch.Tensor:
    """``nn.TransformerEncoder`` (4 post-norm layers).  With a key-padding ``mask`` (uint8 [B,S], 1 = pad)
    the padded rows of the result are 0, as PyTorch's inference fast path returns them (F7)."""
    for i in range(num_layers):
        last = i == num_layers - 1
        x = _encoder_layer(P, f"{pre}layers.{i}.", x, mask, mask if last else None)
    return x


def cross_attention(P: Params, pre: str, q_in, k_in, v_in, mask) -> torch.Tensor:
    """``nn.MultiheadAttention(batch_first=True)(query, key, value, key_padding_mask)`` -> [B,Sq,E]."""
    w, b = P[pre + "in_proj_weight"], P[pre + "in_proj_bias"]
    B, Sq, _ = q_in.shape
    Sk = k_in.shape[1]
    q = ops.linear(q_in.reshape(B * Sq, E), w[0:E], b[0:E]).view(B, Sq, E)
    k = ops.linear(k_in.reshape(B * Sk, E), w[E:2 * E], b[E:2 * E]).view(B, Sk, E)
    v = ops.linear(v_in.reshape(B * Sk, E), w[2 * E:], b[2 * E:]).view(B, Sk, E)
    att = ops.mha_core(q, k, v, mask, HEADS)
    return ops.linear(att.view(B * Sq, E), P[pre + "out_proj.weight"], P[pre + "out_proj.bias"]).view(B, Sq, E)


def mlp_leaky(P: Params, pre: str, x: torch.Tensor, idxs: Sequence[int]) -> torch.Tensor:
    """``nn.Sequential`` of Linear/LeakyReLU pairs ending in a bare Linear (``idxs`` = Linear positions)."""
    for j, i in enumerate(idxs):
        act = ops.ACT_LEAKY_RELU if j < len(idxs) - 1 else ops.ACT_NONE
        x = ops.linear(x, P[f"{pre}{i}.weight"], P[f"{pre}{i}.bias"], act=act)
    return x


# ----------------------------------------------------------------------------------------------------
# a1-a3: mViT
# ----------------------------------------------------------------------------------------------------

def mvit_tokens(P: Params, pre: str, x: torch.Tensor, patch_size: int = 16) -> torch.Tensor:
    """``PatchTransformerEncoder.forward`` (``modules/layers.py:16-24``) as [B,S,E]."""
    tok = ops.patch_embed(x, P[pre + "embedding_convPxP.weight"], P[pre + "embedding_convPxP.bias"],
                          P[pre + "positional_encodings"], patch_size)
    return transformer_encoder(P, pre + "transformer_encoder.", tok)


def range_attention(P: Params, pre: str, x: torch.Tensor, tok: torch.Tensor, n_query: int) -> torch.Tensor:
    """conv3x3 + PixelWiseDotProduct (``miniViT.py:25,31`` / ``ObjCAViT.py:374-375``) -> [B,Q,h,w]."""
    B, C, h, w = x.shape
    x3 = ops.conv3x3(x, P[pre + "conv3x3.weight"], P[pre + "conv3x3.bias"])
    queries = tok[:, 1:n_query + 1, :]
    assert queries.shape[1] == C or True
    ram = ops.batched_matmul_nn(queries, x3.view(B, C, h * w))
    return ram.view(B, queries.shape[1], h, w)


def regress(P: Params, pre: str, tok: torch.Tensor) -> torch.Tensor:
    """regressor MLP on token 0 (``miniViT.py:33`` / ``ObjCAViT.py:378``) -> raw [B,K]."""
    head = tok[:, 0, :]  # [B,E] view, row stride S*E
    return mlp_leaky(P, pre, head, (0, 2, 4))


def mvit_forward(P: Params, pre: str, x: torch.Tensor, n_query: int = 128, norm: str = "linear"):
    tok = mvit_tokens(P, pre + "patch_transformer.", x)
    ram = range_attention(P, pre, x, tok, n_query)
    widths, _, _ = ops.bins(regress(P, pre + "regressor.", tok), norm)
    return widths, ram


# ----------------------------------------------------------------------------------------------------
# a4-a6: ObjCAViT
# ----------------------------------------------------------------------------------------------------

_OFFSETS_CACHE = {}        # young generation: (device index, values) -> (int32 device tensor, upload event | None)
_OFFSETS_OLD = {}          # previous generation: entries not hit since the last rotation
_OFFSETS_GENERATION = 512  # rotation period (insertions)
offsets_capture_refs = None    # a list while a caller (HeadTrainer graph capture) wants to own every offsets tensor handed out


class ObjectBatch:
    """Ragged per-image object rows in CSR form (device offsets) — replaces the reference's python lists."""

    def __init__(self, rows: torch.Tensor, counts: List[int]):
        self.rows = rows                      # [total, *]
        self.counts = counts
        self.B = len(counts)
        self.n_max = max(counts)
        offs = [0]
        for c in counts:
            offs.append(offs[-1] + c)
        self.offsets_host = offs
        self.offsets = self._to_device(offs)

    def _to_device(self, values):
        """int32 device copy of a small host list.  Cached per (device, values): batches with the same object counts reuse the
        tensor (never written by any kernel), so a steady-state step issues no host->device copy here — which is also what lets
        the training step be captured into a CUDA graph (a copy out of a transient pinned buffer must not be replayed).

        Lifetime: a two-generation cache.  Every ``_OFFSETS_GENERATION`` insertions the young generation becomes the old one
        and the previous old one is dropped; a hit in the old generation promotes the entry.  A tensor is therefore freed only
        after it has gone unused for at least one whole generation (hundreds of steps), long after any kernel on any branch
        stream that read it has finished — no per-stream ``record_stream`` bookkeeping is needed.  Captured CUDA graphs bake the
        addresses into kernel arguments and may be replayed arbitrarily later: while ``offsets_capture_refs`` is a list (set by
        ``HeadTrainer`` around a capture) every tensor handed out is appended to it, and the graph entry keeps that list."""
        if not self.rows.is_cuda:
            return torch.tensor(values, dtype=torch.int32)
        dev = self.rows.device
        key = (dev.index, tuple(values))
        hit = _OFFSETS_CACHE.get(key)
        if hit is None:
            hit = _OFFSETS_OLD.pop(key, None)
            if hit is not None:
                _OFFSETS_CACHE[key] = hit
        if hit is not None:
            t, ev = hit
            if ev is not None:
                if ev.query():
                    _OFFSETS_CACHE[key] = (t, None)
                else:
                    torch.cuda.current_stream(dev).wait_event(ev)
        else:
            if len(_OFFSETS_CACHE) >= _OFFSETS_GENERATION:
                _OFFSETS_OLD.clear()
                _OFFSETS_OLD.update(_OFFSETS_CACHE)
                _OFFSETS_CACHE.clear()
            t = torch.tensor(values, dtype=torch.int32).pin_memory().to(dev, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(dev))
            _OFFSETS_CACHE[key] = (t, ev)
        if offsets_capture_refs is not None:
            offsets_capture_refs.append(t)
        return t

    def uniform_offsets(self, n_max: int) -> torch.Tensor:
        """CSR offsets of the dense [B, n_max] padded layout (every image n_max rows)."""
        return self._to_device(list(range(0, (self.B + 1) * n_max, n_max)))


def marshal_objects(object_features: List[torch.Tensor], object_xywh_list: List[Optional[torch.Tensor]], device):
    """lists -> (features [total,F] fp32, boxes [total,4] fp32, counts).  ``None`` boxes become the
    reference's (-1,-1,-1,-1) placeholder (``ObjCAViT.py:312-313``).  Host cost matters here (one serving step is 0.57 ms):
    tensors that already are fp32 on the device go straight into ONE ``torch.cat`` per list, without a no-op ``.to`` each."""
    counts = [f.shape[0] for f in object_features]

    def ready(t):
        return t if (t.dtype == torch.float32 and t.device == device) else t.to(device=device, dtype=torch.float32)

    feats = torch.cat([ready(f) for f in object_features], 0)
    boxes = torch.cat([torch.full((1, 4), -1.0, device=device) if xy is None else ready(xy) for xy in object_xywh_list], 0)
    assert boxes.shape[0] == feats.shape[0], "one box per object embedding expected"
    return feats, boxes, counts


def grid_geometry(h: int, w: int, patch_size: int):
    return math.ceil(h / patch_size), math.ceil(w / patch_size)


def positional_embeddings(P: Params, pre: str, strategy: str, boxes: torch.Tensor, h: int, w: int,
                          patch_size: int = 16, factor: float = 2.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (object pos [total,E], image pos [S,E]) for the 4 strategies (``ObjCAViT.py:318-328,350-360``).
    The image side does not depend on the batch element, so it is produced once for the S patches."""
    gh_t, gw_t = h // patch_size, w // patch_size           # token grid of the patch conv (floor)
    coords = ops.patch_coords(gh_t, gw_t, patch_size, boxes.device)   # [S,4]
    if strategy in ("learned", "learned_bbox_wh"):
        nin = 4 if strategy == "learned_bbox_wh" else 2
        obj = mlp_leaky(P, pre, boxes[:, 0:nin], (0, 2, 4, 6, 8))
        img = mlp_leaky(P, pre, coords[:, 0:nin], (0, 2, 4, 6, 8))
        return obj, img
    table = P[pre + "positional_encodings"]
    gh, gw = grid_geometry(h, w, patch_size)                 # GridRandom uses ceil (ObjCAViT.py:78-79)
    if strategy == "grid_random":
        obj = ops.grid_sample_pos(table, boxes[:, 0:2].contiguous(), gh, gw, 1, h * factor, w * factor)
        img = ops.grid_sample_pos(table, coords[:, 0:2].contiguous(), gh, gw, 2, float(gh), float(gw))
        return obj, img
    if strategy == "grid_random_roi_align":
        obj = ops.ps_roi_align_pos(table, boxes, gh, gw, 1.0 / (patch_size * factor), True)
        img = ops.ps_roi_align_pos(table, coords, gh, gw, 1.0 / patch_size, True)
        return obj, img
    raise SystemExit("Error: ObjCAViT positional embedding strategy not recognised.")  # ObjCAViT.py:284


ObjTokens = Union[ObjectBatch, torch.Tensor]


def saca_forward(P: Params, pre: str, img_tok: torch.Tensor, objs: ObjTokens, no_obj_sa: bool,
                 need_object_output: bool):
    """``SelfAttnCrossAttn.forward`` (``ObjCAViT.py:167-213``).  ``objs``: ragged :class:`ObjectBatch`
    (block 1) or the dense [B,S,E] tensor block 1 produced (block 2, F6)."""
    B, S, _ = img_tok.shape
    att_img = transformer_encoder(P, pre + "image_transformer_encoder.", img_tok)
    if isinstance(objs, ObjectBatch):
        n_max = objs.n_max
        padded, mask = ops.pad_objects(objs.rows, objs.offsets, B, n_max, n_max, 0.0001, False)   # :180-183
        if no_obj_sa:
            att_obj = padded                                                                     # :186
        else:
            att_obj = transformer_encoder(P, pre + "obj_transformer_encoder.", padded, mask)     # :188
        if n_max > S:
            raise ValueError("more objects than image tokens: the reference's F.pad would crop; unsupported")
        # :192-194 objects LEFT-padded, mask RIGHT-padded to S (F5)
        uni = objs.uniform_offsets(n_max)
        att_obj_s, kpm = ops.pad_objects(att_obj.view(B * n_max, E), uni, B, S, n_max, 0.0001, True,
                                         mask_offsets=objs.offsets)
    else:
        assert objs.shape == (B, S, E)
        att_obj_s = objs.contiguous() if no_obj_sa else transformer_encoder(P, pre + "obj_transformer_encoder.",
                                                                            objs.contiguous())
        kpm = None
    final_img = cross_attention(P, pre + "cross_attn_obj_im.", att_img, att_obj_s, att_img, kpm)   # :195-201
    final_obj = None
    if need_object_output:
        final_obj = cross_attention(P, pre + "cross_attn_im_obj.", att_obj_s, att_img, att_obj_s, None)  # :202-207
    return final_img, final_obj


def objcavit_tokens(P: Params, pre: str, x: torch.Tensor, object_features, object_xywh_list, strategy: str,
                    no_obj_sa: bool, use_2_saca: bool, patch_size: int = 16, mutate_inputs: bool = True):
    """``ObjCAViT.forward`` lines 306-371 -> final image tokens [B,S,E]."""
    B, _, h, w = x.shape
    feats, boxes, counts = marshal_objects(object_features, object_xywh_list, x.device)
    obj_pos, img_pos = positional_embeddings(P, pre + "positional_encoder.", strategy, boxes, h, w, patch_size)
    obj_tok = ops.linear(feats, P[pre + "obj_embedding_layer.weight"], P[pre + "obj_embedding_layer.bias"],
                         residual=obj_pos)                                                       # :330
    batch = ObjectBatch(obj_tok, counts)
    if mutate_inputs:  # the reference overwrites the caller's list with the embedded tokens (F11)
        for i in range(B):
            object_features[i] = obj_tok[batch.offsets_host[i]:batch.offsets_host[i + 1]]
    tok = ops.patch_embed(x, P[pre + "image_embedding_convPxP.weight"], P[pre + "image_embedding_convPxP.bias"],
                          img_pos, patch_size)                                                   # :333,362-364
    tok, obj_out = saca_forward(P, pre + "saca_1.", tok, batch, no_obj_sa, need_object_output=use_2_saca)
    if use_2_saca:
        tok, _ = saca_forward(P, pre + "saca_2.", tok, obj_out, no_obj_sa, need_object_output=False)
    return tok


def objcavit_forward(P: Params, pre: str, x, object_features, object_xywh_list, strategy: str, no_obj_sa: bool,
                     use_2_saca: bool, n_query: int = 128, norm: str = "linear"):
    tok = objcavit_tokens(P, pre, x, object_features, object_xywh_list, strategy, no_obj_sa, use_2_saca)
    ram = range_attention(P, pre, x, tok, n_query)
    widths, _, _ = ops.bins(regress(P, pre + "regressor.", tok), norm)
    return widths, ram


# ----------------------------------------------------------------------------------------------------
# a7: tail
# ----------------------------------------------------------------------------------------------------

def head_tail(P: Params, pre: str, ram: torch.Tensor, widths: torch.Tensor, min_depth: float, max_depth: float):
    """``AdaBins.py:77-87``: conv_out + softmax + bin-centre expectation -> (depth [B,1,h,w], edges [B,K+1])."""
    edges, centres = ops.edges_from_widths(widths, min_depth, max_depth)
    depth = ops.tail(ram, P[pre + "0.weight"], P[pre + "0.bias"], centres)
    return depth, edges
