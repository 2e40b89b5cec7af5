"""Training-mode sequencing of the head: forward with saved activations + explicit backward.

The reference trains the head through torch autograd over ATen ops (``modules/GraphBinsLM.py:121-139``); here every stage of the
forward AND of the backward is a C-ABI kernel of ``libobjcavit_b200.so`` (fp32 SIMT parity path), sequenced by this file:

    tape = forward(P, cfg, x, objects...)          -> depth_pred [B,1,h,w], bin_edges [B,K+1]  (+ saved activations)
    backward(P, G, tape, d_depth, d_edges)         -> accumulates d loss / d parameter into G[key] (same keys as P)

``G`` maps a ``state_dict`` key to a zero-initialised (or already holding earlier micro-batches) fp32 buffer of the parameter's
shape — normally views of ONE flat gradient buffer (``grad_sync.FlatGradBuffer``), so the data-parallel step is a single
all-reduce.  Parameters the reference leaves without gradient simply keep their zeros.

Grad-mode semantics of the reference are reproduced: ``nn.TransformerEncoder`` takes the slow path under autograd, so padded
object rows are NOT zeroed (F7), the left-pad / right-mask quirk (F5) is kept.  Dropout (``DropCtx``): p = 0 is the identity (the
configuration gradient parity is defined on); p > 0 applies counter-based masks at the four sites of every
encoder layer and the backward regenerates them — torch's random stream cannot be reproduced, the masks are statistically the same.  The decoder features are constants of the step
(frozen extractor): no gradient w.r.t. ``x`` is produced.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from . import ops
from .engine import E, HEADS, ObjectBatch, grid_geometry, marshal_objects
from .engine import positional_embeddings as engine_positional_embeddings

Params = Dict[str, torch.Tensor]
RELU, LEAKY, NONE = ops.ACT_RELU, ops.ACT_LEAKY_RELU, ops.ACT_NONE


def _split(M: int, N: int, batch: int, K: int, target_blocks: int = 592) -> int:
    """split-K factor that fills the 148 SMs ~4x with 64x64 tiles without cutting K below 256 per split."""
    tiles = ((M + 63) // 64) * ((N + 63) // 64) * batch
    s = max(1, (target_blocks + tiles - 1) // tiles)
    s = min(s, max(1, K // 256), max(1, 65535 // max(batch, 1)), 64)
    return s


class Touched(dict):
    """Gradient map that records which keys the backward pass wrote: parameters it never touches have NO gradient (``.grad is
    None`` in the reference) and the optimizer must leave them alone (no weight decay either)."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.touched = set()

    def __getitem__(self, k):
        self.touched.add(k)
        return super().__getitem__(k)


# ----------------------------------------------------------------------------------------------------------------------
# Weight-gradient stream.  In the backward of a layer only the DATA gradients are on the critical path; the weight / bias gradient
# GEMMs (half of the token-path GEMM time, each far too small to fill 148 SMs) run on a second stream next to them.  The side
# stream waits for everything issued so far on the main stream before each GEMM; tensors it reads are pinned to it with
# record_stream; `join_side` makes the main stream wait (end of the backward, and before any in-place update of a tensor the side
# stream may still read).  Gradient buffers are written by one stream only per parameter, so the result is bit-identical.
# ----------------------------------------------------------------------------------------------------------------------
ASYNC_WGRAD = True
_side_streams = {}
_side_dirty = set()      # keys with work issued since the last join (join_side waits for these only: under CUDA-graph capture a
                         # wait on a stream that is not part of the capture is illegal)


def _side(device, name: str = "") -> "torch.cuda.Stream":
    """The weight-gradient stream that belongs to the CURRENT launching stream (the object branch below has its own, so the two
    branches do not serialise on one queue); ``name`` selects a further queue of the same launching stream."""
    idx = torch.device(device).index or 0
    key = (idx, torch._C._cuda_getCurrentRawStream(idx), name)
    if key not in _side_streams:
        _side_streams[key] = torch.cuda.Stream(device)
    _side_dirty.add(key)
    return _side_streams[key]


def on_side(fn, *tensors, name: str = ""):
    """Run ``fn()`` (kernel launches only) on the weight-gradient stream after everything issued so far; ``tensors`` = what it reads."""
    if not ASYNC_WGRAD:
        return fn()
    dev = tensors[0].device
    side = _side(dev, name)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        fn()
    for t in tensors:
        t.record_stream(side)


def join_side(device, name: Optional[str] = None) -> None:
    """The current stream waits for every weight-gradient stream of the device (``name``: only the queues of that name)."""
    if not ASYNC_WGRAD:
        return
    idx = torch.device(device).index or 0
    cur = torch.cuda.current_stream(device)
    for key in [k for k in _side_dirty if k[0] == idx and (name is None or k[2] == name)]:
        cur.wait_stream(_side_streams[key])
        _side_dirty.discard(key)


# ----------------------------------------------------------------------------------------------------------------------
# Object-branch stream.  Inside SelfAttnCrossAttn the object encoder (B x N_max rows: a few CTAs per kernel) and the image encoder
# (B x S rows) are independent until the cross attention, and both are chains of latency-bound launches; the object chain runs on
# its own stream next to the image chain, forward and backward.  `fork_branch` marks the point the branch starts from,
# `on_branch` runs a region there, `join_branch` orders the launching stream after it.  No kernel changes: bit-identical.
# ----------------------------------------------------------------------------------------------------------------------
import os as _os
PIXEL_LATE = _os.environ.get("OCV_PIXEL_LATE", "1") == "1"   # conv3x3 / unfold / transpose of the pixel branch start after the patch embedding
CHAIN_PRIORITY = _os.environ.get("OCV_CHAIN_PRIORITY", "1") == "1"
PIXEL_QUEUES = max(1, int(_os.environ.get("OCV_PIXEL_QUEUES", "2")))   # the per-image pixel GEMMs alternate over this many queues
UNFOLD_IN_FORWARD = _os.environ.get("OCV_UNFOLD_FWD", "1") == "1"
BRANCH_OBJECTS = int(_os.environ.get("OCV_BRANCH", "63"))  # bit 0: object branch forward, bit 1: backward, bit 2: pixel prologue, bit 3: chamfer loss, bit 4: regressor, bit 5: per-image pixel GEMMs over two queues
_branch_streams = {}


def _branch(device, name: str = "objects") -> "torch.cuda.Stream":
    key = (torch.device(device).index or 0, name)
    if key not in _branch_streams:
        # Branches that carry pieces of the dependency chain (object encoder, regressor, loss) get a high-priority stream, the
        # "pixel" queues (whole-GPU HBM-bound work hoisted off the chain) stay at the default = lowest priority: when SM slots
        # free up, the block scheduler serves the chain first.  (HeadTrainer captures the step on a high-priority stream too; stream
        # priorities are recorded into the kernel nodes of the graph.)
        prio = 0 if (name.startswith("pixel") or not CHAIN_PRIORITY) else -1
        _branch_streams[key] = torch.cuda.Stream(device, priority=prio)
    return _branch_streams[key]


def fork_branch(device, bit: int = 1):
    """-> token marking "everything issued so far on the current stream" (None when branching is off)."""
    if not (BRANCH_OBJECTS & bit):
        return None
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(device))
    return ev


class on_branch:
    """``with on_branch(dev, fork, *inputs):`` — launches inside go to the object-branch stream, ordered after ``fork``;
    ``inputs`` are the tensors of other streams the region reads."""

    def __init__(self, device, fork, *inputs, name: str = "objects"):
        self.device, self.fork, self.inputs, self.ctx, self.name = device, fork, inputs, None, name

    def __enter__(self):
        if self.fork is None:
            return self
        br = _branch(self.device, self.name)
        br.wait_event(self.fork)
        self.ctx = torch.cuda.stream(br)
        self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
            br = _branch(self.device, self.name)
            for t in self.inputs:
                if t is not None:
                    t.record_stream(br)
        return False


def join_branch(device, fork, *outputs, name: str = "objects") -> None:
    """The current stream waits for the branch; ``outputs`` = tensors made there that the current stream reads from now on."""
    if fork is None:
        return
    cur = torch.cuda.current_stream(device)
    cur.wait_stream(_branch(device, name))
    for t in outputs:
        if t is not None:
            t.record_stream(cur)


def _wgrad(dy, x, gw, gb) -> None:
    on_side(lambda: ops.linear_wgrad(dy, x, gw, gb, beta=1.0), dy, x)


class DropCtx:
    """Dropout state of one training forward: probability ``p`` (0 = identity: no dropout kernels are launched) and a base seed;
    every dropout site of the forward draws its own 64-bit site seed with :meth:`site`, recorded on the tape for the backward.
    Sites per encoder layer, in this order (nn.TransformerEncoderLayer): attention probabilities, after out_proj
    (dropout1), after the FFN activation (dropout), after linear2 (dropout2).  The cross-attention MHAs have dropout 0."""

    def __init__(self, p: float = 0.0, seed: int = 0):
        self.p = float(p)
        self.seed = int(seed) & 0xFFFFFFFFFFFF
        self.count = 0

    def site(self) -> int:
        self.count += 1
        return ((self.seed * 0x9E3779B1 + self.count * 0x632BE5AB) ^ (self.count << 48)) & 0xFFFFFFFFFFFFFFFF


NO_DROP = DropCtx(0.0, 0)


# ----------------------------------------------------------------------------------------------------------------------
# MLP chains (positional encoder, regressor)
# ----------------------------------------------------------------------------------------------------------------------

FAST_ROWS = False       # set per forward (precision "bf16"): few-row Linear layers through ocv_linear_rows_f32 instead of tiled GEMMs


def mlp_fwd(P: Params, pre: str, x: torch.Tensor, idxs):
    saved = []
    for j, i in enumerate(idxs):
        act = LEAKY if j < len(idxs) - 1 else NONE
        w = P[f"{pre}{i}.weight"]
        if FAST_ROWS and ops.linear_rows_ok(x, w):        # 8-500 rows: ~5 us per layer against 10-45 us for the tiled fp32 GEMM
            y = ops.linear_rows(x, w, P[f"{pre}{i}.bias"], leaky=act == LEAKY)
        else:
            y = ops.linear(x, w, P[f"{pre}{i}.bias"], act=act)
        saved.append((x, y))
        x = y
    return x, saved


def mlp_bwd(P: Params, G: Params, pre: str, idxs, saved, dy: torch.Tensor, dx_out: Optional[torch.Tensor] = None):
    """dy: gradient of the chain output (dense 2-D).  ``dx_out``: where the gradient w.r.t. the chain input goes (a 2-D view,
    overwritten); ``None`` = the input is a constant of the step (boxes / patch coordinates)."""
    for j in range(len(idxs) - 1, -1, -1):
        i = idxs[j]
        x, y = saved[j]
        if j < len(idxs) - 1:
            dy = ops.act_bwd(dy, y, LEAKY)
        _wgrad(dy, x, G[f"{pre}{i}.weight"], G[f"{pre}{i}.bias"])      # weight-gradient stream: only dgrad is on the chain
        if j > 0:
            dy = ops.linear_dgrad(dy, P[f"{pre}{i}.weight"])
        elif dx_out is not None:
            ops.linear_dgrad(dy, P[f"{pre}{i}.weight"], out=dx_out)


# ----------------------------------------------------------------------------------------------------------------------
# nn.TransformerEncoderLayer (post-norm) / nn.TransformerEncoder
# ----------------------------------------------------------------------------------------------------------------------

NATIVE_LAYER = True      # one C call per encoder layer and direction (csrc/train_layer.cu) instead of ~30 Python-sequenced ones
TOKEN_BF16 = _os.environ.get("OCV_TOKEN_BF16", "1") == "1"   # precision "bf16": the encoder layers' GEMMs on tcgen05 as well


class LazyBf16(dict):
    """bf16 copies of fp32 parameters, cast on first use (one forward + backward); ``HeadTrainer`` passes a mapping of views into ONE
    bf16 mirror of its flat parameter buffer instead (a single cast kernel per step)."""

    def __init__(self, P):
        super().__init__()
        self.P = P

    def __missing__(self, key):
        t = ops.cast_bf16(self.P[key].contiguous())
        self[key] = t
        return t



def encoder_layer_fwd(P: Params, pre: str, x: torch.Tensor, mask, drop: DropCtx = NO_DROP, P16=None, x16=None):
    """-> (out [B,S,E], saved) — with ``P16`` (bf16 weight copies) the tensor-core layer: returns (out, saved, out16)."""
    B, S, _ = x.shape
    if P16 is not None:
        p = drop.p
        sites = (drop.site(), drop.site(), drop.site(), drop.site()) if p > 0 else None
        xc = x.contiguous()
        if x16 is None:
            x16 = ops.cast_bf16(xc)
        out, out16, sv = ops.encoder_layer_fwd_train_bf16([P[pre + k] for k in ops.LAYER_KEYS], [P16[pre + k] for k in ops.LAYER_W16_KEYS],
                                                         xc, x16, mask, p, sites)
        return out, ("bf16", sv, mask, (p, sites)), out16
    if NATIVE_LAYER:
        p = drop.p
        sites = (drop.site(), drop.site(), drop.site(), drop.site()) if p > 0 else None
        xc = x.contiguous()
        out, (qkv, att, y, x1, h, y2) = ops.encoder_layer_fwd_train([P[pre + k] for k in ops.LAYER_KEYS], xc, mask, p, sites)
        return out, (xc.view(B * S, E), qkv, att, y, x1, h, y2, mask, (p, sites))
    x2 = x.view(B * S, E)
    p = drop.p
    qkv = ops.linear(x2, P[pre + "self_attn.in_proj_weight"], P[pre + "self_attn.in_proj_bias"]).view(B, S, 3 * E)
    if p > 0:
        sites = (drop.site(), drop.site(), drop.site(), drop.site())
        att = ops.mha_core(qkv[:, :, 0:E], qkv[:, :, E:2 * E], qkv[:, :, 2 * E:], mask, HEADS, drop=(p, sites[0]))
        sa = ops.linear(att.view(B * S, E), P[pre + "self_attn.out_proj.weight"], P[pre + "self_attn.out_proj.bias"])
        y = ops.dropout(sa, p, sites[1], residual=x2, out=sa)                       # x + dropout1(self-attention)
        x1 = ops.add_layernorm(y, None, P[pre + "norm1.weight"], P[pre + "norm1.bias"])
        h = ops.linear(x1, P[pre + "linear1.weight"], P[pre + "linear1.bias"], act=RELU)
        h = ops.dropout(h, p, sites[2], out=h)                                      # dropout(relu(linear1)): what linear2 sees
        f = ops.linear(h, P[pre + "linear2.weight"], P[pre + "linear2.bias"])
        y2 = ops.dropout(f, p, sites[3], residual=x1, out=f)                        # x1 + dropout2(linear2)
    else:
        sites = None
        att = ops.mha_core(qkv[:, :, 0:E], qkv[:, :, E:2 * E], qkv[:, :, 2 * E:], mask, HEADS)
        y = ops.linear(att.view(B * S, E), P[pre + "self_attn.out_proj.weight"], P[pre + "self_attn.out_proj.bias"], residual=x2)
        x1 = ops.add_layernorm(y, None, P[pre + "norm1.weight"], P[pre + "norm1.bias"])
        h = ops.linear(x1, P[pre + "linear1.weight"], P[pre + "linear1.bias"], act=RELU)
        y2 = ops.linear(h, P[pre + "linear2.weight"], P[pre + "linear2.bias"], residual=x1)
    out = ops.add_layernorm(y2, None, P[pre + "norm2.weight"], P[pre + "norm2.bias"])
    return out.view(B, S, E), (x2, qkv, att, y, x1, h, y2, mask, (p, sites))


def encoder_layer_bwd(P: Params, G: Params, pre: str, saved, dout: torch.Tensor, P16=None) -> torch.Tensor:
    if saved[0] == "bf16":
        _, sv, mask, (p, sites) = saved
        side = _side(dout.device) if ASYNC_WGRAD else None
        return ops.encoder_layer_bwd_train_bf16([P[pre + k] for k in ops.LAYER_KEYS], [P16[pre + k] for k in ops.LAYER_W16_KEYS],
                                                [G[pre + k] for k in ops.LAYER_KEYS], sv, mask, dout, p, sites, side)
    x2, qkv, att, y, x1, h, y2, mask, (p, sites) = saved
    B, S, _ = qkv.shape
    if NATIVE_LAYER:
        side = _side(dout.device) if ASYNC_WGRAD else None
        return ops.encoder_layer_bwd_train([P[pre + k] for k in ops.LAYER_KEYS], [G[pre + k] for k in ops.LAYER_KEYS], x2, qkv, att,
                                           y, x1, h, y2, mask, dout, p, sites, side)
    d = dout.reshape(B * S, E)
    dy2 = ops.layernorm_bwd(d, y2, P[pre + "norm2.weight"], G[pre + "norm2.weight"], G[pre + "norm2.bias"], beta=1.0)
    df = ops.dropout(dy2, p, sites[3]) if p > 0 else dy2                         # gradient of linear2's output
    _wgrad(df, h, G[pre + "linear2.weight"], G[pre + "linear2.bias"])
    dh = ops.linear_dgrad(df, P[pre + "linear2.weight"])
    if p > 0:
        ops.dropout(dh, p, sites[2], out=dh)                                     # h is the post-dropout activation: h > 0 <=> kept & positive
    dh = ops.act_bwd(dh, h, RELU)
    _wgrad(dh, x1, G[pre + "linear1.weight"], G[pre + "linear1.bias"])
    if p == 0:
        join_side(d.device)                                                      # df aliases dy2, which is updated in place next
    ops.linear_dgrad(dh, P[pre + "linear1.weight"], out=dy2, beta=1.0)           # dy2 <- d loss / d x1 (residual + FFN)
    dy = ops.layernorm_bwd(dy2, y, P[pre + "norm1.weight"], G[pre + "norm1.weight"], G[pre + "norm1.bias"], beta=1.0)
    dsa = ops.dropout(dy, p, sites[1]) if p > 0 else dy                          # gradient of out_proj's output
    _wgrad(dsa, att.view(B * S, E), G[pre + "self_attn.out_proj.weight"], G[pre + "self_attn.out_proj.bias"])
    datt = ops.linear_dgrad(dsa, P[pre + "self_attn.out_proj.weight"]).view(B, S, E)
    dqkv = ops.mha_core_bwd(qkv[:, :, 0:E], qkv[:, :, E:2 * E], qkv[:, :, 2 * E:], att, datt, mask, HEADS, packed=True,
                            drop=(p, sites[0]) if p > 0 else None)[3]
    dqkv = dqkv.view(B * S, 3 * E)
    _wgrad(dqkv, x2, G[pre + "self_attn.in_proj_weight"], G[pre + "self_attn.in_proj_bias"])
    if p == 0:
        join_side(d.device)                                                      # dsa aliases dy, which is updated in place next
    ops.linear_dgrad(dqkv, P[pre + "self_attn.in_proj_weight"], out=dy, beta=1.0)  # dy <- d loss / d x (residual + attention)
    return dy.view(B, S, E)


def encoder_fwd(P: Params, pre: str, x: torch.Tensor, mask=None, num_layers: int = 4, drop: DropCtx = NO_DROP, P16=None):
    saved = []
    x16 = None
    for i in range(num_layers):
        if P16 is not None:
            x, sv, x16 = encoder_layer_fwd(P, f"{pre}layers.{i}.", x, mask, drop, P16, x16)
        else:
            x, sv = encoder_layer_fwd(P, f"{pre}layers.{i}.", x, mask, drop)
        saved.append(sv)
    return x, saved


def encoder_bwd(P: Params, G: Params, pre: str, saved, dout: torch.Tensor, P16=None) -> torch.Tensor:
    for i in range(len(saved) - 1, -1, -1):
        dout = encoder_layer_bwd(P, G, f"{pre}layers.{i}.", saved[i], dout, P16)
    return dout


# ----------------------------------------------------------------------------------------------------------------------
# nn.MultiheadAttention used as cross attention
# ----------------------------------------------------------------------------------------------------------------------

def cross_fwd(P: Params, pre: str, q_in, k_in, v_in, mask):
    w, b = P[pre + "in_proj_weight"], P[pre + "in_proj_bias"]
    B, Sq, _ = q_in.shape
    Sk = k_in.shape[1]
    q = ops.linear(q_in.reshape(B * Sq, E), w[0:E], b[0:E]).view(B, Sq, E)
    k = ops.linear(k_in.reshape(B * Sk, E), w[E:2 * E], b[E:2 * E]).view(B, Sk, E)
    v = ops.linear(v_in.reshape(B * Sk, E), w[2 * E:], b[2 * E:]).view(B, Sk, E)
    att = ops.mha_core(q, k, v, mask, HEADS)
    out = ops.linear(att.view(B * Sq, E), P[pre + "out_proj.weight"], P[pre + "out_proj.bias"]).view(B, Sq, E)
    return out, (q_in, k_in, v_in, q, k, v, att, mask)


def cross_bwd(P: Params, G: Params, pre: str, saved, dout: torch.Tensor, acc_q, acc_k, acc_v):
    """``acc_*`` = (tensor [B,S,E] contiguous, beta): where the gradient w.r.t. the query / key / value INPUT is written
    (beta 0) or accumulated (beta 1) — the same tensor may be named twice when one input plays two roles."""
    q_in, k_in, v_in, q, k, v, att, mask = saved
    B, Sq, _ = q.shape
    Sk = k.shape[1]
    w = P[pre + "in_proj_weight"]
    gw, gb = G[pre + "in_proj_weight"], G[pre + "in_proj_bias"]
    d = dout.reshape(B * Sq, E)
    _wgrad(d, att.view(B * Sq, E), G[pre + "out_proj.weight"], G[pre + "out_proj.bias"])
    datt = ops.linear_dgrad(d, P[pre + "out_proj.weight"]).view(B, Sq, E)
    dq, dk, dv = ops.mha_core_bwd(q, k, v, att, datt, mask, HEADS)
    for dz, z_in, lo, acc, rows in ((dq, q_in, 0, acc_q, B * Sq), (dk, k_in, E, acc_k, B * Sk), (dv, v_in, 2 * E, acc_v, B * Sk)):
        dz2 = dz.view(rows, E)
        _wgrad(dz2, z_in.reshape(rows, E), gw[lo:lo + E], gb[lo:lo + E])
        if acc is not None:
            ops.linear_dgrad(dz2, w[lo:lo + E], out=acc[0].view(rows, E), beta=acc[1])


# ----------------------------------------------------------------------------------------------------------------------
# nn.MultiheadAttention as cross attention on the tensor cores (precision "bf16"): projections and their gradients through
# ocv_gemm_bf16_v2, the attention core and its backward through the tcgen05 kernels (csrc/attn_tc.cu).  Sq, Sk <= 256.
# ----------------------------------------------------------------------------------------------------------------------

def _cross_tc_ok(Sq: int, Sk: int) -> bool:
    return TOKEN_BF16 and Sq <= 256 and Sk <= 256


def cross_fwd16(P: Params, P16, pre: str, q_in, k_in, v_in, mask):
    w16, b = P16[pre + "in_proj_weight"], P[pre + "in_proj_bias"]
    B, Sq, _ = q_in.shape
    Sk = k_in.shape[1]
    q_in16 = ops.cast_bf16(q_in.contiguous()).view(B * Sq, E)
    k_in16 = ops.cast_bf16(k_in.contiguous()).view(B * Sk, E)
    v_in16 = q_in16 if v_in is q_in else (k_in16 if v_in is k_in else ops.cast_bf16(v_in.contiguous()).view(-1, E))
    proj = lambda x16, lo: ops.gemm_bf16_v2(x16, w16[lo:lo + E], bias=b[lo:lo + E], want32=False, want16=True)[2]
    q16 = proj(q_in16, 0).view(B, Sq, E)
    k16 = proj(k_in16, E).view(B, Sk, E)
    v16 = proj(v_in16, 2 * E).view(B, Sk, E)
    att16, _, stats = ops.mha_core_tc_bf16(q16, k16, v16, mask, HEADS, want_stats=True)
    out = ops.gemm_bf16_v2(att16.view(B * Sq, E), P16[pre + "out_proj.weight"], bias=P[pre + "out_proj.bias"])[1].view(B, Sq, E)
    return out, ("bf16", q_in16, k_in16, v_in16, q16, k16, v16, att16, stats, mask)


def cross_bwd16(P: Params, P16, G: Params, pre: str, saved, dout: torch.Tensor, acc_q, acc_k, acc_v):
    _, q_in16, k_in16, v_in16, q16, k16, v16, att16, stats, mask = saved
    B, Sq, _ = q16.shape
    Sk = k16.shape[1]
    w16 = P16[pre + "in_proj_weight"]
    gw, gb = G[pre + "in_proj_weight"], G[pre + "in_proj_bias"]
    d = dout.reshape(B * Sq, E).contiguous()
    d16 = ops.cast_bf16(d)
    a16 = att16.view(B * Sq, E)
    go, gob = G[pre + "out_proj.weight"], G[pre + "out_proj.bias"]

    def out_wgrad():
        ops.gemm_bf16_v2(d16, a16, a_mn=True, b_mn=True, residual=go, out32=go)            # dWout += d^T att
        ops.reduce_channels(d, gob, 1, B * Sq, E, 1, beta=1.0)

    on_side(out_wgrad, d16, a16, d)
    datt16 = ops.gemm_bf16_v2(d16, P16[pre + "out_proj.weight"], b_mn=True, want32=False, want16=True)[2].view(B, Sq, E)
    (dq, dk, dv), (dq16, dk16, dv16) = ops.mha_core_bwd_tc_bf16(q16, k16, v16, att16, datt16, stats, mask, HEADS)
    for dz, dz16, z16, lo, acc, rows in ((dq, dq16, q_in16, 0, acc_q, B * Sq), (dk, dk16, k_in16, E, acc_k, B * Sk),
                                          (dv, dv16, v_in16, 2 * E, acc_v, B * Sk)):
        dz2, dz216 = dz.reshape(rows, E), dz16.reshape(rows, E)

        def in_wgrad(dz2=dz2, dz216=dz216, z16=z16, lo=lo):
            ops.gemm_bf16_v2(dz216, z16, a_mn=True, b_mn=True, residual=gw[lo:lo + E], out32=gw[lo:lo + E])
            ops.reduce_channels(dz2.contiguous(), gb[lo:lo + E], 1, rows, E, 1, beta=1.0)

        on_side(in_wgrad, dz216, z16, dz2)
        if acc is not None:
            tgt = acc[0].view(rows, E)
            ops.gemm_bf16_v2(dz216, w16[lo:lo + E], b_mn=True, residual=tgt if acc[1] else None, out32=tgt)


# ----------------------------------------------------------------------------------------------------------------------
# SelfAttnCrossAttn (modules/ObjCAViT.py:167-213)
# ----------------------------------------------------------------------------------------------------------------------

def saca_fwd(P: Params, pre: str, img_tok: torch.Tensor, objs, no_obj_sa: bool, need_object_output: bool, drop: DropCtx = NO_DROP,
             P16=None):
    B, S, _ = img_tok.shape
    dev = img_tok.device
    if isinstance(objs, ObjectBatch):
        n_max = objs.n_max
        if n_max > S:
            raise ValueError("more objects than image tokens: the reference's F.pad would crop; unsupported")
        uni = objs.uniform_offsets(n_max)            # made on the launching stream BEFORE the fork: the branch reads it
    fork = fork_branch(dev)
    att_img, sv_img = encoder_fwd(P, pre + "image_transformer_encoder.", img_tok, drop=drop, P16=P16)
    sv_obj = None
    if isinstance(objs, ObjectBatch):
        with on_branch(dev, fork, objs.rows, objs.offsets, uni):
            padded, mask = ops.pad_objects(objs.rows, objs.offsets, B, n_max, n_max, 0.0001, False)        # :180-183
            att_obj = padded
            if not no_obj_sa:
                att_obj, sv_obj = encoder_fwd(P, pre + "obj_transformer_encoder.", padded, mask, drop=drop, P16=P16)   # :188 (slow path)
            att_obj_s, kpm = ops.pad_objects(att_obj.reshape(B * n_max, E), uni, B, S, n_max, 0.0001, True,
                                             mask_offsets=objs.offsets)                                     # :192-194 (F5)
        geom = ("ragged", objs, uni, n_max)
    else:
        assert objs.shape == (B, S, E)
        with on_branch(dev, fork, objs):
            att_obj_s = objs.contiguous()
            if not no_obj_sa:
                att_obj_s, sv_obj = encoder_fwd(P, pre + "obj_transformer_encoder.", att_obj_s, drop=drop, P16=P16)
        kpm = None
        geom = ("dense",)
    join_branch(dev, fork, att_obj_s, kpm)
    tc = P16 is not None and _cross_tc_ok(S, S)
    if tc:
        final_img, sv_c1 = cross_fwd16(P, P16, pre + "cross_attn_obj_im.", att_img, att_obj_s, att_img, kpm)   # :195-201
    else:
        final_img, sv_c1 = cross_fwd(P, pre + "cross_attn_obj_im.", att_img, att_obj_s, att_img, kpm)          # :195-201
    final_obj, sv_c2 = None, None
    if need_object_output:
        if tc:
            final_obj, sv_c2 = cross_fwd16(P, P16, pre + "cross_attn_im_obj.", att_obj_s, att_img, att_obj_s, None)
        else:
            final_obj, sv_c2 = cross_fwd(P, pre + "cross_attn_im_obj.", att_obj_s, att_img, att_obj_s, None)   # :202-207
    return final_img, final_obj, (sv_img, sv_obj, sv_c1, sv_c2, geom, (B, S))


def saca_bwd(P: Params, G: Params, pre: str, saved, d_final_img: torch.Tensor, d_final_obj: Optional[torch.Tensor], P16=None):
    """-> (d image tokens [B,S,E], d objects: ragged rows [total,E] (block 1) or dense [B,S,E] (block 2))."""
    sv_img, sv_obj, sv_c1, sv_c2, geom, (B, S) = saved
    dev = d_final_img.device
    d_att_img = torch.empty(B, S, E, device=dev, dtype=torch.float32)
    d_att_obj_s = torch.empty(B, S, E, device=dev, dtype=torch.float32)
    xb = (lambda *a: cross_bwd16(P, P16, G, *a)) if sv_c1[0] == "bf16" else (lambda *a: cross_bwd(P, G, *a))
    xb(pre + "cross_attn_obj_im.", sv_c1, d_final_img, (d_att_img, 0.0), (d_att_obj_s, 0.0), (d_att_img, 1.0))
    if sv_c2 is not None and d_final_obj is not None:
        xb(pre + "cross_attn_im_obj.", sv_c2, d_final_obj, (d_att_obj_s, 1.0), (d_att_img, 1.0), (d_att_obj_s, 1.0))
    fork = fork_branch(dev, 2)
    d_img_tok = encoder_bwd(P, G, pre + "image_transformer_encoder.", sv_img, d_att_img, P16)
    with on_branch(dev, fork, d_att_obj_s):
        if geom[0] == "ragged":
            _, objs, uni, n_max = geom
            d_att_obj = ops.unpad_objects(d_att_obj_s, uni, B * n_max, n_max, True).view(B, n_max, E)
            if sv_obj is not None:
                d_att_obj = encoder_bwd(P, G, pre + "obj_transformer_encoder.", sv_obj, d_att_obj, P16)
            d_objs = ops.unpad_objects(d_att_obj.contiguous(), objs.offsets, objs.offsets_host[-1], n_max, False)
        else:
            d_objs = d_att_obj_s
            if sv_obj is not None:
                d_objs = encoder_bwd(P, G, pre + "obj_transformer_encoder.", sv_obj, d_att_obj_s, P16)
    join_branch(dev, fork, d_objs)
    return d_img_tok, d_objs


# ----------------------------------------------------------------------------------------------------------------------
# whole head
# ----------------------------------------------------------------------------------------------------------------------

class Tape:
    """Saved activations of one training forward (plain attributes; freed when the object dies)."""


def forward(P: Params, cfg: dict, x: torch.Tensor, object_features=None, object_xywh_list=None, mutate_inputs: bool = True,
            patch_size: int = 16, n_query: int = 128, drop: DropCtx = NO_DROP, P16=None, xh=None):
    """-> (depth_pred [B,1,h,w], bin_edges [B,K+1], tape).  ``cfg``: kind ('adabins' | 'graphbins'), strategy, no_obj_sa,
    use_2_saca, min_depth, max_depth.  ``xh``: the features already converted to NHWC bf16 [B,h,w,C] (tensor-core mode only: every
    consumer of the features reads that copy, so ``x`` is then used for its shape alone and may be a storage-less placeholder —
    ``features_nhwc_ok`` says when)."""
    t = Tape()
    t.cfg, t.x = cfg, x
    assert xh is None or features_nhwc_ok(cfg, x.shape), "xh= needs the tensor-core path with every reader of x on the NHWC copy"
    t.xh_given = xh
    # precision "bf16": the encoder layers' GEMMs run on tcgen05 with bf16 copies of the weights (P16: given by the trainer as views
    # of one bf16 mirror of its flat parameter buffer, else cast lazily here)
    if cfg.get("pixel") == "bf16" and TOKEN_BF16 and NATIVE_LAYER:
        P16 = P16 if P16 is not None else LazyBf16(P)
    else:
        P16 = None
    t.P16 = P16
    global FAST_ROWS
    FAST_ROWS = cfg.get("pixel") == "bf16"
    B, C, h, w = x.shape
    t.shape = (B, C, h, w)
    # ``cfg["prefix"]``: key prefix of the mViT / ObjCAViT parameters ("" when the inner module itself is the boundary)
    pre = cfg.get("prefix", "adaptive_bins_layer." if cfg["kind"] == "adabins" else "objcavit.")
    if cfg.get("pixel") == "bf16":
        t.late_P, t.late_pre = P, pre
        # the patch-embedding weights in the tensor-core kernel's (ky, kx, c) order: depends on the weights only — its own branch
        # from the very start of the step instead of 25 us on the chain right before the patch embedding
        wkey = pre + ("patch_transformer.embedding_convPxP.weight" if cfg["kind"] == "adabins" else "image_embedding_convPxP.weight")
        t.wp_fork = fork_branch(x.device, 4)
        with on_branch(x.device, t.wp_fork, P[wkey], name="pack"):
            t.wp = ops.pack_patch_weights_bf16(P[wkey])
        _pixel_prologue(P, pre, t, x)
    if cfg["kind"] == "adabins":
        pt = pre + "patch_transformer."
        tok0 = _patch_embed(t, x, P[pt + "embedding_convPxP.weight"], P[pt + "embedding_convPxP.bias"],
                            P[pt + "positional_encodings"], patch_size)
        tok, t.sv_enc = encoder_fwd(P, pt + "transformer_encoder.", tok0, drop=drop, P16=P16)
    else:
        strategy = cfg["strategy"]
        feats, boxes, counts = marshal_objects(object_features, object_xywh_list, x.device)
        coords = ops.patch_coords(h // patch_size, w // patch_size, patch_size, x.device)
        pe = pre + "positional_encoder."
        if strategy in ("learned", "learned_bbox_wh"):
            nin = 4 if strategy == "learned_bbox_wh" else 2
            # ONE chain for the object rows and the patch rows (the same MLP): half the launches of two chains, forward and backward
            t.n_obj_rows = boxes.shape[0]
            pos, t.sv_mlp_pos = mlp_fwd(P, pe, torch.cat([boxes[:, 0:nin], coords[:, 0:nin]], 0), (0, 2, 4, 6, 8))
            obj_pos, img_pos = pos[:t.n_obj_rows], pos[t.n_obj_rows:]
        else:                                # table look-ups (GridRandomPositionalEmbeddings): the inference kernels, boxes kept
            obj_pos, img_pos = engine_positional_embeddings(P, pe, strategy, boxes, h, w, patch_size)
            t.pos_inputs = (boxes, coords, patch_size)
        obj_tok = ops.linear(feats, P[pre + "obj_embedding_layer.weight"], P[pre + "obj_embedding_layer.bias"], residual=obj_pos)
        t.feats = feats
        batch = ObjectBatch(obj_tok, counts)
        if mutate_inputs:                                                            # F11
            for i in range(B):
                object_features[i] = obj_tok[batch.offsets_host[i]:batch.offsets_host[i + 1]]
        tok0 = _patch_embed(t, x, P[pre + "image_embedding_convPxP.weight"], P[pre + "image_embedding_convPxP.bias"],
                            img_pos.contiguous(), patch_size)
        two = bool(cfg.get("use_2_saca"))
        tok, obj_out, t.sv_saca1 = saca_fwd(P, pre + "saca_1.", tok0, batch, bool(cfg.get("no_obj_sa")), two, drop, P16)
        if two:
            tok, _, t.sv_saca2 = saca_fwd(P, pre + "saca_2.", tok, obj_out, bool(cfg.get("no_obj_sa")), False, drop, P16)
    t.pre = pre
    S = tok.shape[1]
    t.tok = tok
    Q = min(n_query, S - 1)
    t.pixel = cfg.get("pixel", "fp32")
    if cfg.get("surface") == "module":
        # mViT.forward / ObjCAViT.forward surface (miniViT.py:22-44, ObjCAViT.py:373-390): (bin_widths_normed, range_attention_maps)
        # are the outputs, the tail belongs to the caller (AdaBins.py:77-87) — fp32 kernels, ram materialised because it IS the output
        t.y_raw, t.sv_reg = mlp_fwd(P, pre + "regressor.", tok[:, 0, :], (0, 2, 4))
        widths, _, _ = ops.bins(t.y_raw, "linear")
        t.Q = Q
        t.x3 = ops.conv3x3(x, P[pre + "conv3x3.weight"], P[pre + "conv3x3.bias"])
        ram = ops.batched_matmul_nn(tok[:, 1:Q + 1, :], t.x3.view(B, C, h * w)).view(B, Q, h, w)
        return widths, ram, t
    # regressor + bins: only the LAST kernel of the pixel path (the tail) needs the bin centres, so in the tensor-core mode the chain
    # runs on a branch next to the logits GEMMs
    reg_fork = fork_branch(x.device, 16) if t.pixel == "bf16" else None
    with on_branch(x.device, reg_fork, tok, name="regressor"):
        t.y_raw, t.sv_reg = mlp_fwd(P, pre + "regressor.", tok[:, 0, :], (0, 2, 4))
        _, edges, t.centres = ops.bins(t.y_raw, "linear", cfg["min_depth"], cfg["max_depth"])
        t.edges_ready = fork_branch(x.device, 8)   # the bin edges exist before the pixel path ends: the trainer runs the chamfer
                                                   # loss from here on a branch, under the pixel forward (trainer.forward_backward)
    t.reg_fork = reg_fork
    t.Q = Q
    if t.pixel == "bf16":
        depth = _pixel_fwd_bf16(P, pre, t, x, tok, Q, edges)
    else:
        t.x3 = ops.conv3x3(x, P[pre + "conv3x3.weight"], P[pre + "conv3x3.bias"])
        t.ram = ops.batched_matmul_nn(tok[:, 1:Q + 1, :], t.x3.view(B, C, h * w)).view(B, Q, h, w)
        depth = ops.tail(t.ram, P["conv_out.0.weight"], P["conv_out.0.bias"], t.centres)
    return depth, edges, t


# ----------------------------------------------------------------------------------------------------------------------
# Pixel path on tensor cores (cfg["pixel"] == "bf16"): conv3x3, logits and every pixel-sized gradient GEMM are tcgen05 kernels on
# bf16 operands (fp32 accumulation), joined by the HBM-bound helpers of csrc/train_pixel_bf16.cu.  The 128 range-attention
# channels are folded away as in the inference kernel: logits[b] = Mb[b] x3[b] + bout with Mb[b] = Wout @ queries[b]  (K x C),
# so neither ram nor d ram is ever formed.  Every GEMM sees its contraction dimension contiguous (K-major operands):
#   logits[b] [K,P]   = Mb[b] [K,C]     @ x3[b] [P,C]^T            (x3 NHWC, as the conv kernel writes it)
#   dx3T[:, b] [C,P]  = Mb[b]^T [C,K]   @ dl[b] [P,K]^T            (dl pixel-major: transposed copy of dlT)
#   dMb[b] [K,C]      = dlT[b] [K,P]    @ x3T[b] [C,P]^T           (split-K over pixels)
#   dW3 [C, 9C]       = dx3T [C, B P]   @ col [9C, B P]^T          (ONE split-K GEMM for the batch; col = im2col of the planar map)
# Gradient precision is bf16 (operands rounded to 8 mantissa bits); the fp32 path stays the parity reference.
# ----------------------------------------------------------------------------------------------------------------------

def _patch_embed(t, x: torch.Tensor, w: torch.Tensor, bias, pos, patch_size: int) -> torch.Tensor:
    """Patch-embedding conv + bias + positional rows -> tokens [B,S,E] fp32.  Tensor-core pixel mode: the tcgen05 patch-embed
    kernel on the NHWC bf16 copy of the features (kept on the tape for the conv3x3 kernel)."""
    if t.cfg.get("pixel") == "bf16":
        cur = torch.cuda.current_stream(x.device)
        if t.xh_ready is not None:                   # t.xh was made on the pixel branch (_pixel_prologue)
            cur.wait_event(t.xh_ready)
            t.xh.record_stream(cur)
        join_branch(x.device, t.wp_fork, t.wp, name="pack")
        tok = ops.patch_embed_bf16(t.xh, t.wp, bias, pos, patch_size, want32=True, want16=False)[0]
        t.wp = None
        if getattr(t, "pixel_rest_pending", False):      # the rest of the pixel branch, ordered after the patch embedding
            t.pixel_rest_pending = False
            t.xh_keep = t.xh
            late = fork_branch(x.device, 4)
            with on_branch(x.device, late, None if getattr(t, "xh_given", None) is not None else x, t.xh, name="pixel"):
                _pixel_prologue_rest(t.late_P, t.late_pre, t, x)
        t.xh = None
        return tok
    return ops.patch_embed(x, w, bias, pos, patch_size)


def features_nhwc_ok(cfg: dict, shape) -> bool:
    """True when no kernel of the step reads the fp32 NCHW features except the layout conversion (so a caller may hand the NHWC
    bf16 copy in as ``forward(..., xh=)``): tensor-core pixel + token path, unfold in the forward, implicit conv3x3 weight gradient."""
    B, C, h, w = shape
    return (cfg.get("pixel") == "bf16" and cfg.get("surface", "head") == "head" and UNFOLD_IN_FORWARD and IMPLICIT_DW3 and C == 128
            and w % 8 == 0)


def _pixel_prologue(P: Params, pre: str, t, x: torch.Tensor) -> None:
    """Everything of the tensor-core pixel path that needs only the features, launched on the "pixel" branch stream at the very
    start of the forward, next to the positional MLPs and the whole token path: NCHW fp32 -> NHWC bf16 (``t.xh``, which the
    patch embedding on the launching stream waits for through ``t.xh_ready``), then conv3x3 -> ``t.x3h``; ``_pixel_fwd_bf16`` joins
    the branch.  (Branching off: the same launches on the current stream.)"""
    dev = x.device
    t.conv_fork = fork_branch(dev, 4)
    t.xh_ready = None
    given = getattr(t, "xh_given", None)
    with on_branch(dev, t.conv_fork, None if given is not None else x, name="pixel"):
        t.xh = xh = given if given is not None else ops.nchw_to_nhwc_bf16(x)
        if t.conv_fork is not None and given is None:
            t.xh_ready = torch.cuda.Event()
            t.xh_ready.record(torch.cuda.current_stream(dev))
        t.pixel_rest_pending = PIXEL_LATE and t.conv_fork is not None
        if not t.pixel_rest_pending:
            _pixel_prologue_rest(P, pre, t, x)


def _pixel_prologue_rest(P: Params, pre: str, t, x: torch.Tensor) -> None:
    """conv3x3 and the two x-only operands of the backward; called inside the "pixel" branch.  With ``PIXEL_LATE`` the forward calls
    it after the patch embedding has been issued (ordered after it), so that these whole-GPU kernels (a persistent 148-CTA conv,
    0.5 ms of HBM-bound copies) run under the encoder chain instead of delaying the positional MLPs and the patch embedding."""
    xh = t.xh_keep if getattr(t, "xh_keep", None) is not None else t.xh
    t.xh_saved = xh if IMPLICIT_DW3 else None            # the backward's conv3x3 weight gradient reads the features in this layout
    t.x3h = ops.conv3x3_nhwc_bf16(xh, ops.pack_conv3x3_bf16(P[pre + "conv3x3.weight"]), P[pre + "conv3x3.bias"])
    # the forward chain needs only x3h from this branch: it waits for THIS point, not for the backward operands made below
    t.x3_ready = None
    if t.conv_fork is not None:
        t.x3_ready = torch.cuda.Event()
        t.x3_ready.record(torch.cuda.current_stream(x.device))
    # B operand of the patch-embedding weight gradient (backward): a function of the features only, made from their NHWC bf16 copy
    # (half the bytes of the fp32 map) under the latency-bound encoder chain; the backward joins the branch where it reads it
    t.unf = ops.patch_unfold_nhwc_bf16(xh, 16) if UNFOLD_IN_FORWARD else None
    # channel-major copy of the conv output (B operand of the dMb GEMMs of the backward): also a function of x only
    Bq, hq, wq, Cq = t.x3h.shape
    t.x3T = ops.transpose_bf16(t.x3h.view(Bq, hq * wq, Cq)) if (UNFOLD_IN_FORWARD and not _fused_tail(hq * wq)) else None
    t.xh_keep = None


# split-K of the conv3x3 weight-gradient GEMM (9 output tiles x split CTAs, each holding an SM's whole register file for ~100 us):
# with 32 the GEMM fills every SM twice over and the chain's small kernels queue behind it
DW3_SPLIT = int(_os.environ.get("OCV_DW3_SPLIT", "32"))
IMPLICIT_DW3 = _os.environ.get("OCV_IMPLICIT_DW3", "1") == "1"   # 0: conv3x3 weight gradient through an im2col buffer + GEMM (round 1)
FUSED_TAIL = _os.environ.get("OCV_FUSED_TAIL", "1") == "1"   # 0: the round-1 path with materialised logits (kept as the general fallback)


def _fused_tail(Pn: int) -> bool:
    """The logits-free kernels (csrc/train_tail_fused_bf16.cu) walk flat 128-pixel tiles that must not straddle images."""
    return FUSED_TAIL and Pn % 128 == 0


def _pixel_fwd_bf16(P: Params, pre: str, t, x: torch.Tensor, tok: torch.Tensor, Q: int, edges=None) -> torch.Tensor:
    B, C, h, w = x.shape
    Pn, S = h * w, tok.shape[1]
    Wout = P["conv_out.0.weight"].view(-1, Q)
    K = Wout.shape[0]
    t.fused_tail = _fused_tail(Pn) and K == 256 and Q == 128
    t.pixel_joined = True
    if t.fused_tail and getattr(t, "x3_ready", None) is not None:
        cur = torch.cuda.current_stream(x.device)
        cur.wait_event(t.x3_ready)                       # conv3x3 done; the unfold behind it on that stream is joined by the backward
        t.x3h.record_stream(cur)
        t.pixel_joined = False
    else:
        join_branch(x.device, t.conv_fork, t.x3h, name="pixel")
    if t.fused_tail:
        # logits never reach HBM: Mb2 = log2e * Wout @ queries[b] (bf16) and blog2 = log2e * bout from one small kernel, then ONE
        # tcgen05 kernel over 128-pixel tiles of x3: logits in TMEM -> softmax in registers -> depth + (max, sum) per pixel
        t.Mb2, t.blog2 = ops.fold_queries_bf16(P["conv_out.0.weight"], P["conv_out.0.bias"], tok[:, 1:Q + 1, :], None)
        join_branch(x.device, t.reg_fork, t.centres, t.y_raw, edges, name="regressor")      # the kernel needs the bin centres
        t.depth, t.smax, t.ssum = ops.tail_fused_fwd_bf16(t.x3h, t.Mb2, t.blog2, t.centres)
        return t.depth.view(B, 1, h, w)
    Mb = torch.empty(B, K, E, device=x.device, dtype=torch.float32)
    ops.gemm_ex(Wout, tok[:, 1:Q + 1, :], K, E, Q, Q, E, Mb, E, batch=B, stride_a=0, stride_b=S * E, stride_c=K * E)
    t.Mb16 = ops.cast_bf16(Mb)
    t.logits = torch.empty(B, K, Pn, device=x.device, dtype=torch.float32)
    x3 = t.x3h.view(B, Pn, C)
    # one GEMM per image, each bound by its 58 MB output write: alternate images over PIXEL_QUEUES queues so that several of them are in flight
    fork = fork_branch(x.device, 32)
    for b in range(B):
        with on_branch(x.device, fork if b % PIXEL_QUEUES else None, t.Mb16, t.x3h, t.logits, name=f"pixel_q{b % PIXEL_QUEUES}"):
            ops.gemm_bf16(t.Mb16[b], x3[b], out=t.logits[b])
    for qi in range(1, min(PIXEL_QUEUES, B)):
        join_branch(x.device, fork, name=f"pixel_q{qi}")
    join_branch(x.device, t.reg_fork, t.centres, t.y_raw, edges, name="regressor")      # the tail needs the bin centres
    depth, t.tail_stats = ops.tail_logits(t.logits, P["conv_out.0.bias"], t.centres, want_stats=True)
    return depth.view(B, 1, h, w)


def _pixel_bwd_bf16_tail(P: Params, G: Params, t, gd: torch.Tensor):
    """-> (dlT [B,K,P] bf16, d loss / d centres [B,K]); accumulates the conv_out bias gradient."""
    dlT, dcent = ops.tail_logits_bwd(t.logits, P["conv_out.0.bias"], t.centres, gd, grad_bias=G["conv_out.0.bias"], beta=1.0,
                                     stats=getattr(t, "tail_stats", None))
    t.logits = t.tail_stats = None
    return dlT, dcent


def _pixel_bwd_bf16_rest(P: Params, G: Params, t, dlT, dtok: torch.Tensor, dMb=None, dx3T=None) -> None:
    """conv_out weight, query (into ``dtok[:, 1:Q+1]``), conv3x3 weight and bias gradients — from dlT (materialised-logits path) or
    from the (dMb, dx3T) the logits-free backward kernel has already produced."""
    pre, x, tok, Q = t.pre, t.x, t.tok, t.Q
    B, C, h, w = t.shape
    Pn, S = h * w, tok.shape[1]
    dev = x.device
    Wout = P["conv_out.0.weight"].view(-1, Q)
    K = Wout.shape[0]
    if dMb is None:
        # d Mb -> conv_out weight and queries: the only part the token-path backward waits for
        x3T = getattr(t, "x3T", None)                                             # [B,C,P]: made in the forward on the pixel branch
        if x3T is None:
            x3T = ops.transpose_bf16(t.x3h.view(B, Pn, C))
        else:
            x3T.record_stream(torch.cuda.current_stream(dev))
            t.x3T = None
        dMb = torch.empty(B, K, E, device=dev, dtype=torch.float32)
        sk = max(1, min(64, Pn // 1024))
        fork = fork_branch(dev, 32)
        for b in range(B):
            with on_branch(dev, fork if b % PIXEL_QUEUES else None, dlT, x3T, dMb, name=f"pixel_q{b % PIXEL_QUEUES}"):
                ops.gemm_bf16(dlT[b], x3T[b], out=dMb[b], split_k=sk)
        for qi in range(1, min(PIXEL_QUEUES, B)):
            join_branch(dev, fork, name=f"pixel_q{qi}")
        del x3T
    queries = tok[:, 1:Q + 1, :]
    ops.gemm_ex(dMb, queries, K, Q, E, E, E, G["conv_out.0.weight"], Q, op_b=1, beta=1.0, batch=B, stride_a=K * E, stride_b=S * E,
                reduce_batch=True)
    ops.gemm_ex(Wout, dMb, Q, E, K, Q, E, dtok[:, 1:Q + 1, :], E, trans_a=True, batch=B, stride_a=0, stride_b=K * E, stride_c=S * E)
    # d x3 (channel-major over the whole batch: the A operand of the weight-gradient GEMM), then conv3x3 weight + bias as one
    # GEMM over all pixels of the batch — nothing downstream needs either, so the whole chain runs on the weight-gradient
    # stream next to the token-path backward
    gw, gb = G[pre + "conv3x3.weight"].view(C, C * 9), G[pre + "conv3x3.bias"]
    Mb16 = getattr(t, "Mb16", None)

    def conv_wgrad():
        dx = dx3T
        if dx is None:
            dl = ops.transpose_bf16(dlT)                                          # [B,P,K]
            MbT16 = ops.transpose_bf16(Mb16)                                      # [B,C,K]
            dx = torch.empty(C, B * Pn, device=dev, dtype=torch.bfloat16)
            for b in range(B):
                ops.gemm_bf16(MbT16[b], dl[b], out=dx[:, b * Pn:(b + 1) * Pn])
            del dl
        xh = getattr(t, "xh_saved", None)
        if xh is not None and w % 8 == 0 and C == 128:
            # implicit GEMM over pixels: dx (channel-major) and the NHWC features are both read as stored, no im2col buffer
            ops.conv3x3_wgrad_bf16(dx, xh, gw, gb, beta=1.0)
            return
        col = ops.im2col3x3_bf16(ops.cast_bf16(x))
        skw = max(1, min(DW3_SPLIT, (B * Pn) // 8192))
        ops.gemm_bf16(dx, col[:C * 9], residual=gw, out=gw, split_k=skw)
        db = ops.gemm_bf16(dx, col[C * 9:], split_k=skw)                      # [C,8]: column 0 = sum over pixels, the rest 0
        ops.reduce_channels(db, gb, 1, 1, C, 8, beta=1.0)

    reads = [r for r in (dlT, Mb16, dx3T, x, getattr(t, "xh_saved", None)) if r is not None]
    on_side(conv_wgrad, *reads, name="conv3x3")       # its own queue: HBM-bound work the layer wgrads must not wait behind


def backward(P: Params, G: Params, t: Tape, d_depth: torch.Tensor, d_edges: Optional[torch.Tensor]) -> None:
    """Accumulate parameter gradients into ``G`` for upstream gradients ``d_depth`` [B,1,h,w] and ``d_edges`` [B,K+1]|None."""
    cfg, pre, x = t.cfg, t.pre, t.x
    B, C, h, w = t.shape
    Pn = h * w
    dev = x.device
    tok = t.tok
    S = tok.shape[1]
    Q = t.Q
    Wout = P["conv_out.0.weight"].view(-1, Q)
    K = Wout.shape[0]
    gd = d_depth.contiguous().view(B, Pn)
    if t.pixel == "bf16":
        dMb = dx3T = dlT = None
        if getattr(t, "fused_tail", False):
            # flash-style: the logits tile is recomputed on the tensor cores, d logits goes straight into the dMb / dx3^T MMAs
            dx3T, dMb, dcent = ops.tail_fused_bwd_bf16(t.x3h, t.Mb2, t.blog2, t.centres, t.depth, t.smax, t.ssum, gd,
                                                       grad_bias=G["conv_out.0.bias"], beta=1.0)
        else:
            dlT, dcent = _pixel_bwd_bf16_tail(P, G, t, gd)
        dtok = torch.zeros(B, S, E, device=dev, dtype=torch.float32)
        de = d_edges.contiguous() if d_edges is not None else None
        fork = fork_branch(dev, 16)      # bins + regressor backward (-> dtok[:, 0]) next to the dMb chain (-> dtok[:, 1:Q+1])
        with on_branch(dev, fork, dcent, de, dtok, name="regressor"):
            dy_raw = ops.bins_bwd(t.y_raw, de, dcent, cfg["min_depth"], cfg["max_depth"])
            mlp_bwd(P, G, pre + "regressor.", (0, 2, 4), t.sv_reg, dy_raw, dx_out=dtok[:, 0, :])
        _pixel_bwd_bf16_rest(P, G, t, dlT, dtok, dMb=dMb, dx3T=dx3T)
        join_branch(dev, fork, name="regressor")
        del dlT, dx3T
        return _token_bwd(P, G, t, dtok)
    ram = t.ram.view(B, Q, Pn)
    # ---- tail: depth = sum_k softmax_k(Wout ram + bout) c_k ----------------------------------------------------------
    logits = torch.empty(B, K, Pn, device=dev, dtype=torch.float32)
    ops.gemm_ex(Wout, ram, K, Pn, Q, Q, Pn, logits, Pn, batch=B, stride_a=0, stride_b=Q * Pn, stride_c=K * Pn)
    dl, pg = ops.tail_bwd(logits, t.centres, gd, bias=P["conv_out.0.bias"])
    dcent = torch.empty(B, K, device=dev, dtype=torch.float32)
    ops.reduce_channels(pg, dcent, B, 1, K, Pn)
    del pg
    ops.reduce_channels(dl, G["conv_out.0.bias"], 1, B, K, Pn, beta=1.0)
    ops.gemm_ex(dl, ram, K, Q, Pn, Pn, Pn, G["conv_out.0.weight"], Q, op_b=1, beta=1.0, batch=B, stride_a=K * Pn, stride_b=Q * Pn,
                reduce_batch=True, split_k=_split(K, Q, B, Pn))
    dram = torch.empty(B, Q, Pn, device=dev, dtype=torch.float32)
    ops.gemm_ex(Wout, dl, Q, Pn, K, Q, Pn, dram, Pn, trans_a=True, batch=B, stride_a=0, stride_b=K * Pn, stride_c=Q * Pn)
    del dl, logits
    # ---- bins + regressor, range attention, conv3x3, tokens --------------------------------------------------------------
    dy_raw = ops.bins_bwd(t.y_raw, d_edges.contiguous() if d_edges is not None else None, dcent, cfg["min_depth"], cfg["max_depth"])
    _ram_bwd(P, G, t, dram, dy_raw)


def _ram_bwd(P: Params, G: Params, t: Tape, dram: torch.Tensor, dy_raw: torch.Tensor) -> None:
    """fp32 path from (d loss / d range_attention_maps [B,Q,P], d loss / d regressor output [B,K]) down to every parameter."""
    pre, x, tok, Q = t.pre, t.x, t.tok, t.Q
    B, C, h, w = t.shape
    Pn, S = h * w, tok.shape[1]
    dev = x.device
    dtok = torch.zeros(B, S, E, device=dev, dtype=torch.float32)
    mlp_bwd(P, G, pre + "regressor.", (0, 2, 4), t.sv_reg, dy_raw, dx_out=dtok[:, 0, :])
    # ---- range attention: ram[b] = queries[b] @ x3[b] ------------------------------------------------------------------
    x3 = t.x3.view(B, C, Pn)
    queries = tok[:, 1:Q + 1, :]
    ops.gemm_ex(dram, x3, Q, E, Pn, Pn, Pn, dtok[:, 1:Q + 1, :], E, op_b=1, batch=B, stride_a=Q * Pn, stride_b=C * Pn,
                stride_c=S * E, split_k=_split(Q, E, B, Pn))
    dx3 = torch.empty(B, C, Pn, device=dev, dtype=torch.float32)
    ops.gemm_ex(queries, dram, E, Pn, Q, E, Pn, dx3, Pn, trans_a=True, batch=B, stride_a=S * E, stride_b=Q * Pn, stride_c=C * Pn)
    del dram
    ops.gemm_ex(dx3, x, C, C * 9, Pn, Pn, 0, G[pre + "conv3x3.weight"], C * 9, op_b=3, beta=1.0, batch=B, stride_a=C * Pn,
                stride_b=C * Pn, reduce_batch=True, split_k=_split(C, C * 9, B, Pn), geom=(C, h, w, 0))
    ops.reduce_channels(dx3, G[pre + "conv3x3.bias"], 1, B, C, Pn, beta=1.0)
    del dx3
    _token_bwd(P, G, t, dtok)


def backward_module(P: Params, G: Params, t: Tape, d_widths: Optional[torch.Tensor], d_ram: Optional[torch.Tensor]) -> None:
    """Backward of the ``surface="module"`` forward: upstream gradients of ``bin_widths_normed`` [B,K] and of
    ``range_attention_maps`` [B,Q,h,w] (either may be None)."""
    B, C, h, w = t.shape
    dev = t.x.device
    K = t.y_raw.shape[1]
    dw = d_widths.contiguous().float() if d_widths is not None else torch.zeros(B, K, device=dev, dtype=torch.float32)
    dy_raw = ops.bins_bwd(t.y_raw, None, None, 0.0, 1.0, grad_widths=dw)
    dram = (d_ram.contiguous().float() if d_ram is not None
            else torch.zeros(B, t.Q, h, w, device=dev, dtype=torch.float32)).view(B, t.Q, h * w)
    _ram_bwd(P, G, t, dram, dy_raw)


def _table_bwd(g_table: torch.Tensor, strategy: str, pos_inputs, d_obj: torch.Tensor, d_img: torch.Tensor, h: int, w: int,
               factor: float = 2.0) -> None:
    """Gradient of the positional table of the grid_random* strategies (``GridRandomPositionalEmbeddings``,
    ``modules/ObjCAViT.py:50-147``): the adjoint of ``engine.positional_embeddings`` — same geometry, same quirks (F13) — for the
    object rows ``d_obj`` [total,E] and the batch-summed image rows ``d_img`` [S,E]."""
    boxes, coords, patch_size = pos_inputs
    gh, gw = grid_geometry(h, w, patch_size)
    fix = ops.table_grad_buffer(gh, gw, E, d_obj.device)
    if strategy == "grid_random":
        ops.grid_sample_pos_bwd(d_obj, boxes[:, 0:2].contiguous(), fix, gh, gw, 1, h * factor, w * factor)
        ops.grid_sample_pos_bwd(d_img, coords[:, 0:2].contiguous(), fix, gh, gw, 2, float(gh), float(gw))
    else:
        ops.ps_roi_align_pos_bwd(d_obj, boxes, fix, gh, gw, 1.0 / (patch_size * factor), True)
        ops.ps_roi_align_pos_bwd(d_img, coords, fix, gh, gw, 1.0 / patch_size, True)
    ops.table_grad_commit(fix, g_table[0:gh * gw], beta=1.0)


def _token_bwd(P: Params, G: Params, t: Tape, dtok: torch.Tensor) -> None:
    """Backward of everything that produced the tokens, from d loss / d tokens [B,S,E]."""
    cfg, pre, x = t.cfg, t.pre, t.x
    B, C, h, w = t.shape
    dev = x.device
    S = t.tok.shape[1]
    if cfg["kind"] == "adabins":
        pt = pre + "patch_transformer."
        dtok0 = encoder_bwd(P, G, pt + "transformer_encoder.", t.sv_enc, dtok, t.P16)
        wkey, bkey = pt + "embedding_convPxP.weight", pt + "embedding_convPxP.bias"
        d_pos_rows = G[pt + "positional_encodings"][0:S]
        ops.reduce_channels(dtok0.contiguous(), d_pos_rows, 1, B, S * E, 1, beta=1.0)
    else:
        d_obj = None
        if cfg.get("use_2_saca"):
            dtok, d_obj = saca_bwd(P, G, pre + "saca_2.", t.sv_saca2, dtok, None, t.P16)
        dtok0, d_obj_tok = saca_bwd(P, G, pre + "saca_1.", t.sv_saca1, dtok, d_obj, t.P16)
        wkey, bkey = pre + "image_embedding_convPxP.weight", pre + "image_embedding_convPxP.bias"
        pe = pre + "positional_encoder."
        learned = cfg["strategy"] in ("learned", "learned_bbox_wh")
        n_obj = t.n_obj_rows if learned else 0
        d_pos = torch.empty(n_obj + S, E, device=dev, dtype=torch.float32)       # [object rows | patch rows], as the forward chain
        d_img_pos = d_pos[n_obj:]
        ops.reduce_channels(dtok0.contiguous(), d_img_pos, 1, B, S * E, 1)
        _wgrad(d_obj_tok, t.feats, G[pre + "obj_embedding_layer.weight"], G[pre + "obj_embedding_layer.bias"])
        if learned:
            d_pos[:n_obj].copy_(d_obj_tok.reshape(n_obj, E))
            mlp_bwd(P, G, pe, (0, 2, 4, 6, 8), t.sv_mlp_pos, d_pos)
        else:
            _table_bwd(G[pe + "positional_encodings"], cfg["strategy"], t.pos_inputs, d_obj_tok.contiguous(), d_img_pos, h, w)
    dt2 = dtok0.contiguous().view(B * S, E)
    Pp = 16
    if t.pixel == "bf16":
        unf = getattr(t, "unf", None)        # made on the pixel branch in the forward
        if not getattr(t, "pixel_joined", True):
            join_branch(dev, t.conv_fork, name="pixel")
            t.pixel_joined = True
        if unf is not None:
            unf.record_stream(torch.cuda.current_stream(dev))
        ops.patch_embed_wgrad_bf16(dt2, x, G[wkey], Pp, unf=unf)
    else:
        ops.gemm_ex(dt2, x, E, C * Pp * Pp, B * S, E, 0, G[wkey], C * Pp * Pp, trans_a=True, op_b=2, beta=1.0,
                    split_k=_split(E, C * Pp * Pp, 1, B * S), geom=(C, h, w, Pp))
    ops.reduce_channels(dt2, G[bkey], 1, B * S, E, 1, beta=1.0)
    join_side(dev)
