"""Thin torch-tensor wrappers over the C ABI (``include/objcavit_b200.h``).

PyTorch is used for device memory (the caching allocator), streams and dtypes only: every wrapper
validates its tensors, allocates outputs/workspaces and passes raw ``data_ptr()``s plus the current
stream to the kernel library.  No wrapper computes anything with torch ops.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib

ACT_NONE, ACT_RELU, ACT_LEAKY_RELU = 0, 1, 2
NORM_CODES = {"linear": 0, "softmax": 1}  # anything else -> sigmoid (miniViT.py:40-41)

# number of kernel launches issued through this module (bench.py reports it as gpu_launches)
launch_count = 0


def _stream() -> int:
    """Raw handle of torch's current stream on the current device.  (torch.cuda.current_stream() builds a Python Stream object
    per call: at ~540 launches per training step that alone was 1.5 ms of host time per step.)"""
    return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _chk(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise _lib.OcvError(f"{name} must be a CUDA tensor: objcavit_b200 has no CPU path")
    if t.dtype != dtype:
        raise _lib.OcvError(f"{name} must be {dtype}, got {t.dtype}")


# bench.py sets `timed_kernel` to an entry-point name to get CUDA-event durations of that kernel on
# the launching stream (list of (start, stop) events in `timed_events`).
timed_kernel = None
timed_events = []
# bench.py --workload train sets `profile_events` to a list to get (entry point, start, stop) CUDA events for EVERY call
profile_events = None


profile_tag = ""          # optional suffix of the next profiled entries (bench / probes: shapes of the generic GEMM calls)


_fn_cache = {}


def _call(name: str, *args, launches: int = 1):
    global launch_count
    launch_count += launches
    if profile_events is None and timed_kernel is None:      # the hot path: cached ctypes function, status checked inline
        fn = _fn_cache.get(name)
        if fn is None:
            fn = _fn_cache[name] = getattr(_lib.lib(), name)
        status = fn(*args)
        if status != 0:
            _lib.check(status, name)
        return
    if profile_events is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(getattr(_lib.lib(), name)(*args), name)
        e1.record()
        profile_events.append((name + profile_tag if profile_tag else name, e0, e1))
        return
    if timed_kernel == name:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(getattr(_lib.lib(), name)(*args), name)
        e1.record()
        timed_events.append((e0, e1))
        return
    _lib.check(getattr(_lib.lib(), name)(*args), name)


def _row_major_2d(t: torch.Tensor, name: str):
    """(rows, cols, ld) of a 2-D view whose last dim is contiguous."""
    if t.dim() != 2 or t.stride(1) != 1:
        raise _lib.OcvError(f"{name}: expected a 2-D tensor with unit inner stride, got {tuple(t.shape)} {t.stride()}")
    return t.shape[0], t.shape[1], t.stride(0)


def linear(a: torch.Tensor, w: torch.Tensor, bias=None, act: int = ACT_NONE, residual=None, out=None):
    """out[M,N] = act(a[M,K] @ w[N,K]^T + bias) + residual   (fp32)."""
    _chk(a, torch.float32, "a"); _chk(w, torch.float32, "w")
    M, K, lda = _row_major_2d(a, "a")
    N, K2, ldb = _row_major_2d(w, "w")
    assert K == K2, (a.shape, w.shape)
    if out is None:
        out = torch.empty(M, N, device=a.device, dtype=torch.float32)
    _, _, ldc = _row_major_2d(out, "out")
    if residual is not None:
        assert residual.shape == out.shape and residual.stride() == out.stride()
    if residual is None and act == ACT_NONE and K >= 512 and ((M + 63) // 64) * ((N + 63) // 64) <= 64:
        split = max(2, min(K // 128, (192 * 64 * 64) // max(M * N, 1), 32))      # long K, few tiles: split-K
        ws = _workspace(split * M * N * 4, a.device, "linear_splitk")
        _call("ocv_linear_splitk_f32", a.data_ptr(), w.data_ptr(), _ptr(bias), out.data_ptr(), M, N, K, lda, ldb, ldc, split,
              ws.data_ptr(), _stream(), launches=2)
        return out
    _call("ocv_linear_f32", a.data_ptr(), w.data_ptr(), _ptr(bias), _ptr(residual), out.data_ptr(),
          M, N, K, lda, ldb, ldc, 1, act, 1, 0, 0, 0, _stream())
    return out


def linear_rows(a: torch.Tensor, w: torch.Tensor, bias=None, leaky: bool = False):
    """out[R,N] = [leaky_relu](a[R,K] @ w[N,K]^T + bias) for few rows (``ocv_linear_rows_f32``; K <= 512, K % 4 == 0)."""
    _chk(a, torch.float32, "a"); _chk(w, torch.float32, "w")
    R, K, lda = _row_major_2d(a, "a")
    N, K2, ldb = _row_major_2d(w, "w")
    assert K == K2 and ldb == K
    out = torch.empty(R, N, device=a.device, dtype=torch.float32)
    _call("ocv_linear_rows_f32", a.data_ptr(), lda, R, w.data_ptr(), _ptr(bias), K, N, int(leaky), out.data_ptr(), N, _stream())
    return out


def linear_rows_ok(a: torch.Tensor, w: torch.Tensor) -> bool:
    K = a.shape[-1]
    return (a.shape[0] <= 1024 and K <= 512 and K % 4 == 0 and a.stride(-1) == 1 and a.stride(0) % 4 == 0 and a.data_ptr() % 16 == 0
            and w.is_contiguous() and w.data_ptr() % 16 == 0)


def batched_matmul_nn(a: torch.Tensor, b: torch.Tensor, out=None):
    """out[n] = a[n] ([M,K], any batch/row stride) @ b[n] ([K,N] contiguous)   (fp32)."""
    _chk(a, torch.float32, "a"); _chk(b, torch.float32, "b")
    Bn, M, K = a.shape
    _, K2, N = b.shape
    assert K == K2 and a.stride(2) == 1 and b.is_contiguous()
    if out is None:
        out = torch.empty(Bn, M, N, device=a.device, dtype=torch.float32)
    _call("ocv_linear_f32", a.data_ptr(), b.data_ptr(), None, None, out.data_ptr(), M, N, K,
          a.stride(1), N, N, 0, ACT_NONE, Bn, a.stride(0), K * N, M * N, _stream())
    return out


def patch_embed(x: torch.Tensor, w: torch.Tensor, bias, pos, P: int = 16):
    """tokens [B,S,E] from NCHW fp32 features (patch conv + bias + positional rows)."""
    _chk(x, torch.float32, "x"); _chk(w, torch.float32, "w")
    assert x.is_contiguous() and w.is_contiguous()
    B, Cc, h, wd = x.shape
    E = w.shape[0]
    S = (h // P) * (wd // P)
    if pos is not None:
        assert pos.is_contiguous() and pos.shape[0] >= S and pos.shape[1] == E
    ws_bytes = _lib.lib().ocv_patch_embed_f32_workspace(B, Cc, h, wd, P, E)
    ws = torch.empty(max(ws_bytes, 4) // 4, device=x.device, dtype=torch.float32)
    out = torch.empty(B, S, E, device=x.device, dtype=torch.float32)
    _call("ocv_patch_embed_f32", x.data_ptr(), w.data_ptr(), _ptr(bias), _ptr(pos), out.data_ptr(),
          B, Cc, h, wd, P, E, ws.data_ptr(), _stream(), launches=2)
    return out


def add_layernorm(x: torch.Tensor, r, gamma, beta, zero_rows=None, out=None):
    _chk(x, torch.float32, "x")
    assert x.is_contiguous() and (r is None or r.is_contiguous())
    E = x.shape[-1]
    rows = x.numel() // E
    if out is None:
        out = torch.empty_like(x)
    _call("ocv_add_layernorm_f32", x.data_ptr(), _ptr(r), gamma.data_ptr(), beta.data_ptr(), out.data_ptr(),
          rows, E, _ptr(zero_rows), _stream())
    return out


def mha_core(q, k, v, key_padding_mask=None, heads: int = 4, drop=None):
    """q [B,Sq,E], k/v [B,Sk,E] views with unit channel stride -> [B,Sq,E].  ``drop`` = (p, seed): dropout on the attention
    probabilities (training mode)."""
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _chk(t, torch.float32, n)
        assert t.dim() == 3 and t.stride(2) == 1
    B, Sq, E = q.shape
    Sk = k.shape[1]
    out = torch.empty(B, Sq, E, device=q.device, dtype=torch.float32)
    if key_padding_mask is not None:
        _chk(key_padding_mask, torch.uint8, "key_padding_mask")
        assert key_padding_mask.is_contiguous() and key_padding_mask.shape == (B, Sk)
    if drop is not None and drop[0] > 0:
        _call("ocv_mha_core_dropout_f32", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), _ptr(key_padding_mask),
              B, Sq, Sk, heads, E // heads, q.stride(0), q.stride(1), k.stride(0), k.stride(1), v.stride(0),
              v.stride(1), out.stride(0), out.stride(1), float(drop[0]), int(drop[1]), _stream())
        return out
    _call("ocv_mha_core_f32", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), _ptr(key_padding_mask),
          B, Sq, Sk, heads, E // heads, q.stride(0), q.stride(1), k.stride(0), k.stride(1), v.stride(0),
          v.stride(1), out.stride(0), out.stride(1), _stream())
    return out


def pad_objects(src: torch.Tensor, offsets: torch.Tensor, B: int, L: int, n_max: int, pad_value: float,
                left_pad: bool, want_mask: bool = True, mask_offsets=None):
    _chk(src, torch.float32, "src"); _chk(offsets, torch.int32, "offsets")
    assert src.is_contiguous() and offsets.numel() == B + 1
    E = src.shape[1]
    dst = torch.empty(B, L, E, device=src.device, dtype=torch.float32)
    mask = torch.empty(B, L, device=src.device, dtype=torch.uint8) if want_mask else None
    _call("ocv_pad_objects_f32", src.data_ptr(), offsets.data_ptr(), _ptr(mask_offsets), dst.data_ptr(), _ptr(mask),
          B, L, n_max, E,
          float(pad_value), int(left_pad), _stream())
    return dst, mask


def patch_coords(gh: int, gw: int, P: int, device) -> torch.Tensor:
    out = torch.empty(gh * gw, 4, device=device, dtype=torch.float32)
    _call("ocv_patch_coords_f32", out.data_ptr(), gh, gw, P, _stream())
    return out


def grid_sample_pos(table: torch.Tensor, pts: torch.Tensor, gh: int, gw: int, norm_mode: int = 0,
                    div0: float = 1.0, div1: float = 1.0, want_idx: bool = False):
    _chk(table, torch.float32, "table"); _chk(pts, torch.float32, "pts")
    assert table.is_contiguous() and pts.is_contiguous() and pts.shape[1] == 2
    N, E = pts.shape[0], table.shape[1]
    assert table.shape[0] >= gh * gw
    out = torch.empty(N, E, device=table.device, dtype=torch.float32)
    idx = torch.empty(N, 4, device=table.device, dtype=torch.int32) if want_idx else None
    _call("ocv_grid_sample_pos_f32", table.data_ptr(), pts.data_ptr(), out.data_ptr(), _ptr(idx), N, gh, gw, E,
          norm_mode, float(div0), float(div1), _stream())
    return (out, idx) if want_idx else out


def ps_roi_align_pos(table: torch.Tensor, boxes: torch.Tensor, gh: int, gw: int, spatial_scale: float,
                     boxes_are_xywh: bool, want_plan: bool = False):
    _chk(table, torch.float32, "table"); _chk(boxes, torch.float32, "boxes")
    assert table.is_contiguous() and boxes.is_contiguous() and boxes.shape[1] == 4
    N, E = boxes.shape[0], table.shape[1]
    assert table.shape[0] >= gh * gw
    out = torch.empty(N, E, device=table.device, dtype=torch.float32)
    plan = torch.empty(N, 2, device=table.device, dtype=torch.int32) if want_plan else None
    _call("ocv_ps_roi_align_pos_f32", table.data_ptr(), boxes.data_ptr(), out.data_ptr(), _ptr(plan), N, gh, gw, E,
          float(spatial_scale), int(boxes_are_xywh), _stream())
    return (out, plan) if want_plan else out


def bins(y_raw: torch.Tensor, norm: str, min_depth: Optional[float] = None, max_depth: Optional[float] = None):
    """-> (widths_normed [B,K], edges [B,K+1] | None, centres [B,K] | None)."""
    _chk(y_raw, torch.float32, "y_raw")
    assert y_raw.is_contiguous()
    B, K = y_raw.shape
    widths = torch.empty_like(y_raw)
    code = NORM_CODES.get(norm, 2)
    if min_depth is None:
        _call("ocv_bins_f32", y_raw.data_ptr(), widths.data_ptr(), None, None, B, K, code, 0.0, 1.0, _stream())
        return widths, None, None
    edges = torch.empty(B, K + 1, device=y_raw.device, dtype=torch.float32)
    centres = torch.empty(B, K, device=y_raw.device, dtype=torch.float32)
    _call("ocv_bins_f32", y_raw.data_ptr(), widths.data_ptr(), edges.data_ptr(), centres.data_ptr(), B, K, code,
          float(min_depth), float(max_depth), _stream())
    return widths, edges, centres


def edges_from_widths(widths: torch.Tensor, min_depth: float, max_depth: float):
    """edges/centres from already-normalised widths (the AdaBins/GraphBins tail, AdaBins.py:79-85)."""
    _chk(widths, torch.float32, "widths")
    assert widths.is_contiguous()
    B, K = widths.shape
    edges = torch.empty(B, K + 1, device=widths.device, dtype=torch.float32)
    centres = torch.empty(B, K, device=widths.device, dtype=torch.float32)
    _call("ocv_bins_f32", widths.data_ptr(), None, edges.data_ptr(), centres.data_ptr(), B, K, 3,
          float(min_depth), float(max_depth), _stream())
    return edges, centres


def conv3x3(x: torch.Tensor, w: torch.Tensor, bias):
    _chk(x, torch.float32, "x"); _chk(w, torch.float32, "w")
    assert x.is_contiguous() and w.is_contiguous()
    B, Cc, h, wd = x.shape
    out = torch.empty_like(x)
    _call("ocv_conv3x3_f32", x.data_ptr(), w.data_ptr(), _ptr(bias), out.data_ptr(), B, Cc, h, wd, _stream())
    return out


def tail(ram: torch.Tensor, w_out: torch.Tensor, b_out, centres: torch.Tensor):
    """depth [B,1,h,w] = sum_k softmax_k(conv1x1(ram)) * centres."""
    _chk(ram, torch.float32, "ram")
    assert ram.is_contiguous() and centres.is_contiguous()
    B, Q, h, wd = ram.shape
    K = w_out.shape[0]
    w2 = w_out.view(K, Q)
    assert w2.is_contiguous()
    depth = torch.empty(B, 1, h, wd, device=ram.device, dtype=torch.float32)
    _call("ocv_tail_f32", ram.data_ptr(), w2.data_ptr(), _ptr(b_out), centres.data_ptr(), depth.data_ptr(),
          B, Q, K, h * wd, _stream())
    return depth


# ---- bf16 tensor-core path --------------------------------------------------------------------------

def nchw_to_nhwc_bf16(x: torch.Tensor, out: Optional[torch.Tensor] = None):
    assert x.is_cuda and x.is_contiguous() and x.dtype in (torch.float32, torch.bfloat16)
    B, Cc, h, wd = x.shape
    if out is None:
        out = torch.empty(B, h, wd, Cc, device=x.device, dtype=torch.bfloat16)
    else:
        assert out.shape == (B, h, wd, Cc) and out.dtype == torch.bfloat16 and out.is_contiguous() and out.device == x.device
    _call("ocv_nchw_to_nhwc_bf16", x.data_ptr(), out.data_ptr(), B, Cc, h, wd, int(x.dtype == torch.bfloat16), _stream())
    return out


def gemm_bf16(a: torch.Tensor, w: torch.Tensor, bias=None, act: int = ACT_NONE, residual=None,
              out_dtype=torch.float32, split_k: int = 1, out=None):
    """out[M,N] = act(a[M,K] @ w[N,K]^T + bias) + residual; a, w bf16; tcgen05 tensor cores."""
    _chk(a, torch.bfloat16, "a"); _chk(w, torch.bfloat16, "w")
    M, K, lda = _row_major_2d(a, "a")
    N, K2, ldb = _row_major_2d(w, "w")
    assert K == K2
    if out is None:
        out = torch.empty(M, N, device=a.device, dtype=out_dtype)
    _, _, ldc = _row_major_2d(out, "out")
    ws = None
    if split_k > 1:
        ws = torch.empty(split_k * M * N, device=a.device, dtype=torch.float32)
    if residual is not None:
        _chk(residual, torch.float32, "residual")
        assert residual.stride(0) == ldc
    _call("ocv_gemm_bf16", a.data_ptr(), w.data_ptr(), _ptr(bias), _ptr(residual), out.data_ptr(), M, N, K,
          lda, ldb, ldc, act, int(out.dtype == torch.bfloat16), split_k, _ptr(ws), _stream(),
          launches=2 if split_k > 1 else 1)
    return out


def pack_conv3x3_bf16(w: torch.Tensor):
    _chk(w, torch.float32, "w")
    assert w.is_contiguous() and tuple(w.shape) == (128, 128, 3, 3)
    out = torch.empty(9, 128, 128, device=w.device, dtype=torch.bfloat16)
    _call("ocv_pack_conv3x3_bf16", w.data_ptr(), out.data_ptr(), _stream())
    return out


LOG2E = 1.4426950408889634


def fold_queries_bf16(w_out: torch.Tensor, b_out, queries: torch.Tensor, b3, logit_scale: float = LOG2E):
    """Mb [B,K,E] bf16, blog [B,K] fp32 from conv_out weights and the per-image queries view [B,Q,E]."""
    _chk(queries, torch.float32, "queries")
    B, Q, E = queries.shape
    assert queries.stride(2) == 1
    K = w_out.shape[0]
    w2 = w_out.view(K, Q)
    Mb = torch.empty(B, K, E, device=queries.device, dtype=torch.bfloat16)
    blog = torch.empty(B, K, device=queries.device, dtype=torch.float32)
    _call("ocv_fold_queries_bf16", w2.data_ptr(), _ptr(b_out), queries.data_ptr(), queries.stride(0), queries.stride(1),
          _ptr(b3), float(logit_scale), Mb.data_ptr(), blog.data_ptr(), B, K, Q, E, _stream())
    return Mb, blog


def pixel_fused_bf16(x_nhwc: torch.Tensor, w3p: torch.Tensor, Mb: torch.Tensor, blog: torch.Tensor,
                     centres: torch.Tensor, max_ctas: int = 0, want_argmax: bool = False):
    """depth [B,1,h,w] fp32 from NHWC bf16 features: conv3x3 -> logits -> softmax -> sum p*centre, one kernel."""
    _chk(x_nhwc, torch.bfloat16, "x_nhwc"); _chk(w3p, torch.bfloat16, "w3p"); _chk(Mb, torch.bfloat16, "Mb")
    _chk(blog, torch.float32, "blog"); _chk(centres, torch.float32, "centres")
    B, h, w, Cc = x_nhwc.shape
    assert Cc == 128 and x_nhwc.is_contiguous() and w3p.is_contiguous() and Mb.is_contiguous()
    assert tuple(Mb.shape) == (B, 256, 128) and tuple(blog.shape) == (B, 256) and tuple(centres.shape) == (B, 256)
    assert blog.is_contiguous() and centres.is_contiguous()
    depth = torch.empty(B, 1, h, w, device=x_nhwc.device, dtype=torch.float32)
    if want_argmax:
        amax = torch.empty(B, 1, h, w, device=x_nhwc.device, dtype=torch.uint8)
        _call("ocv_pixel_fused_bf16_argmax", x_nhwc.data_ptr(), w3p.data_ptr(), Mb.data_ptr(), blog.data_ptr(),
              centres.data_ptr(), depth.data_ptr(), amax.data_ptr(), B, h, w, _stream())
        return depth, amax
    _call("ocv_pixel_fused_bf16", x_nhwc.data_ptr(), w3p.data_ptr(), Mb.data_ptr(), blog.data_ptr(),
          centres.data_ptr(), depth.data_ptr(), B, h, w, int(max_ctas), _stream())
    return depth


# ---- bf16 token path --------------------------------------------------------------------------------

def cast_bf16(x: torch.Tensor, out=None) -> torch.Tensor:
    _chk(x, torch.float32, "x")
    assert x.is_contiguous()
    if out is None:
        out = torch.empty(x.shape, device=x.device, dtype=torch.bfloat16)
    else:
        _chk(out, torch.bfloat16, "out")
        assert out.is_contiguous() and out.numel() == x.numel()
    _call("ocv_cast_f32_to_bf16", x.data_ptr(), out.data_ptr(), x.numel(), _stream())
    return out


def gemm_bf16_ex(a: torch.Tensor, w: torch.Tensor, bias=None, act: int = ACT_NONE, residual=None, ln=None,
                 zero_rows=None, want32: bool = False, want16: bool = True):
    """(out32 | None, out16 | None) = [LayerNorm](act(a @ w^T + bias) + residual); a, w bf16, residual fp32."""
    _chk(a, torch.bfloat16, "a"); _chk(w, torch.bfloat16, "w")
    M, K, lda = _row_major_2d(a, "a")
    N, K2, ldb = _row_major_2d(w, "w")
    assert K == K2
    o32 = torch.empty(M, N, device=a.device, dtype=torch.float32) if want32 else None
    o16 = torch.empty(M, N, device=a.device, dtype=torch.bfloat16) if want16 else None
    ldr = 0
    if residual is not None:
        _chk(residual, torch.float32, "residual")
        _, _, ldr = _row_major_2d(residual, "residual")
    g, be = (ln if ln is not None else (None, None))
    _call("ocv_gemm_bf16_ex", a.data_ptr(), w.data_ptr(), _ptr(bias), _ptr(residual), ldr, _ptr(g), _ptr(be),
          _ptr(zero_rows), _ptr(o32), N, _ptr(o16), N, M, N, K, lda, ldb, act, _stream())
    return o32, o16


def mha_core_bf16(q, k, v, key_padding_mask=None, heads: int = 4):
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _chk(t, torch.bfloat16, n)
        assert t.dim() == 3 and t.stride(2) == 1
    B, Sq, E = q.shape
    Sk = k.shape[1]
    out = torch.empty(B, Sq, E, device=q.device, dtype=torch.bfloat16)
    if key_padding_mask is not None:
        _chk(key_padding_mask, torch.uint8, "key_padding_mask")
        assert key_padding_mask.is_contiguous() and key_padding_mask.shape == (B, Sk)
    _call("ocv_mha_core_bf16", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), _ptr(key_padding_mask),
          B, Sq, Sk, heads, E // heads, q.stride(0), q.stride(1), k.stride(0), k.stride(1), v.stride(0),
          v.stride(1), out.stride(0), out.stride(1), _stream())
    return out


def mha_core_tc_bf16(q, k, v, key_padding_mask=None, heads: int = 4, want32: bool = False, want_stats: bool = False, drop=None):
    """The tcgen05 attention core directly (``ocv_mha_core_tc_bf16``): -> (out bf16 [B,Sq,E], out fp32 | None, stats [B,H,Sq,2] | None).
    ``drop`` = (p, seed): dropout of the attention probabilities (training)."""
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _chk(t, torch.bfloat16, n)
        assert t.dim() == 3 and t.stride(2) == 1
    B, Sq, E = q.shape
    Sk = k.shape[1]
    out = torch.empty(B, Sq, E, device=q.device, dtype=torch.bfloat16)
    out32 = torch.empty(B, Sq, E, device=q.device, dtype=torch.float32) if want32 else None
    stats = torch.empty(B, heads, Sq, 2, device=q.device, dtype=torch.float32) if want_stats else None
    if key_padding_mask is not None:
        _chk(key_padding_mask, torch.uint8, "key_padding_mask")
        assert key_padding_mask.is_contiguous() and key_padding_mask.shape == (B, Sk)
    p_drop, seed = drop if drop is not None else (0.0, 0)
    _call("ocv_mha_core_tc_bf16", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), _ptr(out32), _ptr(stats),
          _ptr(key_padding_mask), B, Sq, Sk, heads, E // heads, q.stride(0), q.stride(1), k.stride(0), k.stride(1), v.stride(0),
          v.stride(1), out.stride(0), out.stride(1), float(p_drop), int(seed), _stream())
    return out, out32, stats


def mha_core_bwd_tc_bf16(q, k, v, out16, dout16, stats, key_padding_mask=None, heads: int = 4, drop=None):
    """Backward of the tcgen05 attention core (``ocv_mha_core_bwd_tc_bf16``): q / k / v / out16 / dout16 bf16 [B,S,E] views,
    ``stats`` from the forward.  -> (dq, dk, dv) fp32 [B,S,E] each and their bf16 copies (dq16, dk16, dv16)."""
    for t, n in ((q, "q"), (k, "k"), (v, "v"), (out16, "out16"), (dout16, "dout16")):
        _chk(t, torch.bfloat16, n)
        assert t.dim() == 3 and t.stride(2) == 1
    _chk(stats, torch.float32, "stats")
    B, Sq, E = q.shape
    Sk = k.shape[1]
    assert k.stride() == v.stride() and out16.stride() == dout16.stride()
    g32 = torch.empty(3, B, max(Sq, Sk), E, device=q.device, dtype=torch.float32)
    g16 = torch.empty(3, B, max(Sq, Sk), E, device=q.device, dtype=torch.bfloat16)
    p_drop, seed = drop if drop is not None else (0.0, 0)
    _call("ocv_mha_core_bwd_tc_bf16", q.data_ptr(), k.data_ptr(), v.data_ptr(), out16.data_ptr(), dout16.data_ptr(), stats.data_ptr(),
          _ptr(key_padding_mask), g32[0].data_ptr(), g32[1].data_ptr(), g32[2].data_ptr(), g16[0].data_ptr(), g16[1].data_ptr(),
          g16[2].data_ptr(), B, Sq, Sk, heads, E // heads, q.stride(0), q.stride(1), k.stride(0), k.stride(1), out16.stride(0),
          out16.stride(1), g32.stride(1), g32.stride(2), float(p_drop), int(seed), _stream())
    return (g32[0, :, :Sq], g32[1, :, :Sk], g32[2, :, :Sk]), (g16[0, :, :Sq], g16[1, :, :Sk], g16[2, :, :Sk])


def pack_patch_weights_bf16(w: torch.Tensor) -> torch.Tensor:
    _chk(w, torch.float32, "w")
    assert w.is_contiguous() and w.dim() == 4 and w.shape[2] == w.shape[3]
    E, Cc, P, _ = w.shape
    out = torch.empty(E, P * P * Cc, device=w.device, dtype=torch.bfloat16)
    _call("ocv_pack_patch_weights_bf16", w.data_ptr(), out.data_ptr(), E, Cc, P, _stream())
    return out


def patch_embed_bf16(x_nhwc: torch.Tensor, wp: torch.Tensor, bias, pos, P: int = 16, want32=True, want16=True):
    """tokens from NHWC bf16 features -> (tok32 [B,S,E] | None, tok16 | None)."""
    _chk(x_nhwc, torch.bfloat16, "x_nhwc"); _chk(wp, torch.bfloat16, "wp")
    assert x_nhwc.is_contiguous() and wp.is_contiguous()
    B, h, w, Cc = x_nhwc.shape
    E = wp.shape[0]
    S = (h // P) * (w // P)
    if pos is not None:
        assert pos.is_contiguous() and pos.shape[0] >= S and pos.shape[1] == E
    ws = torch.empty(max(_lib.lib().ocv_patch_embed_bf16_workspace(B, h, w, P), 4) // 4, device=x_nhwc.device,
                     dtype=torch.float32)
    t32 = torch.empty(B, S, E, device=x_nhwc.device, dtype=torch.float32) if want32 else None
    t16 = torch.empty(B, S, E, device=x_nhwc.device, dtype=torch.bfloat16) if want16 else None
    _call("ocv_patch_embed_bf16", x_nhwc.data_ptr(), wp.data_ptr(), _ptr(bias), _ptr(pos), _ptr(t32), _ptr(t16),
          B, Cc, h, w, P, E, ws.data_ptr(), _stream(), launches=2)
    return t32, t16


def mlp(x: torch.Tensor, weights, biases, residual=None, want32=True, want16=False):
    """Fused Linear/LeakyReLU chain (one launch).  x [R,in] fp32 with unit inner stride."""
    import ctypes as C
    _chk(x, torch.float32, "x")
    R, K0, ld = _row_major_2d(x, "x")
    n = len(weights)
    dims = [K0] + [int(w.shape[0]) for w in weights]
    for i, w in enumerate(weights):
        _chk(w, torch.float32, "w")
        assert w.is_contiguous() and w.shape[1] == dims[i]
    Wp = (C.c_void_p * n)(*[w.data_ptr() for w in weights])
    Bp = (C.c_void_p * n)(*[(b.data_ptr() if b is not None else None) for b in biases])
    Dp = (C.c_int * (n + 1))(*dims)
    o32 = torch.empty(R, dims[-1], device=x.device, dtype=torch.float32) if want32 else None
    o16 = torch.empty(R, dims[-1], device=x.device, dtype=torch.bfloat16) if want16 else None
    if residual is not None:
        assert residual.is_contiguous() and tuple(residual.shape) == (R, dims[-1])
    _call("ocv_mlp_f32", x.data_ptr(), ld, R, n, Wp, Bp, Dp, _ptr(residual), _ptr(o32), _ptr(o16), _stream())
    return o32, o16


# ---- whole head through the native executor (ocv_head_*) ----------------------------------------------

STRATEGIES = {"learned": 0, "learned_bbox_wh": 1, "grid_random": 2, "grid_random_roi_align": 3}


class NativeHead:
    """Owner of one ``ocv_head`` plan plus the torch buffers it borrows (packed weights, workspace)."""

    def __init__(self, kind: int, strategy: str, no_obj_sa: bool, use_2_saca: bool, n_bins: int, obj_dim: int,
                 min_depth: float, max_depth: float):
        import ctypes as C
        if kind == 1 and strategy not in STRATEGIES:
            raise SystemExit("Error: ObjCAViT positional embedding strategy not recognised.")  # ObjCAViT.py:284
        cfg = _lib.HeadConfig(kind, STRATEGIES.get(strategy, 0), int(no_obj_sa), int(use_2_saca), n_bins, obj_dim,
                              float(min_depth), float(max_depth))
        self._h = C.c_void_p()
        _lib.check(_lib.lib().ocv_head_create(C.byref(cfg), C.byref(self._h)), "ocv_head_create")
        self.kind = kind
        self.n_bins = n_bins
        self.packed = None
        self.sig = None
        self.ws = None
        self.lanes = 0
        self._lane_streams = []

    def set_pipeline(self, lanes: int, pixel_ctas: int = 0):
        """Opt into the pipelined serving mode of ``ocv_head_set_pipeline`` (``lanes`` = 0: back to the classic mode)."""
        L = _lib.lib()
        if self.lanes:
            self.join()
        self._check(L.ocv_head_set_pipeline(self._h, int(lanes), int(pixel_ctas)), "ocv_head_set_pipeline")
        self.lanes = int(lanes)
        self.ws = None
        self._lane_streams = []
        self._calls = 0
        if lanes:
            self._pix_stream = torch.cuda.ExternalStream(int(L.ocv_head_stream(self._h, 2, 0)))
            self._lane_streams = [torch.cuda.ExternalStream(int(L.ocv_head_stream(self._h, 0, i))) for i in range(lanes)]

    def join(self):
        """Order the current stream after every outstanding pipelined call (outputs become valid on it)."""
        if self.lanes:
            self._check(_lib.lib().ocv_head_join(self._h, _stream()), "ocv_head_join")

    def _hold(self, x_nhwc, depth, small):
        # tensors allocated on the caller's stream but used on the head's internal streams: tell the caching allocator.
        # The small object inputs / outputs only meet the lane's token stream (they are staged through lane-static buffers).
        tok = self._lane_streams[self._calls % self.lanes]
        self._calls += 1
        x_nhwc.record_stream(tok)
        x_nhwc.record_stream(self._pix_stream)
        depth.record_stream(self._pix_stream)
        for t in small:
            if t is not None:
                t.record_stream(tok)

    def __del__(self):
        try:
            if self._h:
                _lib.lib().ocv_head_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def _check(self, status, what):
        if status == -1:
            key = _lib.lib().ocv_head_missing_key(self._h).decode()
            if key:
                raise _lib.OcvError(f"objcavit_b200 {what}: parameter '{key}' was not registered")
        _lib.check(status, what)

    def bind(self, P, fingerprint=None):
        """(Re)register the parameters and rebuild the derived copies when any storage / version changed.  ``fingerprint``: the
        change detector ``KernelModule._param_state`` already computed for these views (else it is derived here)."""
        sig = fingerprint if fingerprint is not None else tuple((v.data_ptr(), v._version) for v in P.values())
        if sig == self.sig:
            return
        L = _lib.lib()
        if self.lanes:          # lanes may still read the old derived copies
            self.join()
            torch.cuda.current_stream().synchronize()
        for k, v in P.items():
            _chk(v, torch.float32, k)
            assert v.is_contiguous()
            self._check(L.ocv_head_set_param(self._h, k.encode(), v.data_ptr(), v.numel()), "ocv_head_set_param")
        dev = next(iter(P.values())).device
        nbytes = L.ocv_head_packed_bytes(self._h)
        self.packed = torch.empty(nbytes, device=dev, dtype=torch.uint8)
        global launch_count
        launch_count += 2 + sum(1 for _ in P)  # upper bound; packing is outside any timed region
        self._check(L.ocv_head_pack(self._h, self.packed.data_ptr(), _stream()), "ocv_head_pack")
        self.sig = sig

    def forward(self, x_nhwc, feats, boxes, offsets, total, n_max, obj_tokens_out=None):
        global launch_count
        _chk(x_nhwc, torch.bfloat16, "x_nhwc")
        assert x_nhwc.is_contiguous() and x_nhwc.shape[3] == 128
        B, h, w, _ = x_nhwc.shape
        L = _lib.lib()
        need = L.ocv_head_workspace_bytes(self._h, B, h, w, total, n_max)
        if need < 0:
            _lib.check(int(need), "ocv_head_workspace_bytes")
        if self.ws is None or self.ws.numel() < need or self.ws.device != x_nhwc.device:
            if self.lanes and self.ws is not None:      # outstanding lanes still use the old workspace
                self.join()
                torch.cuda.current_stream().synchronize()
            self.ws = torch.empty(need, device=x_nhwc.device, dtype=torch.uint8)
        depth = torch.empty(B, 1, h, w, device=x_nhwc.device, dtype=torch.float32)
        edges = torch.empty(B, self.n_bins + 1, device=x_nhwc.device, dtype=torch.float32)
        if self.kind == 1:
            _chk(feats, torch.float32, "object features"); _chk(boxes, torch.float32, "object boxes")
            _chk(offsets, torch.int32, "offsets")
            assert feats.is_contiguous() and boxes.is_contiguous() and boxes.shape == (total, 4) and feats.shape[0] == total
        st = _stream()
        if timed_kernel == "ocv_head_forward":
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        self._check(L.ocv_head_forward(self._h, x_nhwc.data_ptr(), _ptr(feats), _ptr(boxes), _ptr(offsets), total, n_max,
                                       B, h, w, depth.data_ptr(), edges.data_ptr(), _ptr(obj_tokens_out),
                                       self.ws.data_ptr(), st), "ocv_head_forward")
        if timed_kernel == "ocv_head_forward":
            e1.record()
            timed_events.append((e0, e1))
        launch_count += L.ocv_head_last_launches(self._h)
        if self.lanes:
            self._hold(x_nhwc, depth, (feats, boxes, offsets, obj_tokens_out, edges))
        return depth, edges


def mlp_rows(x: torch.Tensor, weights, biases):
    """Linear/LeakyReLU chain for few rows (the regressor), one launch per layer.  x [R,K0] fp32 (row stride % 4 == 0)."""
    import ctypes as C
    _chk(x, torch.float32, "x")
    R, K0, ld = _row_major_2d(x, "x")
    n = len(weights)
    dims = [K0] + [int(w.shape[0]) for w in weights]
    Wp = (C.c_void_p * n)(*[w.data_ptr() for w in weights])
    Bp = (C.c_void_p * n)(*[(b.data_ptr() if b is not None else None) for b in biases])
    Dp = (C.c_int * (n + 1))(*dims)
    ws = torch.empty(max(_lib.lib().ocv_mlp_rows_f32_workspace(R, n, Dp), 4) // 4, device=x.device, dtype=torch.float32)
    out = torch.empty(R, dims[-1], device=x.device, dtype=torch.float32)
    _call("ocv_mlp_rows_f32", x.data_ptr(), ld, R, n, Wp, Bp, Dp, out.data_ptr(), ws.data_ptr(), _stream(), launches=n)
    return out


# ---- training losses (config 5) ---------------------------------------------------------------------

def _mask_u8(mask):
    if mask is None:
        return None
    if mask.dtype == torch.bool:
        return mask.contiguous().view(torch.uint8)
    _chk(mask, torch.uint8, "depth_mask")
    return mask.contiguous()


def silog(depth_pred: torch.Tensor, depth_gt: torch.Tensor, depth_mask=None, alph: float = 10.0, lam: float = 0.85,
          want_grad: bool = False, grad_scale: float = 1.0):
    """-> loss [1] (and grad w.r.t. depth_pred when ``want_grad``)."""
    _chk(depth_pred, torch.float32, "depth_pred"); _chk(depth_gt, torch.float32, "depth_gt")
    assert depth_pred.is_contiguous() and depth_gt.is_contiguous()
    B, _, h, w = depth_pred.shape
    _, _, H, W = depth_gt.shape
    m = _mask_u8(depth_mask)
    ws = torch.empty(_lib.lib().ocv_silog_workspace(B, H, W), device=depth_pred.device, dtype=torch.uint8)
    loss = torch.empty(1, device=depth_pred.device, dtype=torch.float32)
    _call("ocv_silog_forward", depth_pred.data_ptr(), depth_gt.data_ptr(), _ptr(m), B, h, w, H, W, float(alph), float(lam),
          loss.data_ptr(), ws.data_ptr(), _stream(), launches=2)
    if not want_grad:
        return loss
    grad = torch.empty_like(depth_pred)
    _call("ocv_silog_backward", depth_pred.data_ptr(), depth_gt.data_ptr(), _ptr(m), B, h, w, H, W, float(alph), float(lam),
          float(grad_scale), grad.data_ptr(), ws.data_ptr(), _stream(), launches=2)
    return loss, grad


def bins_chamfer(bin_edges: torch.Tensor, depth_gt: torch.Tensor, depth_mask=None, want_grad: bool = False,
                 grad_scale: float = 1.0):
    """-> loss [1] (and grad w.r.t. bin_edges when ``want_grad``)."""
    _chk(bin_edges, torch.float32, "bin_edges"); _chk(depth_gt, torch.float32, "depth_gt")
    assert bin_edges.is_contiguous() and depth_gt.is_contiguous()
    B, K1 = bin_edges.shape
    HW = depth_gt.numel() // B
    m = _mask_u8(depth_mask)
    ws = torch.empty(_lib.lib().ocv_bins_chamfer_workspace(B, K1 - 1), device=bin_edges.device, dtype=torch.uint8)
    loss = torch.empty(1, device=bin_edges.device, dtype=torch.float32)
    grad = torch.empty_like(bin_edges) if want_grad else None
    _call("ocv_bins_chamfer", bin_edges.data_ptr(), depth_gt.data_ptr(), _ptr(m), B, K1 - 1, HW, float(grad_scale),
          loss.data_ptr(), _ptr(grad), ws.data_ptr(), _stream(), launches=3)
    return (loss, grad) if want_grad else loss


# ---- backward primitives of the fp32 parity path (config 5) --------------------------------------------

_ws_cache = {}


def _workspace(nbytes: int, device, tag: str = "bwd") -> torch.Tensor:
    """Grow-only scratch (uint8) per (device, tag, stream): temporaries are stream-ordered, so reuse within a stream is safe."""
    key = (device.index, tag, _stream())     # one scratch per stream: calls on two streams may overlap
    t = _ws_cache.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.empty(max(int(nbytes), 256), device=device, dtype=torch.uint8)
        _ws_cache[key] = t
    return t


def gemm_ex(a: torch.Tensor, b: torch.Tensor, M: int, N: int, K: int, lda: int, ldb: int, out: torch.Tensor, ldc: int,
            trans_a: bool = False, op_b: int = 0, beta: float = 0.0, batch: int = 1, stride_a: int = 0, stride_b: int = 0,
            stride_c: int = 0, reduce_batch: bool = False, split_k: int = 1, geom=None):
    """out = beta*out + op(a) @ op(b) (raw-stride form of ``ocv_gemm_f32_ex``; see the header for the operand modes)."""
    import ctypes as C
    _chk(a, torch.float32, "a"); _chk(b, torch.float32, "b"); _chk(out, torch.float32, "out")
    L = _lib.lib()
    need = L.ocv_gemm_f32_ex_workspace(M, N, batch, split_k)
    ws = _workspace(need, a.device, "gemm_ex")
    g = (C.c_int * 4)(*geom) if geom is not None else None
    _call("ocv_gemm_f32_ex", int(trans_a), int(op_b), a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, lda, ldb, ldc,
          float(beta), batch, stride_a, stride_b, stride_c, int(reduce_batch), split_k, g, ws.data_ptr(), _stream(), launches=2)
    return out


def _auto_split(M: int, N: int, K: int) -> int:
    """split-K factor of a token-sized SIMT GEMM: enough 64x64 tiles x splits to cover the 148 SMs, at least 64 of K per split."""
    tiles = ((M + 63) // 64) * ((N + 63) // 64)
    return max(1, min((192 + tiles - 1) // tiles, K // 64, 64))


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, out=None, beta: float = 0.0) -> torch.Tensor:
    """dx [M,K] = beta*dx + dy [M,N] @ w [N,K]  (``out``: a 2-D view with unit inner stride, e.g. one token row per image)."""
    M, N, ldy = _row_major_2d(dy, "dy")
    N2, K, ldw = _row_major_2d(w, "w")
    assert N == N2
    if out is None:
        out = torch.empty(M, K, device=dy.device, dtype=torch.float32)
        beta = 0.0
    M2, K2, ldo = _row_major_2d(out, "out")
    assert (M2, K2) == (M, K)
    return gemm_ex(dy, w, M, K, N, ldy, ldw, out, ldo, beta=beta, split_k=_auto_split(M, K, N))


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, dw: torch.Tensor, db=None, beta: float = 1.0):
    """dw [N,K] = beta*dw + dy[M,N]^T @ x[M,K];  db [N] = beta*db + column sums of dy."""
    M, N, ldy = _row_major_2d(dy, "dy")
    M2, K, ldx = _row_major_2d(x, "x")
    assert M == M2 and dw.is_contiguous() and dw.numel() == N * K
    gemm_ex(dy, x, N, K, M, ldy, ldx, dw, K, trans_a=True, beta=beta, split_k=_auto_split(N, K, M))
    if db is not None:
        assert ldy == N, "bias gradient needs a dense dy"
        reduce_channels(dy, db, 1, M, N, 1, beta)


def reduce_channels(x: torch.Tensor, out: torch.Tensor, groups: int, outer: int, C: int, inner: int, beta: float = 0.0):
    _chk(x, torch.float32, "x"); _chk(out, torch.float32, "out")
    assert x.is_contiguous() and out.is_contiguous() and x.numel() == groups * outer * C * inner and out.numel() == groups * C
    _call("ocv_reduce_channels_f32", x.data_ptr(), out.data_ptr(), groups, outer, C, inner, float(beta), _stream())
    return out


def act_bwd(dy: torch.Tensor, out: torch.Tensor, act: int) -> torch.Tensor:
    _chk(dy, torch.float32, "dy"); _chk(out, torch.float32, "out")
    assert dy.is_contiguous() and out.is_contiguous() and dy.shape == out.shape
    dx = torch.empty_like(dy)
    _call("ocv_act_bwd_f32", dy.data_ptr(), out.data_ptr(), dx.data_ptr(), dy.numel(), act, _stream())
    return dx


def layernorm_bwd(dy: torch.Tensor, z: torch.Tensor, gamma: torch.Tensor, dgamma=None, dbeta=None, beta: float = 1.0):
    _chk(dy, torch.float32, "dy"); _chk(z, torch.float32, "z")
    assert dy.is_contiguous() and z.is_contiguous() and dy.shape == z.shape
    E = z.shape[-1]
    rows = z.numel() // E
    dz = torch.empty_like(z)
    ws = _workspace(rows * 8 + 32 * E * 4, z.device, "ln_bwd")
    _call("ocv_layernorm_bwd_f32", dy.data_ptr(), z.data_ptr(), gamma.data_ptr(), dz.data_ptr(), _ptr(dgamma), _ptr(dbeta), rows, E,
          float(beta), ws.data_ptr(), _stream(), launches=3 if dgamma is not None else 1)
    return dz


def layernorm_bwd_dropout(dy: torch.Tensor, z: torch.Tensor, gamma: torch.Tensor, p: float, seed: int, dgamma, dbeta,
                          beta: float = 1.0):
    """-> (dz, dropout(dz)): the data half of the LayerNorm backward fused with the dropout site behind it, then the parameter
    half as its own call (``ocv_layernorm_bwd_dropout_f32`` + ``ocv_layernorm_bwd_params_f32``: what the layer sequencer issues on
    two streams).  Bit-identical to ``layernorm_bwd`` + ``dropout``."""
    _chk(dy, torch.float32, "dy"); _chk(z, torch.float32, "z")
    assert dy.is_contiguous() and z.is_contiguous() and dy.shape == z.shape
    E = z.shape[-1]
    rows = z.numel() // E
    dz, dzd = torch.empty_like(z), torch.empty_like(z)
    ws = _workspace(rows * 8 + 32 * E * 4, z.device, "ln_bwd")
    _call("ocv_layernorm_bwd_dropout_f32", dy.data_ptr(), z.data_ptr(), gamma.data_ptr(), dz.data_ptr(), dzd.data_ptr(), rows, E,
          float(p), int(seed), ws.data_ptr(), _stream())
    _call("ocv_layernorm_bwd_params_f32", dy.data_ptr(), z.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), rows, E, float(beta),
          ws.data_ptr(), _stream(), launches=2)
    return dz, dzd


def mha_core_bwd(q, k, v, out, dout, key_padding_mask=None, heads: int = 4, packed: bool = False, drop=None):
    """Backward of :func:`mha_core`: returns (dq, dk, dv) dense [B,S,E]; ``packed`` (self-attention, Sq == Sk): the three are
    views of one [B,S,3E] buffer laid out like the packed QKV projection output."""
    for t, n in ((q, "q"), (k, "k"), (v, "v"), (out, "out"), (dout, "dout")):
        _chk(t, torch.float32, n)
        assert t.dim() == 3 and t.stride(2) == 1
    B, Sq, E = q.shape
    Sk = k.shape[1]
    assert out.stride() == dout.stride()
    if packed:
        assert Sq == Sk
        dqkv = torch.empty(B, Sq, 3 * E, device=q.device, dtype=torch.float32)
        dq, dk, dv = dqkv[:, :, 0:E], dqkv[:, :, E:2 * E], dqkv[:, :, 2 * E:]
    else:
        dq = torch.empty(B, Sq, E, device=q.device, dtype=torch.float32)
        dk = torch.empty(B, Sk, E, device=q.device, dtype=torch.float32)
        dv = torch.empty(B, Sk, E, device=q.device, dtype=torch.float32)
    ws = _workspace(B * heads * Sq * 12, q.device, "mha_bwd")
    dp, dseed = (float(drop[0]), int(drop[1])) if drop is not None else (0.0, 0)
    _call("ocv_mha_core_bwd_dropout_f32", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), dout.data_ptr(),
          _ptr(key_padding_mask), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), B, Sq, Sk, heads, E // heads, q.stride(0), q.stride(1),
          k.stride(0), k.stride(1), v.stride(0), v.stride(1), out.stride(0), out.stride(1), dq.stride(0), dq.stride(1), dk.stride(0),
          dk.stride(1), dv.stride(0), dv.stride(1), dp, dseed, ws.data_ptr(), _stream(), launches=2)
    return (dq, dk, dv, dqkv) if packed else (dq, dk, dv)


def tail_bwd(logits: torch.Tensor, centres: torch.Tensor, grad_depth: torch.Tensor, bias=None):
    """logits [B,K,P] (overwritten by dl) -> (dl, pg).  ``bias`` [K]: added to the logits before the softmax."""
    _chk(logits, torch.float32, "logits")
    assert logits.is_contiguous() and centres.is_contiguous() and grad_depth.is_contiguous()
    B, K, P = logits.shape
    pg = torch.empty_like(logits)
    _call("ocv_tail_bwd_f32", logits.data_ptr(), pg.data_ptr(), centres.data_ptr(), grad_depth.data_ptr(), _ptr(bias), B, K, P,
          _stream())
    return logits, pg


def bins_bwd(y_raw: torch.Tensor, grad_edges, grad_centres, min_depth: float, max_depth: float, grad_widths=None) -> torch.Tensor:
    _chk(y_raw, torch.float32, "y_raw")
    assert y_raw.is_contiguous()
    B, K = y_raw.shape
    dy = torch.empty_like(y_raw)
    if grad_widths is not None:
        _chk(grad_widths, torch.float32, "grad_widths")
        assert grad_widths.is_contiguous() and grad_widths.shape == y_raw.shape
    _call("ocv_bins_bwd_ex_f32", y_raw.data_ptr(), _ptr(grad_edges), _ptr(grad_centres), _ptr(grad_widths), dy.data_ptr(), B, K,
          float(min_depth), float(max_depth), _stream())
    return dy


def unpad_objects(dpadded: torch.Tensor, offsets: torch.Tensor, total: int, n_max: int, left_pad: bool, out=None, beta: float = 0.0):
    """Adjoint of :func:`pad_objects`: dpadded [B,L,E] -> gradient of the ragged rows [total,E]."""
    _chk(dpadded, torch.float32, "dpadded")
    assert dpadded.is_contiguous() and dpadded.dim() == 3
    B, L, E = dpadded.shape
    if out is None:
        out = torch.empty(total, E, device=dpadded.device, dtype=torch.float32)
        beta = 0.0
    _call("ocv_unpad_objects_f32", dpadded.data_ptr(), offsets.data_ptr(), out.data_ptr(), B, L, n_max, E, int(left_pad), float(beta),
          _stream())
    return out


def sumsq(x: torch.Tensor, scale: float = 1.0, out=None) -> torch.Tensor:
    """out[0] = scale^2 * sum(x^2) (deterministic); stays on the device."""
    _chk(x, torch.float32, "x")
    assert x.is_contiguous()
    if out is None:
        out = torch.empty(1, device=x.device, dtype=torch.float32)
    ws = _workspace(_lib.lib().ocv_sumsq_workspace(x.numel()), x.device, "sumsq")
    _call("ocv_sumsq_f32", x.data_ptr(), x.numel(), float(scale), out.data_ptr(), ws.data_ptr(), _stream(), launches=2)
    return out


def adamw_step(param: torch.Tensor, grad: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor, lr: float, beta1: float,
               beta2: float, eps: float, weight_decay: float, step: int, grad_scale: float = 1.0, grad_sumsq=None,
               max_norm: float = 0.0):
    """In-place AdamW update of a flat parameter buffer (``torch.optim.AdamW`` semantics, optional fused norm clipping)."""
    for t, n in ((param, "param"), (grad, "grad"), (exp_avg, "exp_avg"), (exp_avg_sq, "exp_avg_sq")):
        _chk(t, torch.float32, n)
        assert t.is_contiguous() and t.numel() == param.numel()
    _call("ocv_adamw_step_f32", param.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(), param.numel(), float(lr),
          float(beta1), float(beta2), float(eps), float(weight_decay), int(step), float(grad_scale), _ptr(grad_sumsq), float(max_norm),
          _stream())


# ---- rows either side of the head -------------------------------------------------

def tta_merge(depth: torch.Tensor, depth_mirror, min_depth: float, max_depth: float) -> torch.Tensor:
    """0.5 * (clamp(depth) + clamp(flip_w(depth_mirror))) (``GraphBinsLM.py:159-183``); ``depth_mirror`` None -> clamp only."""
    _chk(depth, torch.float32, "depth")
    assert depth.is_contiguous() and (depth_mirror is None or (depth_mirror.is_contiguous() and depth_mirror.shape == depth.shape))
    B, _, h, w = depth.shape
    out = torch.empty_like(depth)
    _call("ocv_tta_merge_f32", depth.data_ptr(), _ptr(depth_mirror), out.data_ptr(), B, h, w, float(min_depth), float(max_depth),
          _stream())
    return out


def depth_metrics(depth_pred: torch.Tensor, depth_gt: torch.Tensor, min_depth: float, max_depth: float, crop=None, acc=None,
                  want_upsampled: bool = False):
    """-> (sums [9] float64 of this call, up-sampled prediction | None); ``acc`` [9] float64 is incremented in place."""
    import ctypes as C
    _chk(depth_pred, torch.float32, "depth_pred"); _chk(depth_gt, torch.float32, "depth_gt")
    assert depth_pred.is_contiguous() and depth_gt.is_contiguous()
    B, _, h, w = depth_pred.shape
    _, _, H, W = depth_gt.shape
    sums = torch.empty(9, device=depth_pred.device, dtype=torch.float64)
    up = torch.empty_like(depth_gt) if want_upsampled else None
    if acc is not None:
        assert acc.dtype == torch.float64 and acc.is_cuda and acc.numel() == 9 and acc.is_contiguous()
    ws = _workspace(_lib.lib().ocv_depth_metrics_workspace(), depth_pred.device, "metrics")
    c = (C.c_int * 4)(*[int(v) for v in crop]) if crop is not None else None
    _call("ocv_depth_metrics_f32", depth_pred.data_ptr(), depth_gt.data_ptr(), B, h, w, H, W, float(min_depth), float(max_depth), c,
          _ptr(up), sums.data_ptr(), _ptr(acc), ws.data_ptr(), _stream(), launches=2)
    return sums, up


def object_glue(xyxy: torch.Tensor, offsets: torch.Tensor, n_max: int, reverse: bool = True, n_scale: int = 7,
                want_bucket: bool = True):
    """corner boxes (CSR per image) -> (centre/size boxes [total,4], relative-size bucket [total] int32 | None)."""
    _chk(xyxy, torch.float32, "xyxy"); _chk(offsets, torch.int32, "offsets")
    assert xyxy.is_contiguous() and xyxy.shape[1] == 4
    B = offsets.numel() - 1
    xywh = torch.empty_like(xyxy)
    bucket = torch.empty(xyxy.shape[0], device=xyxy.device, dtype=torch.int32) if want_bucket else None
    _call("ocv_object_glue_f32", xyxy.data_ptr(), offsets.data_ptr(), xywh.data_ptr(), _ptr(bucket), B, int(n_max), int(reverse),
          int(n_scale), _stream())
    return xywh, bucket


def conv3x3_nhwc_bf16(x_nhwc: torch.Tensor, w3p: torch.Tensor, bias: torch.Tensor, max_ctas: int = 0) -> torch.Tensor:
    """[B,h,w,128] bf16 NHWC -> conv3x3 (pad 1) + bias -> [B,h,w,128] bf16 NHWC on the tcgen05 implicit-GEMM pipeline."""
    _chk(x_nhwc, torch.bfloat16, "x_nhwc"); _chk(w3p, torch.bfloat16, "w3p"); _chk(bias, torch.float32, "bias")
    assert x_nhwc.is_contiguous() and x_nhwc.shape[3] == 128 and bias.numel() == 128
    B, h, w, _ = x_nhwc.shape
    y = torch.empty_like(x_nhwc)
    _call("ocv_conv3x3_nhwc_bf16", x_nhwc.data_ptr(), w3p.data_ptr(), bias.data_ptr(), y.data_ptr(), B, h, w, int(max_ctas), _stream())
    return y


def dropout(x: torch.Tensor, p: float, seed: int, residual=None, out=None) -> torch.Tensor:
    """out = residual + dropout(x) with the counter-based mask of site ``seed`` (``ocv_dropout_f32``); on a gradient: its backward."""
    _chk(x, torch.float32, "x")
    assert x.is_contiguous() and (residual is None or (residual.is_contiguous() and residual.shape == x.shape))
    if out is None:
        out = torch.empty_like(x)
    _call("ocv_dropout_f32", x.data_ptr(), _ptr(residual), out.data_ptr(), x.numel(), float(p), int(seed), _stream())
    return out


def dropout_add_layernorm(x: torch.Tensor, residual: torch.Tensor, gamma, beta, p: float, seed: int):
    """-> (sum = residual + dropout(x), LayerNorm(sum)) in one kernel; bit-identical to ``dropout`` + ``add_layernorm``."""
    _chk(x, torch.float32, "x"); _chk(residual, torch.float32, "residual")
    assert x.is_contiguous() and residual.is_contiguous() and residual.shape == x.shape
    E = x.shape[-1]
    total, y = torch.empty_like(x), torch.empty_like(x)
    _call("ocv_dropout_add_layernorm_f32", x.data_ptr(), residual.data_ptr(), gamma.data_ptr(), beta.data_ptr(), total.data_ptr(),
          y.data_ptr(), x.numel() // E, E, float(p), int(seed), _stream())
    return total, y


def dropout_act_bwd(dy: torch.Tensor, out: torch.Tensor, act: int, p: float, seed: int) -> torch.Tensor:
    """dx = act'(out) * dropout(dy) in one kernel; bit-identical to ``dropout`` + ``act_bwd``."""
    _chk(dy, torch.float32, "dy"); _chk(out, torch.float32, "out")
    assert dy.is_contiguous() and out.is_contiguous() and dy.shape == out.shape
    dx = torch.empty_like(dy)
    _call("ocv_dropout_act_bwd_f32", dy.data_ptr(), out.data_ptr(), dx.data_ptr(), dy.numel(), act, float(p), int(seed), _stream())
    return dx


# ---- training step, tensor-core pixel path (csrc/train_pixel_bf16.cu) -------------------------------------------------

def tail_fused_fwd_bf16(x3_nhwc: torch.Tensor, Mb2: torch.Tensor, blog2: torch.Tensor, centres: torch.Tensor):
    """-> (depth [B,P], stat_max [B,P], stat_sum [B,P]) from x3 [B,h,w,128] / [B,P,128] bf16 without materialising the logits:
    logits (log2 domain) = x3 @ Mb2[b]^T + blog2[b], softmax over the 256 bins in registers, depth = sum_k p_k centres[b,k]."""
    _chk(x3_nhwc, torch.bfloat16, "x3"); _chk(Mb2, torch.bfloat16, "Mb2"); _chk(blog2, torch.float32, "blog2")
    _chk(centres, torch.float32, "centres")
    assert x3_nhwc.is_contiguous() and Mb2.is_contiguous() and blog2.is_contiguous() and centres.is_contiguous()
    B = Mb2.shape[0]
    Pn = x3_nhwc.numel() // (B * 128)
    assert tuple(Mb2.shape) == (B, 256, 128) and tuple(blog2.shape) == (B, 256) and tuple(centres.shape) == (B, 256)
    out = torch.empty(3, B, Pn, device=x3_nhwc.device, dtype=torch.float32)
    _call("ocv_tail_fused_fwd_bf16", x3_nhwc.data_ptr(), Mb2.data_ptr(), blog2.data_ptr(), centres.data_ptr(), out[0].data_ptr(),
          out[1].data_ptr(), out[2].data_ptr(), B, Pn, _stream())
    return out[0], out[1], out[2]


def tail_fused_bwd_bf16(x3_nhwc: torch.Tensor, Mb2: torch.Tensor, blog2: torch.Tensor, centres: torch.Tensor, depth: torch.Tensor,
                        stat_max: torch.Tensor, stat_sum: torch.Tensor, grad_depth: torch.Tensor, grad_bias=None, beta: float = 1.0):
    """-> (dx3T [128, B*P] bf16, dMb [B,256,128] fp32, grad_centres [B,256]); ``grad_bias`` [256] beta-accumulated in place.  The
    logits are recomputed tile by tile on the tensor cores; nothing of size [B,256,P] touches HBM."""
    _chk(x3_nhwc, torch.bfloat16, "x3"); _chk(Mb2, torch.bfloat16, "Mb2"); _chk(grad_depth, torch.float32, "grad_depth")
    for t in (depth, stat_max, stat_sum):
        _chk(t, torch.float32, "stats")
        assert t.is_contiguous()
    assert x3_nhwc.is_contiguous() and Mb2.is_contiguous() and grad_depth.is_contiguous()
    B = Mb2.shape[0]
    Pn = x3_nhwc.numel() // (B * 128)
    dev = x3_nhwc.device
    dx3T = torch.empty(128, B * Pn, device=dev, dtype=torch.bfloat16)
    dMb = torch.empty(B, 256, 128, device=dev, dtype=torch.float32)
    dcent = torch.empty(B, 256, device=dev, dtype=torch.float32)
    ws = _workspace(_lib.lib().ocv_tail_fused_bwd_workspace(B, Pn), dev, "tail_fused_bwd")
    _call("ocv_tail_fused_bwd_bf16", x3_nhwc.data_ptr(), Mb2.data_ptr(), blog2.data_ptr(), centres.data_ptr(), depth.data_ptr(),
          stat_max.data_ptr(), stat_sum.data_ptr(), grad_depth.data_ptr(), dx3T.data_ptr(), B * Pn, dMb.data_ptr(), dcent.data_ptr(),
          _ptr(grad_bias), float(beta), B, Pn, ws.data_ptr(), _stream(), launches=3)
    return dx3T, dMb, dcent


def tail_logits(logits: torch.Tensor, bias, centres: torch.Tensor, want_stats: bool = False):
    """depth [B,P] = sum_k softmax_k(logits[b,k,p] + bias[k]) centres[b,k]; logits channel-major fp32 [B,K,P].
    ``want_stats``: -> (depth, stats [B,3,P]) — the softmax statistics ``tail_logits_bwd`` would otherwise recompute."""
    _chk(logits, torch.float32, "logits"); _chk(centres, torch.float32, "centres")
    assert logits.is_contiguous() and centres.is_contiguous()
    B, K, Pn = logits.shape
    depth = torch.empty(B, Pn, device=logits.device, dtype=torch.float32)
    if want_stats:
        stats = torch.empty(B, 3, Pn, device=logits.device, dtype=torch.float32)
        _call("ocv_tail_logits_stats_f32", logits.data_ptr(), _ptr(bias), centres.data_ptr(), depth.data_ptr(), stats.data_ptr(), B, K,
              Pn, _stream())
        return depth, stats
    _call("ocv_tail_logits_f32", logits.data_ptr(), _ptr(bias), centres.data_ptr(), depth.data_ptr(), B, K, Pn, _stream())
    return depth


def tail_logits_bwd(logits: torch.Tensor, bias, centres: torch.Tensor, grad_depth: torch.Tensor, grad_bias=None, beta: float = 1.0,
                    stats=None):
    """-> (dlT [B,K,P] bf16, grad_centres [B,K]); ``grad_bias`` [K] is beta-accumulated in place when given.  ``stats``: from
    ``tail_logits(..., want_stats=True)`` on the same logits (one pass over the logits instead of two; same result bit for bit)."""
    _chk(logits, torch.float32, "logits"); _chk(grad_depth, torch.float32, "grad_depth")
    assert logits.is_contiguous() and centres.is_contiguous() and grad_depth.is_contiguous()
    B, K, Pn = logits.shape
    dlT = torch.empty(B, K, Pn, device=logits.device, dtype=torch.bfloat16)
    dcent = torch.empty(B, K, device=logits.device, dtype=torch.float32)
    ws = _workspace(_lib.lib().ocv_tail_logits_bwd_workspace(B, K, Pn), logits.device, "tail_logits_bwd")
    assert stats is None or (stats.is_contiguous() and tuple(stats.shape) == (B, 3, Pn) and stats.dtype == torch.float32)
    _call("ocv_tail_logits_bwd_stats_bf16", logits.data_ptr(), _ptr(bias), centres.data_ptr(), _ptr(stats), grad_depth.data_ptr(),
          dlT.data_ptr(), dcent.data_ptr(), _ptr(grad_bias), float(beta), B, K, Pn, ws.data_ptr(), _stream(), launches=3)
    return dlT, dcent


def transpose_bf16(src: torch.Tensor, out=None) -> torch.Tensor:
    """[Z,R,C] bf16 (unit inner stride) -> [Z,C,R]; ``out``: any [Z,C,R] view with unit inner stride (e.g. a slice of [C, Z*R])."""
    _chk(src, torch.bfloat16, "src")
    Z, R, Cc = src.shape
    assert src.stride(2) == 1
    if out is None:
        out = torch.empty(Z, Cc, R, device=src.device, dtype=torch.bfloat16)
    assert tuple(out.shape) == (Z, Cc, R) and out.stride(2) == 1
    _call("ocv_transpose_bf16", src.data_ptr(), out.data_ptr(), Z, R, Cc, src.stride(0), src.stride(1), out.stride(0), out.stride(1),
          _stream())
    return out


def im2col3x3_bf16(x_planar: torch.Tensor) -> torch.Tensor:
    """[B,C,h,w] bf16 -> col [C*9+8, B*h*w] bf16 (9 shifted copies per channel, a row of ones, 7 zero rows)."""
    _chk(x_planar, torch.bfloat16, "x_planar")
    assert x_planar.is_contiguous()
    B, Cc, h, wd = x_planar.shape
    col = torch.empty(Cc * 9 + 8, B * h * wd, device=x_planar.device, dtype=torch.bfloat16)
    _call("ocv_im2col3x3_bf16", x_planar.data_ptr(), col.data_ptr(), B, Cc, h, wd, _stream())
    return col


def patch_unfold_bf16(x: torch.Tensor, P: int = 16) -> torch.Tensor:
    """unfold(x) [C*P*P, ld] bf16 (ld = B*S rounded up to 8): the K-major B operand of the patch-embedding weight-gradient GEMM.
    Depends on the features only, so the training step launches it off the critical path."""
    _chk(x, torch.float32, "x")
    assert x.is_contiguous()
    B, Cc, h, wd = x.shape
    BS = B * (h // P) * (wd // P)
    ld = (BS + 7) // 8 * 8
    alloc = torch.empty if ld == BS else torch.zeros
    unf = alloc(Cc * P * P, ld, device=x.device, dtype=torch.bfloat16)
    _call("ocv_patch_unfold_bf16", x.data_ptr(), unf.data_ptr(), B, Cc, h, wd, P, ld, _stream())
    return unf


def conv3x3_wgrad_bf16(dyT: torch.Tensor, xh: torch.Tensor, gw: torch.Tensor, gb: Optional[torch.Tensor], beta: float = 1.0) -> None:
    """gw [128,128,3,3] (+ gb [128]) = beta * (.) + conv3x3 weight / bias gradient from dY channel-major ``dyT`` [128, >= B*h*w] bf16 and
    the NHWC bf16 features ``xh`` [B,h,w,128]: implicit GEMM over pixels on tcgen05, no im2col (csrc/conv_wgrad_bf16.cu)."""
    _chk(dyT, torch.bfloat16, "dyT"); _chk(xh, torch.bfloat16, "xh"); _chk(gw, torch.float32, "gw")
    B, h, wd, Cc = xh.shape
    assert Cc == 128 and xh.is_contiguous() and dyT.dim() == 2 and dyT.shape[0] == 128 and dyT.stride(1) == 1 and gw.is_contiguous()
    assert gw.numel() == 128 * 128 * 9 and wd % 8 == 0
    if gb is not None:
        _chk(gb, torch.float32, "gb")
        assert gb.is_contiguous() and gb.numel() == 128
    ws = _workspace(_lib.lib().ocv_conv3x3_wgrad_bf16_workspace(B, h, wd), xh.device, "conv3x3_wgrad")
    _call("ocv_conv3x3_wgrad_bf16", dyT.data_ptr(), dyT.stride(0), xh.data_ptr(), gw.data_ptr(), 0 if gb is None else gb.data_ptr(),
          B, h, wd, float(beta), ws.data_ptr(), _stream())


def patch_unfold_nhwc_bf16(xh: torch.Tensor, P: int = 16) -> torch.Tensor:
    """``patch_unfold_bf16`` from the NHWC bf16 copy of the features ``xh`` [B,h,w,C] (what the tensor-core path already holds)."""
    _chk(xh, torch.bfloat16, "xh")
    assert xh.is_contiguous()
    B, h, wd, Cc = xh.shape
    BS = B * (h // P) * (wd // P)
    ld = (BS + 7) // 8 * 8
    alloc = torch.empty if ld == BS else torch.zeros
    unf = alloc(Cc * P * P, ld, device=xh.device, dtype=torch.bfloat16)
    _call("ocv_patch_unfold_nhwc_bf16", xh.data_ptr(), unf.data_ptr(), B, Cc, h, wd, P, ld, _stream())
    return unf


def patch_embed_wgrad_bf16(dtok: torch.Tensor, x: torch.Tensor, dw: torch.Tensor, P: int = 16, unf: Optional[torch.Tensor] = None):
    """dw [E, C, P, P] += dtok [B*S, E]^T @ unfold(x): the patch-embedding weight gradient as ONE tcgen05 GEMM (bf16 operands).
    ``unf``: a precomputed ``patch_unfold_bf16(x, P)``."""
    _chk(dtok, torch.float32, "dtok"); _chk(dw, torch.float32, "dw")
    assert dtok.is_contiguous() and dw.is_contiguous()
    if unf is None:                       # with a precomputed unfold, x gives the shape only (it may be a storage-less placeholder)
        _chk(x, torch.float32, "x")
        assert x.is_contiguous()
    B, Cc, h, wd = x.shape
    BS, E = dtok.shape
    assert BS == B * (h // P) * (wd // P) and dw.numel() == E * Cc * P * P
    ld = (BS + 7) // 8 * 8
    alloc = torch.empty if ld == BS else torch.zeros
    if unf is None:
        unf = patch_unfold_bf16(x, P)
    assert unf.shape == (Cc * P * P, ld)
    dT = alloc(E, ld, device=x.device, dtype=torch.bfloat16)
    _call("ocv_transpose_cast_f32_bf16", dtok.data_ptr(), dT.data_ptr(), BS, E, E, ld, _stream())
    g = dw.view(E, Cc * P * P)
    gemm_bf16(dT, unf, residual=g, out=g)


# ---- positional-table gradients (grid_random / grid_random_roi_align training) ------------------------------------------

def table_grad_buffer(gh: int, gw: int, E: int, device) -> torch.Tensor:
    """Zeroed fixed-point accumulator [gh*gw, E] int64 for the two scatter kernels below."""
    return torch.zeros(gh * gw, E, device=device, dtype=torch.int64)


def grid_sample_pos_bwd(dout: torch.Tensor, pts: torch.Tensor, fix: torch.Tensor, gh: int, gw: int, norm_mode: int = 0,
                        div0: float = 1.0, div1: float = 1.0) -> None:
    _chk(dout, torch.float32, "dout"); _chk(pts, torch.float32, "pts"); _chk(fix, torch.int64, "fix")
    assert dout.is_contiguous() and pts.is_contiguous() and fix.is_contiguous() and pts.shape == (dout.shape[0], 2)
    N, E = dout.shape
    assert fix.shape == (gh * gw, E)
    _call("ocv_grid_sample_pos_bwd_f32", dout.data_ptr(), pts.data_ptr(), fix.data_ptr(), N, gh, gw, E, norm_mode, float(div0),
          float(div1), _stream())


def ps_roi_align_pos_bwd(dout: torch.Tensor, boxes: torch.Tensor, fix: torch.Tensor, gh: int, gw: int, spatial_scale: float,
                         boxes_are_xywh: bool) -> None:
    _chk(dout, torch.float32, "dout"); _chk(boxes, torch.float32, "boxes"); _chk(fix, torch.int64, "fix")
    assert dout.is_contiguous() and boxes.is_contiguous() and fix.is_contiguous() and boxes.shape == (dout.shape[0], 4)
    N, E = dout.shape
    assert fix.shape == (gh * gw, E)
    _call("ocv_ps_roi_align_pos_bwd_f32", dout.data_ptr(), boxes.data_ptr(), fix.data_ptr(), N, gh, gw, E, float(spatial_scale),
          int(boxes_are_xywh), _stream())


def table_grad_commit(fix: torch.Tensor, dtable: torch.Tensor, beta: float = 1.0) -> None:
    """dtable[:gh*gw] = beta * dtable[:gh*gw] + fix / 2^40  (``dtable``: the leading rows of the table's gradient)."""
    _chk(fix, torch.int64, "fix"); _chk(dtable, torch.float32, "dtable")
    assert fix.is_contiguous() and dtable.is_contiguous() and dtable.numel() == fix.numel()
    _call("ocv_table_grad_commit_f32", fix.data_ptr(), dtable.data_ptr(), fix.numel(), float(beta), _stream())


# ---- native sequencer of one encoder layer of the training step (csrc/train_layer.cu) ---------------------------------------

LAYER_KEYS = ("self_attn.in_proj_weight", "self_attn.in_proj_bias", "self_attn.out_proj.weight", "self_attn.out_proj.bias",
              "linear1.weight", "linear1.bias", "linear2.weight", "linear2.bias", "norm1.weight", "norm1.bias", "norm2.weight",
              "norm2.bias")


def _ptr_array(tensors):
    import ctypes as C
    return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def encoder_layer_fwd_train(W, x: torch.Tensor, mask, drop_p: float, seeds):
    """One nn.TransformerEncoderLayer forward in train() mode, one C call.  W: the 12 parameter tensors in ``LAYER_KEYS`` order;
    x [B,S,128] contiguous.  -> (out [B,S,128], (qkv, att, y, x1, h, y2)) — the activations the backward reads."""
    import ctypes as C
    _chk(x, torch.float32, "x")
    assert x.is_contiguous() and x.dim() == 3 and x.shape[2] == 128 and len(W) == 12
    B, S, E = x.shape
    F_ = W[4].shape[0]
    M, dev = B * S, x.device
    qkv = torch.empty(B, S, 3 * E, device=dev, dtype=torch.float32)
    att, y, x1, y2, out = (torch.empty(M, E, device=dev, dtype=torch.float32) for _ in range(5))
    h = torch.empty(M, F_, device=dev, dtype=torch.float32)
    ws = _workspace(_lib.lib().ocv_encoder_layer_fwd_train_workspace(B, S, F_), dev, "layer_fwd")
    sd = (C.c_ulonglong * 4)(*[int(v) for v in seeds]) if drop_p > 0 else None
    _call("ocv_encoder_layer_fwd_train_f32", _ptr_array(W), x.data_ptr(), _ptr(mask), B, S, F_, float(drop_p), sd, qkv.data_ptr(),
          att.data_ptr(), y.data_ptr(), x1.data_ptr(), h.data_ptr(), y2.data_ptr(), out.data_ptr(), ws.data_ptr(), _stream(),
          launches=11 if drop_p > 0 else 9)
    return out.view(B, S, E), (qkv, att.view(B, S, E), y, x1, h, y2)


def encoder_layer_bwd_train(W, G, x, qkv, att, y, x1, h, y2, mask, dout: torch.Tensor, drop_p: float, seeds, side_stream=None):
    """Explicit backward of the same layer, one C call: -> dx [B,S,128]; parameter gradients are ACCUMULATED into the 12 tensors
    of ``G``, the weight / bias gradient GEMMs on ``side_stream`` (a torch stream; the caller joins it before reading G)."""
    import ctypes as C
    _chk(dout, torch.float32, "dout")
    B, S, E3 = qkv.shape
    E = E3 // 3
    F_ = W[4].shape[0]
    dev = dout.device
    d = dout.reshape(B * S, E)
    assert d.is_contiguous() and x.is_contiguous()
    dx = torch.empty(B * S, E, device=dev, dtype=torch.float32)
    ws = torch.empty(_lib.lib().ocv_encoder_layer_bwd_train_workspace(B, S, F_), device=dev, dtype=torch.uint8)
    sd = (C.c_ulonglong * 4)(*[int(v) for v in seeds]) if drop_p > 0 else None
    _call("ocv_encoder_layer_bwd_train_f32", _ptr_array(W), _ptr_array(G), x.data_ptr(), qkv.data_ptr(), att.data_ptr(), y.data_ptr(),
          x1.data_ptr(), h.data_ptr(), y2.data_ptr(), _ptr(mask), d.data_ptr(), dx.data_ptr(), B, S, F_, float(drop_p), sd,
          ws.data_ptr(), _stream(), None if side_stream is None else side_stream.cuda_stream, launches=29)
    if side_stream is not None:
        # the LayerNorm-parameter sums on the side stream read dout / y2 / y as well (the caller frees dout right after this call)
        for t in (ws, x, att, x1, h, d, y2, y):
            t.record_stream(side_stream)
    return dx.view(B, S, E)


LAYER_W16_KEYS = ("self_attn.in_proj_weight", "self_attn.out_proj.weight", "linear1.weight", "linear2.weight")


def gemm_bf16_v2(a: torch.Tensor, b: torch.Tensor, a_mn: bool = False, b_mn: bool = False, bias=None, act: int = ACT_NONE,
                 residual=None, ln=None, want_pre: bool = False, want32: bool = True, want16: bool = False, drop=None,
                 split_k: int = 1, out32=None):
    """Training-step GEMM on tcgen05 (``ocv_gemm_bf16_v2``): ``[LN](dropout(act(op(a) op(b)^T + bias)) + residual)``.
    ``a``: [M,K] (or [K,M] with ``a_mn``), ``b``: [N,K] (or [K,N] with ``b_mn``), bf16 row-major; ``drop`` = (p, seed).
    -> (pre | None, out32 | None, out16 | None)."""
    _chk(a, torch.bfloat16, "a"); _chk(b, torch.bfloat16, "b")
    r0, c0, lda = _row_major_2d(a, "a")
    r1, c1, ldb = _row_major_2d(b, "b")
    M, K = (c0, r0) if a_mn else (r0, c0)
    N, K2 = (c1, r1) if b_mn else (r1, c1)
    assert K == K2, (a.shape, b.shape, a_mn, b_mn)
    dev = a.device
    o32 = out32 if out32 is not None else (torch.empty(M, N, device=dev, dtype=torch.float32) if want32 else None)
    o16 = torch.empty(M, N, device=dev, dtype=torch.bfloat16) if want16 else None
    pre = torch.empty(M, N, device=dev, dtype=torch.float32) if want_pre else None
    ldr = 0
    if residual is not None:
        _chk(residual, torch.float32, "residual")
        _, _, ldr = _row_major_2d(residual, "residual")
    ws = None
    if split_k > 1:
        ws = _workspace(_lib.lib().ocv_gemm_bf16_workspace(M, N, K, split_k), dev, "gemm_bf16")
    p_drop, seed = drop if drop is not None else (0.0, 0)
    _call("ocv_gemm_bf16_v2", a.data_ptr(), b.data_ptr(), int(a_mn), int(b_mn), _ptr(bias), _ptr(residual), ldr,
          _ptr(ln[0]) if ln else None, _ptr(ln[1]) if ln else None, _ptr(pre), N, _ptr(o32), o32.stride(0) if o32 is not None else 0,
          _ptr(o16), N, M, N, K, lda, ldb, int(act), float(p_drop), int(seed), int(split_k), _ptr(ws), _stream(),
          launches=2 if split_k > 1 else 1)
    return pre, o32, o16


def encoder_layer_fwd_train_bf16(W, W16, x: torch.Tensor, x16: torch.Tensor, mask, drop_p: float, seeds):
    """Tensor-core form of :func:`encoder_layer_fwd_train`.  W16: the 4 bf16 weight copies in ``LAYER_W16_KEYS`` order; x16: bf16
    copy of x.  -> (out fp32 [B,S,128], out bf16, saved) with saved = (x16, qkv, att, att16, y, x1_16, h, h16, y2)."""
    import ctypes as C
    _chk(x, torch.float32, "x"); _chk(x16, torch.bfloat16, "x16")
    assert x.is_contiguous() and x16.is_contiguous() and x.dim() == 3 and x.shape[2] == 128 and len(W) == 12 and len(W16) == 4
    B, S, E = x.shape
    F_ = W[4].shape[0]
    M, dev = B * S, x.device
    f32 = lambda *s: torch.empty(*s, device=dev, dtype=torch.float32)
    b16 = lambda *s: torch.empty(*s, device=dev, dtype=torch.bfloat16)
    qkv, att, y, x1, y2, out, h = f32(B, S, 3 * E), f32(M, E), f32(M, E), f32(M, E), f32(M, E), f32(M, E), f32(M, F_)
    att16, x1_16, out16, h16 = b16(M, E), b16(M, E), b16(M, E), b16(M, F_)
    qkv16 = b16(M, 3 * E)            # operand of the tensor-core attention core, forward and backward
    stats = f32(B, 4, S, 2)          # its softmax statistics (row max, row sum): the backward recomputes P from them
    sd = (C.c_ulonglong * 4)(*[int(v) for v in seeds]) if drop_p > 0 else None
    _call("ocv_encoder_layer_fwd_train_bf16", _ptr_array(W), _ptr_array(W16), x.data_ptr(), x16.data_ptr(), _ptr(mask), B, S, F_,
          float(drop_p), sd, qkv.data_ptr(), qkv16.data_ptr(), stats.data_ptr(), att.data_ptr(), att16.data_ptr(), y.data_ptr(),
          x1.data_ptr(), x1_16.data_ptr(), h.data_ptr(), h16.data_ptr(), y2.data_ptr(), out.data_ptr(), out16.data_ptr(), _stream(),
          launches=6)
    return out.view(B, S, E), out16.view(B, S, E), (x16, qkv, att.view(B, S, E), att16, y, x1_16, h, h16, y2, qkv16, stats)


def encoder_layer_bwd_train_bf16(W, W16, G, saved, mask, dout: torch.Tensor, drop_p: float, seeds, side_stream=None):
    """Explicit backward of the layer on the tensor cores: -> dx [B,S,128]; parameter gradients ACCUMULATED into ``G``."""
    import ctypes as C
    x16, qkv, att, att16, y, x1_16, h, h16, y2, qkv16, stats = saved
    _chk(dout, torch.float32, "dout")
    B, S, E3 = qkv.shape
    E = E3 // 3
    F_ = W[4].shape[0]
    dev = dout.device
    d = dout.reshape(B * S, E)
    assert d.is_contiguous()
    dx = torch.empty(B * S, E, device=dev, dtype=torch.float32)
    ws = torch.empty(_lib.lib().ocv_encoder_layer_bwd_train_bf16_workspace(B, S, F_), device=dev, dtype=torch.uint8)
    sd = (C.c_ulonglong * 4)(*[int(v) for v in seeds]) if drop_p > 0 else None
    _call("ocv_encoder_layer_bwd_train_bf16", _ptr_array(W), _ptr_array(W16), _ptr_array(G), x16.data_ptr(), qkv.data_ptr(),
          qkv16.data_ptr(), stats.data_ptr(), att.data_ptr(), att16.data_ptr(), y.data_ptr(), x1_16.data_ptr(), h.data_ptr(), h16.data_ptr(), y2.data_ptr(), _ptr(mask),
          d.data_ptr(), dx.data_ptr(), B, S, F_, float(drop_p), sd, ws.data_ptr(), _stream(),
          None if side_stream is None else side_stream.cuda_stream, launches=30)
    if side_stream is not None:
        # the LayerNorm-parameter sums on the side queues read dout / y2 / y as well (the caller frees dout right after this call)
        for t in (ws, x16, att16, x1_16, h16, d, y2, y):
            t.record_stream(side_stream)
    return dx.view(B, S, E)
