// Token-path kernels (fp32): residual+LayerNorm, attention core, ragged-object padding, positional
// lookups (grid_sample / ps_roi_align restated), bin-width normalisation.  Contracts: include/objcavit_b200.h
#include "common.cuh"
#include <math.h>

namespace {

// ------------------------------------------------------------------------------------------------
// y = LN(x + r): one warp per row, E <= 512, E % 32 == 0
// ------------------------------------------------------------------------------------------------
__global__ void add_layernorm_kernel(const float* __restrict__ x, const float* __restrict__ r,
                                     const float* __restrict__ gamma, const float* __restrict__ beta,
                                     float* __restrict__ y, int rows, int E, const uint8_t* __restrict__ zero_rows) {
  const int row = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x % 32;
  if (row >= rows) return;
  const int per = E / 32;
  float v[16];
  const float* xr = x + (long long)row * E;
  const float* rr = r ? r + (long long)row * E : nullptr;
  float s = 0.f;
  for (int i = 0; i < per; ++i) {
    const int e = i * 32 + lane;
    v[i] = xr[e] + (rr ? rr[e] : 0.f);
    s += v[i];
  }
  const float mean = warp_sum(s) / E;
  float q = 0.f;
  for (int i = 0; i < per; ++i) {
    const float d = v[i] - mean;
    q += d * d;
  }
  const float rstd = rsqrtf(warp_sum(q) / E + 1e-5f);
  const bool zero = zero_rows && zero_rows[row];
  float* yr = y + (long long)row * E;
  for (int i = 0; i < per; ++i) {
    const int e = i * 32 + lane;
    yr[e] = zero ? 0.f : (v[i] - mean) * rstd * gamma[e] + beta[e];
  }
}

// sum = r + dropout(x) (counter-based mask over the flat [rows, E] index, as ocv_dropout_f32), y = LN(sum): the two dropout +
// residual sites of a training-mode encoder layer fused into the LayerNorm that follows them (same arithmetic, same order:
// bit-identical to ocv_dropout_f32 followed by ocv_add_layernorm_f32); `sum` is kept for the backward.
__global__ void dropout_add_layernorm_kernel(const float* x /* may alias sum */, const float* __restrict__ r, const float* __restrict__ gamma,
                                             const float* __restrict__ beta, float* sum, float* __restrict__ y, int rows,
                                             int E, uint32_t thr, float scale, unsigned long long seed_in,
                                             const unsigned long long* salt) {
  const unsigned long long seed = ocv_salted(seed_in, salt);
  const int row = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x % 32;
  if (row >= rows) return;
  const int per = E / 32;
  float v[16];
  const long long base = (long long)row * E;
  float s = 0.f;
  for (int i = 0; i < per; ++i) {
    const int e = i * 32 + lane;
    const float d = ocv_drop_keep(seed, (unsigned long long)(base + e), thr) ? x[base + e] * scale : 0.f;
    v[i] = r[base + e] + d;
    sum[base + e] = v[i];
    s += v[i];
  }
  const float mean = warp_sum(s) / E;
  float q = 0.f;
  for (int i = 0; i < per; ++i) {
    const float d = v[i] - mean;
    q += d * d;
  }
  const float rstd = rsqrtf(warp_sum(q) / E + 1e-5f);
  for (int i = 0; i < per; ++i) {
    const int e = i * 32 + lane;
    y[base + e] = (v[i] - mean) * rstd * gamma[e] + beta[e];
  }
}

// ------------------------------------------------------------------------------------------------
// attention core, d = 32: one warp per query row, keys in tiles of 32 (lane = key for the scores,
// lane = channel for P@V), online softmax in fp32
// ------------------------------------------------------------------------------------------------
constexpr int ATT_WARPS = 8;
constexpr int ATT_KT = 32;

__global__ void __launch_bounds__(ATT_WARPS * 32)
mha_core_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                float* __restrict__ out, const uint8_t* __restrict__ mask, int Sq, int Sk,
                long long q_bs, int q_rs, long long k_bs, int k_rs, long long v_bs, int v_rs, long long o_bs,
                int o_rs, float drop_p, unsigned long long drop_seed_in, const unsigned long long* drop_salt) {
  const unsigned long long drop_seed = ocv_salted(drop_seed_in, drop_salt);
  __shared__ float Ks[ATT_KT][33];
  __shared__ float Vs[ATT_KT][32];
  __shared__ float Msk[ATT_KT];
  const int b = blockIdx.z, hd = blockIdx.y;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int qi = blockIdx.x * ATT_WARPS + warp;
  const bool active = qi < Sq;
  const float scale = rsqrtf(32.f);
  // every lane keeps the whole (scaled) query vector: broadcast through shuffles of the lane-owned element
  float qown = active ? q[b * q_bs + (long long)qi * q_rs + hd * 32 + lane] * scale : 0.f;
  float m = -INFINITY, l = 0.f, acc = 0.f;
  // attention-probability dropout (training): mask index = flat index into [B, heads, Sq, Sk]
  const uint32_t drop_thr = ocv_drop_threshold(drop_p);
  const float drop_scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  const unsigned long long drop_row = (((unsigned long long)b * gridDim.y + hd) * Sq + (active ? qi : 0)) * Sk;
  for (int k0 = 0; k0 < Sk; k0 += ATT_KT) {
    __syncthreads();
    for (int i = threadIdx.x; i < ATT_KT * 32; i += blockDim.x) {
      const int j = i / 32, c = i % 32;
      const bool ok = (k0 + j) < Sk;
      Ks[j][c] = ok ? k[b * k_bs + (long long)(k0 + j) * k_rs + hd * 32 + c] : 0.f;
      Vs[j][c] = ok ? v[b * v_bs + (long long)(k0 + j) * v_rs + hd * 32 + c] : 0.f;
    }
    if (threadIdx.x < ATT_KT) {
      const int j = k0 + threadIdx.x;
      Msk[threadIdx.x] = (j >= Sk || (mask && mask[(long long)b * Sk + j])) ? 1.f : 0.f;
    }
    __syncthreads();
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < 32; ++c) s = fmaf(__shfl_sync(0xffffffffu, qown, c), Ks[lane][c], s);
    if (Msk[lane] != 0.f) s = -INFINITY;
    const float m_new = fmaxf(m, warp_max(s));
    float p, corr;
    if (m_new == -INFINITY) {
      p = 0.f;
      corr = 1.f;
    } else {
      p = expf(s - m_new);
      corr = expf(m - m_new);
    }
    l = l * corr + warp_sum(p);
    acc *= corr;
    if (drop_p > 0.f) p = ocv_drop_keep(drop_seed, drop_row + k0 + lane, drop_thr) ? p * drop_scale : 0.f;  // after the normaliser
#pragma unroll
    for (int j = 0; j < 32; ++j) acc = fmaf(__shfl_sync(0xffffffffu, p, j), Vs[j][lane], acc);
    m = m_new;
  }
  if (active) out[b * o_bs + (long long)qi * o_rs + hd * 32 + lane] = acc / l;  // all keys masked -> 0/0 = NaN (torch)
}

// ------------------------------------------------------------------------------------------------
// ragged objects -> padded tokens + key-padding mask
// ------------------------------------------------------------------------------------------------
__global__ void pad_objects_kernel(const float* __restrict__ src, const int32_t* __restrict__ offsets,
                                   const int32_t* __restrict__ mask_offsets, float* __restrict__ dst,
                                   uint8_t* __restrict__ mask, int L, int Nmax, int E, float pad_value, int left_pad,
                                   __nv_bfloat16* __restrict__ dst16) {
  pdl_launch();
  pdl_wait();
  const int b = blockIdx.y;
  const int row = blockIdx.x;  // 0..L-1
  const int beg = offsets[b], n = offsets[b + 1] - beg;
  const int left = left_pad ? (L - Nmax) : 0;
  const int j = row - left;
  const bool real = (j >= 0 && j < n);
  float* d = dst + ((long long)b * L + row) * E;
  const float* s = real ? src + (long long)(beg + j) * E : nullptr;
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    const float v = real ? s[e] : pad_value;
    d[e] = v;
    if (dst16) dst16[((long long)b * L + row) * E + e] = __float2bfloat16(v);
  }
  if (mask && threadIdx.x == 0) {
    const int m = mask_offsets ? mask_offsets[b + 1] - mask_offsets[b] : n;
    mask[(long long)b * L + row] = (row < m) ? 0 : 1;
  }
}

__global__ void patch_coords_kernel(float* __restrict__ coords, int gh, int gw, int P) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= gh * gw) return;
  const int y = s / gw, x = s - y * gw;
  float4 c = make_float4((float)(x * P + P / 2), (float)(y * P + P / 2), (float)P, (float)P);
  reinterpret_cast<float4*>(coords)[s] = c;
}

// ------------------------------------------------------------------------------------------------
// F.grid_sample(bilinear, zeros, align_corners=False) on the positional table: one warp per point.
// Explicit _rn ops (no FMA contraction) so that the arithmetic order is the oracle's.
// ------------------------------------------------------------------------------------------------
__global__ void grid_sample_pos_kernel(const float* __restrict__ table, const float* __restrict__ pts,
                                       float* __restrict__ out, int32_t* __restrict__ idx_out, int N, int gh, int gw,
                                       int E, int norm_mode, float div0, float div1) {
  const int n = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x % 32;
  if (n >= N) return;
  float x = pts[2 * n], y = pts[2 * n + 1];
  if (norm_mode == 1) {  // ObjCAViT.py:104-105
    x = __fsub_rn(__fmul_rn(__fdiv_rn(x, div0), 2.f), 1.f);
    y = __fsub_rn(__fmul_rn(__fdiv_rn(y, div1), 2.f), 1.f);
  } else if (norm_mode == 2 && n < 2) {  // ObjCAViT.py:95-96 index the token axis (F13a)
    const float d = n == 0 ? div0 : div1;
    x = __fsub_rn(__fmul_rn(__fdiv_rn(x, d), 2.f), 1.f);
    y = __fsub_rn(__fmul_rn(__fdiv_rn(y, d), 2.f), 1.f);
  }
  const float ix = __fsub_rn(__fmul_rn(__fadd_rn(x, 1.f), (float)gw / 2.f), 0.5f);
  const float iy = __fsub_rn(__fmul_rn(__fadd_rn(y, 1.f), (float)gh / 2.f), 0.5f);
  const float x0 = floorf(ix), y0 = floorf(iy);
  const float x1 = x0 + 1.f, y1 = y0 + 1.f;
  const float wts[4] = {__fmul_rn(__fsub_rn(x1, ix), __fsub_rn(y1, iy)), __fmul_rn(__fsub_rn(ix, x0), __fsub_rn(y1, iy)),
                        __fmul_rn(__fsub_rn(x1, ix), __fsub_rn(iy, y0)), __fmul_rn(__fsub_rn(ix, x0), __fsub_rn(iy, y0))};
  const float cx[4] = {x0, x1, x0, x1}, cy[4] = {y0, y0, y1, y1};
  int idx[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    // NaN / huge coordinates compare false -> out of bounds
    const bool ok = (cx[j] >= 0.f) && (cx[j] < (float)gw) && (cy[j] >= 0.f) && (cy[j] < (float)gh);
    idx[j] = ok ? (int)cy[j] * gw + (int)cx[j] : -1;
  }
  if (idx_out && lane < 4) idx_out[4 * n + lane] = idx[lane];
  for (int e = lane; e < E; e += 32) {
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (idx[j] >= 0) acc = __fadd_rn(acc, __fmul_rn(table[(long long)idx[j] * E + e], wts[j]));
    out[(long long)n * E + e] = acc;
  }
}

// ------------------------------------------------------------------------------------------------
// torchvision ps_roi_align, output 1x1, sampling_ratio = -1 (adaptive), one warp per box.
// ------------------------------------------------------------------------------------------------
__global__ void ps_roi_align_pos_kernel(const float* __restrict__ table, const float* __restrict__ boxes,
                                        float* __restrict__ out, int32_t* __restrict__ plan_out, int N, int gh, int gw,
                                        int E, float scale, int is_xywh) {
  const int n = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x % 32;
  if (n >= N) return;
  float bx1 = boxes[4 * n + 0], by1 = boxes[4 * n + 1], bx2 = boxes[4 * n + 2], by2 = boxes[4 * n + 3];
  if (is_xywh) {  // ObjCAViT.py:115-124: centre/size -> corners, clamp(min=0)
    const float hw = __fdiv_rn(bx2, 2.f), hh = __fdiv_rn(by2, 2.f);
    const float cx = bx1, cy = by1;
    bx1 = fmaxf(__fsub_rn(cx, hw), 0.f);
    by1 = fmaxf(__fsub_rn(cy, hh), 0.f);
    bx2 = fmaxf(__fadd_rn(cx, hw), 0.f);
    by2 = fmaxf(__fadd_rn(cy, hh), 0.f);
  }
  const float rsw = __fsub_rn(__fmul_rn(bx1, scale), 0.5f);
  const float rsh = __fsub_rn(__fmul_rn(by1, scale), 0.5f);
  const float rew = __fsub_rn(__fmul_rn(bx2, scale), 0.5f);
  const float reh = __fsub_rn(__fmul_rn(by2, scale), 0.5f);
  const float rw = __fsub_rn(rew, rsw), rh = __fsub_rn(reh, rsh);
  const int grid_h = (int)ceilf(rh), grid_w = (int)ceilf(rw);
  if (plan_out && lane == 0) {
    plan_out[2 * n] = grid_h;
    plan_out[2 * n + 1] = grid_w;
  }
  const float count = (float)(grid_h * grid_w);
  const int per = (E + 31) / 32;
  float acc[8];
  for (int i = 0; i < per; ++i) acc[i] = 0.f;
  const float H = (float)gh, W = (float)gw;
  for (int iy = 0; iy < grid_h; ++iy) {
    const float y = __fadd_rn(rsh, __fdiv_rn(__fmul_rn((float)iy + 0.5f, rh), (float)grid_h));
    for (int ix = 0; ix < grid_w; ++ix) {
      const float x = __fadd_rn(rsw, __fdiv_rn(__fmul_rn((float)ix + 0.5f, rw), (float)grid_w));
      if (y < -1.f || y > H || x < -1.f || x > W) continue;
      float yy = y <= 0.f ? 0.f : y, xx = x <= 0.f ? 0.f : x;
      int y_low = (int)yy, x_low = (int)xx, y_high, x_high;
      if (y_low >= gh - 1) { y_high = y_low = gh - 1; yy = (float)y_low; } else y_high = y_low + 1;
      if (x_low >= gw - 1) { x_high = x_low = gw - 1; xx = (float)x_low; } else x_high = x_low + 1;
      const float ly = __fsub_rn(yy, (float)y_low), lx = __fsub_rn(xx, (float)x_low);
      const float hy = __fsub_rn(1.f, ly), hx = __fsub_rn(1.f, lx);
      const float w1 = __fmul_rn(hy, hx), w2 = __fmul_rn(hy, lx), w3 = __fmul_rn(ly, hx), w4 = __fmul_rn(ly, lx);
      const float* p1 = table + (long long)(y_low * gw + x_low) * E;
      const float* p2 = table + (long long)(y_low * gw + x_high) * E;
      const float* p3 = table + (long long)(y_high * gw + x_low) * E;
      const float* p4 = table + (long long)(y_high * gw + x_high) * E;
      for (int i = 0; i < per; ++i) {
        const int e = i * 32 + lane;
        if (e < E) {
          float val = __fmul_rn(w1, p1[e]);
          val = __fadd_rn(val, __fmul_rn(w2, p2[e]));
          val = __fadd_rn(val, __fmul_rn(w3, p3[e]));
          val = __fadd_rn(val, __fmul_rn(w4, p4[e]));
          acc[i] = __fadd_rn(acc[i], val);
        }
      }
    }
  }
  for (int i = 0; i < per; ++i) {
    const int e = i * 32 + lane;
    if (e < E) out[(long long)n * E + e] = __fdiv_rn(acc[i], count);  // count == 0 -> 0/0 = NaN (torchvision)
  }
}

// ------------------------------------------------------------------------------------------------
// Backward of the two table lookups w.r.t. the positional table (training with the grid_random* strategies): the same index /
// weight arithmetic as the forward kernels, scattering  w * dout[n,e]  into the table cells.  The scatter target is a 64-bit
// FIXED-POINT accumulator (2^40 per unit): integer atomics commute, so the result does not depend on the order in which warps
// arrive (deterministic), and its resolution (1e-12) is far below fp32.  ocv_table_grad_commit_f32 converts and accumulates it
// into the fp32 gradient.  Coordinates / boxes are inputs of the head: they receive no gradient (as in the reference).
// ------------------------------------------------------------------------------------------------
constexpr double TG_FIX = 1099511627776.0;  // 2^40

__device__ __forceinline__ void table_scatter(long long* __restrict__ fix, int cell, int e, int E, float v) {
  atomicAdd(reinterpret_cast<unsigned long long*>(fix + (long long)cell * E + e), (unsigned long long)__double2ll_rn((double)v * TG_FIX));
}

__global__ void grid_sample_pos_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ pts, long long* __restrict__ fix,
                                           int N, int gh, int gw, int E, int norm_mode, float div0, float div1) {
  const int n = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x % 32;
  if (n >= N) return;
  float x = pts[2 * n], y = pts[2 * n + 1];
  if (norm_mode == 1) {
    x = __fsub_rn(__fmul_rn(__fdiv_rn(x, div0), 2.f), 1.f);
    y = __fsub_rn(__fmul_rn(__fdiv_rn(y, div1), 2.f), 1.f);
  } else if (norm_mode == 2 && n < 2) {
    const float d = n == 0 ? div0 : div1;
    x = __fsub_rn(__fmul_rn(__fdiv_rn(x, d), 2.f), 1.f);
    y = __fsub_rn(__fmul_rn(__fdiv_rn(y, d), 2.f), 1.f);
  }
  const float ix = __fsub_rn(__fmul_rn(__fadd_rn(x, 1.f), (float)gw / 2.f), 0.5f);
  const float iy = __fsub_rn(__fmul_rn(__fadd_rn(y, 1.f), (float)gh / 2.f), 0.5f);
  const float x0 = floorf(ix), y0 = floorf(iy);
  const float x1 = x0 + 1.f, y1 = y0 + 1.f;
  const float wts[4] = {__fmul_rn(__fsub_rn(x1, ix), __fsub_rn(y1, iy)), __fmul_rn(__fsub_rn(ix, x0), __fsub_rn(y1, iy)),
                        __fmul_rn(__fsub_rn(x1, ix), __fsub_rn(iy, y0)), __fmul_rn(__fsub_rn(ix, x0), __fsub_rn(iy, y0))};
  const float cx[4] = {x0, x1, x0, x1}, cy[4] = {y0, y0, y1, y1};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool ok = (cx[j] >= 0.f) && (cx[j] < (float)gw) && (cy[j] >= 0.f) && (cy[j] < (float)gh);
    if (!ok) continue;
    const int cell = (int)cy[j] * gw + (int)cx[j];
    for (int e = lane; e < E; e += 32) table_scatter(fix, cell, e, E, wts[j] * dout[(long long)n * E + e]);
  }
}

__global__ void ps_roi_align_pos_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ boxes, long long* __restrict__ fix,
                                            int N, int gh, int gw, int E, float scale, int is_xywh) {
  const int n = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x % 32;
  if (n >= N) return;
  float bx1 = boxes[4 * n + 0], by1 = boxes[4 * n + 1], bx2 = boxes[4 * n + 2], by2 = boxes[4 * n + 3];
  if (is_xywh) {
    const float hw = __fdiv_rn(bx2, 2.f), hh = __fdiv_rn(by2, 2.f);
    const float cx = bx1, cy = by1;
    bx1 = fmaxf(__fsub_rn(cx, hw), 0.f);
    by1 = fmaxf(__fsub_rn(cy, hh), 0.f);
    bx2 = fmaxf(__fadd_rn(cx, hw), 0.f);
    by2 = fmaxf(__fadd_rn(cy, hh), 0.f);
  }
  const float rsw = __fsub_rn(__fmul_rn(bx1, scale), 0.5f);
  const float rsh = __fsub_rn(__fmul_rn(by1, scale), 0.5f);
  const float rew = __fsub_rn(__fmul_rn(bx2, scale), 0.5f);
  const float reh = __fsub_rn(__fmul_rn(by2, scale), 0.5f);
  const float rw = __fsub_rn(rew, rsw), rh = __fsub_rn(reh, rsh);
  const int grid_h = (int)ceilf(rh), grid_w = (int)ceilf(rw);
  if (grid_h <= 0 || grid_w <= 0) return;  // forward value is 0/0 = NaN there: nothing meaningful to propagate
  const float inv_count = 1.f / (float)(grid_h * grid_w);
  const float H = (float)gh, W = (float)gw;
  for (int iy = 0; iy < grid_h; ++iy) {
    const float y = __fadd_rn(rsh, __fdiv_rn(__fmul_rn((float)iy + 0.5f, rh), (float)grid_h));
    for (int ix = 0; ix < grid_w; ++ix) {
      const float x = __fadd_rn(rsw, __fdiv_rn(__fmul_rn((float)ix + 0.5f, rw), (float)grid_w));
      if (y < -1.f || y > H || x < -1.f || x > W) continue;
      float yy = y <= 0.f ? 0.f : y, xx = x <= 0.f ? 0.f : x;
      int y_low = (int)yy, x_low = (int)xx, y_high, x_high;
      if (y_low >= gh - 1) { y_high = y_low = gh - 1; yy = (float)y_low; } else y_high = y_low + 1;
      if (x_low >= gw - 1) { x_high = x_low = gw - 1; xx = (float)x_low; } else x_high = x_low + 1;
      const float ly = __fsub_rn(yy, (float)y_low), lx = __fsub_rn(xx, (float)x_low);
      const float hy = __fsub_rn(1.f, ly), hx = __fsub_rn(1.f, lx);
      const float w4[4] = {__fmul_rn(hy, hx), __fmul_rn(hy, lx), __fmul_rn(ly, hx), __fmul_rn(ly, lx)};
      const int cell[4] = {y_low * gw + x_low, y_low * gw + x_high, y_high * gw + x_low, y_high * gw + x_high};
      for (int e = lane; e < E; e += 32) {
        const float g = dout[(long long)n * E + e] * inv_count;
#pragma unroll
        for (int j = 0; j < 4; ++j) table_scatter(fix, cell[j], e, E, w4[j] * g);
      }
    }
  }
}

__global__ void table_grad_commit_kernel(const long long* __restrict__ fix, float* __restrict__ dtable, long long n, float beta) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float v = (float)((double)fix[i] / TG_FIX);
  dtable[i] = beta != 0.f ? beta * dtable[i] + v : v;
}

// ------------------------------------------------------------------------------------------------
// bins: one block per image
// ------------------------------------------------------------------------------------------------
__global__ void bins_kernel(const float* __restrict__ y_raw, float* __restrict__ widths, float* __restrict__ edges,
                            float* __restrict__ centres, int K, int norm, float min_depth, float max_depth) {
  __shared__ float sv[1024];
  __shared__ float red[32];
  __shared__ float bcast;
  pdl_launch();
  pdl_wait();
  const int b = blockIdx.x, tid = threadIdx.x;
  const float* yr = y_raw + (long long)b * K;
  const int warp = tid / 32, lane = tid % 32, nwarp = blockDim.x / 32;
  float mx = -INFINITY;
  if (norm == OCV_NORM_SOFTMAX) {
    for (int k = tid; k < K; k += blockDim.x) mx = fmaxf(mx, yr[k]);
    mx = warp_max(mx);
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    if (tid == 0) {
      float m = red[0];
      for (int i = 1; i < nwarp; ++i) m = fmaxf(m, red[i]);
      bcast = m;
    }
    __syncthreads();
    mx = bcast;
    __syncthreads();
  }
  float s = 0.f;
  for (int k = tid; k < K; k += blockDim.x) {
    float y = yr[k];
    if (norm == OCV_NORM_LINEAR) y = fmaxf(y, 0.f) + 0.1f;
    else if (norm == OCV_NORM_SOFTMAX) y = expf(y - mx);
    else if (norm == OCV_NORM_SIGMOID) y = 1.f / (1.f + expf(-y));
    sv[k] = y;
    s += y;
  }
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int i = 0; i < nwarp; ++i) t += red[i];
    bcast = t;
  }
  __syncthreads();
  const float total = (norm == OCV_NORM_NONE) ? 1.f : bcast;
  for (int k = tid; k < K; k += blockDim.x) {
    const float wn = sv[k] / total;
    sv[k] = wn;
    if (widths) widths[(long long)b * K + k] = wn;
  }
  __syncthreads();
  if (tid == 0 && edges) {  // sequential cumsum, the order torch.cumsum uses on the CPU
    float e = min_depth;
    float* eb = edges + (long long)b * (K + 1);
    eb[0] = e;
    const float range = max_depth - min_depth;
    for (int k = 0; k < K; ++k) {
      e += range * sv[k];
      eb[k + 1] = e;
    }
  }
  __syncthreads();
  if (centres && edges) {
    const float* eb = edges + (long long)b * (K + 1);
    for (int k = tid; k < K; k += blockDim.x) centres[(long long)b * K + k] = 0.5f * (eb[k] + eb[k + 1]);
  }
}

}  // namespace

extern "C" int ocv_add_layernorm_f32(const float* x, const float* r, const float* gamma, const float* beta, float* y,
                                     int rows, int E, const uint8_t* zero_rows, void* stream) {
  OCV_CHECK_ARG(x && gamma && beta && y && rows > 0);
  OCV_CHECK_SHAPE(E > 0 && E % 32 == 0 && E <= 512);
  if (int s = ocv_require_sm100()) return s;
  const int wpb = 8;
  add_layernorm_kernel<<<ceil_div(rows, wpb), wpb * 32, 0, ocv_stream(stream)>>>(x, r, gamma, beta, y, rows, E,
                                                                                zero_rows);
  return ocv_launch_status();
}

extern "C" int ocv_dropout_add_layernorm_f32(const float* x, const float* residual, const float* gamma, const float* beta, float* sum,
                                             float* y, int rows, int E, float p, unsigned long long seed, void* stream) {
  OCV_CHECK_ARG(x && residual && gamma && beta && sum && y && rows > 0 && p >= 0.f && p < 1.f);
  OCV_CHECK_SHAPE(E > 0 && E % 32 == 0 && E <= 512);
  if (int s = ocv_require_sm100()) return s;
  const int wpb = 8;
  dropout_add_layernorm_kernel<<<ceil_div(rows, wpb), wpb * 32, 0, ocv_stream(stream)>>>(
      x, residual, gamma, beta, sum, y, rows, E, ocv_drop_threshold(p), 1.f / (1.f - p), seed, ocv_dropout_salt());
  return ocv_launch_status();
}

extern "C" int ocv_mha_core_f32(const float* q, const float* k, const float* v, float* out,
                                const uint8_t* key_padding_mask, int B, int Sq, int Sk, int heads, int d,
                                long long q_bstride, int q_rstride, long long k_bstride, int k_rstride,
                                long long v_bstride, int v_rstride, long long o_bstride, int o_rstride,
                                void* stream) {
  return ocv_mha_core_dropout_f32(q, k, v, out, key_padding_mask, B, Sq, Sk, heads, d, q_bstride, q_rstride, k_bstride, k_rstride,
                                  v_bstride, v_rstride, o_bstride, o_rstride, 0.f, 0ull, stream);
}

extern "C" int ocv_mha_core_dropout_f32(const float* q, const float* k, const float* v, float* out,
                                        const uint8_t* key_padding_mask, int B, int Sq, int Sk, int heads, int d,
                                        long long q_bstride, int q_rstride, long long k_bstride, int k_rstride,
                                        long long v_bstride, int v_rstride, long long o_bstride, int o_rstride,
                                        float drop_p, unsigned long long drop_seed, void* stream) {
  OCV_CHECK_ARG(q && k && v && out && B > 0 && Sq > 0 && Sk > 0 && heads > 0 && drop_p >= 0.f && drop_p < 1.f);
  OCV_CHECK_SHAPE(d == 32 && B <= 65535 && heads <= 65535);
  if (int s = ocv_require_sm100()) return s;
  dim3 grid(ceil_div(Sq, ATT_WARPS), heads, B);
  mha_core_kernel<<<grid, ATT_WARPS * 32, 0, ocv_stream(stream)>>>(q, k, v, out, key_padding_mask, Sq, Sk, q_bstride,
                                                                  q_rstride, k_bstride, k_rstride, v_bstride,
                                                                  v_rstride, o_bstride, o_rstride, drop_p, drop_seed, ocv_dropout_salt());
  return ocv_launch_status();
}

int ocv_pad_objects_impl(const float* src, const int32_t* offsets, const int32_t* mask_offsets, float* dst, void* dst16,
                         uint8_t* mask, int B, int L, int Nmax, int E, float pad_value, int left_pad, void* stream) {
  OCV_CHECK_ARG(src && offsets && dst && B > 0 && L > 0 && E > 0);
  OCV_CHECK_SHAPE(Nmax > 0 && Nmax <= L && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  return ocv_launch_pdl(pad_objects_kernel, dim3(L, B), dim3(128), 0, stream, src, offsets, mask_offsets, dst, mask, L, Nmax,
                        E, pad_value, left_pad, static_cast<__nv_bfloat16*>(dst16));
}

extern "C" int ocv_pad_objects_f32(const float* src, const int32_t* offsets, const int32_t* mask_offsets, float* dst,
                                   uint8_t* mask, int B, int L, int Nmax, int E, float pad_value, int left_pad,
                                   void* stream) {
  return ocv_pad_objects_impl(src, offsets, mask_offsets, dst, nullptr, mask, B, L, Nmax, E, pad_value, left_pad, stream);
}

extern "C" int ocv_patch_coords_f32(float* coords, int gh, int gw, int P, void* stream) {
  OCV_CHECK_ARG(coords && gh > 0 && gw > 0 && P > 0);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(coords) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  patch_coords_kernel<<<ceil_div(gh * gw, 128), 128, 0, ocv_stream(stream)>>>(coords, gh, gw, P);
  return ocv_launch_status();
}

extern "C" int ocv_grid_sample_pos_f32(const float* table, const float* pts, float* out, int32_t* idx_out, int N,
                                       int gh, int gw, int E, int norm_mode, float div0, float div1, void* stream) {
  OCV_CHECK_ARG(table && pts && out && N > 0 && gh > 0 && gw > 0 && E > 0);
  OCV_CHECK_ARG(norm_mode >= 0 && norm_mode <= 2);
  if (int s = ocv_require_sm100()) return s;
  grid_sample_pos_kernel<<<ceil_div(N, 4), 128, 0, ocv_stream(stream)>>>(table, pts, out, idx_out, N, gh, gw, E,
                                                                        norm_mode, div0, div1);
  return ocv_launch_status();
}

extern "C" int ocv_ps_roi_align_pos_f32(const float* table, const float* boxes, float* out, int32_t* plan_out, int N,
                                        int gh, int gw, int E, float spatial_scale, int boxes_are_xywh, void* stream) {
  OCV_CHECK_ARG(table && boxes && out && N > 0 && gh > 0 && gw > 0);
  OCV_CHECK_SHAPE(E > 0 && E <= 256);
  if (int s = ocv_require_sm100()) return s;
  ps_roi_align_pos_kernel<<<ceil_div(N, 4), 128, 0, ocv_stream(stream)>>>(table, boxes, out, plan_out, N, gh, gw, E,
                                                                         spatial_scale, boxes_are_xywh);
  return ocv_launch_status();
}

extern "C" int ocv_bins_f32(const float* y_raw, float* widths_normed, float* edges, float* centres, int B, int K,
                            int norm, float min_depth, float max_depth, void* stream) {
  OCV_CHECK_ARG(y_raw && B > 0);
  OCV_CHECK_SHAPE(K > 0 && K <= 1024 && norm >= 0 && norm <= 3);
  OCV_CHECK_ARG(!(centres && !edges));
  if (int s = ocv_require_sm100()) return s;
  return ocv_launch_pdl(bins_kernel, dim3(B), dim3(256), 0, stream, y_raw, widths_normed, edges, centres, K, norm,
                        min_depth, max_depth);
}

extern "C" int ocv_grid_sample_pos_bwd_f32(const float* dout, const float* pts, long long* table_grad_fix, int N, int gh, int gw, int E,
                                           int norm_mode, float div0, float div1, void* stream) {
  OCV_CHECK_ARG(dout && pts && table_grad_fix && N > 0 && gh > 0 && gw > 0 && E > 0 && norm_mode >= 0 && norm_mode <= 2);
  if (int s = ocv_require_sm100()) return s;
  grid_sample_pos_bwd_kernel<<<ceil_div(N, 4), 128, 0, ocv_stream(stream)>>>(dout, pts, table_grad_fix, N, gh, gw, E, norm_mode, div0,
                                                                            div1);
  return ocv_launch_status();
}

extern "C" int ocv_ps_roi_align_pos_bwd_f32(const float* dout, const float* boxes, long long* table_grad_fix, int N, int gh, int gw,
                                            int E, float spatial_scale, int boxes_are_xywh, void* stream) {
  OCV_CHECK_ARG(dout && boxes && table_grad_fix && N > 0 && gh > 0 && gw > 0 && E > 0);
  if (int s = ocv_require_sm100()) return s;
  ps_roi_align_pos_bwd_kernel<<<ceil_div(N, 4), 128, 0, ocv_stream(stream)>>>(dout, boxes, table_grad_fix, N, gh, gw, E,
                                                                             spatial_scale, boxes_are_xywh);
  return ocv_launch_status();
}

extern "C" int ocv_table_grad_commit_f32(const long long* table_grad_fix, float* table_grad, long long n, float beta, void* stream) {
  OCV_CHECK_ARG(table_grad_fix && table_grad && n > 0);
  if (int s = ocv_require_sm100()) return s;
  table_grad_commit_kernel<<<(unsigned)ceil_div<long long>(n, 256), 256, 0, ocv_stream(stream)>>>(table_grad_fix, table_grad, n, beta);
  return ocv_launch_status();
}
