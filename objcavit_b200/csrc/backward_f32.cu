// Backward kernels of the fp32 parity path that are not GEMMs (head forward + backward):
// channel reductions (bias gradients), LayerNorm backward, activation backward, attention-core backward, the backward of
// the softmax-over-bins expectation (head tail) and of the bin-width normalisation.  All reductions are fixed-order
// (deterministic); gradients are written as  out = beta * out + value  so that parameter gradients can accumulate into
// the flat gradient buffer.  Contract: include/objcavit_b200.h.
#include "common.cuh"
#include <math.h>
#include <mutex>

namespace {

__device__ __forceinline__ float block_sum_f(float v, float* sh) {  // result on every thread; sh >= 32 floats
  v = warp_sum(v);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32, nwarp = (blockDim.x + 31) / 32;
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float t = 0.f;
  for (int i = 0; i < nwarp; ++i) t += sh[i];
  return t;
}

// x viewed [G, outer, C, inner] -> out[g, c] = beta * out[g, c] + sum_{o, i} x[g, o, c, i];  one block per (c, g)
__global__ void __launch_bounds__(512) reduce_channels_kernel(const float* __restrict__ x, float* __restrict__ out, int outer, int C,
                                                              long long inner, float beta) {
  __shared__ float sh[32];
  const int c = blockIdx.x, g = blockIdx.y;
  const float* xg = x + (long long)g * outer * C * inner;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  for (int o = 0; o < outer; ++o) {  // contiguous runs of `inner` floats: no per-element index arithmetic, 128-bit loads when aligned
    const float* run = xg + ((long long)o * C + c) * inner;
    if ((inner & 3) == 0 && (reinterpret_cast<uintptr_t>(run) & 15) == 0) {
      const float4* r4 = reinterpret_cast<const float4*>(run);
      for (long long i = threadIdx.x; i < inner / 4; i += blockDim.x) {
        const float4 v = r4[i];
        a0 += v.x; a1 += v.y; a2 += v.z; a3 += v.w;
      }
    } else {
      for (long long i = threadIdx.x; i < inner; i += blockDim.x) a0 += run[i];
    }
  }
  const float s = block_sum_f((a0 + a1) + (a2 + a3), sh);
  if (threadIdx.x == 0) {
    float* o = out + (long long)g * C + c;
    *o = beta != 0.f ? beta * (*o) + s : s;
  }
}

// inner == 1 (column sums of a [G, rows, C] matrix: Linear bias gradients, positional-row gradients): threads along C so that every
// load is coalesced; 32 row lanes per block combined through shared memory in a fixed order
__global__ void __launch_bounds__(1024) reduce_columns_kernel(const float* __restrict__ x, float* __restrict__ out, int rows, int C,
                                                              float beta) {
  __shared__ float sh[32][33];
  const int c = blockIdx.x * 32 + threadIdx.x, g = blockIdx.y, ry = threadIdx.y;
  const float* xg = x + (long long)g * rows * C;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  if (c < C) {
    int r = ry;
    for (; r + 96 < rows; r += 128) {
      a0 += xg[(long long)r * C + c];
      a1 += xg[(long long)(r + 32) * C + c];
      a2 += xg[(long long)(r + 64) * C + c];
      a3 += xg[(long long)(r + 96) * C + c];
    }
    for (; r < rows; r += 32) a0 += xg[(long long)r * C + c];
  }
  sh[ry][threadIdx.x] = (a0 + a1) + (a2 + a3);
  __syncthreads();
  if (ry == 0 && c < C) {
    float t = 0.f;
    for (int i = 0; i < 32; ++i) t += sh[i][threadIdx.x];
    float* o = out + (long long)g * C + c;
    *o = beta != 0.f ? beta * (*o) + t : t;
  }
}

// dx = dy * act'(.) using the activation OUTPUT (sign-preserving activations)
__global__ void act_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ out, float* __restrict__ dx, long long n,
                               int act) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float o = out[i], g = dy[i];
  dx[i] = act == OCV_ACT_RELU ? (o > 0.f ? g : 0.f) : (act == OCV_ACT_LEAKY_RELU ? (o > 0.f ? g : 0.01f * g) : g);
}

// dx = act'(out) * dropout(dy): the FFN dropout site and the ReLU backward of a training-mode encoder layer in one pass (same
// arithmetic as ocv_dropout_f32 followed by ocv_act_bwd_f32: bit-identical)
__global__ void dropout_act_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ out, float* __restrict__ dx, long long n,
                                       int act, uint32_t thr, float scale, unsigned long long seed_in,
                                       const unsigned long long* salt) {
  const unsigned long long seed = ocv_salted(seed_in, salt);
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float o = out[i];
  const float g = ocv_drop_keep(seed, (unsigned long long)i, thr) ? dy[i] * scale : 0.f;
  dx[i] = act == OCV_ACT_RELU ? (o > 0.f ? g : 0.f) : (act == OCV_ACT_LEAKY_RELU ? (o > 0.f ? g : 0.01f * g) : g);
}

// LayerNorm backward, one warp per row (E <= 512, E % 32 == 0): z = pre-norm input, y = (z - mean) rstd gamma + beta
//   dz = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * gamma;  also stores (mean, rstd) per row for the parameter pass
__global__ void layernorm_bwd_rows_kernel(const float* __restrict__ dy, const float* __restrict__ z, const float* __restrict__ gamma,
                                          float* __restrict__ dz, float* __restrict__ stats, int rows, int E,
                                          float* __restrict__ dz_dropped, uint32_t thr, float scale, unsigned long long seed_in,
                                          const unsigned long long* salt) {
  // dz_dropped (optional) = dropout(dz) over the flat [rows, E] index, as ocv_dropout_f32 would compute it from dz
  const unsigned long long seed = ocv_salted(seed_in, salt);
  const int row = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x % 32;
  if (row >= rows) return;
  const int per = E / 32;
  float v[16], g[16];
  const float* zr = z + (long long)row * E;
  const float* dr = dy + (long long)row * E;
  float s = 0.f;
  for (int i = 0; i < per; ++i) {
    v[i] = zr[i * 32 + lane];
    s += v[i];
  }
  const float mean = warp_sum(s) / E;
  float q = 0.f;
  for (int i = 0; i < per; ++i) {
    v[i] -= mean;
    q += v[i] * v[i];
  }
  const float rstd = rsqrtf(warp_sum(q) / E + 1e-5f);
  float sg = 0.f, sgx = 0.f;
  for (int i = 0; i < per; ++i) {
    v[i] *= rstd;  // xhat
    g[i] = dr[i * 32 + lane] * gamma[i * 32 + lane];
    sg += g[i];
    sgx += g[i] * v[i];
  }
  const float mg = warp_sum(sg) / E, mgx = warp_sum(sgx) / E;
  float* o = dz + (long long)row * E;
  float* od = dz_dropped ? dz_dropped + (long long)row * E : nullptr;
  for (int i = 0; i < per; ++i) {
    const float d = rstd * (g[i] - mg - v[i] * mgx);
    o[i * 32 + lane] = d;
    if (od) od[i * 32 + lane] = ocv_drop_keep(seed, (unsigned long long)((long long)row * E + i * 32 + lane), thr) ? d * scale : 0.f;
  }
  if (lane == 0) {
    stats[2 * row] = mean;
    stats[2 * row + 1] = rstd;
  }
}

// dgamma[e] = beta*dgamma[e] + sum_rows dy * xhat,  dbeta[e] likewise.  Two fixed-order stages: LNG row groups x (32 channels x
// 8 row lanes) blocks write partials, a second kernel adds the LNG partials (4 blocks alone needed 50 us for 1768 rows).
constexpr int LNG = 16;
__global__ void layernorm_bwd_params_kernel(const float* __restrict__ dy, const float* __restrict__ z, const float* __restrict__ stats,
                                            float* __restrict__ part /* [LNG][2][E] */, int rows, int E) {
  __shared__ float sg[8][33], sb[8][33];
  const int e = blockIdx.x * 32 + threadIdx.x, ry = threadIdx.y;
  const int per = (rows + LNG - 1) / LNG, r0 = blockIdx.y * per, r1 = min(rows, r0 + per);
  float ag = 0.f, ab = 0.f;
  if (e < E)
    for (int r = r0 + ry; r < r1; r += 8) {
      const float d = dy[(long long)r * E + e];
      ag += d * (z[(long long)r * E + e] - stats[2 * r]) * stats[2 * r + 1];
      ab += d;
    }
  sg[ry][threadIdx.x] = ag;
  sb[ry][threadIdx.x] = ab;
  __syncthreads();
  if (ry == 0 && e < E) {
    float tg = 0.f, tb = 0.f;
    for (int i = 0; i < 8; ++i) {
      tg += sg[i][threadIdx.x];
      tb += sb[i][threadIdx.x];
    }
    part[((long long)blockIdx.y * 2) * E + e] = tg;
    part[((long long)blockIdx.y * 2 + 1) * E + e] = tb;
  }
}

__global__ void layernorm_bwd_params_final_kernel(const float* __restrict__ part, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                  int E, float beta) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  float tg = 0.f, tb = 0.f;
#pragma unroll
  for (int i = 0; i < LNG; ++i) {
    tg += part[((long long)i * 2) * E + e];
    tb += part[((long long)i * 2 + 1) * E + e];
  }
  dgamma[e] = beta != 0.f ? beta * dgamma[e] + tg : tg;
  dbeta[e] = beta != 0.f ? beta * dbeta[e] + tb : tb;
}

// ---- attention core backward (head dim 32) ------------------------------------------------------------------------------
// Thread-per-row kernels: a thread owns one query (key) row in registers and walks the key (query) rows staged in shared memory,
// so the 32-wide dot products are register FMAs instead of warp reductions.  Four threads share a row (each takes every fourth
// staged row) and are combined by two shuffles; blocks of 32 rows x 4 slices.  Staged rows are padded to 33 floats: the four
// slices of a row read four different banks, the eight rows of a warp read the same address (broadcast).
// (Measured and rejected: 128-bit shared loads — same time, the kernels are bound by the latency of the per-thread FMA
// chains at 6 warps per SM; 8 rows x 16 slices per block — 4x the CTAs each staging the whole key set, 30 % slower.)
// The staged tile is `aj` rows (dynamic shared memory, up to AJ_MAX = 256): at S = 221 the whole key / query set is ONE tile, so
// each CTA pays the global-load latency of staging once instead of eight times (ncu: a third of the stall samples of the
// 64-row-tile version sat on the staging stores), and the second pass of the dq kernel re-uses the resident keys.
constexpr int AR = 32, ASL = 4, AJ_MAX = 256, APAD = 33;

__device__ __forceinline__ void stage_rows(float (*dst)[APAD], const float* __restrict__ src, long long row_stride, int row0, int rows_total,
                                           int aj) {
  for (int idx = threadIdx.x; idx < aj * 32; idx += AR * ASL) {
    const int r = idx >> 5, c = idx & 31;
    dst[r][c] = row0 + r < rows_total ? src[(long long)(row0 + r) * row_stride + c] : 0.f;
  }
}

// pass 1: softmax statistics (m, l), D = dO . O, and dq
__global__ void __launch_bounds__(AR * ASL)
mha_bwd_q_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v, const float* __restrict__ o,
                 const float* __restrict__ dout, const uint8_t* __restrict__ mask, float* __restrict__ dq, float* __restrict__ stats,
                 int Sq, int Sk, int heads, long long q_bs, int q_rs, long long k_bs, int k_rs, long long v_bs, int v_rs,
                 long long o_bs, int o_rs, long long dq_bs, int dq_rs, float drop_p, unsigned long long drop_seed_in,
                 const unsigned long long* drop_salt, int aj) {
  const unsigned long long drop_seed = ocv_salted(drop_seed_in, drop_salt);
  extern __shared__ __align__(16) unsigned char att_smem[];
  float (*Ks)[APAD] = reinterpret_cast<float (*)[APAD]>(att_smem);
  float (*Vs)[APAD] = Ks + aj;
  uint8_t* Ms = reinterpret_cast<uint8_t*>(Vs + aj);
  const bool resident = Sk <= aj;       // one tile: K, V and the mask are staged once for both passes
  const int b = blockIdx.z, hd = blockIdx.y;
  const int r = threadIdx.x / ASL, t = threadIdx.x % ASL;
  const int i = blockIdx.x * AR + r;
  const bool live = i < Sq;
  const float scale = rsqrtf(32.f);
  const uint32_t drop_thr = ocv_drop_threshold(drop_p);
  const float drop_scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  const unsigned long long drop_row = (((unsigned long long)b * heads + hd) * Sq + (live ? i : 0)) * Sk;
  float qi[32], doi[32];
  float D = 0.f;
  {
    const float* qp = q + b * q_bs + (long long)(live ? i : 0) * q_rs + hd * 32;
    const float* dp = dout + b * o_bs + (long long)(live ? i : 0) * o_rs + hd * 32;
    const float* op = o + b * o_bs + (long long)(live ? i : 0) * o_rs + hd * 32;
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      qi[c] = qp[c] * scale;
      doi[c] = dp[c];
      D = fmaf(doi[c], op[c], D);
    }
  }
  const float* kb = k + b * k_bs + hd * 32;
  const float* vb = v + b * v_bs + hd * 32;
  const uint8_t* mb = mask ? mask + (long long)b * Sk : nullptr;
  float m = -INFINITY, l = 0.f;
  for (int j0 = 0; j0 < Sk; j0 += aj) {
    __syncthreads();
    stage_rows(Ks, kb, k_rs, j0, Sk, aj);
    if (resident) stage_rows(Vs, vb, v_rs, j0, Sk, aj);
    for (int x = threadIdx.x; x < aj; x += AR * ASL) Ms[x] = (j0 + x >= Sk) || (mb && mb[j0 + x]);
    __syncthreads();
    for (int jj = t; jj < aj; jj += ASL) {
      if (Ms[jj]) continue;
      float s = 0.f;
#pragma unroll
      for (int c = 0; c < 32; ++c) s = fmaf(qi[c], Ks[jj][c], s);
      const float mn = fmaxf(m, s);
      l = l * __expf(m - mn) + __expf(s - mn);
      m = mn;
    }
  }
#pragma unroll
  for (int off = 1; off < ASL; off <<= 1) {  // combine the four slices of a row
    const float m2 = __shfl_xor_sync(0xffffffffu, m, off), l2 = __shfl_xor_sync(0xffffffffu, l, off);
    const float mn = fmaxf(m, m2);
    l = (m > -INFINITY ? l * __expf(m - mn) : 0.f) + (m2 > -INFINITY ? l2 * __expf(m2 - mn) : 0.f);
    m = mn;
  }
  const float inv_l = l > 0.f ? 1.f / l : 0.f;
  float acc[32];
#pragma unroll
  for (int c = 0; c < 32; ++c) acc[c] = 0.f;
  for (int j0 = 0; j0 < Sk; j0 += aj) {
    if (!resident) {
      __syncthreads();
      stage_rows(Ks, kb, k_rs, j0, Sk, aj);
      stage_rows(Vs, vb, v_rs, j0, Sk, aj);
      for (int x = threadIdx.x; x < aj; x += AR * ASL) Ms[x] = (j0 + x >= Sk) || (mb && mb[j0 + x]);
      __syncthreads();
    }
    for (int jj = t; jj < aj; jj += ASL) {
      if (Ms[jj]) continue;
      float s = 0.f, dP = 0.f;
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        s = fmaf(qi[c], Ks[jj][c], s);
        dP = fmaf(doi[c], Vs[jj][c], dP);
      }
      const float p = __expf(s - m) * inv_l;
      if (drop_p > 0.f) dP = ocv_drop_keep(drop_seed, drop_row + j0 + jj, drop_thr) ? dP * drop_scale : 0.f;
      const float wgt = p * (dP - D);
#pragma unroll
      for (int c = 0; c < 32; ++c) acc[c] = fmaf(wgt, Ks[jj][c], acc[c]);
    }
  }
#pragma unroll
  for (int c = 0; c < 32; ++c) {
    acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], 1);
    acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], 2);
  }
  if (live) {
    float* dst = dq + b * dq_bs + (long long)i * dq_rs + hd * 32;
#pragma unroll
    for (int c = 0; c < 32; ++c)
      if ((c & 3) == t) dst[c] = acc[c] * scale;
    if (t == 0) {
      float* st = stats + (((long long)b * heads + hd) * Sq + i) * 3;
      st[0] = m; st[1] = inv_l; st[2] = D;
    }
  }
}

// pass 2: dk, dv (a thread owns a key row; query rows, dO rows and their statistics are staged)
__global__ void __launch_bounds__(AR * ASL)
mha_bwd_kv_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                  const float* __restrict__ dout, const uint8_t* __restrict__ mask, const float* __restrict__ stats,
                  float* __restrict__ dk, float* __restrict__ dv, int Sq, int Sk, int heads, long long q_bs, int q_rs,
                  long long k_bs, int k_rs, long long v_bs, int v_rs, long long o_bs, int o_rs, long long dk_bs, int dk_rs,
                  long long dv_bs, int dv_rs, float drop_p, unsigned long long drop_seed_in,
                  const unsigned long long* drop_salt, int aj) {
  const unsigned long long drop_seed = ocv_salted(drop_seed_in, drop_salt);
  extern __shared__ __align__(16) unsigned char att_smem[];
  float (*Qs)[APAD] = reinterpret_cast<float (*)[APAD]>(att_smem);
  float (*Ds)[APAD] = Qs + aj;
  float (*St)[3] = reinterpret_cast<float (*)[3]>(Ds + aj);
  const int b = blockIdx.z, hd = blockIdx.y;
  const int r = threadIdx.x / ASL, t = threadIdx.x % ASL;
  const int j = blockIdx.x * AR + r;
  const bool live = j < Sk && !(mask && mask[(long long)b * Sk + j]);
  const float scale = rsqrtf(32.f);
  const uint32_t drop_thr = ocv_drop_threshold(drop_p);
  const float drop_scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  const unsigned long long drop_base = ((unsigned long long)b * heads + hd) * Sq;
  float kj[32], vj[32], ak[32], av[32];
  {
    const float* kp = k + b * k_bs + (long long)(j < Sk ? j : 0) * k_rs + hd * 32;
    const float* vp = v + b * v_bs + (long long)(j < Sk ? j : 0) * v_rs + hd * 32;
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      kj[c] = kp[c] * scale;
      vj[c] = vp[c];
      ak[c] = 0.f;
      av[c] = 0.f;
    }
  }
  const float* qb = q + b * q_bs + hd * 32;
  const float* db = dout + b * o_bs + hd * 32;
  const float* stb = stats + ((long long)b * heads + hd) * Sq * 3;
  for (int i0 = 0; i0 < Sq; i0 += aj) {
    __syncthreads();
    stage_rows(Qs, qb, q_rs, i0, Sq, aj);
    stage_rows(Ds, db, o_rs, i0, Sq, aj);
    for (int idx = threadIdx.x; idx < aj * 3; idx += AR * ASL) St[idx / 3][idx % 3] = i0 + idx / 3 < Sq ? stb[(long long)i0 * 3 + idx] : 0.f;
    __syncthreads();
    if (!live) continue;
    const int n = min(aj, Sq - i0);
    for (int ii = t; ii < n; ii += ASL) {
      float s = 0.f, dP = 0.f;
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        s = fmaf(Qs[ii][c], kj[c], s);
        dP = fmaf(Ds[ii][c], vj[c], dP);
      }
      const float p = __expf(s - St[ii][0]) * St[ii][1];
      float pd = p;
      if (drop_p > 0.f) {
        const float ms = ocv_drop_keep(drop_seed, (drop_base + i0 + ii) * Sk + j, drop_thr) ? drop_scale : 0.f;
        dP *= ms;
        pd *= ms;
      }
      const float wgt = p * (dP - St[ii][2]);
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        av[c] = fmaf(pd, Ds[ii][c], av[c]);
        ak[c] = fmaf(wgt, Qs[ii][c], ak[c]);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 32; ++c) {
    ak[c] += __shfl_xor_sync(0xffffffffu, ak[c], 1);
    ak[c] += __shfl_xor_sync(0xffffffffu, ak[c], 2);
    av[c] += __shfl_xor_sync(0xffffffffu, av[c], 1);
    av[c] += __shfl_xor_sync(0xffffffffu, av[c], 2);
  }
  if (j < Sk) {
    float* dkp = dk + b * dk_bs + (long long)j * dk_rs + hd * 32;
    float* dvp = dv + b * dv_bs + (long long)j * dv_rs + hd * 32;
#pragma unroll
    for (int c = 0; c < 32; ++c)
      if ((c & 3) == t) {
        dkp[c] = ak[c] * scale;
        dvp[c] = av[c];
      }
  }
}

// ---- head tail backward: depth = sum_k softmax_k(l) c_k -------------------------------------------------------------------
// thread = pixel; logits [B,K,P] are overwritten by dl = gd * p * (c - depth); pg [B,K,P] = gd * p (its pixel sum is dcentres)
__global__ void tail_bwd_kernel(float* __restrict__ logits, float* __restrict__ pg, const float* __restrict__ centres,
                                const float* __restrict__ gdepth, const float* __restrict__ bias, int K, long long P) {
  extern __shared__ float cs[];  // [K] centres, [K] logit bias
  float* bs = cs + K;
  const int b = blockIdx.y;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    cs[k] = centres[(long long)b * K + k];
    bs[k] = bias ? bias[k] : 0.f;
  }
  __syncthreads();
  const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (px >= P) return;
  float* l = logits + (long long)b * K * P + px;
  float* g = pg + (long long)b * K * P + px;
  float m = -INFINITY;
  for (int k = 0; k < K; ++k) m = fmaxf(m, l[(long long)k * P] + bs[k]);
  float s = 0.f, a = 0.f;
  for (int k = 0; k < K; ++k) {
    const float e = __expf(l[(long long)k * P] + bs[k] - m);
    s += e;
    a = fmaf(e, cs[k], a);
  }
  const float inv = 1.f / s, d = a * inv, gd = gdepth[(long long)b * P + px];
  for (int k = 0; k < K; ++k) {
    const float p = __expf(l[(long long)k * P] + bs[k] - m) * inv;
    g[(long long)k * P] = gd * p;
    l[(long long)k * P] = gd * p * (cs[k] - d);
  }
}

// ---- bin-width normalisation backward ('linear': w = (relu(y)+0.1)/sum; edges = min + (max-min) cumsum(w); centres = mid-points)
// one block per image, K <= 1024 threads-strided
__global__ void bins_bwd_kernel(const float* __restrict__ y_raw, const float* __restrict__ dedges, const float* __restrict__ dcentres,
                                const float* __restrict__ dwidths, float* __restrict__ dy, int K, float range) {
  __shared__ float de[1025], dw[1024], sh[32];
  const int b = blockIdx.x;
  const float* y = y_raw + (long long)b * K;
  for (int j = threadIdx.x; j <= K; j += blockDim.x) {
    float v = dedges ? dedges[(long long)b * (K + 1) + j] : 0.f;
    if (dcentres) v += 0.5f * ((j > 0 ? dcentres[(long long)b * K + j - 1] : 0.f) + (j < K ? dcentres[(long long)b * K + j] : 0.f));
    de[j] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {  // dw_k = range * sum_{j > k} de_j  (suffix sum, fixed order)
    float run = 0.f;
    for (int k = K - 1; k >= 0; --k) {
      run += de[k + 1];
      dw[k] = range * run + (dwidths ? dwidths[(long long)b * K + k] : 0.f);   // + a gradient given on the normalised widths themselves
    }
  }
  __syncthreads();
  float su = 0.f, sdw = 0.f;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    const float u = fmaxf(y[k], 0.f) + 0.1f;
    su += u;
    sdw += dw[k] * u;
  }
  const float S = block_sum_f(su, sh);
  const float dot = block_sum_f(sdw, sh) / S;  // sum_m dw_m w_m
  for (int k = threadIdx.x; k < K; k += blockDim.x)
    dy[(long long)b * K + k] = y[k] > 0.f ? (dw[k] - dot) / S : 0.f;
}

// ---- elementwise dropout (+ optional residual): out = residual + (keep ? x / (1-p) : 0); the same call on a gradient is its backward
__global__ void dropout_kernel(const float* __restrict__ x, const float* __restrict__ residual, float* __restrict__ out, long long n,
                               uint32_t thr, float scale, unsigned long long seed_in, const unsigned long long* salt) {
  const unsigned long long seed = ocv_salted(seed_in, salt);
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float v = ocv_drop_keep(seed, (unsigned long long)i, thr) ? x[i] * scale : 0.f;
  out[i] = residual ? residual[i] + v : v;
}

// ---- adjoint of ocv_pad_objects_f32: dsrc[offsets[b] + j] = beta * dsrc + dpadded[b, left + j]  (pad rows receive no gradient)
__global__ void unpad_objects_kernel(const float* __restrict__ dpadded, const int32_t* __restrict__ offsets, float* __restrict__ dsrc,
                                     int L, int Nmax, int E, int left_pad, float beta) {
  const int b = blockIdx.y, j = blockIdx.x;
  const int beg = offsets[b], n = offsets[b + 1] - beg;
  if (j >= n) return;
  const int left = left_pad ? (L - Nmax) : 0;
  const float* s = dpadded + ((long long)b * L + left + j) * E;
  float* d = dsrc + (long long)(beg + j) * E;
  for (int e = threadIdx.x; e < E; e += blockDim.x) d[e] = beta != 0.f ? beta * d[e] + s[e] : s[e];
}

}  // namespace

extern "C" int ocv_dropout_f32(const float* x, const float* residual, float* out, long long n, float p, unsigned long long seed,
                               void* stream) {
  OCV_CHECK_ARG(x && out && n > 0 && p >= 0.f && p < 1.f);
  if (int s = ocv_require_sm100()) return s;
  dropout_kernel<<<(unsigned)ceil_div<long long>(n, 256), 256, 0, ocv_stream(stream)>>>(x, residual, out, n, ocv_drop_threshold(p),
                                                                                       1.f / (1.f - p), seed, ocv_dropout_salt());
  return ocv_launch_status();
}

extern "C" int ocv_unpad_objects_f32(const float* dpadded, const int32_t* offsets, float* dsrc, int B, int L, int Nmax, int E,
                                     int left_pad, float beta, void* stream) {
  OCV_CHECK_ARG(dpadded && offsets && dsrc && B > 0 && L > 0 && E > 0);
  OCV_CHECK_SHAPE(Nmax > 0 && Nmax <= L && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  unpad_objects_kernel<<<dim3(Nmax, B), 128, 0, ocv_stream(stream)>>>(dpadded, offsets, dsrc, L, Nmax, E, left_pad, beta);
  return ocv_launch_status();
}

extern "C" int ocv_reduce_channels_f32(const float* x, float* out, int groups, int outer, int C, long long inner, float beta,
                                       void* stream) {
  OCV_CHECK_ARG(x && out && groups > 0 && outer > 0 && C > 0 && inner > 0);
  OCV_CHECK_SHAPE(groups <= 65535);
  if (int s = ocv_require_sm100()) return s;
  if (inner == 1) reduce_columns_kernel<<<dim3(ceil_div(C, 32), groups), dim3(32, 32), 0, ocv_stream(stream)>>>(x, out, outer, C, beta);
  else reduce_channels_kernel<<<dim3(C, groups), inner >= 4096 ? 512 : 128, 0, ocv_stream(stream)>>>(x, out, outer, C, inner, beta);
  return ocv_launch_status();
}

extern "C" int ocv_act_bwd_f32(const float* dy, const float* out, float* dx, long long n, int act, void* stream) {
  OCV_CHECK_ARG(dy && out && dx && n > 0);
  if (int s = ocv_require_sm100()) return s;
  act_bwd_kernel<<<(unsigned)ceil_div<long long>(n, 256), 256, 0, ocv_stream(stream)>>>(dy, out, dx, n, act);
  return ocv_launch_status();
}

extern "C" int ocv_dropout_act_bwd_f32(const float* dy, const float* out, float* dx, long long n, int act, float p,
                                       unsigned long long seed, void* stream) {
  OCV_CHECK_ARG(dy && out && dx && n > 0 && p >= 0.f && p < 1.f);
  if (int s = ocv_require_sm100()) return s;
  dropout_act_bwd_kernel<<<(unsigned)ceil_div<long long>(n, 256), 256, 0, ocv_stream(stream)>>>(
      dy, out, dx, n, act, ocv_drop_threshold(p), 1.f / (1.f - p), seed, ocv_dropout_salt());
  return ocv_launch_status();
}

extern "C" int ocv_layernorm_bwd_f32(const float* dy, const float* z, const float* gamma, float* dz, float* dgamma, float* dbeta,
                                     int rows, int E, float beta, void* workspace /* rows*2 + 32*E floats */, void* stream) {
  OCV_CHECK_ARG(dy && z && gamma && dz && workspace && rows > 0);
  OCV_CHECK_SHAPE(E > 0 && E % 32 == 0 && E <= 512);
  OCV_CHECK_ARG((dgamma == nullptr) == (dbeta == nullptr));
  if (int s = ocv_require_sm100()) return s;
  float* stats = static_cast<float*>(workspace);
  layernorm_bwd_rows_kernel<<<ceil_div(rows, 8), 256, 0, ocv_stream(stream)>>>(dy, z, gamma, dz, stats, rows, E, nullptr, 0u, 1.f, 0ull,
                                                                               nullptr);
  if (int s = ocv_launch_status()) return s;
  if (dgamma) {
    float* part = stats + 2 * (long long)rows;
    layernorm_bwd_params_kernel<<<dim3(ceil_div(E, 32), LNG), dim3(32, 8), 0, ocv_stream(stream)>>>(dy, z, stats, part, rows, E);
    if (int s = ocv_launch_status()) return s;
    layernorm_bwd_params_final_kernel<<<ceil_div(E, 128), 128, 0, ocv_stream(stream)>>>(part, dgamma, dbeta, E, beta);
    return ocv_launch_status();
  }
  return OCV_OK;
}

// Data half of the LayerNorm backward with the dropout site that follows it in the backward of a training-mode encoder layer:
// dz as ocv_layernorm_bwd_f32, dz_dropped = dropout(dz) (bit-identical to ocv_dropout_f32 on dz); row statistics into `workspace`.
extern "C" int ocv_layernorm_bwd_dropout_f32(const float* dy, const float* z, const float* gamma, float* dz, float* dz_dropped, int rows,
                                             int E, float p, unsigned long long seed, void* workspace, void* stream) {
  OCV_CHECK_ARG(dy && z && gamma && dz && dz_dropped && workspace && rows > 0 && p >= 0.f && p < 1.f);
  OCV_CHECK_SHAPE(E > 0 && E % 32 == 0 && E <= 512);
  if (int s = ocv_require_sm100()) return s;
  layernorm_bwd_rows_kernel<<<ceil_div(rows, 8), 256, 0, ocv_stream(stream)>>>(dy, z, gamma, dz, static_cast<float*>(workspace), rows, E,
                                                                               dz_dropped, ocv_drop_threshold(p), 1.f / (1.f - p), seed,
                                                                               ocv_dropout_salt());
  return ocv_launch_status();
}

// The parameter half of ocv_layernorm_bwd_f32 on its own: `workspace` is the one a preceding ocv_layernorm_bwd_f32(dgamma = dbeta =
// NULL) call on the same (dy, z) filled with the row statistics.  Lets a sequencer keep only the data gradient on its chain.
extern "C" int ocv_layernorm_bwd_params_f32(const float* dy, const float* z, float* dgamma, float* dbeta, int rows, int E, float beta,
                                            void* workspace /* rows*2 + 32*E floats */, void* stream) {
  OCV_CHECK_ARG(dy && z && dgamma && dbeta && workspace && rows > 0);
  OCV_CHECK_SHAPE(E > 0 && E % 32 == 0 && E <= 512);
  if (int s = ocv_require_sm100()) return s;
  float* stats = static_cast<float*>(workspace);
  float* part = stats + 2 * (long long)rows;
  layernorm_bwd_params_kernel<<<dim3(ceil_div(E, 32), LNG), dim3(32, 8), 0, ocv_stream(stream)>>>(dy, z, stats, part, rows, E);
  if (int s = ocv_launch_status()) return s;
  layernorm_bwd_params_final_kernel<<<ceil_div(E, 128), 128, 0, ocv_stream(stream)>>>(part, dgamma, dbeta, E, beta);
  return ocv_launch_status();
}

extern "C" int ocv_mha_core_bwd_f32(const float* q, const float* k, const float* v, const float* out, const float* dout,
                                    const uint8_t* key_padding_mask, float* dq, float* dk, float* dv, int B, int Sq, int Sk,
                                    int heads, int d, long long q_bstride, int q_rstride, long long k_bstride, int k_rstride,
                                    long long v_bstride, int v_rstride, long long o_bstride, int o_rstride, long long dq_bstride,
                                    int dq_rstride, long long dk_bstride, int dk_rstride, long long dv_bstride, int dv_rstride,
                                    void* workspace /* B*heads*Sq*3 floats */, void* stream) {
  return ocv_mha_core_bwd_dropout_f32(q, k, v, out, dout, key_padding_mask, dq, dk, dv, B, Sq, Sk, heads, d, q_bstride, q_rstride,
                                      k_bstride, k_rstride, v_bstride, v_rstride, o_bstride, o_rstride, dq_bstride, dq_rstride,
                                      dk_bstride, dk_rstride, dv_bstride, dv_rstride, 0.f, 0ull, workspace, stream);
}

extern "C" int ocv_mha_core_bwd_dropout_f32(const float* q, const float* k, const float* v, const float* out, const float* dout,
                                            const uint8_t* key_padding_mask, float* dq, float* dk, float* dv, int B, int Sq, int Sk,
                                            int heads, int d, long long q_bstride, int q_rstride, long long k_bstride, int k_rstride,
                                            long long v_bstride, int v_rstride, long long o_bstride, int o_rstride,
                                            long long dq_bstride, int dq_rstride, long long dk_bstride, int dk_rstride,
                                            long long dv_bstride, int dv_rstride, float drop_p, unsigned long long drop_seed,
                                            void* workspace /* B*heads*Sq*3 floats */, void* stream) {
  OCV_CHECK_ARG(q && k && v && out && dout && dq && dk && dv && workspace && B > 0 && Sq > 0 && Sk > 0 && heads > 0);
  OCV_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f);
  OCV_CHECK_SHAPE(d == 32 && B <= 65535 && heads <= 65535);
  if (int s = ocv_require_sm100()) return s;
  float* stats = static_cast<float*>(workspace);
  static bool smem_opt_in = false;       // tiles of up to AJ_MAX rows need more than the 48 KB default
  constexpr int SMEM_MAX = 2 * AJ_MAX * APAD * 4 + AJ_MAX * 12;
  if (!smem_opt_in) {
    if (cudaError_t e = cudaFuncSetAttribute(mha_bwd_q_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX)) return (int)e;
    if (cudaError_t e = cudaFuncSetAttribute(mha_bwd_kv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX)) return (int)e;
    smem_opt_in = true;
  }
  auto tile = [](int rows) { return rows >= AJ_MAX ? AJ_MAX : (rows + ASL - 1) / ASL * ASL; };
  const int ajk = tile(Sk), ajq = tile(Sq);
  mha_bwd_q_kernel<<<dim3(ceil_div(Sq, AR), heads, B), AR * ASL, 2 * ajk * APAD * 4 + ajk, ocv_stream(stream)>>>(
      q, k, v, out, dout, key_padding_mask, dq, stats, Sq, Sk, heads, q_bstride, q_rstride, k_bstride, k_rstride, v_bstride,
      v_rstride, o_bstride, o_rstride, dq_bstride, dq_rstride, drop_p, drop_seed, ocv_dropout_salt(), ajk);
  if (int s = ocv_launch_status()) return s;
  mha_bwd_kv_kernel<<<dim3(ceil_div(Sk, AR), heads, B), AR * ASL, 2 * ajq * APAD * 4 + ajq * 12, ocv_stream(stream)>>>(
      q, k, v, dout, key_padding_mask, stats, dk, dv, Sq, Sk, heads, q_bstride, q_rstride, k_bstride, k_rstride, v_bstride,
      v_rstride, o_bstride, o_rstride, dk_bstride, dk_rstride, dv_bstride, dv_rstride, drop_p, drop_seed, ocv_dropout_salt(), ajq);
  return ocv_launch_status();
}

extern "C" int ocv_tail_bwd_f32(float* logits_inout, float* pg_out, const float* centres, const float* grad_depth,
                                const float* logit_bias, int B, int K, long long P, void* stream) {
  OCV_CHECK_ARG(logits_inout && pg_out && centres && grad_depth && B > 0 && P > 0);
  OCV_CHECK_SHAPE(K > 0 && K <= 4096 && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  tail_bwd_kernel<<<dim3((unsigned)ceil_div<long long>(P, 128), B), 128, 2 * K * sizeof(float), ocv_stream(stream)>>>(
      logits_inout, pg_out, centres, grad_depth, logit_bias, K, P);
  return ocv_launch_status();
}

extern "C" int ocv_bins_bwd_ex_f32(const float* y_raw, const float* grad_edges, const float* grad_centres, const float* grad_widths,
                                   float* grad_y, int B, int K, float min_depth, float max_depth, void* stream) {
  OCV_CHECK_ARG(y_raw && grad_y && (grad_edges || grad_centres || grad_widths) && B > 0);
  OCV_CHECK_SHAPE(K > 0 && K <= 1024);
  if (int s = ocv_require_sm100()) return s;
  bins_bwd_kernel<<<B, 256, 0, ocv_stream(stream)>>>(y_raw, grad_edges, grad_centres, grad_widths, grad_y, K, max_depth - min_depth);
  return ocv_launch_status();
}

extern "C" int ocv_bins_bwd_f32(const float* y_raw, const float* grad_edges, const float* grad_centres, float* grad_y, int B, int K,
                                float min_depth, float max_depth, void* stream) {
  return ocv_bins_bwd_ex_f32(y_raw, grad_edges, grad_centres, nullptr, grad_y, B, K, min_depth, max_depth, stream);
}

// ---- LayerNorm backward for callers that have no device memory to lend ------------------------------------------------------
// ocv_layernorm_bwd_f32 needs rows*2 + 32*E floats of scratch (per-row mean / rstd, then the fixed-order partial sums of dgamma and
// dbeta) and expects the caller to bring it.  Some hosts cannot: a plugin that is handed its input and output buffers by a runtime
// whose memory plan was frozen at load time, a callback that runs while the owning framework holds its allocator lock, a real-time
// loop whose policy forbids cudaMalloc after start-up (it synchronises the whole device and may fail late).  For them the scratch
// lives inside the library image: a file-scope device array of fixed size.  Every device that loads the library carries these
// OCV_LN_BWD_NOALLOC_MAX_ROWS*2 + 32*512 floats (192 KB) until the process exits, whether or not the entry point is ever called.
//
// `g_ln_bwd_scratch` below is two things.  In device code it is the array.  In host code the same identifier is only a handle the
// runtime uses to look the array up: it is NOT a device address, and a kernel given `g_ln_bwd_scratch` from the host side would
// fault (or, on a system where the GPU can reach host memory, quietly scribble over the wrong bytes).  So the host asks the runtime
// for the real address once per device, with that device current, and hands the kernels that.
namespace {

constexpr int LN_NOALLOC_MAX_E = 512;                                           // the kernels' own limit
constexpr long long LN_NOALLOC_FLOATS = 2ll * OCV_LN_BWD_NOALLOC_MAX_ROWS + 32ll * LN_NOALLOC_MAX_E;
static __device__ __align__(256) float g_ln_bwd_scratch[LN_NOALLOC_FLOATS];

constexpr int LN_NOALLOC_MAX_DEVICES = 64;
std::mutex g_ln_scratch_mu;
float* g_ln_scratch_addr[LN_NOALLOC_MAX_DEVICES] = {};   // device address of g_ln_bwd_scratch on device i, once resolved

// Device address of the scratch array on the CURRENT device; resolved on the first call per device (this is what may load the
// module on that device, hence "first call before any stream capture").  Returns a status, never allocates.
int ln_noalloc_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= LN_NOALLOC_MAX_DEVICES) return OCV_E_CAPACITY;
  std::lock_guard<std::mutex> lock(g_ln_scratch_mu);
  if (!g_ln_scratch_addr[dev]) {
    float* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_ln_bwd_scratch);
    if (e != cudaSuccess) return (int)e;
    if (!ptr) return OCV_E_DRIVER;
    g_ln_scratch_addr[dev] = ptr;
  }
  *out = g_ln_scratch_addr[dev];
  return OCV_OK;
}

}  // namespace

// Same arithmetic, same kernels, same launch shapes as ocv_layernorm_bwd_f32 — only the scratch pointer differs, so the results
// are bit-identical to it.  Nothing in the scratch has to start at zero: the row kernel writes every (mean, rstd) pair and the
// partial-sum kernel writes every [LNG][2][E] slot before the final kernel reads them, so stale contents of an earlier call are
// overwritten, never accumulated into.  What the shared array does NOT give is isolation: two calls in flight at once on one device
// (different streams, or different host threads) write the same floats and corrupt each other — order them.
extern "C" int ocv_layernorm_bwd_noalloc_f32(const float* dy, const float* z, const float* gamma, float* dz, float* dgamma,
                                             float* dbeta, int rows, int E, float beta, void* stream) {
  OCV_CHECK_ARG(dy && z && gamma && dz && rows > 0);
  OCV_CHECK_SHAPE(E > 0 && E % 32 == 0 && E <= LN_NOALLOC_MAX_E);
  OCV_CHECK_ARG((dgamma == nullptr) == (dbeta == nullptr));
  if (rows > OCV_LN_BWD_NOALLOC_MAX_ROWS) return OCV_E_CAPACITY;   // would run past the built-in array: refuse, do not truncate
  if (int s = ocv_require_sm100()) return s;
  float* stats = nullptr;
  if (int s = ln_noalloc_scratch(&stats)) return s;
  layernorm_bwd_rows_kernel<<<ceil_div(rows, 8), 256, 0, ocv_stream(stream)>>>(dy, z, gamma, dz, stats, rows, E, nullptr, 0u, 1.f, 0ull,
                                                                               nullptr);
  if (int s = ocv_launch_status()) return s;
  if (dgamma) {
    float* part = stats + 2 * (long long)rows;
    layernorm_bwd_params_kernel<<<dim3(ceil_div(E, 32), LNG), dim3(32, 8), 0, ocv_stream(stream)>>>(dy, z, stats, part, rows, E);
    if (int s = ocv_launch_status()) return s;
    layernorm_bwd_params_final_kernel<<<ceil_div(E, 128), 128, 0, ocv_stream(stream)>>>(part, dgamma, dbeta, E, beta);
    return ocv_launch_status();
  }
  return OCV_OK;
}
