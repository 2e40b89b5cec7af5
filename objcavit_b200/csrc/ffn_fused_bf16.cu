// Fused feed-forward block of nn.TransformerEncoderLayer (post-norm) on tcgen05:
//   out = LayerNorm( x + relu(x W1^T + b1) W2^T + b2 ) * gamma + beta          x [M,128], W1 [F,128], W2 [128,F], F = 1024
// replaces linear1 + ReLU + linear2 + residual + norm2 (modules/layers.py:8; modules/ObjCAViT.py:155-161) — previously two GEMM
// launches with the [M,F] hidden activations written to and re-read from HBM by 8x more CTAs than rows need.
// One CTA owns 128 token rows and walks the hidden dimension in chunks of 128:
//   GEMM1(c): H_c[128x128] = X[128x128] W1[c]^T            -> TMEM (double-buffered accumulator)
//   convert : + b1, ReLU, bf16 -> swizzled smem tile       (8 epilogue warps; this is GEMM2's K-major A operand)
//   GEMM2(c): OUT[128x128] += H_c[128x128] W2[:, c]^T      -> TMEM (one accumulator for the whole block)
// GEMM1(c+1) is issued before GEMM2(c), so the tensor pipe works while chunk c is converted.  Weight slabs [128 x 64] stream
// through a 6-slot TMA ring in exactly the MMA issuer's consumption order.  The hidden activations never leave the SM.
// Final epilogue (all 16 warps): + b2 + residual, LayerNorm over the 128-wide row, fp32 and bf16 outputs.
// Warp roles: 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator, 4..11 = convert warps.  Contract: include/objcavit_b200.h.
#include "tc_common.cuh"

namespace {

using namespace tc;

constexpr int E = 128;                  // model width = N of GEMM2 = K of GEMM1
constexpr int CH = 128;                 // hidden chunk
constexpr int SLAB = 128 * 64 * 2;      // 16 KB: [128 rows x 64 k] bf16, 128B swizzle
constexpr int NS = 6;                   // weight ring slots
constexpr int A_BYTES = 2 * SLAB;       // X tile: two K halves
constexpr int H_BYTES = 2 * SLAB;       // one hidden chunk as A operand: two K halves
constexpr int N_CVT = 256;              // convert threads (warps 4..11)
constexpr int THREADS = 512;
constexpr int MAX_F = 2048;
constexpr int SMEM_BYTES = A_BYTES + NS * SLAB + 2 * H_BYTES + 1024 /*align*/ + 512 /*barriers*/ + MAX_F * 4 /*b1*/ +
                           3 * E * 4 /*b2, gamma, beta*/ + 2 * 4 * 128 * 4 /*LayerNorm partials*/;
constexpr uint32_t TM_ACC1 = 0, TM_OUT = 256, TM_COLS = 512;

struct FfnParams {
  const float *b1, *b2, *residual, *gamma, *beta;
  const uint8_t* zero_rows;
  float* C32;
  __nv_bfloat16* C16;
  int M, F, ldr, ldc32, ldc16;
};

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

__global__ void __launch_bounds__(THREADS, 1)
ffn_fused_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW1,
                 const __grid_constant__ CUtensorMap tmW2, FfnParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* a_tile = smem;
  uint8_t* ring = a_tile + A_BYTES;
  uint8_t* h_tile = ring + NS * SLAB;  // [2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(h_tile + 2 * H_BYTES);
  uint64_t* w_full = bars;             // [NS]
  uint64_t* w_empty = w_full + NS;     // [NS]
  uint64_t* a_full = w_empty + NS;
  uint64_t* h_full = a_full + 1;       // [2]
  uint64_t* h_empty = h_full + 2;      // [2]
  uint64_t* acc_full = h_empty + 2;    // [2]
  uint64_t* acc_empty = acc_full + 2;  // [2]
  uint64_t* out_full = acc_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(out_full + 1);
  float* s_b1 = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);
  float* s_vec = s_b1 + MAX_F;         // b2 | gamma | beta
  float* red = s_vec + 3 * E;          // [2][4][128]

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int m0 = blockIdx.x * 128;
  const int n_chunks = p.F / CH;
  pdl_launch();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(&w_full[s], 1);
      mbar_init(&w_empty[s], 1);
    }
    mbar_init(a_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&h_full[i], N_CVT);
      mbar_init(&h_empty[i], 1);
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], N_CVT);
    }
    mbar_init(out_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, TM_COLS);
  // parameter vectors do not depend on the previous kernel of the chain
  for (int i = threadIdx.x; i < p.F; i += THREADS) s_b1[i] = __ldg(p.b1 + i);
  for (int i = threadIdx.x; i < 3 * E; i += THREADS) {
    const int which = i / E, j = i - which * E;
    const float* src = which == 0 ? p.b2 : (which == 1 ? p.gamma : p.beta);
    s_vec[i] = src ? __ldg(src + j) : (which == 1 ? 1.f : 0.f);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer: slabs in the MMA issuer's consumption order =================
    if (elect_one()) {
      uint32_t ring_n = 0;
      auto load_slab = [&](const CUtensorMap* m, int c0, int c1) {
        const int s = ring_n % NS;
        mbar_wait(&w_empty[s], ((ring_n / NS) & 1) ^ 1);
        mbar_arrive_expect_tx(&w_full[s], SLAB);
        tma_load_2d(ring + s * SLAB, m, &w_full[s], c0, c1);
        ++ring_n;
      };
      auto w1 = [&](int c) {
        load_slab(&tmW1, 0, c * CH);
        load_slab(&tmW1, 64, c * CH);
      };
      auto w2 = [&](int c) {
        load_slab(&tmW2, c * CH, 0);
        load_slab(&tmW2, c * CH + 64, 0);
      };
      w1(0);                       // weights first: they do not depend on the previous kernel
      if (n_chunks > 1) w1(1);
      pdl_wait();
      mbar_arrive_expect_tx(a_full, A_BYTES);
      tma_load_2d(a_tile, &tmX, a_full, 0, m0);
      tma_load_2d(a_tile + SLAB, &tmX, a_full, 64, m0);
      for (int c = 0; c < n_chunks; ++c) {
        if (c > 0 && c + 1 < n_chunks) w1(c + 1);
        w2(c);
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 128);
      uint32_t ring_n = 0;
      const uint32_t a_base = smem_u32(a_tile);
      auto gemm1 = [&](int c) {
        const int buf = c & 1;
        mbar_wait(&acc_empty[buf], ((c >> 1) & 1) ^ 1);
        tc_fence_after();
        for (int kh = 0; kh < 2; ++kh) {
          const int s = ring_n % NS;
          mbar_wait(&w_full[s], (ring_n / NS) & 1);
          tc_fence_after();
          const uint32_t b_base = smem_u32(ring + s * SLAB);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem + TM_ACC1 + buf * 128, desc_kmajor_sw128(a_base + kh * SLAB + kk * 32),
                      desc_kmajor_sw128(b_base + kk * 32), idesc, (kh | kk) != 0);
          umma_commit(&w_empty[s]);
          ++ring_n;
        }
        umma_commit(&acc_full[buf]);
      };
      auto gemm2 = [&](int c) {
        const int buf = c & 1;
        mbar_wait(&h_full[buf], (c >> 1) & 1);
        tc_fence_after();
        const uint32_t h_base = smem_u32(h_tile + buf * H_BYTES);
        for (int kh = 0; kh < 2; ++kh) {
          const int s = ring_n % NS;
          mbar_wait(&w_full[s], (ring_n / NS) & 1);
          tc_fence_after();
          const uint32_t b_base = smem_u32(ring + s * SLAB);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem + TM_OUT, desc_kmajor_sw128(h_base + kh * SLAB + kk * 32), desc_kmajor_sw128(b_base + kk * 32), idesc,
                      (c | kh | kk) != 0);
          umma_commit(&w_empty[s]);
          ++ring_n;
        }
        umma_commit(&h_empty[buf]);
      };
      mbar_wait(a_full, 0);
      tc_fence_after();
      gemm1(0);
      for (int c = 0; c < n_chunks; ++c) {
        if (c + 1 < n_chunks) gemm1(c + 1);
        gemm2(c);
      }
      umma_commit(out_full);
    }
  } else if (warp >= 4 && warp < 12) {
    // ================= convert warps: thread = one row x one 64-column half of the chunk =================
    const int e = warp - 4;
    const int q = e & 3;  // == warp % 4: the TMEM lane quarter this warp may read
    const int half = e >> 2;
    const int m = q * 32 + lane;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    for (int c = 0; c < n_chunks; ++c) {
      const int buf = c & 1;
      mbar_wait(&acc_full[buf], (c >> 1) & 1);
      tc_fence_after();
      const uint32_t src = tmem + lane_base + TM_ACC1 + buf * 128 + half * 64;
      uint32_t r0[32], r1[32], packed[32];
      tmem_ld32(src, r0);
      tmem_ld32(src + 32, r1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&acc_empty[buf]);  // the accumulator is in registers: GEMM1(c+2) may overwrite it
      const float* bb = s_b1 + c * CH + half * 64;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        packed[j] = pack_bf16(fmaxf(__uint_as_float(r0[2 * j]) + bb[2 * j], 0.f), fmaxf(__uint_as_float(r0[2 * j + 1]) + bb[2 * j + 1], 0.f));
        packed[16 + j] = pack_bf16(fmaxf(__uint_as_float(r1[2 * j]) + bb[32 + 2 * j], 0.f),
                                   fmaxf(__uint_as_float(r1[2 * j + 1]) + bb[32 + 2 * j + 1], 0.f));
      }
      mbar_wait(&h_empty[buf], ((c >> 1) & 1) ^ 1);  // GEMM2(c-2) has finished reading this staging tile
      uint8_t* row = h_tile + buf * H_BYTES + half * SLAB + m * 128;
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        uint4 v = make_uint4(packed[ch * 4], packed[ch * 4 + 1], packed[ch * 4 + 2], packed[ch * 4 + 3]);
        *reinterpret_cast<uint4*>(row + ((ch ^ (m & 7)) << 4)) = v;
      }
      fence_proxy_async_smem();
      mbar_arrive(&h_full[buf]);
    }
  }
  // ================= final epilogue: all 16 warps =================
  __syncwarp();
  __syncthreads();
  pdl_wait();
  {
    const int quarter = warp & 3, cg = warp >> 2;
    const int row_l = quarter * 32 + lane;
    const int m = m0 + row_l;
    const bool row_ok = m < p.M;
    float v[32];
    // residual row segment straight from global memory (128 contiguous bytes per thread), in flight during the last MMAs
    {
      const float4* rp = reinterpret_cast<const float4*>(p.residual + (long long)(row_ok ? m : 0) * p.ldr + cg * 32);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 t = row_ok ? __ldg(rp + j) : make_float4(0.f, 0.f, 0.f, 0.f);
        v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
      }
    }
    mbar_wait(out_full, 0);
    tc_fence_after();
    uint32_t r[32];
    tmem_ld32(tmem + ((uint32_t)(quarter * 32) << 16) + TM_OUT + cg * 32, r);
    tmem_ld_wait();
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      v[j] += __uint_as_float(r[j]) + s_vec[cg * 32 + j];
      sum += v[j];
    }
    red[cg * 128 + row_l] = sum;
    __syncthreads();
    const float mean = ((red[row_l] + red[128 + row_l]) + (red[256 + row_l] + red[384 + row_l])) * (1.f / E);
    float var = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      v[j] -= mean;
      var = fmaf(v[j], v[j], var);
    }
    red[512 + cg * 128 + row_l] = var;
    __syncthreads();
    const float rstd = rsqrtf(((red[512 + row_l] + red[640 + row_l]) + (red[768 + row_l] + red[896 + row_l])) * (1.f / E) + 1e-5f);
    const bool zero = row_ok && p.zero_rows && p.zero_rows[m];
#pragma unroll
    for (int j = 0; j < 32; ++j)
      v[j] = zero ? 0.f : v[j] * rstd * s_vec[E + cg * 32 + j] + s_vec[2 * E + cg * 32 + j];
    if (row_ok) {
      if (p.C32) {
        float4* d = reinterpret_cast<float4*>(p.C32 + (long long)m * p.ldc32 + cg * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j) d[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
      if (p.C16) {
        uint4* d = reinterpret_cast<uint4*>(p.C16 + (long long)m * p.ldc16 + cg * 32);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          d[j] = make_uint4(pack_bf16(v[8 * j], v[8 * j + 1]), pack_bf16(v[8 * j + 2], v[8 * j + 3]), pack_bf16(v[8 * j + 4], v[8 * j + 5]),
                            pack_bf16(v[8 * j + 6], v[8 * j + 7]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, TM_COLS);
}

}  // namespace

extern "C" int ocv_ffn_fused_bf16(const void* X16, int lda, const void* W1, const float* b1, const void* W2, const float* b2,
                                  const float* residual, int ldr, const float* ln_gamma, const float* ln_beta,
                                  const uint8_t* zero_rows, float* C32, int ldc32, void* C16, int ldc16, int M, int F, void* stream) {
  OCV_CHECK_ARG(X16 && W1 && b1 && W2 && b2 && residual && ln_gamma && ln_beta && (C32 || C16) && M > 0);
  OCV_CHECK_SHAPE(F >= CH && F % CH == 0 && F <= MAX_F && lda % 8 == 0);
  OCV_CHECK_ALIGN(ldr % 4 == 0 && (!C32 || ldc32 % 4 == 0) && (!C16 || ldc16 % 8 == 0));
  OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(residual) | reinterpret_cast<uintptr_t>(C32) | reinterpret_cast<uintptr_t>(C16)) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  CUtensorMap tmX, tmW1, tmW2;
  if (int s = tc::make_tmap_2d_bf16(&tmX, X16, M, E, lda, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  if (int s = tc::make_tmap_2d_bf16(&tmW1, W1, F, E, E, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  if (int s = tc::make_tmap_2d_bf16(&tmW2, W2, E, F, F, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(ffn_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  FfnParams p{};
  p.b1 = b1; p.b2 = b2; p.residual = residual; p.gamma = ln_gamma; p.beta = ln_beta; p.zero_rows = zero_rows;
  p.C32 = C32; p.C16 = static_cast<__nv_bfloat16*>(C16);
  p.M = M; p.F = F; p.ldr = ldr; p.ldc32 = ldc32; p.ldc16 = ldc16;
  return ocv_launch_pdl(ffn_fused_kernel, dim3(ceil_div(M, 128)), dim3(THREADS), SMEM_BYTES, stream, tmX, tmW1, tmW2, p);
}
