// bf16 tensor-core GEMM for the token path and the patch embedding:
//   C[M,N] = act(A[M,K] @ Wt[N,K]^T + bias) (+ residual)
// tcgen05.mma (kind::f16, M=128 x N=128 x K=16 per instruction, fp32 accumulators in TMEM), operands
// staged by TMA with the 128-byte swizzle through a 4-stage mbarrier ring.  Warp roles: warp 0 = TMA
// producer, warp 1 = MMA issuer (one elected lane), warp 2 = TMEM allocator, warps 4-7 = epilogue
// (tcgen05.ld -> bias / activation / residual -> global).  Contract: include/objcavit_b200.h.
#include "tc_common.cuh"

namespace tc {

EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows, uint32_t box_cols, CUtensorMapSwizzle swz) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return OCV_E_DRIVER;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || ((ld * 2) & 15)) return OCV_E_ALIGN;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OCV_OK : OCV_E_DRIVER;
}

}  // namespace tc

namespace {

using namespace tc;

constexpr int GM = 128, GN = 128, GK = 64, STAGES = 4;
constexpr int A_BYTES = GM * GK * 2, B_BYTES = GN * GK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int TP = 132;  // fp32 staging-tile pitch (floats): 16-byte aligned rows, conflict-free 128-bit row accesses
constexpr int TILE_BYTES = GM * TP * 4;
constexpr int GEMM_SMEM = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/ + TILE_BYTES + 3 * GN * 4 /*bias, gamma, beta*/;

struct GemmParams {
  const float* bias;
  const float* residual;
  void* C;
  int M, N, K, ldc, act, c_is_bf16, k_per_split;
  // extended epilogue (ocv_gemm_bf16_ex): LayerNorm over the N == 128 row, second (bf16) output, zeroed rows
  const float* gamma;
  const float* beta;
  const uint8_t* zero_rows;
  __nv_bfloat16* C16;
  int ldc16, ldr;
  // v2 (training): operand majorness, dropout in the epilogue, pre-LayerNorm sum as a third output
  int a_mn, b_mn;                  // 1: the operand is stored [K rows, M|N cols] (contraction index = row) and read MN-major
  uint32_t drop_thr;               // 0 = no dropout; else keep iff hash(seed, m*N+n) >= thr, kept values scaled by drop_scale
  float drop_scale;
  unsigned long long drop_seed;
  const unsigned long long* drop_salt;
  float* Cpre;                     // with LayerNorm: act(A W^T + b) [dropout] + residual BEFORE the normalisation (what its backward reads)
  int ldpre;
};

__global__ void __launch_bounds__(512, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int m0 = blockIdx.y * GM, n0 = blockIdx.x * GN;
  pdl_launch();  // the next kernel of the chain may start its prologue
  int k_begin = 0, k_end = p.K;
  if (p.k_per_split > 0) {
    k_begin = blockIdx.z * p.k_per_split;
    k_end = min(p.K, k_begin + p.k_per_split);
  }
  const int n_k = (k_end > k_begin) ? (k_end - k_begin + GK - 1) / GK : 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, GN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;
  float* tile = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES + 256);
  float* s_bias = tile + GM * TP;  // [128] bias, [128] gamma, [128] beta of this column tile
  const bool split = p.k_per_split > 0;

  if (warp >= 2) {
    // parameters do not depend on the previous kernel: fetch them before the grid dependency resolves
    for (int i = threadIdx.x - 64; i < 3 * GN; i += 448) {
      const int which = i / GN, j = i - which * GN;
      const float* src = which == 0 ? p.bias : (which == 1 ? p.gamma : p.beta);
      s_bias[i] = (src && n0 + j < p.N) ? __ldg(src + n0 + j) : 0.f;
    }
    pdl_wait();
  }
  if (warp >= 2 && p.residual && !split) {
    // residual tile -> shared memory with coalesced 128-bit loads while the main loop runs (warps 2..15 are idle)
    const bool vec = (p.ldr & 3) == 0 && (reinterpret_cast<uintptr_t>(p.residual) & 15) == 0;
    // all loads of a thread are issued before the first store (one L2 round trip, not ten)
    float4 rv[10];
#pragma unroll
    for (int t = 0; t < 10; ++t) {
      const int i = threadIdx.x - 64 + t * 448;
      const int r = i >> 5, c = (i & 31) * 4;
      const int m = m0 + r, n = n0 + c;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < GM * 32 && m < p.M && n < p.N) {
        const float* src = p.residual + (long long)m * p.ldr + n;
        if (vec && n + 3 < p.N) v = __ldg(reinterpret_cast<const float4*>(src));
        else {
          v.x = src[0];
          if (n + 1 < p.N) v.y = src[1];
          if (n + 2 < p.N) v.z = src[2];
          if (n + 3 < p.N) v.w = src[3];
        }
      }
      rv[t] = v;
    }
#pragma unroll
    for (int t = 0; t < 10; ++t) {
      const int i = threadIdx.x - 64 + t * 448;
      if (i < GM * 32) *reinterpret_cast<float4*>(tile + (i >> 5) * TP + (i & 31) * 4) = rv[t];
    }
  }
  if (warp == 0) {
    if (elect_one()) {
      // K-major operand: one box [128 rows x 64 k]; MN-major operand (stored [K rows, cols]): two boxes [64 k rows x 64 cols]
      // (= the two 64-element MN blocks of the canonical MN-major SW128 layout, 8 KB apart)
      auto load_a = [&](int it, int s) {
        uint8_t* dst = smem + s * STAGE_BYTES;
        const int k = k_begin + it * GK;
        if (p.a_mn) {
          tma_load_2d(dst, &tmA, &full[s], m0, k);
          tma_load_2d(dst + A_BYTES / 2, &tmA, &full[s], m0 + 64, k);
        } else {
          tma_load_2d(dst, &tmA, &full[s], k, m0);
        }
      };
      auto load_b = [&](int it, int s) {
        uint8_t* dst = smem + s * STAGE_BYTES + A_BYTES;
        const int k = k_begin + it * GK;
        if (p.b_mn) {
          tma_load_2d(dst, &tmB, &full[s], n0, k);
          tma_load_2d(dst + B_BYTES / 2, &tmB, &full[s], n0 + 64, k);
        } else {
          tma_load_2d(dst, &tmB, &full[s], k, n0);
        }
      };
      // weight tiles of the first stages before the grid dependency resolves, activations after it
      const int n_pre = n_k < STAGES ? n_k : STAGES;
      for (int it = 0; it < n_pre; ++it) {
        mbar_arrive_expect_tx(&full[it], STAGE_BYTES);
        load_b(it, it);
      }
      pdl_wait();
      for (int it = 0; it < n_pre; ++it) load_a(it, it);
      for (int it = n_pre; it < n_k; ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_arrive_expect_tx(&full[s], STAGE_BYTES);
        load_a(it, s);
        load_b(it, s);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(GM, GN, p.a_mn, p.b_mn);
      for (int it = 0; it < n_k; ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t a = smem_u32(smem + s * STAGE_BYTES);
        // K-major: 16 k = +32 B on the start address; MN-major: 16 k rows = +2048 B, MN blocks A_BYTES/2 apart, 8-row groups 1 KB
        const uint64_t da = p.a_mn ? make_smem_desc(a, A_BYTES / 2, 1024, LAYOUT_SW128) : desc_kmajor_sw128(a);
        const uint64_t db = p.b_mn ? make_smem_desc(a + A_BYTES, B_BYTES / 2, 1024, LAYOUT_SW128) : desc_kmajor_sw128(a + A_BYTES);
        const uint64_t sa = p.a_mn ? 128 : 2, sb = p.b_mn ? 128 : 2;   // descriptor start-address units of 16 B
#pragma unroll
        for (int kk = 0; kk < GK / 16; ++kk)
          umma_bf16(tmem_d, da + (uint64_t)kk * sa, db + (uint64_t)kk * sb, idesc, (it | kk) != 0);
        umma_commit(&empty[s]);
      }
      umma_commit(tmem_full);
    }
  }
  // ================= epilogue: ALL 16 warps =================
  // A warp may only read the TMEM lane quarter (warp % 4); the 4 warps of a quarter split the 128 accumulator columns
  // (32 each), so one thread finishes 32 outputs of one row.  With 4 warps per scheduler the global-memory latency of
  // the bias / residual / LayerNorm-parameter loads is hidden; the row-wide LayerNorm statistics are combined through
  // shared memory (two passes: mean, then centred variance).
  __syncwarp();
  __syncthreads();  // residual tile + parameter vectors complete; the two role warps have issued everything
  pdl_wait();       // every thread stores to global memory below
  {
    const int quarter = warp & 3, cg = warp >> 2;
    const int row_l = quarter * 32 + lane;
    const int m = m0 + row_l;
    const int nb = n0 + cg * 32;
    float* red = reinterpret_cast<float*>(smem);  // [2][4][128] floats: the pipeline stages are idle by then
    if (n_k > 0) {
      mbar_wait(tmem_full, 0);
      tc_fence_after();
    }
    uint32_t r[32];
    if (n_k > 0) {
      tmem_ld32(tmem_d + ((uint32_t)(quarter * 32) << 16) + cg * 32, r);
      tmem_ld_wait();
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) r[j] = 0;
    }
    const bool row_ok = m < p.M;
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
    float* trow = tile + row_l * TP + cg * 32;
    if (!split) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int n = nb + 4 * j;
        if (p.bias) {
          const float4 b = *reinterpret_cast<const float4*>(s_bias + cg * 32 + 4 * j);
          v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
        }
        if (p.act == OCV_ACT_RELU) {
#pragma unroll
          for (int t = 0; t < 4; ++t) v[4 * j + t] = fmaxf(v[4 * j + t], 0.f);
        } else if (p.act == OCV_ACT_LEAKY_RELU) {
#pragma unroll
          for (int t = 0; t < 4; ++t) v[4 * j + t] = v[4 * j + t] > 0.f ? v[4 * j + t] : 0.01f * v[4 * j + t];
        }
        if (p.drop_thr) {   // training: dropout of act(A W^T + b), element index m * N + n (the convention of ocv_dropout_f32)
          const unsigned long long seed = ocv_salted(p.drop_seed, p.drop_salt);
          const unsigned long long e0 = (unsigned long long)m * (unsigned long long)p.N + (unsigned long long)n;
#pragma unroll
          for (int t = 0; t < 4; ++t) v[4 * j + t] = ocv_drop_keep(seed, e0 + t, p.drop_thr) ? v[4 * j + t] * p.drop_scale : 0.f;
        }
        if (p.residual) {
          const float4 rr = *reinterpret_cast<const float4*>(trow + 4 * j);
          v[4 * j] += rr.x; v[4 * j + 1] += rr.y; v[4 * j + 2] += rr.z; v[4 * j + 3] += rr.w;
        }
      }
    }
    if (p.gamma && p.Cpre) {  // the pre-normalisation sum is an output of its own (saved for the LayerNorm backward)
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(trow + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      __syncthreads();
      const bool vp = (p.ldpre & 3) == 0 && (reinterpret_cast<uintptr_t>(p.Cpre) & 15) == 0;
      for (int i = threadIdx.x; i < GM * 32; i += 512) {
        const int rr = i >> 5, c = (i & 31) * 4;
        const int mm = m0 + rr, n = n0 + c;
        if (mm >= p.M || n >= p.N) continue;
        const float4 o = *reinterpret_cast<const float4*>(tile + rr * TP + c);
        float* d = p.Cpre + (long long)mm * p.ldpre + n;
        if (vp) *reinterpret_cast<float4*>(d) = o;
        else { d[0] = o.x; d[1] = o.y; d[2] = o.z; d[3] = o.w; }
      }
      __syncthreads();
    }
    if (p.gamma) {  // LayerNorm over the N == 128 row (post-norm encoder layer: LN(residual + A W^T + b))
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) sum += v[j];
      red[cg * 128 + row_l] = sum;
      __syncthreads();
      const float mean = ((red[row_l] + red[128 + row_l]) + (red[256 + row_l] + red[384 + row_l])) * (1.f / GN);
      float var = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        v[j] -= mean;
        var = fmaf(v[j], v[j], var);
      }
      red[512 + cg * 128 + row_l] = var;
      __syncthreads();
      const float rstd = rsqrtf(((red[512 + row_l] + red[640 + row_l]) + (red[768 + row_l] + red[896 + row_l])) * (1.f / GN) + 1e-5f);
      const bool zero = row_ok && p.zero_rows && p.zero_rows[m];
      const float4* g4 = reinterpret_cast<const float4*>(s_bias + GN + cg * 32);
      const float4* e4 = reinterpret_cast<const float4*>(s_bias + 2 * GN + cg * 32);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 g = g4[j], e = e4[j];
        v[4 * j] = zero ? 0.f : v[4 * j] * rstd * g.x + e.x;
        v[4 * j + 1] = zero ? 0.f : v[4 * j + 1] * rstd * g.y + e.y;
        v[4 * j + 2] = zero ? 0.f : v[4 * j + 2] * rstd * g.z + e.z;
        v[4 * j + 3] = zero ? 0.f : v[4 * j + 3] * rstd * g.w + e.w;
      }
    }
    // results -> staging tile (each thread overwrites exactly the residual elements it consumed), then coalesced rows out
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<float4*>(trow + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    __syncthreads();
    float* c32 = nullptr;
    __nv_bfloat16* c16 = nullptr;
    int ld32 = p.ldc, ld16 = p.ldc16;
    if (split) c32 = static_cast<float*>(p.C) + (long long)blockIdx.z * p.M * p.ldc;
    else {
      if (p.C && !p.c_is_bf16) c32 = static_cast<float*>(p.C);
      if (p.C && p.c_is_bf16) { c16 = static_cast<__nv_bfloat16*>(p.C); ld16 = p.ldc; }
      else if (p.C16) c16 = p.C16;
    }
    const bool v32 = c32 && (ld32 & 3) == 0 && (reinterpret_cast<uintptr_t>(c32) & 15) == 0;
    const bool v16 = c16 && (ld16 & 3) == 0 && (reinterpret_cast<uintptr_t>(c16) & 7) == 0;
    for (int i = threadIdx.x; i < GM * 32; i += 512) {
      const int rr = i >> 5, c = (i & 31) * 4;
      const int mm = m0 + rr, n = n0 + c;
      if (mm >= p.M || n >= p.N) continue;
      const float4 o = *reinterpret_cast<const float4*>(tile + rr * TP + c);
      if (c32) {
        float* d = c32 + (long long)mm * ld32 + n;
        if (v32) *reinterpret_cast<float4*>(d) = o;
        else { d[0] = o.x; d[1] = o.y; d[2] = o.z; d[3] = o.w; }
      }
      if (c16) {
        __nv_bfloat16* d = c16 + (long long)mm * ld16 + n;
        if (v16) {
          __nv_bfloat162 h0 = __floats2bfloat162_rn(o.x, o.y), h1 = __floats2bfloat162_rn(o.z, o.w);
          uint2 u;
          u.x = *reinterpret_cast<uint32_t*>(&h0);
          u.y = *reinterpret_cast<uint32_t*>(&h1);
          *reinterpret_cast<uint2*>(d) = u;
        } else {
          d[0] = __float2bfloat16(o.x); d[1] = __float2bfloat16(o.y); d[2] = __float2bfloat16(o.z); d[3] = __float2bfloat16(o.w);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_d, GN);
}

// out = act(sum_z part[z] + bias) + residual   (fixed order -> deterministic)
__global__ void gemm_splitk_reduce_kernel(const float* __restrict__ part, const float* __restrict__ bias,
                                          const float* __restrict__ residual, void* __restrict__ out, int M, int N,
                                          int ldc, int splits, int act, int c_is_bf16) {
  pdl_launch();
  pdl_wait();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)M * N) return;
  const int m = (int)(i / N), n = (int)(i - (long long)m * N);
  float v = 0.f;
  for (int z = 0; z < splits; ++z) v += part[((long long)z * M + m) * N + n];
  if (bias) v += bias[n];
  if (act == OCV_ACT_RELU) v = fmaxf(v, 0.f);
  else if (act == OCV_ACT_LEAKY_RELU) v = v > 0.f ? v : 0.01f * v;
  if (residual) v += residual[(long long)m * ldc + n];
  if (c_is_bf16) static_cast<__nv_bfloat16*>(out)[(long long)m * ldc + n] = __float2bfloat16(v);
  else static_cast<float*>(out)[(long long)m * ldc + n] = v;
}

}  // namespace

extern "C" long long ocv_gemm_bf16_workspace(int M, int N, int K, int split_k) {
  if (split_k <= 1 || M <= 0 || N <= 0) return 0;
  return (long long)split_k * M * N * (long long)sizeof(float);
}

static int gemm_launch(const void* A, const void* Wt, int M, int N, int K, int lda, int ldb, GemmParams p, int split_k,
                       void* workspace, void* C_final, int ldc_final, int c_final_bf16, void* stream);

extern "C" int ocv_gemm_bf16_ex(const void* A, const void* Wt, const float* bias, const float* residual, int ldr,
                                const float* ln_gamma, const float* ln_beta, const uint8_t* zero_rows, float* C32,
                                int ldc32, void* C16, int ldc16, int M, int N, int K, int lda, int ldb, int act,
                                void* stream) {
  OCV_CHECK_ARG(A && Wt && (C32 || C16) && M > 0 && N > 0 && K > 0);
  OCV_CHECK_ARG((ln_gamma == nullptr) == (ln_beta == nullptr));
  OCV_CHECK_SHAPE(K % 8 == 0 && lda % 8 == 0 && ldb % 8 == 0 && N % 4 == 0);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(bias) & 15) == 0);
  if (ln_gamma) {
    OCV_CHECK_SHAPE(N == GN && act == OCV_ACT_NONE);
    OCV_CHECK_ALIGN((!C32 || ldc32 % 4 == 0) && (!C16 || ldc16 % 4 == 0) && (!residual || ldr % 4 == 0));
    OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(bias) | reinterpret_cast<uintptr_t>(residual) |
                      reinterpret_cast<uintptr_t>(ln_gamma) | reinterpret_cast<uintptr_t>(ln_beta) |
                      reinterpret_cast<uintptr_t>(C32)) & 15) == 0 && (reinterpret_cast<uintptr_t>(C16) & 7) == 0);
  }
  GemmParams p{};
  p.bias = bias; p.residual = residual; p.ldr = ldr; p.gamma = ln_gamma; p.beta = ln_beta; p.zero_rows = zero_rows;
  p.M = M; p.N = N; p.K = K; p.act = act;
  if (C32) {
    p.C = C32; p.ldc = ldc32; p.c_is_bf16 = 0;
    p.C16 = static_cast<__nv_bfloat16*>(C16); p.ldc16 = ldc16;
  } else if (ln_gamma) {
    p.C = nullptr; p.C16 = static_cast<__nv_bfloat16*>(C16); p.ldc16 = ldc16;
  } else {
    p.C = C16; p.ldc = ldc16; p.c_is_bf16 = 1;
  }
  return gemm_launch(A, Wt, M, N, K, lda, ldb, p, 1, nullptr, nullptr, 0, 0, stream);
}

extern "C" int ocv_gemm_bf16_v2(const void* A, const void* Bt, int a_mn_major, int b_mn_major, const float* bias,
                                const float* residual, int ldr, const float* ln_gamma, const float* ln_beta, float* Cpre, int ldpre,
                                float* C32, int ldc32, void* C16, int ldc16, int M, int N, int K, int lda, int ldb, int act,
                                float drop_p, unsigned long long drop_seed, int split_k, void* workspace, void* stream) {
  OCV_CHECK_ARG(A && Bt && (C32 || C16) && M > 0 && N > 0 && K > 0 && drop_p >= 0.f && drop_p < 1.f);
  OCV_CHECK_ARG((ln_gamma == nullptr) == (ln_beta == nullptr) && (!Cpre || ln_gamma));
  OCV_CHECK_SHAPE(lda % 8 == 0 && ldb % 8 == 0 && N % 4 == 0 && (a_mn_major ? M % 8 == 0 : K % 8 == 0));
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(bias) & 15) == 0);
  if (split_k < 1) split_k = 1;
  if (split_k > 1) OCV_CHECK_ARG(workspace && C32 && !C16 && !ln_gamma && drop_p == 0.f && (!residual || ldr == ldc32));
  if (ln_gamma) {
    OCV_CHECK_SHAPE(N == GN && act == OCV_ACT_NONE);
    OCV_CHECK_ALIGN((!C32 || ldc32 % 4 == 0) && (!C16 || ldc16 % 4 == 0) && (!residual || ldr % 4 == 0));
    OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(bias) | reinterpret_cast<uintptr_t>(residual) |
                      reinterpret_cast<uintptr_t>(ln_gamma) | reinterpret_cast<uintptr_t>(ln_beta) |
                      reinterpret_cast<uintptr_t>(C32)) & 15) == 0 && (reinterpret_cast<uintptr_t>(C16) & 7) == 0);
  }
  GemmParams p{};
  p.bias = bias; p.residual = residual; p.ldr = ldr; p.gamma = ln_gamma; p.beta = ln_beta;
  p.M = M; p.N = N; p.K = K; p.act = act;
  p.a_mn = a_mn_major ? 1 : 0; p.b_mn = b_mn_major ? 1 : 0;
  p.Cpre = Cpre; p.ldpre = ldpre;
  if (drop_p > 0.f) {
    p.drop_thr = ocv_drop_threshold(drop_p);
    p.drop_scale = 1.f / (1.f - drop_p);
    p.drop_seed = drop_seed;
    p.drop_salt = ocv_dropout_salt();
  }
  if (C32) {
    p.C = C32; p.ldc = ldc32; p.c_is_bf16 = 0;
    p.C16 = static_cast<__nv_bfloat16*>(C16); p.ldc16 = ldc16;
  } else if (ln_gamma) {
    p.C = nullptr; p.C16 = static_cast<__nv_bfloat16*>(C16); p.ldc16 = ldc16;
  } else {
    p.C = C16; p.ldc = ldc16; p.c_is_bf16 = 1;
  }
  return gemm_launch(A, Bt, M, N, K, lda, ldb, p, split_k, workspace, C32, ldc32, 0, stream);
}

extern "C" int ocv_gemm_bf16(const void* A, const void* Wt, const float* bias, const float* residual, void* C, int M,
                             int N, int K, int lda, int ldb, int ldc, int act, int c_is_bf16, int split_k,
                             void* workspace, void* stream) {
  OCV_CHECK_ARG(A && Wt && C && M > 0 && N > 0 && K > 0);
  OCV_CHECK_SHAPE(K % 8 == 0 && lda % 8 == 0 && ldb % 8 == 0 && N % 4 == 0);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(bias) & 15) == 0);
  if (split_k < 1) split_k = 1;
  OCV_CHECK_ARG(split_k == 1 || workspace);
  GemmParams p{};
  p.bias = bias;
  p.residual = residual;
  p.ldr = ldc;
  p.M = M; p.N = N; p.K = K; p.act = act;
  p.C = C; p.ldc = ldc; p.c_is_bf16 = c_is_bf16;
  return gemm_launch(A, Wt, M, N, K, lda, ldb, p, split_k, workspace, C, ldc, c_is_bf16, stream);
}

static int gemm_launch(const void* A, const void* Wt, int M, int N, int K, int lda, int ldb, GemmParams p, int split_k,
                       void* workspace, void* C_final, int ldc_final, int c_final_bf16, void* stream) {
  if (int s = ocv_require_sm100()) return s;
  CUtensorMap tmA, tmB;
  // K-major operand: tensor [rows = M|N, cols = K], box 128 x 64.  MN-major operand: tensor [rows = K, cols = M|N], box 64 x 64.
  if (p.a_mn) { if (int s = tc::make_tmap_2d_bf16(&tmA, A, K, M, lda, GK, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s; }
  else if (int s = tc::make_tmap_2d_bf16(&tmA, A, M, K, lda, GM, GK, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  if (p.b_mn) { if (int s = tc::make_tmap_2d_bf16(&tmB, Wt, K, N, ldb, GK, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s; }
  else if (int s = tc::make_tmap_2d_bf16(&tmB, Wt, N, K, ldb, GN, GK, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  dim3 grid(ceil_div(N, GN), ceil_div(M, GM), split_k);
  if (split_k > 1) {
    p.C = workspace;
    p.ldc = N;
    p.c_is_bf16 = 0;
    p.k_per_split = ceil_div(ceil_div(K, split_k), GK) * GK;
  } else {
    p.k_per_split = 0;
  }
  if (int s = ocv_launch_pdl(gemm_bf16_kernel, grid, dim3(512), GEMM_SMEM, stream, tmA, tmB, p)) return s;
  if (split_k > 1) {
    const long long n = (long long)M * N;
    return ocv_launch_pdl(gemm_splitk_reduce_kernel, dim3((unsigned)ceil_div<long long>(n, 256)), dim3(256), 0, stream,
                          static_cast<const float*>(workspace), p.bias, p.residual, C_final, M, N, ldc_final, split_k,
                          p.act, c_final_bf16);
  }
  return OCV_OK;
}
