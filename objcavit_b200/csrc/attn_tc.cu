// Attention core of nn.MultiheadAttention on tcgen05 (bf16 operands, fp32 accumulation in TMEM), d = 32, 4 heads, Sk <= 448.
//
// One CTA = 128 query rows of one image x one PAIR of heads.  Two 32-wide heads share a 128-byte row, so Q, K and V are TMA
// boxes [128 rows x 64 columns] under the 128-byte swizzle and a head is a +0 / +64 byte start offset of the UMMA descriptors
// (probes/probe_mn_major.cu, cases 4-6).  Per head:
//   S = Q K^T          M128 x N=Sk x K32: A = Q tile, B = all keys (both K-major), one or two MMAs of N <= 256 into TMEM
//   softmax            256 threads: thread = query row (TMEM lane) x half of the keys; row max and row sum merged through shared
//                      memory; masked keys (key_padding_mask, or beyond Sk) are -inf; an all-masked row gives NaN like torch
//   O = P V            P (bf16, un-normalised, optionally dropped out) goes through a ring of [128 x 64-key] K-major tiles; V is
//                      read where TMA put it as the MN-major B operand (N = 32 channels, K = key rows): no transposed copy
//   out = O / rowsum   bf16 (and optionally fp32) rows, optional softmax statistics for a backward pass
// replaces: the attention core inside nn.TransformerEncoderLayer.self_attn and the two cross-attention MHAs
// (modules/ObjCAViT.py:155-169, 188, 195-207; modules/layers.py:8-9, 23).  Contract: include/objcavit_b200.h.
#include "tc_common.cuh"
#include <math.h>

namespace {

using namespace tc;

constexpr int QT = 128, HD = 32, SKMAX = 448, CHK = 64;
constexpr int SLAB = 128 * 128;              // [128 rows x 64 bf16]
constexpr int KV_BOXES = 4;                  // 512 key rows of shared memory each for K and V
constexpr int N_SOFT = 256, THREADS = 128 + N_SOFT;
constexpr uint32_t TM_S = 0, TM_O = 448, TM_COLS = 512;
constexpr int SMEM = SLAB /*Q*/ + 2 * KV_BOXES * SLAB /*K, V*/ + 4 * SLAB /*P ring: 2 halves x 2 slots*/ + 512 * 4 /*mask*/ +
                     4 * 128 * 4 /*row max / sum exchange*/ + 1024 + 256;

struct AttnParams {
  __nv_bfloat16* out16;
  float* out32;
  float* stats;            // [B, heads, Sq, 2] (row max, row sum of 2^(s - max)) in the log2 domain, or NULL
  const uint8_t* mask;     // [B, Sk] 1 = padding, or NULL
  int Sq, Sk, heads;
  long long o_bs;
  int o_rs;
  float qk_scale;          // log2(e) / sqrt(d)
  uint32_t drop_thr;
  float drop_scale;
  unsigned long long drop_seed;
  const unsigned long long* drop_salt;
};

__device__ __forceinline__ void tma_load_3d_(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  tma_load_3d(smem_dst, m, bar, c0, c1, c2);
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void soft_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

__global__ void __launch_bounds__(THREADS, 1)
mha_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
              const __grid_constant__ CUtensorMap tmV, AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* Qs = smem;
  uint8_t* Ks = Qs + SLAB;
  uint8_t* Vs = Ks + KV_BOXES * SLAB;
  uint8_t* Ps = Vs + KV_BOXES * SLAB;                                  // [half][slot]
  uint32_t* valid = reinterpret_cast<uint32_t*>(Ps + 4 * SLAB);       // [16] bit j of word w: key 32 w + j takes part (not masked, < Sk)
  float* red = reinterpret_cast<float*>(valid) + 512;                 // [2 (max | sum)][2 halves][128 rows]
  uint64_t* bars = reinterpret_cast<uint64_t*>(red + 4 * 128);
  uint64_t* qk_full = bars;
  uint64_t* v_full = qk_full + 1;
  uint64_t* m_full = v_full + 1;
  uint64_t* s_full = m_full + 1;
  uint64_t* s_empty = s_full + 1;
  uint64_t* o_full = s_empty + 1;
  uint64_t* o_empty = o_full + 1;
  uint64_t* p_full = o_empty + 1;      // [2][2]
  uint64_t* p_empty = p_full + 4;      // [2][2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(p_empty + 4);

  pdl_launch();
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int q0 = blockIdx.x * QT, hp = blockIdx.y, b = blockIdx.z;
  const int Sk = p.Sk, Skp = (Sk + 15) & ~15;
  const int nchunks = (Skp + CHK - 1) / CHK;
  const int nc0 = (nchunks + 1) / 2;                                   // chunks of key half 0; half 1 takes the rest
  const int nb = (Sk + 127) / 128;                                     // K / V boxes

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(qk_full, 1);
    mbar_init(v_full, 1);
    mbar_init(m_full, 32);
    mbar_init(s_full, 1);
    mbar_init(s_empty, N_SOFT);
    mbar_init(o_full, 1);
    mbar_init(o_empty, 128);
    for (int i = 0; i < 4; ++i) {
      mbar_init(&p_full[i], 128);
      mbar_init(&p_empty[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, TM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      pdl_wait();   // q / k / v are the previous kernel's output
      mbar_arrive_expect_tx(qk_full, (1 + nb) * SLAB);
      tma_load_3d_(Qs, &tmQ, qk_full, hp * 64, q0, b);
      for (int i = 0; i < nb; ++i) tma_load_3d_(Ks + i * SLAB, &tmK, qk_full, hp * 64, i * 128, b);
      mbar_arrive_expect_tx(v_full, nb * SLAB);
      for (int i = 0; i < nb; ++i) tma_load_3d_(Vs + i * SLAB, &tmV, v_full, hp * 64, i * 128, b);
    }
  } else if (warp == 3) {
    pdl_wait();
    for (int w = 0; w < 16; ++w) {
      const int k = w * 32 + lane;
      const bool ok = k < Sk && !(p.mask && p.mask[(long long)b * Sk + k]);
      const uint32_t bits = __ballot_sync(0xffffffffu, ok);
      if (lane == 0) valid[w] = bits;
    }
    mbar_arrive(m_full);
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t qa = smem_u32(Qs), ka = smem_u32(Ks), va = smem_u32(Vs), pa = smem_u32(Ps);
      const uint32_t id_pv = make_idesc_bf16(128, HD, 0, 1);
      uint32_t cons[2] = {0, 0};
      mbar_wait(qk_full, 0);
      tc_fence_after();
      for (int hd = 0; hd < 2; ++hd) {
        const uint32_t hb = hd * 64;                                   // byte offset of the head inside a 128-byte row
        if (hd == 1) {
          mbar_wait(s_empty, 0);                                       // head 0's scores have been read out of TMEM
          tc_fence_after();
        }
        // S = Q K^T: K rows are contiguous across the TMA boxes, so one descriptor spans up to 256 keys
        for (int n0 = 0; n0 < Skp; n0 += 256) {
          const int N = min(256, Skp - n0);
          const uint32_t id = make_idesc_bf16(128, N, 0, 0);
#pragma unroll
          for (int kk = 0; kk < 2; ++kk)
            umma_bf16(tmem + TM_S + n0, desc_kmajor_sw128(qa + hb + kk * 32), desc_kmajor_sw128(ka + n0 * 128 + hb + kk * 32), id, kk);
        }
        umma_commit(s_full);
        if (hd == 0) {
          mbar_wait(v_full, 0);
        } else {
          mbar_wait(o_empty, 0);                                       // head 0's output has left TMEM
        }
        tc_fence_after();
        // O += P V, 64-key tiles of P in the order the two softmax halves produce them
        bool first = true;
        const int cnt[2] = {nc0, nchunks - nc0};
        for (int i = 0; i < nc0; ++i)
          for (int g = 0; g < 2; ++g) {
            if (i >= cnt[g]) continue;
            const uint32_t n = cons[g]++;
            const int slot = n & 1, c = (g ? nc0 : 0) + i;
            mbar_wait(&p_full[g * 2 + slot], (n >> 1) & 1);
            tc_fence_after();
            const uint32_t a0 = pa + (g * 2 + slot) * SLAB;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              umma_bf16(tmem + TM_O, desc_kmajor_sw128(a0 + kk * 32),
                        make_smem_desc(va + (c * CHK + kk * 16) * 128 + hb, SLAB, 1024, LAYOUT_SW128), id_pv, !first);
              first = false;
            }
            umma_commit(&p_empty[g * 2 + slot]);
          }
        umma_commit(o_full);
      }
    }
  } else if (warp >= 4) {
    const int e = warp - 4, qd = e & 3, g = e >> 2;
    const int row = qd * 32 + lane;
    const uint32_t lane_base = (uint32_t)(qd * 32) << 16;
    const int cb = g ? nc0 : 0, ce = g ? nchunks : nc0;
    const float sc = p.qk_scale;
    const unsigned long long dseed = p.drop_thr ? ocv_salted(p.drop_seed, p.drop_salt) : 0ull;
    uint32_t prod = 0;
    pdl_wait();
    mbar_wait(m_full, 0);
    for (int hd = 0; hd < 2; ++hd) {
      const int head = blockIdx.y * 2 + hd;
      mbar_wait(s_full, hd);
      tc_fence_after();
      // ---- pass 1: row maximum over this thread's half of the keys.  TMEM loads have a few hundred cycles of latency each:
      // both 32-column loads of a 64-key chunk are in flight before the one wait (half the round trips)
      float mx = -INFINITY;
      for (int c = cb; c < ce; ++c) {
        const int col0 = c * CHK;
        const bool two = col0 + 32 < Skp;
        uint32_t r0[32], r1[32];
        tmem_ld32(tmem + lane_base + TM_S + col0, r0);
        if (two) tmem_ld32(tmem + lane_base + TM_S + col0 + 32, r1);
        tmem_ld_wait();
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
          if (sub == 1 && !two) break;
          const uint32_t* r = sub ? r1 : r0;
          const uint32_t vb = valid[(col0 >> 5) + sub];          // the same word for every thread: warp-uniform branches
          if (vb == 0xffffffffu) {
#pragma unroll
            for (int j = 0; j < 32; j += 2) mx = fmaxf(mx, fmaxf(__uint_as_float(r[j]), __uint_as_float(r[j + 1])));
          } else if (vb) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if ((vb >> j) & 1u) mx = fmaxf(mx, __uint_as_float(r[j]));
          }
        }
      }
      red[g * 128 + row] = mx;
      soft_bar();
      const float m_row = fmaxf(red[row], red[128 + row]);
      const float m_sc = m_row * sc;
      // ---- pass 2: p = 2^(s*c - m*c); row sum of the un-dropped p; P tiles -> shared memory
      float sum = 0.f;
      const unsigned long long drow = (((unsigned long long)b * p.heads + head) * p.Sq + (unsigned long long)min(q0 + row, p.Sq - 1)) *
                                      (unsigned long long)Sk;
      for (int c = cb; c < ce; ++c) {
        const uint32_t n = prod++;
        const int slot = n & 1;
        uint32_t pk[32];
        uint32_t ra[32], rb2[32];
        const bool two = c * CHK + 32 < Skp;
        tmem_ld32(tmem + lane_base + TM_S + c * CHK, ra);
        if (two) tmem_ld32(tmem + lane_base + TM_S + c * CHK + 32, rb2);
        tmem_ld_wait();
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
          const int col0 = c * CHK + sub * 32;
          if (sub == 0 || two) {
            const uint32_t* r = sub ? rb2 : ra;
            const uint32_t vb = (m_row != -INFINITY) ? valid[col0 >> 5] : 0u;
            if (vb == 0xffffffffu && !p.drop_thr) {          // the common case: no masked key in the group, inference
              float s0 = 0.f, s1 = 0.f;
#pragma unroll
              for (int j = 0; j < 32; j += 2) {
                const float e0 = ex2(fmaf(__uint_as_float(r[j]), sc, -m_sc)), e1 = ex2(fmaf(__uint_as_float(r[j + 1]), sc, -m_sc));
                s0 += e0;
                s1 += e1;
                pk[sub * 16 + j / 2] = pack_bf16(e0, e1);
              }
              sum += s0 + s1;
            } else {
#pragma unroll
              for (int j = 0; j < 32; j += 2) {
                float e0 = 0.f, e1 = 0.f;
                if ((vb >> j) & 1u) e0 = ex2(fmaf(__uint_as_float(r[j]), sc, -m_sc));
                if ((vb >> (j + 1)) & 1u) e1 = ex2(fmaf(__uint_as_float(r[j + 1]), sc, -m_sc));
                sum += e0 + e1;
                if (p.drop_thr) {
                  e0 = ocv_drop_keep(dseed, drow + col0 + j, p.drop_thr) ? e0 * p.drop_scale : 0.f;
                  e1 = ocv_drop_keep(dseed, drow + col0 + j + 1, p.drop_thr) ? e1 * p.drop_scale : 0.f;
                }
                pk[sub * 16 + j / 2] = pack_bf16(e0, e1);
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) pk[sub * 16 + j] = 0u;
          }
        }
        if (c == ce - 1) {                     // last read of the scores by this thread
          tc_fence_before();
          mbar_arrive(s_empty);
        }
        mbar_wait(&p_empty[g * 2 + slot], ((n >> 1) & 1) ^ 1);
        uint8_t* prow = Ps + (g * 2 + slot) * SLAB + row * 128;
#pragma unroll
        for (int ch = 0; ch < 8; ++ch)
          *reinterpret_cast<uint4*>(prow + ((ch ^ (row & 7)) << 4)) = make_uint4(pk[ch * 4], pk[ch * 4 + 1], pk[ch * 4 + 2], pk[ch * 4 + 3]);
        fence_proxy_async_smem();
        mbar_arrive(&p_full[g * 2 + slot]);
      }
      if (cb == ce) {                          // a half without keys (Sk <= 64): nothing read, still release the scores
        tc_fence_before();
        mbar_arrive(s_empty);
      }
      red[256 + g * 128 + row] = sum;
      soft_bar();
      if (g == 0) {
        const float tot = red[256 + row] + red[384 + row];
        mbar_wait(o_full, hd);
        tc_fence_after();
        uint32_t r[32];
        tmem_ld32(tmem + lane_base + TM_O, r);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(o_empty);
        const int qi = q0 + row;
        if (qi < p.Sq) {
          const float inv = 1.f / tot;         // every key masked: 0 * inf = NaN, like torch's softmax over all -inf
          const long long o = (long long)b * p.o_bs + (long long)qi * p.o_rs + head * HD;
          uint4* d16 = reinterpret_cast<uint4*>(p.out16 + o);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            d16[j] = make_uint4(pack_bf16(__uint_as_float(r[8 * j]) * inv, __uint_as_float(r[8 * j + 1]) * inv),
                                pack_bf16(__uint_as_float(r[8 * j + 2]) * inv, __uint_as_float(r[8 * j + 3]) * inv),
                                pack_bf16(__uint_as_float(r[8 * j + 4]) * inv, __uint_as_float(r[8 * j + 5]) * inv),
                                pack_bf16(__uint_as_float(r[8 * j + 6]) * inv, __uint_as_float(r[8 * j + 7]) * inv));
          if (p.out32) {
            float4* d32 = reinterpret_cast<float4*>(p.out32 + o);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              d32[j] = make_float4(__uint_as_float(r[4 * j]) * inv, __uint_as_float(r[4 * j + 1]) * inv,
                                   __uint_as_float(r[4 * j + 2]) * inv, __uint_as_float(r[4 * j + 3]) * inv);
          }
          if (p.stats) {
            float2* st = reinterpret_cast<float2*>(p.stats) + ((long long)b * p.heads + head) * p.Sq + qi;
            *st = make_float2(m_sc, tot);
          }
        }
      }
      soft_bar();   // the exchange buffers are reused by the next head
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, TM_COLS);
}

// rows of [*, 128] bf16 with arbitrary row / batch strides (elements): (64-column slice, row, image)
int make_tmap_rows(CUtensorMap* out, const void* base, int S, int B, long long bs, int rs) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return OCV_E_DRIVER;
  cuuint64_t dims[3] = {128, (cuuint64_t)S, (cuuint64_t)B};
  cuuint64_t strides[2] = {(cuuint64_t)rs * 2, (cuuint64_t)(B > 1 ? bs : (long long)rs * S) * 2};
  cuuint32_t box[3] = {64, 128, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OCV_OK : OCV_E_DRIVER;
}

}  // namespace

// OCV_ATTN_TC: "0" = never take the tcgen05 attention kernels; "1" = take them wherever the shape is covered; unset = the measured
// policy: cross attention of the inference head and every attention core of the training step on tcgen05, the inference encoders'
// self attention on the mma.sync kernel (csrc/attn_bf16.cu), which is 4-6 us faster per launch at S = 300 / 418 (both are
// latency-bound there: profiles/attn_tc_vs_mma_r02.txt).
int ocv_attn_tc_policy() {
  static int policy = -1;
  if (policy < 0) {
    const char* e = getenv("OCV_ATTN_TC");
    policy = !e ? 2 : (e[0] == '0' ? 0 : 1);
  }
  return policy;
}

// true when the tcgen05 kernel covers the shape (else the caller keeps the mma.sync kernel)
bool ocv_mha_tc_supported(int Sq, int Sk, int heads, int d, long long q_bs, int q_rs, long long k_bs, int k_rs, long long v_bs,
                          int v_rs, long long o_bs, int o_rs) {
  return ocv_attn_tc_policy() != 0 && heads == 4 && d == HD && Sk >= 1 && Sk <= SKMAX && Sq >= 1 &&
         ((q_bs | k_bs | v_bs | q_rs | k_rs | v_rs) & 7) == 0 && ((o_bs | o_rs) & 7) == 0;
}

extern "C" int ocv_mha_core_tc_bf16(const void* q, const void* k, const void* v, void* out16, float* out32, float* stats,
                                    const uint8_t* key_padding_mask, int B, int Sq, int Sk, int heads, int d, long long q_bstride,
                                    int q_rstride, long long k_bstride, int k_rstride, long long v_bstride, int v_rstride,
                                    long long o_bstride, int o_rstride, float drop_p, unsigned long long drop_seed, void* stream) {
  OCV_CHECK_ARG(q && k && v && out16 && B > 0 && Sq > 0 && Sk > 0 && drop_p >= 0.f && drop_p < 1.f);
  OCV_CHECK_SHAPE(ocv_mha_tc_supported(Sq, Sk, heads, d, q_bstride, q_rstride, k_bstride, k_rstride, v_bstride, v_rstride, o_bstride,
                                       o_rstride) && B <= 65535);
  const uintptr_t al = reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v) |
                       reinterpret_cast<uintptr_t>(out16) | reinterpret_cast<uintptr_t>(out32);
  OCV_CHECK_ALIGN((al & 15) == 0 && (reinterpret_cast<uintptr_t>(stats) & 7) == 0);
  if (int s = ocv_require_sm100()) return s;
  CUtensorMap tmQ, tmK, tmV;
  if (int s = make_tmap_rows(&tmQ, q, Sq, B, q_bstride, q_rstride)) return s;
  if (int s = make_tmap_rows(&tmK, k, Sk, B, k_bstride, k_rstride)) return s;
  if (int s = make_tmap_rows(&tmV, v, Sk, B, v_bstride, v_rstride)) return s;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(mha_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  AttnParams p{};
  p.out16 = static_cast<__nv_bfloat16*>(out16); p.out32 = out32; p.stats = stats; p.mask = key_padding_mask;
  p.Sq = Sq; p.Sk = Sk; p.heads = heads; p.o_bs = o_bstride; p.o_rs = o_rstride;
  p.qk_scale = 1.4426950408889634f / sqrtf((float)d);
  if (drop_p > 0.f) {
    p.drop_thr = ocv_drop_threshold(drop_p);
    p.drop_scale = 1.f / (1.f - drop_p);
    p.drop_seed = drop_seed;
    p.drop_salt = ocv_dropout_salt();
  }
  return ocv_launch_pdl(mha_tc_kernel, dim3(ceil_div(Sq, QT), heads / 2, B), dim3(THREADS), SMEM, stream, tmQ, tmK, tmV, p);
}

// ======================================================= backward =======================================================
// One CTA = one (image, head), Sq, Sk <= 256 (two 128-row tiles each: the training shapes 221 / 242 / 32).  For every
// (key tile u, query tile t):
//   S = Q_t K_u^T and dPd = dO_t V_u^T        two small MMAs into TMEM (K = 32)
//   epilogue (thread = query row x 64 keys)    P = 2^(c s - m) / sum from the forward's statistics, dropout mask regenerated,
//                                              Pd = P o D', dS = P o (dPd o D' - delta), delta = rowsum(dO o O)
//                                              -> two bf16 tiles [128 q x 128 k] in shared memory
//   dV_u += Pd^T dO_t, dK_u += dS^T Q_t        the tiles read as MN-major A operands (M = keys), dO / Q as MN-major B (N = 32)
//   dQ_t += dS K_u                             dS K-major, K_u as MN-major B
// dV / dK accumulate in TMEM over the query tiles, dQ over the key tiles; nothing of size Sq x Sk touches HBM.
namespace {

constexpr int BW_MAXS = 256;
constexpr uint32_t TB_S = 0, TB_DP = 128, TB_DV = 256, TB_DK = 288, TB_DQ = 320, TB_COLS = 512;
constexpr int BW_SMEM = 8 * SLAB /*Q, K, V, dO: two tiles each*/ + 4 * SLAB /*Pd, dS tiles*/ + 3 * BW_MAXS * 4 /*m, 1/sum, delta*/ +
                        64 /*valid bits*/ + 1024 + 256;

struct AttnBwdParams {
  const __nv_bfloat16* o16;      // forward output [B,Sq,128] (for delta)
  const __nv_bfloat16* do16;     // its gradient (same strides)
  const float* stats;            // [B,heads,Sq,2] from the forward
  const uint8_t* mask;
  float *dq32, *dk32, *dv32;     // fp32 gradients (strides g_bs / g_rs) — and their bf16 copies
  __nv_bfloat16 *dq16, *dk16, *dv16;
  int Sq, Sk, heads;
  long long o_bs, g_bs;
  int o_rs, g_rs;
  float qk_scale, scale;         // log2(e)/sqrt(d), 1/sqrt(d)
  uint32_t drop_thr;
  float drop_scale;
  unsigned long long drop_seed;
  const unsigned long long* drop_salt;
};

__global__ void __launch_bounds__(THREADS, 1)
mha_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO, AttnBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* Qs = smem;                    // [2 tiles]
  uint8_t* Ks = Qs + 2 * SLAB;
  uint8_t* Vs = Ks + 2 * SLAB;
  uint8_t* DOs = Vs + 2 * SLAB;
  uint8_t* Pt = DOs + 2 * SLAB;          // Pd tile: [2 key halves] of [128 q rows x 64 keys]
  uint8_t* St = Pt + 2 * SLAB;           // dS tile
  float* rm = reinterpret_cast<float*>(St + 2 * SLAB);   // [256] row max (log2 domain)
  float* ri = rm + BW_MAXS;                               // [256] 1 / row sum
  float* dl = ri + BW_MAXS;                               // [256] delta
  uint32_t* valid = reinterpret_cast<uint32_t*>(dl + BW_MAXS);   // [8 words]
  uint64_t* bars = reinterpret_cast<uint64_t*>(valid + 16);
  uint64_t* ld_full = bars;
  uint64_t* sd_full = ld_full + 1;
  uint64_t* sd_empty = sd_full + 1;
  uint64_t* tile_full = sd_empty + 1;
  uint64_t* tile_empty = tile_full + 1;
  uint64_t* kv_full = tile_empty + 1;
  uint64_t* kv_empty = kv_full + 1;
  uint64_t* q_full = kv_empty + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_full + 1);

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int head = blockIdx.x, b = blockIdx.y;
  const int hp = head >> 1;
  const uint32_t hb = (head & 1) * 64;
  const int Sq = p.Sq, Sk = p.Sk;
  const int nt = (Sq + 127) / 128, nu = (Sk + 127) / 128;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(ld_full, 1);
    mbar_init(sd_full, 1);
    mbar_init(sd_empty, N_SOFT);
    mbar_init(tile_full, N_SOFT);
    mbar_init(tile_empty, 1);
    mbar_init(kv_full, 1);
    mbar_init(kv_empty, N_SOFT);
    mbar_init(q_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, TB_COLS);
  if (warp == 3) {
    for (int w = 0; w < 8; ++w) {
      const int k = w * 32 + lane;
      const bool ok = k < Sk && !(p.mask && p.mask[(long long)b * Sk + k]);
      const uint32_t bits = __ballot_sync(0xffffffffu, ok);
      if (lane == 0) valid[w] = bits;
    }
  }
  if (warp >= 4) {
    // per-row statistics of the forward and delta = rowsum(dO o O)
    const int i = threadIdx.x - 128;
    if (i < Sq) {
      const float2 st = reinterpret_cast<const float2*>(p.stats)[((long long)b * p.heads + head) * Sq + i];
      rm[i] = st.x;
      ri[i] = 1.f / st.y;
      const long long o = (long long)b * p.o_bs + (long long)i * p.o_rs + head * HD;
      const uint4* po = reinterpret_cast<const uint4*>(p.o16 + o);
      const uint4* pd = reinterpret_cast<const uint4*>(p.do16 + o);
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint4 a = __ldg(po + j), g4 = __ldg(pd + j);
        const __nv_bfloat162* ah = reinterpret_cast<const __nv_bfloat162*>(&a);
        const __nv_bfloat162* gh = reinterpret_cast<const __nv_bfloat162*>(&g4);
#pragma unroll
        for (int t2 = 0; t2 < 4; ++t2) {
          const float2 x = __bfloat1622float2(ah[t2]), y = __bfloat1622float2(gh[t2]);
          acc = fmaf(x.x, y.x, fmaf(x.y, y.y, acc));
        }
      }
      dl[i] = acc;
    } else if (i < BW_MAXS) {
      rm[i] = 0.f;
      ri[i] = 0.f;
      dl[i] = 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      mbar_arrive_expect_tx(ld_full, (2 * nt + 2 * nu) * SLAB);
      for (int t = 0; t < nt; ++t) {
        tma_load_3d(Qs + t * SLAB, &tmQ, ld_full, hp * 64, t * 128, b);
        tma_load_3d(DOs + t * SLAB, &tmDO, ld_full, hp * 64, t * 128, b);
      }
      for (int u = 0; u < nu; ++u) {
        tma_load_3d(Ks + u * SLAB, &tmK, ld_full, hp * 64, u * 128, b);
        tma_load_3d(Vs + u * SLAB, &tmV, ld_full, hp * 64, u * 128, b);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t qa = smem_u32(Qs), ka = smem_u32(Ks), va = smem_u32(Vs), da = smem_u32(DOs), pt = smem_u32(Pt), st = smem_u32(St);
      constexpr uint32_t id_s = make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t id_nn = make_idesc_bf16(128, HD, 1, 1);
      constexpr uint32_t id_kn = make_idesc_bf16(128, HD, 0, 1);
      mbar_wait(ld_full, 0);
      tc_fence_after();
      uint32_t n = 0;
      for (int u = 0; u < nu; ++u) {
        mbar_wait(kv_empty, (u & 1) ^ 1);               // the previous key tile's dV / dK have been drained
        tc_fence_after();
        for (int t = 0; t < nt; ++t, ++n) {
          mbar_wait(sd_empty, (n & 1) ^ 1);
          tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < 2; ++kk) {
            umma_bf16(tmem + TB_S, desc_kmajor_sw128(qa + t * SLAB + hb + kk * 32), desc_kmajor_sw128(ka + u * SLAB + hb + kk * 32), id_s, kk);
            umma_bf16(tmem + TB_DP, desc_kmajor_sw128(da + t * SLAB + hb + kk * 32), desc_kmajor_sw128(va + u * SLAB + hb + kk * 32), id_s, kk);
          }
          umma_commit(sd_full);
          mbar_wait(tile_full, n & 1);
          tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            // dV_u += Pd^T dO_t ; dK_u += dS^T Q_t : A = the tile as MN-major (M = keys: two 64-key blocks SLAB apart), K = query rows
            umma_bf16(tmem + TB_DV, make_smem_desc(pt + kk * 2048, SLAB, 1024, LAYOUT_SW128),
                      make_smem_desc(da + t * SLAB + kk * 2048 + hb, SLAB, 1024, LAYOUT_SW128), id_nn, (t | kk) != 0);
            umma_bf16(tmem + TB_DK, make_smem_desc(st + kk * 2048, SLAB, 1024, LAYOUT_SW128),
                      make_smem_desc(qa + t * SLAB + kk * 2048 + hb, SLAB, 1024, LAYOUT_SW128), id_nn, (t | kk) != 0);
            // dQ_t += dS K_u : A = dS K-major (keys: two slabs x four steps), B = K_u MN-major (N = 32, K = key rows)
            umma_bf16(tmem + TB_DQ + t * HD, desc_kmajor_sw128(st + (kk >> 2) * SLAB + (kk & 3) * 32),
                      make_smem_desc(ka + u * SLAB + kk * 2048 + hb, SLAB, 1024, LAYOUT_SW128), id_kn, (u | kk) != 0);
          }
          umma_commit(tile_empty);
        }
        umma_commit(kv_full);
      }
      umma_commit(q_full);
    }
  } else if (warp >= 4) {
    const int e = warp - 4, qd = e & 3, g = e >> 2;
    const int row = qd * 32 + lane;
    const uint32_t lane_base = (uint32_t)(qd * 32) << 16;
    const float c = p.qk_scale;
    const unsigned long long dseed = p.drop_thr ? ocv_salted(p.drop_seed, p.drop_salt) : 0ull;
    uint32_t n = 0;
    for (int u = 0; u < nu; ++u) {
      for (int t = 0; t < nt; ++t, ++n) {
        const int r = t * 128 + row;
        const bool rok = r < Sq;
        const float m_r = rm[min(r, BW_MAXS - 1)], inv_r = ri[min(r, BW_MAXS - 1)], d_r = dl[min(r, BW_MAXS - 1)];
        const unsigned long long drow = (((unsigned long long)b * p.heads + head) * Sq + (unsigned long long)min(r, Sq - 1)) *
                                        (unsigned long long)Sk;
        mbar_wait(sd_full, n & 1);
        tc_fence_after();
        uint32_t s0[32], s1[32], g0[32], g1[32];
        tmem_ld32(tmem + lane_base + TB_S + g * 64, s0);
        tmem_ld32(tmem + lane_base + TB_S + g * 64 + 32, s1);
        tmem_ld32(tmem + lane_base + TB_DP + g * 64, g0);
        tmem_ld32(tmem + lane_base + TB_DP + g * 64 + 32, g1);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(sd_empty);
        mbar_wait(tile_empty, (n & 1) ^ 1);            // the previous step's gradient MMAs have read the tiles
        uint8_t* prow = Pt + g * SLAB + row * 128;
        uint8_t* srow = St + g * SLAB + row * 128;
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          uint32_t pp[4], ss[4];
#pragma unroll
          for (int j2 = 0; j2 < 4; ++j2) {
            float pd[2], ds[2];
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {
              const int j = ch * 8 + j2 * 2 + h2;                       // key within this thread's 64
              const int k = u * 128 + g * 64 + j;
              const float sv = __uint_as_float(j < 32 ? s0[j] : s1[j - 32]);
              const float gv = __uint_as_float(j < 32 ? g0[j] : g1[j - 32]);
              const bool kv = (valid[k >> 5] >> (k & 31)) & 1u;
              float pr = (rok && kv) ? ex2(fmaf(sv, c, -m_r)) * inv_r : 0.f;
              float keep = 1.f;
              if (p.drop_thr) keep = ocv_drop_keep(dseed, drow + k, p.drop_thr) ? p.drop_scale : 0.f;
              pd[h2] = pr * keep;
              ds[h2] = pr * (gv * keep - d_r);
            }
            pp[j2] = pack_bf16(pd[0], pd[1]);
            ss[j2] = pack_bf16(ds[0], ds[1]);
          }
          *reinterpret_cast<uint4*>(prow + ((ch ^ (row & 7)) << 4)) = make_uint4(pp[0], pp[1], pp[2], pp[3]);
          *reinterpret_cast<uint4*>(srow + ((ch ^ (row & 7)) << 4)) = make_uint4(ss[0], ss[1], ss[2], ss[3]);
        }
        fence_proxy_async_smem();
        mbar_arrive(tile_full);
      }
      // dV_u (warpgroup 0) and dK_u (warpgroup 1): lane = key row
      mbar_wait(kv_full, u & 1);
      tc_fence_after();
      {
        uint32_t r32[32];
        tmem_ld32(tmem + lane_base + (g == 0 ? TB_DV : TB_DK), r32);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(kv_empty);
        const int key = u * 128 + row;
        if (key < Sk) {
          const float f = g == 0 ? 1.f : p.scale;
          const long long o = (long long)b * p.g_bs + (long long)key * p.g_rs + head * HD;
          float4* d32 = reinterpret_cast<float4*>((g == 0 ? p.dv32 : p.dk32) + o);
          uint4* d16 = reinterpret_cast<uint4*>((g == 0 ? p.dv16 : p.dk16) + o);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            d32[j] = make_float4(f * __uint_as_float(r32[4 * j]), f * __uint_as_float(r32[4 * j + 1]), f * __uint_as_float(r32[4 * j + 2]),
                                 f * __uint_as_float(r32[4 * j + 3]));
#pragma unroll
          for (int j = 0; j < 4; ++j)
            d16[j] = make_uint4(pack_bf16(f * __uint_as_float(r32[8 * j]), f * __uint_as_float(r32[8 * j + 1])),
                                pack_bf16(f * __uint_as_float(r32[8 * j + 2]), f * __uint_as_float(r32[8 * j + 3])),
                                pack_bf16(f * __uint_as_float(r32[8 * j + 4]), f * __uint_as_float(r32[8 * j + 5])),
                                pack_bf16(f * __uint_as_float(r32[8 * j + 6]), f * __uint_as_float(r32[8 * j + 7])));
        }
      }
    }
    // dQ: warpgroup g drains query tile g
    mbar_wait(q_full, 0);
    tc_fence_after();
    if (g < nt) {
      uint32_t r32[32];
      tmem_ld32(tmem + lane_base + TB_DQ + g * HD, r32);
      tmem_ld_wait();
      const int r = g * 128 + row;
      if (r < Sq) {
        const float f = p.scale;
        const long long o = (long long)b * p.g_bs + (long long)r * p.g_rs + head * HD;
        float4* d32 = reinterpret_cast<float4*>(p.dq32 + o);
        uint4* d16 = reinterpret_cast<uint4*>(p.dq16 + o);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          d32[j] = make_float4(f * __uint_as_float(r32[4 * j]), f * __uint_as_float(r32[4 * j + 1]), f * __uint_as_float(r32[4 * j + 2]),
                               f * __uint_as_float(r32[4 * j + 3]));
#pragma unroll
        for (int j = 0; j < 4; ++j)
          d16[j] = make_uint4(pack_bf16(f * __uint_as_float(r32[8 * j]), f * __uint_as_float(r32[8 * j + 1])),
                              pack_bf16(f * __uint_as_float(r32[8 * j + 2]), f * __uint_as_float(r32[8 * j + 3])),
                              pack_bf16(f * __uint_as_float(r32[8 * j + 4]), f * __uint_as_float(r32[8 * j + 5])),
                              pack_bf16(f * __uint_as_float(r32[8 * j + 6]), f * __uint_as_float(r32[8 * j + 7])));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, TB_COLS);
}

}  // namespace

bool ocv_mha_bwd_tc_supported(int Sq, int Sk, int heads, int d) {
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("OCV_ATTN_BWD_TC");
    enabled = (e && e[0] == '0') ? 0 : 1;
  }
  return enabled && heads == 4 && d == HD && Sq >= 1 && Sq <= BW_MAXS && Sk >= 1 && Sk <= BW_MAXS;
}

extern "C" int ocv_mha_core_bwd_tc_bf16(const void* q16, const void* k16, const void* v16, const void* o16, const void* do16,
                                        const float* stats, const uint8_t* key_padding_mask, float* dq32, float* dk32, float* dv32,
                                        void* dq16, void* dk16, void* dv16, int B, int Sq, int Sk, int heads, int d,
                                        long long q_bstride, int q_rstride, long long kv_bstride, int kv_rstride, long long o_bstride,
                                        int o_rstride, long long g_bstride, int g_rstride, float drop_p, unsigned long long drop_seed,
                                        void* stream) {
  OCV_CHECK_ARG(q16 && k16 && v16 && o16 && do16 && stats && dq32 && dk32 && dv32 && dq16 && dk16 && dv16 && B > 0 && drop_p >= 0.f &&
                drop_p < 1.f);
  OCV_CHECK_SHAPE(ocv_mha_bwd_tc_supported(Sq, Sk, heads, d) && B <= 65535);
  OCV_CHECK_SHAPE(((q_bstride | kv_bstride | o_bstride | g_bstride | q_rstride | kv_rstride | o_rstride | g_rstride) & 7) == 0);
  const uintptr_t al = reinterpret_cast<uintptr_t>(q16) | reinterpret_cast<uintptr_t>(k16) | reinterpret_cast<uintptr_t>(v16) |
                       reinterpret_cast<uintptr_t>(o16) | reinterpret_cast<uintptr_t>(do16) | reinterpret_cast<uintptr_t>(dq32) |
                       reinterpret_cast<uintptr_t>(dk32) | reinterpret_cast<uintptr_t>(dv32) | reinterpret_cast<uintptr_t>(dq16) |
                       reinterpret_cast<uintptr_t>(dk16) | reinterpret_cast<uintptr_t>(dv16);
  OCV_CHECK_ALIGN((al & 15) == 0 && (reinterpret_cast<uintptr_t>(stats) & 7) == 0);
  if (int s = ocv_require_sm100()) return s;
  CUtensorMap tmQ, tmK, tmV, tmDO;
  if (int s = make_tmap_rows(&tmQ, q16, Sq, B, q_bstride, q_rstride)) return s;
  if (int s = make_tmap_rows(&tmK, k16, Sk, B, kv_bstride, kv_rstride)) return s;
  if (int s = make_tmap_rows(&tmV, v16, Sk, B, kv_bstride, kv_rstride)) return s;
  if (int s = make_tmap_rows(&tmDO, do16, Sq, B, o_bstride, o_rstride)) return s;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(mha_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BW_SMEM);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  AttnBwdParams p{};
  p.o16 = static_cast<const __nv_bfloat16*>(o16); p.do16 = static_cast<const __nv_bfloat16*>(do16); p.stats = stats;
  p.mask = key_padding_mask;
  p.dq32 = dq32; p.dk32 = dk32; p.dv32 = dv32;
  p.dq16 = static_cast<__nv_bfloat16*>(dq16); p.dk16 = static_cast<__nv_bfloat16*>(dk16); p.dv16 = static_cast<__nv_bfloat16*>(dv16);
  p.Sq = Sq; p.Sk = Sk; p.heads = heads; p.o_bs = o_bstride; p.o_rs = o_rstride; p.g_bs = g_bstride; p.g_rs = g_rstride;
  p.scale = 1.f / sqrtf((float)d);
  p.qk_scale = 1.4426950408889634f * p.scale;
  if (drop_p > 0.f) {
    p.drop_thr = ocv_drop_threshold(drop_p);
    p.drop_scale = 1.f / (1.f - drop_p);
    p.drop_seed = drop_seed;
    p.drop_salt = ocv_dropout_salt();
  }
  mha_bwd_tc_kernel<<<dim3(heads, B), THREADS, BW_SMEM, ocv_stream(stream)>>>(tmQ, tmK, tmV, tmDO, p);
  return ocv_launch_status();
}
