// Attention core of nn.MultiheadAttention on the bf16 token path: flash-style, one CTA = 64 query rows of one
// (image, head), keys streamed in tiles of 64 through shared memory, scores and P@V on the tensor cores
// (mma.sync m16n8k16 bf16 -> fp32; the P fragments are re-used in registers as the A operand of P@V), online
// softmax in fp32 with ex2.  d = 32 (4 heads x 32 in every reference config).  S <= 1200 here, so this stage is
// <1 % of the head's FLOPs: the warp-level MMA keeps it simple; the large contractions use tcgen05.
// Contract: include/objcavit_b200.h (ocv_mha_core_bf16).
#include "common.cuh"
#include <math.h>

namespace {

constexpr int QT = 64, KT = 64, D = 32, PITCH = 40;  // 80-byte rows: 16B-aligned and ldmatrix conflict-free

__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], const void* p) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], const void* p) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// rows [r0, r0+64) x 32 channels of a [*, rs]-pitched bf16 matrix -> smem tile (zero rows past `rows`)
__device__ __forceinline__ void load_tile(__nv_bfloat16 (*dst)[PITCH], const __nv_bfloat16* src, int r0, int rows, int rs) {
  for (int i = threadIdx.x; i < 64 * 4; i += 128) {
    const int r = i >> 2, c = (i & 3) * 8;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r0 + r < rows) v = *reinterpret_cast<const uint4*>(src + (long long)(r0 + r) * rs + c);
    *reinterpret_cast<uint4*>(&dst[r][c]) = v;
  }
}

__global__ void __launch_bounds__(128)
mha_bf16_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k,
                const __nv_bfloat16* __restrict__ v, __nv_bfloat16* __restrict__ out,
                const uint8_t* __restrict__ mask, int Sq, int Sk, long long q_bs, int q_rs, long long k_bs, int k_rs,
                long long v_bs, int v_rs, long long o_bs, int o_rs) {
  // The whole K and V of this (image, head) are staged ONCE (Sk <= ~1300 keys x 32 dims: <= 200 KB): one global-memory
  // round trip per CTA instead of one per key tile — the stage is latency-bound, not bandwidth- or FLOP-bound.
  extern __shared__ __align__(16) uint8_t att_smem[];
  const int Skp = (Sk + KT - 1) / KT * KT;
  __nv_bfloat16 (*Qs)[PITCH] = reinterpret_cast<__nv_bfloat16 (*)[PITCH]>(att_smem);
  __nv_bfloat16 (*Kall)[PITCH] = Qs + QT;
  __nv_bfloat16 (*Vall)[PITCH] = Kall + Skp;
  float* Mall = reinterpret_cast<float*>(Vall + Skp);
  pdl_launch();
  pdl_wait();  // q, k, v come from the previous kernel of the chain
  const int b = blockIdx.z, hd = blockIdx.y, q0 = blockIdx.x * QT;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const __nv_bfloat16* qb = q + b * q_bs + hd * D;
  const __nv_bfloat16* kb = k + b * k_bs + hd * D;
  const __nv_bfloat16* vb = v + b * v_bs + hd * D;
  // cp.async (LDGSTS): every 16-byte piece of Q, K and V is in flight at once, no staging registers
  auto cp16 = [](void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
  };
  for (int i = threadIdx.x; i < QT * 4; i += 128) {
    const int r = i >> 2, c = (i & 3) * 8;
    if (q0 + r < Sq) cp16(&Qs[r][c], qb + (long long)(q0 + r) * q_rs + c);
    else *reinterpret_cast<uint4*>(&Qs[r][c]) = make_uint4(0, 0, 0, 0);
  }
  for (int i = threadIdx.x; i < Skp * 4; i += 128) {
    const int r = i >> 2, c = (i & 3) * 8;
    if (r < Sk) {
      cp16(&Kall[r][c], kb + (long long)r * k_rs + c);
      cp16(&Vall[r][c], vb + (long long)r * v_rs + c);
    } else {
      *reinterpret_cast<uint4*>(&Kall[r][c]) = make_uint4(0, 0, 0, 0);
      *reinterpret_cast<uint4*>(&Vall[r][c]) = make_uint4(0, 0, 0, 0);
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int j = threadIdx.x; j < Skp; j += blockDim.x)
    Mall[j] = (j >= Sk || (mask && mask[(long long)b * Sk + j])) ? 1.f : 0.f;
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  uint32_t qa[2][4];
  {
    const int mi = lane >> 3;
    const int row = warp * 16 + (mi & 1) * 8 + (lane & 7);
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) ldsm4(qa[ks], &Qs[row][ks * 16 + (mi >> 1) * 8]);
  }
  const float scale = rsqrtf((float)D) * 1.4426950408889634f;
  float m_i[2] = {-INFINITY, -INFINITY}, l_i[2] = {0.f, 0.f};
  float o[4][4] = {};
  for (int k0 = 0; k0 < Sk; k0 += KT) {
    __nv_bfloat16 (*Ks)[PITCH] = Kall + k0;
    __nv_bfloat16 (*Vs)[PITCH] = Vall + k0;
    const float* Msk = Mall + k0;
    float s[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
      uint32_t kf[4];
      ldsm4(kf, &Ks[8 * j + (lane & 7)][(lane >> 3) * 8]);
      mma16816(s[j], qa[0], kf[0], kf[1]);
      mma16816(s[j], qa[1], kf[2], kf[3]);
    }
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const bool dead = Msk[8 * j + 2 * t + (e & 1)] != 0.f;
        s[j][e] = dead ? -INFINITY : s[j][e] * scale;
        mx[e >> 1] = fmaxf(mx[e >> 1], s[j][e]);
      }
    float corr[2], msafe[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float m_new = fmaxf(m_i[r], mx[r]);
      msafe[r] = (m_new == -INFINITY) ? 0.f : m_new;
      corr[r] = ex2f(m_i[r] - msafe[r]);   // m_i = -inf -> 0
      m_i[r] = m_new;
      l_i[r] *= corr[r];
    }
#pragma unroll
    for (int jd = 0; jd < 4; ++jd) {
      o[jd][0] *= corr[0]; o[jd][1] *= corr[0];
      o[jd][2] *= corr[1]; o[jd][3] *= corr[1];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        s[j][e] = ex2f(s[j][e] - msafe[e >> 1]);
        l_i[e >> 1] += s[j][e];
      }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      uint32_t pa[4] = {pack2(s[2 * kk][0], s[2 * kk][1]), pack2(s[2 * kk][2], s[2 * kk][3]),
                        pack2(s[2 * kk + 1][0], s[2 * kk + 1][1]), pack2(s[2 * kk + 1][2], s[2 * kk + 1][3])};
      const int mi = lane >> 3;
      const int row = 16 * kk + (mi & 1) * 8 + (lane & 7);
#pragma unroll
      for (int jp = 0; jp < 2; ++jp) {
        uint32_t vf[4];
        ldsm4t(vf, &Vs[row][(2 * jp + (mi >> 1)) * 8]);
        mma16816(o[2 * jp], pa, vf[0], vf[1]);
        mma16816(o[2 * jp + 1], pa, vf[2], vf[3]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_i[r] += __shfl_xor_sync(0xffffffffu, l_i[r], 1);
    l_i[r] += __shfl_xor_sync(0xffffffffu, l_i[r], 2);
  }
  __nv_bfloat16* ob = out + b * o_bs + hd * D;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int row = q0 + warp * 16 + g + 8 * r;
    if (row < Sq) {
      const float inv = 1.f / l_i[r];  // every key masked -> 0 * inf = NaN, like torch's softmax over all -inf
#pragma unroll
      for (int jd = 0; jd < 4; ++jd)
        *reinterpret_cast<uint32_t*>(ob + (long long)row * o_rs + 8 * jd + 2 * t) =
            pack2(o[jd][2 * r] * inv, o[jd][2 * r + 1] * inv);
    }
  }
}

}  // namespace

extern "C" int ocv_mha_core_bf16(const void* q, const void* k, const void* v, void* out, const uint8_t* key_padding_mask,
                                 int B, int Sq, int Sk, int heads, int d, long long q_bstride, int q_rstride,
                                 long long k_bstride, int k_rstride, long long v_bstride, int v_rstride,
                                 long long o_bstride, int o_rstride, void* stream) {
  OCV_CHECK_ARG(q && k && v && out && B > 0 && Sq > 0 && Sk > 0 && heads > 0);
  OCV_CHECK_SHAPE(d == D && B <= 65535 && heads <= 65535);
  const uintptr_t al = reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v);
  OCV_CHECK_ALIGN((al & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 3) == 0);
  OCV_CHECK_ALIGN(((q_bstride | k_bstride | v_bstride | q_rstride | k_rstride | v_rstride) & 7) == 0 &&
                  ((o_bstride | o_rstride) & 1) == 0);
  if (int s = ocv_require_sm100()) return s;
  // OCV_ATTN_TC=1: the tcgen05 kernel for every covered shape (d = 32, 4 heads, Sk <= 448); by default this entry point keeps the
  // mma.sync kernel, which is the faster one for self attention at the reference sizes (see ocv_attn_tc_policy)
  if (ocv_attn_tc_policy() == 1 &&
      ocv_mha_tc_supported(Sq, Sk, heads, d, q_bstride, q_rstride, k_bstride, k_rstride, v_bstride, v_rstride, o_bstride, o_rstride) &&
      (reinterpret_cast<uintptr_t>(out) & 15) == 0)
    return ocv_mha_core_tc_bf16(q, k, v, out, nullptr, nullptr, key_padding_mask, B, Sq, Sk, heads, d, q_bstride, q_rstride, k_bstride,
                                k_rstride, v_bstride, v_rstride, o_bstride, o_rstride, 0.f, 0ull, stream);
  dim3 grid(ceil_div(Sq, QT), heads, B);
  const int Skp = ceil_div(Sk, KT) * KT;
  const size_t smem = (size_t)(QT + 2 * Skp) * PITCH * sizeof(__nv_bfloat16) + (size_t)Skp * sizeof(float);
  OCV_CHECK_SHAPE(smem <= 227 * 1024);
  static size_t smem_set = 0;
  if (smem > 48 * 1024 && smem > smem_set) {
    cudaError_t e = cudaFuncSetAttribute(mha_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    smem_set = smem;
  }
  return ocv_launch_pdl(mha_bf16_kernel, grid, dim3(128), smem, stream, static_cast<const __nv_bfloat16*>(q),
                        static_cast<const __nv_bfloat16*>(k), static_cast<const __nv_bfloat16*>(v),
                        static_cast<__nv_bfloat16*>(out), key_padding_mask, Sq, Sk, q_bstride, q_rstride, k_bstride,
                        k_rstride, v_bstride, v_rstride, o_bstride, o_rstride);
}
