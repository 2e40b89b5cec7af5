// Fused pixel kernel of the depth head (bf16 tensor-core path), one persistent tcgen05 kernel:
//
//   conv3x3 (implicit GEMM over an NHWC bf16 halo tile)  ->  logits = Mb[b] @ conv  (Mb = log2e * Wout @ Q[b])
//   ->  per-pixel softmax over the 256 bins  ->  depth = sum_k p_k * centre_k
//
// replaces conv3x3 + PixelWiseDotProduct + conv_out + Softmax + sum(out*centers)
//   (modules/ObjCAViT.py:374-375, modules/GraphBins.py:109,119; modules/miniViT.py:25,31, modules/AdaBins.py:77,87).
// Neither the [B,128,h,w] range-attention maps nor the [B,256,h,w] probability volume ever exist in HBM.
//
// Tile = 8 x 16 pixels (M = 128 TMEM lanes).  Per tile the CTA
//   1. TMA-loads the (10 x 18 pixel) x 128-channel halo once (two 64-channel planes, 128B swizzle, OOB = conv zero padding);
//      the 9 taps are 9 *shifted smem descriptors* into that halo (start + (dy*10+dx)*128 B, group stride 1280 B) —
//      the swizzle is a function of the absolute smem address, so row shifts need no re-load (probes/probe_desc_shift.cu);
//   2. streams the 18 (tap, k-half) weight slabs [128 x 64] through an mbarrier ring and issues 72 tcgen05.mma
//      (M128 N128 K16) into one of two conv accumulators in TMEM;
//   3. the epilogue warps read the accumulator (tcgen05.ld), round to bf16 and write it as the K-major A operand
//      of the logits GEMM (swizzled smem staging);
//   4. 16 more tcgen05.mma (4 slabs of Mb[b]) give the 128 x 256 logits in TMEM;
//   5. the epilogue warps run an online softmax over the 256 columns of their own TMEM lane (one thread = one
//      pixel) and store one fp32 depth per pixel.
// Warp roles: 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator, 4..7 = convert + softmax epilogue.
// The conv accumulator is double buffered, so conv(i+1) runs on the tensor pipe while the epilogue of tile i
// and the conversion of tile i+1 run on the SIMT pipes.
#include "tc_common.cuh"

namespace {

using namespace tc;

constexpr int C = 128;                 // channels in = channels out of conv3x3 = E
constexpr int NB = 256;                // bins
constexpr int TW = 8, TH = 16;         // tile (pixels): M index m = ty*8 + tx
constexpr int HW_ = TW + 2, HH = TH + 2;
constexpr int HALO_PLANE = ((HW_ * HH * 128 + 1023) / 1024) * 1024;   // 23040 -> 23552 B per 64-channel plane
constexpr int HALO_BYTES = 2 * HALO_PLANE;
constexpr int HALO_TX = 2 * HW_ * HH * 128;                            // bytes TMA delivers per halo
constexpr int SLAB = 128 * 64 * 2;     // 16 KB: [128 rows x 64 k] bf16, 128B swizzle
constexpr int NS = 6;                  // slab ring stages
constexpr int STG_BYTES = 2 * SLAB;    // bf16 conv tile as A operand: two K-halves
constexpr int N_EPI = 128;
constexpr int SMEM_BYTES = 2 * HALO_BYTES + NS * SLAB + STG_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t TM_CONV0 = 0, TM_LOGITS = 256, TM_COLS = 512;

struct PixParams {
  const float* blog;     // [B,256]  log2e * (bout + Wout Q b3)
  const float* centres;  // [B,256]
  float* depth;          // [B,h*w]
  int B, h, w, tiles_x, tiles_y, n_tiles;
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr, uint32_t sbo_bytes) {
  return make_smem_desc(saddr, 16, sbo_bytes, LAYOUT_SW128);
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

struct TileCoord { int b, y0, x0; };
__device__ __forceinline__ TileCoord tile_coord(const PixParams& p, int t) {
  const int per = p.tiles_x * p.tiles_y;
  const int b = t / per, r = t - b * per;
  const int ty = r / p.tiles_x;
  return {b, ty * TH, (r - ty * p.tiles_x) * TW};
}

__global__ void __launch_bounds__(256, 1)
pixel_fused_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                   const __grid_constant__ CUtensorMap tmMb, PixParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* halo = smem;                                // [2][2 planes]
  uint8_t* ring = halo + 2 * HALO_BYTES;               // [NS] slabs
  uint8_t* stg = ring + NS * SLAB;                     // [2 k-halves] bf16 conv tile
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg + STG_BYTES);
  uint64_t* w_full = bars;                  // [NS]
  uint64_t* w_empty = w_full + NS;          // [NS]
  uint64_t* halo_full = w_empty + NS;       // [2]
  uint64_t* halo_empty = halo_full + 2;     // [2]
  uint64_t* conv_full = halo_empty + 2;     // [2]
  uint64_t* conv_empty = conv_full + 2;     // [2]
  uint64_t* stg_full = conv_empty + 2;
  uint64_t* stg_empty = stg_full + 1;
  uint64_t* log_full = stg_empty + 1;
  uint64_t* log_empty = log_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(log_empty + 1);

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  // contiguous chunk of tiles per CTA (neighbouring tiles share halo rows in L2 and the per-image operands)
  const int per_cta = (p.n_tiles + gridDim.x - 1) / gridDim.x;
  const int t_begin = min(p.n_tiles, (int)blockIdx.x * per_cta);
  const int t_end = min(p.n_tiles, t_begin + per_cta);
  const int n_my = t_end - t_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmMb);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(&w_full[s], 1);
      mbar_init(&w_empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&halo_full[i], 1);
      mbar_init(&halo_empty[i], 1);
      mbar_init(&conv_full[i], 1);
      mbar_init(&conv_empty[i], N_EPI);
    }
    mbar_init(stg_full, N_EPI);
    mbar_init(stg_empty, 1);
    mbar_init(log_full, 1);
    mbar_init(log_empty, N_EPI);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, TM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one() && n_my > 0) {
      uint32_t ring_n = 0;  // slabs issued so far
      auto load_halo = [&](int it) {
        const int buf = it & 1;
        mbar_wait(&halo_empty[buf], ((it >> 1) & 1) ^ 1);
        const TileCoord tc_ = tile_coord(p, t_begin + it);
        mbar_arrive_expect_tx(&halo_full[buf], HALO_TX);
        uint8_t* dst = halo + buf * HALO_BYTES;
        tma_load_4d(dst, &tmX, &halo_full[buf], 0, tc_.x0 - 1, tc_.y0 - 1, tc_.b);
        tma_load_4d(dst + HALO_PLANE, &tmX, &halo_full[buf], 64, tc_.x0 - 1, tc_.y0 - 1, tc_.b);
      };
      auto load_slab = [&](const CUtensorMap* m, int c0, int c1) {
        const int s = ring_n % NS;
        mbar_wait(&w_empty[s], ((ring_n / NS) & 1) ^ 1);
        mbar_arrive_expect_tx(&w_full[s], SLAB);
        tma_load_2d(ring + s * SLAB, m, &w_full[s], c0, c1);
        ++ring_n;
      };
      auto load_logit_slabs = [&](int it) {
        const int b = tile_coord(p, t_begin + it).b;
        for (int nh = 0; nh < 2; ++nh)
          for (int kh = 0; kh < 2; ++kh) load_slab(&tmMb, kh * 64, b * NB + nh * 128);
      };
      load_halo(0);
      for (int it = 0; it < n_my; ++it) {
        if (it + 1 < n_my) load_halo(it + 1);
        for (int tap = 0; tap < 9; ++tap)
          for (int kh = 0; kh < 2; ++kh) load_slab(&tmW, kh * 64, tap * C);
        if (it > 0) load_logit_slabs(it - 1);
      }
      load_logit_slabs(n_my - 1);
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one() && n_my > 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 128);
      uint32_t ring_n = 0;
      auto logits_mma = [&](int j) {
        mbar_wait(stg_full, j & 1);
        mbar_wait(log_empty, (j & 1) ^ 1);
        tc_fence_after();
        const uint32_t a_base = smem_u32(stg);
        for (int nh = 0; nh < 2; ++nh)
          for (int kh = 0; kh < 2; ++kh) {
            const int s = ring_n % NS;
            mbar_wait(&w_full[s], (ring_n / NS) & 1);
            tc_fence_after();
            const uint32_t b_base = smem_u32(ring + s * SLAB);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_bf16(tmem + TM_LOGITS + nh * 128, desc_sw128(a_base + kh * SLAB + kk * 32, 1024),
                        desc_sw128(b_base + kk * 32, 1024), idesc, (kh | kk) != 0);
            umma_commit(&w_empty[s]);
            ++ring_n;
          }
        umma_commit(stg_empty);
        umma_commit(log_full);
      };
      for (int it = 0; it < n_my; ++it) {
        const int buf = it & 1;
        mbar_wait(&conv_empty[buf], ((it >> 1) & 1) ^ 1);
        mbar_wait(&halo_full[buf], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t h_base = smem_u32(halo + buf * HALO_BYTES);
        const uint32_t acc = tmem + TM_CONV0 + buf * 128;
        for (int tap = 0; tap < 9; ++tap) {
          const int dy = tap / 3, dx = tap - dy * 3;
          for (int kh = 0; kh < 2; ++kh) {
            const int s = ring_n % NS;
            mbar_wait(&w_full[s], (ring_n / NS) & 1);
            tc_fence_after();
            const uint32_t a_base = h_base + kh * HALO_PLANE + (dy * HW_ + dx) * 128;
            const uint32_t b_base = smem_u32(ring + s * SLAB);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_bf16(acc, desc_sw128(a_base + kk * 32, HW_ * 128), desc_sw128(b_base + kk * 32, 1024), idesc,
                        (tap | kh | kk) != 0);
            umma_commit(&w_empty[s]);
            ++ring_n;
          }
        }
        umma_commit(&halo_empty[buf]);
        umma_commit(&conv_full[buf]);
        if (it > 0) logits_mma(it - 1);
      }
      logits_mma(n_my - 1);
    }
  } else if (warp >= 4) {
    // ================= convert + softmax epilogue (one thread = one pixel = one TMEM lane) =================
    const int q = warp - 4;
    const int m = q * 32 + lane;
    const int ty = m >> 3, tx = m & 7;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    auto convert = [&](int it) {
      const int buf = it & 1;
      mbar_wait(&conv_full[buf], (it >> 1) & 1);
      tc_fence_after();
      uint32_t packed[64];
#pragma unroll
      for (int c4 = 0; c4 < 4; ++c4) {
        uint32_t r[32];
        tmem_ld32(tmem + lane_base + TM_CONV0 + buf * 128 + c4 * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) packed[c4 * 16 + j] = pack_bf16(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
      }
      tc_fence_before();
      mbar_arrive(&conv_empty[buf]);
      mbar_wait(stg_empty, (it & 1) ^ 1);
      uint8_t* row = stg + m * 128;
#pragma unroll
      for (int kh = 0; kh < 2; ++kh)
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          uint4 v = make_uint4(packed[(kh * 8 + ch) * 4], packed[(kh * 8 + ch) * 4 + 1], packed[(kh * 8 + ch) * 4 + 2],
                               packed[(kh * 8 + ch) * 4 + 3]);
          *reinterpret_cast<uint4*>(row + kh * SLAB + ((ch ^ (m & 7)) << 4)) = v;
        }
      fence_proxy_async_smem();
      mbar_arrive(stg_full);
    };
    auto softmax_store = [&](int j) {
      const TileCoord tc_ = tile_coord(p, t_begin + j);
      const float4* bl = reinterpret_cast<const float4*>(p.blog + (long long)tc_.b * NB);
      const float4* ce = reinterpret_cast<const float4*>(p.centres + (long long)tc_.b * NB);
      mbar_wait(log_full, j & 1);
      tc_fence_after();
      float mx = -INFINITY, s = 0.f, a = 0.f;
#pragma unroll 1
      for (int c = 0; c < NB / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(tmem + lane_base + TM_LOGITS + c * 32, r);
        tmem_ld_wait();
        float v[32];
        float cm = -INFINITY;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 b4 = __ldg(bl + c * 8 + j4);
          v[4 * j4 + 0] = __uint_as_float(r[4 * j4 + 0]) + b4.x;
          v[4 * j4 + 1] = __uint_as_float(r[4 * j4 + 1]) + b4.y;
          v[4 * j4 + 2] = __uint_as_float(r[4 * j4 + 2]) + b4.z;
          v[4 * j4 + 3] = __uint_as_float(r[4 * j4 + 3]) + b4.w;
          cm = fmaxf(cm, fmaxf(fmaxf(v[4 * j4], v[4 * j4 + 1]), fmaxf(v[4 * j4 + 2], v[4 * j4 + 3])));
        }
        const float m_new = fmaxf(mx, cm);
        const float sc = ex2(mx - m_new);
        s *= sc;
        a *= sc;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 c4 = __ldg(ce + c * 8 + j4);
          const float e0 = ex2(v[4 * j4 + 0] - m_new), e1 = ex2(v[4 * j4 + 1] - m_new);
          const float e2 = ex2(v[4 * j4 + 2] - m_new), e3 = ex2(v[4 * j4 + 3] - m_new);
          s += (e0 + e1) + (e2 + e3);
          a = fmaf(e0, c4.x, a);
          a = fmaf(e1, c4.y, a);
          a = fmaf(e2, c4.z, a);
          a = fmaf(e3, c4.w, a);
        }
        mx = m_new;
      }
      tc_fence_before();
      mbar_arrive(log_empty);
      const int y = tc_.y0 + ty, x = tc_.x0 + tx;
      if (y < p.h && x < p.w) p.depth[((long long)tc_.b * p.h + y) * p.w + x] = a / s;
    };
    for (int it = 0; it < n_my; ++it) {
      convert(it);
      if (it > 0) softmax_store(it - 1);
    }
    if (n_my > 0) softmax_store(n_my - 1);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, TM_COLS);
}

int make_tmap_features(CUtensorMap* out, const void* X, int B, int h, int w) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return OCV_E_DRIVER;
  if (reinterpret_cast<uintptr_t>(X) & 15) return OCV_E_ALIGN;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)w * C * 2, (cuuint64_t)h * w * C * 2};
  cuuint32_t box[4] = {64, HW_, HH, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(X), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OCV_OK : OCV_E_DRIVER;
}

}  // namespace

extern "C" int ocv_pixel_fused_bf16(const void* X, const void* W3p, const void* Mb, const float* blog,
                                    const float* centres, float* depth, int B, int h, int w, int max_ctas,
                                    void* stream) {
  OCV_CHECK_ARG(X && W3p && Mb && blog && centres && depth && B > 0 && h > 0 && w > 0);
  OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(blog) | reinterpret_cast<uintptr_t>(centres)) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  CUtensorMap tmX, tmW, tmMb;
  if (int s = make_tmap_features(&tmX, X, B, h, w)) return s;
  if (int s = tc::make_tmap_2d_bf16(&tmW, W3p, 9 * C, C, C, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  if (int s = tc::make_tmap_2d_bf16(&tmMb, Mb, (uint64_t)B * NB, C, C, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaError_t e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(pixel_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) { n_sm = 0; return (int)e; }
  }
  PixParams p{};
  p.blog = blog; p.centres = centres; p.depth = depth;
  p.B = B; p.h = h; p.w = w;
  p.tiles_x = ceil_div(w, TW);
  p.tiles_y = ceil_div(h, TH);
  const long long nt = (long long)p.tiles_x * p.tiles_y * B;
  OCV_CHECK_SHAPE(nt < (1ll << 30));
  p.n_tiles = (int)nt;
  int grid = n_sm;
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (grid > p.n_tiles) grid = p.n_tiles;
  pixel_fused_kernel<<<grid, 256, SMEM_BYTES, ocv_stream(stream)>>>(tmX, tmW, tmMb, p);
  return ocv_launch_status();
}
