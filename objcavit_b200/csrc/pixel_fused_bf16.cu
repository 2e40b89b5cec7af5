// Fused pixel kernel of the depth head (bf16 tensor-core path), one persistent tcgen05 kernel:
//
//   conv3x3 (implicit GEMM over an NHWC bf16 halo tile)  ->  logits = Mb[b] @ conv  (Mb = log2e * Wout @ Q[b])
//   ->  per-pixel softmax over the 256 bins  ->  depth = sum_k p_k * centre_k
//
// replaces conv3x3 + PixelWiseDotProduct + conv_out + Softmax + sum(out*centers)
//   (modules/ObjCAViT.py:374-375, modules/GraphBins.py:109,119; modules/miniViT.py:25,31, modules/AdaBins.py:77,87).
// Neither the [B,128,h,w] range-attention maps nor the [B,256,h,w] probability volume ever exist in HBM.
//
// Work unit = a PAIR of 8 x 16 pixel tiles (a 16 x 16 pixel block: tile A = columns 0..7, tile B = columns 8..15), M = 128
// TMEM lanes each.  The kernel is bound by the L2 -> SM stream of the conv weights (288 KB per pass, they do not fit in
// shared memory next to the operands), so every weight slab is used for BOTH tiles: 8 MMAs per 16 KB slab instead of 4.
// Per pair the CTA
//   1. TMA-loads the (18 x 18 pixel) x 128-channel halo as two 64-channel planes (128B swizzle, OOB = conv zero padding);
//      the 9 taps x 2 tiles are 18 *shifted smem descriptors* into a plane (start + (dy*18+dx [+8])*128 B, 8-row group
//      stride 18*128 B): the swizzle is a function of the absolute smem address, so shifts need no re-load
//      (probes/probe_desc_shift.cu).  The k loop runs plane-major (all taps of channels 0..63, then of 64..127), so each
//      plane buffer is free half a pair early and the next pair's plane lands while this pair still computes: two plane
//      buffers give full double buffering;
//   2. streams the 18 (k-half, tap) weight slabs [128 x 64] through a 4-slot mbarrier ring and issues 144 tcgen05.mma
//      (M128 N128 K16) into the two conv accumulators in TMEM;
//   3. the epilogue warps read the accumulators (tcgen05.ld), round to bf16 and write them as the K-major A operands of
//      the logits GEMMs (two swizzled staging tiles);
//   4. per tile 16 more tcgen05.mma (4 slabs of Mb[b]) give the 128 x 256 logits in TMEM;
//   5. the epilogue warps run an online softmax over the 256 columns of their TMEM lane and store one fp32 depth per pixel.
// Warp roles: 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator, 4..11 = convert + softmax epilogue: two warpgroups,
// each thread owns one pixel row and one half of the columns (conv channels 0..63 | 64..127, bins 0..127 | 128..255);
// the two softmax halves are merged through shared memory.
// MMA order per pair i:  conv(i) plane 0 | logits B(i-1) | conv(i) plane 1 | logits A(i)   — the logits GEMM of one tile
// runs while the epilogue converts / soft-maxes the other, so the tensor pipe idles only while tile A is converted.
#include "tc_common.cuh"
#include <cstdlib>

namespace {

using namespace tc;

constexpr int C = 128;                 // channels in = channels out of conv3x3 = E
constexpr int NB = 256;                // bins
constexpr int TW = 8, TH = 16;         // tile (pixels): M index m = ty*8 + tx
constexpr int PW = 2 * TW;             // pair block width
constexpr int HW_ = PW + 2, HH = TH + 2;                               // halo 18 x 18
constexpr int PLANE_TX = HW_ * HH * 128;                               // bytes TMA delivers per 64-channel plane
constexpr int PLANE = ((PLANE_TX + 1023) / 1024) * 1024;               // 41472 -> 41984
constexpr int SLAB = 128 * 64 * 2;     // 16 KB: [128 rows x 64 k] bf16, 128B swizzle
constexpr int NS = 4;                  // slab ring slots
constexpr int STG_BYTES = 2 * SLAB;    // bf16 conv tile as A operand: two K-halves
constexpr int N_EPI = 256;             // epilogue threads (warps 4..11)
constexpr int THREADS = 128 + N_EPI;
constexpr int MERGE_BYTES = 2 * 5 * 128 * 4;
constexpr int SMEM_BYTES = 2 * PLANE + NS * SLAB + 2 * STG_BYTES + MERGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t TM_CONVA = 0, TM_CONVB = 128, TM_LOGITS = 256, TM_COLS = 512;

struct PixParams {
  const float* blog;     // [B,256]  log2e * (bout + Wout Q b3)
  const float* centres;  // [B,256]
  float* depth;          // [B,h*w]
  uint8_t* argmax;       // [B,h*w] bin with the largest logit (ARGMAX instantiation only: the parity tests' bin-agreement check)
  __nv_bfloat16* conv_out;  // CONV_ONLY instantiation: [B,h,w,128] bf16 NHWC = conv3x3(X) + conv_bias (the kernel stops there)
  const float* conv_bias;   // [128]
  int B, h, w, tiles_x, tiles_y, n_tiles;   // tiles = 16x16 pair blocks (8 x th blocks for the transposed kernel)
  int th;                                   // transposed kernel: block height (even, <= 32)
  long long* dbg_out;                       // measurement aid (OCV_PIXEL_DEBUG & 8): per CTA, cycles the MMA issuer waited per barrier kind
  int debug;                                // measurement aid (OCV_PIXEL_DEBUG): 1 = weight / Mb slabs loaded for the first block only, 2 = no softmax math, 4 = planes first block only
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

// packed fp32x2 arithmetic (FADD2 / FMUL2 / FFMA2, 3-input FMNMX3): the softmax epilogue is issue-bound, these halve it
__device__ __forceinline__ uint64_t pk2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ uint64_t pk2u(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

struct TileCoord { int b, y0, x0; };
__device__ __forceinline__ TileCoord tile_coord(const PixParams& p, int t) {
  const int per = p.tiles_x * p.tiles_y;
  const int b = t / per, r = t - b * per;
  const int ty = r / p.tiles_x;
  return {b, ty * TH, (r - ty * p.tiles_x) * PW};
}

// MODE 0: the fused head tail; 1: + per-pixel bin argmax (parity tests); 2: CONV_ONLY — the same implicit-GEMM conv3x3 pipeline
// with a bias + bf16 NHWC store epilogue and no logits / softmax stage (ocv_conv3x3_nhwc_bf16: the decoder's final conv emitting
// the head's input layout).
template <int MODE>
__global__ void __launch_bounds__(THREADS, 1)
pixel_fused_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                   const __grid_constant__ CUtensorMap tmMb, PixParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* plane = smem;                               // [2] 64-channel halo planes
  uint8_t* ring = plane + 2 * PLANE;                   // [NS] slabs
  uint8_t* stg = ring + NS * SLAB;                     // [2 tiles][2 k-halves] bf16 conv tiles
  float* merge = reinterpret_cast<float*>(stg + 2 * STG_BYTES);   // [2 parity][3][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg + 2 * STG_BYTES + MERGE_BYTES);
  uint64_t* w_full = bars;                  // [NS]
  uint64_t* w_empty = w_full + NS;          // [NS]
  uint64_t* pl_full = w_empty + NS;         // [2]
  uint64_t* pl_empty = pl_full + 2;         // [2]
  uint64_t* stg_full = pl_empty + 2;        // [2]  (tile A, tile B)
  uint64_t* stg_empty = stg_full + 2;       // [2]
  uint64_t* conv_full = stg_empty + 2;
  uint64_t* conv_empty = conv_full + 1;
  uint64_t* log_full = conv_empty + 1;
  uint64_t* log_empty = log_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(log_empty + 1);

  constexpr bool ARGMAX = MODE == 1, CONV_ONLY = MODE == 2;
  pdl_launch();
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  // contiguous chunk of pair blocks per CTA (neighbours share halo rows in L2 and the per-image operands)
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
      mbar_init(&pl_full[i], 1);
      mbar_init(&pl_empty[i], 1);
      mbar_init(&stg_full[i], N_EPI);
      mbar_init(&stg_empty[i], 1);
    }
    mbar_init(conv_full, 1);
    mbar_init(conv_empty, N_EPI);
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
    // slab order == the MMA issuer's consumption order; the plane loads are slotted in where their buffer is known free
    if (elect_one() && n_my > 0) {
      uint32_t ring_n = 0;  // slabs issued so far
      bool waited = false;
      auto load_plane = [&](int it, int kh) {
        mbar_wait(&pl_empty[kh], (it & 1) ^ 1);
        const TileCoord tc_ = tile_coord(p, t_begin + it);
        mbar_arrive_expect_tx(&pl_full[kh], PLANE_TX);
        tma_load_4d(plane + kh * PLANE, &tmX, &pl_full[kh], kh * 64, tc_.x0 - 1, tc_.y0 - 1, tc_.b);
      };
      auto load_slab = [&](const CUtensorMap* m, int c0, int c1) {
        const int s = ring_n % NS;
        mbar_wait(&w_empty[s], ((ring_n / NS) & 1) ^ 1);
        mbar_arrive_expect_tx(&w_full[s], SLAB);
        tma_load_2d(ring + s * SLAB, m, &w_full[s], c0, c1);
        ++ring_n;
      };
      auto load_logit_slabs = [&](int it) {
        if (!waited) {  // Mb is produced by the previous kernel of the chain; features and conv weights are not
          pdl_wait();
          waited = true;
        }
        const int b = tile_coord(p, t_begin + it).b;
        for (int nh = 0; nh < 2; ++nh)
          for (int kh = 0; kh < 2; ++kh) load_slab(&tmMb, kh * 64, b * NB + nh * 128);
      };
      if (CONV_ONLY) pdl_wait();  // here X is typically the previous kernel's output (layout conversion of the decoder features)
      load_plane(0, 0);
      for (int it = 0; it < n_my; ++it) {
        for (int tap = 0; tap < 5; ++tap) load_slab(&tmW, 0, tap * C);
        load_plane(it, 1);
        for (int tap = 5; tap < 9; ++tap) load_slab(&tmW, 0, tap * C);
        if (!CONV_ONLY && it > 0) load_logit_slabs(it - 1);            // logits B(it-1)
        for (int tap = 0; tap < 5; ++tap) load_slab(&tmW, 64, tap * C);
        if (it + 1 < n_my) load_plane(it + 1, 0);
        for (int tap = 5; tap < 9; ++tap) load_slab(&tmW, 64, tap * C);
        if (!CONV_ONLY) load_logit_slabs(it);                          // logits A(it)
      }
      if (!CONV_ONLY) load_logit_slabs(n_my - 1);                      // logits B(n_my-1)
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one() && n_my > 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 128);
      uint32_t ring_n = 0, log_n = 0;
      // logits of tile `tile` (0 = A, 1 = B) of pair `it`: 4 slabs of Mb[b] against the staged bf16 conv tile
      auto logits_mma = [&](int it, int tile) {
        mbar_wait(&stg_full[tile], it & 1);
        mbar_wait(log_empty, (log_n & 1) ^ 1);
        tc_fence_after();
        const uint32_t a_base = smem_u32(stg + tile * STG_BYTES);
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
        umma_commit(&stg_empty[tile]);
        umma_commit(log_full);
        ++log_n;
      };
      auto conv_half = [&](int it, int kh) {
        mbar_wait(&pl_full[kh], it & 1);
        tc_fence_after();
        const uint32_t h_base = smem_u32(plane + kh * PLANE);
        for (int tap = 0; tap < 9; ++tap) {
          const int dy = tap / 3, dx = tap - dy * 3;
          const int s = ring_n % NS;
          mbar_wait(&w_full[s], (ring_n / NS) & 1);
          tc_fence_after();
          const uint32_t a_base = h_base + (dy * HW_ + dx) * 128;
          const uint32_t b_base = smem_u32(ring + s * SLAB);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem + TM_CONVA, desc_sw128(a_base + kk * 32, HW_ * 128), desc_sw128(b_base + kk * 32, 1024), idesc,
                      (kh | tap | kk) != 0);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem + TM_CONVB, desc_sw128(a_base + TW * 128 + kk * 32, HW_ * 128),
                      desc_sw128(b_base + kk * 32, 1024), idesc, (kh | tap | kk) != 0);
          umma_commit(&w_empty[s]);
          ++ring_n;
        }
        umma_commit(&pl_empty[kh]);
      };
      for (int it = 0; it < n_my; ++it) {
        mbar_wait(conv_empty, (it & 1) ^ 1);   // both conv accumulators drained by the epilogue (pair it-1)
        tc_fence_after();
        conv_half(it, 0);
        if (!CONV_ONLY && it > 0) logits_mma(it - 1, 1);
        conv_half(it, 1);
        umma_commit(conv_full);
        if (!CONV_ONLY) logits_mma(it, 0);
      }
      if (!CONV_ONLY) logits_mma(n_my - 1, 1);
    }
  } else if (warp >= 4) {
    // ================= convert + softmax epilogue =================
    // thread = one pixel row (TMEM lane) x one column half
    const int e = warp - 4;
    const int q = e & 3;            // == warp % 4: the TMEM lane quarter this warp may read
    const int half = e >> 2;
    const int m = q * 32 + lane;
    const int ty = m >> 3, tx = m & 7;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    uint32_t job = 0;
    if (n_my > 0) pdl_wait();       // blog / centres are produced by the previous kernels of the chain
    // conv accumulator of `tile`, channels [64*half, 64*half+64) -> bf16 -> K-half `half` of the staging tile
    auto convert = [&](int it, int tile) {
      uint32_t packed[32];
      {
        const uint32_t src = tmem + lane_base + (tile ? TM_CONVB : TM_CONVA) + half * 64;
        uint32_t r0[32], r1[32];
        tmem_ld32(src, r0);
        tmem_ld_wait();
        tmem_ld32(src + 32, r1);  // in flight while the first 32 columns are packed
#pragma unroll
        for (int j = 0; j < 16; ++j) packed[j] = pack_bf16(__uint_as_float(r0[2 * j]), __uint_as_float(r0[2 * j + 1]));
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) packed[16 + j] = pack_bf16(__uint_as_float(r1[2 * j]), __uint_as_float(r1[2 * j + 1]));
      }
      mbar_wait(&stg_empty[tile], (it & 1) ^ 1);
      uint8_t* row = stg + tile * STG_BYTES + half * SLAB + m * 128;
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        uint4 v = make_uint4(packed[ch * 4], packed[ch * 4 + 1], packed[ch * 4 + 2], packed[ch * 4 + 3]);
        *reinterpret_cast<uint4*>(row + ((ch ^ (m & 7)) << 4)) = v;
      }
      fence_proxy_async_smem();
      mbar_arrive(&stg_full[tile]);
    };
    // CONV_ONLY: conv accumulator + bias -> bf16 -> 128 contiguous bytes of the NHWC output per thread
    auto convert_store = [&](int it, int tile) {
      const TileCoord tc_ = tile_coord(p, t_begin + it);
      const uint32_t src = tmem + lane_base + (tile ? TM_CONVB : TM_CONVA) + half * 64;
      uint32_t r0[32], r1[32];
      tmem_ld32(src, r0);
      tmem_ld32(src + 32, r1);
      tmem_ld_wait();
      const int y = tc_.y0 + ty, x = tc_.x0 + tile * TW + tx;
      if (y < p.h && x < p.w) {
        const float4* b4 = reinterpret_cast<const float4*>(p.conv_bias + half * 64);
        uint4* dst = reinterpret_cast<uint4*>(p.conv_out + (((long long)tc_.b * p.h + y) * p.w + x) * C + half * 64);
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          const uint32_t* r = ch < 4 ? r0 + ch * 8 : r1 + (ch - 4) * 8;
          const float4 ba = __ldg(b4 + 2 * ch), bb = __ldg(b4 + 2 * ch + 1);
          dst[ch] = make_uint4(pack_bf16(__uint_as_float(r[0]) + ba.x, __uint_as_float(r[1]) + ba.y),
                               pack_bf16(__uint_as_float(r[2]) + ba.z, __uint_as_float(r[3]) + ba.w),
                               pack_bf16(__uint_as_float(r[4]) + bb.x, __uint_as_float(r[5]) + bb.y),
                               pack_bf16(__uint_as_float(r[6]) + bb.z, __uint_as_float(r[7]) + bb.w));
        }
      }
    };
    auto softmax_store = [&](int it, int tile) {
      const TileCoord tc_ = tile_coord(p, t_begin + it);
      const float4* bl = reinterpret_cast<const float4*>(p.blog + (long long)tc_.b * NB + half * 128);
      const float4* ce = reinterpret_cast<const float4*>(p.centres + (long long)tc_.b * NB + half * 128);
      mbar_wait(log_full, job & 1);
      tc_fence_after();
      float mx = -INFINITY, best = -INFINITY;
      int bidx = 0;
      uint64_t s2a = 0, s2b = 0, a2a = 0, a2b = 0;   // two (sum, weighted sum) accumulator pairs: shorter dependency chains
      const uint32_t src = tmem + lane_base + TM_LOGITS + half * 128;
      uint32_t rb[2][32];
      tmem_ld32(src, rb[0]);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_ld_wait();
        if (c < 3) tmem_ld32(src + (c + 1) * 32, rb[(c + 1) & 1]);  // next 32 bins in flight while these are processed
        const uint32_t* r = rb[c & 1];
        uint64_t v2[16];
        float cm = -INFINITY;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 b4 = __ldg(bl + c * 8 + j4);
          v2[2 * j4] = add2(pk2u(r[4 * j4], r[4 * j4 + 1]), pk2(b4.x, b4.y));
          v2[2 * j4 + 1] = add2(pk2u(r[4 * j4 + 2], r[4 * j4 + 3]), pk2(b4.z, b4.w));
          float x0, x1, x2, x3;
          upk2(v2[2 * j4], x0, x1);
          upk2(v2[2 * j4 + 1], x2, x3);
          cm = max3(cm, x0, x1);
          cm = max3(cm, x2, x3);
          if (ARGMAX) {
            const int col = half * 128 + c * 32 + 4 * j4;
            if (x0 > best) { best = x0; bidx = col; }
            if (x1 > best) { best = x1; bidx = col + 1; }
            if (x2 > best) { best = x2; bidx = col + 2; }
            if (x3 > best) { best = x3; bidx = col + 3; }
          }
        }
        const float m_new = fmaxf(mx, cm);
        const float sc = ex2(mx - m_new);
        const uint64_t sc2 = pk2(sc, sc), nm2 = pk2(-m_new, -m_new);
        s2a = mul2(s2a, sc2); s2b = mul2(s2b, sc2);
        a2a = mul2(a2a, sc2); a2b = mul2(a2b, sc2);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 c4 = __ldg(ce + c * 8 + j4);
          float t0, t1, t2, t3;
          upk2(add2(v2[2 * j4], nm2), t0, t1);
          upk2(add2(v2[2 * j4 + 1], nm2), t2, t3);
          const uint64_t e01 = pk2(ex2(t0), ex2(t1)), e23 = pk2(ex2(t2), ex2(t3));
          s2a = add2(s2a, e01);
          s2b = add2(s2b, e23);
          a2a = fma2(e01, pk2(c4.x, c4.y), a2a);
          a2b = fma2(e23, pk2(c4.z, c4.w), a2b);
        }
        mx = m_new;
      }
      float s, a;
      {
        float p0, p1, q0, q1;
        upk2(add2(s2a, s2b), p0, p1);
        upk2(add2(a2a, a2b), q0, q1);
        s = p0 + p1;
        a = q0 + q1;
      }
      tc_fence_before();
      mbar_arrive(log_empty);
      // merge the two bin halves of this pixel (upper half -> shared memory -> lower half)
      float* mg = merge + (job & 1) * 5 * 128;
      if (half == 1) {
        mg[m] = mx;
        mg[128 + m] = s;
        mg[256 + m] = a;
        if (ARGMAX) {
          mg[384 + m] = best;
          mg[512 + m] = __int_as_float(bidx);
        }
      }
      epi_bar();
      if (half == 0) {
        const float mx1 = mg[m], s1 = mg[128 + m], a1 = mg[256 + m];
        const float mm = fmaxf(mx, mx1);
        const float f0 = ex2(mx - mm), f1 = ex2(mx1 - mm);
        const float st = s * f0 + s1 * f1, at = a * f0 + a1 * f1;
        const int y = tc_.y0 + ty, x = tc_.x0 + tile * TW + tx;
        if (y < p.h && x < p.w) {
          p.depth[((long long)tc_.b * p.h + y) * p.w + x] = at / st;
          if (ARGMAX) {  // first maximum, like torch.argmax: the upper half wins only when strictly larger
            const int win = mg[384 + m] > best ? __float_as_int(mg[512 + m]) : bidx;
            p.argmax[((long long)tc_.b * p.h + y) * p.w + x] = (uint8_t)win;
          }
        }
      }
      ++job;
    };
    for (int it = 0; it < n_my; ++it) {
      mbar_wait(conv_full, it & 1);
      tc_fence_after();
      if (CONV_ONLY) {
        convert_store(it, 0);
        convert_store(it, 1);
        tc_fence_before();
        mbar_arrive(conv_empty);
        continue;
      }
      convert(it, 0);
      convert(it, 1);
      tc_fence_before();
      mbar_arrive(conv_empty);
      softmax_store(it, 0);   // logits A(it) are issued right after conv(it)
      softmax_store(it, 1);   // logits B(it) are issued in the middle of conv(it+1) (or last): before conv_full(it+1)
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, TM_COLS);
}

// =====================================================================================================================
// Transposed variant of the fused head tail (MODE 0 / 1), the one ocv_pixel_fused_bf16 launches.
//
// probes/probe_cta2.cu (B200, 148 CTAs, resident operands, tight issue loop): one tcgen05.mma M128 N128 K16 takes 100 clk, N256
// takes 164 clk — about 36 clk per instruction + 0.5 clk per column — whatever the accumulator pattern, cta_group or the TMA fill
// traffic next to it.  The conv3x3 has only 128 output channels, so with pixels on M every MMA is an N128 one (176 per 256
// pixels = 17.6k clk).  Here the conv is computed TRANSPOSED, conv^T[co, px] = W[co, ci] . X[px, ci]^T: the weight slab is the
// A operand (M = 128 output channels) and the halo plane is the B operand with N = 8 x th <= 256 PIXELS per instruction (a block
// of 8 x th pixels: its 8-pixel rows are the operand's 8-row groups, one constant stride of 10 halo pixels apart, so all nine
// taps are still shifted descriptors into one plane).  The block height th is chosen per launch so that the image height is
// covered without waste (240 rows: th = 30, N = 240).  The accumulator comes out channel-major (TMEM lane = channel), which
// is exactly the MN-major A operand of the logits GEMM: the epilogue thread of channel c writes its pixels as one 128-byte row
// per 64-pixel block, and logits[px, bin] = sum_c conv^T[c, px] Mb[bin, c] runs as M128 (pixels) x N256 (all bins) x K16 —
// 8 instructions per 128-pixel tile instead of 16.  Per 240 pixels: 72 + 16 MMAs of N <= 256 (13.9k clk) instead of 176 of N128.
// Everything else (TMA producer / MMA issuer / 8 epilogue warps, plane double buffering by channel half, slab ring, softmax)
// is the pipeline of pixel_fused_kernel above.
constexpr int TWT = 8;                       // block width (pixels) = rows of one 8-row operand group
constexpr int THT_MAX = 32;                  // block height <= 32: N = 8 th <= 256
constexpr int HWT = TWT + 2;                 // halo width
constexpr int PLANE_T = ((HWT * (THT_MAX + 2) * 128 + 1023) / 1024) * 1024;   // 43520 -> 44032
constexpr int STG_T = 2 * SLAB;              // one 128-pixel tile as MN-major A operand: 2 blocks of [128 channels x 64 px]
constexpr int SMEM_T_BYTES = 2 * PLANE_T + NS * SLAB + 2 * STG_T + MERGE_BYTES + 1024 + 256;
constexpr uint32_t TM_CONVT = 0;             // conv^T accumulator: 128 channel lanes x (8 th) pixel columns
static_assert(SMEM_T_BYTES <= 232448, "shared memory budget");
static_assert(NS == 4, "the logits operand (256 bins x 64 channels = two adjacent ring slots) relies on an even slot count");

__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t saddr, uint32_t lbo_bytes) {
  return make_smem_desc(saddr, lbo_bytes, 1024, LAYOUT_SW128);
}

__device__ __forceinline__ TileCoord block_coord_t(const PixParams& p, int t) {
  const int per = p.tiles_x * p.tiles_y;
  const int b = t / per, r = t - b * per;
  const int ty = r / p.tiles_x;
  return {b, ty * p.th, (r - ty * p.tiles_x) * TWT};
}

template <int MODE>
__global__ void __launch_bounds__(THREADS, 1)
pixel_fused_t_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                     const __grid_constant__ CUtensorMap tmMb, PixParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* plane = smem;                               // [2] 64-channel halo planes of the 10 x (th+2) pixel halo
  uint8_t* ring = plane + 2 * PLANE_T;                 // [NS] slabs
  uint8_t* stg = ring + NS * SLAB;                     // [2 tiles] bf16 conv^T tiles (MN-major A operands of the logits GEMM)
  float* merge = reinterpret_cast<float*>(stg + 2 * STG_T);
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg + 2 * STG_T + MERGE_BYTES);
  uint64_t* w_full = bars;                  // [NS]
  uint64_t* w_empty = w_full + NS;          // [NS]
  uint64_t* pl_full = w_empty + NS;         // [2]
  uint64_t* pl_empty = pl_full + 2;         // [2]
  uint64_t* stg_full = pl_empty + 2;        // [2]
  uint64_t* stg_empty = stg_full + 2;       // [2]
  uint64_t* conv_full = stg_empty + 2;
  uint64_t* conv_empty = conv_full + 1;
  uint64_t* log_full = conv_empty + 1;
  uint64_t* log_empty = log_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(log_empty + 1);

  constexpr bool ARGMAX = MODE == 1;
  pdl_launch();
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int per_cta = (p.n_tiles + gridDim.x - 1) / gridDim.x;
  const int t_begin = min(p.n_tiles, (int)blockIdx.x * per_cta);
  const int t_end = min(p.n_tiles, t_begin + per_cta);
  const int n_my = t_end - t_begin;
  const int nt = 4 * p.th;                  // pixels per logits tile (half a block), <= 128
  const uint32_t plane_tx = (uint32_t)(HWT * (p.th + 2) * 128);

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
      mbar_init(&pl_full[i], 1);
      mbar_init(&pl_empty[i], 1);
      mbar_init(&stg_full[i], N_EPI / 2);   // one warpgroup converts one tile
      mbar_init(&stg_empty[i], 1);
    }
    mbar_init(conv_full, 1);
    mbar_init(conv_empty, N_EPI);
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
    // ================= TMA producer: unit order == the MMA issuer's consumption order =================
    if (elect_one() && n_my > 0) {
      uint32_t ring_n = 0;
      bool waited = false;
      auto load_plane = [&](int it, int kh) {
        mbar_wait(&pl_empty[kh], (it & 1) ^ 1);
        const TileCoord tc_ = block_coord_t(p, t_begin + it);
        if ((p.debug & 4) && it > 0) {
          mbar_arrive(&pl_full[kh]);
          return;
        }
        mbar_arrive_expect_tx(&pl_full[kh], plane_tx);
        tma_load_4d(plane + kh * PLANE_T, &tmX, &pl_full[kh], kh * 64, tc_.x0 - 1, tc_.y0 - 1, tc_.b);
      };
      auto load_slab = [&](const CUtensorMap* m, int c0, int c1) {
        const int s = ring_n % NS;
        mbar_wait(&w_empty[s], ((ring_n / NS) & 1) ^ 1);
        if ((p.debug & 1) && ring_n >= 26) {
          mbar_arrive(&w_full[s]);
        } else {
          mbar_arrive_expect_tx(&w_full[s], SLAB);
          tma_load_2d(ring + s * SLAB, m, &w_full[s], c0, c1);
        }
        ++ring_n;
      };
      // Mb[b] as two [256 bins x 64 channels] operands: each is two adjacent ring slots (the unit count before it is even)
      auto load_mb = [&](int it) {
        if (!waited) {
          pdl_wait();
          waited = true;
        }
        const int b = block_coord_t(p, t_begin + it).b;
        for (int kh = 0; kh < 2; ++kh)
          for (int nh = 0; nh < 2; ++nh) load_slab(&tmMb, kh * 64, b * NB + nh * 128);
      };
      load_plane(0, 0);
      for (int it = 0; it < n_my; ++it) {
        for (int tap = 0; tap < 5; ++tap) load_slab(&tmW, 0, tap * C);
        load_plane(it, 1);
        for (int tap = 5; tap < 9; ++tap) load_slab(&tmW, 0, tap * C);
        load_slab(&tmW, 64, 0);
        if (it > 0) load_mb(it - 1);                                   // logits of tile 1 of the previous block
        for (int tap = 1; tap < 5; ++tap) load_slab(&tmW, 64, tap * C);
        if (it + 1 < n_my) load_plane(it + 1, 0);
        for (int tap = 5; tap < 9; ++tap) load_slab(&tmW, 64, tap * C);
        load_mb(it);                                                   // logits of tile 0
      }
      load_mb(n_my - 1);
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one() && n_my > 0) {
      const uint32_t idesc_conv = make_idesc_bf16(128, 8 * p.th);
      constexpr uint32_t idesc_log = make_idesc_bf16(128, NB, 1, 0);   // A = conv^T tile, MN-major; B = Mb, K-major
      uint32_t ring_n = 0, log_n = 0;
      long long wt[6] = {0, 0, 0, 0, 0, 0};   // waited on: 0 w_full (conv) 1 w_full (Mb) 2 pl_full 3 conv_empty 4 stg_full 5 log_empty
      const bool timing = p.dbg_out != nullptr;
      const long long t_start = clock64();
      unsigned long long g_start;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_start));
      auto twait = [&](uint64_t* bar, uint32_t parity, int slot) {
        if (!timing) {
          mbar_wait(bar, parity);
          return;
        }
        const long long t = clock64();
        mbar_wait(bar, parity);
        wt[slot] += clock64() - t;
      };
      auto conv_tap = [&](uint32_t h_base, int tap, bool first) {
        const int dy = tap / 3, dx = tap - dy * 3;
        const int s = ring_n % NS;
        twait(&w_full[s], (ring_n / NS) & 1, 0);
        tc_fence_after();
        const uint32_t a_base = smem_u32(ring + s * SLAB);             // weights [128 co x 64 ci]
        const uint32_t b_base = h_base + (dy * HWT + dx) * 128;        // pixels of the shifted block, 8-row groups HWT*128 B apart
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16(tmem + TM_CONVT, desc_sw128(a_base + kk * 32, 1024), desc_sw128(b_base + kk * 32, HWT * 128), idesc_conv,
                    (!first || kk != 0) ? 1u : 0u);
        umma_commit(&w_empty[s]);
        ++ring_n;
      };
      auto logits_mma = [&](int it, int tile) {
        twait(&stg_full[tile], it & 1, 4);
        twait(log_empty, (log_n & 1) ^ 1, 5);
        tc_fence_after();
        const uint32_t a_base = smem_u32(stg + tile * STG_T);
        for (int kh = 0; kh < 2; ++kh) {
          const int s = ring_n % NS;   // even
          twait(&w_full[s], (ring_n / NS) & 1, 1);
          twait(&w_full[s + 1], (ring_n / NS) & 1, 1);
          tc_fence_after();
          const uint32_t b_base = smem_u32(ring + s * SLAB);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem + TM_LOGITS, desc_mn_sw128(a_base + (kh * 64 + kk * 16) * 128, SLAB), desc_sw128(b_base + kk * 32, 1024),
                      idesc_log, (kh | kk) != 0);
          umma_commit(&w_empty[s]);
          umma_commit(&w_empty[s + 1]);
          ring_n += 2;
        }
        umma_commit(&stg_empty[tile]);
        umma_commit(log_full);
        ++log_n;
      };
      for (int it = 0; it < n_my; ++it) {
        twait(conv_empty, (it & 1) ^ 1, 3);
        tc_fence_after();
        twait(&pl_full[0], it & 1, 2);
        tc_fence_after();
        const uint32_t h0 = smem_u32(plane);
        for (int tap = 0; tap < 9; ++tap) conv_tap(h0, tap, tap == 0);
        umma_commit(&pl_empty[0]);
        twait(&pl_full[1], it & 1, 2);
        tc_fence_after();
        const uint32_t h1 = smem_u32(plane + PLANE_T);
        conv_tap(h1, 0, false);
        if (it > 0) logits_mma(it - 1, 1);
        for (int tap = 1; tap < 9; ++tap) conv_tap(h1, tap, false);
        umma_commit(&pl_empty[1]);
        umma_commit(conv_full);
        logits_mma(it, 0);
      }
      logits_mma(n_my - 1, 1);
      if (timing) {
        long long* o = p.dbg_out + (long long)blockIdx.x * 16;
        for (int i = 0; i < 6; ++i) o[i] = wt[i];
        o[6] = clock64() - t_start;
        o[7] = n_my;
        unsigned long long g_end;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_end));
        o[8] = (long long)(g_end - g_start);
      }
    }
  } else if (warp >= 4) {
    // ================= convert + softmax epilogue =================
    const int e = warp - 4;
    const int q = e & 3;            // == warp % 4: the TMEM lane quarter this warp may read
    const int half = e >> 2;        // convert: the tile this warpgroup stages; softmax: the half of the bins
    const int m = q * 32 + lane;    // convert: channel; softmax: pixel of the tile
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    uint32_t job = 0;
    if (n_my > 0) pdl_wait();
    // conv^T accumulator, channel m, the nt pixels of tile `half` (128 columns are read: the surplus is never used) -> bf16 ->
    // row m of the two 64-pixel blocks of the staging tile
    auto convert = [&](int it) {
      const uint32_t src = tmem + lane_base + TM_CONVT + (uint32_t)(half * nt);
      mbar_wait(&stg_empty[half], (it & 1) ^ 1);
#pragma unroll
      for (int blk = 0; blk < 2; ++blk) {
        uint32_t r0[32], r1[32];
        tmem_ld32(src + blk * 64, r0);
        tmem_ld32(src + blk * 64 + 32, r1);
        tmem_ld_wait();
        uint8_t* row = stg + half * STG_T + blk * SLAB + m * 128;
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          const uint32_t* r = ch < 4 ? r0 + ch * 8 : r1 + (ch - 4) * 8;
          const uint4 v = make_uint4(pack_bf16(__uint_as_float(r[0]), __uint_as_float(r[1])),
                                     pack_bf16(__uint_as_float(r[2]), __uint_as_float(r[3])),
                                     pack_bf16(__uint_as_float(r[4]), __uint_as_float(r[5])),
                                     pack_bf16(__uint_as_float(r[6]), __uint_as_float(r[7])));
          *reinterpret_cast<uint4*>(row + ((ch ^ (m & 7)) << 4)) = v;
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(&stg_full[half]);
    };
    auto softmax_store = [&](int it, int tile) {
      const TileCoord tc_ = block_coord_t(p, t_begin + it);
      const float4* bl = reinterpret_cast<const float4*>(p.blog + (long long)tc_.b * NB + half * 128);
      const float4* ce = reinterpret_cast<const float4*>(p.centres + (long long)tc_.b * NB + half * 128);
      mbar_wait(log_full, job & 1);
      tc_fence_after();
      float mx = -INFINITY, best = -INFINITY;
      int bidx = 0;
      uint64_t s2a = 0, s2b = 0, a2a = 0, a2b = 0;
      const uint32_t src = tmem + lane_base + TM_LOGITS + half * 128;
      uint32_t rb[2][32];
      tmem_ld32(src, rb[0]);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_ld_wait();
        if (c < 3) tmem_ld32(src + (c + 1) * 32, rb[(c + 1) & 1]);
        if (p.debug & 2) continue;
        const uint32_t* r = rb[c & 1];
        uint64_t v2[16];
        float cm = -INFINITY;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 b4 = __ldg(bl + c * 8 + j4);
          v2[2 * j4] = add2(pk2u(r[4 * j4], r[4 * j4 + 1]), pk2(b4.x, b4.y));
          v2[2 * j4 + 1] = add2(pk2u(r[4 * j4 + 2], r[4 * j4 + 3]), pk2(b4.z, b4.w));
          float x0, x1, x2, x3;
          upk2(v2[2 * j4], x0, x1);
          upk2(v2[2 * j4 + 1], x2, x3);
          cm = max3(cm, x0, x1);
          cm = max3(cm, x2, x3);
          if (ARGMAX) {
            const int col = half * 128 + c * 32 + 4 * j4;
            if (x0 > best) { best = x0; bidx = col; }
            if (x1 > best) { best = x1; bidx = col + 1; }
            if (x2 > best) { best = x2; bidx = col + 2; }
            if (x3 > best) { best = x3; bidx = col + 3; }
          }
        }
        const float m_new = fmaxf(mx, cm);
        const float sc = ex2(mx - m_new);
        const uint64_t sc2 = pk2(sc, sc), nm2 = pk2(-m_new, -m_new);
        s2a = mul2(s2a, sc2); s2b = mul2(s2b, sc2);
        a2a = mul2(a2a, sc2); a2b = mul2(a2b, sc2);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 c4 = __ldg(ce + c * 8 + j4);
          float t0, t1, t2, t3;
          upk2(add2(v2[2 * j4], nm2), t0, t1);
          upk2(add2(v2[2 * j4 + 1], nm2), t2, t3);
          const uint64_t e01 = pk2(ex2(t0), ex2(t1)), e23 = pk2(ex2(t2), ex2(t3));
          s2a = add2(s2a, e01);
          s2b = add2(s2b, e23);
          a2a = fma2(e01, pk2(c4.x, c4.y), a2a);
          a2b = fma2(e23, pk2(c4.z, c4.w), a2b);
        }
        mx = m_new;
      }
      float s, a;
      {
        float p0, p1, q0, q1;
        upk2(add2(s2a, s2b), p0, p1);
        upk2(add2(a2a, a2b), q0, q1);
        s = p0 + p1;
        a = q0 + q1;
      }
      tc_fence_before();
      mbar_arrive(log_empty);
      float* mg = merge + (job & 1) * 5 * 128;
      if (half == 1) {
        mg[m] = mx;
        mg[128 + m] = s;
        mg[256 + m] = a;
        if (ARGMAX) {
          mg[384 + m] = best;
          mg[512 + m] = __int_as_float(bidx);
        }
      }
      epi_bar();
      if (half == 0) {
        const float mx1 = mg[m], s1 = mg[128 + m], a1 = mg[256 + m];
        const float mm = fmaxf(mx, mx1);
        const float f0 = ex2(mx - mm), f1 = ex2(mx1 - mm);
        const float st = s * f0 + s1 * f1, at = a * f0 + a1 * f1;
        const int n = tile * nt + m;                       // pixel of the block: n = y * 8 + x
        const int y = tc_.y0 + (n >> 3), x = tc_.x0 + (n & 7);
        if (m < nt && y < p.h && x < p.w) {
          p.depth[((long long)tc_.b * p.h + y) * p.w + x] = at / st;
          if (ARGMAX) {
            const int win = mg[384 + m] > best ? __float_as_int(mg[512 + m]) : bidx;
            p.argmax[((long long)tc_.b * p.h + y) * p.w + x] = (uint8_t)win;
          }
        }
      }
      ++job;
    };
    for (int it = 0; it < n_my; ++it) {
      mbar_wait(conv_full, it & 1);
      tc_fence_after();
      convert(it);
      tc_fence_before();
      mbar_arrive(conv_empty);
      softmax_store(it, 0);
      softmax_store(it, 1);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, TM_COLS);
}

// block height of the transposed kernel for an image of h rows: even, <= 32, least (blocks x MMA time per block)
int pick_block_height(int h) {
  int best = 32;
  long long best_cost = -1;
  for (int th = 32; th >= 16; th -= 2) {
    const long long cost = (long long)ceil_div(h, th) * (72 * (36 + 4 * th) + 16 * 164);
    if (best_cost < 0 || cost < best_cost) { best = th; best_cost = cost; }
  }
  return best;
}

int make_tmap_features_t(CUtensorMap* out, const void* X, int B, int h, int w, int th) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return OCV_E_DRIVER;
  if (reinterpret_cast<uintptr_t>(X) & 15) return OCV_E_ALIGN;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)w * C * 2, (cuuint64_t)h * w * C * 2};
  cuuint32_t box[4] = {64, HWT, (cuuint32_t)(th + 2), 1};  // one 64-channel plane of the 10 x (th+2) pixel halo
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(X), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OCV_OK : OCV_E_DRIVER;
}

int make_tmap_features(CUtensorMap* out, const void* X, int B, int h, int w) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return OCV_E_DRIVER;
  if (reinterpret_cast<uintptr_t>(X) & 15) return OCV_E_ALIGN;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)w * C * 2, (cuuint64_t)h * w * C * 2};
  cuuint32_t box[4] = {64, HW_, HH, 1};  // one 64-channel plane of the 18 x 18 pixel halo
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(X), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OCV_OK : OCV_E_DRIVER;
}

}  // namespace

static int pixel_launch(const void* X, const void* W3p, const void* Mb, const float* blog, const float* centres, float* depth,
                        uint8_t* argmax, int B, int h, int w, int max_ctas, void* stream, void* conv_out = nullptr,
                        const float* conv_bias = nullptr) {
  const bool conv_only = conv_out != nullptr;
  OCV_CHECK_ARG(X && W3p && B > 0 && h > 0 && w > 0);
  if (conv_only) {
    OCV_CHECK_ARG(conv_bias);
    OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(conv_out) | reinterpret_cast<uintptr_t>(conv_bias)) & 15) == 0);
  } else {
    OCV_CHECK_ARG(Mb && blog && centres && depth);
    OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(blog) | reinterpret_cast<uintptr_t>(centres)) & 15) == 0);
  }
  if (int s = ocv_require_sm100()) return s;
  // the fused tail runs on the transposed kernel (N <= 256 pixels per MMA); OCV_PIXEL_T=0 selects the pixels-on-M kernel
  // (read per call: the parity tests run both kernels in one process)
  const char* env_t = getenv("OCV_PIXEL_T");
  const bool transposed = !conv_only && !(env_t && env_t[0] == '0');
  const int th = transposed ? pick_block_height(h) : 0;
  CUtensorMap tmX, tmW, tmMb;
  if (int s = transposed ? make_tmap_features_t(&tmX, X, B, h, w, th) : make_tmap_features(&tmX, X, B, h, w)) return s;
  if (int s = tc::make_tmap_2d_bf16(&tmW, W3p, 9 * C, C, C, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  if (conv_only) tmMb = tmW;  // unused by that instantiation
  else if (int s = tc::make_tmap_2d_bf16(&tmMb, Mb, (uint64_t)B * NB, C, C, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaError_t e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(pixel_fused_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(pixel_fused_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(pixel_fused_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(pixel_fused_t_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_T_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(pixel_fused_t_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_T_BYTES);
    if (e != cudaSuccess) { n_sm = 0; return (int)e; }
  }
  PixParams p{};
  p.blog = blog; p.centres = centres; p.depth = depth; p.argmax = argmax;
  p.conv_out = static_cast<__nv_bfloat16*>(conv_out); p.conv_bias = conv_bias;
  p.B = B; p.h = h; p.w = w;
  p.th = th;
#ifdef OCV_PROBES   // measurement build only (probes/time_pixel_dbg.py): with a knob set the results are wrong on purpose
  if (const char* e = getenv("OCV_PIXEL_DEBUG")) p.debug = atoi(e);
#endif
  p.tiles_x = transposed ? ceil_div(w, TWT) : ceil_div(w, PW);
  p.tiles_y = transposed ? ceil_div(h, th) : ceil_div(h, TH);
  const long long nt = (long long)p.tiles_x * p.tiles_y * B;
  OCV_CHECK_SHAPE(nt < (1ll << 30));
  p.n_tiles = (int)nt;
  int grid = n_sm;
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (grid > p.n_tiles) grid = p.n_tiles;
#ifdef OCV_PROBES
  if (transposed && (p.debug & 8)) {   // measurement aid: synchronous launch + wait breakdown of the MMA issuer on stderr
    if (argmax) return OCV_E_ARG;      // the timing instantiation has no argmax output
    if (grid > 148) grid = 148;        // one 16-counter row per CTA in the fixed buffer below
    static long long* dbg = nullptr;
    if (!dbg) cudaMalloc(&dbg, 148 * 16 * sizeof(long long));
    cudaMemset(dbg, 0, 148 * 16 * sizeof(long long));
    p.dbg_out = dbg;
    int st = ocv_launch_pdl(pixel_fused_t_kernel<0>, dim3(grid), dim3(THREADS), SMEM_T_BYTES, stream, tmX, tmW, tmMb, p);
    cudaStreamSynchronize(static_cast<cudaStream_t>(stream));
    long long hbuf[148 * 16];
    cudaMemcpy(hbuf, dbg, sizeof(hbuf), cudaMemcpyDeviceToHost);
    double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int c = 0; c < grid && c < 148; ++c)
      for (int i = 0; i < 9; ++i) acc[i] += (double)hbuf[c * 16 + i] / grid;
    fprintf(stderr, "[ocv pixel_t] MMA issuer loop %.1f us at %.0f MHz; ", acc[8] * 1e-3, acc[6] / (acc[8] * 1e-3));
    fprintf(stderr, "[ocv pixel_t] per CTA: %.0f blocks, %.0f clk total = %.0f clk/block; waits/block: w_full(conv) %.0f  w_full(Mb) %.0f  "
            "pl_full %.0f  conv_empty %.0f  stg_full %.0f  log_empty %.0f\n", acc[7], acc[6], acc[6] / acc[7], acc[0] / acc[7], acc[1] / acc[7],
            acc[2] / acc[7], acc[3] / acc[7], acc[4] / acc[7], acc[5] / acc[7]);
    return st;
  }
#endif
  if (transposed) {
    if (argmax) return ocv_launch_pdl(pixel_fused_t_kernel<1>, dim3(grid), dim3(THREADS), SMEM_T_BYTES, stream, tmX, tmW, tmMb, p);
    return ocv_launch_pdl(pixel_fused_t_kernel<0>, dim3(grid), dim3(THREADS), SMEM_T_BYTES, stream, tmX, tmW, tmMb, p);
  }
  if (conv_only) return ocv_launch_pdl(pixel_fused_kernel<2>, dim3(grid), dim3(THREADS), SMEM_BYTES, stream, tmX, tmW, tmMb, p);
  if (argmax) return ocv_launch_pdl(pixel_fused_kernel<1>, dim3(grid), dim3(THREADS), SMEM_BYTES, stream, tmX, tmW, tmMb, p);
  return ocv_launch_pdl(pixel_fused_kernel<0>, dim3(grid), dim3(THREADS), SMEM_BYTES, stream, tmX, tmW, tmMb, p);
}

extern "C" int ocv_pixel_fused_bf16(const void* X, const void* W3p, const void* Mb, const float* blog,
                                    const float* centres, float* depth, int B, int h, int w, int max_ctas,
                                    void* stream) {
  return pixel_launch(X, W3p, Mb, blog, centres, depth, nullptr, B, h, w, max_ctas, stream);
}

extern "C" int ocv_pixel_fused_bf16_argmax(const void* X, const void* W3p, const void* Mb, const float* blog,
                                           const float* centres, float* depth, uint8_t* argmax_bin, int B, int h, int w,
                                           void* stream) {
  OCV_CHECK_ARG(argmax_bin);
  return pixel_launch(X, W3p, Mb, blog, centres, depth, argmax_bin, B, h, w, 0, stream);
}

extern "C" int ocv_conv3x3_nhwc_bf16(const void* X, const void* W3p, const float* bias, void* Y, int B, int h, int w, int max_ctas,
                                     void* stream) {
  OCV_CHECK_ARG(Y && bias);
  return pixel_launch(X, W3p, nullptr, nullptr, nullptr, nullptr, nullptr, B, h, w, max_ctas, stream, Y, bias);
}
