// Training step, pixel path WITHOUT materialised logits (bf16 tensor-core mode): two tcgen05 kernels over flat 128-pixel tiles
// of the conv3x3 output x3 [B*P, 128] (bf16 NHWC, written by ocv_conv3x3_nhwc_bf16).
//
//   forward   ocv_tail_fused_fwd_bf16:  logits[px, k] = x3[px, :] . Mb2[b, k, :]  (Mb2 = log2e * Wout @ Q[b], 256 x 128)
//             in TMEM -> online softmax over the 256 bins in registers -> depth[px] = sum_k p_k c_k, and the two softmax
//             statistics of the pixel (max and sum of 2^(l - max)); 12 bytes written per pixel.
//   backward  ocv_tail_fused_bwd_bf16:  recomputes the logits tile TRANSPOSED (logits^T[k, px] = Mb2 . x3^T: TMEM lane = bin),
//             forms dl[k, px] = g p_k (c_k - depth) from the saved statistics, and feeds the bf16 tile straight into two more
//             MMAs:  dMb[b][k, c] += dl^T-tile . x3-tile  (accumulated in TMEM over all tiles of an image) and
//             dx3^T[c, px] = ln2 * Mb2^T . dl  (stored channel-major: the A operand of the conv3x3 weight-gradient GEMM).
//             Per-bin sums over pixels (d centres = sum g p_k, d conv_out.bias = sum dl_k) are thread-local because a thread owns
//             a BIN: no cross-lane reduction.  The x3 / Mb2 / dl tiles each exist ONCE in shared memory: the MMAs that contract
//             over a tile's row index read it as an MN-major operand (probes/probe_mn_major.cu), no transposed copies.
//
// replaces (training, bf16 mode): 8 logits GEMMs writing [B,256,P] fp32 (463 MB at B=8 208x272), ocv_tail_logits_stats_f32,
// ocv_tail_logits_bwd_stats_bf16 (+ its dlT [B,256,P] bf16), two batched transposes and 16 per-image GEMMs; HBM traffic of the
// pair: 2 x 116 MB of x3 + 116 MB of dx3^T + 5 planes of 1.8 MB against ~2.9 GB.
// reference sites: modules/GraphBins.py:109-119 (conv_out, softmax, sum out*centers), modules/layers.py:31-36 and their autograd.
// Contract: include/objcavit_b200.h.
#include "tc_common.cuh"

namespace {

using namespace tc;

constexpr int C = 128;              // channels of x3 = E
constexpr int NB = 256;             // bins
constexpr int TP = 128;             // pixels per tile
constexpr int SLAB = 128 * 128;     // [128 rows x 64 bf16] under the 128B swizzle = 16 KB
constexpr int DLSLAB = 256 * 128;   // [256 bin rows x 64 px]
constexpr int N_EPI = 256;
constexpr int THREADS = 128 + N_EPI;
constexpr float LN2 = 0.6931471805599453f;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr, uint32_t lbo_bytes) { return make_smem_desc(saddr, lbo_bytes, 1024, LAYOUT_SW128); }
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

struct TailParams {
  const float* blog2;      // [B,256]  log2e * conv_out bias
  const float* centres;    // [B,256]
  float* depth;            // [B*P]   (fwd: out; bwd: in)
  float* smax;             // [B*P]   max_k of the log2-domain logits
  float* ssum;             // [B*P]   sum_k 2^(l_k - max)
  const float* gdepth;     // bwd: d loss / d depth [B*P]
  __nv_bfloat16* dx3T;     // bwd: [128][ld_dx] bf16
  long long ld_dx;
  float* part_dmb;         // bwd: [n_cta][runs][256][128] fp32
  float* part_sums;        // bwd: [n_cta][runs][2 (px half)][2 (sum g p, sum dl)][256]
  int* part_img;           // bwd: [n_cta][runs] image of the run, -1 = unused
  int B, P, tiles_per_img, n_tiles, runs;
};

// ======================================================= forward ========================================================
constexpr int F_SMEM = 2 * 2 * SLAB /*x3 x2*/ + 4 * SLAB /*Mb*/ + 2 * 3 * 128 * 4 /*merge*/ + 1024 + 256;
constexpr uint32_t F_TM_COLS = 512;   // two logits buffers of 256 columns

__global__ void __launch_bounds__(THREADS, 1)
tail_fwd_kernel(const __grid_constant__ CUtensorMap tmX3, const __grid_constant__ CUtensorMap tmMb, TailParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* x3s = smem;                        // [2 bufs][2 k-halves]
  uint8_t* mbs = x3s + 4 * SLAB;              // [2 bin halves][2 k-halves]
  float* merge = reinterpret_cast<float*>(mbs + 4 * SLAB);   // [2][3][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(merge) + 2 * 3 * 128 * 4);
  uint64_t* x_full = bars;            // [2]
  uint64_t* x_empty = x_full + 2;     // [2]
  uint64_t* mb_full = x_empty + 2;
  uint64_t* mb_empty = mb_full + 1;
  uint64_t* log_full = mb_empty + 1;  // [2]
  uint64_t* log_empty = log_full + 2; // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(log_empty + 2);

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int per_cta = (p.n_tiles + gridDim.x - 1) / gridDim.x;
  const int t_begin = min(p.n_tiles, (int)blockIdx.x * per_cta);
  const int t_end = min(p.n_tiles, t_begin + per_cta);
  const int n_my = t_end - t_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX3);
    tma_prefetch_desc(&tmMb);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&x_full[i], 1);
      mbar_init(&x_empty[i], 1);
      mbar_init(&log_full[i], 1);
      mbar_init(&log_empty[i], N_EPI);
    }
    mbar_init(mb_full, 1);
    mbar_init(mb_empty, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, F_TM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (elect_one() && n_my > 0) {
      int cur_b = -1, img_n = 0;
      for (int it = 0; it < n_my; ++it) {
        const int t = t_begin + it, b = t / p.tiles_per_img, px0 = (t - b * p.tiles_per_img) * TP;
        if (b != cur_b) {
          mbar_wait(mb_empty, (img_n & 1) ^ 1);
          mbar_arrive_expect_tx(mb_full, 4 * SLAB);
          for (int nh = 0; nh < 2; ++nh)
            for (int kh = 0; kh < 2; ++kh) tma_load_2d(mbs + (nh * 2 + kh) * SLAB, &tmMb, mb_full, kh * 64, b * NB + nh * 128);
          cur_b = b;
          ++img_n;
        }
        const int buf = it & 1;
        mbar_wait(&x_empty[buf], ((it >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&x_full[buf], 2 * SLAB);
        for (int kh = 0; kh < 2; ++kh)
          tma_load_2d(x3s + (buf * 2 + kh) * SLAB, &tmX3, &x_full[buf], kh * 64, (int)((long long)b * p.P + px0));
      }
    }
  } else if (warp == 1) {
    if (elect_one() && n_my > 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 128);
      int cur_b = -1, img_n = 0;
      for (int it = 0; it < n_my; ++it) {
        const int t = t_begin + it, b = t / p.tiles_per_img;
        if (b != cur_b) {
          if (cur_b >= 0) umma_commit(mb_empty);   // arrives when every MMA issued so far (all of them read the old Mb) is done
          mbar_wait(mb_full, img_n & 1);
          cur_b = b;
          ++img_n;
        }
        const int buf = it & 1;
        mbar_wait(&x_full[buf], (it >> 1) & 1);
        mbar_wait(&log_empty[buf], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t a0 = smem_u32(x3s + buf * 2 * SLAB), b0 = smem_u32(mbs);
        for (int nh = 0; nh < 2; ++nh)
          for (int kh = 0; kh < 2; ++kh)
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_bf16(tmem + buf * 256 + nh * 128, desc_kmajor_sw128(a0 + kh * SLAB + kk * 32),
                        desc_kmajor_sw128(b0 + (nh * 2 + kh) * SLAB + kk * 32), idesc, (kh | kk) != 0);
        umma_commit(&x_empty[buf]);
        umma_commit(&log_full[buf]);
      }
    }
  } else if (warp >= 4) {
    // softmax epilogue: thread = one pixel (TMEM lane) x one half of the bins; the halves are merged through shared memory
    const int e = warp - 4, q = e & 3, half = e >> 2;
    const int m = q * 32 + lane;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    for (int it = 0; it < n_my; ++it) {
      const int t = t_begin + it, b = t / p.tiles_per_img, px0 = (t - b * p.tiles_per_img) * TP;
      const int buf = it & 1;
      const float4* bl = reinterpret_cast<const float4*>(p.blog2 + (long long)b * NB + half * 128);
      const float4* ce = reinterpret_cast<const float4*>(p.centres + (long long)b * NB + half * 128);
      mbar_wait(&log_full[buf], (it >> 1) & 1);
      tc_fence_after();
      float mx = -INFINITY, s = 0.f, a = 0.f;
      const uint32_t src = tmem + lane_base + buf * 256 + half * 128;
      uint32_t rb[2][32];
      tmem_ld32(src, rb[0]);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_ld_wait();
        if (c < 3) tmem_ld32(src + (c + 1) * 32, rb[(c + 1) & 1]);
        const uint32_t* r = rb[c & 1];
        float v[32];
        float cm = -INFINITY;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 b4 = __ldg(bl + c * 8 + j4);
          v[4 * j4] = __uint_as_float(r[4 * j4]) + b4.x;
          v[4 * j4 + 1] = __uint_as_float(r[4 * j4 + 1]) + b4.y;
          v[4 * j4 + 2] = __uint_as_float(r[4 * j4 + 2]) + b4.z;
          v[4 * j4 + 3] = __uint_as_float(r[4 * j4 + 3]) + b4.w;
          cm = fmaxf(fmaxf(cm, fmaxf(v[4 * j4], v[4 * j4 + 1])), fmaxf(v[4 * j4 + 2], v[4 * j4 + 3]));
        }
        const float m_new = fmaxf(mx, cm);
        const float sc = ex2(mx - m_new);
        float s0 = 0.f, s1 = 0.f, a0 = 0.f, a1 = 0.f;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 c4 = __ldg(ce + c * 8 + j4);
          const float e0 = ex2(v[4 * j4] - m_new), e1 = ex2(v[4 * j4 + 1] - m_new), e2 = ex2(v[4 * j4 + 2] - m_new),
                      e3 = ex2(v[4 * j4 + 3] - m_new);
          s0 += e0 + e2;
          s1 += e1 + e3;
          a0 = fmaf(e0, c4.x, fmaf(e2, c4.z, a0));
          a1 = fmaf(e1, c4.y, fmaf(e3, c4.w, a1));
        }
        s = fmaf(s, sc, s0 + s1);
        a = fmaf(a, sc, a0 + a1);
        mx = m_new;
      }
      tc_fence_before();
      mbar_arrive(&log_empty[buf]);
      float* mg = merge + buf * 3 * 128;
      if (half == 1) {
        mg[m] = mx;
        mg[128 + m] = s;
        mg[256 + m] = a;
      }
      epi_bar();
      if (half == 0) {
        const float mx1 = mg[m], s1 = mg[128 + m], a1 = mg[256 + m];
        const float mm = fmaxf(mx, mx1);
        const float f0 = ex2(mx - mm), f1 = ex2(mx1 - mm);
        const float st = s * f0 + s1 * f1, at = a * f0 + a1 * f1;
        const long long idx = (long long)b * p.P + px0 + m;
        p.depth[idx] = at / st;
        p.smax[idx] = mm;
        p.ssum[idx] = st;
      }
      epi_bar();   // the merge buffer of this parity is reused two tiles later; cheap, and keeps the two halves in step
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, F_TM_COLS);
}

// ======================================================= backward =======================================================
constexpr int B_SMEM = 2 * 2 * SLAB /*x3 x2*/ + 4 * SLAB /*Mb*/ + 2 * DLSLAB /*dl^T*/ + 2 * 128 * 16 /*per-pixel vectors*/ + 1024 + 256;
constexpr uint32_t TM_LOG = 0, TM_DX = 128, TM_DMB = 256, B_TM_COLS = 512;

__global__ void __launch_bounds__(THREADS, 1)
tail_bwd_kernel(const __grid_constant__ CUtensorMap tmX3, const __grid_constant__ CUtensorMap tmMb, TailParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* x3s = smem;                        // [2 bufs][2 channel halves]: [128 px rows x 64 ch]
  uint8_t* mbs = x3s + 4 * SLAB;              // [2 bin halves][2 channel halves]: [128 bin rows x 64 ch]
  uint8_t* dls = mbs + 4 * SLAB;              // [2 px halves]: [256 bin rows x 64 px]
  float4* vec = reinterpret_cast<float4*>(dls + 2 * DLSLAB);   // [2 bufs][128 px] {max2, g / sum, depth, -}
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(vec) + 2 * 128 * 16);
  uint64_t* x_full = bars;             // [2]
  uint64_t* x_empty = x_full + 2;      // [2]
  uint64_t* v_full = x_empty + 2;      // [2]
  uint64_t* v_empty = v_full + 2;      // [2]
  uint64_t* mb_full = v_empty + 2;
  uint64_t* mb_empty = mb_full + 1;
  uint64_t* log_full = mb_empty + 1;
  uint64_t* log_empty = log_full + 1;
  uint64_t* dl_full = log_empty + 1;   // [2 bin halves]
  uint64_t* dl_empty = dl_full + 2;    // [2]
  uint64_t* dx_full = dl_empty + 2;
  uint64_t* dx_empty = dx_full + 1;
  uint64_t* dmb_full = dx_empty + 1;
  uint64_t* dmb_empty = dmb_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dmb_empty + 1);

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int per_cta = (p.n_tiles + gridDim.x - 1) / gridDim.x;
  const int t_begin = min(p.n_tiles, (int)blockIdx.x * per_cta);
  const int t_end = min(p.n_tiles, t_begin + per_cta);
  const int n_my = t_end - t_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX3);
    tma_prefetch_desc(&tmMb);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&x_full[i], 1);
      mbar_init(&x_empty[i], 1);
      mbar_init(&v_full[i], 32);
      mbar_init(&v_empty[i], N_EPI);
      mbar_init(&dl_full[i], N_EPI);
      mbar_init(&dl_empty[i], 1);
    }
    mbar_init(mb_full, 1);
    mbar_init(mb_empty, 1);
    mbar_init(log_full, 1);
    mbar_init(log_empty, N_EPI);
    mbar_init(dx_full, 1);
    mbar_init(dx_empty, N_EPI);
    mbar_init(dmb_full, 1);
    mbar_init(dmb_empty, N_EPI);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, B_TM_COLS);
  // unused run slots of this CTA are marked so the reduction skips them
  if (warp == 3) for (int r = lane; r < p.runs; r += 32) p.part_img[(long long)blockIdx.x * p.runs + r] = -1;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer: x3 tiles (double buffered), Mb2 of the current image (resident) =================
    if (elect_one() && n_my > 0) {
      int cur_b = -1, img_n = 0;
      for (int it = 0; it < n_my; ++it) {
        const int t = t_begin + it, b = t / p.tiles_per_img, px0 = (t - b * p.tiles_per_img) * TP;
        if (b != cur_b) {
          mbar_wait(mb_empty, (img_n & 1) ^ 1);
          mbar_arrive_expect_tx(mb_full, 4 * SLAB);
          for (int nh = 0; nh < 2; ++nh)
            for (int kh = 0; kh < 2; ++kh) tma_load_2d(mbs + (nh * 2 + kh) * SLAB, &tmMb, mb_full, kh * 64, b * NB + nh * 128);
          cur_b = b;
          ++img_n;
        }
        const int buf = it & 1;
        mbar_wait(&x_empty[buf], ((it >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&x_full[buf], 2 * SLAB);
        for (int kh = 0; kh < 2; ++kh)
          tma_load_2d(x3s + (buf * 2 + kh) * SLAB, &tmX3, &x_full[buf], kh * 64, (int)((long long)b * p.P + px0));
      }
    }
  } else if (warp == 3) {
    // ================= per-pixel vectors of the tile: {max2, g / sum, depth} =================
    for (int it = 0; it < n_my; ++it) {
      const int t = t_begin + it, b = t / p.tiles_per_img, px0 = (t - b * p.tiles_per_img) * TP;
      const int buf = it & 1;
      const long long idx = (long long)b * p.P + px0 + lane * 4;
      const float4 mx = __ldg(reinterpret_cast<const float4*>(p.smax + idx));
      const float4 sm = __ldg(reinterpret_cast<const float4*>(p.ssum + idx));
      const float4 dp = __ldg(reinterpret_cast<const float4*>(p.depth + idx));
      const float4 g = __ldg(reinterpret_cast<const float4*>(p.gdepth + idx));
      mbar_wait(&v_empty[buf], ((it >> 1) & 1) ^ 1);
      float4* dst = vec + buf * 128 + lane * 4;
      dst[0] = make_float4(mx.x, g.x / sm.x, dp.x, 0.f);
      dst[1] = make_float4(mx.y, g.y / sm.y, dp.y, 0.f);
      dst[2] = make_float4(mx.z, g.z / sm.z, dp.z, 0.f);
      dst[3] = make_float4(mx.w, g.w / sm.w, dp.w, 0.f);
      mbar_arrive(&v_full[buf]);
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (elect_one() && n_my > 0) {
      constexpr uint32_t id_kk = make_idesc_bf16(128, 128, 0, 0);   // logits^T: both K-major
      constexpr uint32_t id_kn = make_idesc_bf16(128, 128, 0, 1);   // dMb: A = dl^T K-major, B = x3 MN-major
      constexpr uint32_t id_nn = make_idesc_bf16(128, 128, 1, 1);   // dx3^T: A = Mb2 MN-major, B = dl^T MN-major
      int cur_b = -1, img_n = 0, flush_n = 0;
      uint32_t log_n = 0;
      bool dmb_acc = false, ahead = false;
      const uint32_t mb0 = smem_u32(mbs), dl0 = smem_u32(dls);
      // logits^T half [128 bins x 128 px] = Mb2[nh] (A, K = channels) . x3^T (B, K = channels); the single TMEM buffer is free
      // again once the epilogue has READ the previous half (it arrives on log_empty before it starts computing)
      auto logits_mma = [&](uint32_t x0, int nh) {
        mbar_wait(log_empty, (log_n & 1) ^ 1);
        tc_fence_after();
        for (int kh = 0; kh < 2; ++kh)
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem + TM_LOG, desc_kmajor_sw128(mb0 + (nh * 2 + kh) * SLAB + kk * 32),
                      desc_kmajor_sw128(x0 + kh * SLAB + kk * 32), id_kk, (kh | kk) != 0);
        umma_commit(log_full);
        ++log_n;
      };
      // the two gradient GEMMs of a bin half, once the epilogue has written its dl^T rows
      auto grad_mma = [&](uint32_t x0, int it, int nh) {
        mbar_wait(&dl_full[nh], it & 1);
        tc_fence_after();
        // dMb[nh] [128 bins x 128 ch] += dl^T (A, K = pixels: two 64-pixel slabs) . x3 (B MN-major: N = channels, K = pixel rows)
        for (int ph = 0; ph < 2; ++ph)
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem + TM_DMB + nh * 128, desc_kmajor_sw128(dl0 + ph * DLSLAB + nh * 128 * 128 + kk * 32),
                      desc_mn(x0 + (ph * 4 + kk) * 2048, SLAB), id_kn, dmb_acc || (ph | kk) != 0);
        // dx3^T [128 ch x 128 px] (+)= Mb2[nh]^T (A MN-major: M = channels, K = bin rows) . dl^T (B MN-major: N = pixels, K = bin rows)
        if (nh == 0) {
          mbar_wait(dx_empty, (it & 1) ^ 1);
          tc_fence_after();
        }
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)
          umma_bf16(tmem + TM_DX, desc_mn(mb0 + nh * 2 * SLAB + kk * 2048, SLAB),
                    desc_mn(dl0 + (nh * 128 + kk * 16) * 128, DLSLAB), id_nn, (nh | kk) != 0);
        umma_commit(&dl_empty[nh]);
      };
      // Order per tile:  L(it,0) [issued ahead, at the end of tile it-1, when the image does not change]  L(it,1)  G(it,0)
      // L(it+1,0)  G(it,1): the logits of the NEXT half are always in flight while the epilogue turns the current half into dl.
      for (int it = 0; it < n_my; ++it) {
        const int t = t_begin + it, b = t / p.tiles_per_img;
        const int buf = it & 1;
        const uint32_t x0 = smem_u32(x3s + buf * 2 * SLAB);
        if (b != cur_b) {
          if (cur_b >= 0) {
            umma_commit(dmb_full);                    // dMb of the finished image is complete in TMEM
            umma_commit(mb_empty);                    // and nothing reads its Mb2 any more
            mbar_wait(dmb_empty, flush_n & 1);        // the epilogue has drained the accumulators
            tc_fence_after();
            ++flush_n;
          }
          mbar_wait(mb_full, img_n & 1);
          tc_fence_after();
          cur_b = b;
          ++img_n;
          dmb_acc = false;
        }
        if (!ahead) {
          mbar_wait(&x_full[buf], (it >> 1) & 1);
          tc_fence_after();
          logits_mma(x0, 0);
        }
        ahead = false;
        logits_mma(x0, 1);
        grad_mma(x0, it, 0);
        if (it + 1 < n_my && (t + 1) / p.tiles_per_img == b) {   // next tile, same image (same Mb2): its first logits half now
          mbar_wait(&x_full[buf ^ 1], ((it + 1) >> 1) & 1);
          tc_fence_after();
          logits_mma(smem_u32(x3s + (buf ^ 1) * 2 * SLAB), 0);
          ahead = true;
        }
        grad_mma(x0, it, 1);
        dmb_acc = true;
        umma_commit(dx_full);
        umma_commit(&x_empty[buf]);
      }
      umma_commit(dmb_full);   // last image of this CTA
    }
  } else if (warp >= 4) {
    // ================= epilogue: thread = one BIN (TMEM lane) x one half of the tile's pixels =================
    const int e = warp - 4, q = e & 3, ph = e >> 2;
    const int row = q * 32 + lane;                       // TMEM lane: bin within the half (logits, dMb) / channel (dx3^T)
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    float s_pg[2] = {0.f, 0.f}, s_dl[2] = {0.f, 0.f};    // per-bin sums over this CTA's pixels of the current image
    int cur_b = -1, run = -1, flush_n = 0;
    uint32_t log_n = 0;
    auto flush = [&](int b) {
      // dMb of image b (partial: this CTA's tiles) -> workspace; per-bin sums next to it
      mbar_wait(dmb_full, flush_n & 1);
      tc_fence_after();
      const long long slot = (long long)blockIdx.x * p.runs + run;
      float* dst = p.part_dmb + slot * (NB * C);
#pragma unroll 1
      for (int nh = 0; nh < 2; ++nh) {
        uint32_t r0[32], r1[32];
        tmem_ld32(tmem + lane_base + TM_DMB + nh * 128 + ph * 64, r0);
        tmem_ld32(tmem + lane_base + TM_DMB + nh * 128 + ph * 64 + 32, r1);
        tmem_ld_wait();
        float4* o = reinterpret_cast<float4*>(dst + (long long)(nh * 128 + row) * C + ph * 64);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          o[j] = make_float4(__uint_as_float(r0[4 * j]), __uint_as_float(r0[4 * j + 1]), __uint_as_float(r0[4 * j + 2]),
                             __uint_as_float(r0[4 * j + 3]));
#pragma unroll
        for (int j = 0; j < 8; ++j)
          o[8 + j] = make_float4(__uint_as_float(r1[4 * j]), __uint_as_float(r1[4 * j + 1]), __uint_as_float(r1[4 * j + 2]),
                                 __uint_as_float(r1[4 * j + 3]));
        float* ps = p.part_sums + (slot * 2 + ph) * 2 * NB;
        ps[nh * 128 + row] = s_pg[nh];
        ps[NB + nh * 128 + row] = s_dl[nh];
        s_pg[nh] = s_dl[nh] = 0.f;
      }
      if (threadIdx.x == 128) p.part_img[slot] = b;
      tc_fence_before();
      mbar_arrive(dmb_empty);
      ++flush_n;
    };
    for (int it = 0; it < n_my; ++it) {
      const int t = t_begin + it, b = t / p.tiles_per_img, px0 = (t - b * p.tiles_per_img) * TP;
      if (b != cur_b) {
        if (cur_b >= 0) flush(cur_b);
        cur_b = b;
        ++run;
      }
      const int buf = it & 1;
      mbar_wait(&v_full[buf], (it >> 1) & 1);
      const float4* vv = vec + buf * 128 + ph * 64;
#pragma unroll 1
      for (int nh = 0; nh < 2; ++nh) {
        const int k = nh * 128 + row;
        const float bl = __ldg(p.blog2 + (long long)b * NB + k), ck = __ldg(p.centres + (long long)b * NB + k);
        mbar_wait(log_full, log_n & 1);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        tmem_ld32(tmem + lane_base + TM_LOG + ph * 64, r0);
        tmem_ld32(tmem + lane_base + TM_LOG + ph * 64 + 32, r1);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(log_empty);          // the issuer may overwrite the logits with the next half
        ++log_n;
        mbar_wait(&dl_empty[nh], (it & 1) ^ 1);   // last tile's MMAs have finished reading these dl^T rows
        uint8_t* drow = dls + ph * DLSLAB + k * 128;
        float spg = 0.f, sdl = 0.f;
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          float d[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int px = ch * 8 + j;
            const float4 v = vv[px];                                    // broadcast: every lane of the warp reads the same pixel
            const float l = __uint_as_float(px < 32 ? r0[px] : r1[px - 32]);
            const float pg = ex2(l + bl - v.x) * v.y;                   // g * p_k
            d[j] = pg * (ck - v.z);                                     // d loss / d logit_k
            spg += pg;
            sdl += d[j];
          }
          *reinterpret_cast<uint4*>(drow + ((ch ^ (k & 7)) << 4)) =
              make_uint4(pack_bf16(d[0], d[1]), pack_bf16(d[2], d[3]), pack_bf16(d[4], d[5]), pack_bf16(d[6], d[7]));
        }
        s_pg[nh] += spg;
        s_dl[nh] += sdl;
        fence_proxy_async_smem();
        mbar_arrive(&dl_full[nh]);
      }
      mbar_arrive(&v_empty[buf]);
      // dx3^T tile: lane = channel, columns = pixels -> ln2 * acc -> bf16 -> 128 contiguous bytes of the channel-major output
      mbar_wait(dx_full, it & 1);
      tc_fence_after();
      {
        uint32_t r0[32], r1[32];
        tmem_ld32(tmem + lane_base + TM_DX + ph * 64, r0);
        tmem_ld32(tmem + lane_base + TM_DX + ph * 64 + 32, r1);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(dx_empty);
        uint4* o = reinterpret_cast<uint4*>(p.dx3T + (long long)row * p.ld_dx + (long long)b * p.P + px0 + ph * 64);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          o[j] = make_uint4(pack_bf16(LN2 * __uint_as_float(r0[8 * j]), LN2 * __uint_as_float(r0[8 * j + 1])),
                            pack_bf16(LN2 * __uint_as_float(r0[8 * j + 2]), LN2 * __uint_as_float(r0[8 * j + 3])),
                            pack_bf16(LN2 * __uint_as_float(r0[8 * j + 4]), LN2 * __uint_as_float(r0[8 * j + 5])),
                            pack_bf16(LN2 * __uint_as_float(r0[8 * j + 6]), LN2 * __uint_as_float(r0[8 * j + 7])));
#pragma unroll
        for (int j = 0; j < 4; ++j)
          o[4 + j] = make_uint4(pack_bf16(LN2 * __uint_as_float(r1[8 * j]), LN2 * __uint_as_float(r1[8 * j + 1])),
                                pack_bf16(LN2 * __uint_as_float(r1[8 * j + 2]), LN2 * __uint_as_float(r1[8 * j + 3])),
                                pack_bf16(LN2 * __uint_as_float(r1[8 * j + 4]), LN2 * __uint_as_float(r1[8 * j + 5])),
                                pack_bf16(LN2 * __uint_as_float(r1[8 * j + 6]), LN2 * __uint_as_float(r1[8 * j + 7])));
      }
    }
    if (cur_b >= 0) flush(cur_b);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, B_TM_COLS);
}

// dMb[b] = sum of the partial accumulators of image b in (CTA, run) order; d centres[b] likewise; fixed order -> deterministic
__global__ void tail_bwd_reduce_kernel(const float* __restrict__ part_dmb, const float* __restrict__ part_sums,
                                       const int* __restrict__ part_img, int n_slots, float* __restrict__ dMb,
                                       float* __restrict__ dcent, float* __restrict__ dsum /*[B,256]*/) {
  const int k = blockIdx.x, b = blockIdx.y, c = threadIdx.x;
  float acc = 0.f, s1 = 0.f, s2 = 0.f;
  for (int s = 0; s < n_slots; ++s) {
    if (part_img[s] != b) continue;
    acc += part_dmb[((long long)s * NB + k) * C + c];
    if (c == 0) {
      const float* ps = part_sums + (long long)s * 4 * NB;
      s1 += ps[k] + ps[2 * NB + k];
      s2 += ps[NB + k] + ps[3 * NB + k];
    }
  }
  dMb[((long long)b * NB + k) * C + c] = acc;
  if (c == 0) {
    dcent[(long long)b * NB + k] = s1;
    dsum[(long long)b * NB + k] = s2;
  }
}

// grad_bias[k] = beta * grad_bias[k] + sum_b dsum[b, k]   (fixed order)
__global__ void tail_bwd_bias_kernel(const float* __restrict__ dsum, float* __restrict__ gbias, int B, float beta) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= NB) return;
  float s = 0.f;
  for (int b = 0; b < B; ++b) s += dsum[(long long)b * NB + k];
  gbias[k] = (beta != 0.f ? beta * gbias[k] : 0.f) + s;
}

int n_sm_cached() {
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n_sm = 0;
  }
  return n_sm;
}

struct Plan { int grid, per_cta, runs, n_tiles, tiles_per_img; };
Plan make_plan(int B, long long P) {
  Plan pl{};
  pl.tiles_per_img = (int)(P / TP);
  pl.n_tiles = B * pl.tiles_per_img;
  const int n_sm = n_sm_cached();
  pl.grid = pl.n_tiles < n_sm ? pl.n_tiles : n_sm;
  pl.per_cta = (pl.n_tiles + pl.grid - 1) / pl.grid;
  pl.runs = (pl.per_cta + pl.tiles_per_img - 1) / pl.tiles_per_img + 1;   // images a CTA's contiguous tile range can touch
  return pl;
}

}  // namespace

extern "C" int ocv_tail_fused_fwd_bf16(const void* x3_nhwc_bf16, const void* Mb2_bf16, const float* blog2, const float* centres,
                                       float* depth, float* stat_max, float* stat_sum, int B, long long P, void* stream) {
  OCV_CHECK_ARG(x3_nhwc_bf16 && Mb2_bf16 && blog2 && centres && depth && stat_max && stat_sum && B > 0 && P > 0);
  OCV_CHECK_SHAPE(P % TP == 0 && (long long)B * P < (1ll << 31));
  OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(blog2) | reinterpret_cast<uintptr_t>(centres)) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  CUtensorMap tmX3, tmMb;
  if (int s = tc::make_tmap_2d_bf16(&tmX3, x3_nhwc_bf16, (uint64_t)B * P, C, C, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  if (int s = tc::make_tmap_2d_bf16(&tmMb, Mb2_bf16, (uint64_t)B * NB, C, C, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(tail_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  const Plan pl = make_plan(B, P);
  if (pl.grid <= 0) return OCV_E_DRIVER;
  TailParams p{};
  p.blog2 = blog2; p.centres = centres; p.depth = depth; p.smax = stat_max; p.ssum = stat_sum;
  p.B = B; p.P = (int)P; p.tiles_per_img = pl.tiles_per_img; p.n_tiles = pl.n_tiles;
  tail_fwd_kernel<<<pl.grid, THREADS, F_SMEM, ocv_stream(stream)>>>(tmX3, tmMb, p);
  return ocv_launch_status();
}

extern "C" long long ocv_tail_fused_bwd_workspace(int B, long long P) {
  if (B <= 0 || P <= 0 || P % TP != 0) return 0;
  const Plan pl = make_plan(B, P);
  const long long slots = (long long)pl.grid * pl.runs;
  return slots * (NB * C + 4 * NB) * 4 + ((slots * 4 + 255) / 256) * 256 + (long long)B * NB * 4 + 1024;
}

extern "C" int ocv_tail_fused_bwd_bf16(const void* x3_nhwc_bf16, const void* Mb2_bf16, const float* blog2, const float* centres,
                                       const float* depth, const float* stat_max, const float* stat_sum, const float* grad_depth,
                                       void* dx3T_bf16, long long ld_dx3T, float* dMb, float* grad_centres, float* grad_bias,
                                       float bias_beta, int B, long long P, void* workspace, void* stream) {
  OCV_CHECK_ARG(x3_nhwc_bf16 && Mb2_bf16 && blog2 && centres && depth && stat_max && stat_sum && grad_depth && dx3T_bf16 && dMb &&
                grad_centres && workspace && B > 0 && P > 0);
  OCV_CHECK_SHAPE(P % TP == 0 && (long long)B * P < (1ll << 31) && ld_dx3T >= (long long)B * P && ld_dx3T % 8 == 0);
  OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(dx3T_bf16) | reinterpret_cast<uintptr_t>(depth) | reinterpret_cast<uintptr_t>(stat_max) |
                    reinterpret_cast<uintptr_t>(stat_sum) | reinterpret_cast<uintptr_t>(grad_depth) |
                    reinterpret_cast<uintptr_t>(workspace)) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  CUtensorMap tmX3, tmMb;
  if (int s = tc::make_tmap_2d_bf16(&tmX3, x3_nhwc_bf16, (uint64_t)B * P, C, C, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  if (int s = tc::make_tmap_2d_bf16(&tmMb, Mb2_bf16, (uint64_t)B * NB, C, C, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(tail_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, B_SMEM);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  const Plan pl = make_plan(B, P);
  if (pl.grid <= 0) return OCV_E_DRIVER;
  const long long slots = (long long)pl.grid * pl.runs;
  uint8_t* w = static_cast<uint8_t*>(workspace);
  TailParams p{};
  p.blog2 = blog2; p.centres = centres; p.depth = const_cast<float*>(depth); p.smax = const_cast<float*>(stat_max);
  p.ssum = const_cast<float*>(stat_sum); p.gdepth = grad_depth;
  p.dx3T = static_cast<__nv_bfloat16*>(dx3T_bf16); p.ld_dx = ld_dx3T;
  p.part_dmb = reinterpret_cast<float*>(w);
  p.part_sums = p.part_dmb + slots * NB * C;
  p.part_img = reinterpret_cast<int*>(p.part_sums + slots * 4 * NB);
  float* dsum = reinterpret_cast<float*>(w + slots * (NB * C + 4 * NB) * 4 + ((slots * 4 + 255) / 256) * 256);
  p.B = B; p.P = (int)P; p.tiles_per_img = pl.tiles_per_img; p.n_tiles = pl.n_tiles; p.runs = pl.runs;
  tail_bwd_kernel<<<pl.grid, THREADS, B_SMEM, ocv_stream(stream)>>>(tmX3, tmMb, p);
  if (int s = ocv_launch_status()) return s;
  tail_bwd_reduce_kernel<<<dim3(NB, B), C, 0, ocv_stream(stream)>>>(p.part_dmb, p.part_sums, p.part_img, (int)slots, dMb, grad_centres, dsum);
  if (int s = ocv_launch_status()) return s;
  if (grad_bias) {
    tail_bwd_bias_kernel<<<1, NB, 0, ocv_stream(stream)>>>(dsum, grad_bias, B, bias_beta);
    return ocv_launch_status();
  }
  return OCV_OK;
}
