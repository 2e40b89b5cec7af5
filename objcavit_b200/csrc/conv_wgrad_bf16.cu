// conv3x3 weight (and bias) gradient of the training step as an IMPLICIT GEMM on tcgen05 — no im2col buffer.
//
//   dW[co, ci, ky, kx] = sum_{b,y,x} dY[b,y,x,co] * X[b, y+ky-1, x+kx-1, ci]          db[co] = sum_{b,y,x} dY[b,y,x,co]
//
// reference: autograd of nn.Conv2d(128, 128, 3, padding=1) (modules/ObjCAViT.py:374, modules/miniViT.py:25).
// replaces (bf16 training path): fp32 -> planar bf16 cast of the features, ocv_im2col3x3_bf16 (1.04 GB written per step at
// B=8 208x272) and the 128 x 1160 x (B P) GEMM that read it back: 2.7 GB of HBM traffic -> 0.7 GB.
//
// The contraction index is the PIXEL.  Both operands already exist with the pixel as the row index of 128-byte rows, which is
// what tcgen05 reads as K without any copy:
//   A = dY^T [128 co, B P] channel-major bf16 (what ocv_tail_fused_bwd_bf16 writes), addressed as (x, image row, channel): TMA box
//       [128 channels x 64 pixels of one row] = the K-major A operand, 4 K-steps of 16 pixels, zero beyond the end of the row;
//   B = X NHWC bf16 [B,h,w,128]: a TMA box of one image row, 66 pixels x 64 channels per plane (two planes = 128 channels), rows =
//       pixels = K, 64 channels contiguous = N: the MN-major B operand (probes/probe_mn_major.cu case 0).  The three horizontal taps
//       are the SAME tile read through a start address shifted by 0 / 1 / 2 rows (case 7 / 8), image borders are TMA zero fill.
// A CTA owns one vertical tap ky (grid = 3 x chunks): it streams strips of up to 64 pixels of row y with row y+ky-1 of X and
// accumulates its three [128 co x 128 ci] taps in TMEM (384 columns) over its whole chunk of the batch; a fourth accumulator
// (N = 16 against a tile of ones) is the bias gradient.  Partials are written once per CTA and summed in fixed order by a second
// kernel (deterministic), which also scatters them into the [co][ci][ky][kx] weight layout.
// Contract: include/objcavit_b200.h.
#include "tc_common.cuh"
#include <stdlib.h>

namespace {

using namespace tc;

constexpr int C = 128;
constexpr int SLAB = 128 * 128;          // A tile: [128 co x 64 pixels] bf16
constexpr int HALO = 66;                 // pixels of one X row tile: 64 + one either side
constexpr int PLANE = 9 * 1024;          // 66 x 128 B = 8448 -> next multiple of 1024 (swizzle phase)
constexpr int STAGE = SLAB + 2 * PLANE;  // 34816
constexpr int NSTG = 5;
constexpr int ONES = 2048;               // [16 rows x 128 B] of bf16 ones
constexpr int THREADS = 192;             // warp 0 TMA, warp 1 MMA, warps 2..5 write the partials (warp % 4 = TMEM lane quarter)
constexpr int SMEM_BYTES = NSTG * STAGE + ONES + 1024 + 256;
constexpr uint32_t TM_BIAS = 384, TM_COLS = 512;

struct WgradParams {
  float* partial;     // [grid][3 kx][128 co][128 ci] fp32, then [grid][128] bias sums behind them
  float* pbias;
  int B, h, w, nsx, sw;            // strips per row, strip width (16 / 32 / 48 / 64 pixels = K-steps used)
  long long P;                     // h * w
  int n_chunks, strips_total;
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

__global__ void __launch_bounds__(THREADS, 1)
conv3x3_wgrad_kernel(const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmX, WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* stages = smem;                               // [NSTG] { A slab | X plane (ci 0..63) | X plane (ci 64..127) }
  uint8_t* ones = smem + NSTG * STAGE;
  uint64_t* full = reinterpret_cast<uint64_t*>(ones + ONES);
  uint64_t* empty = full + NSTG;
  uint64_t* done = empty + NSTG;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int ky = blockIdx.x % 3, chunk = blockIdx.x / 3;
  const int per = (p.strips_total + p.n_chunks - 1) / p.n_chunks;
  const int s_begin = min(p.strips_total, chunk * per), s_end = min(p.strips_total, s_begin + per);
  const int n_my = s_end - s_begin;
  const int nk = p.sw / 16;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmD);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < NSTG; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, TM_COLS);
  for (int i = threadIdx.x; i < ONES / 4; i += THREADS) reinterpret_cast<uint32_t*>(ones)[i] = 0x3F803F80u;   // bf16 1.0 pairs
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      for (int i = 0; i < n_my; ++i) {
        const int st = i % NSTG;
        mbar_wait(&empty[st], ((i / NSTG) & 1) ^ 1);
        const int s = s_begin + i;
        const int row = s / p.nsx, xi = s - row * p.nsx;      // row = b * h + y
        const int b = row / p.h, y = row - b * p.h;
        const int x0 = xi * p.sw;
        uint8_t* sa = stages + st * STAGE;
        mbar_arrive_expect_tx(&full[st], SLAB + 2 * HALO * 128);
        // dY^T is addressed as (x, image row, channel): a strip that runs past the end of the image row is zero-filled there
        tma_load_3d(sa, &tmD, &full[st], x0, row, 0);
        tma_load_4d(sa + SLAB, &tmX, &full[st], 0, x0 - 1, y + ky - 1, b);
        tma_load_4d(sa + SLAB + PLANE, &tmX, &full[st], 64, x0 - 1, y + ky - 1, b);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 128, 0, 1);    // A K-major (dY^T), B MN-major (X rows)
      constexpr uint32_t idesc_b = make_idesc_bf16(128, 16, 0, 1);
      const uint64_t ones_desc = make_smem_desc(smem_u32(ones), 1024, 1024, LAYOUT_SW128);
      for (int i = 0; i < n_my; ++i) {
        const int st = i % NSTG;
        mbar_wait(&full[st], (i / NSTG) & 1);
        tc_fence_after();
        const uint32_t a0 = smem_u32(stages + st * STAGE), x0 = a0 + SLAB;
        for (int kk = 0; kk < nk; ++kk) {
          const uint64_t da = desc_kmajor_sw128(a0 + kk * 32);
          const uint32_t acc = (i | kk) != 0;
#pragma unroll
          for (int kx = 0; kx < 3; ++kx)
            umma_bf16(tmem + kx * 128, da, make_smem_desc(x0 + (kk * 16 + kx) * 128, PLANE, 1024, LAYOUT_SW128), idesc, acc);
          umma_bf16(tmem + TM_BIAS, da, ones_desc, idesc_b, acc);
        }
        umma_commit(&empty[st]);
      }
      umma_commit(done);
    }
  } else {
    // ================= partials: TMEM -> global, once per CTA =================
    const int q = warp % 4;
    const int co = q * 32 + lane;
    mbar_wait(done, 0);
    tc_fence_after();
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    float* dst = p.partial + ((long long)blockIdx.x * 3 * C + co) * C;
    for (int kx = 0; kx < 3; ++kx)
      for (int cb = 0; cb < 4; ++cb) {
        uint32_t r[32];
        if (n_my > 0) {
          tmem_ld32(tmem + lane_base + kx * 128 + cb * 32, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = 0;
        }
        float4* o = reinterpret_cast<float4*>(dst + (long long)kx * C * C + cb * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          o[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                             __uint_as_float(r[4 * j + 3]));
      }
    {
      uint32_t r[32];
      float v = 0.f;
      if (n_my > 0) {
        tmem_ld32(tmem + lane_base + TM_BIAS, r);   // columns 0..15 hold the same sum; the rest of the 32 is never written
        tmem_ld_wait();
        v = __uint_as_float(r[0]);
      }
      p.pbias[(long long)blockIdx.x * C + co] = v;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, TM_COLS);
}

// gw[co][ci][ky][kx] = beta * gw + sum_chunks partial[chunk*3 + ky][kx][co][ci]  (fixed order);  gb[co] = beta * gb + sum_chunks bias
__global__ void conv3x3_wgrad_reduce_kernel(const float* __restrict__ partial, const float* __restrict__ pbias, float* __restrict__ gw,
                                            float* __restrict__ gb, int n_chunks, float beta) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;     // (tap, co, ci), ci fastest: coalesced partial reads
  if (idx < 9 * C * C) {
    const int tap = idx / (C * C), r = idx - tap * C * C;
    const int co = r / C, ci = r - co * C;
    const int ky = tap / 3, kx = tap - ky * 3;
    float acc = 0.f;
    for (int ch = 0; ch < n_chunks; ++ch) acc += partial[(((long long)(ch * 3 + ky) * 3 + kx) * C + co) * C + ci];
    float* d = gw + ((long long)co * C + ci) * 9 + tap;
    *d = beta != 0.f ? beta * *d + acc : acc;
  } else if (gb && idx < 9 * C * C + C) {
    const int co = idx - 9 * C * C;
    float acc = 0.f;
    for (int ch = 0; ch < n_chunks; ++ch) acc += pbias[(long long)(ch * 3) * C + co];   // the ky = 0 CTAs (every CTA sums all its pixels)
    gb[co] = beta != 0.f ? beta * gb[co] + acc : acc;
  }
}

int n_chunks_for(int strips) {
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n_sm = 148;
  }
  int n = n_sm / 3;
  if (n > strips) n = strips;
  return n < 1 ? 1 : n;
}

}  // namespace

extern "C" long long ocv_conv3x3_wgrad_bf16_workspace(int B, int h, int w) {
  if (B <= 0 || h <= 0 || w <= 0) return 0;
  const int n = n_chunks_for(1 << 30) * 3;
  return (long long)n * (3LL * C * C + C) * 4 + 256;
}

extern "C" int ocv_conv3x3_wgrad_bf16(const void* dyT_bf16, long long ld_dy, const void* x_nhwc_bf16, float* gw, float* gb, int B, int h,
                                      int w, float beta, void* workspace, void* stream) {
  OCV_CHECK_ARG(dyT_bf16 && x_nhwc_bf16 && gw && workspace && B > 0 && h > 0 && w > 0);
  OCV_CHECK_SHAPE(ld_dy >= (long long)B * h * w && ld_dy % 8 == 0 && w % 8 == 0 && (long long)B * h * w < (1ll << 31));
  OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(dyT_bf16) | reinterpret_cast<uintptr_t>(x_nhwc_bf16)) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(workspace) & 255) == 0);
  if (int s = ocv_require_sm100()) return s;
  // strip width: a multiple of 16 (<= 64).  Every strip costs a fixed stage (a 64-pixel dY box + two haloed feature planes through
  // TMA and one mbarrier round trip) whatever its width, every 16 pixels one K-step (4 MMAs).  Measured on the B200 at 8x208x272
  // (profiles/wgrad_sw_r03.txt): 16 -> 188 us, 32 -> 136, 48 -> 128, 64 -> 129, i.e. about 7 units per strip against 4 per K-step;
  // the earlier "fewest K-steps" rule chose 16 there.  Ties go to the wider strip.
  int sw = 64, best = -1;
  for (int cand = 64; cand >= 16; cand -= 16) {
    const int strips = ceil_div(w, cand);
    const int cost = 7 * strips + 4 * strips * (cand / 16);
    if (best < 0 || cost < best) { best = cost; sw = cand; }
  }
#ifdef OCV_PROBES   // measurement build only (probes/time_conv_wgrad.py): force a strip width
  if (const char* e = getenv("OCV_WGRAD_SW")) {
    const int v = atoi(e);
    if (v == 16 || v == 32 || v == 48 || v == 64) sw = v;
  }
#endif
  CUtensorMap tmD, tmX;
  {
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) return OCV_E_DRIVER;
    {  // dY^T [128, ld] seen as [co][image row][x]: box = 64 pixels of one row x all 128 channels
      cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)B * h, (cuuint64_t)C};
      cuuint64_t strides[2] = {(cuuint64_t)w * 2, (cuuint64_t)ld_dy * 2};
      cuuint32_t box[3] = {64, 1, 128};
      cuuint32_t estr[3] = {1, 1, 1};
      CUresult r = enc(&tmD, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(dyT_bf16), dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) return OCV_E_DRIVER;
    }
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)w * C * 2, (cuuint64_t)h * w * C * 2};
    cuuint32_t box[4] = {64, HALO, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&tmX, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x_nhwc_bf16), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return OCV_E_DRIVER;
  }
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(conv3x3_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  WgradParams p{};
  p.B = B; p.h = h; p.w = w; p.sw = sw; p.nsx = ceil_div(w, sw); p.P = (long long)h * w;
  p.strips_total = B * h * p.nsx;
  p.n_chunks = n_chunks_for(p.strips_total);
  p.partial = static_cast<float*>(workspace);
  p.pbias = p.partial + (long long)p.n_chunks * 3 * 3 * C * C;
  const int grid = p.n_chunks * 3;
  conv3x3_wgrad_kernel<<<grid, THREADS, SMEM_BYTES, ocv_stream(stream)>>>(tmD, tmX, p);
  if (int s = ocv_launch_status()) return s;
  const int total = 9 * C * C + C;
  conv3x3_wgrad_reduce_kernel<<<ceil_div(total, 256), 256, 0, ocv_stream(stream)>>>(p.partial, p.pbias, gw, gb, p.n_chunks, beta);
  return ocv_launch_status();
}
