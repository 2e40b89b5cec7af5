// Pixel-path kernels, fp32 SIMT (parity path): conv3x3 (NCHW) and the fused tail
// (1x1 conv -> softmax over bins -> sum p*centre).  Contracts: include/objcavit_b200.h
#include "common.cuh"
#include <math.h>

namespace {

// ------------------------------------------------------------------------------------------------
// conv3x3, C_in = C_out = 128.  CTA: 128 out-channels x (4 rows x 32 cols); thread: 8 out-ch x 8 px.
// ------------------------------------------------------------------------------------------------
constexpr int CV_ROWS = 4, CV_COLS = 32, CV_CC = 8, CV_XS = 36;  // XS: padded row pitch of the halo tile

__global__ void __launch_bounds__(256)
conv3x3_f32_kernel(const float* __restrict__ X, const float* __restrict__ W, const float* __restrict__ bias,
                   float* __restrict__ Y, int C, int h, int w) {
  __shared__ __align__(16) float Xs[CV_CC][CV_ROWS + 2][CV_XS];
  __shared__ __align__(16) float Ws[CV_CC * 9][128];
  const int tid = threadIdx.x;
  const int b = blockIdx.z;
  const int y0 = blockIdx.y * CV_ROWS, x0 = blockIdx.x * CV_COLS;
  const int og = tid / 16, pg = tid % 16;
  const int prow = pg / 4, px = (pg % 4) * 8;
  const float* Xb = X + (long long)b * C * h * w;
  float acc[8][8] = {};
  for (int c0 = 0; c0 < C; c0 += CV_CC) {
    __syncthreads();
    // halo tile: 8 channels x 6 rows x 34 cols (zero padded)
    for (int i = tid; i < CV_CC * (CV_ROWS + 2) * (CV_COLS + 2); i += 256) {
      const int c = i / ((CV_ROWS + 2) * (CV_COLS + 2));
      const int r = (i / (CV_COLS + 2)) % (CV_ROWS + 2);
      const int cc = i % (CV_COLS + 2);
      const int yy = y0 + r - 1, xx = x0 + cc - 1;
      float v = 0.f;
      if (yy >= 0 && yy < h && xx >= 0 && xx < w) v = Xb[((long long)(c0 + c) * h + yy) * w + xx];
      Xs[c][r][cc] = v;
    }
    // weight tile: Ws[(c,tap)][o] = W[o][c0+c][tap]; thread <-> o so the smem stores are conflict-free
    {
      const int o = tid % 128;
      for (int r = tid / 128; r < CV_CC * 9; r += 2) Ws[r][o] = W[((long long)o * C + c0) * 9 + r];
    }
    __syncthreads();
#pragma unroll 1
    for (int c = 0; c < CV_CC; ++c) {
#pragma unroll
      for (int dy = 0; dy < 3; ++dy) {
        float xv[10];
        const float* xr = &Xs[c][prow + dy][px];
        const float4 a0 = *reinterpret_cast<const float4*>(xr);
        const float4 a1 = *reinterpret_cast<const float4*>(xr + 4);
        xv[0] = a0.x; xv[1] = a0.y; xv[2] = a0.z; xv[3] = a0.w;
        xv[4] = a1.x; xv[5] = a1.y; xv[6] = a1.z; xv[7] = a1.w;
        xv[8] = xr[8]; xv[9] = xr[9];
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
          const float* wr = &Ws[c * 9 + dy * 3 + dx][og * 8];
          const float4 w0 = *reinterpret_cast<const float4*>(wr);
          const float4 w1 = *reinterpret_cast<const float4*>(wr + 4);
          const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(wv[i], xv[j + dx], acc[i][j]);
        }
      }
    }
  }
  const int yy = y0 + prow;
  if (yy < h) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int o = og * 8 + i;
      const float bo = bias ? bias[o] : 0.f;
      float* yr = Y + (((long long)b * 128 + o) * h + yy) * w;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int xx = x0 + px + j;
        if (xx < w) yr[xx] = acc[i][j] + bo;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// tail: logits = Wout @ ram + bout -> softmax over K=256 -> sum p*centre.  CTA: 64 px, 256 threads.
// ------------------------------------------------------------------------------------------------
constexpr int TL_PX = 64, TL_K = 256, TL_CC = 16, TL_LS = TL_PX + 1;

__global__ void __launch_bounds__(256)
tail_f32_kernel(const float* __restrict__ ram, const float* __restrict__ Wout, const float* __restrict__ bout,
                const float* __restrict__ centres, float* __restrict__ depth, int Q, long long P) {
  extern __shared__ __align__(16) float smem[];
  float* Ls = smem;                      // [256][65] logits
  float* Rs = Ls + TL_K * TL_LS;         // [TL_CC][64] ram chunk
  float* Wsm = Rs + TL_CC * TL_PX;       // [TL_CC][256] Wout chunk (transposed)
  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const long long p0 = (long long)blockIdx.x * TL_PX;
  const int bg = tid / 8, pg = tid % 8;
  const float* rb = ram + (long long)b * Q * P;
  float acc[8][8] = {};
  for (int c0 = 0; c0 < Q; c0 += TL_CC) {
    __syncthreads();
    for (int i = tid; i < TL_CC * TL_PX; i += 256) {
      const int c = i / TL_PX, p = i % TL_PX;
      Rs[i] = (p0 + p < P && c0 + c < Q) ? rb[(long long)(c0 + c) * P + p0 + p] : 0.f;
    }
    for (int j = 0; j < TL_CC; ++j) Wsm[j * TL_K + tid] = (c0 + j < Q) ? Wout[(long long)tid * Q + c0 + j] : 0.f;
    __syncthreads();
#pragma unroll
    for (int c = 0; c < TL_CC; ++c) {
      const float4 w0 = *reinterpret_cast<const float4*>(&Wsm[c * TL_K + bg * 8]);
      const float4 w1 = *reinterpret_cast<const float4*>(&Wsm[c * TL_K + bg * 8 + 4]);
      const float4 r0 = *reinterpret_cast<const float4*>(&Rs[c * TL_PX + pg * 8]);
      const float4 r1 = *reinterpret_cast<const float4*>(&Rs[c * TL_PX + pg * 8 + 4]);
      const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
      const float rv[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(wv[i], rv[j], acc[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int k = bg * 8 + i;
    const float bk = bout ? bout[k] : 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) Ls[k * TL_LS + pg * 8 + j] = acc[i][j] + bk;
  }
  __syncthreads();
  const int warp = tid / 32, lane = tid % 32;
  float cen[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) cen[i] = centres[(long long)b * TL_K + lane + 32 * i];
  for (int pp = 0; pp < TL_PX / 8; ++pp) {
    const int p = warp * (TL_PX / 8) + pp;
    float v[8], mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      v[i] = Ls[(lane + 32 * i) * TL_LS + p];
      mx = fmaxf(mx, v[i]);
    }
    mx = warp_max(mx);
    float s = 0.f, sc = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float e = expf(v[i] - mx);
      s += e;
      sc = fmaf(e, cen[i], sc);
    }
    s = warp_sum(s);
    sc = warp_sum(sc);
    if (lane == 0 && p0 + p < P) depth[(long long)b * P + p0 + p] = sc / s;
  }
}

constexpr int TL_SMEM = (TL_K * TL_LS + TL_CC * TL_PX + TL_CC * TL_K) * (int)sizeof(float);

}  // namespace

extern "C" int ocv_conv3x3_f32(const float* X, const float* W, const float* bias, float* Y, int B, int C, int h, int w,
                               void* stream) {
  OCV_CHECK_ARG(X && W && Y && B > 0 && h > 0 && w > 0);
  OCV_CHECK_SHAPE(C == 128 && C % CV_CC == 0 && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  dim3 grid(ceil_div(w, CV_COLS), ceil_div(h, CV_ROWS), B);
  conv3x3_f32_kernel<<<grid, 256, 0, ocv_stream(stream)>>>(X, W, bias, Y, C, h, w);
  return ocv_launch_status();
}

extern "C" int ocv_tail_f32(const float* ram, const float* Wout, const float* bout, const float* centres, float* depth,
                            int B, int Q, int K, long long P, void* stream) {
  OCV_CHECK_ARG(ram && Wout && centres && depth && B > 0 && P > 0 && Q > 0);
  OCV_CHECK_SHAPE(K == TL_K && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tail_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TL_SMEM);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  dim3 grid((unsigned)ceil_div<long long>(P, TL_PX), B);
  tail_f32_kernel<<<grid, 256, TL_SMEM, ocv_stream(stream)>>>(ram, Wout, bout, centres, depth, Q, P);
  return ocv_launch_status();
}
