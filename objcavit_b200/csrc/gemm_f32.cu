// fp32 SIMT dense stages of the token path (parity path): generic linear + patch embedding.
// See include/objcavit_b200.h for the contract of each entry point.
#include "common.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16, NT = 256;

struct PatchGeom {  // im2col view of an NCHW feature map for the PxP/stride-P patch-embed conv
  int C, h, w, P, gw, S;
};

// A_MODE 0: A is a dense [M,K] matrix.  A_MODE 1: A[m][k] is gathered from the NCHW feature map.
template <int A_MODE>
__device__ __forceinline__ float load_a(const float* __restrict__ A, int m, int k, int lda, const PatchGeom& g) {
  if (A_MODE == 0) return A[(long long)m * lda + k];
  const int b = m / g.S, s = m - b * g.S;
  const int ty = s / g.gw, tx = s - ty * g.gw;
  const int pp = g.P * g.P;
  const int c = k / pp, r = k - c * pp;
  const int ky = r / g.P, kx = r - ky * g.P;
  return A[(((long long)b * g.C + c) * g.h + (ty * g.P + ky)) * g.w + tx * g.P + kx];
}

template <int A_MODE, int TRANSB>
__global__ void __launch_bounds__(NT)
sgemm_kernel(const float* __restrict__ A, const float* __restrict__ Bm, const float* __restrict__ bias,
             const float* __restrict__ residual, float* __restrict__ C, int M, int N, int K, int lda, int ldb,
             int ldc, int act, long long strideA, long long strideB, long long strideC, int k_per_split,
             PatchGeom g) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  int k_begin = 0, k_end = K;
  if (k_per_split > 0) {  // split-K: blockIdx.z is the split, partials go to C + z*M*N
    k_begin = blockIdx.z * k_per_split;
    k_end = min(K, k_begin + k_per_split);
    C += (long long)blockIdx.z * M * ldc;
  } else {
    A += blockIdx.z * strideA;
    Bm += blockIdx.z * strideB;
    C += blockIdx.z * strideC;
    if (residual) residual += blockIdx.z * strideC;
  }
  const int ty = tid / 16, tx = tid % 16;
  float acc[4][4] = {};
  // Register double buffering: the global loads of tile t+1 are issued before the FMAs of tile t (token-sized problems have
  // fewer tiles than SMs, so no other CTA hides the latency).
  const int r4 = tid / 4, k4 = (tid % 4) * 4;      // (row, 4 consecutive k) of a K-contiguous operand
  const int k16 = tid / 16, c16 = (tid % 16) * 4;  // (k, 4 consecutive n) of Bm [K,N]
  float ra[4], rb[4];
  auto fetch = [&](int k0) {
    const int m = m0 + r4;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + k4 + j;
      ra[j] = (m < M && k < k_end) ? load_a<A_MODE>(A, m, k, lda, g) : 0.f;
    }
    if (TRANSB) {  // Bm [N,K]
      const int n = n0 + r4;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = k0 + k4 + j;
        rb[j] = (n < N && k < k_end) ? Bm[(long long)n * ldb + k] : 0.f;
      }
    } else {  // Bm [K,N]
      const int k = k0 + k16;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = n0 + c16 + j;
        rb[j] = (n < N && k < k_end) ? Bm[(long long)k * ldb + n] : 0.f;
      }
    }
  };
  fetch(k_begin);
  for (int k0 = k_begin; k0 < k_end; k0 += BK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      As[k4 + j][r4] = ra[j];
      if (TRANSB) Bs[k4 + j][r4] = rb[j];
      else Bs[k16][c16 + j] = rb[j];
    }
    __syncthreads();
    if (k0 + BK < k_end) fetch(k0 + BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j];
      if (k_per_split == 0) {
        if (bias) v += bias[n];
        if (act == OCV_ACT_RELU) v = fmaxf(v, 0.f);
        else if (act == OCV_ACT_LEAKY_RELU) v = v > 0.f ? v : 0.01f * v;
        if (residual) v += residual[(long long)m * ldc + n];
      }
      C[(long long)m * ldc + n] = v;
    }
  }
}

// tokens[m,n] = sum_z part[z,m,n] + bias[n] + pos[m % S, n]   (fixed summation order -> deterministic)
__global__ void splitk_reduce_kernel(const float* __restrict__ part, const float* __restrict__ bias,
                                     const float* __restrict__ pos, float* __restrict__ out, int M, int N, int S,
                                     int splits) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)M * N) return;
  const int m = (int)(i / N), n = (int)(i - (long long)m * N);
  float v = 0.f;
  for (int z = 0; z < splits; ++z) v += part[(long long)z * M * N + i];
  if (bias) v += bias[n];
  if (pos) v += pos[(long long)(m % S) * N + n];
  out[i] = v;
}

int patch_splits(int M, int N, int K) {
  const int tiles = ceil_div(M, BM) * ceil_div(N, BN);
  int s = ceil_div(296, tiles);
  if (s < 1) s = 1;
  if (s > 64) s = 64;
  while (s > 1 && (K / s) < 256) --s;
  return s;
}

}  // namespace

extern "C" int ocv_linear_f32(const float* A, const float* Bm, const float* bias, const float* residual, float* C,
                              int M, int N, int K, int lda, int ldb, int ldc, int transB, int act, int batch,
                              long long strideA, long long strideB, long long strideC, void* stream) {
  OCV_CHECK_ARG(A && Bm && C);
  OCV_CHECK_ARG(M > 0 && N > 0 && K > 0 && batch > 0);
  OCV_CHECK_SHAPE(batch <= 65535);
  if (int s = ocv_require_sm100()) return s;
  dim3 grid(ceil_div(N, BN), ceil_div(M, BM), batch);
  PatchGeom g{};
  if (transB)
    sgemm_kernel<0, 1><<<grid, NT, 0, ocv_stream(stream)>>>(A, Bm, bias, residual, C, M, N, K, lda, ldb, ldc, act,
                                                          strideA, strideB, strideC, 0, g);
  else
    sgemm_kernel<0, 0><<<grid, NT, 0, ocv_stream(stream)>>>(A, Bm, bias, residual, C, M, N, K, lda, ldb, ldc, act,
                                                          strideA, strideB, strideC, 0, g);
  return ocv_launch_status();
}

extern "C" long long ocv_patch_embed_f32_workspace(int B, int C, int h, int w, int P, int E) {
  if (B <= 0 || C <= 0 || P <= 0 || E <= 0 || h < P || w < P) return 0;
  const int S = (h / P) * (w / P);
  const int M = B * S, K = C * P * P;
  return (long long)patch_splits(M, E, K) * M * E * (long long)sizeof(float);
}

extern "C" int ocv_patch_embed_f32(const float* X, const float* W, const float* bias, const float* pos,
                                   float* tokens, int B, int C, int h, int w, int P, int E, void* workspace,
                                   void* stream) {
  OCV_CHECK_ARG(X && W && tokens && workspace);
  OCV_CHECK_ARG(B > 0 && C > 0 && P > 0 && E > 0);
  OCV_CHECK_SHAPE(h >= P && w >= P);
  if (int s = ocv_require_sm100()) return s;
  PatchGeom g{C, h, w, P, w / P, (h / P) * (w / P)};
  const int M = B * g.S, K = C * P * P;
  const int splits = patch_splits(M, E, K);
  int kps = ceil_div(ceil_div(K, splits), BK) * BK;
  dim3 grid(ceil_div(E, BN), ceil_div(M, BM), splits);
  float* part = static_cast<float*>(workspace);
  sgemm_kernel<1, 1><<<grid, NT, 0, ocv_stream(stream)>>>(X, W, nullptr, nullptr, part, M, E, K, 0, K, E, 0, 0, 0, 0,
                                                        kps, g);
  if (int s = ocv_launch_status()) return s;
  const long long n = (long long)M * E;
  splitk_reduce_kernel<<<(unsigned)ceil_div<long long>(n, 256), 256, 0, ocv_stream(stream)>>>(part, bias, pos, tokens,
                                                                                           M, E, g.S, splits);
  return ocv_launch_status();
}
