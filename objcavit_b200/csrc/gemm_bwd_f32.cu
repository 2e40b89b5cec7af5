// fp32 SIMT GEMM for the backward pass of the parity path:
//   C[M,N] = beta * C + sum over (batch, k) of  op(A)[M,K] @ op(B)[K,N]
// with the operand forms the gradients need: transposed A (weight gradients: dW = dY^T X), transposed B (data gradients
// through nn.Linear weights), and two GATHERED B operands so that the im2col matrices of the two convolutions never exist:
//   B_PATCH : B[m][k] = X[b, c, ty*P+ky, tx*P+kx]        (patch-embedding weight gradient, k = (c, ky, kx))
//   B_IM2COL: B[p][j] = X[b, c, y+ky-1, x+kx-1] or 0     (conv3x3 weight gradient, j = (c, ky, kx), p = (y, x))
// Long reductions are split over K (and over the batch when the result is a sum over images); the partial tiles go to a
// caller workspace and are added in a fixed order (deterministic), then merged into C with beta (0 = overwrite, 1 =
// accumulate into a gradient buffer).  Contract: include/objcavit_b200.h (ocv_gemm_f32_ex).
#include "common.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16, NT = 256;

struct Gather {  // geometry of the gathered B operand
  int C, h, w, P, gw, S;
};

enum { B_N = 0, B_T = 1, B_PATCH = 2, B_IM2COL = 3 };

__device__ __forceinline__ float load_b(int mode, const float* __restrict__ Bm, int k, int n, int ldb, const Gather& g) {
  if (mode == B_N) return Bm[(long long)k * ldb + n];
  if (mode == B_T) return Bm[(long long)n * ldb + k];
  if (mode == B_PATCH) {  // k = token row m (b folded in), n = (c, ky, kx)
    const int b = k / g.S, s = k - b * g.S;
    const int ty = s / g.gw, tx = s - ty * g.gw;
    const int pp = g.P * g.P;
    const int c = n / pp, r = n - c * pp;
    const int ky = r / g.P, kx = r - ky * g.P;
    return Bm[(((long long)b * g.C + c) * g.h + (ty * g.P + ky)) * g.w + tx * g.P + kx];
  }
  // B_IM2COL: k = pixel p of ONE image (the batch offset is applied through strideB), n = (c, ky, kx)
  const int y = k / g.w, x = k - y * g.w;
  const int c = n / 9, r = n - c * 9;
  const int yy = y + r / 3 - 1, xx = x + (r % 3) - 1;
  if (yy < 0 || yy >= g.h || xx < 0 || xx >= g.w) return 0.f;
  return Bm[((long long)c * g.h + yy) * g.w + xx];
}

template <int TA, int BMODE>
__global__ void __launch_bounds__(NT)
sgemm_ex_kernel(const float* __restrict__ A, const float* __restrict__ Bm, float* __restrict__ C, int M, int N, int K, int lda,
                int ldb, int ldc, long long strideA, long long strideB, long long strideC, int splits, int k_per_split,
                int to_partials, Gather g) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int batch = blockIdx.z / splits, split = blockIdx.z - batch * splits;
  const int k_begin = split * k_per_split, k_end = min(K, k_begin + k_per_split);
  A += batch * strideA;
  Bm += batch * strideB;
  if (to_partials) C += (long long)blockIdx.z * M * N;  // partial plane [z][M][N] (dense)
  else C += batch * strideC;
  const int ldo = to_partials ? N : ldc;
  const int ty = tid / 16, tx = tid % 16;
  float acc[4][4] = {};
  // Register double buffering: the global loads of tile t+1 are issued before the FMAs of tile t, so a CTA that is alone on
  // its SM (token-sized problems: fewer tiles than SMs) still overlaps memory latency with compute.
  const int r4 = tid / 4, k4 = (tid % 4) * 4;     // "k fastest" operand: (row, 4 consecutive k)
  const int k16 = tid / 16, c16 = (tid % 16) * 4;  // "row fastest" operand: (k, 4 consecutive rows)
  float ra[4], rb[4];
  auto fetch = [&](int k0) {
    if (TA == 0) {  // A[m][k], k contiguous
      const int m = m0 + r4;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = k0 + k4 + j;
        ra[j] = (m < M && k < k_end) ? A[(long long)m * lda + k] : 0.f;
      }
    } else {  // A^T: element (m, k) at A[k*lda + m], m contiguous
      const int k = k0 + k16;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int m = m0 + c16 + j;
        ra[j] = (m < M && k < k_end) ? A[(long long)k * lda + m] : 0.f;
      }
    }
    if (BMODE == B_T || BMODE == B_IM2COL) {  // k (pixel) fastest
      const int n = n0 + r4;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = k0 + k4 + j;
        rb[j] = (n < N && k < k_end) ? load_b(BMODE, Bm, k, n, ldb, g) : 0.f;
      }
    } else {  // n fastest
      const int k = k0 + k16;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = n0 + c16 + j;
        rb[j] = (n < N && k < k_end) ? load_b(BMODE, Bm, k, n, ldb, g) : 0.f;
      }
    }
  };
  fetch(k_begin);
  for (int k0 = k_begin; k0 < k_end; k0 += BK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (TA == 0) As[k4 + j][r4] = ra[j];
      else As[k16][c16 + j] = ra[j];
      if (BMODE == B_T || BMODE == B_IM2COL) Bs[k4 + j][r4] = rb[j];
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
      if (n < N) C[(long long)m * ldo + n] = acc[i][j];
    }
  }
}

// C[b] = beta * C[b] + sum_{z in group b} part[z]   (fixed order)
__global__ void gemm_ex_reduce_kernel(const float* __restrict__ part, float* __restrict__ C, int M, int N, int ldc,
                                      long long strideC, int groups, int per_group, float beta) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long mn = (long long)M * N;
  if (i >= mn * groups) return;
  const int grp = (int)(i / mn);
  const long long r = i - (long long)grp * mn;
  const int m = (int)(r / N), n = (int)(r - (long long)m * N);
  float v = 0.f;
  for (int z = 0; z < per_group; ++z) v += part[((long long)grp * per_group + z) * mn + r];
  float* c = C + grp * strideC + (long long)m * ldc + n;
  *c = beta != 0.f ? beta * (*c) + v : v;
}

// C[m][n] = sum_z part[z][m][n] + bias[n]   (fixed order)
__global__ void linear_splitk_reduce_kernel(const float* __restrict__ part, const float* __restrict__ bias, float* __restrict__ C,
                                            int M, int N, int ldc, int splits) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long mn = (long long)M * N;
  if (i >= mn) return;
  const int m = (int)(i / N), n = (int)(i - (long long)m * N);
  float v = bias ? bias[n] : 0.f;
  for (int z = 0; z < splits; ++z) v += part[(long long)z * mn + i];
  C[(long long)m * ldc + n] = v;
}

template <int TA>
int launch_ex(int opB, dim3 grid, void* stream, const float* A, const float* Bm, float* C, int M, int N, int K, int lda, int ldb,
              int ldc, long long sA, long long sB, long long sC, int splits, int kps, int to_part, Gather g) {
  cudaStream_t st = ocv_stream(stream);
  switch (opB) {
    case B_N: sgemm_ex_kernel<TA, B_N><<<grid, NT, 0, st>>>(A, Bm, C, M, N, K, lda, ldb, ldc, sA, sB, sC, splits, kps, to_part, g); break;
    case B_T: sgemm_ex_kernel<TA, B_T><<<grid, NT, 0, st>>>(A, Bm, C, M, N, K, lda, ldb, ldc, sA, sB, sC, splits, kps, to_part, g); break;
    case B_PATCH: sgemm_ex_kernel<TA, B_PATCH><<<grid, NT, 0, st>>>(A, Bm, C, M, N, K, lda, ldb, ldc, sA, sB, sC, splits, kps, to_part, g); break;
    default: sgemm_ex_kernel<TA, B_IM2COL><<<grid, NT, 0, st>>>(A, Bm, C, M, N, K, lda, ldb, ldc, sA, sB, sC, splits, kps, to_part, g); break;
  }
  return ocv_launch_status();
}

}  // namespace

extern "C" long long ocv_gemm_f32_ex_workspace(int M, int N, int batch, int split_k) {
  if (M <= 0 || N <= 0 || batch <= 0) return 0;
  if (split_k < 1) split_k = 1;
  return (long long)batch * split_k * M * N * (long long)sizeof(float);
}

extern "C" int ocv_gemm_f32_ex(int transA, int opB, const float* A, const float* Bm, float* C, int M, int N, int K, int lda, int ldb,
                               int ldc, float beta, int batch, long long strideA, long long strideB, long long strideC,
                               int reduce_batch, int split_k, const int* gather_geom /* host: C,h,w,P (opB >= 2) */,
                               void* workspace, void* stream) {
  OCV_CHECK_ARG(A && Bm && C && M > 0 && N > 0 && K > 0 && batch > 0);
  OCV_CHECK_ARG(opB >= 0 && opB <= 3 && (opB < 2 || gather_geom));
  if (split_k < 1) split_k = 1;
  OCV_CHECK_SHAPE((long long)batch * split_k <= 65535);
  const bool partials = split_k > 1 || reduce_batch || beta != 0.f;
  OCV_CHECK_ARG(!partials || workspace);
  if (int s = ocv_require_sm100()) return s;
  Gather g{};
  if (opB >= 2) {
    g.C = gather_geom[0]; g.h = gather_geom[1]; g.w = gather_geom[2]; g.P = gather_geom[3];
    OCV_CHECK_SHAPE(g.C > 0 && g.h > 0 && g.w > 0 && (opB == B_IM2COL || g.P > 0));
    if (opB == B_PATCH) { g.gw = g.w / g.P; g.S = (g.h / g.P) * g.gw; }
  }
  const int kps = ceil_div(ceil_div(K, split_k), BK) * BK;
  const int splits = ceil_div(K, kps);
  dim3 grid(ceil_div(N, BN), ceil_div(M, BM), batch * splits);
  float* dst = partials ? static_cast<float*>(workspace) : C;
  int s = transA ? launch_ex<1>(opB, grid, stream, A, Bm, dst, M, N, K, lda, ldb, ldc, strideA, strideB, strideC, splits, kps,
                                partials ? 1 : 0, g)
                 : launch_ex<0>(opB, grid, stream, A, Bm, dst, M, N, K, lda, ldb, ldc, strideA, strideB, strideC, splits, kps,
                                partials ? 1 : 0, g);
  if (s) return s;
  if (!partials) return OCV_OK;
  const int groups = reduce_batch ? 1 : batch;
  const int per_group = reduce_batch ? batch * splits : splits;
  const long long n = (long long)M * N * groups;
  gemm_ex_reduce_kernel<<<(unsigned)ceil_div<long long>(n, 256), 256, 0, ocv_stream(stream)>>>(
      static_cast<const float*>(workspace), C, M, N, ldc, strideC, groups, per_group, beta);
  return ocv_launch_status();
}

// nn.Linear forward for problems with a long K and few output tiles (linear2 of the encoder layers: K = 1024, N = 128):
// split-K partials + fixed-order reduction with the bias, so that 8-56 tiles become 100+ CTAs.  workspace: split_k*M*N floats.
extern "C" int ocv_linear_splitk_f32(const float* A, const float* W, const float* bias, float* C, int M, int N, int K, int lda,
                                     int ldb, int ldc, int split_k, void* workspace, void* stream) {
  OCV_CHECK_ARG(A && W && C && workspace && M > 0 && N > 0 && K > 0 && split_k >= 1 && split_k <= 65535);
  if (int s = ocv_require_sm100()) return s;
  const int kps = ceil_div(ceil_div(K, split_k), BK) * BK;
  const int splits = ceil_div(K, kps);
  dim3 grid(ceil_div(N, BN), ceil_div(M, BM), splits);
  Gather g{};
  sgemm_ex_kernel<0, B_T><<<grid, NT, 0, ocv_stream(stream)>>>(A, W, static_cast<float*>(workspace), M, N, K, lda, ldb, ldc, 0, 0, 0,
                                                              splits, kps, 1, g);
  if (int s = ocv_launch_status()) return s;
  const long long n = (long long)M * N;
  linear_splitk_reduce_kernel<<<(unsigned)ceil_div<long long>(n, 256), 256, 0, ocv_stream(stream)>>>(
      static_cast<const float*>(workspace), bias, C, M, N, ldc, splits);
  return ocv_launch_status();
}
