// Small token-path kernels shared by the fp32 and bf16 paths: fused MLP chains (positional-encoder MLP, bin-width
// regressor) and the fp32 -> bf16 cast used for operand copies.  Contracts: include/objcavit_b200.h
#include "common.cuh"

namespace {

constexpr int MLP_MAXW = 512;    // widest layer
constexpr int MLP_MAXL = 8;
constexpr int MLP_THREADS = 512;

struct MlpParams {
  const float* W[MLP_MAXL];
  const float* b[MLP_MAXL];
  int dims[MLP_MAXL + 1];
  int n_layers;
  const float* x;
  long long x_ld;
  const float* residual;  // [R, dims[n]] added to the last layer's output (may be NULL)
  float* out32;
  __nv_bfloat16* out16;
  int R;
};

// One CTA = ROWS rows through the whole chain; activations ping-pong in shared memory.  A layer (K -> N) is spread over
// the 512 threads as (output neuron n) x (K split): thread (n, part) reads its slice of row n of the nn.Linear weight
// with independent, unrolled 128-bit loads (the chain is latency-bound, not FLOP-bound: every load in flight counts),
// keeps ROWS accumulators, and the K-split partials are reduced through shared memory in a fixed order.
// LeakyReLU(0.01) between layers, none after the last.
template <int ROWS>
__global__ void __launch_bounds__(MLP_THREADS)
mlp_chain_kernel(MlpParams p) {
  __shared__ __align__(16) float buf[2][ROWS][MLP_MAXW];
  __shared__ float red[MLP_THREADS][ROWS];
  pdl_launch();
  pdl_wait();  // x (and the residual) may come from the previous kernel of the chain
  const int r0 = blockIdx.x * ROWS;
  const int in0 = p.dims[0];
  for (int i = threadIdx.x; i < ROWS * in0; i += MLP_THREADS) {
    const int r = i / in0, k = i - r * in0;
    buf[0][r][k] = (r0 + r < p.R) ? p.x[(long long)(r0 + r) * p.x_ld + k] : 0.f;
  }
  __syncthreads();
  int cur = 0;
  for (int l = 0; l < p.n_layers; ++l) {
    const int K = p.dims[l], N = p.dims[l + 1];
    const bool last = l == p.n_layers - 1;
    const float* __restrict__ W = p.W[l];
    const float* __restrict__ bias = p.b[l];
    for (int n0 = 0; n0 < N; n0 += MLP_THREADS) {
      const int Nc = min(N - n0, MLP_THREADS);
      // K split: power of two, every part a multiple of 4 wide (float4 path) and at least 8 wide
      int ks = 1;
      while (ks * 2 * Nc <= MLP_THREADS && (K % (ks * 2 * 4)) == 0 && K / (ks * 2) >= 8) ks *= 2;
      const int part = threadIdx.x / Nc, nl = threadIdx.x - part * Nc;
      const bool active = part < ks;
      float acc[ROWS];
#pragma unroll
      for (int r = 0; r < ROWS; ++r) acc[r] = 0.f;
      if (active) {
        const int kw = K / ks, k0 = part * kw;
        const float* wrow = W + (long long)(n0 + nl) * K + k0;
        if ((kw & 3) == 0 && (K & 3) == 0) {
          const float4* w4 = reinterpret_cast<const float4*>(wrow);
#pragma unroll 8
          for (int k4 = 0; k4 < kw / 4; ++k4) {
            const float4 w = __ldg(w4 + k4);
#pragma unroll
            for (int r = 0; r < ROWS; ++r) {
              const float4 a = *reinterpret_cast<const float4*>(&buf[cur][r][k0 + 4 * k4]);
              acc[r] = fmaf(w.x, a.x, acc[r]);
              acc[r] = fmaf(w.y, a.y, acc[r]);
              acc[r] = fmaf(w.z, a.z, acc[r]);
              acc[r] = fmaf(w.w, a.w, acc[r]);
            }
          }
        } else {
          for (int k = 0; k < kw; ++k) {
            const float w = __ldg(wrow + k);
#pragma unroll
            for (int r = 0; r < ROWS; ++r) acc[r] = fmaf(w, buf[cur][r][k0 + k], acc[r]);
          }
        }
      }
      if (ks > 1) {
        if (active) {
#pragma unroll
          for (int r = 0; r < ROWS; ++r) red[threadIdx.x][r] = acc[r];
        }
        __syncthreads();
        if (part == 0) {
          for (int q = 1; q < ks; ++q) {
#pragma unroll
            for (int r = 0; r < ROWS; ++r) acc[r] += red[q * Nc + nl][r];
          }
        }
      }
      if (part == 0) {
        const int n = n0 + nl;
        const float bv = bias ? __ldg(bias + n) : 0.f;
#pragma unroll
        for (int r = 0; r < ROWS; ++r) {
          float v = acc[r] + bv;
          if (!last) {
            buf[cur ^ 1][r][n] = v > 0.f ? v : 0.01f * v;
          } else if (r0 + r < p.R) {
            const long long o = (long long)(r0 + r) * N + n;
            if (p.residual) v += p.residual[o];
            if (p.out32) p.out32[o] = v;
            if (p.out16) p.out16[o] = __float2bfloat16(v);
          }
        }
      }
      if (ks > 1) __syncthreads();  // red[] is reused by the next chunk / layer
    }
    __syncthreads();
    cur ^= 1;
  }
}

// One Linear (+ LeakyReLU) layer for FEW rows (the bin-width regressor: one row per image).  The chain kernel above makes
// every CTA stream all the weights; here the output neurons are spread over the grid instead (8 per CTA, one warp each, lanes
// split K), so each weight is read once chip-wide, and the layers are separate PDL-chained launches.  The warp's weight row
// is fetched before the grid dependency resolves.
constexpr int MLPL_ROWS = 16, MLPL_MAXK = 512;
__global__ void __launch_bounds__(256)
mlp_layer_kernel(const float* __restrict__ x, long long x_ld, const float* __restrict__ W, const float* __restrict__ bias,
                 float* __restrict__ out, int out_ld, int R, int K, int N, int leaky) {
  __shared__ __align__(16) float xs[MLPL_ROWS][MLPL_MAXK];
  pdl_launch();
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int n = blockIdx.x * 8 + warp;
  const int r0 = blockIdx.y * MLPL_ROWS;
  float4 w[MLPL_MAXK / 128];
#pragma unroll
  for (int j = 0; j < MLPL_MAXK / 128; ++j) {
    const int k = (lane + 32 * j) * 4;
    w[j] = (n < N && k < K) ? __ldg(reinterpret_cast<const float4*>(W + (long long)n * K + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const float bv = (bias && n < N) ? __ldg(bias + n) : 0.f;
  pdl_wait();
  for (int i = threadIdx.x; i < MLPL_ROWS * (K / 4); i += 256) {
    const int r = i / (K / 4), k = (i - r * (K / 4)) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r0 + r < R) v = *reinterpret_cast<const float4*>(x + (long long)(r0 + r) * x_ld + k);
    *reinterpret_cast<float4*>(&xs[r][k]) = v;
  }
  __syncthreads();
  float acc[MLPL_ROWS];
#pragma unroll
  for (int r = 0; r < MLPL_ROWS; ++r) acc[r] = 0.f;
#pragma unroll
  for (int j = 0; j < MLPL_MAXK / 128; ++j) {
    const int k = (lane + 32 * j) * 4;
    if (k < K) {
#pragma unroll
      for (int r = 0; r < MLPL_ROWS; ++r) {
        const float4 a = *reinterpret_cast<const float4*>(&xs[r][k]);
        acc[r] = fmaf(w[j].x, a.x, acc[r]);
        acc[r] = fmaf(w[j].y, a.y, acc[r]);
        acc[r] = fmaf(w[j].z, a.z, acc[r]);
        acc[r] = fmaf(w[j].w, a.w, acc[r]);
      }
    }
  }
  float mine = 0.f;
#pragma unroll
  for (int r = 0; r < MLPL_ROWS; ++r) {
    const float v = warp_sum(acc[r]);
    if (lane == r) mine = v;
  }
  if (lane < MLPL_ROWS && r0 + lane < R && n < N) {
    float v = mine + bv;
    if (leaky) v = v > 0.f ? v : 0.01f * v;
    out[(long long)(r0 + lane) * out_ld + n] = v;
  }
}

__global__ void cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  pdl_launch();
  pdl_wait();
  const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    const float4 v = *reinterpret_cast<const float4*>(src + i);
    __nv_bfloat162 h0 = __floats2bfloat162_rn(v.x, v.y), h1 = __floats2bfloat162_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&h0);
    u.y = *reinterpret_cast<uint32_t*>(&h1);
    *reinterpret_cast<uint2*>(dst + i) = u;
  } else {
    for (long long j = i; j < n; ++j) dst[j] = __float2bfloat16(src[j]);
  }
}

}  // namespace

extern "C" int ocv_mlp_f32(const float* x, long long x_ld, int R, int n_layers, const float* const* weights,
                           const float* const* biases, const int* dims, const float* residual, float* out32,
                           void* out16, void* stream) {
  OCV_CHECK_ARG(x && weights && biases && dims && (out32 || out16) && R > 0);
  OCV_CHECK_SHAPE(n_layers >= 1 && n_layers <= MLP_MAXL);
  MlpParams p{};
  for (int l = 0; l <= n_layers; ++l) {
    OCV_CHECK_SHAPE(dims[l] > 0 && dims[l] <= MLP_MAXW);
    p.dims[l] = dims[l];
  }
  for (int l = 0; l < n_layers; ++l) {
    OCV_CHECK_ARG(weights[l] != nullptr);
    p.W[l] = weights[l];
    p.b[l] = biases[l];
  }
  if (int s = ocv_require_sm100()) return s;
  p.n_layers = n_layers; p.x = x; p.x_ld = x_ld; p.residual = residual; p.out32 = out32;
  p.out16 = static_cast<__nv_bfloat16*>(out16); p.R = R;
  // few rows (the regressor: one row per image): one row per CTA so that every image gets its own SM
  if (R <= 64) return ocv_launch_pdl(mlp_chain_kernel<1>, dim3(R), dim3(MLP_THREADS), 0, stream, p);
  return ocv_launch_pdl(mlp_chain_kernel<4>, dim3(ceil_div(R, 4)), dim3(MLP_THREADS), 0, stream, p);
}

extern "C" int ocv_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
  OCV_CHECK_ARG(src && dst && n > 0);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 7) == 0);
  if (int s = ocv_require_sm100()) return s;
  return ocv_launch_pdl(cast_bf16_kernel, dim3((unsigned)ceil_div<long long>(ceil_div<long long>(n, 4), 256)), dim3(256), 0,
                        stream, src, static_cast<__nv_bfloat16*>(dst), n);
}

// one Linear (+ LeakyReLU 0.01) layer for few rows: the training step keeps every layer's output for its backward
extern "C" int ocv_linear_rows_f32(const float* x, long long x_ld, int R, const float* W, const float* bias, int K, int N, int leaky,
                                   float* out, int out_ld, void* stream) {
  OCV_CHECK_ARG(x && W && out && R > 0 && N > 0);
  OCV_CHECK_SHAPE(K > 0 && K <= MLPL_MAXK && K % 4 == 0 && out_ld >= N);
  OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(W)) & 15) == 0 && x_ld % 4 == 0);
  if (int s = ocv_require_sm100()) return s;
  return ocv_launch_pdl(mlp_layer_kernel, dim3(ceil_div(N, 8), ceil_div(R, MLPL_ROWS)), dim3(256), 0, stream, x, x_ld, W, bias, out,
                        out_ld, R, K, N, leaky ? 1 : 0);
}

extern "C" long long ocv_mlp_rows_f32_workspace(int R, int n_layers, const int* dims) {
  if (R <= 0 || n_layers < 1 || !dims) return 0;
  int mx = 0;
  for (int l = 1; l < n_layers; ++l) mx = dims[l] > mx ? dims[l] : mx;
  return 2ll * R * mx * (long long)sizeof(float);
}

extern "C" int ocv_mlp_rows_f32(const float* x, long long x_ld, int R, int n_layers, const float* const* weights,
                                const float* const* biases, const int* dims, float* out32, void* workspace, void* stream) {
  OCV_CHECK_ARG(x && weights && biases && dims && out32 && R > 0 && (n_layers == 1 || workspace));
  OCV_CHECK_SHAPE(n_layers >= 1 && n_layers <= MLP_MAXL);
  for (int l = 0; l < n_layers; ++l) {
    OCV_CHECK_SHAPE(dims[l] > 0 && dims[l] <= MLPL_MAXK && dims[l] % 4 == 0 && dims[l + 1] > 0);
    OCV_CHECK_ARG(weights[l] != nullptr);
    OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(weights[l]) & 15) == 0);
  }
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(x) & 15) == 0 && x_ld % 4 == 0);
  if (int s = ocv_require_sm100()) return s;
  int mx = 0;
  for (int l = 1; l < n_layers; ++l) mx = dims[l] > mx ? dims[l] : mx;
  float* buf[2] = {static_cast<float*>(workspace), static_cast<float*>(workspace) + (size_t)R * mx};
  const float* in = x;
  long long in_ld = x_ld;
  for (int l = 0; l < n_layers; ++l) {
    const bool last = l == n_layers - 1;
    float* o = last ? out32 : buf[l & 1];
    const int N = dims[l + 1];
    if (int s = ocv_launch_pdl(mlp_layer_kernel, dim3(ceil_div(N, 8), ceil_div(R, MLPL_ROWS)), dim3(256), 0, stream, in, in_ld,
                               weights[l], biases[l], o, N, R, dims[l], N, last ? 0 : 1))
      return s;
    in = o;
    in_ld = N;
  }
  return OCV_OK;
}
