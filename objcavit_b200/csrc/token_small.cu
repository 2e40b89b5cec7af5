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
