// Small token-path kernels shared by the fp32 and bf16 paths: fused MLP chains (positional-encoder MLP, bin-width
// regressor) and the fp32 -> bf16 cast used for operand copies.  Contracts: include/objcavit_b200.h
#include "common.cuh"

namespace {

constexpr int MLP_ROWS = 8;      // rows per CTA
constexpr int MLP_MAXW = 512;    // widest layer
constexpr int MLP_MAXL = 8;

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

// One CTA = MLP_ROWS rows through the whole chain; activations ping-pong in shared memory; one warp per output
// neuron (lanes split K, weights read coalesced), LeakyReLU(0.01) between layers, none after the last.
__global__ void __launch_bounds__(256)
mlp_chain_kernel(MlpParams p) {
  __shared__ float buf[2][MLP_ROWS][MLP_MAXW];
  const int r0 = blockIdx.x * MLP_ROWS;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32, nwarp = blockDim.x / 32;
  const int in0 = p.dims[0];
  for (int i = threadIdx.x; i < MLP_ROWS * in0; i += blockDim.x) {
    const int r = i / in0, k = i - r * in0;
    buf[0][r][k] = (r0 + r < p.R) ? p.x[(long long)(r0 + r) * p.x_ld + k] : 0.f;
  }
  __syncthreads();
  int cur = 0;
  for (int l = 0; l < p.n_layers; ++l) {
    const int K = p.dims[l], N = p.dims[l + 1];
    const bool last = l == p.n_layers - 1;
    const float* W = p.W[l];
    const float* bias = p.b[l];
    for (int n = warp; n < N; n += nwarp) {
      float acc[MLP_ROWS];
#pragma unroll
      for (int r = 0; r < MLP_ROWS; ++r) acc[r] = 0.f;
      for (int k = lane; k < K; k += 32) {
        const float w = W[(long long)n * K + k];
#pragma unroll
        for (int r = 0; r < MLP_ROWS; ++r) acc[r] = fmaf(w, buf[cur][r][k], acc[r]);
      }
#pragma unroll
      for (int r = 0; r < MLP_ROWS; ++r) acc[r] = warp_sum(acc[r]);
      if (lane < MLP_ROWS) {
        float v = 0.f;
#pragma unroll
        for (int r = 0; r < MLP_ROWS; ++r) v = (lane == r) ? acc[r] : v;
        v += bias ? bias[n] : 0.f;
        if (!last) {
          v = v > 0.f ? v : 0.01f * v;
          buf[cur ^ 1][lane][n] = v;
        } else if (r0 + lane < p.R) {
          const long long o = (long long)(r0 + lane) * N + n;
          if (p.residual) v += p.residual[o];
          if (p.out32) p.out32[o] = v;
          if (p.out16) p.out16[o] = __float2bfloat16(v);
        }
      }
    }
    __syncthreads();
    cur ^= 1;
  }
}

__global__ void cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
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
  mlp_chain_kernel<<<ceil_div(R, MLP_ROWS), 256, 0, ocv_stream(stream)>>>(p);
  return ocv_launch_status();
}

extern "C" int ocv_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
  OCV_CHECK_ARG(src && dst && n > 0);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 7) == 0);
  if (int s = ocv_require_sm100()) return s;
  cast_bf16_kernel<<<(unsigned)ceil_div<long long>(ceil_div<long long>(n, 4), 256), 256, 0, ocv_stream(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst), n);
  return ocv_launch_status();
}
