// Operand preparation for the tensor-core pixel path: feature-map layout change (NCHW -> NHWC bf16),
// conv3x3 weight packing, and the per-image fold of the 128 range-attention queries into the 1x1
// conv_out weights (logits = Wout @ (Q @ conv) + bout  ==  (Wout @ Q) @ conv + bout').
// Contracts: include/objcavit_b200.h
#include "common.cuh"

namespace {

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

// block (32,8): tile of 32 x-positions x C channels of one image row
template <typename T>
__global__ void nchw_to_nhwc_bf16_kernel(const T* __restrict__ src, __nv_bfloat16* __restrict__ dst, int C, int h,
                                         int w) {
  extern __shared__ float tile[];  // [32][C+1]
  const int b = blockIdx.z, y = blockIdx.y, x0 = blockIdx.x * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const T* s = src + (long long)b * C * h * w + (long long)y * w;
  for (int c = ty; c < C; c += 8) {
    const int x = x0 + tx;
    tile[tx * (C + 1) + c] = (x < w) ? to_f32<T>(s[(long long)c * h * w + x]) : 0.f;
  }
  __syncthreads();
  __nv_bfloat16* d = dst + (((long long)b * h + y) * w + x0) * C;
  const int tid = ty * 32 + tx;
  // C/2 bf16x2 words per pixel
  const int words = C / 2;
  for (int i = tid; i < 32 * words; i += 256) {
    const int px = i / words, cw = i % words;
    if (x0 + px < w) {
      const __nv_bfloat162 v = __floats2bfloat162_rn(tile[px * (C + 1) + 2 * cw], tile[px * (C + 1) + 2 * cw + 1]);
      reinterpret_cast<__nv_bfloat162*>(d + (long long)px * C)[cw] = v;
    }
  }
}

// W [128 o][128 c][3][3] fp32 -> W3p [9 tap][128 o][128 c] bf16
__global__ void pack_conv3x3_kernel(const float* __restrict__ W, __nv_bfloat16* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 9 * 128 * 128) return;
  const int tap = i / (128 * 128), o = (i / 128) % 128, c = i % 128;
  out[i] = __float2bfloat16(W[((long long)o * 128 + c) * 9 + tap]);
}

// Mb[b,k,c] = sum_q Wout[k,q] * Q[b,q,c];  blog[b,k] = bout[k] + sum_c Mb[b,k,c] * b3[c]
// block = FQ_BINS bins of one image, thread = channel c (E == blockDim.x).  The image's whole query block Q[b] (Q x E fp32)
// and the FQ_BINS rows of Wout are staged in shared memory with cp.async — every byte in flight at once: the stage is
// latency-bound (a 256x128x128 GEMM per image), so one memory round trip per CTA instead of one per query row.
constexpr int FQ_BINS = 32;
__global__ void fold_queries_kernel(const float* __restrict__ Wout, const float* __restrict__ bout,
                                    const float* __restrict__ queries, long long q_bs, int q_rs,
                                    const float* __restrict__ b3, float scale, __nv_bfloat16* __restrict__ Mb,
                                    float* __restrict__ blog, int K, int Q, int E) {
  extern __shared__ __align__(16) float fq_smem[];
  float* ws = fq_smem;                 // [FQ_BINS][Q]
  float* qs = ws + FQ_BINS * Q;        // [Q][E]
  float* red = qs + (size_t)Q * E;     // [FQ_BINS][nwarp]
  pdl_launch();
  const int b = blockIdx.y, k0 = blockIdx.x * FQ_BINS, c = threadIdx.x;
  auto cp16 = [](void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
  };
  for (int i = threadIdx.x; i < FQ_BINS * Q / 4; i += blockDim.x) {  // conv_out weights: independent of the chain
    const int kk = (i * 4) / Q, q = (i * 4) % Q;
    if (k0 + kk < K) cp16(ws + i * 4, Wout + (long long)(k0 + kk) * Q + q);
    else *reinterpret_cast<float4*>(ws + i * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  pdl_wait();  // the queries are the previous kernels' tokens
  const float* qb = queries + b * q_bs;
  for (int i = threadIdx.x; i < Q * E / 4; i += blockDim.x) {
    const int q = (i * 4) / E, e = (i * 4) % E;
    cp16(qs + i * 4, qb + (long long)q * q_rs + e);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  float acc[FQ_BINS];
#pragma unroll
  for (int i = 0; i < FQ_BINS; ++i) acc[i] = 0.f;
  for (int q = 0; q < Q; q += 4) {
    const float q0 = qs[(q + 0) * E + c], q1 = qs[(q + 1) * E + c], q2 = qs[(q + 2) * E + c], q3 = qs[(q + 3) * E + c];
#pragma unroll
    for (int i = 0; i < FQ_BINS; ++i) {
      const float4 w = *reinterpret_cast<const float4*>(ws + i * Q + q);  // broadcast
      acc[i] = fmaf(w.x, q0, acc[i]);
      acc[i] = fmaf(w.y, q1, acc[i]);
      acc[i] = fmaf(w.z, q2, acc[i]);
      acc[i] = fmaf(w.w, q3, acc[i]);
    }
  }
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32, nwarp = blockDim.x / 32;
  const float bc = b3 ? b3[c] : 0.f;
#pragma unroll
  for (int i = 0; i < FQ_BINS; ++i) {
    if (k0 + i < K) Mb[((long long)b * K + k0 + i) * E + c] = __float2bfloat16(acc[i] * scale);
    const float s = warp_sum(acc[i] * bc);
    if (lane == 0) red[i * nwarp + warp] = s;
  }
  __syncthreads();
  if (threadIdx.x < FQ_BINS && k0 + threadIdx.x < K) {
    float s = bout ? bout[k0 + threadIdx.x] : 0.f;
    for (int w = 0; w < nwarp; ++w) s += red[threadIdx.x * nwarp + w];
    blog[(long long)b * K + k0 + threadIdx.x] = s * scale;
  }
}

// fast path of the layout conversion: fp32 [B,128,P] -> bf16 [B,P,128] over tiles of 64 flat pixels (P % 64 == 0, 16-byte aligned):
// 8 independent 16-byte loads per thread (4 pixels of one channel), bf16 tile in shared memory (row pitch 272 B), 256 contiguous
// bytes per pixel out as 16-byte stores.
constexpr int CV_TP = 64, CV_PITCH = 136;   // bf16 elements per tile row (128 + 8: 16-byte aligned rows, conflict-light columns)
__global__ void __launch_bounds__(256)
nchw_to_nhwc_c128_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long P) {
  __shared__ __align__(16) __nv_bfloat16 tile[CV_TP * CV_PITCH];
  const int b = blockIdx.y;
  const long long p0 = (long long)blockIdx.x * CV_TP;
  const float* s = src + (long long)b * 128 * P + p0;
  const int t = threadIdx.x, q = t & 15, c0 = t >> 4;      // q: group of 4 pixels, c0 + 16 j: channel
  float4 v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = __ldcs(reinterpret_cast<const float4*>(s + (long long)(c0 + 16 * j) * P) + q);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c0 + 16 * j;
    tile[(4 * q + 0) * CV_PITCH + c] = __float2bfloat16(v[j].x);
    tile[(4 * q + 1) * CV_PITCH + c] = __float2bfloat16(v[j].y);
    tile[(4 * q + 2) * CV_PITCH + c] = __float2bfloat16(v[j].z);
    tile[(4 * q + 3) * CV_PITCH + c] = __float2bfloat16(v[j].w);
  }
  __syncthreads();
  uint4* d = reinterpret_cast<uint4*>(dst + ((long long)b * P + p0) * 128);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int idx = k * 256 + t, px = idx >> 4, part = idx & 15;
    d[idx] = *reinterpret_cast<const uint4*>(tile + px * CV_PITCH + part * 8);
  }
}

}  // namespace

extern "C" int ocv_nchw_to_nhwc_bf16(const void* src, void* dst_bf16, int B, int C, int h, int w, int src_is_bf16,
                                     void* stream) {
  OCV_CHECK_ARG(src && dst_bf16 && B > 0 && h > 0 && w > 0);
  OCV_CHECK_SHAPE(C > 0 && C % 2 == 0 && C <= 256 && B <= 65535 && h <= 65535);
  if (int s = ocv_require_sm100()) return s;
  const long long Pn = (long long)h * w;
  if (!src_is_bf16 && C == 128 && Pn % CV_TP == 0 && Pn / CV_TP < (1ll << 31) &&
      ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst_bf16)) & 15) == 0) {
    nchw_to_nhwc_c128_kernel<<<dim3((unsigned)(Pn / CV_TP), B), 256, 0, ocv_stream(stream)>>>(
        static_cast<const float*>(src), static_cast<__nv_bfloat16*>(dst_bf16), Pn);
    return ocv_launch_status();
  }
  dim3 grid(ceil_div(w, 32), h, B), block(32, 8);
  const size_t smem = 32 * (C + 1) * sizeof(float);
  if (src_is_bf16)
    nchw_to_nhwc_bf16_kernel<__nv_bfloat16><<<grid, block, smem, ocv_stream(stream)>>>(
        static_cast<const __nv_bfloat16*>(src), static_cast<__nv_bfloat16*>(dst_bf16), C, h, w);
  else
    nchw_to_nhwc_bf16_kernel<float><<<grid, block, smem, ocv_stream(stream)>>>(
        static_cast<const float*>(src), static_cast<__nv_bfloat16*>(dst_bf16), C, h, w);
  return ocv_launch_status();
}

extern "C" int ocv_pack_conv3x3_bf16(const float* W, void* W3p, void* stream) {
  OCV_CHECK_ARG(W && W3p);
  if (int s = ocv_require_sm100()) return s;
  pack_conv3x3_kernel<<<ceil_div(9 * 128 * 128, 256), 256, 0, ocv_stream(stream)>>>(W, static_cast<__nv_bfloat16*>(W3p));
  return ocv_launch_status();
}

extern "C" int ocv_fold_queries_bf16(const float* Wout, const float* bout, const float* queries, long long q_bstride,
                                     int q_rstride, const float* b3, float logit_scale, void* Mb, float* blog, int B, int K,
                                     int Q, int E, void* stream) {
  OCV_CHECK_ARG(Wout && queries && Mb && blog && B > 0 && K > 0 && Q > 0);
  OCV_CHECK_SHAPE(E % 32 == 0 && E >= 32 && E <= 1024 && B <= 65535 && Q % 4 == 0);
  OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(Wout) | reinterpret_cast<uintptr_t>(queries)) & 15) == 0 &&
                  (q_bstride % 4) == 0 && (q_rstride % 4) == 0);
  if (int s = ocv_require_sm100()) return s;
  const size_t smem = ((size_t)FQ_BINS * Q + (size_t)Q * E + FQ_BINS * (E / 32)) * sizeof(float);
  OCV_CHECK_SHAPE(smem <= 200 * 1024);
  static size_t smem_set = 0;
  if (smem > 48 * 1024 && smem > smem_set) {
    cudaError_t e = cudaFuncSetAttribute(fold_queries_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    smem_set = smem;
  }
  return ocv_launch_pdl(fold_queries_kernel, dim3(ceil_div(K, FQ_BINS), B), dim3(E), smem, stream, Wout, bout, queries,
                        q_bstride, q_rstride, b3, logit_scale, static_cast<__nv_bfloat16*>(Mb), blog, K, Q, E);
}
