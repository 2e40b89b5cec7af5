// Helper kernels of the tensor-core pixel path of the TRAINING step (config 5, `pixel="bf16"`): everything here is HBM-bound
// streaming work between the tcgen05 GEMMs (csrc/tc_gemm.cu) and the tcgen05 conv kernel (csrc/pixel_fused_bf16.cu, CONV_ONLY):
//   * tail forward / backward on channel-major logits [B,K,P] (what the GEMM  Mb[b] (K x C) @ x3[b]^T (C x P)  writes):
//       depth = sum_k softmax_k(l + bias) c_k;   dl = g p (c - depth) as bf16, channel-major [B,K,P] (the K-major operand of the
//       pixel-contraction GEMM dMb = dl x3^T);   per-(image, bin) sums of g p (= d loss / d centres) and of dl (= d loss / d bias)
//       through fixed-order block partials (deterministic).
//   * batched bf16 transpose [R,C] -> [C,R] (dl -> pixel-major for dx3^T = Mb^T dl;  x3 NHWC -> channel-major for dMb).
//   * im2col of the planar bf16 feature map for the 3x3 weight gradient: col[(c,ky,kx), b*P + p] = x[b,c,y+ky-1,x+kx-1], plus a
//       row of ones (row C*9) so that the same GEMM yields the bias gradient, plus 7 zero rows (GEMM N alignment).
// replaces (backward of): conv_out + Softmax + sum (modules/AdaBins.py:77,87; modules/GraphBins.py:109,119) and the weight
// gradient of self.conv3x3 (modules/miniViT.py:25; modules/ObjCAViT.py:374) that torch autograd computes in the reference.
// Algorithmic bytes: tail fwd 4K B/pixel read; tail bwd 4K read (second pass through L2) + 2K written; transpose 2+2 B/element;
// im2col 2 x 9C B/pixel written.   Contract: include/objcavit_b200.h.
#include "common.cuh"
#include <cuda_bf16.h>
#include <math.h>

namespace {

constexpr int TPB = 128;   // pixels per block (thread = pixel)

// depth[b,p] = sum_k softmax_k(l[b,k,p] + bias[k]) c[b,k]  — one pass, online softmax
__global__ void __launch_bounds__(TPB)
tail_logits_kernel(const float* __restrict__ logits, const float* __restrict__ bias, const float* __restrict__ centres,
                   float* __restrict__ depth, int K, long long P) {
  extern __shared__ float cs[];  // [K] centres, [K] bias
  float* bs = cs + K;
  const int b = blockIdx.y;
  for (int k = threadIdx.x; k < K; k += TPB) {
    cs[k] = centres[(long long)b * K + k];
    bs[k] = bias ? bias[k] : 0.f;
  }
  __syncthreads();
  const long long px = (long long)blockIdx.x * TPB + threadIdx.x;
  if (px >= P) return;
  const float* l = logits + (long long)b * K * P + px;
  float m = -INFINITY, s = 0.f, a = 0.f;
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    const float v = l[(long long)k * P] + bs[k];
    if (v > m) {
      const float r = __expf(m - v);
      s *= r;
      a *= r;
      m = v;
    }
    const float e = __expf(v - m);
    s += e;
    a = fmaf(e, cs[k], a);
  }
  depth[(long long)b * P + px] = a / s;
}

// dlT[b,k,p] (bf16) = g p (c - depth);   part[b][blk][0][k] = sum over the block's pixels of g p,  part[..][1][k] = of dl
__global__ void __launch_bounds__(TPB)
tail_logits_bwd_kernel(const float* __restrict__ logits, const float* __restrict__ bias, const float* __restrict__ centres,
                       const float* __restrict__ gdepth, __nv_bfloat16* __restrict__ dlT, float* __restrict__ part, int K,
                       long long P) {
  extern __shared__ float sm[];  // [K] centres, [K] bias, [TPB/32][2][K] per-warp sums
  float* cs = sm;
  float* bs = sm + K;
  float* ws = sm + 2 * K;
  const int b = blockIdx.y, warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  for (int k = threadIdx.x; k < K; k += TPB) {
    cs[k] = centres[(long long)b * K + k];
    bs[k] = bias ? bias[k] : 0.f;
  }
  __syncthreads();
  const long long px = (long long)blockIdx.x * TPB + threadIdx.x;
  const bool live = px < P;
  const float* l = logits + (long long)b * K * P + (live ? px : 0);
  float m = -INFINITY, s = 0.f, a = 0.f;
  if (live) {
#pragma unroll 4
    for (int k = 0; k < K; ++k) {
      const float v = l[(long long)k * P] + bs[k];
      if (v > m) {
        const float r = __expf(m - v);
        s *= r;
        a *= r;
        m = v;
      }
      const float e = __expf(v - m);
      s += e;
      a = fmaf(e, cs[k], a);
    }
  }
  const float inv = live ? 1.f / s : 0.f, d = a * inv, gd = live ? gdepth[(long long)b * P + px] : 0.f;
  __nv_bfloat16* o = dlT + (long long)b * K * P + (live ? px : 0);
  float* wsum = ws + warp * 2 * K;
  for (int k = 0; k < K; ++k) {
    float pg = 0.f, dl = 0.f;
    if (live) {
      pg = gd * __expf(l[(long long)k * P] + bs[k] - m) * inv;
      dl = pg * (cs[k] - d);
      o[(long long)k * P] = __float2bfloat16(dl);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      pg += __shfl_xor_sync(0xffffffffu, pg, off);
      dl += __shfl_xor_sync(0xffffffffu, dl, off);
    }
    if (lane == 0) {
      wsum[k] = pg;
      wsum[K + k] = dl;
    }
  }
  __syncthreads();
  float* dst = part + ((long long)b * gridDim.x + blockIdx.x) * 2 * K;
  for (int i = threadIdx.x; i < 2 * K; i += TPB) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < TPB / 32; ++w) t += ws[w * 2 * K + i];
    dst[i] = t;
  }
}

// dcentres[b,k] = sum_blk part[b][blk][0][k];   dbias[k] = beta*dbias[k] + sum_b sum_blk part[b][blk][1][k]   (fixed order)
__global__ void tail_logits_bwd_final_kernel(const float* __restrict__ part, float* __restrict__ dcentres, float* __restrict__ dbias,
                                             int B, int nblk, int K, float beta) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  float tb = 0.f;
  for (int b = 0; b < B; ++b) {
    float tc = 0.f, td = 0.f;
    const float* p = part + (long long)b * nblk * 2 * K;
    for (int i = 0; i < nblk; ++i) {
      tc += p[(long long)i * 2 * K + k];
      td += p[(long long)i * 2 * K + K + k];
    }
    dcentres[(long long)b * K + k] = tc;
    tb += td;
  }
  if (dbias) dbias[k] = (beta != 0.f ? beta * dbias[k] : 0.f) + tb;
}

// dst[z][c][r] = src[z][r][c]   (bf16; R % 64 == 0, C % 64 == 0; 64 x 64 tiles, 32-bit accesses on both sides)
__global__ void __launch_bounds__(256)
transpose_bf16_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, long long src_bstride, long long src_ld,
                      long long dst_bstride, long long dst_ld) {
  __shared__ unsigned short tile[64][66];
  const int tx = threadIdx.x, ty = threadIdx.y;  // (32, 8)
  const long long r0 = (long long)blockIdx.y * 64, c0 = (long long)blockIdx.x * 64;
  const uint32_t* s = src + ((long long)blockIdx.z * src_bstride + r0 * src_ld + c0) / 2;
#pragma unroll
  for (int r = ty; r < 64; r += 8) {
    const uint32_t v = s[(r * src_ld) / 2 + tx];
    tile[r][2 * tx] = (unsigned short)(v & 0xffffu);
    tile[r][2 * tx + 1] = (unsigned short)(v >> 16);
  }
  __syncthreads();
  uint32_t* d = dst + ((long long)blockIdx.z * dst_bstride + c0 * dst_ld + r0) / 2;
#pragma unroll
  for (int c = ty; c < 64; c += 8)
    d[(c * dst_ld) / 2 + tx] = (uint32_t)tile[2 * tx][c] | ((uint32_t)tile[2 * tx + 1][c] << 16);
}

// col[row, b*P + p], row = c*9 + ky*3 + kx < C*9: x[b,c,y+ky-1,x+kx-1] (0 outside);  row C*9: 1;  rows above: 0.   8 pixels per thread.
__global__ void __launch_bounds__(256)
im2col3x3_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, int B, int C, int h, int w) {
  const int row = blockIdx.y;
  const long long P = (long long)h * w, total = P * B;
  const long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (g >= total) return;
  __align__(16) __nv_bfloat16 v[8];
  if (row >= C * 9) {
    const __nv_bfloat16 f = __float2bfloat16(row == C * 9 ? 1.f : 0.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = f;
  } else {
    const int c = row / 9, t = row - c * 9, dy = t / 3 - 1, dx = t - (t / 3) * 3 - 1;
    const int b = (int)(g / P);                      // P % 8 == 0: the 8 pixels of a thread belong to one image
    const long long p = g - (long long)b * P;
    int y = (int)(p / w), xx = (int)(p - (long long)y * w);
    const __nv_bfloat16* plane = x + ((long long)b * C + c) * P;
    const __nv_bfloat16 z = __float2bfloat16(0.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int sy = y + dy, sx = xx + dx;
      v[i] = (sy >= 0 && sy < h && sx >= 0 && sx < w) ? plane[(long long)sy * w + sx] : z;
      if (++xx == w) { xx = 0; ++y; }
    }
  }
  *reinterpret_cast<uint4*>(col + (long long)row * total + g) = *reinterpret_cast<const uint4*>(v);
}

}  // namespace

extern "C" int ocv_tail_logits_f32(const float* logits, const float* logit_bias, const float* centres, float* depth, int B, int K,
                                   long long P, void* stream) {
  OCV_CHECK_ARG(logits && centres && depth && B > 0 && P > 0);
  OCV_CHECK_SHAPE(K > 0 && K <= 4096 && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  tail_logits_kernel<<<dim3((unsigned)ceil_div<long long>(P, TPB), B), TPB, 2 * K * sizeof(float), ocv_stream(stream)>>>(
      logits, logit_bias, centres, depth, K, P);
  return ocv_launch_status();
}

extern "C" long long ocv_tail_logits_bwd_workspace(int B, int K, long long P) {
  if (B <= 0 || K <= 0 || P <= 0) return 0;
  return (long long)B * ceil_div<long long>(P, TPB) * 2 * K * (long long)sizeof(float);
}

extern "C" int ocv_tail_logits_bwd_bf16(const float* logits, const float* logit_bias, const float* centres, const float* grad_depth,
                                        void* dlT_bf16, float* grad_centres, float* grad_bias, float bias_beta, int B, int K,
                                        long long P, void* workspace, void* stream) {
  OCV_CHECK_ARG(logits && centres && grad_depth && dlT_bf16 && grad_centres && workspace && B > 0 && P > 0);
  OCV_CHECK_SHAPE(K > 0 && K <= 2048 && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  const int nblk = (int)ceil_div<long long>(P, TPB);
  const size_t smem = (size_t)(2 * K + (TPB / 32) * 2 * K) * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(tail_logits_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  float* part = static_cast<float*>(workspace);
  tail_logits_bwd_kernel<<<dim3(nblk, B), TPB, smem, ocv_stream(stream)>>>(logits, logit_bias, centres, grad_depth,
                                                                         static_cast<__nv_bfloat16*>(dlT_bf16), part, K, P);
  if (int s = ocv_launch_status()) return s;
  tail_logits_bwd_final_kernel<<<ceil_div(K, 128), 128, 0, ocv_stream(stream)>>>(part, grad_centres, grad_bias, B, nblk, K, bias_beta);
  return ocv_launch_status();
}

extern "C" int ocv_transpose_bf16(const void* src, void* dst, int batch, int R, int C, long long src_bstride, long long src_ld,
                                  long long dst_bstride, long long dst_ld, void* stream) {
  OCV_CHECK_ARG(src && dst && batch > 0 && R > 0 && C > 0);
  OCV_CHECK_SHAPE(R % 64 == 0 && C % 64 == 0 && batch <= 65535 && R / 64 <= 65535 && src_ld >= C && dst_ld >= R);
  OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 3) == 0 && src_ld % 2 == 0 &&
                  dst_ld % 2 == 0 && src_bstride % 2 == 0 && dst_bstride % 2 == 0);
  if (int s = ocv_require_sm100()) return s;
  transpose_bf16_kernel<<<dim3(C / 64, R / 64, batch), dim3(32, 8), 0, ocv_stream(stream)>>>(
      static_cast<const uint32_t*>(src), static_cast<uint32_t*>(dst), src_bstride, src_ld, dst_bstride, dst_ld);
  return ocv_launch_status();
}

extern "C" int ocv_im2col3x3_bf16(const void* x_planar_bf16, void* col_bf16, int B, int C, int h, int w, void* stream) {
  OCV_CHECK_ARG(x_planar_bf16 && col_bf16 && B > 0 && C > 0 && h > 0 && w > 0);
  OCV_CHECK_SHAPE(((long long)h * w) % 8 == 0 && C * 9 + 8 <= 65535);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(col_bf16) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  const long long groups = (long long)B * h * w / 8;
  im2col3x3_bf16_kernel<<<dim3((unsigned)ceil_div<long long>(groups, 256), C * 9 + 8), 256, 0, ocv_stream(stream)>>>(
      static_cast<const __nv_bfloat16*>(x_planar_bf16), static_cast<__nv_bfloat16*>(col_bf16), B, C, h, w);
  return ocv_launch_status();
}
