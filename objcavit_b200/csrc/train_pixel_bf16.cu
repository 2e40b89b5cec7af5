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

// Online softmax statistics of one pixel's K logits (stride P apart): m = max, s = sum exp(l - m), a = sum exp(l - m) c.
// Four independent chains (k mod 4) keep four loads and four exp's in flight; merged in fixed order.  K % 4 == 0.
__device__ __forceinline__ void softmax_stats(const float* __restrict__ l, long long P, const float* __restrict__ bs,
                                              const float* __restrict__ cs, int K, float& m, float& s, float& a) {
  float mm[4], ss[4], aa[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) { mm[c] = -INFINITY; ss[c] = 0.f; aa[c] = 0.f; }
#pragma unroll 2
  for (int k = 0; k < K; k += 4) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float v = l[(long long)(k + c) * P] + bs[k + c];
      const float mn = fmaxf(mm[c], v);
      const float r = __expf(mm[c] - mn), e = __expf(v - mn);
      ss[c] = fmaf(ss[c], r, e);
      aa[c] = fmaf(aa[c], r, e * cs[k + c]);
      mm[c] = mn;
    }
  }
  m = fmaxf(fmaxf(mm[0], mm[1]), fmaxf(mm[2], mm[3]));
  s = 0.f;
  a = 0.f;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float r = __expf(mm[c] - m);
    s = fmaf(ss[c], r, s);
    a = fmaf(aa[c], r, a);
  }
}

// depth[b,p] = sum_k softmax_k(l[b,k,p] + bias[k]) c[b,k]  — one pass, online softmax
__global__ void __launch_bounds__(TPB)
tail_logits_kernel(const float* __restrict__ logits, const float* __restrict__ bias, const float* __restrict__ centres,
                   float* __restrict__ depth, int K, long long P, float* __restrict__ stats /* [B,3,P] (m, s, a) or NULL */) {
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
  float m, s, a;
  softmax_stats(l, P, bs, cs, K, m, s, a);
  depth[(long long)b * P + px] = a / s;
  if (stats) {                       // what the backward's first pass would recompute from the same logits: saved, bit for bit
    float* st = stats + (long long)b * 3 * P + px;
    st[0] = m;
    st[P] = s;
    st[2 * P] = a;
  }
}

// dlT[b,k,p] (bf16) = g p (c - depth);   part[b][blk][0][k] = sum over the block's pixels of g p,  part[..][1][k] = of dl.
// Two passes over the block's [K x 128] logits (online-softmax statistics on four independent chains; outputs): the loads of a
// pass are independent (deep memory-level parallelism).  The per-bin pixel sums are warp-reduced by a transposed
// butterfly, kept in registers (lane j owns bins j, j+32, ...), and the four warps are added in fixed order.  K % 32 == 0, K <= 512.
template <int KK>  // K / 32
__global__ void __launch_bounds__(TPB, 4)
tail_logits_bwd_kernel(const float* __restrict__ logits, const float* __restrict__ bias, const float* __restrict__ centres,
                       const float* __restrict__ gdepth, __nv_bfloat16* __restrict__ dlT, float* __restrict__ part, long long P,
                       const float* __restrict__ stats /* [B,3,P] from the forward, or NULL: recompute */) {
  constexpr int K = KK * 32;
  __shared__ float cs[K], bs[K], wacc[TPB / 32][2 * K];
  const int b = blockIdx.y, warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  for (int k = threadIdx.x; k < K; k += TPB) {
    cs[k] = centres[(long long)b * K + k];
    bs[k] = bias ? bias[k] : 0.f;
  }
  __syncthreads();
  const long long px = (long long)blockIdx.x * TPB + threadIdx.x;
  const bool live = px < P;
  const float* l = logits + (long long)b * K * P + (live ? px : 0);
  float m, s, a;
  if (stats) {                       // one pass over the logits instead of two
    const float* st = stats + (long long)b * 3 * P + (live ? px : 0);
    m = st[0];
    s = st[P];
    a = st[2 * P];
  } else {
    softmax_stats(l, P, bs, cs, K, m, s, a);
  }
  const float inv = 1.f / s, d = a * inv, gd = live ? gdepth[(long long)b * P + px] : 0.f;
  __nv_bfloat16* o = dlT + (long long)b * K * P + (live ? px : 0);
#pragma unroll 1
  for (int kk = 0; kk < KK; ++kk) {
    float pg[32], dl[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const int k = kk * 32 + j;
      pg[j] = gd * __expf(l[(long long)k * P] + bs[k] - m) * inv;
      dl[j] = pg[j] * (cs[k] - d);
      if (live) o[(long long)k * P] = __float2bfloat16(dl[j]);
    }
    // transposed butterfly: 31 shuffles per array leave the sum over the 32 pixels of bin kk*32 + lane in element 0 of lane
    // `lane` (a plain butterfly per bin would be 160); fixed order -> deterministic
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      const bool upper = (lane & off) != 0;
#pragma unroll
      for (int i = 0; i < off; ++i) {
        const float sp = upper ? pg[i] : pg[i + off], kp = upper ? pg[i + off] : pg[i];
        pg[i] = kp + __shfl_xor_sync(0xffffffffu, sp, off);
        const float sd = upper ? dl[i] : dl[i + off], kd = upper ? dl[i + off] : dl[i];
        dl[i] = kd + __shfl_xor_sync(0xffffffffu, sd, off);
      }
    }
    wacc[warp][kk * 32 + lane] = pg[0];
    wacc[warp][K + kk * 32 + lane] = dl[0];
  }
  __syncthreads();
  float* dst = part + ((long long)b * gridDim.x + blockIdx.x) * 2 * K;
  for (int i = threadIdx.x; i < 2 * K; i += TPB) {  // fixed order over the warps
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < TPB / 32; ++w) t += wacc[w][i];
    dst[i] = t;
  }
}

// dcentres[b,k] = sum_blk part[b][blk][0][k];  pb[b][k] = sum_blk part[b][blk][1][k].   Block = (32 bins, 8 slices of the block
// list), slices added in fixed order: deterministic, and parallel enough that the 2 K nblk B floats stream at L2 speed.
__global__ void __launch_bounds__(256)
tail_logits_bwd_final_kernel(const float* __restrict__ part, float* __restrict__ dcentres, float* __restrict__ pb, int nblk, int K) {
  __shared__ float sc[8][32], sd[8][32];
  const int b = blockIdx.y, k = blockIdx.x * 32 + threadIdx.x, sl = threadIdx.y;
  float tc = 0.f, td = 0.f;
  if (k < K) {
    const float* p = part + (long long)b * nblk * 2 * K + k;
#pragma unroll 4
    for (int i = sl; i < nblk; i += 8) {
      tc += p[(long long)i * 2 * K];
      td += p[(long long)i * 2 * K + K];
    }
  }
  sc[sl][threadIdx.x] = tc;
  sd[sl][threadIdx.x] = td;
  __syncthreads();
  if (sl == 0 && k < K) {
    float a = 0.f, c = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a += sc[i][threadIdx.x];
      c += sd[i][threadIdx.x];
    }
    dcentres[(long long)b * K + k] = a;
    pb[(long long)b * K + k] = c;
  }
}

// dbias[k] = beta*dbias[k] + sum_b pb[b][k]
__global__ void tail_logits_bwd_bias_kernel(const float* __restrict__ pb, float* __restrict__ dbias, int B, int K, float beta) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  float t = 0.f;
  for (int b = 0; b < B; ++b) t += pb[(long long)b * K + k];
  dbias[k] = (beta != 0.f ? beta * dbias[k] : 0.f) + t;
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
im2col3x3_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, int B, int C, int h, int w, int row0) {
  const int row = row0 + blockIdx.y;
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

// Patch unfold for the patch-embedding weight gradient: out[(c*Pp*Pp + py*Pp + px), b*S + sy*gw + sx] = x[b,c,sy*Pp+py,sx*Pp+px]
// (bf16; the rows are the weight's (c,py,px) order, the columns the token order): dW [E, C Pp Pp] = dtok^T [E, B S] @ out^T is then a
// K-major GEMM.  One block per (strip group, c, b): `R` patch rows of the plane are staged in shared memory (coalesced reads),
// then every (py,px) row gets its R*gw consecutive columns from one warp (coalesced writes).
__global__ void __launch_bounds__(256)
patch_unfold_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int C, int h, int w, int Pp, int R, long long ld) {
  extern __shared__ __nv_bfloat16 strip[];  // [R*Pp][w]
  const int b = blockIdx.z, c = blockIdx.y, sy0 = blockIdx.x * R;
  const int gh = h / Pp, gw = w / Pp, S = gh * gw;
  const int rows = min(R, gh - sy0);
  const float* src = x + (((long long)b * C + c) * h + (long long)sy0 * Pp) * w;
  const int n = rows * Pp * w;
  for (int i = threadIdx.x; i < n; i += 256) strip[i] = __float2bfloat16(src[i]);
  __syncthreads();
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int run = rows * gw;
  for (int pr = warp; pr < Pp * Pp; pr += 8) {
    const int py = pr / Pp, px = pr - py * Pp;
    __nv_bfloat16* dst = out + ((long long)c * Pp * Pp + pr) * ld + (long long)b * S + (long long)sy0 * gw;
    for (int e = lane; e < run; e += 32) {
      const int sy = e / gw, sx = e - sy * gw;
      dst[e] = strip[(sy * Pp + py) * w + sx * Pp + px];
    }
  }
}

// dst[c][r] (bf16, row pitch dst_ld) = src[r][c] (fp32, row pitch src_ld): any R, C (32 x 32 tiles with bounds checks)
__global__ void transpose_cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int R, int C, long long src_ld,
                                      long long dst_ld) {
  __shared__ float tile[32][33];
  const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  for (int rr = threadIdx.y; rr < 32; rr += 8)
    tile[rr][threadIdx.x] = (r0 + rr < R && c0 + threadIdx.x < C) ? src[(long long)(r0 + rr) * src_ld + c0 + threadIdx.x] : 0.f;
  __syncthreads();
  for (int cc = threadIdx.y; cc < 32; cc += 8)
    if (c0 + cc < C && r0 + threadIdx.x < R) dst[(long long)(c0 + cc) * dst_ld + r0 + threadIdx.x] = __float2bfloat16(tile[threadIdx.x][cc]);
}

// Vector path of the same (w % 8 == 0): a thread takes 8 consecutive pixels of one (c, ky) and writes the three kx rows from ONE
// aligned 16-byte load of the source row plus its two neighbours (3 loads and 3 stores instead of 24 + 3).
__global__ void __launch_bounds__(256)
im2col3x3_vec_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, int B, int C, int h, int w) {
  const int c = blockIdx.y / 3, ky = blockIdx.y - c * 3;
  const long long P = (long long)h * w, total = P * B;
  const long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (g >= total) return;
  const int b = (int)(g / P);
  const long long p = g - (long long)b * P;
  const int y = (int)(p / w), xx = (int)(p - (long long)y * w), sy = y + ky - 1;
  __align__(16) unsigned short v[10];  // source pixels xx-1 .. xx+8 of row sy (zeros outside)
#pragma unroll
  for (int i = 0; i < 10; ++i) v[i] = 0;
  if (sy >= 0 && sy < h) {
    const unsigned short* src = reinterpret_cast<const unsigned short*>(x) + ((long long)b * C + c) * P + (long long)sy * w + xx;
    const uint4 mid = *reinterpret_cast<const uint4*>(src);
    const unsigned short* mv = reinterpret_cast<const unsigned short*>(&mid);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[1 + i] = mv[i];
    if (xx > 0) v[0] = src[-1];
    if (xx + 8 < w) v[9] = src[8];
  }
  unsigned short* dst = reinterpret_cast<unsigned short*>(col) + ((long long)(c * 9 + ky * 3)) * total + g;
#pragma unroll
  for (int kx = 0; kx < 3; ++kx) {
    __align__(16) unsigned short o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = v[i + kx];
    *reinterpret_cast<uint4*>(dst + (long long)kx * total) = *reinterpret_cast<const uint4*>(o);
  }
}

}  // namespace

extern "C" int ocv_tail_logits_f32(const float* logits, const float* logit_bias, const float* centres, float* depth, int B, int K,
                                   long long P, void* stream) {
  OCV_CHECK_ARG(logits && centres && depth && B > 0 && P > 0);
  OCV_CHECK_SHAPE(K > 0 && K % 4 == 0 && K <= 4096 && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  tail_logits_kernel<<<dim3((unsigned)ceil_div<long long>(P, TPB), B), TPB, 2 * K * sizeof(float), ocv_stream(stream)>>>(
      logits, logit_bias, centres, depth, K, P, nullptr);
  return ocv_launch_status();
}

// Same, and the per-pixel softmax statistics (max, sum, centre-weighted sum) into stats [B,3,P] for ocv_tail_logits_bwd_stats_bf16.
extern "C" int ocv_tail_logits_stats_f32(const float* logits, const float* logit_bias, const float* centres, float* depth, float* stats,
                                         int B, int K, long long P, void* stream) {
  OCV_CHECK_ARG(logits && centres && depth && stats && B > 0 && P > 0);
  OCV_CHECK_SHAPE(K > 0 && K % 4 == 0 && K <= 4096 && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  tail_logits_kernel<<<dim3((unsigned)ceil_div<long long>(P, TPB), B), TPB, 2 * K * sizeof(float), ocv_stream(stream)>>>(
      logits, logit_bias, centres, depth, K, P, stats);
  return ocv_launch_status();
}

extern "C" long long ocv_tail_logits_bwd_workspace(int B, int K, long long P) {
  if (B <= 0 || K <= 0 || P <= 0) return 0;
  return ((long long)B * ceil_div<long long>(P, TPB) * 2 * K + (long long)B * K) * (long long)sizeof(float);
}

extern "C" int ocv_tail_logits_bwd_stats_bf16(const float* logits, const float* logit_bias, const float* centres, const float* stats,
                                              const float* grad_depth, void* dlT_bf16, float* grad_centres, float* grad_bias,
                                              float bias_beta, int B, int K, long long P, void* workspace, void* stream);

extern "C" int ocv_tail_logits_bwd_bf16(const float* logits, const float* logit_bias, const float* centres, const float* grad_depth,
                                        void* dlT_bf16, float* grad_centres, float* grad_bias, float bias_beta, int B, int K,
                                        long long P, void* workspace, void* stream) {
  return ocv_tail_logits_bwd_stats_bf16(logits, logit_bias, centres, nullptr, grad_depth, dlT_bf16, grad_centres, grad_bias, bias_beta, B,
                                        K, P, workspace, stream);
}

// `stats` = what ocv_tail_logits_stats_f32 saved in the forward (NULL: recomputed in a first pass over the logits)
extern "C" int ocv_tail_logits_bwd_stats_bf16(const float* logits, const float* logit_bias, const float* centres, const float* stats,
                                              const float* grad_depth, void* dlT_bf16, float* grad_centres, float* grad_bias,
                                              float bias_beta, int B, int K, long long P, void* workspace, void* stream) {
  OCV_CHECK_ARG(logits && centres && grad_depth && dlT_bf16 && grad_centres && workspace && B > 0 && P > 0);
  OCV_CHECK_SHAPE(K > 0 && K % 32 == 0 && K <= 512 && B <= 65535);
  if (int s = ocv_require_sm100()) return s;
  const int nblk = (int)ceil_div<long long>(P, TPB);
  float* part = static_cast<float*>(workspace);
  __nv_bfloat16* o = static_cast<__nv_bfloat16*>(dlT_bf16);
  const dim3 grid(nblk, B);
  cudaStream_t st = ocv_stream(stream);
  switch (K / 32) {
    case 2: tail_logits_bwd_kernel<2><<<grid, TPB, 0, st>>>(logits, logit_bias, centres, grad_depth, o, part, P, stats); break;
    case 4: tail_logits_bwd_kernel<4><<<grid, TPB, 0, st>>>(logits, logit_bias, centres, grad_depth, o, part, P, stats); break;
    case 8: tail_logits_bwd_kernel<8><<<grid, TPB, 0, st>>>(logits, logit_bias, centres, grad_depth, o, part, P, stats); break;
    case 16: tail_logits_bwd_kernel<16><<<grid, TPB, 0, st>>>(logits, logit_bias, centres, grad_depth, o, part, P, stats); break;
    default: return OCV_E_SHAPE;  // n_bins of the reference's params files: 256 (64 / 128 / 512 also built)
  }
  if (int s = ocv_launch_status()) return s;
  float* pb = part + (long long)B * nblk * 2 * K;
  tail_logits_bwd_final_kernel<<<dim3(K / 32, B), dim3(32, 8), 0, st>>>(part, grad_centres, pb, nblk, K);
  if (int s = ocv_launch_status()) return s;
  if (grad_bias) {
    tail_logits_bwd_bias_kernel<<<ceil_div(K, 128), 128, 0, st>>>(pb, grad_bias, B, K, bias_beta);
    return ocv_launch_status();
  }
  return OCV_OK;
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
  const unsigned gx = (unsigned)ceil_div<long long>(groups, 256);
  const __nv_bfloat16* xs = static_cast<const __nv_bfloat16*>(x_planar_bf16);
  __nv_bfloat16* cd = static_cast<__nv_bfloat16*>(col_bf16);
  if (w % 8 == 0 && (reinterpret_cast<uintptr_t>(x_planar_bf16) & 15) == 0) {
    im2col3x3_vec_kernel<<<dim3(gx, C * 3), 256, 0, ocv_stream(stream)>>>(xs, cd, B, C, h, w);
    if (int s = ocv_launch_status()) return s;
    im2col3x3_bf16_kernel<<<dim3(gx, 8), 256, 0, ocv_stream(stream)>>>(xs, cd, B, C, h, w, C * 9);   // ones row + zero rows
  } else {
    im2col3x3_bf16_kernel<<<dim3(gx, C * 9 + 8), 256, 0, ocv_stream(stream)>>>(xs, cd, B, C, h, w, 0);
  }
  return ocv_launch_status();
}

extern "C" int ocv_patch_unfold_bf16(const float* x, void* out_bf16, int B, int C, int h, int w, int P, long long ld, void* stream) {
  OCV_CHECK_ARG(x && out_bf16 && B > 0 && C > 0 && h > 0 && w > 0 && P > 0);
  OCV_CHECK_SHAPE(h % P == 0 && w % P == 0 && ld >= (long long)B * (h / P) * (w / P) && B <= 65535 && C <= 65535);
  if (int s = ocv_require_sm100()) return s;
  const int gh = h / P;
  const size_t strip = (size_t)P * w * sizeof(__nv_bfloat16);
  OCV_CHECK_SHAPE(strip <= 200 * 1024);
  int R = (int)((96 * 1024) / strip);
  R = R < 1 ? 1 : (R > gh ? gh : R);
  const size_t smem = strip * R;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(patch_unfold_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  patch_unfold_bf16_kernel<<<dim3(ceil_div(gh, R), C, B), 256, smem, ocv_stream(stream)>>>(x, static_cast<__nv_bfloat16*>(out_bf16), C, h, w,
                                                                                      P, R, ld);
  return ocv_launch_status();
}

// unfold from the NHWC bf16 copy of the features (what the tensor-core path already holds): half the bytes of the fp32 NCHW read,
// 64-byte channel chunks in, runs of (rows x gw) tokens out.  Block = (group of patch rows, py, 32-channel chunk, image): the pixel
// rows y = sy*P + py of the group go to shared memory as [token][px][16 channel-pair words] (+1 pad word per token: a warp reading
// one (px, channel pair) for 32 consecutive tokens hits 32 banks), then every warp writes token-contiguous rows of the output.
__global__ void __launch_bounds__(256)
patch_unfold_nhwc_kernel(const __nv_bfloat16* __restrict__ xh, __nv_bfloat16* __restrict__ out, int C, int h, int w, int Pp, int TG,
                         long long ld) {
  extern __shared__ uint32_t utile[];  // [tokens of the group][Pp * 16 + 1]
  const int gh = h / Pp, gw = w / Pp, S = gh * gw;
  const int sy0 = blockIdx.x * TG, py = blockIdx.y;
  const int chunks = C / 32;
  const int b = blockIdx.z / chunks, c0 = (blockIdx.z - b * chunks) * 32;
  const int rows = min(TG, gh - sy0);
  const int ntok = rows * gw;
  const int tstride = Pp * 16 + 1;
  // load: one 16-byte piece (8 channels) per thread and step
  const int npix = rows * gw * Pp;      // pixels of the group's rows that belong to whole patches
  const int n16 = npix * 4;
  for (int i0 = threadIdx.x; i0 < n16; i0 += 256 * 8) {    // 8 independent 16-byte loads in flight per thread
    uint4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * 256;
      if (i < n16) {
        const int pix = i >> 2, part = i & 3;
        const int r = pix / (gw * Pp), xx = pix - r * gw * Pp;
        v[u] = __ldcs(reinterpret_cast<const uint4*>(xh + (((long long)b * h + (long long)(sy0 + r) * Pp + py) * w + xx) * C + c0 + part * 8));
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * 256;
      if (i < n16) {
        const int pix = i >> 2, part = i & 3;
        const int r = pix / (gw * Pp), xx = pix - r * gw * Pp;
        const int sx = xx / Pp, px = xx - sx * Pp;
        uint32_t* d = utile + (r * gw + sx) * tstride + px * 16 + part * 4;
        d[0] = v[u].x; d[1] = v[u].y; d[2] = v[u].z; d[3] = v[u].w;
      }
    }
  }
  __syncthreads();
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  unsigned short* o16 = reinterpret_cast<unsigned short*>(out);
  const long long col0 = (long long)b * S + (long long)sy0 * gw;
  for (int item = warp; item < Pp * 16; item += 8) {      // (px, channel pair)
    const int px = item >> 4, cw = item & 15;
    const long long row_lo = ((long long)(c0 + 2 * cw) * Pp + py) * Pp + px;
    unsigned short* d_lo = o16 + row_lo * ld + col0;
    unsigned short* d_hi = d_lo + (long long)Pp * Pp * ld;
    for (int t = lane; t < ntok; t += 32) {
      const uint32_t v = utile[t * tstride + px * 16 + cw];
      d_lo[t] = (unsigned short)(v & 0xFFFFu);
      d_hi[t] = (unsigned short)(v >> 16);
    }
  }
}

extern "C" int ocv_patch_unfold_nhwc_bf16(const void* xh_bf16, void* out_bf16, int B, int C, int h, int w, int P, long long ld, void* stream) {
  OCV_CHECK_ARG(xh_bf16 && out_bf16 && B > 0 && C > 0 && h > 0 && w > 0 && P > 0);
  OCV_CHECK_SHAPE(h >= P && w >= P && C % 32 == 0 && ld >= (long long)B * (h / P) * (w / P) && P <= 65535 && (long long)B * (C / 32) <= 65535);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(xh_bf16) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  const int gh = h / P, gw = w / P;
  const size_t per_row = (size_t)gw * (P * 16 + 1) * sizeof(uint32_t);   // one row of patches
  OCV_CHECK_SHAPE(per_row <= 200 * 1024);
  int tg_max = (int)((200 * 1024) / per_row);
  tg_max = tg_max > gh ? gh : tg_max;
  const int groups = ceil_div(gh, tg_max);
  const int TG = ceil_div(gh, groups);
  const size_t smem = per_row * TG;
  static size_t smem_set = 0;
  if (smem > 48 * 1024 && smem > smem_set) {
    cudaError_t e = cudaFuncSetAttribute(patch_unfold_nhwc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    smem_set = smem;
  }
  patch_unfold_nhwc_kernel<<<dim3(groups, P, B * (C / 32)), 256, smem, ocv_stream(stream)>>>(
      static_cast<const __nv_bfloat16*>(xh_bf16), static_cast<__nv_bfloat16*>(out_bf16), C, h, w, P, TG, ld);
  return ocv_launch_status();
}

extern "C" int ocv_transpose_cast_f32_bf16(const float* src, void* dst_bf16, int R, int C, long long src_ld, long long dst_ld,
                                           void* stream) {
  OCV_CHECK_ARG(src && dst_bf16 && R > 0 && C > 0 && src_ld >= C && dst_ld >= R);
  OCV_CHECK_SHAPE(ceil_div(R, 32) <= 65535);
  if (int s = ocv_require_sm100()) return s;
  transpose_cast_kernel<<<dim3(ceil_div(C, 32), ceil_div(R, 32)), dim3(32, 8), 0, ocv_stream(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst_bf16), R, C, src_ld, dst_ld);
  return ocv_launch_status();
}
