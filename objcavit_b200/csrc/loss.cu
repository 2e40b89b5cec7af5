// Training losses of the head, forward and the gradients w.r.t. the head's two outputs:
//   SILog  (losses/SILogLoss.py:43-56):  bilinear up-sample of depth_pred to the GT size (align_corners=True), boolean mask,
//           g = log p - log t,  loss = alph * sqrt( sum g^2 / n - lam (sum g)^2 / n^2 )
//   BinsChamfer (losses/BinsChamferLoss.py:21-37 + pytorch3d chamfer_distance defaults): bidirectional squared-L2 chamfer
//           between the 256 bin centres and the valid GT depths of each image, point-mean, batch-mean.
// Both are HBM-bound streaming passes over the GT map: one read of gt + mask (+ the 4-tap gather of pred), fp32 math,
// reductions through per-block partials in a caller workspace added up in a FIXED order (deterministic, no float atomics).
// 1-D chamfer: the centres are sorted (cumsum of positive widths), so a target's nearest centre is a binary search over a
// 256-entry shared-memory table, and a centre's nearest target is the last target below / first target above it, which
// only needs a per-interval min / max (integer atomics on positive floats are order-independent).
// Contract: include/objcavit_b200.h (ocv_silog_*, ocv_bins_chamfer_*).
#include "common.cuh"
#include <math.h>

namespace {

constexpr int LB = 256;        // threads per block
constexpr int L_MAXBLK = 1024; // blocks per launch (upper bound of the partial arrays)

struct Bilinear {
  int y0, y1, x0, x1;
  float ly, lx;
};

// torch upsample_bilinear2d, align_corners=True: src = dst * (in-1)/(out-1)
__device__ __forceinline__ Bilinear bilinear_src(int Y, int X, float sy, float sx, int h, int w) {
  Bilinear r;
  const float fy = sy * (float)Y, fx = sx * (float)X;
  r.y0 = (int)fy;
  r.x0 = (int)fx;
  r.y1 = r.y0 + (r.y0 < h - 1 ? 1 : 0);
  r.x1 = r.x0 + (r.x0 < w - 1 ? 1 : 0);
  r.ly = fy - (float)r.y0;
  r.lx = fx - (float)r.x0;
  return r;
}

__device__ __forceinline__ float bilinear_val(const float* p, const Bilinear& r, int w) {
  const float h0 = 1.f - r.ly, w0 = 1.f - r.lx;
  return h0 * (w0 * p[r.y0 * w + r.x0] + r.lx * p[r.y0 * w + r.x1]) + r.ly * (w0 * p[r.y1 * w + r.x0] + r.lx * p[r.y1 * w + r.x1]);
}

__device__ __forceinline__ double block_sum(double v, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0)
    for (int i = 0; i < LB / 32; ++i) t += sh[i];
  return t;  // valid on thread 0
}

// pass 1: per-block partial (n, sum g, sum g^2)
__global__ void __launch_bounds__(LB)
silog_partial_kernel(const float* __restrict__ pred, const float* __restrict__ gt, const uint8_t* __restrict__ mask,
                     int B, int h, int w, int H, int W, float sy, float sx, double* __restrict__ part) {
  __shared__ double sh[LB / 32];
  pdl_launch();
  pdl_wait();
  const long long total = (long long)B * H * W;
  double n = 0.0, s1 = 0.0, s2 = 0.0;
  for (long long i = (long long)blockIdx.x * LB + threadIdx.x; i < total; i += (long long)gridDim.x * LB) {
    if (mask && !mask[i]) continue;
    const int b = (int)(i / ((long long)H * W));
    const int r = (int)(i - (long long)b * H * W);
    const int Y = r / W, X = r - Y * W;
    const Bilinear bl = bilinear_src(Y, X, sy, sx, h, w);
    const float p = bilinear_val(pred + (long long)b * h * w, bl, w);
    const float g = logf(p) - logf(gt[i]);
    n += 1.0;
    s1 += (double)g;
    s2 += (double)g * (double)g;
  }
  const double tn = block_sum(n, sh), t1 = block_sum(s1, sh), t2 = block_sum(s2, sh);
  if (threadIdx.x == 0) {
    part[3 * blockIdx.x] = tn;
    part[3 * blockIdx.x + 1] = t1;
    part[3 * blockIdx.x + 2] = t2;
  }
}

// pass 2 (one warp): fixed-order sum of the partials -> stats (n, sum g, sum g^2, Dg), loss
__global__ void silog_final_kernel(const double* __restrict__ part, int nblk, float alph, float lam, double* __restrict__ stats,
                                   float* __restrict__ loss) {
  pdl_launch();
  pdl_wait();
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  // one warp, fixed order: lane l adds partials l, l+32, ...; then a fixed shuffle tree (deterministic)
  double n = 0.0, s1 = 0.0, s2 = 0.0;
  for (int i = threadIdx.x; i < nblk; i += 32) {
    n += part[3 * i];
    s1 += part[3 * i + 1];
    s2 += part[3 * i + 2];
  }
  for (int o = 16; o > 0; o >>= 1) {
    n += __shfl_xor_sync(0xffffffffu, n, o);
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  if (threadIdx.x != 0) return;
  const double Dg = s2 / n - ((double)lam / (n * n)) * (s1 * s1);
  stats[0] = n; stats[1] = s1; stats[2] = s2; stats[3] = Dg;
  *loss = (float)((double)alph * sqrt(Dg));
}

// backward pass 1: gradient w.r.t. the UP-SAMPLED prediction, per GT pixel (0 where masked out)
//   dL/dg_i = alph / sqrt(Dg) * (g_i / n - lam * sum g / n^2),  dg_i/dp_i = 1 / p_i
__global__ void __launch_bounds__(LB)
silog_grad_up_kernel(const float* __restrict__ pred, const float* __restrict__ gt, const uint8_t* __restrict__ mask, int B,
                     int h, int w, int H, int W, float sy, float sx, const double* __restrict__ stats, float alph, float lam,
                     float grad_scale, float* __restrict__ gup) {
  pdl_launch();
  pdl_wait();
  const long long total = (long long)B * H * W;
  const double n = stats[0], s1 = stats[1], Dg = stats[3];
  const float c0 = (float)((double)alph / sqrt(Dg) / n) * grad_scale;
  const float c1 = (float)((double)lam * s1 / n);
  for (long long i = (long long)blockIdx.x * LB + threadIdx.x; i < total; i += (long long)gridDim.x * LB) {
    float gv = 0.f;
    if (!mask || mask[i]) {
      const int b = (int)(i / ((long long)H * W));
      const int r = (int)(i - (long long)b * H * W);
      const int Y = r / W, X = r - Y * W;
      const Bilinear bl = bilinear_src(Y, X, sy, sx, h, w);
      const float p = bilinear_val(pred + (long long)b * h * w, bl, w);
      const float g = logf(p) - logf(gt[i]);
      gv = c0 * (g - c1) / p;
    }
    gup[i] = gv;
  }
}

// backward pass 2: adjoint of the bilinear up-sampling as a GATHER (deterministic): every source pixel collects the
// up-sampled pixels whose 2x2 footprint contains it
__global__ void __launch_bounds__(LB)
silog_grad_pred_kernel(const float* __restrict__ gup, int B, int h, int w, int H, int W, float sy, float sx,
                       float* __restrict__ gpred) {
  pdl_launch();
  pdl_wait();
  const long long total = (long long)B * h * w;
  const long long i = (long long)blockIdx.x * LB + threadIdx.x;
  if (i >= total) return;
  const int b = (int)(i / ((long long)h * w));
  const int r = (int)(i - (long long)b * h * w);
  const int y = r / w, x = r - y * w;
  // destination rows whose source row index y0 is y-1 or y: fy in [y-1, y+1)
  int Ylo = 0, Yhi = H - 1, Xlo = 0, Xhi = W - 1;
  if (sy > 0.f) {
    Ylo = max(0, (int)floorf((float)(y - 1) / sy) - 1);
    Yhi = min(H - 1, (int)ceilf((float)(y + 1) / sy) + 1);
  }
  if (sx > 0.f) {
    Xlo = max(0, (int)floorf((float)(x - 1) / sx) - 1);
    Xhi = min(W - 1, (int)ceilf((float)(x + 1) / sx) + 1);
  }
  const float* g = gup + (long long)b * H * W;
  float acc = 0.f;
  for (int Y = Ylo; Y <= Yhi; ++Y) {
    const float fy = sy * (float)Y;
    const int y0 = (int)fy;
    const int y1 = y0 + (y0 < h - 1 ? 1 : 0);
    const float ly = fy - (float)y0;
    float wy = 0.f;
    if (y0 == y) wy += 1.f - ly;
    if (y1 == y) wy += ly;
    if (wy == 0.f) continue;
    for (int X = Xlo; X <= Xhi; ++X) {
      const float fx = sx * (float)X;
      const int x0 = (int)fx;
      const int x1 = x0 + (x0 < w - 1 ? 1 : 0);
      const float lx = fx - (float)x0;
      float wx = 0.f;
      if (x0 == x) wx += 1.f - lx;
      if (x1 == x) wx += lx;
      if (wx != 0.f) acc = fmaf(wy * wx, g[(long long)Y * W + X], acc);
    }
  }
  gpred[i] = acc;
}

inline int grid_for(long long n) {
  long long g = (n + LB - 1) / LB;
  return (int)(g < 1 ? 1 : (g > L_MAXBLK ? L_MAXBLK : g));
}

inline float scale_ac(int in, int out) { return out > 1 ? (float)(in - 1) / (float)(out - 1) : 0.f; }

}  // namespace

extern "C" long long ocv_silog_workspace(int B, int H, int W) {
  if (B <= 0 || H <= 0 || W <= 0) return 0;
  // partials [L_MAXBLK][3] doubles + stats [4] doubles + the up-sampled gradient map [B,H,W] floats
  return (long long)(3 * L_MAXBLK + 4) * 8 + (long long)B * H * W * 4;
}

extern "C" int ocv_silog_forward(const float* depth_pred, const float* depth_gt, const uint8_t* depth_mask, int B, int h, int w,
                                 int H, int W, float alph, float lam, float* loss, void* workspace, void* stream) {
  OCV_CHECK_ARG(depth_pred && depth_gt && loss && workspace && B > 0 && h > 0 && w > 0 && H > 0 && W > 0);
  if (int s = ocv_require_sm100()) return s;
  double* part = static_cast<double*>(workspace);
  double* stats = part + 3 * L_MAXBLK;
  const int nblk = grid_for((long long)B * H * W);
  if (int s = ocv_launch_pdl(silog_partial_kernel, dim3(nblk), dim3(LB), 0, stream, depth_pred, depth_gt, depth_mask, B, h, w, H,
                             W, scale_ac(h, H), scale_ac(w, W), part))
    return s;
  return ocv_launch_pdl(silog_final_kernel, dim3(1), dim3(32), 0, stream, (const double*)part, nblk, alph, lam, stats, loss);
}

extern "C" int ocv_silog_backward(const float* depth_pred, const float* depth_gt, const uint8_t* depth_mask, int B, int h, int w,
                                  int H, int W, float alph, float lam, float grad_scale, float* grad_pred, void* workspace,
                                  void* stream) {
  OCV_CHECK_ARG(depth_pred && depth_gt && grad_pred && workspace && B > 0 && h > 0 && w > 0 && H > 0 && W > 0);
  if (int s = ocv_require_sm100()) return s;
  double* stats = static_cast<double*>(workspace) + 3 * L_MAXBLK;  // written by ocv_silog_forward on the same workspace
  float* gup = reinterpret_cast<float*>(stats + 4);
  const float sy = scale_ac(h, H), sx = scale_ac(w, W);
  if (int s = ocv_launch_pdl(silog_grad_up_kernel, dim3(grid_for((long long)B * H * W)), dim3(LB), 0, stream, depth_pred, depth_gt,
                             depth_mask, B, h, w, H, W, sy, sx, (const double*)stats, alph, lam, grad_scale, gup))
    return s;
  const long long n = (long long)B * h * w;
  return ocv_launch_pdl(silog_grad_pred_kernel, dim3((unsigned)((n + LB - 1) / LB)), dim3(LB), 0, stream, (const float*)gup, B, h, w,
                        H, W, sy, sx, grad_pred);
}

// ------------------------------------------------------------------------------------------------------------------
// BinsChamfer
// ------------------------------------------------------------------------------------------------------------------
namespace {

constexpr int CH_MAXK = 1024;          // bins
constexpr int CH_BPI = 64;             // blocks per image (upper bound)
constexpr double CH_FIX = 16777216.0;  // 2^24: fixed-point scale of the per-centre target sums (integer atomics are deterministic)

struct ChamferWs {
  int* cnt;                  // [B,K]    targets whose nearest centre is k
  unsigned long long* sumt;  // [B,K]    their sum, fixed point
  unsigned int* tmin;        // [B,K+1]  smallest target in interval j = (c_{j-1}, c_j]   (float bits; +inf when empty)
  unsigned int* tmax;        // [B,K+1]  largest target in interval j                      (float bits; 0 when empty)
  double* part;              // [B,CH_BPI,2]  per-block (sum d^2, T)
};

inline size_t ch_align(size_t v) { return (v + 255) / 256 * 256; }

inline ChamferWs chamfer_carve(void* ws, int B, int K, size_t* bytes) {
  uint8_t* p = static_cast<uint8_t*>(ws);
  size_t off = 0;
  ChamferWs W;
  W.cnt = reinterpret_cast<int*>(p + off); off = ch_align(off + (size_t)B * K * 4);
  W.sumt = reinterpret_cast<unsigned long long*>(p + off); off = ch_align(off + (size_t)B * K * 8);
  W.tmin = reinterpret_cast<unsigned int*>(p + off); off = ch_align(off + (size_t)B * (K + 1) * 4);
  W.tmax = reinterpret_cast<unsigned int*>(p + off); off = ch_align(off + (size_t)B * (K + 1) * 4);
  W.part = reinterpret_cast<double*>(p + off); off = ch_align(off + (size_t)B * CH_BPI * 2 * 8);
  if (bytes) *bytes = off;
  return W;
}

__global__ void chamfer_init_kernel(ChamferWs W, int B, int K) {
  pdl_launch();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B * K) {
    W.cnt[i] = 0;
    W.sumt[i] = 0ull;
  }
  if (i < B * (K + 1)) {
    W.tmin[i] = 0x7f800000u;
    W.tmax[i] = 0u;
  }
  if (i < B * CH_BPI * 2) W.part[i] = 0.0;
}

// grid (bpi, B): targets of image b -> nearest centre by binary search over the shared-memory table
__global__ void __launch_bounds__(LB)
chamfer_targets_kernel(const float* __restrict__ edges, const float* __restrict__ gt, const uint8_t* __restrict__ mask, int K,
                       long long HW, ChamferWs W) {
  __shared__ float c[CH_MAXK];
  __shared__ int s_cnt[CH_MAXK];
  __shared__ unsigned long long s_sum[CH_MAXK];
  __shared__ unsigned int s_min[CH_MAXK + 1], s_max[CH_MAXK + 1];
  __shared__ double sh[LB / 32];
  pdl_launch();
  pdl_wait();
  const int b = blockIdx.y;
  const float* e = edges + (long long)b * (K + 1);
  for (int k = threadIdx.x; k <= K; k += LB) {
    if (k < K) {
      c[k] = 0.5f * (e[k + 1] + e[k]);   // BinsChamferLoss.py:26
      s_cnt[k] = 0;
      s_sum[k] = 0ull;
    }
    s_min[k] = 0x7f800000u;
    s_max[k] = 0u;
  }
  __syncthreads();
  const float* g = gt + (long long)b * HW;
  const uint8_t* m = mask ? mask + (long long)b * HW : nullptr;
  double d2 = 0.0, T = 0.0;
  for (long long i = (long long)blockIdx.x * LB + threadIdx.x; i < HW; i += (long long)gridDim.x * LB) {
    if (m && !m[i]) continue;
    const float t = g[i];
    int lo = 0, hi = K;          // j = number of centres < t
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (c[mid] < t) lo = mid + 1; else hi = mid;
    }
    const int j = lo;
    const int ka = max(j - 1, 0), kb = min(j, K - 1);
    const float da = (c[ka] - t) * (c[ka] - t), db = (c[kb] - t) * (c[kb] - t);
    const int kn = db < da ? kb : ka;
    d2 += (double)fminf(da, db);
    T += 1.0;
    atomicAdd(&s_cnt[kn], 1);
    atomicAdd(&s_sum[kn], (unsigned long long)((double)t * CH_FIX + 0.5));
    atomicMin(&s_min[j], __float_as_uint(t));   // t > 0: the bit patterns order like the floats
    atomicMax(&s_max[j], __float_as_uint(t));
  }
  __syncthreads();
  for (int k = threadIdx.x; k <= K; k += LB) {
    if (k < K && s_cnt[k]) {
      atomicAdd(&W.cnt[(long long)b * K + k], s_cnt[k]);
      atomicAdd(&W.sumt[(long long)b * K + k], s_sum[k]);
    }
    if (s_max[k]) {
      atomicMin(&W.tmin[(long long)b * (K + 1) + k], s_min[k]);
      atomicMax(&W.tmax[(long long)b * (K + 1) + k], s_max[k]);
    }
  }
  const double td = block_sum(d2, sh), tT = block_sum(T, sh);
  if (threadIdx.x == 0) {
    W.part[((long long)b * CH_BPI + blockIdx.x) * 2] = td;
    W.part[((long long)b * CH_BPI + blockIdx.x) * 2 + 1] = tT;
  }
}

// one block: per image the centre -> nearest-target direction, both means, the loss and d loss / d bin_edges
__global__ void __launch_bounds__(LB)
chamfer_final_kernel(const float* __restrict__ edges, int B, int K, int bpi, ChamferWs W, float grad_scale,
                     float* __restrict__ loss, float* __restrict__ grad_edges) {
  __shared__ float below[CH_MAXK], above[CH_MAXK], gc[CH_MAXK];
  __shared__ double sh[LB / 32], sp[2 * CH_BPI];
  __shared__ double s_T, s_dy;
  pdl_launch();
  pdl_wait();
  double total = 0.0;
  for (int b = 0; b < B; ++b) {
    const float* e = edges + (long long)b * (K + 1);
    // stage everything thread 0 scans into shared memory with parallel (independent) loads first: a single thread walking
    // global memory paid one L2 round trip per element (190 us per call at B = 8, K = 256)
    for (int i = threadIdx.x; i < 2 * bpi; i += LB) sp[i] = W.part[(long long)b * CH_BPI * 2 + i];
    for (int k = threadIdx.x; k < K; k += LB) {
      const unsigned int lo = W.tmax[(long long)b * (K + 1) + k], hi = W.tmin[(long long)b * (K + 1) + k + 1];
      below[k] = lo ? __uint_as_float(lo) : 0.f;
      above[k] = hi != 0x7f800000u ? __uint_as_float(hi) : INFINITY;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double T = 0.0, dyp = 0.0;
      for (int i = 0; i < bpi; ++i) {
        dyp += sp[2 * i];
        T += sp[2 * i + 1];
      }
      s_T = T;
      s_dy = dyp;
      float run = 0.f;  // largest target <= c_k (0 = none: targets are positive)
      for (int k = 0; k < K; ++k) {
        run = fmaxf(run, below[k]);
        below[k] = run;
      }
      float up = INFINITY;  // smallest target > c_k
      for (int k = K - 1; k >= 0; --k) {
        up = fminf(up, above[k]);
        above[k] = up;
      }
    }
    __syncthreads();
    const double T = s_T;
    double dx = 0.0;
    for (int k = threadIdx.x; k < K; k += LB) {
      const float c = 0.5f * (e[k + 1] + e[k]);
      float g = 0.f;
      if (T > 0.0) {
        const float dlo = below[k] > 0.f ? (c - below[k]) * (c - below[k]) : INFINITY;
        const float dhi = (above[k] - c) * (above[k] - c);   // inf when there is no target above
        const float tn = dhi < dlo ? above[k] : below[k];
        dx += (double)fminf(dlo, dhi);
        // d/dc of  (1/K) sum_k (c_k - t*_k)^2  +  (1/T) sum_t (c_nn(t) - t)^2 , each averaged over the batch
        const double cnt = (double)W.cnt[(long long)b * K + k];
        const double st = (double)W.sumt[(long long)b * K + k] / CH_FIX;
        g = (float)((2.0 * ((double)c - (double)tn) / K + 2.0 * (cnt * (double)c - st) / T) / B) * grad_scale;
      }
      gc[k] = g;
    }
    const double tx = block_sum(dx, sh);
    if (threadIdx.x == 0 && T > 0.0) total += tx / K + s_dy / T;
    __syncthreads();
    if (grad_edges) {  // c_k = (e_k + e_{k+1}) / 2
      for (int j = threadIdx.x; j <= K; j += LB)
        grad_edges[(long long)b * (K + 1) + j] = 0.5f * ((j > 0 ? gc[j - 1] : 0.f) + (j < K ? gc[j] : 0.f));
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *loss = (float)(total / B);
}

}  // namespace

extern "C" long long ocv_bins_chamfer_workspace(int B, int K) {
  if (B <= 0 || K <= 0) return 0;
  size_t bytes = 0;
  chamfer_carve(nullptr, B, K, &bytes);
  return (long long)bytes;
}

extern "C" int ocv_bins_chamfer(const float* bin_edges, const float* depth_gt, const uint8_t* depth_mask, int B, int K,
                                long long HW, float grad_scale, float* loss, float* grad_edges, void* workspace, void* stream) {
  OCV_CHECK_ARG(bin_edges && depth_gt && loss && workspace && B > 0 && HW > 0);
  OCV_CHECK_SHAPE(K >= 2 && K <= CH_MAXK && B <= 65535);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(workspace) & 255) == 0);
  if (int s = ocv_require_sm100()) return s;
  ChamferWs W = chamfer_carve(workspace, B, K, nullptr);
  const int n_init = B * (K + 1) > B * CH_BPI * 2 ? B * (K + 1) : B * CH_BPI * 2;
  if (int s = ocv_launch_pdl(chamfer_init_kernel, dim3((n_init + 255) / 256), dim3(256), 0, stream, W, B, K)) return s;
  long long want = (HW + (long long)LB * 8 - 1) / ((long long)LB * 8);
  const int bpi = (int)(want < 1 ? 1 : (want > CH_BPI ? CH_BPI : want));
  if (int s = ocv_launch_pdl(chamfer_targets_kernel, dim3(bpi, B), dim3(LB), 0, stream, bin_edges, depth_gt, depth_mask, K, HW, W))
    return s;
  return ocv_launch_pdl(chamfer_final_kernel, dim3(1), dim3(LB), 0, stream, bin_edges, B, K, bpi, W, grad_scale, loss, grad_edges);
}
