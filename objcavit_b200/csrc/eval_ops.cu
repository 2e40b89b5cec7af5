// The rows either side of the head, all HBM-bound streaming / index work:
//   N2  evaluation epilogue: clamp + un-mirror + average of the two test-time passes (modules/GraphBinsLM.py:159-183)
//   N2/N3 metrics: bilinear up-sampling to the GT size (align_corners=True) + nan_to_num + validity / crop mask
//       (metrics/MetricsPreprocess.py:17-45) fused with the eight masked reductions of metrics/*.py (abs_rel, sq_rel, rmse,
//       rmse_log, log10, delta<1.25^k) — one pass over the ground truth instead of 16 torchmetrics updates on cloned tensors
//   N4  object front-end glue: per-image reversal + xyxy -> centre/size boxes (modules/Yolov7Wrapper.py:118-123) and the
//       relative-size bucket of every object against the next one (modules/ObjectLanguageStrategy.py:61-76), on device.
// Algorithmic bytes: merge 12 B / low-res pixel; metrics 4 B / GT pixel (+ 4 when the up-sampled map is wanted; the
// low-res prediction is read through L1/L2).  Reductions are two-stage, fixed order (deterministic), accumulated in fp64.
// Contract: include/objcavit_b200.h.
#include "common.cuh"
#include <limits.h>
#include <math.h>

namespace {

constexpr int NT = 256;
constexpr int MAXBLK = 1184;  // 8 x 148
constexpr int NACC = 9;       // abs_rel, sq_rel, sq_diff, sq_log_diff, abs_log10_diff, d1, d2, d3, valid count

__global__ void tta_merge_kernel(const float* __restrict__ a, const float* __restrict__ b_mirror, float* __restrict__ out,
                                 long long rows, int w, float lo, float hi) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * w) return;
  const long long r = i / w;
  const int x = (int)(i - r * w);
  // torch.clamp keeps NaN; fminf/fmaxf would drop it
  float va = a[i];
  va = va != va ? va : fminf(fmaxf(va, lo), hi);
  if (!b_mirror) {
    out[i] = va;
    return;
  }
  float vb = b_mirror[r * w + (w - 1 - x)];
  vb = vb != vb ? vb : fminf(fmaxf(vb, lo), hi);
  out[i] = 0.5f * (va + vb);
}

struct MetricArgs {
  int B, h, w, H, W;
  float sy, sx, lo, hi;
  int y0, y1, x0, x1;  // crop rectangle [y0,y1) x [x0,x1); y0 < 0 = no crop
};

__global__ void __launch_bounds__(NT)
metrics_partial_kernel(const float* __restrict__ pred, const float* __restrict__ gt, float* __restrict__ up_out, MetricArgs a,
                       double* __restrict__ part) {
  __shared__ double sh[NT / 32][NACC];
  float acc[NACC];
#pragma unroll
  for (int k = 0; k < NACC; ++k) acc[k] = 0.f;
  double dacc[NACC];
#pragma unroll
  for (int k = 0; k < NACC; ++k) dacc[k] = 0.0;
  const long long HW = (long long)a.H * a.W, total = HW * a.B;
  int since = 0;
  for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < total; i += (long long)gridDim.x * NT) {
    const int b = (int)(i / HW);
    const long long r = i - (long long)b * HW;
    const int Y = (int)(r / a.W), X = (int)(r - (long long)Y * a.W);
    // torch upsample_bilinear2d, align_corners=True
    const float fy = a.sy * (float)Y, fx = a.sx * (float)X;
    const int yy0 = (int)fy, xx0 = (int)fx;
    const int yy1 = yy0 + (yy0 < a.h - 1 ? 1 : 0), xx1 = xx0 + (xx0 < a.w - 1 ? 1 : 0);
    const float ly = fy - (float)yy0, lx = fx - (float)xx0, h0 = 1.f - ly, w0 = 1.f - lx;
    const float* p = pred + (long long)b * a.h * a.w;
    float v = h0 * (w0 * p[yy0 * a.w + xx0] + lx * p[yy0 * a.w + xx1]) + ly * (w0 * p[yy1 * a.w + xx0] + lx * p[yy1 * a.w + xx1]);
    // nan_to_num(nan=min_depth, posinf=max_depth, neginf=max_depth)
    if (v != v) v = a.lo;
    else if (isinf(v)) v = a.hi;
    if (up_out) up_out[i] = v;
    const float t = gt[i];
    bool ok = t > a.lo && t <= a.hi;
    if (a.y0 >= 0) ok = ok && Y >= a.y0 && Y < a.y1 && X >= a.x0 && X < a.x1;
    if (ok) {
      const float d = t - v;
      acc[0] += fabsf(d) / t;
      acc[1] += d * d / t;
      acc[2] += d * d;
      const float lg = logf(t) - logf(v);
      acc[3] += lg * lg;
      acc[4] += fabsf(log10f(t) - log10f(v));
      const float th = fmaxf(t / v, v / t);
      acc[5] += th < 1.25f ? 1.f : 0.f;
      acc[6] += th < 1.5625f ? 1.f : 0.f;
      acc[7] += th < 1.953125f ? 1.f : 0.f;
      acc[8] += 1.f;
    }
    if (++since == 8) {  // bound the fp32 running sums: spill to fp64 every 8 pixels
#pragma unroll
      for (int k = 0; k < NACC; ++k) {
        dacc[k] += (double)acc[k];
        acc[k] = 0.f;
      }
      since = 0;
    }
  }
#pragma unroll
  for (int k = 0; k < NACC; ++k) {
    double v = dacc[k] + (double)acc[k];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x % 32 == 0) sh[threadIdx.x / 32][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < NACC) {
    double t = 0.0;
    for (int wv = 0; wv < NT / 32; ++wv) t += sh[wv][threadIdx.x];
    part[(long long)blockIdx.x * NACC + threadIdx.x] = t;
  }
}

// sums[k] (this call) and acc[k] += sums[k] (running state across calls).  One warp per accumulator: lane l adds the partials of
// blocks l, l+32, ... in order, then a fixed shuffle tree — deterministic, and 32 independent load streams instead of one thread
// walking 1184 blocks (which cost 60 us).
__global__ void __launch_bounds__(NACC * 32)
metrics_final_kernel(const double* __restrict__ part, int nblk, double* __restrict__ sums, double* __restrict__ acc) {
  const int k = threadIdx.x / 32, lane = threadIdx.x % 32;
  double t = 0.0;
  for (int i = lane; i < nblk; i += 32) t += part[(long long)i * NACC + k];
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if (lane == 0) {
    if (sums) sums[k] = t;
    if (acc) acc[k] += t;
  }
}

// one thread per object: reversed order inside the image, xyxy -> (cx, cy, w, h), relative-size bucket vs the next object
__global__ void object_glue_kernel(const float* __restrict__ xyxy, const int32_t* __restrict__ offsets, float* __restrict__ xywh,
                                   int32_t* __restrict__ bucket, int B, int reverse, int n_scale) {
  const int b = blockIdx.y;
  const int beg = offsets[b], n = offsets[b + 1] - beg;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  auto box = [&](int k, float& cx, float& cy, float& bw, float& bh) {
    const float* s = xyxy + (long long)(beg + (reverse ? n - 1 - k : k)) * 4;
    cx = __fdiv_rn(__fadd_rn(s[0], s[2]), 2.f);
    cy = __fdiv_rn(__fadd_rn(s[1], s[3]), 2.f);
    bw = __fsub_rn(s[2], s[0]);
    bh = __fsub_rn(s[3], s[1]);
  };
  float cx, cy, bw, bh;
  box(j, cx, cy, bw, bh);
  float* d = xywh + (long long)(beg + j) * 4;
  d[0] = cx; d[1] = cy; d[2] = bw; d[3] = bh;
  if (!bucket) return;
  int idx = -1;  // a single object has no clause (ObjectLanguageStrategy.py:64-65)
  if (n > 1) {
    float nx, ny, nw, nh;
    box((j + 1) % n, nx, ny, nw, nh);
    const float ratio = __fdiv_rn(__fmul_rn(bw, bh), __fmul_rn(nw, nh));  // fp32 tensor arithmetic in the reference
    if (!(ratio > 0.f)) {
      idx = INT_MIN;  // math.log raises on zero / negative / NaN area ratios in the reference: flagged, never a silent bucket
    } else {
      double rel = log((double)ratio) + 1.0;  // math.log on the python float
      rel /= 2.0;
      rel *= (double)(n_scale - 3);
      rel = rint(rel) + 1.0;                  // np.round: half to even
      idx = (int)fmin(fmax(rel, 0.0), (double)(n_scale - 1));
    }
  }
  bucket[beg + j] = idx;
}

}  // namespace

extern "C" int ocv_tta_merge_f32(const float* depth, const float* depth_mirror, float* out, int B, int h, int w, float min_depth,
                                 float max_depth, void* stream) {
  OCV_CHECK_ARG(depth && out && B > 0 && h > 0 && w > 0);
  if (int s = ocv_require_sm100()) return s;
  const long long n = (long long)B * h * w;
  tta_merge_kernel<<<(unsigned)ceil_div<long long>(n, NT), NT, 0, ocv_stream(stream)>>>(depth, depth_mirror, out, (long long)B * h, w,
                                                                                      min_depth, max_depth);
  return ocv_launch_status();
}

extern "C" long long ocv_depth_metrics_workspace(void) { return (long long)MAXBLK * NACC * sizeof(double); }

extern "C" int ocv_depth_metrics_f32(const float* depth_pred, const float* depth_gt, int B, int h, int w, int H, int W,
                                     float min_depth, float max_depth, const int* crop_host, float* upsampled_out, double* sums,
                                     double* acc, void* workspace, void* stream) {
  OCV_CHECK_ARG(depth_pred && depth_gt && workspace && (sums || acc) && B > 0 && h > 0 && w > 0 && H > 0 && W > 0);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(workspace) & 7) == 0);
  if (int s = ocv_require_sm100()) return s;
  MetricArgs a;
  a.B = B; a.h = h; a.w = w; a.H = H; a.W = W;
  a.sy = H > 1 ? (float)(h - 1) / (float)(H - 1) : 0.f;
  a.sx = W > 1 ? (float)(w - 1) / (float)(W - 1) : 0.f;
  a.lo = min_depth; a.hi = max_depth;
  a.y0 = -1; a.y1 = a.x0 = a.x1 = 0;
  if (crop_host) { a.y0 = crop_host[0]; a.y1 = crop_host[1]; a.x0 = crop_host[2]; a.x1 = crop_host[3]; }
  const long long total = (long long)B * H * W;
  const int nblk = (int)(ceil_div<long long>(total, (long long)NT * 8) < MAXBLK ? ceil_div<long long>(total, (long long)NT * 8) : MAXBLK);
  double* part = static_cast<double*>(workspace);
  metrics_partial_kernel<<<nblk, NT, 0, ocv_stream(stream)>>>(depth_pred, depth_gt, upsampled_out, a, part);
  if (int s = ocv_launch_status()) return s;
  metrics_final_kernel<<<1, NACC * 32, 0, ocv_stream(stream)>>>(part, nblk, sums, acc);
  return ocv_launch_status();
}

extern "C" int ocv_object_glue_f32(const float* xyxy, const int32_t* offsets, float* xywh, int32_t* size_bucket, int B, int n_max,
                                   int reverse, int n_scale, void* stream) {
  OCV_CHECK_ARG(xyxy && offsets && xywh && B > 0 && n_max > 0);
  OCV_CHECK_SHAPE(B <= 65535 && (!size_bucket || n_scale >= 4));
  if (int s = ocv_require_sm100()) return s;
  object_glue_kernel<<<dim3(ceil_div(n_max, 64), B), 64, 0, ocv_stream(stream)>>>(xyxy, offsets, xywh, size_bucket, B, reverse, n_scale);
  return ocv_launch_status();
}
