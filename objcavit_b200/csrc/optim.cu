// Optimizer step of the training path on FLAT fp32 buffers: the head's parameters, gradients and AdamW
// moments each live in one contiguous buffer (objcavit_b200/grad_sync.py), so gradient-norm clipping and the AdamW update are
// two launches for the whole head instead of ~150 per-tensor torch launches.  Both kernels are HBM-bound streams:
//   sumsq : reads n floats                       (4 B / element)
//   adamw : reads param, grad, m, v; writes param, m, v   (28 B / element)
// replaces: Lightning's gradient_clip_val=0.1 -> torch.nn.utils.clip_grad_norm_ (main.py:35-38,64) and torch.optim.AdamW.step
// (modules/GraphBinsLM.py:467).  Contract: include/objcavit_b200.h.
#include "common.cuh"
#include <math.h>

namespace {

constexpr int SUMSQ_BLOCKS = 592;  // 4 x 148 SMs; fixed so that the reduction order never depends on the device state
constexpr int NT = 256;

__global__ void __launch_bounds__(NT) sumsq_partial_kernel(const float* __restrict__ x, long long n, float* __restrict__ part) {
  __shared__ float sh[NT / 32];
  float acc = 0.f;
  const long long n4 = n / 4;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < n4; i += (long long)gridDim.x * NT) {
    const float4 v = x4[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0)
    for (long long i = n4 * 4 + threadIdx.x; i < n; i += NT) acc += x[i] * x[i];
  acc = warp_sum(acc);
  if (threadIdx.x % 32 == 0) sh[threadIdx.x / 32] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < NT / 32; ++i) t += sh[i];
    part[blockIdx.x] = t;
  }
}

__global__ void sumsq_final_kernel(const float* __restrict__ part, int nparts, float scale2, float* __restrict__ out) {
  // one warp, fixed order: lane-strided partial sums, then the shuffle tree
  float t = 0.f;
  for (int i = threadIdx.x; i < nparts; i += 32) t += part[i];
  t = warp_sum(t);
  if (threadIdx.x == 0) out[0] = t * scale2;
}

struct AdamArgs {
  float lr, beta1, beta2, eps, wd, bc1, bc2_sqrt, grad_scale, max_norm;
};

__device__ __forceinline__ void adamw_one(float& p, float g, float& m, float& v, const AdamArgs& a, float gs) {
  g *= gs;
  p *= 1.f - a.lr * a.wd;                       // decoupled weight decay
  m += (g - m) * (1.f - a.beta1);               // lerp, as torch
  v = a.beta2 * v + (1.f - a.beta2) * g * g;
  const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
  p -= (a.lr / a.bc1) * (m / denom);
}

__global__ void __launch_bounds__(NT) adamw_kernel(float* __restrict__ param, const float* __restrict__ grad, float* __restrict__ m,
                                                   float* __restrict__ v, long long n, AdamArgs a,
                                                   const float* __restrict__ grad_sumsq) {
  float gs = a.grad_scale;
  if (grad_sumsq) {  // clip_grad_norm_: coef = max_norm / (norm + 1e-6), clamped to 1
    const float norm = sqrtf(grad_sumsq[0]);
    gs *= fminf(a.max_norm / (norm + 1e-6f), 1.f);
  }
  const long long n4 = n / 4;
  float4* p4 = reinterpret_cast<float4*>(param);
  const float4* g4 = reinterpret_cast<const float4*>(grad);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  for (long long i = (long long)blockIdx.x * NT + threadIdx.x; i < n4; i += (long long)gridDim.x * NT) {
    float4 p = p4[i], mm = m4[i], vv = v4[i];
    const float4 g = g4[i];
    adamw_one(p.x, g.x, mm.x, vv.x, a, gs);
    adamw_one(p.y, g.y, mm.y, vv.y, a, gs);
    adamw_one(p.z, g.z, mm.z, vv.z, a, gs);
    adamw_one(p.w, g.w, mm.w, vv.w, a, gs);
    p4[i] = p; m4[i] = mm; v4[i] = vv;
  }
  if (blockIdx.x == 0)
    for (long long i = n4 * 4 + threadIdx.x; i < n; i += NT) adamw_one(param[i], grad[i], m[i], v[i], a, gs);
}

}  // namespace

extern "C" long long ocv_sumsq_workspace(long long n) { return n > 0 ? (long long)SUMSQ_BLOCKS * sizeof(float) : 0; }

extern "C" int ocv_sumsq_f32(const float* x, long long n, float scale, float* out, void* workspace, void* stream) {
  OCV_CHECK_ARG(x && out && workspace && n > 0);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(x) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  float* part = static_cast<float*>(workspace);
  sumsq_partial_kernel<<<SUMSQ_BLOCKS, NT, 0, ocv_stream(stream)>>>(x, n, part);
  if (int s = ocv_launch_status()) return s;
  sumsq_final_kernel<<<1, 32, 0, ocv_stream(stream)>>>(part, SUMSQ_BLOCKS, scale * scale, out);
  return ocv_launch_status();
}

extern "C" int ocv_adamw_step_f32(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                                  float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale,
                                  const float* grad_sumsq, float max_norm, void* stream) {
  OCV_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && n > 0 && step >= 1);
  OCV_CHECK_ALIGN(((reinterpret_cast<uintptr_t>(param) | reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(exp_avg) |
                    reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  AdamArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = weight_decay;
  a.bc1 = (float)(1.0 - pow((double)beta1, (double)step));
  a.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
  a.grad_scale = grad_scale; a.max_norm = max_norm;
  const long long blocks = ceil_div<long long>(n / 4 + 1, NT);
  const int grid = (int)(blocks < 148 * 8 ? blocks : 148 * 8);
  adamw_kernel<<<grid, NT, 0, ocv_stream(stream)>>>(param, grad, exp_avg, exp_avg_sq, n, a, grad_sumsq);
  return ocv_launch_status();
}
