// Shared helpers for the objcavit_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/objcavit_b200.h"

#define OCV_CHECK_ARG(cond) do { if (!(cond)) return OCV_E_ARG; } while (0)
#define OCV_CHECK_SHAPE(cond) do { if (!(cond)) return OCV_E_SHAPE; } while (0)
#define OCV_CHECK_ALIGN(cond) do { if (!(cond)) return OCV_E_ALIGN; } while (0)

// returns the cudaError of the launch just issued (does not synchronise)
static inline int ocv_launch_status() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? OCV_OK : (int)e;
}

// OCV_OK iff the current device is sm_100 (cached per process).
int ocv_require_sm100();

// pad_objects with an optional second (bf16) copy of the padded rows (token_ops.cu); used by the head executor
int ocv_pad_objects_impl(const float* src, const int32_t* offsets, const int32_t* mask_offsets, float* dst, void* dst16,
                         uint8_t* mask, int B, int L, int Nmax, int E, float pad_value, int left_pad, void* stream);

static inline cudaStream_t ocv_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

template <typename T>
static inline T ceil_div(T a, T b) { return (a + b - 1) / b; }
