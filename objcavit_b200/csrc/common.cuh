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

// csrc/attn_tc.cu: does the tcgen05 attention kernel cover this shape / these strides (elements)?
bool ocv_mha_tc_supported(int Sq, int Sk, int heads, int d, long long q_bs, int q_rs, long long k_bs, int k_rs, long long v_bs, int v_rs,
                          long long o_bs, int o_rs);

bool ocv_mha_bwd_tc_supported(int Sq, int Sk, int heads, int d);
int ocv_attn_tc_policy();   // 0 never, 1 everywhere covered, 2 (default) cross attention + training step

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

// ---- dropout (training mode) ---------------------------------------------------------------------------------------
// Counter-based: element `idx` of dropout site `seed` is kept iff the upper 32 bits of a splitmix64 hash of (seed, idx) are
// >= p * 2^32.  Stateless, so the backward pass regenerates exactly the forward mask from (seed, idx); the statistics match
// torch's Bernoulli(1-p) masks, the random stream itself is (necessarily) a different one.
__host__ __device__ __forceinline__ uint32_t ocv_drop_threshold(float p) {
  const double t = (double)p * 4294967296.0;
  return t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
}
__device__ __forceinline__ bool ocv_drop_keep(unsigned long long seed, unsigned long long idx, uint32_t thresh) {
  unsigned long long z = seed + idx * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (uint32_t)(z >> 32) >= thresh;
}

// Per-step salt of the dropout masks, read from DEVICE memory by every dropout kernel (nullptr = none: seeds are used as passed).
// A CUDA-graph replay of the training step freezes the seeds (kernel arguments); the salt is what changes between replays.
// Set with ocv_set_dropout_salt (csrc/api.cu); the effective seed is seed ^ *salt.
const unsigned long long* ocv_dropout_salt();
__device__ __forceinline__ unsigned long long ocv_salted(unsigned long long seed, const unsigned long long* salt) {
  return salt ? (seed ^ __ldg(salt)) : seed;
}

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------------------
// The head is a chain of ~50 short dependent kernels: launch latency and each kernel's prologue (barrier init, TMEM
// allocation, tensor-map fetch, weight tiles, bias vectors) would otherwise sit on the critical path.  Every kernel of
// the chain calls pdl_launch() first (the next kernel in the stream may start its own prologue) and pdl_wait() before
// it touches anything an earlier kernel produced (and before it writes anything an earlier kernel may still read).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool ocv_pdl_enabled();  // false when OCV_NO_PDL=1 is set in the environment (api.cu)

template <typename... KArgs, typename... Args>
static inline int ocv_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, void* stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = ocv_stream(stream);
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = ocv_pdl_enabled() ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
  if (e != cudaSuccess) return (int)e;
  return ocv_launch_status();
}
