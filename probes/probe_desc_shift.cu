// Probe (test tool, not product): does a tcgen05 SW128 K-major A descriptor whose start address is shifted by
// whole 128-byte rows (not 1024-B aligned) read the rows TMA wrote, and which base_offset / SBO values
// are legal?  Result decides the halo-tile layout of the fused pixel kernel (DESIGN.md "conv3x3").
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../include -o probe_desc_shift.bin probe_desc_shift.cu ../objcavit_b200/csrc/tc_gemm.cu ../objcavit_b200/csrc/api.cu
#include "../objcavit_b200/csrc/tc_common.cuh"
#include <vector>
#include <cstdlib>

using namespace tc;

struct Cfg { int shift, sbo, bo; };
constexpr int NCFG_MAX = 64;
constexpr int ROWS = 512;

__device__ __forceinline__ uint64_t desc_full(uint32_t saddr, uint32_t sbo_bytes, uint32_t base_off) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) |
         (1ull << 46) | ((uint64_t)(base_off & 7) << 49) | ((uint64_t)LAYOUT_SW128 << 61);
}

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Cfg* cfgs, int ncfg,
             float* out /*[ncfg][128][16]*/) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* sA = smem;                    // 512 rows x 128 B = 64 KB
  uint8_t* sB = smem + ROWS * 128;       // 16 rows x 128 B = 2 KB
  uint64_t* bar_load = reinterpret_cast<uint64_t*>(sB + 2048);
  uint64_t* bar_mma = bar_load + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (threadIdx.x == 0) {
    mbar_init(bar_load, 1);
    mbar_init(bar_mma, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar_load, ROWS * 128 + 2048);
    tma_load_2d(sA, &tmA, bar_load, 0, 0);
    tma_load_2d(sA + 256 * 128, &tmA, bar_load, 0, 256);
    tma_load_2d(sB, &tmB, bar_load, 0, 0);
  }
  mbar_wait(bar_load, 0);
  __syncthreads();
  for (int c = 0; c < ncfg; ++c) {
    const Cfg cf = cfgs[c];
    if (threadIdx.x == 0) {
      tc_fence_after();
      constexpr uint32_t idesc = make_idesc_bf16(128, 16);
      const uint32_t a0 = smem_u32(sA) + cf.shift * 128;
      for (int kk = 0; kk < 4; ++kk) {
        const uint64_t da = desc_full(a0 + kk * 32, cf.sbo, cf.bo);
        const uint64_t db = desc_full(smem_u32(sB) + kk * 32, 1024, 0);
        umma_bf16(tmem_d, da, db, idesc, kk != 0);
      }
      umma_commit(bar_mma);
    }
    mbar_wait(bar_mma, c & 1);
    tc_fence_after();
    uint32_t r[32];
    tmem_ld32(tmem_d + ((uint32_t)(warp * 32) << 16), r);
    tmem_ld_wait();
    float* o = out + ((long long)c * 128 + warp * 32 + lane) * 16;
    for (int j = 0; j < 16; ++j) o[j] = __uint_as_float(r[j]);
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) tmem_dealloc(tmem_d, 32);
}

int main() {
  std::vector<__nv_bfloat16> hA(ROWS * 64), hB(16 * 64);
  std::vector<float> fA(ROWS * 64), fB(16 * 64);
  for (int r = 0; r < ROWS; ++r)
    for (int c = 0; c < 64; ++c) {
      fA[r * 64 + c] = (float)(((r * 7 + c * 3 + (r / 8) * 5) % 17) - 8);
      hA[r * 64 + c] = __float2bfloat16(fA[r * 64 + c]);
    }
  srand(1);
  for (int i = 0; i < 16 * 64; ++i) {
    fB[i] = (float)((rand() % 7) - 3);
    hB[i] = __float2bfloat16(fB[i]);
  }
  std::vector<Cfg> cfgs;
  const int shifts[] = {0, 1, 2, 3, 7, 9};
  const int sbos[] = {1024, 1280, 2048};
  for (int s : shifts)
    for (int sbo : sbos)
      for (int mode = 0; mode < 3; ++mode) {
        int bo = mode == 0 ? 0 : (mode == 1 ? (s & 7) : ((8 - (s & 7)) & 7));
        if (mode > 0 && bo == 0) continue;
        cfgs.push_back({s, sbo, bo});
      }
  const int ncfg = (int)cfgs.size();
  __nv_bfloat16 *dA, *dB;
  Cfg* dC;
  float* dO;
  cudaMalloc(&dA, hA.size() * 2);
  cudaMalloc(&dB, hB.size() * 2);
  cudaMalloc(&dC, ncfg * sizeof(Cfg));
  cudaMalloc(&dO, (size_t)ncfg * 128 * 16 * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dC, cfgs.data(), ncfg * sizeof(Cfg), cudaMemcpyHostToDevice);
  CUtensorMap tmA, tmB;
  if (make_tmap_2d_bf16(&tmA, dA, ROWS, 64, 64, 256, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return 2;
  if (make_tmap_2d_bf16(&tmB, dB, 16, 64, 64, 16, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return 2;
  const int smem = ROWS * 128 + 2048 + 1024 + 256;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe_kernel<<<1, 128, smem>>>(tmA, tmB, dC, ncfg, dO);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
  std::vector<float> hO((size_t)ncfg * 128 * 16);
  cudaMemcpy(hO.data(), dO, hO.size() * 4, cudaMemcpyDeviceToHost);
  for (int c = 0; c < ncfg; ++c) {
    const Cfg cf = cfgs[c];
    int bad = 0;
    for (int m = 0; m < 128; ++m) {
      const int row = cf.shift + (m / 8) * (cf.sbo / 128) + (m % 8);
      for (int n = 0; n < 16; ++n) {
        float want = 0.f;
        for (int k = 0; k < 64; ++k) want += fA[row * 64 + k] * fB[n * 64 + k];
        if (want != hO[((size_t)c * 128 + m) * 16 + n]) ++bad;
      }
    }
    printf("shift %d sbo %4d base_offset %d : %s (%d/2048 mismatches)\n", cf.shift, cf.sbo, cf.bo, bad ? "WRONG" : "ok", bad);
  }
  return 0;
}
