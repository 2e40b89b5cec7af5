// Probe (test tool, not product): tcgen05 MN-major shared-memory operands under the 128-byte swizzle.
// A tile that TMA wrote as rows of 128 B (64 bf16) — the K-major layout of one GEMM — is read by ANOTHER MMA as its MN-major
// operand (the row index becomes K, the 64 contiguous elements become M or N), so a backward kernel can contract over the row
// dimension of a tile it already holds without a transposed copy.  Canonical MN-major SW128 layout (cute make_umma_desc<MN>):
// ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units: 64 MN elements contiguous, LBO = byte stride between 64-element MN blocks,
// SBO = byte stride between groups of 8 K rows.  Cases:
//   0: D[m,n] = sum_k A[m,k]  B[k,n]   A K-major,  B MN-major (N = 128: two 64-wide blocks)
//   1: D[m,n] = sum_k A[k,m]  B[k,n]   A MN-major, B MN-major
//   2: D[m,n] = sum_k A[m,k]  B[k,n]   A K-major,  B MN-major with N = 256 (four blocks)
//   3: D[m,n] = sum_k A[k,m]  B[n,k]   A MN-major, B K-major
//   4: D[m,n] = sum_{k<32} A[m,32+k] B[n,32+k]   both K-major, start address +64 B inside the 128-byte row (the second of two
//      32-wide heads packed in one swizzled row), K = 32
//   5: D[m,n] = sum_k A[m,k] B[k,32+n], n < 32   B MN-major with N = 32 (half a 64-element block) at +64 B
//   6: D[m,n] = sum_k A[k,m] B[k,32+n], n < 32   A MN-major, B MN-major N = 32 at +64 B
//   7: D[m,n] = sum_{k<112} A[m,k] B[k+1,n]     B MN-major with the start address shifted by ONE K row (+128 B): a window that
//      slides over the rows of a tile (the implicit conv3x3 weight gradient reads its 3 horizontal taps this way)
//   8: the same with a shift of 2 rows
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../include -o probe_mn_major.bin probe_mn_major.cu ../objcavit_b200/csrc/tc_gemm.cu ../objcavit_b200/csrc/api.cu
#include "../objcavit_b200/csrc/tc_common.cuh"
#include <vector>
#include <cstdlib>

using namespace tc;

constexpr int SLAB = 128 * 128;   // [128 rows x 64 bf16] = 16 KB

__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr, uint32_t lbo, uint32_t sbo) { return make_smem_desc(saddr, lbo, sbo, LAYOUT_SW128); }

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmX, float* out /*[9][128][256]*/) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* sP = smem;               // P [128 x 128]: 2 slabs (column halves)
  uint8_t* sX = smem + 2 * SLAB;    // X [128 x 256]: 4 slabs (column quarters)
  uint64_t* bar_load = reinterpret_cast<uint64_t*>(sX + 4 * SLAB);
  uint64_t* bar_mma = bar_load + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (threadIdx.x == 0) {
    mbar_init(bar_load, 1);
    mbar_init(bar_mma, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar_load, 6 * SLAB);
    for (int s = 0; s < 2; ++s) tma_load_2d(sP + s * SLAB, &tmP, bar_load, s * 64, 0);
    for (int s = 0; s < 4; ++s) tma_load_2d(sX + s * SLAB, &tmX, bar_load, s * 64, 0);
  }
  mbar_wait(bar_load, 0);
  __syncthreads();
  for (int c = 0; c < 9; ++c) {
    if (threadIdx.x == 0) {
      tc_fence_after();
      const uint32_t p0 = smem_u32(sP), x0 = smem_u32(sX);
      for (int kk = 0; kk < 8; ++kk) {   // K = 128 in steps of 16
        const uint64_t a_k = desc_kmajor_sw128(p0 + (kk / 4) * SLAB + (kk % 4) * 32);     // K-major: column block kk/4, 32 B per step
        const uint64_t a_mn = desc_mn(p0 + kk * 2048, SLAB, 1024);                         // MN-major: 16 rows per step
        const uint64_t b_mn = desc_mn(x0 + kk * 2048, SLAB, 1024);
        const uint64_t b_k = desc_kmajor_sw128(x0 + (kk / 4) * SLAB + (kk % 4) * 32);
        if (c == 0) umma_bf16(tmem_d, a_k, b_mn, make_idesc_bf16(128, 128, 0, 1), kk != 0);
        if (c == 1) umma_bf16(tmem_d, a_mn, b_mn, make_idesc_bf16(128, 128, 1, 1), kk != 0);
        if (c == 2) umma_bf16(tmem_d, a_k, b_mn, make_idesc_bf16(128, 256, 0, 1), kk != 0);
        if (c == 3) umma_bf16(tmem_d, a_mn, b_k, make_idesc_bf16(128, 128, 1, 0), kk != 0);
        if (c == 4 && kk < 2) umma_bf16(tmem_d, desc_kmajor_sw128(p0 + 64 + kk * 32), desc_kmajor_sw128(x0 + 64 + kk * 32),
                                        make_idesc_bf16(128, 128, 0, 0), kk != 0);
        if (c == 5) umma_bf16(tmem_d, a_k, desc_mn(x0 + 64 + kk * 2048, SLAB, 1024), make_idesc_bf16(128, 32, 0, 1), kk != 0);
        if (c == 6) umma_bf16(tmem_d, a_mn, desc_mn(x0 + 64 + kk * 2048, SLAB, 1024), make_idesc_bf16(128, 32, 1, 1), kk != 0);
        if (c >= 7 && kk < 7) umma_bf16(tmem_d, a_k, desc_mn(x0 + (c - 6) * 128 + kk * 2048, SLAB, 1024), make_idesc_bf16(128, 128, 0, 1), kk != 0);
      }
      umma_commit(bar_mma);
    }
    mbar_wait(bar_mma, c & 1);
    tc_fence_after();
    for (int cb = 0; cb < 8; ++cb) {
      uint32_t r[32];
      tmem_ld32(tmem_d + ((uint32_t)(warp * 32) << 16) + cb * 32, r);
      tmem_ld_wait();
      float* o = out + ((long long)c * 128 + warp * 32 + lane) * 256 + cb * 32;
      for (int j = 0; j < 32; ++j) o[j] = __uint_as_float(r[j]);
    }
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) tmem_dealloc(tmem_d, 256);
}

int main() {
  std::vector<__nv_bfloat16> hP(128 * 128), hX(128 * 256);
  std::vector<float> fP(128 * 128), fX(128 * 256);
  srand(7);
  for (int i = 0; i < 128 * 128; ++i) { fP[i] = (float)((rand() % 9) - 4); hP[i] = __float2bfloat16(fP[i]); }
  for (int i = 0; i < 128 * 256; ++i) { fX[i] = (float)((rand() % 7) - 3); hX[i] = __float2bfloat16(fX[i]); }
  __nv_bfloat16 *dP, *dX;
  float* dO;
  cudaMalloc(&dP, hP.size() * 2);
  cudaMalloc(&dX, hX.size() * 2);
  cudaMalloc(&dO, 9 * 128 * 256 * 4);
  cudaMemset(dO, 0, 9 * 128 * 256 * 4);
  cudaMemcpy(dP, hP.data(), hP.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dX, hX.data(), hX.size() * 2, cudaMemcpyHostToDevice);
  CUtensorMap tmP, tmX;
  if (make_tmap_2d_bf16(&tmP, dP, 128, 128, 128, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return 2;
  if (make_tmap_2d_bf16(&tmX, dX, 128, 256, 256, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return 2;
  const int smem = 6 * SLAB + 1024 + 256;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe_kernel<<<1, 128, smem>>>(tmP, tmX, dO);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
  std::vector<float> hO(9 * 128 * 256);
  cudaMemcpy(hO.data(), dO, hO.size() * 4, cudaMemcpyDeviceToHost);
  const char* names[9] = {"A K-major,  B MN-major, N=128", "A MN-major, B MN-major, N=128", "A K-major,  B MN-major, N=256",
                          "A MN-major, B K-major,  N=128", "K-major +64 B, K=32", "B MN-major N=32 +64 B", "A MN, B MN N=32 +64 B",
                          "B MN-major, K rows shifted by 1", "B MN-major, K rows shifted by 2"};
  int rc = 0;
  for (int c = 0; c < 9; ++c) {
    const int N = c == 2 ? 256 : ((c == 5 || c == 6) ? 32 : 128);
    int bad = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < N; ++n) {
        float want = 0.f;
        for (int k = 0; k < (c == 4 ? 32 : (c >= 7 ? 112 : 128)); ++k) {
          float a = (c == 0 || c == 2 || c == 5) ? fP[m * 128 + k] : fP[k * 128 + m];
          float b = c == 3 ? fX[n * 256 + k] : fX[k * 256 + n];
          if (c == 4) { a = fP[m * 128 + 32 + k]; b = fX[n * 256 + 32 + k]; }
          if (c == 5 || c == 6) b = fX[k * 256 + 32 + n];
          if (c >= 7) { a = fP[m * 128 + k]; b = fX[(k + c - 6) * 256 + n]; }
          want += a * b;
        }
        if (want != hO[((size_t)c * 128 + m) * 256 + n]) ++bad;
      }
    printf("case %d (%s): %s (%d/%d mismatches)\n", c, names[c], bad ? "WRONG" : "ok", bad, 128 * N);
    if (bad) rc = 1;
  }
  return rc;
}
