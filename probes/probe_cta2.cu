// Probe (test tool, not product): tcgen05.mma.cta_group::2 mechanics + the shared-memory cost of SS MMAs.
//   part 1 (correctness): D[256,128] = A[256,64] B[128,64]^T on a CTA pair: each CTA TMA-loads ITS 128 rows of A and ITS 64 rows
//           of B (cp.async.bulk.tensor ... cta_group::2, completion bytes counted on the leader's mbarrier), the leader issues
//           4 x tcgen05.mma.cta_group::2 (M256 N128 K16), tcgen05.commit ... multicast::cluster arrives on both CTAs' barriers,
//           each CTA reads its 128 TMEM lanes.
//   part 2 (throughput): clk per MMA for back-to-back MMAs on resident operands, per variant
//           v1: cta_group::1 M128 N128    v2: cta_group::2 M256 N128    v3: cta_group::1 M128 N256    v4: cta_group::2 M256 N256
//           each alone and next to a background stream of TMA fills (bg = slabs of 16 KB in flight per CTA), reporting the
//           fill rate the background reached: tells whether operand reads + fills share one ~128 B/clk port.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../include -o probe_cta2.bin probe_cta2.cu ../objcavit_b200/csrc/tc_gemm.cu ../objcavit_b200/csrc/api.cu
#include "../objcavit_b200/csrc/tc_common.cuh"
#include <vector>
#include <cstdlib>

using namespace tc;

constexpr int SLAB = 128 * 128;  // [128 rows x 64 bf16] = 16 KB
constexpr int THREADS = 192;     // warp 0: TMA, warp 1: MMA, warps 2-5: TMEM readers (warp % 4 = lane quarter)

__device__ __forceinline__ uint32_t cta_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t base, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(ncols) : "memory");
}
// TMA load into THIS CTA's shared memory, completion bytes on the LEADER CTA's barrier (same offset, peer bit cleared)
__device__ __forceinline__ void tma_load_2d_cg2(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_cg2(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_cg2(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"(mask)
               : "memory");
}

struct Stats { long long clk; long long mmas; long long bg_bytes; };

// variant: 1,2,3,4 as in the header; iters = MMA groups of 4 (K = 64); bg = background fills in flight (0 = none)
template <int NACC, int RUN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, float* out, int variant, int iters,
             int bg, Stats* stats) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint8_t* sA = smem;                // 16 KB
  uint8_t* sB = smem + SLAB;         // up to 32 KB (N = 256 single CTA)
  uint8_t* sBG = smem + 3 * SLAB;    // 8 background slots of 16 KB
  uint64_t* bar_load = reinterpret_cast<uint64_t*>(sBG + 8 * SLAB);
  uint64_t* bar_mma = bar_load + 1;
  uint64_t* bar_bg = bar_mma + 1;    // [8]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_bg + 8);
  volatile int* stop = reinterpret_cast<volatile int*>(tmem_slot + 1);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const uint32_t rank = cta_rank();
  const bool pair = variant == 0 || variant == 2 || variant == 4;
  const int Nn = (variant == 3 || variant == 4) ? 256 : 128;
  const int b_rows = pair ? Nn / 2 : Nn;   // rows of B this CTA holds
  if (threadIdx.x == 0) {
    mbar_init(bar_load, 1);
    mbar_init(bar_mma, 1);
    for (int i = 0; i < 8; ++i) mbar_init(&bar_bg[i], 1);
    *stop = 0;
    fence_barrier_init();
  }
  if (warp == 2) {
    if (pair) tmem_alloc2(tmem_slot, 512);
    else tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;

  if (warp == 0 && elect_one()) {
    const uint32_t bytes_own = SLAB + b_rows * 128;
    if (pair) {
      if (rank == 0) mbar_arrive_expect_tx(bar_load, 2 * bytes_own);
      tma_load_2d_cg2(sA, &tmA, bar_load, 0, rank * 128);
      // B rows of this CTA: rank * b_rows .. ; the B map's box has 64 rows
      for (int r0 = 0; r0 < b_rows; r0 += 64) tma_load_2d_cg2(sB + r0 * 128, &tmB, bar_load, 0, rank * b_rows + r0);
    } else {
      mbar_arrive_expect_tx(bar_load, bytes_own);
      tma_load_2d(sA, &tmA, bar_load, 0, rank * 128);
      for (int r0 = 0; r0 < b_rows; r0 += 64) tma_load_2d(sB + r0 * 128, &tmB, bar_load, 0, r0);
    }
    // background fills: keep `bg` slab loads in flight until the MMA thread says stop
    if (bg > 0) {
      long long n = 0;
      uint32_t ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int i = 0; i < bg; ++i) {
        mbar_arrive_expect_tx(&bar_bg[i], SLAB);
        tma_load_2d(sBG + i * SLAB, &tmA, &bar_bg[i], 0, (i & 1) * 128);
      }
      int i = 0;
      while (!*stop) {
        mbar_wait(&bar_bg[i], ph[i]);
        ph[i] ^= 1;
        ++n;
        mbar_arrive_expect_tx(&bar_bg[i], SLAB);
        tma_load_2d(sBG + i * SLAB, &tmA, &bar_bg[i], 0, (i & 1) * 128);
        i = (i + 1) % bg;
      }
      for (int j = 0; j < bg; ++j) mbar_wait(&bar_bg[j], ph[j]);
      if (stats) stats[blockIdx.x].bg_bytes = n * SLAB;
    }
  }
  if (warp == 1 && elect_one()) {
    if (!pair || rank == 0) {
      mbar_wait(bar_load, 0);
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
      const uint32_t idesc = make_idesc_bf16(pair ? 256 : 128, Nn);
      const long long t0 = clock64();
      // accumulator of MMA number n = 4 it + kk: (n / RUN) % NACC; descriptors precomputed, body fully unrolled, so the issuing
      // thread spends a handful of instructions per MMA
      uint64_t da[4], db[4];
      for (int kk = 0; kk < 4; ++kk) { da[kk] = desc_kmajor_sw128(a0 + kk * 32); db[kk] = desc_kmajor_sw128(b0 + kk * 32); }
      for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const int acc = RUN == 1 ? (kk % NACC) : (it & (NACC - 1));
          const bool first = RUN == 1 ? (it == 0 && kk < NACC) : (it < NACC && kk == 0);
          const uint32_t d = tmem_d + (uint32_t)(acc * Nn);
          if (pair) umma_bf16_cg2(d, da[kk], db[kk], idesc, !first);
          else umma_bf16(d, da[kk], db[kk], idesc, !first);
        }
      }
      if (pair) umma_commit_cg2(bar_mma, 3);
      else umma_commit(bar_mma);
      mbar_wait(bar_mma, 0);
      const long long t1 = clock64();
      if (stats) { stats[blockIdx.x].clk = t1 - t0; stats[blockIdx.x].mmas = 4LL * iters; }
    }
  }
  if (warp >= 2) {
    mbar_wait(bar_mma, 0);   // pair: the multicast commit arrives here in both CTAs
    tc_fence_after();
    *stop = 1;               // ends this CTA's background loop
  }
  __syncthreads();
  if (warp >= 2 && out && iters == 1) {
    const int q = warp % 4;
    for (int cb = 0; cb < Nn / 32; ++cb) {
      uint32_t r[32];
      tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + cb * 32, r);
      tmem_ld_wait();
      float* o = out + ((long long)(blockIdx.x * 128 + q * 32 + lane)) * 256 + cb * 32;
      for (int j = 0; j < 32; ++j) o[j] = __uint_as_float(r[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 2) {
    if (pair) tmem_dealloc2(tmem_d, 512);
    else tmem_dealloc(tmem_d, 512);
  }
}

int main() {
  // A [256 x 64], B [256 x 64] (N up to 256)
  std::vector<__nv_bfloat16> hA(256 * 64), hB(256 * 64);
  std::vector<float> fA(256 * 64), fB(256 * 64);
  srand(11);
  for (int i = 0; i < 256 * 64; ++i) { fA[i] = (float)((rand() % 9) - 4); hA[i] = __float2bfloat16(fA[i]); }
  for (int i = 0; i < 256 * 64; ++i) { fB[i] = (float)((rand() % 7) - 3); hB[i] = __float2bfloat16(fB[i]); }
  __nv_bfloat16 *dA, *dB;
  float* dO;
  Stats* dS;
  cudaMalloc(&dA, hA.size() * 2);
  cudaMalloc(&dB, hB.size() * 2);
  cudaMalloc(&dO, 296 * 128 * 256 * 4);
  cudaMalloc(&dS, 296 * sizeof(Stats));
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  CUtensorMap tmA, tmB;
  if (make_tmap_2d_bf16(&tmA, dA, 256, 64, 64, 128, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return 2;
  if (make_tmap_2d_bf16(&tmB, dB, 256, 64, 64, 64, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return 2;
  const int smem = 11 * SLAB + 1024 + 256;
  typedef void (*KernelFn)(const CUtensorMap, const CUtensorMap, float*, int, int, int, Stats*);
  KernelFn fns[3][2] = {{probe_kernel<1, 1>, probe_kernel<1, 4>}, {probe_kernel<2, 1>, probe_kernel<2, 4>}, {probe_kernel<4, 1>, probe_kernel<4, 4>}};
  for (auto& row : fns) for (auto f : row) cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  int rc = 0;
  // ---- part 1: correctness of every variant (iters = 1: K = 64)
  for (int v = 0; v <= 4; ++v) {
    const bool pair = v == 0 || v == 2 || v == 4;
    const int Nn = (v == 3 || v == 4) ? 256 : 128;
    cudaMemset(dO, 0, 2 * 128 * 256 * 4);
    probe_kernel<1, 4><<<2, THREADS, smem>>>(tmA, tmB, dO, v, 1, 0, nullptr);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("variant %d: CUDA error: %s\n", v, cudaGetErrorString(e)); return 1; }
    std::vector<float> hO(2 * 128 * 256);
    cudaMemcpy(hO.data(), dO, hO.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int m = 0; m < 256; ++m)
      for (int n = 0; n < Nn; ++n) {
        float want = 0.f;
        for (int k = 0; k < 64; ++k) want += fA[m * 64 + k] * fB[n * 64 + k];
        if (want != hO[(size_t)m * 256 + n]) ++bad;
      }
    printf("variant %d (%s M%d N%d): %s (%d/%d mismatches)\n", v, pair ? "cta_group::2" : "cta_group::1", pair ? 256 : 128, Nn,
           bad ? "WRONG" : "ok", bad, 256 * Nn);
    if (bad) rc = 1;
  }
  // ---- part 2: clk per MMA on a full machine (148 CTAs): accumulators used round-robin (nacc) with `run` consecutive MMAs on
  // each, without / with 4 background TMA fills in flight per CTA
  const int iters = 4096;
  const int grid = 148;
  for (int v = 1; v <= 4; ++v) {
    const bool pair = v == 2 || v == 4;
    const int Nn = (v == 3 || v == 4) ? 256 : 128;
    for (int nacc = 1; nacc * Nn <= 512; nacc *= 2)
      for (int run : {1, 4})
        for (int bg : {0, 4}) {
          if (nacc == 1 && run == 1) continue;
          cudaMemset(dS, 0, 296 * sizeof(Stats));
          fns[nacc == 1 ? 0 : nacc == 2 ? 1 : 2][run == 1 ? 0 : 1]<<<grid, THREADS, smem>>>(tmA, tmB, nullptr, v, iters, bg, dS);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("timing v%d bg%d: CUDA error: %s\n", v, bg, cudaGetErrorString(e)); return 1; }
          std::vector<Stats> hS(grid);
          cudaMemcpy(hS.data(), dS, grid * sizeof(Stats), cudaMemcpyDeviceToHost);
          double clk = 0, fill = 0;
          int n = 0;
          for (int c = 0; c < grid; ++c)
            if (hS[c].mmas) { clk += (double)hS[c].clk / hS[c].mmas; ++n; }
          clk /= n;
          for (int c = 0; c < grid; ++c) fill += (double)hS[c].bg_bytes;
          const double window = clk * 4.0 * iters;
          const double flop_per_sm = 2.0 * 128 * Nn * 16;                // per SM per MMA (a pair MMA is 2 SMs' worth)
          const double opnd = (128 * 32 + (pair ? Nn / 2 : Nn) * 32);    // operand bytes read from THIS SM's shared memory per MMA
          printf("v%d %s N%d acc %d run %d bg %d: %6.1f clk/MMA = %4.0f FLOP/clk/SM (%.2f of 8192), operands %3.0f B/clk/SM, fills %4.1f B/clk/SM\n", v,
                 pair ? "cg2 M256" : "cg1 M128", Nn, nacc, run, bg, clk, flop_per_sm / clk, flop_per_sm / clk / 8192.0, opnd / clk,
                 fill / grid / window);
        }
  }
  return rc;
}
