// Patch embedding on the bf16 path: the PxP / stride-P conv of PatchTransformerEncoder / ObjCAViT as a tcgen05 GEMM
// whose A operand is read straight out of the NHWC bf16 feature map by TMA — non-overlapping patches are a pure
// re-indexing, so no im2col buffer exists:
//   5-D tensor map over X: (kx*C + c [P*C contiguous], tx [gw], ky [P], ty [gh], b [B]); one box = (64 k, gw, 1, TYB, 1)
//   = the 64-wide K slice of TYB full rows of patches = up to 128 GEMM rows, landing in smem in the K-major
//   128B-swizzled layout tcgen05.mma consumes.  K order is (ky, kx, c); the weights are re-packed once to match.
// Split-K over ky (K = P*P*C = 32768 is long, M = B*S is short); fp32 partials are reduced in a fixed order
// (deterministic) together with the bias and positional rows, producing fp32 tokens and their bf16 copy.
// replaces: embedding_convPxP + positional add + permute (modules/layers.py:17-22); image_embedding_convPxP and the
//   image positional add (modules/ObjCAViT.py:333,362-364).
#include "tc_common.cuh"

namespace {

using namespace tc;

constexpr int PE_E = 128, PE_GK = 64, PE_STAGES = 4;
constexpr int PE_A_BYTES = 128 * PE_GK * 2, PE_B_BYTES = PE_E * PE_GK * 2, PE_STAGE = 2 * PE_A_BYTES + PE_B_BYTES;
constexpr int PE_SMEM = PE_STAGES * PE_STAGE + 1024 + 256;

struct PeParams {
  float* part;  // [splits][B*S][128]
  int B, gh, gw, S, tyb, tiles_per_img, n_boxes, PC /*P*C*/, P, merged, kit_total, kit_per_split;
};

__device__ __forceinline__ void tma_load_5d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                            int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

// One CTA = TWO boxes of up to 128 GEMM rows (two TMEM accumulators) x one contiguous range of k-iterations (split-K).
// The weight slab of a k-iteration is loaded once and used by both boxes: the kernel is bound by the L2->SM stream
// (A from HBM once, weights from L2 once per CTA), so halving the weight re-reads is what matters.
// Box t covers patch rows [t*tyb, (t+1)*tyb): of the (image, patch-row) axis flattened when h == gh*P ("merged": the
// image stride is then a multiple of the patch-row stride and boxes run across image boundaries without waste), else of
// one image.
__global__ void __launch_bounds__(512, 1)
patch_embed_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, PeParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024 - (base_u32 & 1023)) & 1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + PE_STAGES * PE_STAGE);
  uint64_t* empty = full + PE_STAGES;
  uint64_t* tmem_full = empty + PE_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  pdl_launch();
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int box0 = blockIdx.x * 2;
  const bool has2 = box0 + 1 < p.n_boxes;
  const int kit0 = blockIdx.y * p.kit_per_split;
  const int n_k = max(0, min(p.kit_total, kit0 + p.kit_per_split) - kit0);
  const int kin = p.PC / PE_GK;                  // k-iterations per ky
  const int rows = p.tyb * p.gw;                 // GEMM rows one box delivers (<= 128)
  // TMA coordinates (patch-row, image) of a box
  auto box_coord = [&](int t, int& c3, int& c4) {
    if (p.merged) { c3 = t * p.tyb; c4 = 0; }
    else { c4 = t / p.tiles_per_img; c3 = (t - c4 * p.tiles_per_img) * p.tyb; }
  };
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < PE_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 2 * PE_E);
  // rows the box does not cover are read by the MMA as stale smem: harmless (rows are independent and those outputs
  // are never stored) but keep them finite.  Rows are contiguous 128-byte ranges under the 128B swizzle.
  if (rows < 128) {
    const int tail16 = (128 - rows) * 8;  // uint4 per A buffer
    for (int i = threadIdx.x; i < PE_STAGES * 2 * tail16; i += blockDim.x) {
      const int buf = i / tail16, j = i - buf * tail16;
      reinterpret_cast<uint4*>(smem + (buf >> 1) * PE_STAGE + (buf & 1) * PE_A_BYTES + rows * 128)[j] = make_uint4(0, 0, 0, 0);
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;
  if (warp == 0) {
    if (elect_one()) {
      int c3a, c4a, c3b = 0, c4b = 0;
      box_coord(box0, c3a, c4a);
      if (has2) box_coord(box0 + 1, c3b, c4b);
      const uint32_t tx = (has2 ? 2 : 1) * rows * PE_GK * 2 + PE_B_BYTES;
      for (int it = 0; it < n_k; ++it) {
        const int s = it % PE_STAGES;
        mbar_wait(&empty[s], ((it / PE_STAGES) & 1) ^ 1);
        const int kit = kit0 + it;
        const int ky = kit / kin, kc = (kit - ky * kin) * PE_GK;
        mbar_arrive_expect_tx(&full[s], tx);
        uint8_t* a = smem + s * PE_STAGE;
        tma_load_5d(a, &tmX, &full[s], kc, 0, ky, c3a, c4a);
        if (has2) tma_load_5d(a + PE_A_BYTES, &tmX, &full[s], kc, 0, ky, c3b, c4b);
        tma_load_2d(a + 2 * PE_A_BYTES, &tmW, &full[s], ky * p.PC + kc, 0);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, PE_E);
      for (int it = 0; it < n_k; ++it) {
        const int s = it % PE_STAGES;
        mbar_wait(&full[s], (it / PE_STAGES) & 1);
        tc_fence_after();
        const uint32_t a = smem_u32(smem + s * PE_STAGE);
        const uint64_t da0 = desc_kmajor_sw128(a), da1 = desc_kmajor_sw128(a + PE_A_BYTES);
        const uint64_t db = desc_kmajor_sw128(a + 2 * PE_A_BYTES);
#pragma unroll
        for (int kk = 0; kk < PE_GK / 16; ++kk)
          umma_bf16(tmem_d, da0 + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), idesc, (it | kk) != 0);
        if (has2) {
#pragma unroll
          for (int kk = 0; kk < PE_GK / 16; ++kk)
            umma_bf16(tmem_d + PE_E, da1 + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), idesc, (it | kk) != 0);
        }
        umma_commit(&empty[s]);
      }
      umma_commit(tmem_full);
    }
  }
  __syncwarp();
  pdl_wait();  // the partial planes may still be read by the previous step's reduce kernel
  // epilogue, all 16 warps: warp -> TMEM lane quarter (warp % 4) x 32-column group (warp / 4), both accumulators
  {
    const int quarter = warp & 3, cg = warp >> 2;
    const int r = quarter * 32 + lane;
    if (n_k > 0) {
      mbar_wait(tmem_full, 0);
      tc_fence_after();
    }
    const int tyl = r / p.gw, tx = r - tyl * p.gw;
    for (int j = 0; j < (has2 ? 2 : 1); ++j) {
      int c3, c4;
      box_coord(box0 + j, c3, c4);
      const int pr = c3 + tyl;  // patch row (flattened over images when merged)
      bool ok = r < rows;
      long long m;
      if (p.merged) {
        ok = ok && pr < p.gh * p.B;
        m = (long long)pr * p.gw + tx;   // == b*S + ty*gw + tx
      } else {
        ok = ok && pr < p.gh;
        m = (long long)c4 * p.S + pr * p.gw + tx;
      }
      uint32_t v[32];
      if (n_k > 0) {
        tmem_ld32(tmem_d + ((uint32_t)(quarter * 32) << 16) + j * PE_E + cg * 32, v);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int q = 0; q < 32; ++q) v[q] = 0;
      }
      if (ok) {
        float* dst = p.part + ((long long)blockIdx.y * p.B * p.S + m) * PE_E + cg * 32;
#pragma unroll
        for (int q = 0; q < 32; q += 4)
          *reinterpret_cast<float4*>(dst + q) = make_float4(__uint_as_float(v[q]), __uint_as_float(v[q + 1]),
                                                            __uint_as_float(v[q + 2]), __uint_as_float(v[q + 3]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_d, 2 * PE_E);
}

// tokens[m,:] = sum_z part[z,m,:] + bias + pos[m % S,:]  ->  fp32 and bf16 copies
__global__ void patch_reduce_kernel(const float* __restrict__ part, const float* __restrict__ bias,
                                    const float* __restrict__ pos, float* __restrict__ tok32,
                                    __nv_bfloat16* __restrict__ tok16, int M, int S, int splits) {
  pdl_launch();
  pdl_wait();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // float4 index
  if (i >= (long long)M * (PE_E / 4)) return;
  const int m = (int)(i / (PE_E / 4)), n4 = (int)(i % (PE_E / 4));
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int z = 0; z < splits; ++z) {
    const float4 v = reinterpret_cast<const float4*>(part)[(long long)z * M * (PE_E / 4) + i];
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  if (bias) {
    const float4 v = reinterpret_cast<const float4*>(bias)[n4];
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  if (pos) {
    const float4 v = reinterpret_cast<const float4*>(pos)[(long long)(m % S) * (PE_E / 4) + n4];
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  if (tok32) reinterpret_cast<float4*>(tok32)[i] = acc;
  if (tok16) {
    __nv_bfloat162 h0 = __floats2bfloat162_rn(acc.x, acc.y), h1 = __floats2bfloat162_rn(acc.z, acc.w);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&h0);
    u.y = *reinterpret_cast<uint32_t*>(&h1);
    reinterpret_cast<uint2*>(tok16)[i] = u;
  }
}

// W [E][C][P][P] fp32 -> Wp [E][(ky, kx, c)] bf16
__global__ void pack_patch_weights_kernel(const float* __restrict__ W, __nv_bfloat16* __restrict__ out, int E, int C, int P) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long K = (long long)C * P * P;
  if (i >= E * K) return;
  const int e = (int)(i / K);
  const int r = (int)(i - e * K);
  const int ky = r / (P * C), kx = (r / C) % P, c = r % C;
  out[i] = __float2bfloat16(W[(((long long)e * C + c) * P + ky) * P + kx]);
}

struct PePlan {
  int gh, gw, tyb, tiles_per_img, n_boxes, merged, kit_total, splits, kit_per_split;
};

int pe_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

// One wave of CTAs: (pairs of boxes) x (k splits) <= number of SMs, every CTA the same number of k-iterations.
PePlan pe_plan(int B, int C, int h, int w, int P) {
  PePlan q{};
  q.gh = h / P; q.gw = w / P;
  q.merged = (h == q.gh * P) ? 1 : 0;  // boxes may then run across images
  const int pr_total = q.merged ? q.gh * B : q.gh;
  q.tyb = 128 / q.gw < pr_total ? 128 / q.gw : pr_total;
  q.tiles_per_img = ceil_div(q.gh, q.tyb);
  q.n_boxes = q.merged ? ceil_div(q.gh * B, q.tyb) : B * q.tiles_per_img;
  q.kit_total = P * (P * C / PE_GK);
  const int pairs = ceil_div(q.n_boxes, 2);
  int splits = pe_sm_count() / pairs;
  if (splits < 1) splits = 1;
  if (splits > 64) splits = 64;
  if (splits > q.kit_total) splits = q.kit_total;
  q.kit_per_split = ceil_div(q.kit_total, splits);
  q.splits = ceil_div(q.kit_total, q.kit_per_split);
  return q;
}

}  // namespace

// The 5-D view of the NHWC feature map the kernel's A operand is read through: (kx*C + c, tx, ky, ty, b); merged: the (ty, b) axes
// collapse into one axis of gh*B patch rows.  Also used by the head executor to re-point a CAPTURED launch at a new feature
// tensor (csrc/head_exec.cu: the lane graphs must not depend on the caller's input address).
int ocv_patch_embed_xmap(CUtensorMap* out, const void* X, int B, int C, int h, int w, int P) {
  tc::EncodeTiledFn enc = tc::get_encode_tiled();
  if (!enc) return OCV_E_DRIVER;
  if (reinterpret_cast<uintptr_t>(X) & 15) return OCV_E_ALIGN;
  const PePlan q = pe_plan(B, C, h, w, P);
  const int gh = h / P, gw = w / P;
  cuuint64_t dims[5] = {(cuuint64_t)P * C, (cuuint64_t)gw, (cuuint64_t)P, (cuuint64_t)(q.merged ? gh * B : gh),
                        (cuuint64_t)(q.merged ? 1 : B)};
  cuuint64_t strides[4] = {(cuuint64_t)P * C * 2, (cuuint64_t)w * C * 2, (cuuint64_t)P * w * C * 2,
                           (cuuint64_t)h * w * C * 2};
  cuuint32_t box[5] = {PE_GK, (cuuint32_t)gw, 1, (cuuint32_t)q.tyb, 1};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(X), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OCV_OK : OCV_E_DRIVER;
}

// true iff `func` (cudaKernelNodeParams::func of a captured launch) is the patch-embedding GEMM; its argument 0 is the feature map
bool ocv_patch_embed_is_kernel(const void* func) { return func == reinterpret_cast<const void*>(&patch_embed_kernel); }

extern "C" int ocv_pack_patch_weights_bf16(const float* W, void* Wp, int E, int C, int P, void* stream) {
  OCV_CHECK_ARG(W && Wp && E > 0 && C > 0 && P > 0);
  if (int s = ocv_require_sm100()) return s;
  const long long n = (long long)E * C * P * P;
  pack_patch_weights_kernel<<<(unsigned)ceil_div<long long>(n, 256), 256, 0, ocv_stream(stream)>>>(
      W, static_cast<__nv_bfloat16*>(Wp), E, C, P);
  return ocv_launch_status();
}

extern "C" long long ocv_patch_embed_bf16_workspace(int B, int h, int w, int P) {
  if (B <= 0 || P <= 0 || h < P || w < P) return 0;
  const long long S = (long long)(h / P) * (w / P);
  if (w / P > 128) return 0;
  const PePlan q = pe_plan(B, 4096 /*upper bound on the split count for any C*/, h, w, P);
  return (long long)q.splits * B * S * PE_E * (long long)sizeof(float);
}

extern "C" int ocv_patch_embed_bf16(const void* X, const void* Wp, const float* bias, const float* pos, float* tok32,
                                    void* tok16, int B, int C, int h, int w, int P, int E, void* workspace,
                                    void* stream) {
  OCV_CHECK_ARG(X && Wp && (tok32 || tok16) && workspace && B > 0);
  OCV_CHECK_SHAPE(E == PE_E && P > 0 && C > 0 && (P * C) % PE_GK == 0 && h >= P && w >= P);
  const int gh = h / P, gw = w / P;
  OCV_CHECK_SHAPE(gw <= 128 && B * ceil_div(gh, 1) <= (1 << 30));
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(X) & 15) == 0);
  if (int s = ocv_require_sm100()) return s;
  tc::EncodeTiledFn enc = tc::get_encode_tiled();
  if (!enc) return OCV_E_DRIVER;
  const PePlan q = pe_plan(B, C, h, w, P);
  CUtensorMap tmX, tmW;
  if (int s = ocv_patch_embed_xmap(&tmX, X, B, C, h, w, P)) return s;
  const int K = P * P * C;
  if (int s = tc::make_tmap_2d_bf16(&tmW, Wp, E, K, K, PE_E, PE_GK, CU_TENSOR_MAP_SWIZZLE_128B)) return s;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(patch_embed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PE_SMEM);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  PeParams p{};
  p.part = static_cast<float*>(workspace);
  p.B = B; p.gh = gh; p.gw = gw; p.S = gh * gw; p.tyb = q.tyb; p.tiles_per_img = q.tiles_per_img; p.n_boxes = q.n_boxes;
  p.PC = P * C; p.P = P; p.merged = q.merged; p.kit_total = q.kit_total; p.kit_per_split = q.kit_per_split;
  const int splits = q.splits;
  dim3 grid(ceil_div(q.n_boxes, 2), splits);
  if (int s = ocv_launch_pdl(patch_embed_kernel, grid, dim3(512), PE_SMEM, stream, tmX, tmW, p)) return s;
  const int M = B * p.S;
  const long long n4 = (long long)M * (PE_E / 4);
  return ocv_launch_pdl(patch_reduce_kernel, dim3((unsigned)ceil_div<long long>(n4, 256)), dim3(256), 0, stream,
                        (const float*)p.part, bias, pos, tok32, static_cast<__nv_bfloat16*>(tok16), M, p.S, splits);
}
