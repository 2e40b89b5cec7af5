// Native host-side sequencing of the whole depth head on the bf16 tensor-core path: ONE C call issues every kernel of
//   ObjCAViT.forward (modules/ObjCAViT.py:306-390) | mViT.forward (modules/miniViT.py:22-44)
//   + the head tail (modules/GraphBins.py:107-119 | modules/AdaBins.py:76-87)
// into caller-provided workspace, with the object branch (positional MLP, Linear 512->128, object encoder) on a side
// stream next to the image branch (patch embedding, image encoder) and an event join before the cross-attention.
// Everything is launch-only (no allocation, no synchronisation), so a caller may capture the call in a CUDA graph.
// Contract: include/objcavit_b200.h (ocv_head_*).
#include "tc_common.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>

namespace {

constexpr int E = 128, HEADS = 4, FFN = 1024, P = 16, QN = 128;
typedef __nv_bfloat16 bf16;

struct Param {
  const float* ptr;
  long long numel;
};

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

bool ends_with(const std::string& s, const char* suf) {
  const size_t n = strlen(suf);
  return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

// bump allocator over the caller's workspace
struct Arena {
  uint8_t* base;
  size_t off = 0;
  explicit Arena(void* b) : base(static_cast<uint8_t*>(b)) {}
  template <typename T>
  T* take(size_t n) {
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off = align_up(off + n * sizeof(T));
    return p;
  }
};

// scratch of one encoder / attention branch over `rows` token rows
struct Scratch {
  float* x32[3];
  bf16* x16[3];
  bf16 *qkv, *att, *h, *kk;
  void carve(Arena& a, size_t rows) {
    for (int i = 0; i < 3; ++i) {
      x32[i] = a.take<float>(rows * E);
      x16[i] = a.take<bf16>(rows * E);
    }
    qkv = a.take<bf16>(rows * 3 * E);
    att = a.take<bf16>(rows * E);
    h = a.take<bf16>(rows * FFN);
    kk = a.take<bf16>(rows * E);
  }
};

struct Workspace {
  Scratch img, obj;
  float *obj_pos, *obj_tok, *pad32, *obj_s32, *y_raw, *centres, *blog, *pe_ws, *pts, *reg_ws;
  float *in_feat, *in_xywh, *edges;  // pipelined mode: lane-static copies of the object inputs / the bin edges (graph replay)
  int32_t* in_off;
  bf16 *feat16, *pad16, *obj_s16, *Mb;
  uint8_t *mask, *kpm;
  int32_t* uni;
  size_t bytes;
};

}  // namespace

struct ocv_head {
  ocv_head_config cfg;
  std::unordered_map<std::string, Param> params;
  std::unordered_map<std::string, size_t> bf16_off;  // key -> byte offset of the bf16 copy in `packed`
  size_t off_patch = 0, off_w3p = 0, off_imgpos = 0, off_coords = 0, off_pts = 0, packed_bytes = 0;
  uint8_t* packed = nullptr;
  int imgpos_h = -1, imgpos_w = -1;
  std::string pre, missing;
  cudaStream_t side = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  int launches = 0;
  bool timing = false;
  cudaEvent_t tev[2] = {nullptr, nullptr};
  // pipelined mode (ocv_head_set_pipeline): consecutive calls rotate over `lanes` lanes, each with its own token stream, side
  // stream, events and workspace slice; the fused pixel kernels of all lanes run back to back on ONE pixel stream with a capped
  // grid, so that the latency-bound token chains of the following calls run next to it on the SMs it leaves free.
  struct Lane {
    cudaStream_t tok = nullptr, side = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t tok_done = nullptr, pix_done = nullptr;
    bool used = false;
    // the token path of a lane is replayed from a CUDA graph once the same call signature has been seen twice
    cudaGraphExec_t gexec = nullptr;
    // The captured launches bake in ONE caller address: the feature map (argument 0 of the patch-embedding GEMM, a TMA
    // descriptor).  A serving loop hands over a freshly allocated tensor every call, so the lane keeps the source graph alive,
    // remembers that kernel node and re-points it with cudaGraphExecKernelNodeSetParams when the address differs.
    cudaGraph_t graph = nullptr;
    cudaGraphNode_t pe_node = nullptr;
    cudaKernelNodeParams pe_params = {};
    const void* graph_x = nullptr;
    unsigned long long sig[8] = {0, 0, 0, 0, 0, 0, 0, 0}, cand[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int graph_launches = 0;
    bool graph_failed = false;
  } lane[4];
  bool use_graphs = true;
  int lanes = 0;  // 0 = classic: everything on the caller's stream
  int pixel_ctas = 0;
  cudaStream_t pix = nullptr;
  cudaEvent_t in_ready = nullptr;
  unsigned long long calls = 0;

  const float* f32(const std::string& key) {
    auto it = params.find(key);
    if (it == params.end()) {
      if (missing.empty()) missing = key;
      return nullptr;
    }
    return it->second.ptr;
  }
  const bf16* w16(const std::string& key) {
    auto it = bf16_off.find(key);
    if (it == bf16_off.end() || !packed) {
      if (missing.empty()) missing = key;
      return nullptr;
    }
    return reinterpret_cast<const bf16*>(packed + it->second);
  }
};

namespace {

// Streams of the pipelined mode are handed to the caller's caching allocator (torch: Tensor.record_stream), which may still
// record events on them after the head is gone: they are never destroyed, only returned to a process-wide pool.
std::mutex g_pool_mu;
std::vector<cudaStream_t> g_stream_pool;

cudaError_t acquire_stream(cudaStream_t* out) {
  {
    std::lock_guard<std::mutex> lk(g_pool_mu);
    if (!g_stream_pool.empty()) {
      *out = g_stream_pool.back();
      g_stream_pool.pop_back();
      return cudaSuccess;
    }
  }
  return cudaStreamCreateWithFlags(out, cudaStreamNonBlocking);
}

void release_stream(cudaStream_t s) {
  if (!s) return;
  std::lock_guard<std::mutex> lk(g_pool_mu);
  g_stream_pool.push_back(s);
}

constexpr int IMGPOS_ROWS = 4096;  // rows reserved for the cached image positional rows (S <= 1200 in every config)

bool needs_bf16(const std::string& k) {
  if (k.find("_encoder_layers.") != std::string::npos) return false;  // never-executed prototypes
  return ends_with(k, "in_proj_weight") || ends_with(k, "out_proj.weight") || ends_with(k, "linear1.weight") ||
         ends_with(k, "linear2.weight") || ends_with(k, "obj_embedding_layer.weight");
}

void plan_packed(ocv_head* H) {
  size_t off = 0;
  H->bf16_off.clear();
  std::vector<std::string> keys;
  for (auto& kv : H->params)
    if (needs_bf16(kv.first)) keys.push_back(kv.first);
  std::sort(keys.begin(), keys.end());
  for (auto& k : keys) {
    H->bf16_off[k] = off;
    off = align_up(off + (size_t)H->params[k].numel * 2);
  }
  H->off_patch = off;
  off = align_up(off + (size_t)E * P * P * E * 2);
  H->off_w3p = off;
  off = align_up(off + (size_t)9 * E * E * 2);
  H->off_imgpos = off;
  off = align_up(off + (size_t)IMGPOS_ROWS * E * 4);
  H->off_coords = off;
  off = align_up(off + (size_t)IMGPOS_ROWS * 4 * 4);
  H->off_pts = off;
  off = align_up(off + (size_t)IMGPOS_ROWS * 2 * 4);
  H->packed_bytes = off;
}

}  // namespace

extern "C" int ocv_head_create(const ocv_head_config* cfg, ocv_head** out) {
  OCV_CHECK_ARG(cfg && out);
  OCV_CHECK_ARG(cfg->kind == 0 || cfg->kind == 1);
  OCV_CHECK_ARG(cfg->strategy >= 0 && cfg->strategy <= 3);
  OCV_CHECK_SHAPE(cfg->n_bins == 256 && (cfg->kind == 0 || (cfg->obj_dim > 0 && cfg->obj_dim % 8 == 0)));
  ocv_head* H = new (std::nothrow) ocv_head();
  if (!H) return OCV_E_ARG;
  H->cfg = *cfg;
  H->pre = cfg->kind == 0 ? "adaptive_bins_layer." : "objcavit.";
  *out = H;
  return OCV_OK;
}

extern "C" void ocv_head_destroy(ocv_head* H) {
  if (!H) return;
  for (auto& e : H->ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : H->tev)
    if (e) cudaEventDestroy(e);
  if (H->side) cudaStreamDestroy(H->side);
  for (auto& l : H->lane) {
    for (auto& e : l.ev)
      if (e) cudaEventDestroy(e);
    if (l.gexec) cudaGraphExecDestroy(l.gexec);
    if (l.graph) cudaGraphDestroy(l.graph);
    if (l.tok_done) cudaEventDestroy(l.tok_done);
    if (l.pix_done) cudaEventDestroy(l.pix_done);
    release_stream(l.tok);
    release_stream(l.side);
  }
  if (H->in_ready) cudaEventDestroy(H->in_ready);
  release_stream(H->pix);
  delete H;
}

extern "C" int ocv_head_set_param(ocv_head* H, const char* key, const float* ptr, long long numel) {
  OCV_CHECK_ARG(H && key && ptr && numel > 0);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(ptr) & 15) == 0);
  H->params[key] = Param{ptr, numel};
  H->packed = nullptr;  // derived copies are stale
  for (auto& l : H->lane) {  // captured launch sequences hold the old pointers
    if (l.gexec) cudaGraphExecDestroy(l.gexec);
    l.gexec = nullptr;
    if (l.graph) cudaGraphDestroy(l.graph);
    l.graph = nullptr;
    l.pe_node = nullptr;
    l.cand[5] = 0;
    l.graph_failed = false;
  }
  H->imgpos_h = H->imgpos_w = -1;
  return OCV_OK;
}

extern "C" long long ocv_head_packed_bytes(const ocv_head* H) {
  if (!H) return OCV_E_ARG;
  plan_packed(const_cast<ocv_head*>(H));
  return (long long)H->packed_bytes;
}

extern "C" const char* ocv_head_missing_key(const ocv_head* H) { return H ? H->missing.c_str() : ""; }
extern "C" int ocv_head_last_launches(const ocv_head* H) { return H ? H->launches : 0; }

extern "C" int ocv_head_pack(ocv_head* H, void* packed, void* stream) {
  OCV_CHECK_ARG(H && packed);
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(packed) & 255) == 0);
  plan_packed(H);
  H->missing.clear();
  uint8_t* base = static_cast<uint8_t*>(packed);
  for (auto& kv : H->bf16_off) {
    const Param& p = H->params[kv.first];
    if (int s = ocv_cast_f32_to_bf16(p.ptr, base + kv.second, p.numel, stream)) return s;
  }
  const std::string pw = H->pre + (H->cfg.kind == 0 ? "patch_transformer.embedding_convPxP.weight" : "image_embedding_convPxP.weight");
  const float* wp = H->f32(pw);
  const float* w3 = H->f32(H->pre + "conv3x3.weight");
  if (!wp || !w3) return OCV_E_ARG;
  if (H->params[pw].numel != (long long)E * E * P * P) return OCV_E_SHAPE;
  if (int s = ocv_pack_patch_weights_bf16(wp, base + H->off_patch, E, E, P, stream)) return s;
  if (int s = ocv_pack_conv3x3_bf16(w3, base + H->off_w3p, stream)) return s;
  H->packed = base;
  H->imgpos_h = H->imgpos_w = -1;
  return OCV_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// workspace
// ------------------------------------------------------------------------------------------------------------------
namespace {

Workspace carve(const ocv_head* H, void* base, int B, int h, int w, int total, int n_max) {
  Workspace W{};
  Arena a(base);
  const size_t S = (size_t)(h / P) * (w / P), M = (size_t)B * S;
  W.img.carve(a, M);
  W.y_raw = a.take<float>((size_t)B * 256);
  W.centres = a.take<float>((size_t)B * 256);
  W.blog = a.take<float>((size_t)B * 256);
  W.Mb = a.take<bf16>((size_t)B * 256 * E);
  W.reg_ws = a.take<float>((size_t)2 * B * 256);
  W.pe_ws = a.take<float>((size_t)ocv_patch_embed_bf16_workspace(B, h, w, P) / 4 + 4);
  W.edges = a.take<float>((size_t)B * 257);
  if (H->cfg.kind == 1) {
    W.in_feat = a.take<float>((size_t)total * H->cfg.obj_dim);
    W.in_xywh = a.take<float>((size_t)total * 4);
    W.in_off = a.take<int32_t>((size_t)B + 1);
    const size_t Mo = (size_t)B * n_max;
    W.obj.carve(a, H->cfg.use_2_saca ? M : Mo);
    W.obj_pos = a.take<float>((size_t)total * E);
    W.obj_tok = a.take<float>((size_t)total * E);
    W.pts = a.take<float>((size_t)total * 2);
    W.feat16 = a.take<bf16>((size_t)total * H->cfg.obj_dim);
    W.pad32 = a.take<float>(Mo * E);
    W.pad16 = a.take<bf16>(Mo * E);
    W.mask = a.take<uint8_t>(Mo);
    W.obj_s32 = a.take<float>(M * E);
    W.obj_s16 = a.take<bf16>(M * E);
    W.kpm = a.take<uint8_t>(M);
    W.uni = a.take<int32_t>((size_t)B + 1);
  }
  W.bytes = a.off;
  return W;
}

__global__ void uniform_offsets_kernel(int32_t* o, int n, int step) {
  pdl_launch();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = i * step;
}

// ------------------------------------------------------------------------------------------------------------------
// building blocks (each returns an OCV status; `H->launches` counts kernels)
// ------------------------------------------------------------------------------------------------------------------
struct Ctx {
  ocv_head* H;
  int status = OCV_OK;
  bool ok() const { return status == OCV_OK; }
  void run(int s, int launches = 1) {
    if (status == OCV_OK) {
      status = s;
      H->launches += launches;
    }
  }
};

// nn.TransformerEncoderLayer (post-norm, ReLU) x4: x (index `cur` of the scratch ring) -> returns the index of the result
int encoder(Ctx& c, const std::string& pre, Scratch& s, int cur, int B, int S, const uint8_t* mask, void* st) {
  ocv_head* H = c.H;
  const int M = B * S;
  for (int l = 0; l < 4 && c.ok(); ++l) {
    const std::string lp = pre + "layers." + std::to_string(l) + ".";
    const int y = (cur + 1) % 3, z = (cur + 2) % 3;
    const bf16 *w_in = H->w16(lp + "self_attn.in_proj_weight"), *w_out = H->w16(lp + "self_attn.out_proj.weight");
    const bf16 *w_1 = H->w16(lp + "linear1.weight"), *w_2 = H->w16(lp + "linear2.weight");
    const float *b_in = H->f32(lp + "self_attn.in_proj_bias"), *b_out = H->f32(lp + "self_attn.out_proj.bias");
    const float *b_1 = H->f32(lp + "linear1.bias"), *b_2 = H->f32(lp + "linear2.bias");
    const float *g1 = H->f32(lp + "norm1.weight"), *e1 = H->f32(lp + "norm1.bias");
    const float *g2 = H->f32(lp + "norm2.weight"), *e2 = H->f32(lp + "norm2.bias");
    if (!w_in || !w_out || !w_1 || !w_2 || !b_in || !b_out || !b_1 || !b_2 || !g1 || !e1 || !g2 || !e2) {
      c.status = OCV_E_ARG;
      break;
    }
    c.run(ocv_gemm_bf16_ex(s.x16[cur], w_in, b_in, nullptr, 0, nullptr, nullptr, nullptr, nullptr, 0, s.qkv, 3 * E, M,
                           3 * E, E, E, E, OCV_ACT_NONE, st));
    c.run(ocv_mha_core_bf16(s.qkv, s.qkv + E, s.qkv + 2 * E, s.att, mask, B, S, S, HEADS, E / HEADS, (long long)S * 3 * E,
                            3 * E, (long long)S * 3 * E, 3 * E, (long long)S * 3 * E, 3 * E, (long long)S * E, E, st));
    c.run(ocv_gemm_bf16_ex(s.att, w_out, b_out, s.x32[cur], E, g1, e1, nullptr, s.x32[y], E, s.x16[y], E, M, E, E, E, E,
                           OCV_ACT_NONE, st));
    c.run(ocv_gemm_bf16_ex(s.x16[y], w_1, b_1, nullptr, 0, nullptr, nullptr, nullptr, nullptr, 0, s.h, FFN, M, FFN, E, E,
                           E, OCV_ACT_RELU, st));
    c.run(ocv_gemm_bf16_ex(s.h, w_2, b_2, s.x32[y], E, g2, e2, l == 3 ? mask : nullptr, s.x32[z], E, s.x16[z], E, M, E,
                           FFN, FFN, FFN, OCV_ACT_NONE, st));
    cur = z;
  }
  return cur;
}

// nn.MultiheadAttention(query = qv, key = k, value = qv) (both reference call sites, ObjCAViT.py:195-207) -> out32/out16.
// The K projection runs on `st_k` (the stream that produced k16); q/v projection, attention and out-proj on `st`.
void cross_attention(Ctx& c, const std::string& pre, Scratch& s, const bf16* qv16, const bf16* k16, const uint8_t* kpm,
                     float* out32, bf16* out16, int B, int S, void* st) {
  ocv_head* H = c.H;
  const int M = B * S;
  const bf16 *w = H->w16(pre + "in_proj_weight"), *wo = H->w16(pre + "out_proj.weight");
  const float *b = H->f32(pre + "in_proj_bias"), *bo = H->f32(pre + "out_proj.bias");
  if (!w || !wo || !b || !bo) {
    c.status = OCV_E_ARG;
    return;
  }
  c.run(ocv_gemm_bf16_ex(qv16, w, b, nullptr, 0, nullptr, nullptr, nullptr, nullptr, 0, s.qkv, 3 * E, M, 3 * E, E, E, E,
                         OCV_ACT_NONE, st));
  c.run(ocv_gemm_bf16_ex(k16, w + (size_t)E * E, b + E, nullptr, 0, nullptr, nullptr, nullptr, nullptr, 0, s.kk, E, M, E, E,
                         E, E, OCV_ACT_NONE, st));
  // patch x object cross attention: the tcgen05 attention core (north star); OCV_ATTN_TC=0 keeps the mma.sync kernel
  if (ocv_attn_tc_policy() != 0 &&
      ocv_mha_tc_supported(S, S, HEADS, E / HEADS, (long long)S * 3 * E, 3 * E, (long long)S * E, E, (long long)S * 3 * E, 3 * E,
                           (long long)S * E, E))
    c.run(ocv_mha_core_tc_bf16(s.qkv, s.kk, s.qkv + 2 * E, s.att, nullptr, nullptr, kpm, B, S, S, HEADS, E / HEADS,
                               (long long)S * 3 * E, 3 * E, (long long)S * E, E, (long long)S * 3 * E, 3 * E, (long long)S * E, E, 0.f,
                               0ull, st));
  else
  c.run(ocv_mha_core_bf16(s.qkv, s.kk, s.qkv + 2 * E, s.att, kpm, B, S, S, HEADS, E / HEADS, (long long)S * 3 * E, 3 * E,
                          (long long)S * E, E, (long long)S * 3 * E, 3 * E, (long long)S * E, E, st));
  c.run(ocv_gemm_bf16_ex(s.att, wo, bo, nullptr, 0, nullptr, nullptr, nullptr, out32, E, out16, E, M, E, E, E, E,
                         OCV_ACT_NONE, st));
}

// nn.Sequential of Linear/LeakyReLU ending in a bare Linear at module indices idx[0..n)
void mlp(Ctx& c, const std::string& pre, const int* idx, int n, const int* dims, const float* x, long long x_ld, int R,
         const float* residual, float* out32, void* st) {
  const float* Ws[8];
  const float* bs[8];
  for (int i = 0; i < n; ++i) {
    Ws[i] = c.H->f32(pre + std::to_string(idx[i]) + ".weight");
    bs[i] = c.H->f32(pre + std::to_string(idx[i]) + ".bias");
    if (!Ws[i] || !bs[i]) {
      c.status = OCV_E_ARG;
      return;
    }
  }
  c.run(ocv_mlp_f32(x, x_ld, R, n, Ws, bs, dims, residual, out32, nullptr, st));
}

}  // namespace

extern "C" long long ocv_head_workspace_bytes(const ocv_head* H, int B, int h, int w, int total_obj, int n_max) {
  if (!H || B <= 0 || h < P || w < P) return OCV_E_ARG;
  return (long long)carve(H, nullptr, B, h, w, total_obj, n_max).bytes * (H->lanes > 0 ? H->lanes : 1);
}

// ------------------------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------------------------
namespace {

const int MLP_POS_IDX[5] = {0, 2, 4, 6, 8};
const int MLP_REG_IDX[3] = {0, 2, 4};

// object-side (rows = boxes [N,4]) or image-side (rows = patch coords [S,4]) positional embeddings -> out [N,E]
void positional(Ctx& c, const float* rows, int N, bool image_side, int h, int w, float* pts_tmp, float* out, void* st) {
  ocv_head* H = c.H;
  const std::string pre = H->pre + "positional_encoder.";
  const int strat = H->cfg.strategy;
  if (strat <= 1) {  // learned / learned_bbox_wh (ObjCAViT.py:257-282): the MLP sees the first 2 | 4 columns
    const int nin = strat == 1 ? 4 : 2;
    const int dims[6] = {nin, 32, 64, 128, 256, E};
    mlp(c, pre, MLP_POS_IDX, 5, dims, rows, 4, N, nullptr, out, st);
    return;
  }
  const float* table = H->f32(pre + "positional_encodings");
  if (!table) {
    c.status = OCV_E_ARG;
    return;
  }
  const int gh = (h + P - 1) / P, gw = (w + P - 1) / P;  // GridRandom uses ceil (ObjCAViT.py:78-79)
  if ((long long)gh * gw > H->params[pre + "positional_encodings"].numel / E) {
    c.status = OCV_E_SHAPE;
    return;
  }
  if (strat == 2) {
    if (c.ok()) {
      cudaError_t e = cudaMemcpy2DAsync(pts_tmp, 8, rows, 16, 8, N, cudaMemcpyDeviceToDevice, ocv_stream(st));
      if (e != cudaSuccess) c.status = (int)e;
    }
    if (image_side) c.run(ocv_grid_sample_pos_f32(table, pts_tmp, out, nullptr, N, gh, gw, E, 2, (float)gh, (float)gw, st));
    else c.run(ocv_grid_sample_pos_f32(table, pts_tmp, out, nullptr, N, gh, gw, E, 1, h * 2.f, w * 2.f, st));
  } else {
    c.run(ocv_ps_roi_align_pos_f32(table, rows, out, nullptr, N, gh, gw, E, image_side ? 1.f / P : 1.f / (P * 2.f), 1, st));
  }
}

}  // namespace

extern "C" int ocv_head_forward(ocv_head* H, const void* X, const float* obj_feat, const float* obj_xywh,
                                const int32_t* obj_offsets, int total, int n_max, int B, int h, int w, float* depth,
                                float* bin_edges, float* obj_tokens_out, void* workspace, void* stream) {
  OCV_CHECK_ARG(H && X && depth && bin_edges && workspace && B > 0 && h >= P && w >= P);
  OCV_CHECK_ARG(H->packed != nullptr);  // ocv_head_pack first
  const int S = (h / P) * (w / P), M = B * S;
  OCV_CHECK_SHAPE(S >= QN + 1 && S <= IMGPOS_ROWS);  // token 0 + 128 queries (miniViT.py:27)
  const bool ocv = H->cfg.kind == 1;
  if (ocv) {
    OCV_CHECK_ARG(obj_feat && obj_xywh && obj_offsets && total > 0 && n_max > 0);
    OCV_CHECK_SHAPE(n_max <= S);  // F.pad with a negative amount would crop: not supported
  }
  if (int s = ocv_require_sm100()) return s;
  const bool piped = H->lanes > 0;
  if (ocv && !piped && !H->side) {
    cudaError_t e = cudaStreamCreateWithFlags(&H->side, cudaStreamNonBlocking);
    for (int i = 0; i < 4 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&H->ev[i], cudaEventDisableTiming);
    if (e != cudaSuccess) return (int)e;
  }
  H->launches = 0;
  H->missing.clear();
  Ctx c{H};
  const std::string& pre = H->pre;
  void* s0 = stream;
  void* s1 = H->side;
  cudaEvent_t* evs = H->ev;
  ocv_head::Lane* lane = nullptr;
  if (piped) {
    lane = &H->lane[H->calls++ % (unsigned)H->lanes];
    s0 = lane->tok;
    s1 = lane->side;
    evs = lane->ev;
    // batch-invariant cached operands are produced on the CALLER's stream, before any lane may read them
    if (ocv && (H->imgpos_h != h || H->imgpos_w != w)) {
      float* imgpos = reinterpret_cast<float*>(H->packed + H->off_imgpos);
      float* coords = reinterpret_cast<float*>(H->packed + H->off_coords);
      for (int l = 0; l < H->lanes && c.ok(); ++l)  // lanes still reading the previous (h, w) rows finish first
        if (H->lane[l].used) {
          cudaError_t e = cudaStreamWaitEvent(ocv_stream(stream), H->lane[l].tok_done, 0);
          if (e != cudaSuccess) c.status = (int)e;
        }
      c.run(ocv_patch_coords_f32(coords, h / P, w / P, P, stream));
      positional(c, coords, S, true, h, w, reinterpret_cast<float*>(H->packed + H->off_pts), imgpos, stream);
      if (c.ok()) {
        H->imgpos_h = h;
        H->imgpos_w = w;
      }
    }
    if (c.ok()) {  // the lane starts when the caller's inputs are ready and its previous pixel kernel has released the workspace
      cudaError_t e = cudaEventRecord(H->in_ready, ocv_stream(stream));
      if (e == cudaSuccess) e = cudaStreamWaitEvent(ocv_stream(s0), H->in_ready, 0);
      if (e == cudaSuccess && lane->used) e = cudaStreamWaitEvent(ocv_stream(s0), lane->pix_done, 0);
      if (e != cudaSuccess) c.status = (int)e;
    }
    const size_t slice = carve(H, nullptr, B, h, w, total, n_max).bytes;
    workspace = static_cast<uint8_t*>(workspace) + slice * ((H->calls - 1) % (unsigned)H->lanes);
  }
  Workspace W = carve(H, workspace, B, h, w, total, n_max);
  // pipelined mode: the token path reads lane-static copies of the small object inputs and writes the bin edges / embedded object
  // tokens into lane-static buffers, so that its launch sequence depends only on (X, shapes, workspace) and can be replayed
  float* edges_out = bin_edges;
  if (piped && c.ok()) {
    cudaError_t e = cudaSuccess;
    if (ocv) {
      e = cudaMemcpyAsync(W.in_feat, obj_feat, (size_t)total * H->cfg.obj_dim * 4, cudaMemcpyDeviceToDevice, ocv_stream(s0));
      if (e == cudaSuccess) e = cudaMemcpyAsync(W.in_xywh, obj_xywh, (size_t)total * 16, cudaMemcpyDeviceToDevice, ocv_stream(s0));
      if (e == cudaSuccess) e = cudaMemcpyAsync(W.in_off, obj_offsets, (size_t)(B + 1) * 4, cudaMemcpyDeviceToDevice, ocv_stream(s0));
      obj_feat = W.in_feat;
      obj_xywh = W.in_xywh;
      obj_offsets = W.in_off;
    }
    if (e != cudaSuccess) c.status = (int)e;
    edges_out = W.edges;
  }
  float* obj_tok = (obj_tokens_out && !piped) ? obj_tokens_out : W.obj_tok;
  auto token_path = [&]() -> int {
  auto fork = [&](cudaEvent_t ev, void* from, void* to) {
    if (!c.ok()) return;
    cudaError_t e = cudaEventRecord(ev, ocv_stream(from));
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ocv_stream(to), ev, 0);
    if (e != cudaSuccess) c.status = (int)e;
  };

  // ---- object branch (side stream): ObjCAViT.py:309-330, 180-194 ------------------------------------------------
  if (ocv) {
    fork(evs[0], s0, s1);
    positional(c, obj_xywh, total, false, h, w, W.pts, W.obj_pos, s1);
    const bf16* w_emb = H->w16(pre + "obj_embedding_layer.weight");
    const float* b_emb = H->f32(pre + "obj_embedding_layer.bias");
    if (!w_emb || !b_emb) return OCV_E_ARG;
    const int D = H->cfg.obj_dim;
    c.run(ocv_cast_f32_to_bf16(obj_feat, W.feat16, (long long)total * D, s1));
    c.run(ocv_gemm_bf16_ex(W.feat16, w_emb, b_emb, W.obj_pos, E, nullptr, nullptr, nullptr, obj_tok, E, nullptr, 0, total, E,
                           D, D, D, OCV_ACT_NONE, s1));
    // pad_sequence: objects -> [B,n_max,E] (pad 0.0001), mask True on padding (:180-183)
    c.run(ocv_pad_objects_impl(obj_tok, obj_offsets, nullptr, W.obj.x32[0], W.obj.x16[0], W.mask, B, n_max, n_max, E,
                               0.0001f, 0, s1));
    int ocur = 0;
    if (!H->cfg.no_obj_sa) ocur = encoder(c, pre + "saca_1.obj_transformer_encoder.", W.obj, 0, B, n_max, W.mask, s1);
    // :192-194 objects LEFT-padded to S, mask RIGHT-padded (F5)
    if (c.ok()) {
      c.run(ocv_launch_pdl(uniform_offsets_kernel, dim3(ceil_div(B + 1, 128)), dim3(128), 0, s1, W.uni, B + 1, n_max));
    }
    c.run(ocv_pad_objects_impl(W.obj.x32[ocur], W.uni, obj_offsets, W.obj_s32, W.obj_s16, W.kpm, B, S, n_max, E, 0.0001f, 1,
                               s1));
    if (c.ok()) {
      cudaError_t e = cudaEventRecord(evs[1], ocv_stream(s1));
      if (e != cudaSuccess) c.status = (int)e;
    }
  }

  // ---- image branch (caller's stream) ---------------------------------------------------------------------------
  const float* pos = nullptr;
  const float* pe_bias = nullptr;
  if (!ocv) {
    pos = H->f32(pre + "patch_transformer.positional_encodings");
    pe_bias = H->f32(pre + "patch_transformer.embedding_convPxP.bias");
    if (!pos || !pe_bias) return OCV_E_ARG;
    OCV_CHECK_SHAPE(H->params[pre + "patch_transformer.positional_encodings"].numel >= (long long)S * E);
  } else {
    pe_bias = H->f32(pre + "image_embedding_convPxP.bias");
    if (!pe_bias) return OCV_E_ARG;
    float* imgpos = reinterpret_cast<float*>(H->packed + H->off_imgpos);
    float* coords = reinterpret_cast<float*>(H->packed + H->off_coords);
    if (H->imgpos_h != h || H->imgpos_w != w) {  // batch-invariant: once per (h, w) and weight version
      c.run(ocv_patch_coords_f32(coords, h / P, w / P, P, s0));
      positional(c, coords, S, true, h, w, reinterpret_cast<float*>(H->packed + H->off_pts), imgpos, s0);
      if (c.ok()) {
        H->imgpos_h = h;
        H->imgpos_w = w;
      }
    }
    pos = imgpos;
  }
  c.run(ocv_patch_embed_bf16(X, H->packed + H->off_patch, pe_bias, pos, W.img.x32[0], W.img.x16[0], B, E, h, w, P, E, W.pe_ws,
                             s0), 2);
  int cur = encoder(c, ocv ? pre + "saca_1.image_transformer_encoder." : pre + "patch_transformer.transformer_encoder.",
                    W.img, 0, B, S, nullptr, s0);
  if (ocv) {
    if (c.ok()) {
      cudaError_t e = cudaStreamWaitEvent(ocv_stream(s0), evs[1], 0);
      if (e != cudaSuccess) c.status = (int)e;
    }
    const bool two = H->cfg.use_2_saca != 0;
    if (two) {  // final_object_features feed block 2 (F6): q = objects, k = image tokens, v = objects (:202-207)
      fork(evs[2], s0, s1);
      cross_attention(c, pre + "saca_1.cross_attn_im_obj.", W.obj, W.obj_s16, W.img.x16[cur], nullptr, W.obj.x32[0],
                      W.obj.x16[0], B, S, s1);
      if (c.ok()) {
        cudaError_t e = cudaEventRecord(evs[3], ocv_stream(s1));
        if (e != cudaSuccess) c.status = (int)e;
      }
    }
    int nxt = (cur + 1) % 3;  // :195-201 q = image, k = padded objects, v = image
    cross_attention(c, pre + "saca_1.cross_attn_obj_im.", W.img, W.img.x16[cur], W.obj_s16, W.kpm, W.img.x32[nxt],
                    W.img.x16[nxt], B, S, s0);
    cur = nxt;
    if (two) {
      if (c.ok()) {  // the side stream still reads the block-1 image tokens as keys: let it finish before they are recycled
        cudaError_t e = cudaStreamWaitEvent(ocv_stream(s0), evs[3], 0);
        if (e != cudaSuccess) c.status = (int)e;
      }
      int ocur = 0;
      if (!H->cfg.no_obj_sa) ocur = encoder(c, pre + "saca_2.obj_transformer_encoder.", W.obj, 0, B, S, nullptr, s1);
      if (c.ok()) {
        cudaError_t e = cudaEventRecord(evs[1], ocv_stream(s1));
        if (e != cudaSuccess) c.status = (int)e;
      }
      cur = encoder(c, pre + "saca_2.image_transformer_encoder.", W.img, cur, B, S, nullptr, s0);
      if (c.ok()) {
        cudaError_t e = cudaStreamWaitEvent(ocv_stream(s0), evs[1], 0);
        if (e != cudaSuccess) c.status = (int)e;
      }
      nxt = (cur + 1) % 3;
      cross_attention(c, pre + "saca_2.cross_attn_obj_im.", W.img, W.img.x16[cur], W.obj.x16[ocur], nullptr, W.img.x32[nxt],
                      W.img.x16[nxt], B, S, s0);
      cur = nxt;
    }
  }

  // ---- regressor, bins, folded queries, fused pixel kernel ---------------------------------------------------------
  const float* tok = W.img.x32[cur];
  {
    const int dims[4] = {E, 256, 256, H->cfg.n_bins};
    const float* Ws[3];
    const float* bs[3];
    for (int i = 0; i < 3; ++i) {
      Ws[i] = H->f32(pre + "regressor." + std::to_string(MLP_REG_IDX[i]) + ".weight");
      bs[i] = H->f32(pre + "regressor." + std::to_string(MLP_REG_IDX[i]) + ".bias");
      if (!Ws[i] || !bs[i]) return OCV_E_ARG;
    }
    c.run(ocv_mlp_rows_f32(tok, (long long)S * E, B, 3, Ws, bs, dims, W.y_raw, W.reg_ws, s0), 3);
  }
  c.run(ocv_bins_f32(W.y_raw, W.blog /*normalised widths: scratch, overwritten below*/, edges_out, W.centres, B, H->cfg.n_bins,
                     OCV_NORM_LINEAR, H->cfg.min_depth, H->cfg.max_depth, s0));
  const float *w_out = H->f32("conv_out.0.weight"), *b_out = H->f32("conv_out.0.bias"), *b3 = H->f32(pre + "conv3x3.bias");
  if (!w_out || !b_out || !b3) return OCV_E_ARG;
  c.run(ocv_fold_queries_bf16(w_out, b_out, tok + E, (long long)S * E, E, b3, 1.4426950408889634f, W.Mb, W.blog, B,
                              H->cfg.n_bins, QN, E, s0));
  return c.status;
  };  // token_path

  if (!piped) {
    if (int st = token_path()) return st;
  } else if (c.ok()) {
    // signature of the launch sequence: everything the token path's kernels were given
    const unsigned long long sig[8] = {0ull /* the input address is NOT part of it: see Lane::pe_node */, (unsigned long long)(uintptr_t)workspace,
                                       ((unsigned long long)B << 40) | ((unsigned long long)h << 20) | (unsigned long long)w,
                                       ((unsigned long long)total << 32) | (unsigned long long)n_max,
                                       (unsigned long long)(uintptr_t)H->packed, 1ull, 0ull, 0ull};
    bool replayed = false;
    if (H->use_graphs && !lane->graph_failed) {
      if (lane->gexec && memcmp(sig, lane->sig, sizeof(sig)) == 0) {
        cudaError_t e = cudaSuccess;
        if (X != lane->graph_x) {   // same shapes, new input tensor: re-point the one node that reads it
          CUtensorMap tmX;
          if (int st = ocv_patch_embed_xmap(&tmX, X, B, E, h, w, P)) return st;
          cudaKernelNodeParams kp = lane->pe_params;
          void* args[3] = {&tmX, lane->pe_params.kernelParams[1], lane->pe_params.kernelParams[2]};
          kp.kernelParams = args;
          e = cudaGraphExecKernelNodeSetParams(lane->gexec, lane->pe_node, &kp);
          if (e != cudaSuccess) return (int)e;
          lane->graph_x = X;
        }
        e = cudaGraphLaunch(lane->gexec, ocv_stream(s0));
        if (e != cudaSuccess) return (int)e;
        H->launches += lane->graph_launches;
        replayed = true;
        if (getenv("OCV_DEBUG")) fprintf(stderr, "[ocv] replay\n");
      } else if (memcmp(sig, lane->cand, sizeof(sig)) == 0) {  // second sighting: capture (the capture itself launches nothing)
        if (lane->gexec) {
          cudaGraphExecDestroy(lane->gexec);
          lane->gexec = nullptr;
        }
        if (lane->graph) {
          cudaGraphDestroy(lane->graph);
          lane->graph = nullptr;
        }
        lane->pe_node = nullptr;
        cudaGraph_t g = nullptr;
        cudaError_t e = cudaStreamBeginCapture(ocv_stream(s0), cudaStreamCaptureModeThreadLocal);
        int st = OCV_OK;
        if (e == cudaSuccess) {
          const int before = H->launches;
          st = token_path();
          e = cudaStreamEndCapture(ocv_stream(s0), &g);
          if (st == OCV_OK && e == cudaSuccess) e = cudaGraphInstantiate(&lane->gexec, g, 0);
          if (st == OCV_OK && e == cudaSuccess) {   // find the patch-embedding launch (the only reader of X inside the graph)
            size_t nn = 0;
            e = cudaGraphGetNodes(g, nullptr, &nn);
            std::vector<cudaGraphNode_t> nodes(nn);
            if (e == cudaSuccess && nn) e = cudaGraphGetNodes(g, nodes.data(), &nn);
            for (size_t i = 0; i < nn && e == cudaSuccess; ++i) {
              cudaGraphNodeType ty;
              if (cudaGraphNodeGetType(nodes[i], &ty) != cudaSuccess || ty != cudaGraphNodeTypeKernel) continue;
              cudaKernelNodeParams kp;
              if (cudaGraphKernelNodeGetParams(nodes[i], &kp) != cudaSuccess) continue;
              if (ocv_patch_embed_is_kernel(kp.func)) {
                lane->pe_node = nodes[i];
                lane->pe_params = kp;   // kernelParams points into the graph's own argument storage: the graph stays alive
              }
            }
            if (e == cudaSuccess && !lane->pe_node) e = cudaErrorInvalidValue;   // cannot re-point: do not replay at all
          }
          if (e == cudaSuccess && st == OCV_OK) {
            lane->graph = g;
            lane->graph_x = X;
          } else if (g) {
            cudaGraphDestroy(g);
          }
          lane->graph_launches = H->launches - before;
          H->launches = before;
        }
        if (getenv("OCV_DEBUG")) fprintf(stderr, "[ocv] lane capture: status %d cuda %d (%s) nodes %d\n", st, (int)e, cudaGetErrorString(e), lane->graph_launches);
        if (st != OCV_OK || e != cudaSuccess || !lane->gexec) {
          cudaGetLastError();
          if (lane->gexec) cudaGraphExecDestroy(lane->gexec);
          lane->gexec = nullptr;
          lane->graph_failed = true;  // fall back to eager sequencing for this lane
          c.status = OCV_OK;
        } else {
          memcpy(lane->sig, sig, sizeof(sig));
          e = cudaGraphLaunch(lane->gexec, ocv_stream(s0));
          if (e != cudaSuccess) return (int)e;
          H->launches += lane->graph_launches;
          replayed = true;
        }
      } else {
        memcpy(lane->cand, sig, sizeof(sig));
      }
    }
    if (!replayed)
      if (int st = token_path()) return st;
    cudaError_t e = cudaMemcpyAsync(bin_edges, W.edges, (size_t)B * (H->cfg.n_bins + 1) * 4, cudaMemcpyDeviceToDevice, ocv_stream(s0));
    if (e == cudaSuccess && ocv && obj_tokens_out)
      e = cudaMemcpyAsync(obj_tokens_out, W.obj_tok, (size_t)total * E * 4, cudaMemcpyDeviceToDevice, ocv_stream(s0));
    if (e != cudaSuccess) return (int)e;
  }
  if (!c.ok()) return c.status;

  void* sp = s0;
  if (piped && c.ok()) {  // hand over to the shared pixel stream
    sp = H->pix;
    cudaError_t e = cudaEventRecord(lane->tok_done, ocv_stream(s0));
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ocv_stream(sp), lane->tok_done, 0);
    if (e != cudaSuccess) c.status = (int)e;
  }
  if (H->timing && c.ok()) cudaEventRecord(H->tev[0], ocv_stream(sp));
  c.run(ocv_pixel_fused_bf16(X, H->packed + H->off_w3p, W.Mb, W.blog, W.centres, depth, B, h, w, piped ? H->pixel_ctas : 0, sp));
  if (H->timing && c.ok()) cudaEventRecord(H->tev[1], ocv_stream(sp));
  if (piped && c.ok()) {
    cudaError_t e = cudaEventRecord(lane->pix_done, ocv_stream(sp));
    if (e != cudaSuccess) c.status = (int)e;
    lane->used = true;
  }
  return c.status;
}

extern "C" int ocv_head_set_pipeline(ocv_head* H, int lanes, int pixel_ctas) {
  OCV_CHECK_ARG(H && lanes >= 0 && lanes <= 4 && pixel_ctas >= 0);
  if (lanes > 0) {
    cudaError_t e = cudaSuccess;
    if (!H->pix) e = acquire_stream(&H->pix);
    if (e == cudaSuccess && !H->in_ready) e = cudaEventCreateWithFlags(&H->in_ready, cudaEventDisableTiming);
    for (int l = 0; l < lanes && e == cudaSuccess; ++l) {
      ocv_head::Lane& ln = H->lane[l];
      if (ln.tok) continue;
      e = acquire_stream(&ln.tok);
      if (e == cudaSuccess) e = acquire_stream(&ln.side);
      for (int i = 0; i < 4 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&ln.ev[i], cudaEventDisableTiming);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ln.tok_done, cudaEventDisableTiming);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ln.pix_done, cudaEventDisableTiming);
    }
    if (e != cudaSuccess) return (int)e;
  }
  H->lanes = lanes;
  H->pixel_ctas = pixel_ctas;
  H->calls = 0;
  H->use_graphs = getenv("OCV_NO_GRAPHS") == nullptr;
  return OCV_OK;
}

extern "C" int ocv_head_join(ocv_head* H, void* stream) {
  OCV_CHECK_ARG(H);
  for (auto& ln : H->lane) {
    if (!ln.used) continue;
    cudaError_t e = cudaStreamWaitEvent(ocv_stream(stream), ln.pix_done, 0);
    if (e != cudaSuccess) return (int)e;
  }
  return OCV_OK;
}

extern "C" void* ocv_head_stream(const ocv_head* H, int which, int lane) {
  if (!H || lane < 0 || lane >= 4) return nullptr;
  return which == 0 ? (void*)H->lane[lane].tok : which == 1 ? (void*)H->lane[lane].side : (void*)H->pix;
}

extern "C" int ocv_head_enable_timing(ocv_head* H, int on) {
  OCV_CHECK_ARG(H);
  if (on && !H->tev[0]) {
    for (auto& e : H->tev) {
      cudaError_t r = cudaEventCreate(&e);
      if (r != cudaSuccess) return (int)r;
    }
  }
  H->timing = on != 0;
  return OCV_OK;
}

extern "C" float ocv_head_pixel_ms(ocv_head* H) {
  if (!H || !H->tev[0]) return -1.f;
  float ms = -1.f;
  if (cudaEventSynchronize(H->tev[1]) != cudaSuccess) return -1.f;
  if (cudaEventElapsedTime(&ms, H->tev[0], H->tev[1]) != cudaSuccess) return -1.f;
  return ms;
}
