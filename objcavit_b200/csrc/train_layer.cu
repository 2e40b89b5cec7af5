// Native sequencer of one nn.TransformerEncoderLayer of the TRAINING step (post-norm, ReLU, E = 128, 4 heads; dropout at the four
// sites of SURVEY.md F8): the forward with every activation the backward needs, and the explicit backward.  It launches exactly the
// kernels `engine_train.py` used to sequence call by call from Python (same entry points, same order, same split-K choices, so
// the results are bit-identical); the point is the HOST: ~30 C-ABI calls per layer and direction become one, which matters
// because the bf16 step is bound by host issue time.  The weight / bias gradient GEMMs go to `side_stream`
// (nothing downstream of them inside the layer), ordered against the main stream with events.
// replaces: nn.TransformerEncoderLayer forward + autograd backward (modules/layers.py:8; modules/ObjCAViT.py:155-161).
// Contract: include/objcavit_b200.h.
#include "common.cuh"
#include "objcavit_b200.h"

namespace {

constexpr int LE = 128, LH = 4, LD = 32;
enum { W_IN = 0, B_IN, W_OUT, B_OUT, W_1, B_1, W_2, B_2, G_1, BE_1, G_2, BE_2, N_LAYER_TENSORS };

inline int auto_split(int M, int N, int K) {  // objcavit_b200/ops.py:_auto_split
  const int tiles = ((M + 63) / 64) * ((N + 63) / 64);
  int s = (192 + tiles - 1) / tiles;
  if (s > K / 64) s = K / 64;
  if (s > 64) s = 64;
  return s < 1 ? 1 : s;
}

inline int linear_split(int M, int N, int K) {  // objcavit_b200/ops.py:linear — long K, few tiles
  if (K < 512 || ((M + 63) / 64) * ((N + 63) / 64) > 64) return 0;
  long long s = K / 128;
  const long long cap = (192LL * 64 * 64) / ((long long)M * N > 0 ? (long long)M * N : 1);
  if (s > cap) s = cap;
  if (s > 32) s = 32;
  return (int)(s < 2 ? 2 : s);
}

inline size_t al(size_t v) { return (v + 255) / 256 * 256; }

// events ordering the two streams (never destroyed: the process owns them for its lifetime)
cudaEvent_t next_event() {
  static cudaEvent_t pool[16];
  static bool made = false;
  static int cur = 0;
  if (!made) {
    for (auto& e : pool) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    made = true;
  }
  cur = (cur + 1) % 16;
  return pool[cur];
}

inline int order_after(cudaStream_t later, cudaStream_t earlier) {  // work issued on `later` from now on waits for `earlier` so far
  if (later == earlier) return OCV_OK;
  cudaEvent_t e = next_event();
  if (cudaError_t r = cudaEventRecord(e, earlier)) return (int)r;
  if (cudaError_t r = cudaStreamWaitEvent(later, e, 0)) return (int)r;
  return OCV_OK;
}

#define TRY(x) do { if (int s_ = (x)) return s_; } while (0)

int linear(const float* a, const float* w, const float* b, const float* res, float* c, int M, int N, int K, int act, void* st) {
  return ocv_linear_f32(a, w, b, res, c, M, N, K, K, K, N, 1, act, 1, 0, 0, 0, st);
}

// dw [N,K] += dy[M,N]^T x[M,K];  db [N] += column sums of dy      (objcavit_b200/ops.py:linear_wgrad)
int wgrad(const float* dy, const float* x, float* dw, float* db, int M, int N, int K, void* ws, void* st) {
  TRY(ocv_gemm_f32_ex(1, 0, dy, x, dw, N, K, M, N, K, K, 1.f, 1, 0, 0, 0, 0, auto_split(N, K, M), nullptr, ws, st));
  return ocv_reduce_channels_f32(dy, db, 1, M, N, 1, 1.f, st);
}

// dx [M,K] = beta*dx + dy[M,N] w[N,K]                               (objcavit_b200/ops.py:linear_dgrad)
int dgrad(const float* dy, const float* w, float* dx, int M, int N, int K, float beta, void* ws, void* st) {
  return ocv_gemm_f32_ex(0, 0, dy, w, dx, M, K, N, N, K, K, beta, 1, 0, 0, 0, 0, auto_split(M, K, N), nullptr, ws, st);
}

long long gemm_ws_floats(int M, int F) {  // the largest partial buffer any GEMM of the layer needs (per stream)
  long long need = 0;
  auto upd = [&](int m, int n, int k) {
    const long long v = (long long)auto_split(m, n, k) * m * n;
    if (v > need) need = v;
  };
  upd(LE, F, M); upd(F, LE, M); upd(LE, LE, M); upd(3 * LE, LE, M);   // weight gradients
  upd(M, F, LE); upd(M, LE, F); upd(M, LE, LE); upd(M, LE, 3 * LE);   // data gradients
  return need;
}

}  // namespace

extern "C" long long ocv_encoder_layer_fwd_train_workspace(int B, int S, int F) {
  if (B <= 0 || S <= 0 || F <= 0) return 0;
  const int M = B * S, sp = linear_split(M, LE, F);
  return (long long)al((size_t)(sp > 0 ? sp : 1) * M * LE * sizeof(float));
}

extern "C" int ocv_encoder_layer_fwd_train_f32(const float* const* W, const float* x, const uint8_t* mask, int B, int S, int F,
                                               float drop_p, const unsigned long long* seeds, float* qkv, float* att, float* y,
                                               float* x1, float* h, float* y2, float* out, void* workspace, void* stream) {
  OCV_CHECK_ARG(W && x && qkv && att && y && x1 && h && y2 && out && workspace && B > 0 && S > 0 && F > 0);
  OCV_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || seeds));
  for (int i = 0; i < N_LAYER_TENSORS; ++i) OCV_CHECK_ARG(W[i] != nullptr);
  const int M = B * S;
  const long long ME = (long long)M * LE;
  const bool drop = drop_p > 0.f;
  TRY(linear(x, W[W_IN], W[B_IN], nullptr, qkv, M, 3 * LE, LE, OCV_ACT_NONE, stream));
  TRY(ocv_mha_core_dropout_f32(qkv, qkv + LE, qkv + 2 * LE, att, mask, B, S, S, LH, LD, (long long)S * 3 * LE, 3 * LE,
                               (long long)S * 3 * LE, 3 * LE, (long long)S * 3 * LE, 3 * LE, (long long)S * LE, LE,
                               drop ? drop_p : 0.f, drop ? seeds[0] : 0ull, stream));
  if (drop) {                                                                         // y = x + dropout1(self-attention), x1 = LN(y)
    TRY(linear(att, W[W_OUT], W[B_OUT], nullptr, y, M, LE, LE, OCV_ACT_NONE, stream));
    TRY(ocv_dropout_add_layernorm_f32(y, x, W[G_1], W[BE_1], y, x1, M, LE, drop_p, seeds[1], stream));
  } else {
    TRY(linear(att, W[W_OUT], W[B_OUT], x, y, M, LE, LE, OCV_ACT_NONE, stream));
    TRY(ocv_add_layernorm_f32(y, nullptr, W[G_1], W[BE_1], x1, M, LE, nullptr, stream));
  }
  TRY(linear(x1, W[W_1], W[B_1], nullptr, h, M, F, LE, OCV_ACT_RELU, stream));
  if (drop) {
    TRY(ocv_dropout_f32(h, nullptr, h, (long long)M * F, drop_p, seeds[2], stream));   // dropout(relu(linear1)): what linear2 sees
    const int sp = linear_split(M, LE, F);
    if (sp > 0) TRY(ocv_linear_splitk_f32(h, W[W_2], W[B_2], y2, M, LE, F, F, F, LE, sp, workspace, stream));
    else TRY(linear(h, W[W_2], W[B_2], nullptr, y2, M, LE, F, OCV_ACT_NONE, stream));
    return ocv_dropout_add_layernorm_f32(y2, x1, W[G_2], W[BE_2], y2, out, M, LE, drop_p, seeds[3], stream);   // x1 + dropout2(linear2), LN
  }
  TRY(linear(h, W[W_2], W[B_2], x1, y2, M, LE, F, OCV_ACT_NONE, stream));
  return ocv_add_layernorm_f32(y2, nullptr, W[G_2], W[BE_2], out, M, LE, nullptr, stream);
}

extern "C" long long ocv_encoder_layer_bwd_train_workspace(int B, int S, int F) {
  if (B <= 0 || S <= 0 || F <= 0) return 0;
  const long long M = (long long)B * S;
  size_t bytes = 0;
  bytes += 4 * al((size_t)M * LE * 4);            // dy2, df, dsa, datt
  bytes += 2 * al((size_t)M * F * 4);             // dh, dh2
  bytes += al((size_t)M * 3 * LE * 4);            // dqkv
  bytes += al((size_t)B * LH * S * 3 * 4);        // attention statistics
  bytes += 2 * al((size_t)(M * 2 + 32 * LE) * 4); // LayerNorm row statistics + parameter partials, one set per LayerNorm
  bytes += 2 * al((size_t)gemm_ws_floats((int)M, F) * 4);   // GEMM partials: main stream, weight-gradient stream
  return (long long)bytes;
}

extern "C" int ocv_encoder_layer_bwd_train_f32(const float* const* W, float* const* G, const float* x, const float* qkv,
                                               const float* att, const float* y, const float* x1, const float* h, const float* y2,
                                               const uint8_t* mask, const float* dout, float* dx, int B, int S, int F, float drop_p,
                                               const unsigned long long* seeds, void* workspace, void* stream, void* side_stream) {
  OCV_CHECK_ARG(W && G && x && qkv && att && y && x1 && h && y2 && dout && dx && workspace && B > 0 && S > 0 && F > 0);
  OCV_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || seeds));
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(workspace) & 255) == 0);
  for (int i = 0; i < N_LAYER_TENSORS; ++i) OCV_CHECK_ARG(W[i] != nullptr && G[i] != nullptr);
  const int M = B * S;
  const long long ME = (long long)M * LE, MF = (long long)M * F;
  const bool drop = drop_p > 0.f;
  cudaStream_t st = ocv_stream(stream), sd = side_stream ? ocv_stream(side_stream) : st;
  uint8_t* p = static_cast<uint8_t*>(workspace);
  auto take = [&](size_t bytes) { void* r = p; p += al(bytes); return r; };
  float* dy2 = static_cast<float*>(take(ME * 4));
  float* dfb = static_cast<float*>(take(ME * 4));
  float* dsab = static_cast<float*>(take(ME * 4));
  float* datt = static_cast<float*>(take(ME * 4));
  float* dh = static_cast<float*>(take(MF * 4));
  float* dh2 = static_cast<float*>(take(MF * 4));
  float* dqkv = static_cast<float*>(take(ME * 3 * 4));
  void* stats = take((size_t)B * LH * S * 3 * 4);
  void* lnws2 = take((size_t)((long long)M * 2 + 32 * LE) * 4);   // norm2 / norm1: their parameter reductions run on the
  void* lnws1 = take((size_t)((long long)M * 2 + 32 * LE) * 4);   // weight-gradient stream while the chain goes on
  const size_t gws = (size_t)gemm_ws_floats(M, F) * 4;
  void* ws_main = take(gws);
  void* ws_side = take(gws);
  float* dy = dx;  // d loss / d (x + attention branch), becomes d loss / d x after the last accumulation

  const float* df = dy2;                                                           // gradient of linear2's output
  if (drop) {
    TRY(ocv_layernorm_bwd_dropout_f32(dout, y2, W[G_2], dy2, dfb, M, LE, drop_p, seeds[3], lnws2, st));
    df = dfb;
  } else {
    TRY(ocv_layernorm_bwd_f32(dout, y2, W[G_2], dy2, nullptr, nullptr, M, LE, 1.f, lnws2, st));
  }
  TRY(order_after(sd, st));
  TRY(ocv_layernorm_bwd_params_f32(dout, y2, G[G_2], G[BE_2], M, LE, 1.f, lnws2, sd));
  TRY(wgrad(df, h, G[W_2], G[B_2], M, LE, F, ws_side, sd));
  TRY(dgrad(df, W[W_2], dh, M, LE, F, 0.f, ws_main, st));
  if (drop) TRY(ocv_dropout_act_bwd_f32(dh, h, dh2, MF, OCV_ACT_RELU, drop_p, seeds[2], st));   // h is post-dropout: h > 0 <=> kept & positive
  else TRY(ocv_act_bwd_f32(dh, h, dh2, MF, OCV_ACT_RELU, st));
  TRY(order_after(sd, st));
  TRY(wgrad(dh2, x1, G[W_1], G[B_1], M, F, LE, ws_side, sd));
  if (!drop) TRY(order_after(st, sd));                                             // df aliases dy2, updated in place next
  TRY(dgrad(dh2, W[W_1], dy2, M, F, LE, 1.f, ws_main, st));                        // dy2 <- d loss / d x1 (residual + FFN)
  const float* dsa = dy;                                                           // gradient of out_proj's output
  if (drop) {
    TRY(ocv_layernorm_bwd_dropout_f32(dy2, y, W[G_1], dy, dsab, M, LE, drop_p, seeds[1], lnws1, st));
    dsa = dsab;
  } else {
    TRY(ocv_layernorm_bwd_f32(dy2, y, W[G_1], dy, nullptr, nullptr, M, LE, 1.f, lnws1, st));
  }
  TRY(order_after(sd, st));
  TRY(ocv_layernorm_bwd_params_f32(dy2, y, G[G_1], G[BE_1], M, LE, 1.f, lnws1, sd));
  TRY(wgrad(dsa, att, G[W_OUT], G[B_OUT], M, LE, LE, ws_side, sd));
  TRY(dgrad(dsa, W[W_OUT], datt, M, LE, LE, 0.f, ws_main, st));
  const long long qs = (long long)S * 3 * LE;
  TRY(ocv_mha_core_bwd_dropout_f32(qkv, qkv + LE, qkv + 2 * LE, att, datt, mask, dqkv, dqkv + LE, dqkv + 2 * LE, B, S, S, LH, LD, qs,
                                   3 * LE, qs, 3 * LE, qs, 3 * LE, (long long)S * LE, LE, qs, 3 * LE, qs, 3 * LE, qs, 3 * LE,
                                   drop ? drop_p : 0.f, drop ? seeds[0] : 0ull, stats, st));
  TRY(order_after(sd, st));
  TRY(wgrad(dqkv, x, G[W_IN], G[B_IN], M, 3 * LE, LE, ws_side, sd));
  if (!drop) TRY(order_after(st, sd));                                             // dsa aliases dy, updated in place next
  return dgrad(dqkv, W[W_IN], dy, M, 3 * LE, LE, 1.f, ws_main, st);                // dy <- d loss / d x (residual + attention)
}

// ======================================================================================================================
// Tensor-core form of the same layer (training step, precision "bf16"): every GEMM of the forward and of the backward —
// data gradients AND weight gradients — runs on tcgen05 (ocv_gemm_bf16_v2: bf16 operands, fp32 accumulation in TMEM);
// bias + ReLU + the dropout sites + residual + LayerNorm are the GEMM's epilogue, the transposed contractions of the backward
// (dX = dY W, dW = dY^T X) read the SAME row-major tensors as MN-major operands, so no transposed copies exist.  The attention core
// and the LayerNorm / activation backward stay the fp32 kernels (softmax statistics, row reductions); their outputs get one bf16
// copy each (ocv_cast_f32_to_bf16) as the next GEMM's A operand.  fp32 master weights; W16 = bf16 copies in nn.Linear layout.
// ======================================================================================================================
namespace {

enum { W16_IN = 0, W16_OUT, W16_1, W16_2, N_LAYER_W16 };

int gemm16(const void* A, const void* B, int a_mn, int b_mn, const float* bias, const float* res, const float* g, const float* be,
           float* pre, float* c32, void* c16, int M, int N, int K, int lda, int ldb, int act, float drop_p, unsigned long long seed,
           int split, void* ws, void* st) {
  return ocv_gemm_bf16_v2(A, B, a_mn, b_mn, bias, res, N, g, be, pre, N, c32, N, c16, N, M, N, K, lda, ldb, act, drop_p, seed, split, ws, st);
}

inline int wsplit(int Mtok) {  // split-K of the weight-gradient GEMMs (K = token rows): only for long token counts — at the
  int s = Mtok / 4096;         // config-5 size (1768 rows = 28 k-steps) one CTA per output tile is faster than GEMM + reduction
  return s < 1 ? 1 : (s > 8 ? 8 : s);
}

// second weight-gradient queue (bias-gradient column sums, LayerNorm parameter reductions): these 2-12 us kernels would otherwise
// queue behind the weight-gradient GEMMs on the one side stream and become the tail of the step.  One per device, never destroyed.
cudaStream_t second_side_stream() {
  static cudaStream_t s[16] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
  if (!s[dev] && cudaStreamCreateWithFlags(&s[dev], cudaStreamNonBlocking) != cudaSuccess) s[dev] = nullptr;
  return s[dev];
}

// dW [N,K] += dY^T X (both operands MN-major) on `st`; dB [N] += column sums of dY (fp32) on `st2`
int wgrad16(const void* dy16, const float* dy32, const void* x16, float* dw, float* db, int Mtok, int N, int K, void* ws, void* st,
            void* st2) {
  const int sp = wsplit(Mtok);
  TRY(ocv_gemm_bf16_v2(dy16, x16, 1, 1, nullptr, dw, K, nullptr, nullptr, nullptr, 0, dw, K, nullptr, 0, N, K, Mtok, N, K, OCV_ACT_NONE,
                       0.f, 0ull, sp, ws, st));
  return ocv_reduce_channels_f32(dy32, db, 1, Mtok, N, 1, 1.f, st2);
}

}  // namespace

extern "C" long long ocv_encoder_layer_bwd_train_bf16_workspace(int B, int S, int F) {
  if (B <= 0 || S <= 0 || F <= 0) return 0;
  const long long M = (long long)B * S;
  size_t bytes = 0;
  bytes += 4 * al((size_t)M * LE * 4);            // dy2, df, dsa, datt
  bytes += 2 * al((size_t)M * F * 4);             // dh, dh2
  bytes += al((size_t)M * 3 * LE * 4);            // dqkv
  bytes += 3 * al((size_t)M * LE * 2) + al((size_t)M * F * 2) + al((size_t)M * 3 * LE * 2);   // bf16 copies: df, dsa, datt, dh2, dqkv
  bytes += al((size_t)B * LH * S * 3 * 4);        // attention statistics
  bytes += 2 * al((size_t)(M * 2 + 32 * LE) * 4); // LayerNorm row statistics + parameter partials
  bytes += 2 * al((size_t)8 * F * 3 * LE * 4);    // split-K partials of the weight-gradient GEMMs (side stream), generous
  return (long long)bytes;
}

extern "C" int ocv_encoder_layer_fwd_train_bf16(const float* const* W, const void* const* W16, const float* x, const void* x16,
                                                const uint8_t* mask, int B, int S, int F, float drop_p,
                                                const unsigned long long* seeds, float* qkv, void* qkv16, float* att_stats, float* att,
                                                void* att16, float* y, float* x1, void* x1_16, float* h, void* h16, float* y2,
                                                float* out, void* out16, void* stream) {
  OCV_CHECK_ARG(W && W16 && x && x16 && qkv && att && att16 && y && x1 && x1_16 && h && h16 && y2 && out && out16 && B > 0 && S > 0 && F > 0);
  OCV_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || seeds));
  for (int i = 0; i < N_LAYER_TENSORS; ++i) OCV_CHECK_ARG(W[i] != nullptr);
  for (int i = 0; i < N_LAYER_W16; ++i) OCV_CHECK_ARG(W16[i] != nullptr);
  const int M = B * S;
  const bool drop = drop_p > 0.f;
  const float p = drop ? drop_p : 0.f;
  // qkv = x W_in^T + b  (fp32: the attention backward reads it; bf16: the tensor-core attention core below)
  const bool tc_att = ocv_mha_tc_supported(S, S, LH, LD, (long long)S * 3 * LE, 3 * LE, (long long)S * 3 * LE, 3 * LE, (long long)S * 3 * LE,
                                           3 * LE, (long long)S * LE, LE) && qkv16 != nullptr;
  TRY(gemm16(x16, W16[W16_IN], 0, 0, W[B_IN], nullptr, nullptr, nullptr, nullptr, qkv, tc_att ? qkv16 : nullptr, M, 3 * LE, LE, LE, LE,
             OCV_ACT_NONE, 0.f, 0ull, 1, nullptr, stream));
  if (tc_att) {
    // attention core on tcgen05: same dropout masks as the fp32 kernel (its backward regenerates them), fp32 + bf16 output
    const __nv_bfloat16* q16 = static_cast<const __nv_bfloat16*>(qkv16);
    TRY(ocv_mha_core_tc_bf16(q16, q16 + LE, q16 + 2 * LE, att16, att, att_stats, mask, B, S, S, LH, LD, (long long)S * 3 * LE, 3 * LE,
                             (long long)S * 3 * LE, 3 * LE, (long long)S * 3 * LE, 3 * LE, (long long)S * LE, LE, p,
                             drop ? seeds[0] : 0ull, stream));
  } else {
    TRY(ocv_mha_core_dropout_f32(qkv, qkv + LE, qkv + 2 * LE, att, mask, B, S, S, LH, LD, (long long)S * 3 * LE, 3 * LE,
                                 (long long)S * 3 * LE, 3 * LE, (long long)S * 3 * LE, 3 * LE, (long long)S * LE, LE, p,
                                 drop ? seeds[0] : 0ull, stream));
    TRY(ocv_cast_f32_to_bf16(att, att16, (long long)M * LE, stream));
  }
  // y = x + dropout1(att W_out^T + b); x1 = LN1(y)      — one GEMM, three outputs (y, x1, bf16 x1)
  TRY(gemm16(att16, W16[W16_OUT], 0, 0, W[B_OUT], x, W[G_1], W[BE_1], y, x1, x1_16, M, LE, LE, LE, LE, OCV_ACT_NONE, p,
             drop ? seeds[1] : 0ull, 1, nullptr, stream));
  // h = dropout(relu(x1 W_1^T + b))                      — fp32 (activation backward) and bf16 (next GEMM, weight gradient)
  TRY(gemm16(x1_16, W16[W16_1], 0, 0, W[B_1], nullptr, nullptr, nullptr, nullptr, h, h16, M, F, LE, LE, LE, OCV_ACT_RELU, p,
             drop ? seeds[2] : 0ull, 1, nullptr, stream));
  // y2 = x1 + dropout2(h W_2^T + b); out = LN2(y2)
  return gemm16(h16, W16[W16_2], 0, 0, W[B_2], x1, W[G_2], W[BE_2], y2, out, out16, M, LE, F, F, F, OCV_ACT_NONE, p,
                drop ? seeds[3] : 0ull, 1, nullptr, stream);
}

extern "C" int ocv_encoder_layer_bwd_train_bf16(const float* const* W, const void* const* W16, float* const* G, const void* x16,
                                                const float* qkv, const void* qkv16, const float* att_stats, const float* att,
                                                const void* att16, const float* y,
                                                const void* x1_16, const float* h, const void* h16, const float* y2,
                                                const uint8_t* mask, const float* dout, float* dx, int B, int S, int F, float drop_p,
                                                const unsigned long long* seeds, void* workspace, void* stream, void* side_stream) {
  OCV_CHECK_ARG(W && W16 && G && x16 && qkv && att && att16 && y && x1_16 && h && h16 && y2 && dout && dx && workspace && B > 0 && S > 0 && F > 0);
  OCV_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || seeds));
  OCV_CHECK_ALIGN((reinterpret_cast<uintptr_t>(workspace) & 255) == 0);
  for (int i = 0; i < N_LAYER_TENSORS; ++i) OCV_CHECK_ARG(W[i] != nullptr && G[i] != nullptr);
  for (int i = 0; i < N_LAYER_W16; ++i) OCV_CHECK_ARG(W16[i] != nullptr);
  const int M = B * S;
  const long long ME = (long long)M * LE, MF = (long long)M * F;
  const bool drop = drop_p > 0.f;
  cudaStream_t st = ocv_stream(stream), sd = side_stream ? ocv_stream(side_stream) : st;
  cudaStream_t sd2 = sd;
  if (side_stream) {
    if (cudaStream_t s2 = second_side_stream()) sd2 = s2;
  }
  // sd2 joins sd at every point sd is ordered after the main stream (and once more at the end): a caller that joins `side_stream`
  // has joined both
  auto fork_side = [&]() -> int {
    TRY(order_after(sd, st));
    if (sd2 != sd) TRY(order_after(sd2, st));
    return OCV_OK;
  };
  uint8_t* p = static_cast<uint8_t*>(workspace);
  auto take = [&](size_t bytes) { void* r = p; p += al(bytes); return r; };
  float* dy2 = static_cast<float*>(take(ME * 4));
  float* dfb = static_cast<float*>(take(ME * 4));
  float* dsab = static_cast<float*>(take(ME * 4));
  float* datt = static_cast<float*>(take(ME * 4));
  float* dh = static_cast<float*>(take(MF * 4));
  float* dh2 = static_cast<float*>(take(MF * 4));
  float* dqkv = static_cast<float*>(take(ME * 3 * 4));
  void* df16 = take(ME * 2);
  void* dsa16 = take(ME * 2);
  void* dh2_16 = take(MF * 2);
  void* dqkv16 = take(ME * 3 * 2);
  void* datt16 = take(ME * 2);
  void* stats = take((size_t)B * LH * S * 3 * 4);
  void* lnws2 = take((size_t)((long long)M * 2 + 32 * LE) * 4);
  void* lnws1 = take((size_t)((long long)M * 2 + 32 * LE) * 4);
  void* ws_side = take((size_t)8 * F * 3 * LE * 4);
  float* dy = dx;  // d loss / d (x + attention branch), becomes d loss / d x after the last accumulation

  const float* df = dy2;                                                           // gradient of linear2's output
  if (drop) {
    TRY(ocv_layernorm_bwd_dropout_f32(dout, y2, W[G_2], dy2, dfb, M, LE, drop_p, seeds[3], lnws2, st));
    df = dfb;
  } else {
    TRY(ocv_layernorm_bwd_f32(dout, y2, W[G_2], dy2, nullptr, nullptr, M, LE, 1.f, lnws2, st));
  }
  TRY(ocv_cast_f32_to_bf16(df, df16, ME, st));
  TRY(fork_side());
  TRY(ocv_layernorm_bwd_params_f32(dout, y2, G[G_2], G[BE_2], M, LE, 1.f, lnws2, sd2));
  TRY(wgrad16(df16, df, h16, G[W_2], G[B_2], M, LE, F, ws_side, sd, sd2));                                   // dW2 [128,F] += df^T h
  // dh [M,F] = df [M,128] W_2 [128,F]   (B = W_2 as stored: contraction index = row -> MN-major)
  TRY(gemm16(df16, W16[W16_2], 0, 1, nullptr, nullptr, nullptr, nullptr, nullptr, dh, nullptr, M, F, LE, LE, F, OCV_ACT_NONE, 0.f, 0ull,
             1, nullptr, st));
  if (drop) TRY(ocv_dropout_act_bwd_f32(dh, h, dh2, MF, OCV_ACT_RELU, drop_p, seeds[2], st));   // h is post-dropout: h > 0 <=> kept & positive
  else TRY(ocv_act_bwd_f32(dh, h, dh2, MF, OCV_ACT_RELU, st));
  TRY(ocv_cast_f32_to_bf16(dh2, dh2_16, MF, st));
  TRY(fork_side());
  TRY(wgrad16(dh2_16, dh2, x1_16, G[W_1], G[B_1], M, F, LE, ws_side, sd, sd2));                        // dW1 [F,128] += dh2^T x1
  if (!drop) {
    TRY(order_after(st, sd));
    TRY(order_after(st, sd2));
  }     // the side stream's bias-gradient sums read df = dy2, which is updated in place next
  // dy2 += dh2 [M,F] W_1 [F,128]        (residual + FFN branch; in place: the tile is read before it is written)
  TRY(gemm16(dh2_16, W16[W16_1], 0, 1, nullptr, dy2, nullptr, nullptr, nullptr, dy2, nullptr, M, LE, F, F, LE, OCV_ACT_NONE, 0.f, 0ull,
             1, nullptr, st));
  const float* dsa = dy;                                                           // gradient of out_proj's output
  if (drop) {
    TRY(ocv_layernorm_bwd_dropout_f32(dy2, y, W[G_1], dy, dsab, M, LE, drop_p, seeds[1], lnws1, st));
    dsa = dsab;
  } else {
    TRY(ocv_layernorm_bwd_f32(dy2, y, W[G_1], dy, nullptr, nullptr, M, LE, 1.f, lnws1, st));
  }
  TRY(ocv_cast_f32_to_bf16(dsa, dsa16, ME, st));
  TRY(fork_side());
  TRY(ocv_layernorm_bwd_params_f32(dy2, y, G[G_1], G[BE_1], M, LE, 1.f, lnws1, sd2));
  TRY(wgrad16(dsa16, dsa, att16, G[W_OUT], G[B_OUT], M, LE, LE, ws_side, sd, sd2));                         // dWout += dsa^T att
  // attention core backward on tcgen05 when the forward ran there (its statistics and bf16 qkv were kept) and the shape fits
  const bool tc_att = qkv16 && att_stats && ocv_mha_bwd_tc_supported(S, S, LH, LD) &&
                      ocv_mha_tc_supported(S, S, LH, LD, (long long)S * 3 * LE, 3 * LE, (long long)S * 3 * LE, 3 * LE, (long long)S * 3 * LE,
                                           3 * LE, (long long)S * LE, LE);   // i.e. the forward ran on the tensor-core kernel too
  const long long qs = (long long)S * 3 * LE;
  if (tc_att) {
    TRY(gemm16(dsa16, W16[W16_OUT], 0, 1, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, datt16, M, LE, LE,
               LE, LE, OCV_ACT_NONE, 0.f, 0ull, 1, nullptr, st));                                       // datt = dsa W_out
    const __nv_bfloat16* q16 = static_cast<const __nv_bfloat16*>(qkv16);
    __nv_bfloat16* g16 = static_cast<__nv_bfloat16*>(dqkv16);
    TRY(ocv_mha_core_bwd_tc_bf16(q16, q16 + LE, q16 + 2 * LE, att16, datt16, att_stats, mask, dqkv, dqkv + LE, dqkv + 2 * LE, g16, g16 + LE,
                                 g16 + 2 * LE, B, S, S, LH, LD, qs, 3 * LE, qs, 3 * LE, (long long)S * LE, LE, qs, 3 * LE,
                                 drop ? drop_p : 0.f, drop ? seeds[0] : 0ull, st));
  } else {
    TRY(gemm16(dsa16, W16[W16_OUT], 0, 1, nullptr, nullptr, nullptr, nullptr, nullptr, datt, nullptr, M, LE, LE, LE, LE, OCV_ACT_NONE,
               0.f, 0ull, 1, nullptr, st));                                                              // datt = dsa W_out
    TRY(ocv_mha_core_bwd_dropout_f32(qkv, qkv + LE, qkv + 2 * LE, att, datt, mask, dqkv, dqkv + LE, dqkv + 2 * LE, B, S, S, LH, LD, qs,
                                     3 * LE, qs, 3 * LE, qs, 3 * LE, (long long)S * LE, LE, qs, 3 * LE, qs, 3 * LE, qs, 3 * LE,
                                     drop ? drop_p : 0.f, drop ? seeds[0] : 0ull, stats, st));
    TRY(ocv_cast_f32_to_bf16(dqkv, dqkv16, ME * 3, st));
  }
  TRY(fork_side());
  TRY(wgrad16(dqkv16, dqkv, x16, G[W_IN], G[B_IN], M, 3 * LE, LE, ws_side, sd, sd2));                  // dWin [384,128] += dqkv^T x
  if (sd2 != sd) TRY(order_after(sd, sd2));                                        // whoever joins `side_stream` has joined both queues
  if (!drop) TRY(order_after(st, sd));                                             // dsa aliases dy, updated in place next
  // dy += dqkv [M,384] W_in [384,128]   (residual + attention branch)
  return gemm16(dqkv16, W16[W16_IN], 0, 1, nullptr, dy, nullptr, nullptr, nullptr, dy, nullptr, M, LE, 3 * LE, 3 * LE, LE, OCV_ACT_NONE,
                0.f, 0ull, 1, nullptr, st);
}
