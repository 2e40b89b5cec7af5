/*
 * objcavit_b200 — C ABI of the B200 (sm_100a) ObjCAViT / AdaBins depth-head kernels.
 *
 * This is the drop-in boundary.  The reference has no native layer: its head is a
 * chain of PyTorch nn.Module calls.  Each entry point below replaces the library calls of one stage of
 * that chain; the "replaces" line names the reference site (path:line under the reference tree).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host; tensors are dense row-major in
 *     the stated layout; fp32 unless stated; `stream` is a cudaStream_t passed as void*.
 *   - return value: 0 = OK; negative = OCV_E_* (bad argument / unsupported shape / wrong arch);
 *     positive = the cudaError_t of the failed launch.  Never throws, never calls exit().
 *   - no entry point synchronises the device or allocates device memory; workspaces are caller-owned.
 *   - all entry points are re-entrant; there is no global mutable state apart from the per-process
 *     cache of function attributes (max dynamic smem) that is set once.
 *   - there is NO CPU fallback: on a device that is not sm_100 every compute entry returns OCV_E_ARCH.
 */
#ifndef OBJCAVIT_B200_H_
#define OBJCAVIT_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OCV_OK 0
#define OCV_E_ARG (-1)      /* null pointer / non-positive size            */
#define OCV_E_SHAPE (-2)    /* shape the kernel does not support            */
#define OCV_E_ALIGN (-3)    /* pointer or stride alignment                  */
#define OCV_E_ARCH (-4)     /* device is not sm_100                         */
#define OCV_E_DRIVER (-5)   /* driver entry point (tensor map encode) unavailable */
#define OCV_E_CAPACITY (-6) /* larger than the fixed scratch built into a *_noalloc_* entry point */

/* activation codes for ocv_linear_f32 */
#define OCV_ACT_NONE 0
#define OCV_ACT_RELU 1
#define OCV_ACT_LEAKY_RELU 2 /* slope 0.01 (nn.LeakyReLU default) */

/* bin-width normalisation (mViT/ObjCAViT `norm` argument) */
#define OCV_NORM_LINEAR 0
#define OCV_NORM_SOFTMAX 1
#define OCV_NORM_SIGMOID 2
#define OCV_NORM_NONE 3 /* input is already normalised widths: only edges / centres are produced */

/* ---- library / device --------------------------------------------------------------------------- */
int ocv_abi_version(void);                       /* bumps when a signature changes */
int ocv_device_check(int device);                /* OCV_OK iff `device` is compute capability 10.x */
const char* ocv_status_string(int status);

/* ---- generic fp32 dense stage --------------------------------------------------------------------
 * C[b] = act(A[b] @ op(Bm[b]) + bias) + residual[b]
 *   A: [M,K] lda; Bm: transB ? [N,K] ldb (nn.Linear weight layout) : [K,N] ldb; C,residual: [M,N] ldc.
 *   batch strides in elements (0 = shared operand).  bias [N] or NULL; residual NULL or same layout as C.
 * replaces: nn.Linear / F.linear everywhere on the token path (modules/ObjCAViT.py:257-282,289,299-303;
 *   modules/miniViT.py:16-20), the in/out projections and FFN of nn.TransformerEncoderLayer
 *   (modules/layers.py:8; modules/ObjCAViT.py:155-164), torch.matmul in PixelWiseDotProduct
 *   (modules/layers.py:35, transB=0 gives the [n,q,h*w] result directly, no permute copies). */
int ocv_linear_f32(const float* A, const float* Bm, const float* bias, const float* residual, float* C,
                   int M, int N, int K, int lda, int ldb, int ldc, int transB, int act,
                   int batch, long long strideA, long long strideB, long long strideC, void* stream);

/* ---- patch embedding -----------------------------------------------------------------------------
 * tokens[b,s,e] = sum_{c,ky,kx} W[e,c,ky,kx] * X[b,c,ty*P+ky,tx*P+kx] + bias[e] + pos[s,e]   (s = ty*gw+tx)
 *   X [B,C,h,w] NCHW; W [E,C,P,P]; pos [>=S,E] row-major or NULL; tokens [B,S,E]; gh=h/P, gw=w/P (floor).
 *   workspace: ocv_patch_embed_f32_workspace() bytes (split-K partials; deterministic reduction).
 * replaces: embedding_convPxP + positional add + permute (modules/layers.py:17-22) and
 *   image_embedding_convPxP (+ the image positional add) (modules/ObjCAViT.py:333,362-364). */
long long ocv_patch_embed_f32_workspace(int B, int C, int h, int w, int P, int E);
int ocv_patch_embed_f32(const float* X, const float* W, const float* bias, const float* pos, float* tokens,
                        int B, int C, int h, int w, int P, int E, void* workspace, void* stream);

/* ---- transformer pieces --------------------------------------------------------------------------
 * y = LayerNorm(x + r) * gamma + beta over the last dim E (eps 1e-5); r may be NULL.  rows = B*S.
 * zero_rows (uint8 [rows], may be NULL): rows whose flag is non-zero are written as 0 — the padded rows
 *   of nn.TransformerEncoder's inference fast path.
 * replaces: norm1/norm2 of nn.TransformerEncoderLayer (post-norm). */
int ocv_add_layernorm_f32(const float* x, const float* r, const float* gamma, const float* beta, float* y,
                          int rows, int E, const uint8_t* zero_rows, void* stream);

/* Scaled-dot-product attention core for nn.MultiheadAttention:
 *   out[b,i,h*d:(h+1)*d] = softmax_j( q[b,i,h]·k[b,j,h] / sqrt(d) + (mask[b,j] ? -inf : 0) ) @ v[b,j,h]
 *   q/k/v/out are [B,S,*] views with explicit row strides (elements) so they can alias a packed QKV
 *   buffer; heads*d contiguous channels starting at the pointer.  d must be 32.  mask uint8 [B,Sk] or NULL.
 * replaces: the attention inside F.multi_head_attention_forward (modules/ObjCAViT.py:195-207 and every
 *   encoder layer). */
int ocv_mha_core_f32(const float* q, const float* k, const float* v, float* out, const uint8_t* key_padding_mask,
                     int B, int Sq, int Sk, int heads, int d,
                     long long q_bstride, int q_rstride, long long k_bstride, int k_rstride,
                     long long v_bstride, int v_rstride, long long o_bstride, int o_rstride, void* stream);

/* ---- object-token bookkeeping (bit-exact index work) ----------------------------------------------
 * Ragged -> padded: src [total,E] rows of image b are [offsets[b],offsets[b+1]) (CSR, int32, device).
 *   dst [B,L,E] filled with pad_value, image b's rows copied to [left, left+n_b) with left = L-Nmax when
 *   left_pad else 0;   mask [B,L] uint8 (optional) = 1 except [0,m_b) (i.e. ALWAYS right-padded)
 *   where m_b comes from mask_offsets (CSR, optional; NULL -> m_b = n_b).
 * replaces: pad_sequence x2 + F.pad x2 (modules/ObjCAViT.py:180-183,192-194). */
int ocv_pad_objects_f32(const float* src, const int32_t* offsets, const int32_t* mask_offsets, float* dst,
                        uint8_t* mask, int B, int L, int Nmax, int E, float pad_value, int left_pad, void* stream);

/* patch-centre coordinates coords[s] = (16x+8, 16y+8, 16, 16), s = y*gw+x  -> [S,4]
 * replaces: modules/ObjCAViT.py:336-347 (batch-invariant, so computed once, not per image). */
int ocv_patch_coords_f32(float* coords, int gh, int gw, int P, void* stream);

/* GridRandomPositionalEmbeddings.  table [rows,E]; grid = first gh*gw rows as [gh,gw,E].
 * centre mode: bilinear grid_sample (zeros padding, align_corners=False) at points pts [N,2] (x,y) -> out [N,E];
 *   norm_mode 0: pts already normalised to [-1,1];
 *   norm_mode 1 ("obj", modules/ObjCAViT.py:104-105): x' = (x/div0)*2-1, y' = (y/div1)*2-1
 *               (the caller passes div0 = full-res HEIGHT, div1 = full-res WIDTH: quirk F13b is the caller's);
 *   norm_mode 2 ("img", modules/ObjCAViT.py:95-96, quirk F13a): point 0 -> both coords (c/div0)*2-1,
 *               point 1 -> both coords (c/div1)*2-1, every other point keeps its raw pixel coordinates.
 *   idx_out [N,4] int32 (nw,ne,sw,se flat index or -1) optional: the bit-exact index mapping.
 * roi mode: ps_roi_align(output 1x1, sampling_ratio=-1) -> out [N,E].  boxes [N,4]: corner boxes (xyxy) or, when
 *   boxes_are_xywh, centre/size boxes converted as modules/ObjCAViT.py:115-124 (corners, clamp(min=0)).
 *   Zero-area boxes give NaN exactly like torchvision.  plan_out [N,2] int32 (grid_h,grid_w) optional.
 * replaces: F.grid_sample (modules/ObjCAViT.py:99,109) / torchvision.ops.ps_roi_align (:128,144). */
int ocv_grid_sample_pos_f32(const float* table, const float* pts, float* out, int32_t* idx_out,
                            int N, int gh, int gw, int E, int norm_mode, float div0, float div1, void* stream);
int ocv_ps_roi_align_pos_f32(const float* table, const float* boxes, float* out, int32_t* plan_out,
                             int N, int gh, int gw, int E, float spatial_scale, int boxes_are_xywh, void* stream);

/* ---- bins -----------------------------------------------------------------------------------------
 * y_raw [B,K] (regressor output) -> widths_normed [B,K] (norm: relu+0.1 | softmax | sigmoid, then /sum),
 *   edges [B,K+1] = cumsum(pad(min_depth, (max-min)*widths)), centres [B,K] = mid-points.  K <= 1024.
 * replaces: modules/miniViT.py:34-42 ≡ modules/ObjCAViT.py:379-388 and modules/AdaBins.py:79-85. */
int ocv_bins_f32(const float* y_raw, float* widths_normed, float* edges, float* centres,
                 int B, int K, int norm, float min_depth, float max_depth, void* stream);

/* ---- pixel path, fp32 (parity path) ---------------------------------------------------------------
 * 3x3 / pad 1 / stride 1 convolution, NCHW fp32, C_in = C_out = 128.
 * replaces: self.conv3x3 (modules/miniViT.py:25; modules/ObjCAViT.py:374). */
int ocv_conv3x3_f32(const float* X, const float* W, const float* bias, float* Y,
                    int B, int C, int h, int w, void* stream);

/* depth[b,p] = sum_k softmax_k( Wout[k,:]·ram[b,:,p] + bout[k] ) * centres[b,k];  ram [B,Q,P] (NCHW).
 * The [B,K,h,w] probability volume is never written to memory.
 * replaces: conv_out (1x1 conv + Softmax) and torch.sum(out*centers) (modules/AdaBins.py:77,87). */
int ocv_tail_f32(const float* ram, const float* Wout, const float* bout, const float* centres, float* depth,
                 int B, int Q, int K, long long P, void* stream);

/* ---- pixel path, bf16 tensor-core (throughput path) -----------------------------------------------
 * NCHW fp32|bf16 -> NHWC bf16 feature map (the layout the tcgen05 kernels consume).  src_is_bf16 selects
 * the source dtype. */
int ocv_nchw_to_nhwc_bf16(const void* src, void* dst_bf16, int B, int C, int h, int w, int src_is_bf16, void* stream);

/* Generic bf16 tensor-core GEMM (tcgen05.mma, TMA-fed, fp32 accumulation in TMEM):
 *   C[M,N] = act(A[M,K] @ Wt[N,K]^T + bias) (+ residual),  A/Wt bf16 row-major (K contiguous),
 *   C fp32 or bf16 (c_is_bf16), ldc in elements.  K % 64 == 0, N % 16 == 0, N <= 256 per tile column.
 *   split_k > 1 writes fp32 partials to `workspace` ([split_k,M,N]) reduced deterministically.
 * replaces: the same sites as ocv_linear_f32 on the throughput path, and the patch-embed GEMM. */
long long ocv_gemm_bf16_workspace(int M, int N, int K, int split_k);
int ocv_gemm_bf16(const void* A, const void* Wt, const float* bias, const float* residual, void* C,
                  int M, int N, int K, int lda, int ldb, int ldc, int act, int c_is_bf16, int split_k,
                  void* workspace, void* stream);

/* ---- bf16 token path ------------------------------------------------------------------------------
 * Attention core on bf16 operands (tensor-core scores and P@V, fp32 online softmax); same argument meaning as
 * ocv_mha_core_f32 with q/k/v/out bf16 (strides in elements, multiples of 8; q/k/v 16-byte aligned). */
int ocv_mha_core_bf16(const void* q, const void* k, const void* v, void* out, const uint8_t* key_padding_mask,
                      int B, int Sq, int Sk, int heads, int d,
                      long long q_bstride, int q_rstride, long long k_bstride, int k_rstride,
                      long long v_bstride, int v_rstride, long long o_bstride, int o_rstride, void* stream);
/* The same attention core on tcgen05 (csrc/attn_tc.cu; ocv_mha_core_bf16 dispatches to it when the shape is covered: 4 heads x 32,
 * Sk <= 448, strides multiples of 8 elements, out 16-byte aligned).  One CTA = 128 query rows x a pair of heads; S = Q K^T and
 * O = P V as tcgen05.mma with fp32 accumulators in TMEM, V read as an MN-major operand.  Extras for the training step: out32 (fp32
 * copy of the output, same strides; or NULL), stats [B,heads,Sq,2] = (row max, row sum of 2^(s - max)) of the scaled log2-domain
 * scores (or NULL), and dropout of the attention probabilities after the normaliser (drop_p, drop_seed: the mask of element
 * ((b*heads + h)*Sq + q)*Sk + k, as ocv_mha_core_dropout_f32). */
int ocv_mha_core_tc_bf16(const void* q, const void* k, const void* v, void* out16, float* out32, float* stats,
                         const uint8_t* key_padding_mask, int B, int Sq, int Sk, int heads, int d, long long q_bstride,
                         int q_rstride, long long k_bstride, int k_rstride, long long v_bstride, int v_rstride, long long o_bstride,
                         int o_rstride, float drop_p, unsigned long long drop_seed, void* stream);
/* Backward of the attention core on tcgen05 (one CTA per (image, head); Sq, Sk <= 256): from q / k / v (bf16), the forward output
 * o16 and its gradient do16 (bf16, strides o_*), and the forward's statistics: dq / dk / dv in fp32 (strides g_*) and bf16 copies.
 * S and dO V^T per 128 x 128 tile in TMEM, P and dS recomputed in registers (dropout masks regenerated), dV += Pd^T dO,
 * dK += dS^T Q, dQ += dS K with the bf16 tiles read as MN-major operands.  Nothing of size Sq x Sk reaches HBM. */
int ocv_mha_core_bwd_tc_bf16(const void* q16, const void* k16, const void* v16, const void* o16, const void* do16, const float* stats,
                             const uint8_t* key_padding_mask, float* dq32, float* dk32, float* dv32, void* dq16, void* dk16,
                             void* dv16, int B, int Sq, int Sk, int heads, int d, long long q_bstride, int q_rstride,
                             long long kv_bstride, int kv_rstride, long long o_bstride, int o_rstride, long long g_bstride,
                             int g_rstride, float drop_p, unsigned long long drop_seed, void* stream);

/* Patch embedding from the NHWC bf16 feature map (tcgen05 GEMM, A operand fetched by a 5-D TMA box, split-K over ky):
 *   Wp = ocv_pack_patch_weights_bf16(W [E,C,P,P] fp32) -> bf16 [E, (ky,kx,c)];
 *   tokens[b,s,:] = patch(b,s) . Wp^T + bias + pos[s,:]  -> tok32 fp32 [B,S,E] and/or tok16 bf16 [B,S,E].
 *   E == 128, (P*C) % 64 == 0, w/P <= 128.  workspace: ocv_patch_embed_bf16_workspace() bytes.
 * replaces: the same sites as ocv_patch_embed_f32. */
int ocv_pack_patch_weights_bf16(const float* W, void* Wp, int E, int C, int P, void* stream);
long long ocv_patch_embed_bf16_workspace(int B, int h, int w, int P);
int ocv_patch_embed_bf16(const void* X_nhwc_bf16, const void* Wp, const float* bias, const float* pos, float* tok32,
                         void* tok16, int B, int C, int h, int w, int P, int E, void* workspace, void* stream);

/* Fused MLP chain (fp32 SIMT, one launch): x [R, dims[0]] (row pitch x_ld) -> Linear/LeakyReLU(0.01) x (n_layers-1)
 *   -> Linear (+ residual [R, dims[n_layers]]) -> out32 fp32 and/or out16 bf16.  weights/biases/dims are HOST arrays
 *   (of device pointers / ints); weights[l] is [dims[l+1], dims[l]] (nn.Linear layout); every width <= 512.
 * replaces: the positional_encoder nn.Sequential (modules/ObjCAViT.py:257-282) and the regressor
 *   (modules/miniViT.py:16-20,33; modules/ObjCAViT.py:299-303,378). */
int ocv_mlp_f32(const float* x, long long x_ld, int R, int n_layers, const float* const* weights_host,
                const float* const* biases_host, const int* dims_host, const float* residual, float* out32,
                void* out16, void* stream);
/* One Linear (+ LeakyReLU(0.01) when `leaky`) layer over FEW rows (R x K, K <= 512, K % 4 == 0): output neurons spread over the
 * grid, every weight read once chip-wide (~5 us against 10-45 us for a tiled GEMM at 8-500 rows).  The training step runs the
 * regressor and positional MLPs through it layer by layer, keeping each layer's output for the backward.
 * replaces: nn.Linear (+ nn.LeakyReLU) of self.regressor / self.positional_encoder (modules/ObjCAViT.py:257-282, 299-303). */
int ocv_linear_rows_f32(const float* x, long long x_ld, int R, const float* W, const float* bias, int K, int N, int leaky, float* out,
                        int out_ld, void* stream);
/* Same chain for FEW rows (R ~ batch size: the regressor): one launch per layer, output neurons spread over the grid so each
 * weight is read once.  Every dims[l] % 4 == 0 and <= 512 for l < n_layers; x 16-byte aligned, x_ld % 4 == 0.
 * workspace: ocv_mlp_rows_f32_workspace() bytes (ping-pong activations). */
long long ocv_mlp_rows_f32_workspace(int R, int n_layers, const int* dims_host);
int ocv_mlp_rows_f32(const float* x, long long x_ld, int R, int n_layers, const float* const* weights_host,
                     const float* const* biases_host, const int* dims_host, float* out32, void* workspace, void* stream);
int ocv_cast_f32_to_bf16(const float* src, void* dst_bf16, long long n, void* stream);

/* Training-step form of the tcgen05 GEMM (csrc/tc_gemm.cu): C = [LayerNorm]( dropout(act(op(A) op(B)^T + bias)) + residual ).
 *   a_mn_major = 0: A is [M,K] row-major (lda);  1: A is stored [K,M] row-major (the contraction index is the ROW) and is read as
 *   an MN-major shared-memory operand — no transposed copy.  Same for B: 0 = [N,K] (nn.Linear weight layout), 1 = [K,N].
 *   So dX = dY W is (A = dY, B = W [N_out rows = contraction, K_in cols], b_mn_major = 1) and dW = dY^T X is (A = dY, a_mn_major = 1,
 *   B = X, b_mn_major = 1).  drop_p > 0: keep mask of element m*N+n from (drop_seed ^ salt), as ocv_dropout_f32.
 *   Cpre (with ln_gamma): the pre-normalisation sum, a third output.  split_k > 1: fp32 output only, fixed-order reduction. */
int ocv_gemm_bf16_v2(const void* A, const void* Bt, int a_mn_major, int b_mn_major, const float* bias, const float* residual, int ldr,
                     const float* ln_gamma, const float* ln_beta, float* Cpre, int ldpre, float* C32, int ldc32, void* C16, int ldc16,
                     int M, int N, int K, int lda, int ldb, int act, float drop_p, unsigned long long drop_seed, int split_k,
                     void* workspace, void* stream);
/* Extended epilogue of the same kernel (the bf16 token path):
 *   y = act(A @ Wt^T + bias) + residual (fp32, row pitch ldr);  if ln_gamma: y = LayerNorm_128(y)*gamma+beta (eps 1e-5,
 *   requires N == 128, act none), rows flagged in zero_rows (uint8 [M]) are written as 0.
 *   y is written as fp32 to C32 (row pitch ldc32) and/or as bf16 to C16 (ldc16); either may be NULL, not both.
 * replaces: out_proj + dropout-free residual + norm1, linear2 + residual + norm2 of nn.TransformerEncoderLayer. */
int ocv_gemm_bf16_ex(const void* A, const void* Wt, const float* bias, const float* residual, int ldr,
                     const float* ln_gamma, const float* ln_beta, const uint8_t* zero_rows, float* C32, int ldc32,
                     void* C16, int ldc16, int M, int N, int K, int lda, int ldb, int act, void* stream);

/* Fused pixel kernel (one persistent tcgen05 kernel, objcavit_b200/csrc/pixel_fused_bf16.cu):
 *   conv3x3 (implicit GEMM over an NHWC bf16 halo tile, 9 shifted smem descriptors) -> per-image logits GEMM with
 *   Mb[b] = s * Wout @ Q[b] (bins x channels, bf16) -> per-pixel softmax over K=256 bins -> sum p*centre.
 *   X [B,h,w,128] bf16 NHWC (16-byte aligned); W3p: packed conv weights from ocv_pack_conv3x3_bf16;
 *   Mb [B,256,128] bf16 and blog [B,256] fp32 from ocv_fold_queries_bf16 with logit_scale = log2(e)
 *   (the kernel exponentiates with ex2); centres [B,256] fp32; depth [B,h*w] fp32.
 *   max_ctas: 0 = one persistent CTA per SM; >0 caps the grid (tests exercise multi-tile CTAs with it).
 * replaces: conv3x3 + PixelWiseDotProduct + conv_out + Softmax + sum (modules/ObjCAViT.py:374-375;
 *   modules/GraphBins.py:109,119; modules/miniViT.py:25,31; modules/AdaBins.py:77,87). */
int ocv_pack_conv3x3_bf16(const float* W /*[128,128,3,3]*/, void* W3p /*bf16 [9,128,128]*/, void* stream);
/* Mb[b,k,c] = s * sum_q Wout[k,q] Q[b,q,c];  blog[b,k] = s * (bout[k] + sum_{q,c} Wout[k,q] Q[b,q,c] b3[c]).
 * (conv_out(Q @ (conv + b3)) == (Wout Q) @ conv + blog: the 128 range-attention channels are folded away.) */
int ocv_fold_queries_bf16(const float* Wout /*[K,Q]*/, const float* bout /*[K]*/, const float* queries /*[B,Q,E] ld*/,
                          long long q_bstride, int q_rstride, const float* b3 /*[E]*/, float logit_scale,
                          void* Mb /*bf16 [B,K,E]*/, float* blog /*[B,K]*/, int B, int K, int Q, int E, void* stream);
int ocv_pixel_fused_bf16(const void* X, const void* W3p, const void* Mb, const float* blog, const float* centres,
                         float* depth, int B, int h, int w, int max_ctas, void* stream);

/* The conv3x3 stage of the same kernel on its own (CONV_ONLY instantiation): Y [B,h,w,128] bf16 NHWC = conv3x3(X, pad 1) + bias,
 * fp32 accumulation in TMEM.  X, W3p as above; bias [128] fp32; Y 16-byte aligned.
 * replaces: the decoder's final convolution `DenseFeatureExtractor.conv3` (3x3, 128 -> 128, no activation;
 *   modules/DenseFeatureExtractor.py:97,116) so that the feature map is born in the layout / dtype the head consumes
 *   (no fp32 NCHW round trip between decoder and head). */
int ocv_conv3x3_nhwc_bf16(const void* X, const void* W3p, const float* bias, void* Y, int B, int h, int w, int max_ctas,
                          void* stream);

/* Same kernel, additionally writing the bin with the largest logit per pixel (uint8 [B,h*w], first maximum): used by the
 * parity tests for the north star's "identical bin argmax" criterion; not on the throughput path. */
int ocv_pixel_fused_bf16_argmax(const void* X, const void* W3p, const void* Mb, const float* blog, const float* centres,
                                float* depth, uint8_t* argmax_bin, int B, int h, int w, void* stream);

/* ---- backward pass of the fp32 parity path (config 5) ----------------------------------------------------------------
 * Every gradient entry writes  out = beta * out + value  where stated, so parameter gradients can accumulate in place.
 *
 * C[M,N] = beta*C + sum_{batch (if reduce_batch), k} op(A)[M,K] @ op(B)[K,N]:
 *   transA: element (m,k) of op(A) is A[k*lda+m] (weight gradients dW = dY^T X) instead of A[m*lda+k];
 *   opB: 0 = B[k*ldb+n];  1 = B[n*ldb+k];
 *        2 = gathered patches of an NCHW map, B[m][k=(c,ky,kx)] = X[b,c,ty*P+ky,tx*P+kx] (patch-embedding weight gradient);
 *        3 = gathered 3x3 neighbourhoods of ONE image (batch through strideB), B[p][j=(c,ky,kx)] = X[c,y+ky-1,x+kx-1] or 0
 *            (conv3x3 weight gradient);  gather_geom_host = {C, h, w, P} for opB >= 2.
 *   split_k > 1 / reduce_batch / beta != 0 go through `workspace` (ocv_gemm_f32_ex_workspace bytes), added in a fixed order.
 * replaces: the autograd of nn.Linear / nn.Conv2d / torch.matmul on the path. */
/* nn.Linear forward C[M,N] = A[M,K] @ W[N,K]^T + bias with split-K (fixed-order reduction): for long-K / few-tile problems
 * (linear2 of nn.TransformerEncoderLayer in the training forward).  workspace: split_k * M * N floats. */
int ocv_linear_splitk_f32(const float* A, const float* W, const float* bias, float* C, int M, int N, int K, int lda, int ldb,
                          int ldc, int split_k, void* workspace, void* stream);
long long ocv_gemm_f32_ex_workspace(int M, int N, int batch, int split_k);
int ocv_gemm_f32_ex(int transA, int opB, const float* A, const float* Bm, float* C, int M, int N, int K, int lda, int ldb, int ldc,
                    float beta, int batch, long long strideA, long long strideB, long long strideC, int reduce_batch, int split_k,
                    const int* gather_geom_host, void* workspace, void* stream);
/* x viewed [groups, outer, C, inner] -> out[g,c] = beta*out[g,c] + sum_{o,i} x[g,o,c,i]   (bias / centre gradients) */
int ocv_reduce_channels_f32(const float* x, float* out, int groups, int outer, int C, long long inner, float beta, void* stream);
/* dx = dy * act'(.) from the activation OUTPUT (ReLU / LeakyReLU 0.01 keep the sign) */
int ocv_act_bwd_f32(const float* dy, const float* out, float* dx, long long n, int act, void* stream);
/* LayerNorm backward: z = pre-norm input [rows,E]; dz [rows,E]; dgamma/dbeta [E] (optional pair, beta-accumulated);
 * workspace rows*2 + 32*E floats (row statistics + fixed-order partials of the parameter sums). */
int ocv_layernorm_bwd_f32(const float* dy, const float* z, const float* gamma, float* dz, float* dgamma, float* dbeta, int rows,
                          int E, float beta, void* workspace, void* stream);
/* Data half of the LayerNorm backward fused with the dropout site behind it (backward of a train()-mode encoder layer):
 * dz as ocv_layernorm_bwd_f32, dz_dropped = dropout(dz) — bit-identical to ocv_dropout_f32(dz); row statistics into `workspace`
 * (rows*2 + 32*E floats) for a following ocv_layernorm_bwd_params_f32. */
int ocv_layernorm_bwd_dropout_f32(const float* dy, const float* z, const float* gamma, float* dz, float* dz_dropped, int rows, int E,
                                  float p, unsigned long long seed, void* workspace, void* stream);
/* The parameter half of ocv_layernorm_bwd_f32 alone: call ocv_layernorm_bwd_f32 with dgamma = dbeta = NULL first (data gradient +
 * row statistics into `workspace`), then this on the same dy / z / workspace — possibly on another stream, ordered after it. */
int ocv_layernorm_bwd_params_f32(const float* dy, const float* z, float* dgamma, float* dbeta, int rows, int E, float beta,
                                 void* workspace, void* stream);
/* ocv_layernorm_bwd_f32 for hosts that may not allocate device memory.  (No counterpart in the reference: there torch's caching
 * allocator provides every temporary.  This exists for the integration cases listed below, nothing in the package calls it.)
 *   Who needs it: code that receives its input and output buffers from someone else and has no way to obtain more — an operator
 *   plugin inside an inference runtime whose memory plan is fixed when the engine is built, a callback entered while the host
 *   framework holds its allocator lock, a control loop whose rules ban cudaMalloc once it is running (cudaMalloc synchronises the
 *   device and can fail at an arbitrary later time).  Such a caller cannot produce the `workspace` argument of
 *   ocv_layernorm_bwd_f32, so this variant takes none: the rows*2 + 32*E floats of scratch come from an array compiled into the
 *   library.  The entry point calls no cudaMalloc / cudaFree and no library routine that allocates.
 *   Cost and limit: the array holds OCV_LN_BWD_NOALLOC_MAX_ROWS rows at E <= 512 = 192 KB, resident on every GPU that loads the
 *   library until the process ends.  rows > OCV_LN_BWD_NOALLOC_MAX_ROWS returns OCV_E_CAPACITY and launches nothing.
 *   First call: the first call on each device (with that device current) looks up the array's device address; make that call
 *   during set-up, before any stream capture begins on that device.
 *   Sharing: all calls on one device use the same array, and it keeps whatever the previous call left in it.  No part of it needs
 *   zeroing (every slot is written before it is read within a call), but calls that overlap on the GPU — two streams, two
 *   threads — race on it and give wrong gradients.  Serialise them (one stream, or events).
 *   Arguments, shapes and results are those of ocv_layernorm_bwd_f32 (bit-identical output). */
#define OCV_LN_BWD_NOALLOC_MAX_ROWS 16384
int ocv_layernorm_bwd_noalloc_f32(const float* dy, const float* z, const float* gamma, float* dz, float* dgamma, float* dbeta,
                                  int rows, int E, float beta, void* stream);
/* Backward of ocv_mha_core_f32 (same view conventions; `out` = its forward result): dq, dk, dv.  workspace B*heads*Sq*3 floats. */
int ocv_mha_core_bwd_f32(const float* q, const float* k, const float* v, const float* out, const float* dout,
                         const uint8_t* key_padding_mask, float* dq, float* dk, float* dv, int B, int Sq, int Sk, int heads, int d,
                         long long q_bstride, int q_rstride, long long k_bstride, int k_rstride, long long v_bstride, int v_rstride,
                         long long o_bstride, int o_rstride, long long dq_bstride, int dq_rstride, long long dk_bstride,
                         int dk_rstride, long long dv_bstride, int dv_rstride, void* workspace, void* stream);
/* Backward of depth = sum_k softmax_k(l + bias) c_k per pixel: logits l [B,K,P] are overwritten by dl = g p (c - depth);
 * pg [B,K,P] = g p (its sum over pixels is d loss / d centres).   g = grad_depth [B,P];  logit_bias [K] or NULL (the conv_out
 * bias, so that the logits operand can be the bare GEMM Wout @ ram). */
int ocv_tail_bwd_f32(float* logits_inout, float* pg_out, const float* centres, const float* grad_depth, const float* logit_bias,
                     int B, int K, long long P, void* stream);
/* Dropout of training mode (nn.TransformerEncoderLayer: dropout 0.1 on the attention probabilities, after out_proj, after the
 * FFN activation and after linear2).  Counter-based masks: element idx of site `seed` is kept iff
 * hash(seed, idx) >= p 2^32 (csrc/common.cuh), kept values are scaled by 1/(1-p).  Same statistics as torch's Bernoulli masks, a
 * different random stream (torch's Philox offsets cannot be reproduced); the backward pass regenerates the forward mask.
 *   ocv_dropout_f32: out = residual + dropout(x) (residual optional, in place allowed); applied to a gradient it is its own backward.
 *   ocv_mha_core_dropout_f32 / ocv_mha_core_bwd_dropout_f32: the attention core with dropout on the probabilities, mask index =
 *   flat index into [B, heads, Sq, Sk] (so ocv_dropout_f32 on a tensor of ones of that shape reproduces the mask: tests). */
int ocv_dropout_f32(const float* x, const float* residual, float* out, long long n, float p, unsigned long long seed, void* stream);
/* Fused forms used by the native encoder-layer sequencer (bit-identical to the two-call sequences they replace):
 *   ocv_dropout_add_layernorm_f32: sum = residual + dropout(x) over the flat [rows,E] index, y = LayerNorm(sum) * gamma + beta
 *     (`sum` is kept: the LayerNorm backward reads it);  replaces dropout1 + norm1 / dropout2 + norm2 of
 *     nn.TransformerEncoderLayer in train() mode.
 *   ocv_dropout_act_bwd_f32: dx = act'(out) * dropout(dy)  (the FFN dropout site folded into the activation backward). */
int ocv_dropout_add_layernorm_f32(const float* x, const float* residual, const float* gamma, const float* beta, float* sum, float* y,
                                  int rows, int E, float p, unsigned long long seed, void* stream);
int ocv_dropout_act_bwd_f32(const float* dy, const float* out, float* dx, long long n, int act, float p, unsigned long long seed,
                            void* stream);
/* Per-step salt of every dropout mask: `device_salt` points at ONE 64-bit word in device memory (or NULL = none, the default);
 * each dropout kernel uses seed ^ *device_salt.  It exists for CUDA-graph replay of the training step, where the site seeds are
 * frozen kernel arguments: the caller changes the word between replays.  Process-wide (one process per GPU). */
int ocv_set_dropout_salt(const unsigned long long* device_salt);
/* Clear the salt only if `owner` is the pointer currently installed (a trainer going away must not disarm another one's). */
int ocv_release_dropout_salt(const unsigned long long* owner);
int ocv_mha_core_dropout_f32(const float* q, const float* k, const float* v, float* out, const uint8_t* key_padding_mask,
                             int B, int Sq, int Sk, int heads, int d, long long q_bstride, int q_rstride, long long k_bstride,
                             int k_rstride, long long v_bstride, int v_rstride, long long o_bstride, int o_rstride, float drop_p,
                             unsigned long long drop_seed, void* stream);
int ocv_mha_core_bwd_dropout_f32(const float* q, const float* k, const float* v, const float* out, const float* dout,
                                 const uint8_t* key_padding_mask, float* dq, float* dk, float* dv, int B, int Sq, int Sk, int heads,
                                 int d, long long q_bstride, int q_rstride, long long k_bstride, int k_rstride, long long v_bstride,
                                 int v_rstride, long long o_bstride, int o_rstride, long long dq_bstride, int dq_rstride,
                                 long long dk_bstride, int dk_rstride, long long dv_bstride, int dv_rstride, float drop_p,
                                 unsigned long long drop_seed, void* workspace, void* stream);

/* ---- native sequencer of one encoder layer of the training step (objcavit_b200/csrc/train_layer.cu) ----------------------
 * One call = the forward (or the explicit backward) of one nn.TransformerEncoderLayer (post-norm, ReLU, E = 128, 4 heads, FFN
 * width F; modules/layers.py:8, modules/ObjCAViT.py:155-161) in train() mode, built from the entry points above in the order
 * autograd would run them.  W / G: HOST arrays of 12 DEVICE pointers (parameters / their gradient buffers, accumulated) in the
 * order in_proj_weight, in_proj_bias, out_proj.weight, out_proj.bias, linear1.weight, linear1.bias, linear2.weight,
 * linear2.bias, norm1.weight, norm1.bias, norm2.weight, norm2.bias.  x [B*S,128]; mask uint8 [B,S] or NULL; drop_p + seeds (HOST
 * array of 4 site seeds: attention probabilities, dropout1, dropout, dropout2; ignored when drop_p == 0).
 * Forward outputs = what the backward reads: qkv [B*S,384], att, y (pre-norm1 sum), x1, h [B*S,F] (post ReLU / dropout), y2
 * (pre-norm2 sum), out.  Backward: dout [B*S,128] -> dx [B*S,128]; weight / bias gradients run on side_stream (NULL = same
 * stream), ordered with events; the caller makes its stream wait for side_stream before it reads G. */
long long ocv_encoder_layer_fwd_train_workspace(int B, int S, int F);
int ocv_encoder_layer_fwd_train_f32(const float* const* W, const float* x, const uint8_t* mask, int B, int S, int F, float drop_p,
                                    const unsigned long long* seeds, float* qkv, float* att, float* y, float* x1, float* h, float* y2,
                                    float* out, void* workspace, void* stream);
long long ocv_encoder_layer_bwd_train_workspace(int B, int S, int F);
int ocv_encoder_layer_bwd_train_f32(const float* const* W, float* const* G, const float* x, const float* qkv, const float* att,
                                    const float* y, const float* x1, const float* h, const float* y2, const uint8_t* mask,
                                    const float* dout, float* dx, int B, int S, int F, float drop_p, const unsigned long long* seeds,
                                    void* workspace, void* stream, void* side_stream);
/* Tensor-core form of the two calls above (training step, precision "bf16"; csrc/train_layer.cu): all GEMMs of the layer — forward,
 * data gradients, weight gradients — through ocv_gemm_bf16_v2, the dropout / residual / LayerNorm sites in its epilogue.
 * W16[4]: bf16 copies of in_proj_weight, out_proj.weight, linear1.weight, linear2.weight (nn.Linear layout).  x16 / att16 / x1_16 /
 * h16 / out16: bf16 copies of the corresponding activations ([B*S, 128 | F]) — outputs of the forward, inputs of the backward. */
int ocv_encoder_layer_fwd_train_bf16(const float* const* W, const void* const* W16, const float* x, const void* x16,
                                     const uint8_t* mask, int B, int S, int F, float drop_p, const unsigned long long* seeds,
                                     float* qkv, void* qkv16 /* [B*S,384] bf16, or NULL: fp32 attention core */,
                                     float* att_stats /* [B,4,S,2] softmax statistics of the tcgen05 attention core, or NULL */,
                                     float* att, void* att16, float* y, float* x1, void* x1_16, float* h, void* h16, float* y2,
                                     float* out, void* out16, void* stream);
long long ocv_encoder_layer_bwd_train_bf16_workspace(int B, int S, int F);
int ocv_encoder_layer_bwd_train_bf16(const float* const* W, const void* const* W16, float* const* G, const void* x16, const float* qkv,
                                     const void* qkv16, const float* att_stats /* both from the forward, or NULL: fp32 attention backward */,
                                     const float* att, const void* att16, const float* y, const void* x1_16, const float* h,
                                     const void* h16, const float* y2, const uint8_t* mask, const float* dout, float* dx, int B, int S,
                                     int F, float drop_p, const unsigned long long* seeds, void* workspace, void* stream,
                                     void* side_stream);


/* Backward of ocv_grid_sample_pos_f32 / ocv_ps_roi_align_pos_f32 w.r.t. the positional table (training with the grid_random /
 * grid_random_roi_align strategies; what autograd does through F.grid_sample / torchvision.ops.ps_roi_align in
 * modules/ObjCAViT.py:99,109,128,144).  Same argument meaning as the forward; dout [N,E].  The kernels scatter into
 * table_grad_fix [gh*gw, E] int64 (fixed point, 2^40 per unit; zero it first; integer atomics -> deterministic);
 * ocv_table_grad_commit_f32 then does table_grad[i] = beta*table_grad[i] + fix[i] / 2^40. */
int ocv_grid_sample_pos_bwd_f32(const float* dout, const float* pts, long long* table_grad_fix, int N, int gh, int gw, int E,
                                int norm_mode, float div0, float div1, void* stream);
int ocv_ps_roi_align_pos_bwd_f32(const float* dout, const float* boxes, long long* table_grad_fix, int N, int gh, int gw, int E,
                                 float spatial_scale, int boxes_are_xywh, void* stream);
int ocv_table_grad_commit_f32(const long long* table_grad_fix, float* table_grad, long long n, float beta, void* stream);

/* ---- training step, tensor-core pixel path (precision "bf16"; objcavit_b200/csrc/train_pixel_bf16.cu) ----------------------
 * The pixel path of the training step as tcgen05 GEMMs (ocv_conv3x3_nhwc_bf16, ocv_gemm_bf16) joined by HBM-bound helpers:
 *   ocv_tail_logits_f32: depth [B,P] = sum_k softmax_k(logits[b,k,p] + bias[k]) centres[b,k] on channel-major fp32 logits
 *     [B,K,P] (one pass, online softmax).  replaces conv_out's Softmax + torch.sum(out*centers) (modules/AdaBins.py:77,87).
 *   ocv_tail_logits_bwd_bf16: its backward — dlT [B,K,P] bf16 = g p (c - depth); grad_centres [B,K] = sum_p g p;
 *     grad_bias [K] = bias_beta*grad_bias + sum_{b,p} dl  (fixed-order block partials in `workspace`: deterministic).
 *   ocv_transpose_bf16: dst[z][c][r] = src[z][r][c], bf16, R % 64 == 0, C % 64 == 0 (strides in elements, even).
 *   ocv_im2col3x3_bf16: col [C*9+8, B*P] bf16 from the planar map x [B,C,h,w] bf16: row c*9+ky*3+kx = the map shifted by
 *     (ky-1, kx-1) with zero padding, row C*9 = 1 (its GEMM column is the conv bias gradient), 7 zero rows.  (h*w) % 8 == 0.
 * (backward of) self.conv3x3 (modules/miniViT.py:25; modules/ObjCAViT.py:374) as ONE GEMM dW[o, :] = dx3T[o, B*P] @ col^T. */
/* Pixel path of the training step WITHOUT materialised logits (csrc/train_tail_fused_bf16.cu); P % 128 == 0.
 *   x3 [B*P,128] bf16 NHWC = conv3x3(features) + bias (ocv_conv3x3_nhwc_bf16); Mb2 [B,256,128] bf16 = log2e * Wout @ Q[b] and
 *   blog2 [B,256] = log2e * conv_out bias (ocv_fold_queries_bf16 with b3 = NULL, logit_scale = log2e); centres [B,256].
 * fwd: depth [B*P], stat_max / stat_sum [B*P] (max and sum of 2^(l - max) over the 256 log2-domain logits of a pixel).
 *      replaces self.conv_out + softmax + sum(out * centers) (modules/GraphBins.py:109-119) on top of PixelWiseDotProduct.
 * bwd: from grad_depth [B*P]: dx3T [128, ld_dx3T] bf16 (channel-major d loss / d x3), dMb [B,256,128] fp32 (natural-logit units:
 *      d loss / d (Wout @ Q[b])), grad_centres [B,256], grad_bias [256] = bias_beta * grad_bias + sum over pixels of d logits
 *      (conv_out bias; may be NULL).  Deterministic (fixed-order reduction of per-CTA partial accumulators in `workspace`). */
int ocv_tail_fused_fwd_bf16(const void* x3_nhwc_bf16, const void* Mb2_bf16, const float* blog2, const float* centres, float* depth,
                            float* stat_max, float* stat_sum, int B, long long P, void* stream);
long long ocv_tail_fused_bwd_workspace(int B, long long P);
int ocv_tail_fused_bwd_bf16(const void* x3_nhwc_bf16, const void* Mb2_bf16, const float* blog2, const float* centres,
                            const float* depth, const float* stat_max, const float* stat_sum, const float* grad_depth,
                            void* dx3T_bf16, long long ld_dx3T, float* dMb, float* grad_centres, float* grad_bias, float bias_beta,
                            int B, long long P, void* workspace, void* stream);
int ocv_tail_logits_f32(const float* logits, const float* logit_bias, const float* centres, float* depth, int B, int K,
                        long long P, void* stream);
/* ..._stats_...: the forward also saves the per-pixel softmax statistics (max, sum, centre-weighted sum; stats [B,3,P] fp32), the
 * backward reads them instead of recomputing them in a first pass over the 4K-byte-per-pixel logits (same values, bit for bit). */
int ocv_tail_logits_stats_f32(const float* logits, const float* logit_bias, const float* centres, float* depth, float* stats, int B,
                              int K, long long P, void* stream);
int ocv_tail_logits_bwd_stats_bf16(const float* logits, const float* logit_bias, const float* centres, const float* stats,
                                   const float* grad_depth, void* dlT_bf16, float* grad_centres, float* grad_bias, float bias_beta,
                                   int B, int K, long long P, void* workspace, void* stream);
long long ocv_tail_logits_bwd_workspace(int B, int K, long long P);
int ocv_tail_logits_bwd_bf16(const float* logits, const float* logit_bias, const float* centres, const float* grad_depth,
                             void* dlT_bf16, float* grad_centres, float* grad_bias, float bias_beta, int B, int K, long long P,
                             void* workspace, void* stream);
int ocv_transpose_bf16(const void* src, void* dst, int batch, int R, int C, long long src_bstride, long long src_ld,
                       long long dst_bstride, long long dst_ld, void* stream);
int ocv_im2col3x3_bf16(const void* x_planar_bf16, void* col_bf16, int B, int C, int h, int w, void* stream);
/* Operands of the patch-embedding weight gradient as a K-major GEMM  dW [E, C P P] = dtokT [E, B S] @ unfold [C P P, B S]^T
 * (backward of embedding_convPxP, modules/layers.py:11-12,17; image_embedding_convPxP, modules/ObjCAViT.py:243,333):
 *   ocv_patch_unfold_bf16: out[(c*P*P + py*P + px), b*S + sy*gw + sx] = x[b,c,sy*P+py,sx*P+px]; x fp32 NCHW, out bf16 with row
 *     pitch ld (elements) >= B*S; columns >= B*S are left untouched (zero them once when ld is padded).
 *   ocv_transpose_cast_f32_bf16: dst[c][r] = bf16(src[r][c]) for any R, C (pitches in elements). */
int ocv_patch_unfold_bf16(const float* x, void* out_bf16, int B, int C, int h, int w, int P, long long ld, void* stream);
/* conv3x3 weight + bias gradient as an implicit GEMM over pixels (csrc/conv_wgrad_bf16.cu), no im2col buffer:
 *   gw[co][ci][ky][kx] = beta*gw + sum_{b,y,x} dY[b,y,x,co] X[b,y+ky-1,x+kx-1,ci],  gb[co] = beta*gb + sum dY   (gb may be NULL)
 * dyT_bf16: dY channel-major [128, ld_dy] bf16 (column = (b*h + y)*w + x; what ocv_tail_fused_bwd_bf16 writes), x_nhwc_bf16 [B,h,w,128],
 * w % 8 == 0, workspace of ocv_conv3x3_wgrad_bf16_workspace bytes (256-byte aligned).  Deterministic (fixed-order partial sums).
 * replaces autograd of nn.Conv2d(128,128,3,padding=1) w.r.t. its parameters: modules/ObjCAViT.py:374, modules/miniViT.py:25. */
long long ocv_conv3x3_wgrad_bf16_workspace(int B, int h, int w);
int ocv_conv3x3_wgrad_bf16(const void* dyT_bf16, long long ld_dy, const void* x_nhwc_bf16, float* gw, float* gb, int B, int h, int w,
                           float beta, void* workspace, void* stream);
/* the same operand from the NHWC bf16 copy of the features (xh [B,h,w,C], C % 32 == 0, 16-byte aligned): half the input bytes */
int ocv_patch_unfold_nhwc_bf16(const void* xh_bf16, void* out_bf16, int B, int C, int h, int w, int P, long long ld, void* stream);
int ocv_transpose_cast_f32_bf16(const float* src, void* dst_bf16, int R, int C, long long src_ld, long long dst_ld, void* stream);
/* Adjoint of ocv_pad_objects_f32: dsrc[offsets[b]+j, :] = beta * dsrc + dpadded[b, left+j, :] for j < n_b (left = L-Nmax when
 * left_pad else 0); pad rows carry no gradient. */
int ocv_unpad_objects_f32(const float* dpadded, const int32_t* offsets, float* dsrc, int B, int L, int Nmax, int E, int left_pad,
                          float beta, void* stream);
/* Backward of ocv_bins_f32 (norm 'linear'): grad_edges [B,K+1] and/or grad_centres [B,K] -> grad_y [B,K] (regressor output). */
int ocv_bins_bwd_f32(const float* y_raw, const float* grad_edges, const float* grad_centres, float* grad_y, int B, int K,
                     float min_depth, float max_depth, void* stream);
/* Same with an additional gradient given directly on the normalised widths (grad_widths [B,K] or NULL): the backward of the
 * `bin_widths_normed` output of mViT.forward / ObjCAViT.forward (miniViT.py:34-42, ObjCAViT.py:379-388) at the module surface. */
int ocv_bins_bwd_ex_f32(const float* y_raw, const float* grad_edges, const float* grad_centres, const float* grad_widths,
                        float* grad_y, int B, int K, float min_depth, float max_depth, void* stream);

/* ---- training losses (config 5), forward + gradients w.r.t. the head outputs ---------------------------------------
 * SILog: depth_pred [B,1,h,w] is up-sampled bilinearly (align_corners=True) to the GT size [B,1,H,W]; mask uint8 (1 = valid)
 *   or NULL; loss[0] = alph * sqrt(sum g^2 / n - lam (sum g)^2 / n^2), g = log p - log t over valid pixels of the WHOLE batch.
 *   ocv_silog_backward must follow ocv_silog_forward on the same workspace (it reuses the reduction statistics) and writes
 *   grad_pred [B,1,h,w] = grad_scale * d loss / d depth_pred.  Deterministic (fixed-order reductions, gather adjoint).
 * replaces: losses/SILogLoss.py:43-56 (and its autograd). */
long long ocv_silog_workspace(int B, int H, int W);
int ocv_silog_forward(const float* depth_pred, const float* depth_gt, const uint8_t* depth_mask, int B, int h, int w, int H, int W,
                      float alph, float lam, float* loss, void* workspace, void* stream);
int ocv_silog_backward(const float* depth_pred, const float* depth_gt, const uint8_t* depth_mask, int B, int h, int w, int H,
                       int W, float alph, float lam, float grad_scale, float* grad_pred, void* workspace, void* stream);
/* BinsChamfer: centres = mid-points of bin_edges [B,K+1]; targets = depth_gt[b][mask[b]] (HW values per image);
 *   loss[0] = mean_b (1/K) sum_k min_t (c_k - t)^2 + mean_b (1/T_b) sum_t min_k (c_k - t)^2   (pytorch3d chamfer_distance
 *   defaults: squared L2, point mean, batch mean, both directions);  grad_edges [B,K+1] (optional) = grad_scale * d loss / d bin_edges.
 *   Centres must be ascending (they are: cumulative sums of positive widths).  Targets must be > 0.  Deterministic.
 * replaces: losses/BinsChamferLoss.py:21-37 + pytorch3d.loss.chamfer_distance (third party, not vendored: parity unpinned). */
long long ocv_bins_chamfer_workspace(int B, int K);
int ocv_bins_chamfer(const float* bin_edges, const float* depth_gt, const uint8_t* depth_mask, int B, int K, long long HW,
                     float grad_scale, float* loss, float* grad_edges, void* workspace, void* stream);

/* ---- rows either side of the head -------------------------------------------------------------------
 * Test-time-augmentation merge: out = 0.5 * (clamp(depth, lo, hi) + clamp(flip_w(depth_mirror), lo, hi)); NaN stays NaN like
 *   torch.clamp; depth_mirror NULL -> out = clamp(depth).  All [B,1,h,w].
 * replaces: modules/GraphBinsLM.py:159-183 (clamp x2, flip, average). */
int ocv_tta_merge_f32(const float* depth, const float* depth_mirror, float* out, int B, int h, int w, float min_depth,
                      float max_depth, void* stream);
/* Evaluation metrics in one pass over the ground truth: depth_pred [B,1,h,w] is up-sampled bilinearly (align_corners=True) to
 *   [B,1,H,W], nan -> min_depth, +-inf -> max_depth, mask = gt > min_depth && gt <= max_depth && inside crop_host = {y0,y1,x0,x1}
 *   (NULL = no crop; the Garg / Eigen rectangles are the caller's), then the NINE fp64 sums
 *   {sum |gt-p|/gt, sum (gt-p)^2/gt, sum (gt-p)^2, sum (ln gt - ln p)^2, sum |log10 gt - log10 p|, #(thr<1.25), #(thr<1.25^2),
 *    #(thr<1.25^3), #valid}  with thr = max(gt/p, p/gt)
 *   are written to sums[9] (this call, optional) and ADDED to acc[9] (running state over batches, optional; sum-reduce it
 *   across ranks).  upsampled_out [B,1,H,W] optional.  Deterministic (two-stage fixed-order reduction).
 *   workspace: ocv_depth_metrics_workspace() bytes.
 * replaces: metrics/MetricsPreprocess.py:17-45 + the update() of metrics/{AbsRel,SqRel,RMSE,RMSELog,Log10,AccThresh}.py
 *   (modules/GraphBinsLM.py:198-222). */
long long ocv_depth_metrics_workspace(void);
int ocv_depth_metrics_f32(const float* depth_pred, const float* depth_gt, int B, int h, int w, int H, int W, float min_depth,
                          float max_depth, const int* crop_host, float* upsampled_out, double* sums, double* acc, void* workspace,
                          void* stream);
/* Object front-end glue: detections of image b are rows [offsets[b], offsets[b+1]) of xyxy [total,4] (corner boxes); written in
 *   REVERSED order per image when `reverse` (the reference's reversed(det)), as centre/size boxes xywh [total,4]; size_bucket
 *   [total] int32 (optional) = index into the relative-size scale (n_scale entries, 7 in the reference) comparing each object's
 *   box area with the NEXT object's: clip(round((ln(a_i/a_next)+1)/2*(n_scale-3))+1, 0, n_scale-1); -1 for a lone object;
 *   INT_MIN where the reference would raise (non-positive / NaN area ratio).  Bit-exact index work.
 * replaces: modules/Yolov7Wrapper.py:118-123 (reversed + xyxy2xywh) and modules/ObjectLanguageStrategy.py:61-81. */
int ocv_object_glue_f32(const float* xyxy, const int32_t* offsets, float* xywh, int32_t* size_bucket, int B, int n_max, int reverse,
                        int n_scale, void* stream);

/* ---- optimizer step on flat buffers (config 5) ---------------------------------------------------------------------------
 * out[0] = scale^2 * sum x^2 (fixed-order two-stage reduction; deterministic).  workspace: ocv_sumsq_workspace() bytes.
 * ocv_adamw_step_f32: torch.optim.AdamW semantics on flat param / grad / exp_avg / exp_avg_sq buffers (16-byte aligned):
 *   g = grad * grad_scale * clip, clip = grad_sumsq ? min(1, max_norm / (sqrt(grad_sumsq[0]) + 1e-6)) : 1   (grad_sumsq: DEVICE scalar,
 *   so that no host synchronisation sits between the all-reduce, the clipping and the update);  step >= 1 (bias correction).
 * replaces: clip_grad_norm_ (gradient_clip_val, main.py:35-38,64) + torch.optim.AdamW.step (modules/GraphBinsLM.py:467). */
long long ocv_sumsq_workspace(long long n);
int ocv_sumsq_f32(const float* x, long long n, float scale, float* out, void* workspace, void* stream);
int ocv_adamw_step_f32(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, float lr, float beta1,
                       float beta2, float eps, float weight_decay, int step, float grad_scale, const float* grad_sumsq,
                       float max_norm, void* stream);

/* ---- whole head, one call (native host-side sequencing) -------------------------------------------------
 * An ocv_head is a HOST-side plan: the head's configuration, the device pointers of its parameters keyed by the
 * reference's own state_dict names, and two non-blocking side streams used to run the object branch next to the image
 * branch.  It owns no device memory: packed weights and the workspace are caller buffers.  One ocv_head serves
 * one caller thread at a time (it caches the batch-invariant image positional rows for the last (h, w)).
 *
 * replaces, as ONE call: ObjCAViT.forward (modules/ObjCAViT.py:306-390) or mViT.forward (modules/miniViT.py:22-44)
 *   followed by the head tail of GraphBins.forward / AdaBins.forward (modules/GraphBins.py:107-119; modules/AdaBins.py:76-87),
 *   on the bf16 tensor-core path.  Quirks F5/F6/F7/F13 of SURVEY.md are reproduced.
 *
 * kind: 0 = AdaBins mViT head (keys "adaptive_bins_layer.*", "conv_out.0.*"), 1 = ObjCAViT head ("objcavit.*", "conv_out.0.*").
 * strategy (kind 1): 0 learned, 1 learned_bbox_wh, 2 grid_random, 3 grid_random_roi_align. */
typedef struct ocv_head ocv_head;
typedef struct ocv_head_config {
  int kind, strategy, no_obj_sa, use_2_saca;
  int n_bins;      /* 256 */
  int obj_dim;     /* 512 */
  float min_depth, max_depth;
  int reserved[8]; /* zero */
} ocv_head_config;

int ocv_head_create(const ocv_head_config* cfg, ocv_head** out);
void ocv_head_destroy(ocv_head* head);
/* Register one fp32 parameter (device pointer, borrowed) under its state_dict key relative to the head module, e.g.
 * "objcavit.saca_1.image_transformer_encoder.layers.0.linear1.weight".  Unknown keys are accepted (the never-executed
 * prototype layers of SURVEY.md F9 are simply not used). */
int ocv_head_set_param(ocv_head* head, const char* key, const float* device_ptr, long long numel);
/* Derived operand copies (bf16 GEMM weights, packed conv weights): size, then fill `packed` (device, 256-byte aligned).
 * Must be called again whenever a parameter changes.  Returns OCV_E_ARG naming nothing if a needed key is missing;
 * ocv_head_missing_key() then returns that key. */
long long ocv_head_packed_bytes(const ocv_head* head);
int ocv_head_pack(ocv_head* head, void* packed, void* stream);
const char* ocv_head_missing_key(const ocv_head* head);
/* Workspace for one forward: total_obj = sum_i N_i, n_max = max_i N_i (0, 0 for kind 0). */
long long ocv_head_workspace_bytes(const ocv_head* head, int B, int h, int w, int total_obj, int n_max);
/* X: [B,h,w,128] bf16 NHWC.  obj_feat [total_obj, obj_dim] fp32, obj_xywh [total_obj,4] fp32 (images without detections
 * carry the (-1,-1,-1,-1) placeholder row, modules/ObjCAViT.py:312-313), obj_offsets int32 [B+1] CSR (DEVICE).
 * Outputs: depth [B,h*w] fp32, bin_edges [B,n_bins+1] fp32, obj_tokens_out [total_obj,128] fp32 or NULL (the embedded
 * object tokens the reference writes back into the caller's list, F11).  Launches only; no sync, no allocation. */
int ocv_head_forward(ocv_head* head, const void* X_nhwc_bf16, const float* obj_feat, const float* obj_xywh,
                     const int32_t* obj_offsets, int total_obj, int n_max, int B, int h, int w, float* depth,
                     float* bin_edges, float* obj_tokens_out, void* workspace, void* stream);
/* Pipelined serving mode (opt-in; lanes = 0 restores the classic mode where everything runs on the caller's stream).
 * With lanes = L in 1..4, call i of ocv_head_forward runs its token path on lane i % L (the head's own streams, ordered after
 * the caller's stream at call time) and its fused pixel kernel on ONE shared pixel stream with at most pixel_ctas CTAs
 * (0 = one per SM), so the latency-bound token chains of the next calls overlap the tensor-bound pixel kernel of this one.
 * `workspace` must then hold L x ocv_head_workspace_bytes' single-lane size (ocv_head_workspace_bytes already reports the
 * multiplied size once the mode is set).  The call returns WITHOUT ordering the caller's stream after the work:
 *   - outputs (depth, bin_edges, obj_tokens_out) are valid on `stream` only after ocv_head_join(head, stream);
 *   - X and the object inputs must stay untouched until then; set_param / pack require a join first.
 * ocv_head_stream: the head's internal streams (which: 0 token lane, 1 side lane, 2 pixel stream) so that a caching
 * allocator can be told about them (torch: Tensor.record_stream). */
int ocv_head_set_pipeline(ocv_head* head, int lanes, int pixel_ctas);
int ocv_head_join(ocv_head* head, void* stream);
void* ocv_head_stream(const ocv_head* head, int which, int lane);
/* number of kernel launches the last ocv_head_forward issued */
int ocv_head_last_launches(const ocv_head* head);
/* Measurement aid (bench.py): when enabled, ocv_head_forward brackets the fused pixel kernel with CUDA events on the
 * caller's stream; ocv_head_pixel_ms() waits for the last pair and returns its duration in ms (the one entry point that
 * blocks the calling thread), or -1 if none was recorded. */
int ocv_head_enable_timing(ocv_head* head, int on);
float ocv_head_pixel_ms(ocv_head* head);

#ifdef __cplusplus
}
#endif
#endif /* OBJCAVIT_B200_H_ */
