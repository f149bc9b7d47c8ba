/* applestar_b200 -- C ABI of the sm_100a kernels behind the DI-star / Applestar policy-value hot path.
 *
 * The reference has no native code: each entry point below replaces the stock ATen call
 * sequence at the cited reference lines.  Conventions: raw device pointers + sizes, caller-owned memory and
 * workspaces, `stream` is a cudaStream_t passed as void*, return 0 = ok, <0 = argument error (AS_E*),
 * >0 = cudaError_t.  No entry point allocates or frees device memory unless it says so.
 */
#ifndef APPLESTAR_B200_H
#define APPLESTAR_B200_H
#include <stdint.h>
#include "../applestar_b200/csrc/loss.h"

#ifdef __cplusplus
extern "C" {
#endif

#define AS_EINVAL (-1)
#define AS_ESIZE (-2)
#define AS_EALIGN (-3)
#define AS_EUNSUPPORTED (-4)
#define AS_EDRIVER (-5)

/* ABI version of this library (bumped when a signature changes). */
int as_abi_version(void);

/* ScatterConnection forward -- replaces scatter_connection(), distar/agent/default/model/module_utils.py:11-34
 * (called from model/encoder.py:40-42).  emb [R,N,D] fp32, xs/ys [R,N] (idx_bytes = 1: uint8, 8: int64),
 * entity_num [R] int64 or NULL (rows >= entity_num scatter 0.0, encoder.py:38), out [R,D,H,W] fp32 fully written.
 * cover = 0: 'add' (ascending-entity-order fp32 sums, bit-exact vs the CPU reference), 1: 'cover' (last wins).
 * N <= 512, D % 8 == 0, H*W % 4 == 0. */
int as_scatter_connection_fwd(const float* emb, const void* xs, const void* ys, int idx_bytes,
                              const int64_t* entity_num, float* out, int64_t R, int N, int D, int H, int W, int cover,
                              void* stream);

/* ScatterConnection backward ('add'): d_emb[r,e,:] = e < entity_num[r] ? d_out[r,:,cell(e)] : 0
 * -- replaces autograd of scatter_add_ at module_utils.py:29. */
int as_scatter_connection_bwd(const float* d_out, const void* xs, const void* ys, int idx_bytes,
                              const int64_t* entity_num, float* d_emb, int64_t R, int N, int D, int H, int W,
                              void* stream);

/* bf16 tensor-core GEMM (tcgen05 / TMEM / TMA):  C[b][m][n] = act(alpha * sum_k A[b][m][k] B[b][n][k] + bias[n]).
 * Replaces every nn.Linear / torch.matmul on the path (fc_block ctools/torch_utils/network/nn_module.py:231-270,
 * Attention module_utils.py:95-110, LSTM gates lstm.py:140-141) and their autograd matmuls.
 * C2 (optional, NULL to skip): bf16 copy of a fp32 C written by the same epilogue.  b_k_shift: rows added to the
 * contraction coordinate of an MN-major B (used by the convolution wgrad taps; 0 otherwise).
 * a_mn / b_mn = 0: operand stored [rows][K contiguous]; 1: stored [K rows][M-or-N contiguous] (transposed use
 * without a copy).  batch = nb0*nb1 with per-operand element strides.  c_bf16: 0 fp32 / 1 bf16 output.
 * splits > 1: fp32 partial sums to ws[splits][batch][M][N] (then as_splitk_reduce); the value must equal
 * as_gemm_splits(K, splits).  Pointers 16-byte aligned, lda/ldb multiples of 8 elements. */
int as_gemm_bf16(const void* A, const void* B, void* C, void* C2, const float* bias, int M, int N, int K, int a_mn,
                 int b_mn,
                 int64_t lda, int64_t ldb, int64_t ldc, int nb0, int nb1, int64_t a_sb0, int64_t a_sb1, int64_t b_sb0,
                 int64_t b_sb1, int64_t c_sb0, int64_t c_sb1, int c_bf16, int relu, float alpha, int splits, float* ws,
                 int b_k_shift, void* stream);
/* Number of non-empty K splits the kernel will produce when `want` are requested. */
int as_gemm_splits(int K, int want);
/* as_gemm_bf16_masked: C = alpha * A B^T (same operand conventions as as_gemm_bf16, no batch / bias) with the ReLU
 * backward of the layer that feeds on C folded into the epilogue: C[m][n] is written as zero where mask[m][n] <= 0
 * (mask = that layer's bf16 output, C's layout).  Used for d hidden of the Transformer FFN (module_utils.py:121-123:
 * ReLU on both layers): autograd's threshold_backward pass over the [rows, 1024] gradient disappears.
 * Needs N % 32 == 0, ldc % 8 == 0 and 16-byte aligned C / mask. */
int as_gemm_bf16_masked(const void* A, const void* B, void* C, const void* mask, int M, int N, int K, int a_mn, int b_mn, int64_t lda,
                        int64_t ldb, int64_t ldc, int c_bf16, float alpha, void* stream);

/* The same ReLU backward from SIGN BITS instead of the activation itself (1/16 of the bytes; the epilogue of a K = 256
 * GEMM cannot hide a second [M, N] bf16 stream): as_gemm_bf16_relu_bits is the forward layer, C bf16 = relu(A B^T + bias)
 * with the sign bits (C[m][n] > 0) written alongside as bits[m][n / 32] (N % 32 == 0; within a word, column 2k of the 32 is
 * bit 15-k and column 2k+1 is bit 31-k: the order the packed bf16 pairs give for free); as_gemm_bf16_bitmasked computes
 * C = alpha * A B^T zeroed where the bit is clear. */
int as_gemm_bf16_relu_bits(const void* A, const void* B, void* C, const float* bias, uint32_t* bits, int M, int N, int K, int64_t lda,
                           int64_t ldb, int64_t ldc, void* stream);
int as_gemm_bf16_bitmasked(const void* A, const void* B, void* C, const uint32_t* bits, int M, int N, int K, int a_mn, int b_mn,
                           int64_t lda, int64_t ldb, int64_t ldc, int c_bf16, float alpha, void* stream);

/* out[i] = (accumulate ? out[i] : 0) + sum_s ws[s][i], s ascending (deterministic). */
int as_splitk_reduce(const float* ws, float* out, int64_t n, int splits, int accumulate, void* stream);

/* 3x3 / 1x1 convolution as implicit GEMM over zero-padded NHWC bf16 pixels -- replaces nn.Conv2d inside
 * conv2d_block (nn_module.py:119-174; ResBlock res_block.py:61-65, GatedResBlock module_utils.py:224-231).
 * x [R,(H+2),(W+2),Cin] bf16 with a zero border, w [Cout][taps*Cin] bf16 (tap-major), y [R,(H+2),(W+2),Cout]
 * (border written as zero).  taps = 9 (3x3, pad 1) or 1.  Cin % 8 == 0. */
int as_conv_nhwc_bf16(const void* x, const void* w, void* y, const float* bias, int64_t R, int H, int W, int Cin,
                      int Cout, int taps, int w_mn, int y_bf16, int relu, void* stream);
/* y = act(conv(x) + bias + addend): the same convolution with a bf16 map of y's shape added in the epilogue.  Used in the
 * backward pass where an activation feeds a convolution AND a residual / gate branch (res_block.py:60-64,
 * module_utils.py:222-231): the data gradient of the convolution is accumulated onto the branch's gradient in the GEMM
 * epilogue instead of by a separate add over the two maps.  Cout % 32 == 0. */
int as_conv_nhwc_bf16_add(const void* x, const void* w, void* y, const float* bias, const void* addend, int64_t R, int H,
                          int W, int Cin, int Cout, int taps, int relu, void* stream);
/* as_conv_nhwc_bf16_masked: data gradient of a convolution whose INPUT is the ReLU output `mask` of the previous layer
 * (conv1 -> ReLU -> conv2 in ResBlock / GatedResBlock / the gate chain, res_block.py:61-65, module_utils.py:221-230):
 * y = (conv(x, w) + addend) * (mask > 0), i.e. the previous layer's threshold_backward is applied in the GEMM epilogue
 * and that layer's backward starts from an already masked gradient.  addend may be NULL; mask has y's layout. */
int as_conv_nhwc_bf16_masked(const void* x, const void* w, void* y, const void* addend, const void* mask, int64_t R, int H, int W,
                             int Cin, int Cout, int taps, void* stream);
/* The same pair from sign bits (layout as in as_gemm_bf16_relu_bits, one row of Cout / 32 words per padded pixel):
 * as_conv_nhwc_bf16_relu_bits is the ReLU convolution that also writes the bits of its output, as_conv_nhwc_bf16_bitmasked
 * the data gradient y = (conv(x, w) + addend) zeroed where the bit is clear.  Cout % 32 == 0. */
int as_conv_nhwc_bf16_relu_bits(const void* x, const void* w, void* y, const float* bias, uint32_t* bits, int64_t R, int H, int W, int Cin,
                                int Cout, int taps, void* stream);
int as_conv_nhwc_bf16_bitmasked(const void* x, const void* w, void* y, const void* addend, const uint32_t* bits, int64_t R, int H, int W,
                                int Cin, int Cout, int taps, void* stream);

/* Weight gradient of the NHWC convolution above: dw[tap][co][ci] = sum_p dy[p][co] * x[p + shift(tap)][ci] as one
 * tap-batched, split-K tcgen05 GEMM with both operands read MN-major.  dy [R,(H+2),(W+2),Cout] bf16 (zero border),
 * x [R,(H+2),(W+2),Cin] bf16; dw fp32 [taps][Cout][Cin]; ws: splits*taps*Cout*Cin floats when splits > 1
 * (splits must equal as_gemm_splits(R*(H+2)*(W+2), want)).  Replaces autograd of nn.Conv2d (nn_module.py:166-168). */
int as_conv_wgrad_nhwc_bf16(const void* dy, const void* x, float* dw, float* ws, int64_t R, int H, int W, int Cin,
                            int Cout, int taps, int splits, void* stream);
/* Tap pairing of the 3x3 weight gradient: when the M side of dW = dy^T x has 64 channels (Cout == 64, or Cin == 64 with
 * the operands swapped), two taps share one 128-row tcgen05.mma.  as_conv_wgrad_pair_mode tells which form
 * as_conv_wgrad_nhwc_bf16 will use (0 none, 1 dy side, 2 swapped; 11 / 12: the same with all nine taps taken from one
 * shared-memory slab per 64-pixel block, one CTA per split -- maps up to 125 pixels wide) and as_conv_wgrad_ws_floats the
 * split-K workspace size it then needs. */
int as_conv_wgrad_pair_mode(int Cin, int Cout, int taps, int splits, int W);
int64_t as_conv_wgrad_ws_floats(int Cin, int Cout, int taps, int splits, int W);

/* Layout conversion between the reference's NCHW fp32 maps and the padded NHWC bf16 maps the convolutions use:
 * dst [R,H+2,W+2,Cp] (border and channels >= C zero; relu != 0 applies ReLU on the way) / its inverse (interior,
 * first C channels). */
int as_nchw_to_pnhwc(const float* src, void* dst, int64_t R, int C, int Cp, int H, int W, int relu, void* stream);
int as_pnhwc_to_nchw(const void* src, float* dst, int64_t R, int C, int Cp, int H, int W, void* stream);

/* bf16 element-wise tails (n % 8 == 0): mode 0 y = a + b (skip add, action_arg_head.py:429); mode 1
 * y = relu(a + b) (ResBlock tail, res_block.py:64); as_relu_bwd_bf16: dz = dy * (y > 0). */
int as_add_act_bf16(const void* a, const void* b, void* y, int64_t n, int mode, void* stream);
int as_relu_bwd_bf16(const void* dy, const void* y, void* dz, int64_t n, void* stream);
int as_relu_bwd_f32(const float* dy, const float* y, float* dz, int64_t n, void* stream);

/* GatedResBlock tail y = relu(tanh(r * sigmoid(g)) * UpdateSP + x) (module_utils.py:227-229), bf16 maps, fp32 math.
 * Backward also yields d UpdateSP (device float) through `scratch` of as_gated_scratch_floats() floats. */
int as_gated_fwd_bf16(const void* r, const void* g, const void* x, const float* update_sp, void* y, int64_t n,
                      void* stream);
int64_t as_gated_scratch_floats(void);
int as_gated_bwd_bf16(const void* dy, const void* y, const void* r, const void* g, const float* update_sp, void* dr,
                      void* dg, void* dx, float* d_update_sp, float* scratch, int64_t n, void* stream);

/* 2x bilinear upsampling (F.interpolate(scale_factor=2, mode='bilinear'), action_arg_head.py:440) on padded NHWC
 * bf16 maps: in [R,H+2,W+2,C] -> out [R,2H+2,2W+2,C]; backward is the exact transpose in gather form. C % 8 == 0. */
int as_upsample2_fwd_bf16(const void* in, void* out, int64_t R, int H, int W, int C, void* stream);
int as_upsample2_bwd_bf16(const void* dout, void* din, int64_t R, int H, int W, int C, void* stream);

/* Location-head input relu(cat[reshape(project_embed), map_skip[-1]]) (action_arg_head.py:419-422) as a padded NHWC
 * bf16 map with Cp (>= 4 + Cs) channels, and its backward (d_emb fp32 [R,4*H*W], d_skip bf16 like skip). */
int as_loc_cat_fwd(const float* emb, const void* skip, void* out, int64_t R, int H, int W, int Cs, int Cp, void* stream);
int as_loc_cat_bwd(const void* dcat, const float* emb, const void* skip, float* d_emb, void* d_skip, int64_t R, int H,
                   int W, int Cs, int Cp, void* stream);

/* Final location-head convolution 3x3, 32 -> 1 (upsample.2, action_arg_head.py:443) fused with the flatten and the
 * temperature division (:445-446): in padded NHWC bf16 [R,H+2,W+2,32], w fp32 [9][32], logits fp32 [R,H*W].
 * Backward: d_in (bf16, zero border), dw [9][32], db [1]; scratch of as_conv_c32_1_scratch_floats() floats. */
int as_conv_c32_1_fwd(const void* in, const float* w, const float* bias, float* logits, int64_t R, int H, int W,
                      float inv_temperature, void* stream);
int64_t as_conv_c32_1_scratch_floats(void);
int as_conv_c32_1_bwd(const float* dlogits, const void* in, const float* w, void* d_in, float* dw, float* db,
                      float* scratch, int64_t R, int H, int W, float inv_temperature, void* stream);

/* out[c] = sum_r x[r][c] for bf16 x [rows, C] (bias gradients); scratch of as_colsum_scratch_floats(C) floats. */
int64_t as_colsum_scratch_floats(int C);
int as_colsum_bf16(const void* x, float* out, float* scratch, int64_t rows, int C, void* stream);
int as_colsum_f32(const float* x, float* out, float* scratch, int64_t rows, int C, void* stream);
/* ReLU backward and bias gradient of a convolution in one pass over bf16 rows [rows, C] (C % 8 == 0, 16-byte aligned):
 * dz = dy * (y > 0), colsum[c] = sum_r dz[r][c] -- the `dy * (y > 0)` and `.sum` passes autograd runs for
 * conv2d_block's ReLU + bias (nn_module.py:119-174).  scratch: as_colsum_scratch_floats(C) floats. */
int as_relu_bwd_colsum_bf16(const void* dy, const void* y, void* dz, float* colsum, float* scratch, int64_t rows, int C,
                            void* stream);
/* Backward prologue of fc_block (nn_module.py:231-270) in one pass over the fp32 upstream gradient [rows, C],
 * C % 8 == 0: dz16 = bf16(dy * (y > 0)) for the dgrad / wgrad GEMMs (y32 or y16 = the layer's ReLU output, both NULL =
 * no activation) and colsum[c] = sum_r dy*(y>0) = the bias gradient (colsum NULL = no bias).  Replaces autograd's
 * threshold_backward, the dtype cast and the `.sum(0)`.  scratch: as_colsum_scratch_floats(C) floats. */
int as_linear_bwd_prep(const float* dy, const float* y32, const void* y16, void* dz16, float* colsum, float* scratch,
                       int64_t rows, int C, void* stream);
/* The same when the layer's output existed as an fp32 and a bf16 copy and BOTH received a gradient (dy = dy32 + dy16):
 * the sum is formed inside the pass instead of by a separate add over the two tensors. */
int as_linear_bwd_prep2(const float* dy32, const void* dy16, const float* y32, const void* y16, void* dz16, float* colsum,
                        float* scratch, int64_t rows, int C, void* stream);   /* fp32 rows, same scratch size */
/* Same column sums for hosts that may not allocate device memory at all (a process that gave its whole device
 * budget to an arena at start-up, or a step loop that cannot afford the device-wide synchronisation cudaMalloc /
 * cudaFree bring).  This entry point has no counterpart in the reference.  It allocates nothing: the partial sums
 * go to a 1 MiB scratch area built into the library (present on every GPU that loads it until the process exits),
 * which limits it to C <= 256 (AS_ESIZE beyond that: use as_colsum_bf16 with your own scratch).  The first call on a
 * device looks the area's address up (with that device current); make that first call before any stream capture.
 * All calls share the one area, which keeps the previous call's partial sums: nothing needs clearing (each call
 * overwrites what it reads), but two calls that overlap on the GPU (different streams) race on it. */
int as_colsum_bf16_static(const void* x, float* out, int64_t rows, int C, void* stream);

/* Spatial input features -- replaces the one-hot embeddings, the effect index_put and torch.cat of
 * SpatialEncoder.forward (obs_encoder/spatial_encoder.py:52-71): planes[7] = height_map, visibility_map, creep,
 * player_relative, alerts, pathable, buildable (uint8 [R,H,W]); effects[6] = int16 [R,n_eff] flat pixel ids
 * (zero padding sets pixel 0, as in the reference); scatter_map fp32 [R,32,H,W].  out: zero-padded NHWC bf16
 * [R,H+2,W+2,64] (56 real channels in the reference order, 8 zero), the A operand of the 1x1 project GEMM. */
int as_spatial_features(const void* const* planes, const void* const* effects, int n_eff, const float* scatter_map,
                        void* out, int64_t R, int H, int W, int Cs, void* stream);
/* The same with ScatterConnection ('add', module_utils.py:11-34 via encoder.py:37-42) fused in: the 32 scatter
 * channels of a pixel are summed straight from emb [R,N,32] fp32 at the clamped (xs, ys) of the entities below
 * entity_num, in ascending entity order (bit-identical to as_scatter_connection_fwd), so the dense fp32 map is never
 * written.  idx_bytes 1 (uint8) or 8 (int64); N <= 512. */
int as_spatial_features_scatter(const void* const* planes, const void* const* effects, int n_eff, const float* emb, const void* xs,
                                const void* ys, int idx_bytes, const int64_t* entity_num, int N, void* out, int64_t R, int H,
                                int W, void* stream);
/* ScatterConnection backward reading the gradient of that feature map directly (channels ch0 .. ch0+D-1). */
int as_scatter_connection_bwd_pnhwc(const void* d_feat, const void* xs, const void* ys, int idx_bytes,
                                    const int64_t* entity_num, float* d_emb, int64_t R, int N, int D, int H, int W,
                                    int Cp, int ch0, void* stream);

/* 2x2 max-pool on padded NHWC bf16 maps -- replaces F.max_pool2d (spatial_encoder.py:79).  in [R,H+2,W+2,C] ->
 * out [R,H/2+2,W/2+2,C]; backward routes to the first maximum of the window (torch semantics), border zero. */
int as_maxpool2_fwd_bf16(const void* in, void* out, int64_t R, int H, int W, int C, void* stream);
int as_maxpool2_bwd_bf16(const void* in, const void* dout, void* din, int64_t R, int H, int W, int C, int relu_input, void* stream);
/* relu_input != 0: `in` is a ReLU layer's output; windows whose maximum is 0 route nothing, so din already carries that
 * layer's ReLU backward mask. */

/* LayerNorm(x (+ residual)) * gamma + beta, eps 1e-5, optional fused ReLU -- replaces nn.LayerNorm at
 * module_utils.py:135,139 (Transformer), lstm.py:140-151 (LN-LSTM), res_block.py:68-141 (ResFCBlock/ResFCBlock2).
 * x/res [rows, C] fp32 (res may be NULL); outputs y32 fp32 and/or y16 bf16 (either may be NULL);
 * stats [rows, 2] = (mean, rstd) saved for backward (may be NULL).  C <= 1536. */
int as_layer_norm_fwd(const float* x, const float* res, const float* gamma, const float* beta, float* y32, void* y16,
                      float* stats, int64_t rows, int C, int relu, void* stream);
/* Backward: upstream gradient = dy32 (fp32, may be NULL) + dy16 (bf16, may be NULL).  dx (also the residual's
 * gradient) may be NULL; dgamma/dbeta need `scratch` of as_layer_norm_bwd_scratch_floats(C) floats. */
int64_t as_layer_norm_bwd_scratch_floats(int C);
int as_layer_norm_bwd(const float* dy32, const void* dy16, const float* x, const float* res, const float* gamma,
                      const float* beta, const float* stats, float* dx, float* dgamma, float* dbeta, float* scratch,
                      int64_t rows, int C, int relu, void* stream);

/* as_layer_norm_bwd_fused: backward of y = LN(res + x) where x is the output of a linear layer (post-LN Transformer:
 * module_utils.py:133-139, x = attention projection or FFN output) whose own backward follows.  One pass yields the
 * residual-path gradient dx (fp32, may be NULL), the linear layer's pre-activation gradient dz16 = bf16(dx * (x > 0))
 * (mask_x != 0 when that layer has a ReLU), dgamma / dbeta and dbias = column sums of dz, replacing autograd's
 * threshold_backward + cast + bias `.sum(0)` passes.  C = 256 or 384. */
int64_t as_layer_norm_bwd_fused_scratch_floats(int C);
int as_layer_norm_bwd_fused(const float* dy32, const void* dy16, const float* x, const float* res, const float* gamma,
                            const float* beta, const float* stats, float* dx, void* dz16, int mask_x, float* dgamma, float* dbeta,
                            float* dbias, float* scratch, int64_t rows, int C, int relu, void* stream);

/* LayerNorm-LSTM recurrence over a whole unroll, one CTA per trajectory -- replaces the Python time loop of
 * LSTMLayer / LayerNormLSTMCell (model/lstm.py:138-167; 65 x 3 cell calls per learner step).
 * gi [T,B,4H] = LN_i(x W_ih^T) for all t (computed by as_gemm_bf16 + as_layer_norm_fwd), h0/c0 [B,H],
 * whh bf16 [4H,H]; outputs y [T,B,H] (h_t) and c_all [T,B,H]; u_all [T,B,4H], z_all [T,B,H], stats [T,B,4] are
 * saved for backward (pass NULL for inference).  H % 8 == 0, H <= 1024. */
int as_lnlstm_fwd(const float* gi, const float* h0, const float* c0, const void* whh, const float* ln_h_w,
                  const float* ln_h_b, const float* ln_c_w, const float* ln_c_b, float* y, float* c_all, float* u_all,
                  float* z_all, float* stats, int T, int B, int H, void* stream);
/* Reverse-time twin.  whh_t = W_hh^T bf16 [H,4H]; dy [T,B,H] / dhT / dcT [B,H] may be NULL.  Outputs: dgi [T,B,4H],
 * du [T,B,4H] (dW_hh = sum_{t,b} du h_{t-1}^T is a GEMM the caller runs), dh0/dc0 [B,H] (may be NULL), and dparam
 * [B, 10H] = per-trajectory partial gradients of (ln_h_w, ln_h_b, ln_c_w, ln_c_b). */
int as_lnlstm_bwd(const float* gi, const float* h0, const float* c0, const float* y, const float* c_all,
                  const float* u_all, const float* z_all, const float* stats, const void* whh_t, const float* ln_h_w,
                  const float* ln_h_b, const float* ln_c_w, const float* ln_c_b, const float* dy, const float* dhT,
                  const float* dcT, float* dgi, float* du, float* dh0, float* dc0, float* dparam, int T, int B, int H,
                  void* stream);
/* 1 when both directions run for this H on clusters of 8 CTAs (one or two trajectories per cluster) that keep their
 * W_hh / W_hh^T row slices in shared memory for the whole unroll and exchange u / dh slices through distributed
 * shared memory once per step (H = 384, the core LSTM); 0 when one CTA per trajectory streams the weights from L2.
 * Same arguments and results either way. */
int as_lnlstm_cluster_ok(int H);

/* Masked attention softmax with warp-shuffle reductions -- replaces masked_fill_ + F.softmax at
 * module_utils.py:100-104.  scores [groups, Nq, Nk] fp32 or bf16 (scores_bf16 = 1; already scaled by 1/sqrt(d)); key j
 * of group g is valid iff j < key_len[g / groups_per_len] (key_len NULL = no mask); probs bf16 with masked keys
 * exactly 0.  Nk <= 1024.  The arithmetic is fp32 either way. */
int as_softmax_fwd(const void* scores, const int64_t* key_len, void* probs, int64_t groups, int Nq, int Nk,
                   int groups_per_len, int scores_bf16, void* stream);
/* dscores (bf16) = probs * (dprobs - rowsum(dprobs * probs)); dprobs fp32 or bf16 (dprobs_bf16 = 1). */
int as_softmax_bwd(const void* probs, const void* dprobs, void* dscores, int64_t rows, int Nk, int dprobs_bf16,
                   void* stream);

/* Mean of the valid entity rows: out[r,c] = sum_{n < len[r]} x[r,n,c] / len[r] over x [R,N,C] fp32 (C % 4 == 0,
 * C <= 1024) -- the mask multiply + sum + divide of entity_encoder.py:83-85; backward writes
 * dx[r,n,c] = n < len[r] ? dout[r,c] / len[r] : 0. */
int as_masked_mean_fwd(const float* x, const int64_t* len, float* out, int64_t R, int N, int C, void* stream);
int as_masked_mean_bwd(const float* dout, const int64_t* len, float* dx, int64_t R, int N, int C, void* stream);

/* Teacher-forced SelectedUnits pointer loop (head/action_arg_head.py:168-216) without the per-step one-hot tensors.
 * as_su_prefix_fwd: emb[r,i,:] = mean (sum when selected_units_num[r] == 0) of the key rows of the distinct units picked
 * at steps <= i before the end token (= entity_num[r]) -- the `one_hot`/`einsum`/division of :198-206 for all S steps;
 * cnt [R,S] and isnew [R,S] are saved for as_su_prefix_bwd, which accumulates d key (caller zeroes dkey first).
 * key [R,N1,32] fp32, selected_units [R,Smax] int64, S <= min(Smax, 64).
 * as_su_mask_logits: out = valid ? x * scale : fill over x [R,S,N1], valid = the logits mask of step i (:179-187):
 * n <= entity_num, not the end token at step 0, not picked before step i.  Forward: fill = -1e9 * scale; backward:
 * the same call on the gradient with fill = 0. */
int as_su_prefix_fwd(const float* key, const int64_t* selected_units, const int64_t* entity_num, const int64_t* su_num, float* emb,
                     float* cnt, uint8_t* isnew, int64_t R, int N1, int D, int Smax, int S, void* stream);
int as_su_prefix_bwd(const float* demb, const int64_t* selected_units, const int64_t* su_num, const float* cnt, const uint8_t* isnew,
                     float* dkey, int64_t R, int N1, int D, int Smax, int S, void* stream);
int as_su_mask_logits(const float* x, const int64_t* selected_units, const int64_t* entity_num, float* out, int64_t R, int N1,
                      int Smax, int S, float scale, float fill, void* stream);

/* Front of the spatial encoder in one kernel per direction: ScatterConnection 'add' (model/encoder.py:37-42,
 * module_utils.py:11-34) + the 56-channel input (one-hot planes, height / 256, effect planes with the pixel-0 padding
 * quirk, scatter channels: obs_encoder/spatial_encoder.py:52-71) + 1x1 project conv 56 -> 32 + ReLU (:72) + the first
 * 2x2 max-pool (:79), evaluated sparsely and exactly in fp32 (two table rows per pixel).
 * planes[7] uint8 [R,H,W] (height, visibility, creep, player_relative, alerts, pathable, buildable), effects[6] int16
 * [R,n_eff], emb [R,N,32] fp32 projected entity embeddings, xs / ys [R,N] (idx_bytes 1 or 8), weight [32,56] fp32, bias
 * [32] -> out [R, H/2+2, W/2+2, 32] bf16, zero border.  H % 38 == 0, W even and <= 160, N <= 512.  `cells` =
 * as_spatial_front_cells_bytes(R) bytes (256-byte aligned) that receive the weight tables and the per-observation sorted
 * cell list; `codes` (may be NULL at inference) = R * (H/2) * (W/2) * 16 bytes: per pooled pixel and channel the 4-bit
 * index of the window pixel that won the max (15: none survived the ReLU).  The backward reads both buffers.
 * Backward: dpool (layout of out) -> dweight [32,56] columns 0..23 (columns 24..55 = g_entity^T emb, left to the caller),
 * dbias [32], g_entity [R,N,32] = gradient of the pre-activation at each entity's cell (rows >= entity_num zero), from
 * which d emb = g_entity W[:,24:56]; scratch = as_spatial_front_bwd_scratch_floats(R, H) floats. */
int as_spatial_front_fwd(const void* const* planes, const void* const* effects, int n_eff, const float* emb, const void* xs,
                         const void* ys, int idx_bytes, const int64_t* entity_num, int N, const float* weight, const float* bias,
                         void* cells, void* out, void* codes, int64_t R, int H, int W, void* stream);
int64_t as_spatial_front_cells_bytes(int64_t R);
int64_t as_spatial_front_bwd_scratch_floats(int64_t R, int H);
int as_spatial_front_bwd(const void* const* planes, const void* const* effects, int n_eff, const float* emb, const void* xs,
                         const void* ys, int idx_bytes, const int64_t* entity_num, int N, const float* weight, const float* bias,
                         const void* cells, const void* codes, const void* dpool, float* dweight, float* dbias, float* g_entity,
                         float* scratch, int64_t R, int H, int W, void* stream);

/* Pointer keys of the SelectedUnits head (head/action_arg_head.py:118-129): out [R, N+1, D] = key [R, N, D] with one zero
 * row appended and the end-flag embedding [D] placed at row entity_num[r].  Backward: dkey = dout without the end-flag row
 * (zero there), dend[d] = sum_r dout[r, entity_num[r], d] in row order (deterministic).  fp32, D % 4 == 0, D <= 256. */
int as_pointer_keys_fwd(const float* key, const float* end_embedding, const int64_t* entity_num, float* out, int64_t R, int N, int D,
                        void* stream);
int as_pointer_keys_bwd(const float* dout, const int64_t* entity_num, float* dkey, float* dend, int64_t R, int N, int D, void* stream);

/* Trajectory staging (rl_training/rl_dataloader.py:206-245 pads the teacher's selected-units logits to [T,B,64,N+1] with
 * -1e9: 134 MB of the 456 MB learner batch, almost all padding).  The host sends only the rows the teacher emitted plus
 * one int32 per dense row; as_expand_rows_f32 rebuilds the dense tensor on the device:
 * out[r,:] = row_of[r] >= 0 ? rows[row_of[r],:] : fill. */
int as_expand_rows_f32(const float* rows, const int32_t* row_of, float* out, int64_t total, int K, float fill, void* stream);

/* Supervised loss rows (sl_training/sl_loss.py:16-36,127-285): cross entropy of R rows of K fp32 logits (row pitch ld)
 * against int64 labels.  loss[r] = (1-smoothing)*(lse - x[label]) + smoothing*(lse - mean(x)); lse[r] and the arg-max of
 * every row (first maximum) are also written.  excl != NULL selects the SelectedUnits label mask of sl_loss.py:190-204:
 * rows are (sample b, step s) with s < S, excl = labels [R/S, S], excl_len[b] = selected_units_num - 1 (>= 0); logit k of
 * row (b, s) counts as -1e9 when k is one of excl[b, :excl_len[b]] other than this step's own label.
 * as_ce_rows_bwd: dlogits[r, :] = g[r] * d loss[r] / d logits[r, :] (zero at excluded positions and for rows with g == 0;
 * every element of the R x K block is written). */
int as_ce_rows_fwd(const float* logits, int64_t ld, const int64_t* labels, const int64_t* excl, const int64_t* excl_len, int S, int K,
                   float smoothing, float* loss, float* lse, int32_t* argmax, int64_t R, void* stream);
int as_ce_rows_bwd(const float* logits, int64_t ld, const int64_t* labels, const int64_t* excl, const int64_t* excl_len, int S, int K,
                   float smoothing, const float* lse, const float* g, float* dlogits, int64_t R, void* stream);

/* Scalar encoder front (obs_encoder/scalar_encoder.py:91-132, :41-53).  dtype codes: 0 uint8, 1 int8, 2 int16, 3 float16,
 * 4 int64, 5 float32 (the narrow storage types of lib/features.py:44-54 are read directly).
 * as_embed_relu_fwd: out[r,:] = relu(table[min(idx[r], V-1), :]) (clamp + nn.Embedding + ReLU, :112-116).
 * as_one_hot_bf16: out[r,:] = e_{min(idx[r], V-1)} as bf16 [R, Vpad]: operand of the GEMMs that replace an embedding's
 *   index_put backward and the fc over a one-hot action (head/action_type_head.py:60-61, action_arg_head.py:48-50).
 * as_time_encode: out[r,2k] = sin(t[r] pos[2k]), out[r,2k+1] = cos(t[r] pos[2k+1]) (:91-97).
 * as_bo_features: build-order token features [one_hot(order, A) | eye(L)[position] | bits_NB(x) | bits_NB(y) | 0] as
 *   bf16 [tokens, Kpad] (x = location % map_w, y = location / map_w, bits MSB first): operand of the embedding GEMM.
 * as_cast_any_bf16: [rows, K] of any of the dtypes above -> bf16 [rows, Kpad] zero padded (GEMM operand of an fc field). */
int as_embed_relu_fwd(const void* idx, int dtype, const float* table, float* out, int64_t R, int V, int D, void* stream);
int as_one_hot_bf16(const void* idx, int dtype, void* out, int64_t R, int V, int Vpad, void* stream);
int as_time_encode(const void* t, int dtype, const float* position_array, float* out, int64_t R, int D, void* stream);
int as_bo_features(const void* order, int order_dtype, const void* location, int location_dtype, void* out, int64_t tokens, int L,
                   int A, int NB, int map_w, int Kpad, void* stream);
int as_cast_any_bf16(const void* src, int dtype, void* dst, int64_t rows, int K, int Kpad, void* stream);

/* bf16 entity stream of the post-LN Transformer (module_utils.py:133-139): LayerNorm over 256 channels with the linear
 * output z, the residual, the result and every gradient in bf16 (statistics and reductions fp32).
 * as_ln256_bf16_fwd: y = LN(z + res) * gamma + beta (+ReLU), stats [rows,2] = (mean, rstd); res may be NULL.
 * as_ln256_bf16_bwd: g = dy_a + dy_b (dy_b may be NULL: the gradients of y's two consumers arrive separately) ->
 * dres (residual path, may be NULL), dz = dx * (z > 0 when mask_z), dgamma, dbeta, dbias = column sums of dz;
 * scratch = as_ln256_bf16_bwd_scratch_floats() floats.  as_masked_mean256_bf16_*: entity_encoder.py:83-85 on that stream. */
int as_ln256_bf16_fwd(const void* z, const void* res, const float* gamma, const float* beta, void* y, float* stats, int64_t rows,
                      int relu, void* stream);
int64_t as_ln256_bf16_bwd_scratch_floats(void);
int as_ln256_bf16_bwd(const void* dy_a, const void* dy_b, const void* z, const void* res, const float* gamma, const float* beta,
                      const float* stats, void* dres, void* dz, int mask_z, float* dgamma, float* dbeta, float* dbias,
                      float* scratch, int64_t rows, int relu, void* stream);
int as_masked_mean256_bf16_fwd(const void* x, const int64_t* len, float* out, int64_t R, int N, void* stream);
int as_masked_mean256_bf16_bwd(const float* dout, const int64_t* len, void* dx, int64_t R, int N, void* stream);

/* Tail of the location head (head/action_arg_head.py:440-446): bilinear x2 upsampling (F.interpolate, align_corners
 * False) of x [R,h+2,w+2,32] (padded NHWC bf16) + 3x3 convolution 32 -> 1 (zero padding) + flatten + 1/temperature ->
 * out [R, 2h * 2w] fp32, WITHOUT forming the upsampled map: the two linear steps commute into nine numbers per
 * low-resolution pixel (csrc/upconv.cu).  w9 [9][32] fp32 (tap-major), bias [1] or NULL, w <= 80.
 * Backward: dout [R, 2h * 2w] -> dx (x's layout, zero border; multiplied by (x > 0) when relu_input: x is a ReLU output
 * with no other consumer), dw9 [9][32], dbias [1]; scratch = as_upconv_c32_1_scratch_floats(R, h) floats. */
int as_upconv_c32_1_fwd(const void* x, const float* w9, const float* bias, float* out, int64_t R, int h, int w, float inv_temp,
                        void* stream);
int64_t as_upconv_c32_1_scratch_floats(int64_t R, int h);
int as_upconv_c32_1_bwd(const float* dout, const void* x, const float* w9, void* dx, float* dw9, float* dbias, float* scratch,
                        int64_t R, int h, int w, float inv_temp, int relu_input, void* stream);

/* bf16 GEMM-operand layouts of a Conv2d weight [Cout, Cin, k, k] fp32 (ctools/torch_utils/network/nn_module.py:119-174
 * holds it in that layout) for as_conv_nhwc_bf16: w_fwd [Cout][tap][cin_pad] (forward) and w_dgrad
 * [cin_pad][flipped tap][Cout] (data gradient), channels beyond Cin zero.  taps = k * k. */
int as_conv_weight_layouts(const float* weight, void* w_fwd, void* w_dgrad, int cout, int cin, int cin_pad, int taps, void* stream);

/* Spatial-encoder FC weight (obs_encoder/spatial_encoder.py:88-89, [out, C*H*W] over the NCHW flatten) re-indexed for
 * the zero-padded NHWC activation layout: w16 [out, (H+2)*(W+2)*C] bf16 (border columns zero); the backward maps the
 * weight gradient of that layout back to [out, C*H*W] fp32. */
int as_spatial_fc_weight_fwd(const float* weight, void* w16, int out_features, int C, int H, int W, void* stream);
int as_spatial_fc_weight_bwd(const float* dw_padded, float* dweight, int out_features, int C, int H, int W, void* stream);

/* Pointer logits of the TargetUnit head (head/action_arg_head.py:343-363: `query . key^T`, entity mask, 1/temperature)
 * and the key product of the SelectedUnits head (:190-196):
 *   out[r,s,n] = n < key_len[r] ? scale * sum_k query[r,s,k] key[r,n,k] : fill      (key_len NULL: no mask)
 * query [R,S,32], key [R,N,32], out [R,S,N] fp32, S <= 64, N <= 1024.  The backward gives dquery and / or dkey (either
 * may be NULL) from dout; masked positions carry no gradient. */
int as_pointer_logits_fwd(const float* query, const float* key, const int64_t* key_len, float* out, int64_t R, int S, int N, int D,
                          float scale, float fill, void* stream);
int as_pointer_logits_bwd(const float* dout, const float* query, const float* key, const int64_t* key_len, float* dquery, float* dkey,
                          int64_t R, int S, int N, int D, float scale, void* stream);

/* as_attn_fwd: fused attention forward of the entity Transformer (module_utils.py:71-111): out[r,n,h*D:(h+1)*D] =
 * softmax_j(q.k_j * scale, keys j >= key_len[r] filled with -1e9) V over the packed bf16 qkv [R, N, 3*H*D] (q | k | v
 * column blocks).  Scores stay in tensor memory; probs [R,H,N,N] bf16 (optional, NULL at inference) receives the
 * normalised probabilities the backward products read.  Supported when as_attn_fwd_ok(N, D) != 0 (D == 128, N a
 * multiple of 128 up to 512); otherwise AS_EUNSUPPORTED and the caller composes as_gemm_bf16 + as_softmax_fwd. */
int as_attn_fwd_ok(int N, int D);
int as_attn_fwd(const void* qkv, const int64_t* key_len, void* out, void* probs, int64_t R, int N, int H, int D, float scale,
                void* stream);
/* as_attn_bwd_dq: query side of the attention backward in one launch with the same pipeline: dP = dout V^T stays in
 * tensor memory, dscores = probs o (dP - rowsum(dout o out)) is written once as bf16 [R,H,N,N] (for the dK product)
 * and dq = scale * dscores K goes to the q column block of dqkv [R,N,3*H*D].  dK = scale dscores^T Q and
 * dV = probs^T dout remain as_gemm_bf16 calls. */
int as_attn_bwd_dq(const void* qkv, const void* out, const void* dout, const void* probs, void* dscores, void* dqkv, int64_t R, int N,
                   int H, int D, float scale, void* stream);

/* as_attn_small_fwd / _bwd: the same attention for short sequences and narrow heads (N <= 32, D in {8, 16}: the
 * build-order Transformer, obs_encoder/scalar_encoder.py:19-53) -- one warp per (row, head), fp32 in registers,
 * probabilities recomputed in the backward pass.  dqkv receives dq | dk | dv in the layout of qkv. */
int as_attn_small_ok(int N, int D);
int as_attn_small_fwd(const void* qkv, const int64_t* key_len, void* out, int64_t R, int N, int H, int D, float scale, void* stream);
int as_attn_small_bwd(const void* qkv, const int64_t* key_len, const void* dout, void* dqkv, int64_t R, int N, int H, int D,
                      float scale, void* stream);

/* as_su_sample: the whole SelectedUnits SAMPLING loop (head/action_arg_head.py:262-313: up to 64 autoregressive picks
 * per row) in one launch.  query_fc1 o (ae0 + embed_fc2(.)) is folded by the caller into c0 = Wq1 ae0 + bq1 [R,256],
 * c1 = Wq1 bf2 [256] and mt = (Wq1 Wf2)^T [256,256]; wf1t/wq2t/wih_t/whh_t are the transposed fp32 weights of
 * embed_fc1, query_fc2 and the H=32 LN-LSTM cell (lstm.py:138-153).  noise [S,R,N1] is Gumbel(0,1) (arg-max of
 * logits + noise == the reference's multinomial draw).  Writes logits [R,S,N1] (masked -1e9, / temperature), units
 * [R,S], num [R] (selected_units_num, :283), extra [R,N1] (:307-309), emb [R,32] (mean picked key after the last
 * step) and steps_row [R]: the reference's loop length (:305 early break) is max(1, max_r steps_row[r]). */
#ifndef AS_SU_SAMPLE_ARGS_DEFINED
#define AS_SU_SAMPLE_ARGS_DEFINED
typedef struct as_su_sample_args {
  const float *key, *c0, *c1, *mt, *wf1t, *bf1, *wq2t, *bq2, *wih_t, *whh_t;
  const float *ln_i_w, *ln_i_b, *ln_h_w, *ln_h_b, *ln_c_w, *ln_c_b;
  const int64_t* entity_num;
  const uint8_t* su_mask;
  const float* noise;
  float* logits;
  int64_t* units;
  int64_t* num;
  float* extra;
  float* emb;
  int* steps_row;
  int* ended_rows;   /* two device ints, 0 at launch, or NULL: [0] rows that have raised their end flag, [1] the latest end step;
                        once [0] == R a CTA leaves its loop as soon as it has computed [1] steps */
  int R, N1, S, extra_units;
  float temperature;
} as_su_sample_args;
#endif
int as_su_sample(const as_su_sample_args* args, void* stream);

/* bf16 feature rows for the entity embedding GEMM -- replaces the 28 embedding lookups + torch.cat at
 * obs_encoder/entity_encoder.py:60-78.  Per field f: data[f] = device array [rows] of element type dtype[f]
 * (0 u8, 1 i8, 2 i16, 3 f16, 4 i64, 5 f32), kind[f] (0 one-hot clamped to num-1, 1 num-bit binary MSB first,
 * 2 raw value), col[f] = first output column.  out [rows, pitch] bf16 (pitch % 8 == 0, zero padded).
 * *neg_flag (device int) is set to 1 if a one-hot index is negative (the reference raises, :68-71). */
int as_entity_features(const void* const* data, const int* kind, const int* num, const int* col, const int* dtype,
                       int n_fields, void* out, int64_t rows, int pitch, int* neg_flag, void* stream);

/* Gather scattered fp32 tensors into one flat buffer: table (device, 3 x int64 per chunk) = {source pointer or 0 for
 * zero fill, destination offset in elements, element count}.  Used once per all-reduce bucket to move the parameter
 * gradients autograd produced into the flat gradient buffer (instead of one accumulate launch per parameter,
 * rl_learner.py:118-125 / dist_helper.py:421-431 work per parameter tensor). */
int as_pack_f32(const int64_t* table, int chunks, float* dst, void* stream);

/* fp32 -> bf16 cast (operand staging for the GEMMs). */
int as_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream);

/* Gumbel-max sampling: out[r] = argmax_k(logit[r,k] + noise[r,k]), lowest index on ties -- replaces
 * softmax + torch.multinomial at action_type_head.py:57-58 and action_arg_head.py:46-47,79-80,147-148,361-362,447-449. */
int as_gumbel_argmax(const float* logit, const float* noise, int64_t* out, int64_t rows, int K, void* stream);

/* Fused RL loss -- replaces ReinforcementLoss.compute_loss (rl_training/rl_loss.py:33-185) and the scans of
 * rl_training/as_rl_utils.py (vtrace_advantages :284-312, upgo_returns :265-281, td_lambda_loss :221-243,
 * entropy_loss :52-72, kl_loss :75-103).  forward fills args->out (means + total_loss), args->dvalue and the
 * scratch buffers; backward writes args->dlogit scaled by grad_scale. */
int as_rl_loss_forward(const AsLossArgs* args, void* stream);
int as_rl_loss_backward(const AsLossArgs* args, float grad_scale, void* stream);

/* Global-norm clip + Adam on flat fp32 buffers -- replaces GradClip 'pytorch_norm'
 * (ctools/torch_utils/grad_clip.py:140-143) + torch.optim.Adam.step (agent/default/rl_learner.py:73-80,125,132).
 * scratch: as_clip_adam_scratch_bytes() bytes of device memory.  shadow (bf16 copy of p) may be NULL.
 * norm_out (device float, may be NULL) receives the pre-clip gradient norm.  step_dev (device int, may be NULL):
 * when given, the kernel increments it and derives the bias corrections from it (so a captured CUDA graph of the
 * step replays with the right step count); otherwise the host `step` (>= 1) is used. */
int64_t as_clip_adam_scratch_bytes(void);
int as_clip_adam(float* p, const float* g, float* m, float* v, void* shadow, int64_t n, float max_norm, float lr,
                 float beta1, float beta2, float eps, float weight_decay, int step, int* step_dev, float* scratch,
                 float* norm_out, void* stream);
/* The same with torch.optim.Adam's per-parameter behaviour: the flat buffers hold n_params tensors that start at
 * multiples of 8 elements (n % 8 == 0); block_param[i / 8] = tensor of element i; a tensor with param_active == 0
 * (its .grad was None this step, e.g. the policy while only_update_baseline is set: rl_learner.py:149-161) keeps its
 * weights and moments, the others use their own step count param_step[p] + 1 for the bias corrections; param_step is
 * incremented for the active tensors afterwards.  step_dev (may be NULL) is a global step counter, incremented. */
int as_clip_adam_per_param(float* p, const float* g, float* m, float* v, void* shadow, int64_t n, float max_norm, float lr,
                           float beta1, float beta2, float eps, float weight_decay, const uint16_t* block_param,
                           const int32_t* param_active, int32_t* param_step, int n_params, int* step_dev, float* scratch,
                           float* norm_out, void* stream);

/* ---- Diagnostics (tools/ only; no counterpart in the reference).  They set or read PROCESS-GLOBAL state and are not for
 * production callers: as_debug_gemm_mode changes how every later GEMM / convolution launch behaves (bits 0-2: epilogue
 * stages skipped for profiling, bit 3: register-store epilogue, bit 4: per-tap convolution mode); the *_profile entry
 * points install a device buffer that the issuing warps fill with cycle counters (NULL uninstalls it);
 * as_debug_lstm_max_clusters reports how many 8-CTA clusters of the recurrent kernel fit the device. */
int as_debug_gemm_mode(int mode);
int as_debug_gemm_profile(void* buf);
int as_debug_attn_profile(void* buf);
int as_debug_lstm_profile(void* buf);
int as_debug_lstm_max_clusters(int H);

#ifdef __cplusplus
}
#endif
#endif
