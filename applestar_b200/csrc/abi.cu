// extern "C" surface of libapplestar_b200.so (declarations + reference citations: include/applestar_b200.h).
#include "common.cuh"
#include "gemm.h"
#include "loss.h"
#include "su_sample.h"

namespace as {
extern int g_gemm_debug;
extern int g_gemm_no_tma_store;
extern int g_gemm_no_slab;
int gemm_set_profile(long long*);
int attn_fwd_ok(int, int);
int attn_set_profile(long long*);
int attn_small_ok(int, int);
int attn_small_fwd(const void*, const long long*, void*, long long, int, int, int, float, cudaStream_t);
int attn_small_bwd(const void*, const long long*, const void*, void*, long long, int, int, int, float, cudaStream_t);
int attn_fwd(const void*, const long long*, void*, void*, long long, int, int, int, float, cudaStream_t);
int attn_bwd_dq(const void*, const void*, const void*, const void*, void*, void*, long long, int, int, int, float, cudaStream_t);
int scatter_fwd(const float*, const void*, const void*, int, const long long*, float*, long long, int, int, int, int,
                int, cudaStream_t);
int scatter_bwd(const float*, const void*, const void*, int, const long long*, float*, long long, int, int, int, int,
                cudaStream_t);
int cast_f32_bf16(const float*, void*, long long, cudaStream_t);
int pack_f32(const long long*, int, float*, cudaStream_t);
int relu_bwd_f32(const float*, const float*, float*, long long, cudaStream_t);
int gumbel_argmax(const float*, const float*, long long*, long long, int, cudaStream_t);
int loss_forward(const AsLossArgs&, cudaStream_t);
int loss_backward(const AsLossArgs&, float, cudaStream_t);
long long clip_adam_scratch_bytes();
int clip_adam_per_param(float*, const float*, float*, float*, void*, long long, float, float, float, float, float, float,
                        const unsigned short*, const int*, int*, int, int*, float*, float*, cudaStream_t);
int clip_adam(float*, const float*, float*, float*, void*, long long, float, float, float, float, float, float, int,
              int*, float*, float*, cudaStream_t);
int layer_norm_fwd(const float*, const float*, const float*, const float*, float*, void*, float*, long long, int, int,
                   cudaStream_t);
long long layer_norm_bwd_scratch_floats(int);
long long layer_norm_bwd_fused_scratch_floats(int);
int layer_norm_bwd_fused(const float*, const void*, const float*, const float*, const float*, const float*, const float*, float*,
                         void*, int, float*, float*, float*, float*, long long, int, int, cudaStream_t);
int layer_norm_bwd(const float*, const void*, const float*, const float*, const float*, const float*, const float*,
                   float*, float*, float*, float*, long long, int, int, cudaStream_t);
int softmax_fwd(const void*, const long long*, void*, long long, int, int, int, int, cudaStream_t);
int softmax_bwd(const void*, const void*, void*, long long, int, int, cudaStream_t);
int entity_features(const void* const*, const int*, const int*, const int*, const int*, int, void*, long long, int,
                    int*, cudaStream_t);
int nchw_to_pnhwc(const float*, void*, long long, int, int, int, int, int, cudaStream_t);
int pnhwc_to_nchw(const void*, float*, long long, int, int, int, int, cudaStream_t);
int add_act(const void*, const void*, void*, long long, int, cudaStream_t);
int relu_bwd(const void*, const void*, void*, long long, cudaStream_t);
int gated_fwd(const void*, const void*, const void*, const float*, void*, long long, cudaStream_t);
long long gated_scratch_floats();
int gated_bwd(const void*, const void*, const void*, const void*, const float*, void*, void*, void*, float*, float*,
              long long, cudaStream_t);
int up2_fwd(const void*, void*, long long, int, int, int, cudaStream_t);
int up2_bwd(const void*, void*, long long, int, int, int, cudaStream_t);
int loc_cat_fwd(const float*, const void*, void*, long long, int, int, int, int, cudaStream_t);
int loc_cat_bwd(const void*, const float*, const void*, float*, void*, long long, int, int, int, int, cudaStream_t);
int conv_c32_1_fwd(const void*, const float*, const float*, float*, long long, int, int, float, cudaStream_t);
long long conv_c32_1_scratch_floats();
int conv_c32_1_bwd(const float*, const void*, const float*, void*, float*, float*, float*, long long, int, int, float,
                   cudaStream_t);
long long colsum_scratch_floats(int);
int colsum_bf16(const void*, float*, float*, long long, int, cudaStream_t);
int colsum_bf16_static(const void*, float*, long long, int, cudaStream_t);
int relu_bwd_colsum(const void*, const void*, void*, float*, float*, long long, int, cudaStream_t);
int su_prefix_fwd(const float*, const long long*, const long long*, const long long*, float*, float*, unsigned char*, long long,
                  int, int, int, int, cudaStream_t);
int su_prefix_bwd(const float*, const long long*, const long long*, const float*, const unsigned char*, float*, long long, int, int,
                  int, int, cudaStream_t);
int su_mask_logits(const float*, const long long*, const long long*, float*, long long, int, int, int, float, float, cudaStream_t);
int conv_weight_layouts(const float*, void*, void*, int, int, int, int, cudaStream_t);
int expand_rows(const float*, const int*, float*, long long, int, float, cudaStream_t);
int pointer_keys_fwd(const float*, const float*, const long long*, float*, long long, int, int, cudaStream_t);
int pointer_keys_bwd(const float*, const long long*, float*, float*, long long, int, int, cudaStream_t);
int ce_rows_fwd(const float*, long long, const long long*, const long long*, const long long*, int, int, float, float*, float*, int*, long long,
                cudaStream_t);
int ce_rows_bwd(const float*, long long, const long long*, const long long*, const long long*, int, int, float, const float*, const float*,
                float*, long long, cudaStream_t);
int embed_relu_fwd(const void*, int, const float*, float*, long long, int, int, cudaStream_t);
int one_hot_bf16(const void*, int, void*, long long, int, int, cudaStream_t);
int time_encode(const void*, int, const float*, float*, long long, int, cudaStream_t);
int bo_features(const void*, int, const void*, int, void*, long long, int, int, int, int, int, cudaStream_t);
int cast_any_bf16(const void*, int, void*, long long, int, int, cudaStream_t);
int ln256_bf16_fwd(const void*, const void*, const float*, const float*, void*, float*, long long, int, cudaStream_t);
long long ln256_bf16_bwd_scratch_floats();
int ln256_bf16_bwd(const void*, const void*, const void*, const void*, const float*, const float*, const float*, void*, void*, int,
                   float*, float*, float*, float*, long long, int, cudaStream_t);
int masked_mean256_bf16_fwd(const void*, const long long*, float*, long long, int, cudaStream_t);
int masked_mean256_bf16_bwd(const float*, const long long*, void*, long long, int, cudaStream_t);
int upconv_fwd(const void*, const float*, const float*, float*, long long, int, int, float, cudaStream_t);
long long upconv_scratch_floats(long long, int);
int upconv_bwd(const float*, const void*, const float*, void*, float*, float*, float*, long long, int, int, float, int, cudaStream_t);
int spatial_front_fwd(const void* const*, const void* const*, int, const float*, const void*, const void*, int, const long long*, int,
                      const float*, const float*, void*, void*, void*, long long, int, int, cudaStream_t);
long long spatial_front_cells_bytes(long long);
long long spatial_front_bwd_scratch_floats(long long, int);
int spatial_front_bwd(const void* const*, const void* const*, int, const float*, const void*, const void*, int, const long long*, int,
                      const float*, const float*, const void*, const void*, const void*, float*, float*, float*, float*, long long, int,
                      int, cudaStream_t);
int spatial_fc_weight_fwd(const float*, void*, int, int, int, int, cudaStream_t);
int spatial_fc_weight_bwd(const float*, float*, int, int, int, int, cudaStream_t);
int pointer_logits_fwd(const float*, const float*, const long long*, float*, long long, int, int, int, float, float, cudaStream_t);
int pointer_logits_bwd(const float*, const float*, const float*, const long long*, float*, float*, long long, int, int, int, float,
                       cudaStream_t);
int masked_mean_fwd(const float*, const long long*, float*, long long, int, int, cudaStream_t);
int masked_mean_bwd(const float*, const long long*, float*, long long, int, int, cudaStream_t);
int linear_bwd_prep(const float*, const void*, const float*, const void*, void*, float*, float*, long long, int, cudaStream_t);
int colsum_f32(const float*, float*, float*, long long, int, cudaStream_t);
int lnlstm_fwd(const float*, const float*, const float*, const void*, const float*, const float*, const float*,
               const float*, float*, float*, float*, float*, float*, int, int, int, cudaStream_t);
bool lnlstm_cluster_ok(int);
int lnlstm_cluster_max_active(int);
int lnlstm_set_profile(long long*);
int lnlstm_bwd(const float*, const float*, const float*, const float*, const float*, const float*, const float*,
               const float*, const void*, const float*, const float*, const float*, const float*, const float*,
               const float*, const float*, float*, float*, float*, float*, float*, int, int, int, cudaStream_t);
int spatial_features(const void* const*, const void* const*, int, const float*, const float*, const void*, const void*, int,
                     const long long*, int, void*, long long, int, int, int, cudaStream_t);
int scatter_bwd_pnhwc(const void*, const void*, const void*, int, const long long*, float*, long long, int, int, int,
                      int, int, int, cudaStream_t);
int maxpool2_fwd(const void*, void*, long long, int, int, int, cudaStream_t);
int maxpool2_bwd(const void*, const void*, void*, long long, int, int, int, int, cudaStream_t);
}  // namespace as

#define ST(s) reinterpret_cast<cudaStream_t>(s)

namespace as {
const Knobs& knobs() {
  static const Knobs k = [] {
    Knobs v{};
    v.wgrad_no_pair = getenv("AS_WGRAD_NO_PAIR") != nullptr;
    v.wgrad_no_slab = getenv("AS_WGRAD_NO_SLAB") != nullptr;
    v.slab_no_resident = getenv("AS_SLAB_NO_RESIDENT") != nullptr;
    v.slab_mc = getenv("AS_SLAB_MC") != nullptr;       // CTA-pair weight multicast (this GPU's own weight buffer only); slower, off
    v.slab_no_halfk = getenv("AS_SLAB_NO_HALFK") != nullptr;
    v.debug_tmap = getenv("AS_DEBUG_TMAP") != nullptr;
    v.no_plain_epilogue = getenv("AS_NO_PLAIN_EPILOGUE") != nullptr;
    const char* e = getenv("AS_SLAB_BSLOTS");
    v.slab_bslots = e ? atoi(e) : 0;
    return v;
  }();
  return k;
}
}  // namespace as

extern "C" {

int as_abi_version(void) { return 5; }

// diagnostics only (not part of the public header): 0 normal, 1 GEMM epilogue skips global stores, 2 only drains TMEM
// bit 3: register-store epilogue, bit 4: per-tap convolution mode
int as_debug_gemm_mode(int mode) {
  as::g_gemm_debug = mode & 7; as::g_gemm_no_tma_store = (mode >> 3) & 1;
  as::g_gemm_no_slab = (mode >> 4) & 1;
  return 0;
}

int as_scatter_connection_fwd(const float* emb, const void* xs, const void* ys, int idx_bytes,
                              const int64_t* entity_num, float* out, int64_t R, int N, int D, int H, int W, int cover,
                              void* stream) {
  return as::scatter_fwd(emb, xs, ys, idx_bytes, reinterpret_cast<const long long*>(entity_num), out, R, N, D, H, W,
                         cover, ST(stream));
}

int as_scatter_connection_bwd(const float* d_out, const void* xs, const void* ys, int idx_bytes,
                              const int64_t* entity_num, float* d_emb, int64_t R, int N, int D, int H, int W,
                              void* stream) {
  return as::scatter_bwd(d_out, xs, ys, idx_bytes, reinterpret_cast<const long long*>(entity_num), d_emb, R, N, D, H,
                         W, ST(stream));
}

int as_gemm_splits(int K, int want) {
  const int num_kb = (K + 63) / 64;
  if (want <= 1) return 1;
  const int per = (num_kb + want - 1) / want;
  return (num_kb + per - 1) / per;
}

int as_gemm_bf16(const void* A, const void* B, void* C, void* C2, const float* bias, int M, int N, int K, int a_mn,
                 int b_mn,
                 int64_t lda, int64_t ldb, int64_t ldc, int nb0, int nb1, int64_t a_sb0, int64_t a_sb1, int64_t b_sb0,
                 int64_t b_sb1, int64_t c_sb0, int64_t c_sb1, int c_bf16, int relu, float alpha, int splits, float* ws,
                 int b_k_shift, void* stream) {
  as::GemmCall g;
  g.A = A; g.B = B; g.C = C; g.C2 = C2; g.bias = bias;
  g.M = M; g.N = N; g.K = K;
  g.b_k_shift = b_k_shift;
  g.a_mn = a_mn != 0; g.b_mn = b_mn != 0;
  g.lda = lda; g.ldb = ldb; g.ldc = ldc;
  g.nb0 = nb0; g.nb1 = nb1;
  g.a_sb0 = a_sb0; g.a_sb1 = a_sb1; g.b_sb0 = b_sb0; g.b_sb1 = b_sb1; g.c_sb0 = c_sb0; g.c_sb1 = c_sb1;
  g.c_bf16 = c_bf16; g.relu = relu; g.alpha = alpha;
  g.splits = splits; g.ws = ws;
  return as::gemm_run(g, ST(stream));
}

int as_gemm_bf16_masked(const void* A, const void* B, void* C, const void* mask, int M, int N, int K, int a_mn, int b_mn, int64_t lda,
                        int64_t ldb, int64_t ldc, int c_bf16, float alpha, void* stream) {
  as::GemmCall g;
  g.A = A; g.B = B; g.C = C; g.mask = mask;
  g.M = M; g.N = N; g.K = K;
  g.a_mn = a_mn != 0; g.b_mn = b_mn != 0;
  g.lda = lda; g.ldb = ldb; g.ldc = ldc;
  g.c_bf16 = c_bf16; g.alpha = alpha;
  return as::gemm_run(g, ST(stream));
}

int as_gemm_bf16_relu_bits(const void* A, const void* B, void* C, const float* bias, uint32_t* bits, int M, int N, int K, int64_t lda,
                           int64_t ldb, int64_t ldc, void* stream) {
  as::GemmCall g;
  g.A = A; g.B = B; g.C = C; g.bias = bias; g.bits_out = bits;
  g.M = M; g.N = N; g.K = K;
  g.lda = lda; g.ldb = ldb; g.ldc = ldc;
  g.c_bf16 = 1; g.relu = 1;
  return as::gemm_run(g, ST(stream));
}

int as_gemm_bf16_bitmasked(const void* A, const void* B, void* C, const uint32_t* bits, int M, int N, int K, int a_mn, int b_mn,
                           int64_t lda, int64_t ldb, int64_t ldc, int c_bf16, float alpha, void* stream) {
  as::GemmCall g;
  g.A = A; g.B = B; g.C = C; g.bits_in = bits;
  g.M = M; g.N = N; g.K = K;
  g.a_mn = a_mn != 0; g.b_mn = b_mn != 0;
  g.lda = lda; g.ldb = ldb; g.ldc = ldc;
  g.c_bf16 = c_bf16; g.alpha = alpha;
  return as::gemm_run(g, ST(stream));
}

int as_splitk_reduce(const float* ws, float* out, int64_t n, int splits, int accumulate, void* stream) {
  if (!ws || !out || splits <= 0) return AS_EINVAL;
  return as::splitk_reduce(ws, out, n, splits, accumulate, ST(stream));
}

static int conv_nhwc_impl(const void* x, const void* w, void* y, const float* bias, const void* resid, int64_t R, int H, int W,
                          int Cin, int Cout, int taps, int w_mn, int y_bf16, int relu, void* stream, const void* mask = nullptr,
                          uint32_t* bits_out = nullptr, const uint32_t* bits_in = nullptr) {
  if (taps != 1 && taps != 9) return AS_EUNSUPPORTED;
  if (R <= 0 || H <= 0 || W <= 0 || Cin <= 0 || Cout <= 0) return AS_ESIZE;
  if (Cin % 8 != 0) return AS_EUNSUPPORTED;
  const int pw = W + 2, ph = H + 2;
  const int64_t P = R * (int64_t)ph * pw;
  if (P > 2147483647LL) return AS_ESIZE;
  as::GemmCall g;
  g.A = x; g.B = w; g.C = y; g.bias = bias; g.resid = resid; g.mask = mask; g.bits_out = bits_out; g.bits_in = bits_in;
  g.M = (int)P; g.N = Cout; g.K = Cin;
  g.a_mn = false; g.b_mn = w_mn != 0;
  g.lda = Cin; g.ldb = g.b_mn ? Cout : (int64_t)taps * Cin; g.ldc = Cout;
  g.c_bf16 = y_bf16; g.relu = relu; g.alpha = 1.f;
  g.taps = taps;
  if (taps == 9) {
    for (int ky = 0; ky < 3; ++ky)
      for (int kx = 0; kx < 3; ++kx) g.shift[ky * 3 + kx] = (ky - 1) * pw + (kx - 1);
  } else {
    g.shift[0] = 0;
  }
  g.img_w = pw; g.img_h = ph;
  return as::gemm_run(g, ST(stream));
}
int as_conv_nhwc_bf16(const void* x, const void* w, void* y, const float* bias, int64_t R, int H, int W, int Cin,
                      int Cout, int taps, int w_mn, int y_bf16, int relu, void* stream) {
  return conv_nhwc_impl(x, w, y, bias, nullptr, R, H, W, Cin, Cout, taps, w_mn, y_bf16, relu, stream);
}
int as_conv_nhwc_bf16_add(const void* x, const void* w, void* y, const float* bias, const void* addend, int64_t R, int H,
                          int W, int Cin, int Cout, int taps, int relu, void* stream) {
  if (!addend) return AS_EINVAL;
  return conv_nhwc_impl(x, w, y, bias, addend, R, H, W, Cin, Cout, taps, 0, 1, relu, stream);
}

int as_conv_nhwc_bf16_masked(const void* x, const void* w, void* y, const void* addend, const void* mask, int64_t R, int H, int W,
                             int Cin, int Cout, int taps, void* stream) {
  if (!mask) return AS_EINVAL;
  return conv_nhwc_impl(x, w, y, nullptr, addend, R, H, W, Cin, Cout, taps, 0, 1, 0, stream, mask);
}

int as_conv_nhwc_bf16_relu_bits(const void* x, const void* w, void* y, const float* bias, uint32_t* bits, int64_t R, int H, int W, int Cin,
                                int Cout, int taps, void* stream) {
  if (!bits || (Cout % 32) != 0) return AS_EINVAL;
  return conv_nhwc_impl(x, w, y, bias, nullptr, R, H, W, Cin, Cout, taps, 0, 1, 1, stream, nullptr, bits, nullptr);
}
int as_conv_nhwc_bf16_bitmasked(const void* x, const void* w, void* y, const void* addend, const uint32_t* bits, int64_t R, int H, int W,
                                int Cin, int Cout, int taps, void* stream) {
  if (!bits || (Cout % 32) != 0) return AS_EUNSUPPORTED;
  return conv_nhwc_impl(x, w, y, nullptr, addend, R, H, W, Cin, Cout, taps, 0, 1, 0, stream, nullptr, nullptr, bits);
}

// 0: one tap per tile; 1 / 2: tap pairs on a 64-channel M side (dy / swapped x), needs split-K partials
// 11 / 12: the same pairing with all nine taps walked from one shared-memory slab per pixel block (wgrad_slab.cu)
int as_conv_wgrad_pair_mode(int Cin, int Cout, int taps, int splits, int W) {
  if (taps != 9 || splits <= 1 || as::knobs().wgrad_no_pair) return 0;
  int mode = 0;
  if (Cout == 64) mode = 1;
  else if (Cin == 64 && Cout < 64) mode = 2;
  if (mode && !as::knobs().wgrad_no_slab && as::wgrad_slab_ok(mode, Cin, Cout, W + 2)) mode += 10;
  // 64 -> 128 channels: the plain form has a full M = 128; the slab form (operands swapped, two column halves) still wins by
  // reading the pixels once
  if (!mode && Cin == 64 && Cout == 128 && !as::knobs().wgrad_no_slab && as::wgrad_slab_ok(2, Cin, Cout, W + 2)) mode = 12;
  return mode;
}
int64_t as_conv_wgrad_ws_floats(int Cin, int Cout, int taps, int splits, int W) {
  if (splits <= 1) return 0;
  const int pair = as_conv_wgrad_pair_mode(Cin, Cout, taps, splits, W) % 10;
  if (pair) return (int64_t)splits * 5 * 128 * (pair == 1 ? Cin : Cout);
  return (int64_t)splits * taps * Cout * Cin;
}
int as_conv_wgrad_nhwc_bf16(const void* dy, const void* x, float* dw, float* ws, int64_t R, int H, int W, int Cin,
                            int Cout, int taps, int splits, void* stream) {
  if (taps != 1 && taps != 9) return AS_EUNSUPPORTED;
  if (R <= 0 || H <= 0 || W <= 0 || Cin <= 0 || Cout <= 0) return AS_ESIZE;
  if ((Cin % 8) || (Cout % 8)) return AS_EALIGN;
  const int pw = W + 2, ph = H + 2;
  const int64_t P = R * (int64_t)ph * pw;
  if (P > 2147483647LL) return AS_ESIZE;
  const int pmode = as_conv_wgrad_pair_mode(Cin, Cout, taps, splits, W);
  const int pair = pmode % 10;
  if (pmode >= 10) {
    const int n = pair == 1 ? Cin : Cout;
    int rc = as::wgrad_slab(pair, dy, x, ws, P, pw, Cin, Cout, splits, ST(stream));
    if (rc) return rc;
    return as::wgrad_pair_reduce(ws, dw, splits, 5LL * 128 * n, taps, Cout, Cin, pair == 2, pair == 1, ST(stream));
  }
  if (pair) {
    // two taps per 128-row MMA (gemm_tcgen05.cu, wg_pair).  mode 1: M = Cout = 64, the second tap's dy rows are read
    // shift[b] - shift[a] earlier; mode 2 (Cout < 64 = Cin): operands swapped, M = Cin = 64, both x chunks shifted
    int shift[10];
    for (int ky = 0; ky < 3; ++ky)
      for (int kx = 0; kx < 3; ++kx) shift[ky * 3 + kx] = (ky - 1) * pw + (kx - 1);
    shift[9] = shift[8];
    as::GemmCall g;
    g.K = (int)P; g.a_mn = true; g.b_mn = true; g.nb0 = 5; g.nb1 = 1; g.wgrad_taps = 5; g.wg_pair = 1;
    g.M = 64;
    if (pair == 1) {
      g.A = dy; g.B = x; g.N = Cin; g.lda = Cout; g.ldb = Cin;
      for (int b = 0; b < 5; ++b) { g.pair_a[b][0] = 0; g.pair_a[b][1] = -(shift[2 * b + 1] - shift[2 * b]); g.pair_b[b] = shift[2 * b]; }
    } else {
      g.A = x; g.B = dy; g.N = Cout; g.lda = Cin; g.ldb = Cout;
      for (int b = 0; b < 5; ++b) { g.pair_a[b][0] = shift[2 * b]; g.pair_a[b][1] = shift[2 * b + 1]; g.pair_b[b] = 0; }
    }
    g.C = dw; g.ldc = g.N; g.c_sb0 = 128LL * g.N; g.c_bf16 = 0;
    g.splits = splits; g.ws = ws;
    int rc = as::gemm_run(g, ST(stream));
    if (rc) return rc;
    return as::wgrad_pair_reduce(ws, dw, splits, 5LL * 128 * g.N, taps, Cout, Cin, pair == 2, 0, ST(stream));
  }
  as::GemmCall g;
  g.A = dy; g.B = x;
  g.M = Cout; g.N = Cin; g.K = (int)P;
  g.a_mn = true; g.b_mn = true;
  g.lda = Cout; g.ldb = Cin;
  g.nb0 = taps; g.nb1 = 1;
  g.wgrad_taps = taps;
  if (taps == 9) {
    for (int ky = 0; ky < 3; ++ky)
      for (int kx = 0; kx < 3; ++kx) g.shift[ky * 3 + kx] = (ky - 1) * pw + (kx - 1);
  } else {
    g.shift[0] = 0;
  }
  g.C = dw; g.ldc = Cin; g.c_sb0 = (int64_t)Cout * Cin; g.c_bf16 = 0;
  g.splits = splits; g.ws = ws;
  int rc = as::gemm_run(g, ST(stream));
  if (rc) return rc;
  if (splits > 1) return as::splitk_reduce(ws, dw, (int64_t)taps * Cout * Cin, splits, 0, ST(stream));
  return AS_OK;
}

int as_nchw_to_pnhwc(const float* src, void* dst, int64_t R, int C, int Cp, int H, int W, int relu, void* stream) {
  return as::nchw_to_pnhwc(src, dst, R, C, Cp, H, W, relu, ST(stream));
}
int as_pnhwc_to_nchw(const void* src, float* dst, int64_t R, int C, int Cp, int H, int W, void* stream) {
  return as::pnhwc_to_nchw(src, dst, R, C, Cp, H, W, ST(stream));
}
int as_add_act_bf16(const void* a, const void* b, void* y, int64_t n, int mode, void* stream) {
  return as::add_act(a, b, y, n, mode, ST(stream));
}
int as_relu_bwd_bf16(const void* dy, const void* y, void* dz, int64_t n, void* stream) {
  return as::relu_bwd(dy, y, dz, n, ST(stream));
}
int as_gated_fwd_bf16(const void* r, const void* g, const void* x, const float* update_sp, void* y, int64_t n,
                      void* stream) {
  return as::gated_fwd(r, g, x, update_sp, y, n, ST(stream));
}
int64_t as_gated_scratch_floats(void) { return as::gated_scratch_floats(); }
int as_gated_bwd_bf16(const void* dy, const void* y, const void* r, const void* g, const float* update_sp, void* dr,
                      void* dg, void* dx, float* d_update_sp, float* scratch, int64_t n, void* stream) {
  return as::gated_bwd(dy, y, r, g, update_sp, dr, dg, dx, d_update_sp, scratch, n, ST(stream));
}
int as_upsample2_fwd_bf16(const void* in, void* out, int64_t R, int H, int W, int C, void* stream) {
  return as::up2_fwd(in, out, R, H, W, C, ST(stream));
}
int as_upsample2_bwd_bf16(const void* dout, void* din, int64_t R, int H, int W, int C, void* stream) {
  return as::up2_bwd(dout, din, R, H, W, C, ST(stream));
}
int as_loc_cat_fwd(const float* emb, const void* skip, void* out, int64_t R, int H, int W, int Cs, int Cp, void* stream) {
  return as::loc_cat_fwd(emb, skip, out, R, H, W, Cs, Cp, ST(stream));
}
int as_loc_cat_bwd(const void* dcat, const float* emb, const void* skip, float* d_emb, void* d_skip, int64_t R, int H,
                   int W, int Cs, int Cp, void* stream) {
  return as::loc_cat_bwd(dcat, emb, skip, d_emb, d_skip, R, H, W, Cs, Cp, ST(stream));
}
int as_conv_c32_1_fwd(const void* in, const float* w, const float* bias, float* logits, int64_t R, int H, int W,
                      float inv_temperature, void* stream) {
  return as::conv_c32_1_fwd(in, w, bias, logits, R, H, W, inv_temperature, ST(stream));
}
int64_t as_conv_c32_1_scratch_floats(void) { return as::conv_c32_1_scratch_floats(); }
int as_conv_c32_1_bwd(const float* dlogits, const void* in, const float* w, void* d_in, float* dw, float* db,
                      float* scratch, int64_t R, int H, int W, float inv_temperature, void* stream) {
  return as::conv_c32_1_bwd(dlogits, in, w, d_in, dw, db, scratch, R, H, W, inv_temperature, ST(stream));
}
int64_t as_colsum_scratch_floats(int C) { return as::colsum_scratch_floats(C); }
int as_colsum_bf16(const void* x, float* out, float* scratch, int64_t rows, int C, void* stream) {
  return as::colsum_bf16(x, out, scratch, rows, C, ST(stream));
}

int as_spatial_features(const void* const* planes, const void* const* effects, int n_eff, const float* scatter_map,
                        void* out, int64_t R, int H, int W, int Cs, void* stream) {
  return as::spatial_features(planes, effects, n_eff, scatter_map, nullptr, nullptr, nullptr, 0, nullptr, 0, out, R, H, W, Cs,
                              ST(stream));
}
int as_spatial_features_scatter(const void* const* planes, const void* const* effects, int n_eff, const float* emb, const void* xs,
                                const void* ys, int idx_bytes, const int64_t* entity_num, int N, void* out, int64_t R, int H,
                                int W, void* stream) {
  return as::spatial_features(planes, effects, n_eff, nullptr, emb, xs, ys, idx_bytes,
                              reinterpret_cast<const long long*>(entity_num), N, out, R, H, W, 32, ST(stream));
}
int as_scatter_connection_bwd_pnhwc(const void* d_feat, const void* xs, const void* ys, int idx_bytes,
                                    const int64_t* entity_num, float* d_emb, int64_t R, int N, int D, int H, int W,
                                    int Cp, int ch0, void* stream) {
  return as::scatter_bwd_pnhwc(d_feat, xs, ys, idx_bytes, reinterpret_cast<const long long*>(entity_num), d_emb, R, N,
                               D, H, W, Cp, ch0, ST(stream));
}
int as_maxpool2_fwd_bf16(const void* in, void* out, int64_t R, int H, int W, int C, void* stream) {
  return as::maxpool2_fwd(in, out, R, H, W, C, ST(stream));
}
int as_maxpool2_bwd_bf16(const void* in, const void* dout, void* din, int64_t R, int H, int W, int C, int relu_input, void* stream) {
  return as::maxpool2_bwd(in, dout, din, R, H, W, C, relu_input, ST(stream));
}

int as_colsum_f32(const float* x, float* out, float* scratch, int64_t rows, int C, void* stream) {
  return as::colsum_f32(x, out, scratch, rows, C, ST(stream));
}
int as_relu_bwd_colsum_bf16(const void* dy, const void* y, void* dz, float* colsum, float* scratch, int64_t rows, int C,
                            void* stream) {
  return as::relu_bwd_colsum(dy, y, dz, colsum, scratch, rows, C, ST(stream));
}
int as_su_prefix_fwd(const float* key, const int64_t* selected_units, const int64_t* entity_num, const int64_t* su_num, float* emb,
                     float* cnt, uint8_t* isnew, int64_t R, int N1, int D, int Smax, int S, void* stream) {
  return as::su_prefix_fwd(key, reinterpret_cast<const long long*>(selected_units), reinterpret_cast<const long long*>(entity_num),
                           reinterpret_cast<const long long*>(su_num), emb, cnt, isnew, R, N1, D, Smax, S, ST(stream));
}
int as_su_prefix_bwd(const float* demb, const int64_t* selected_units, const int64_t* su_num, const float* cnt, const uint8_t* isnew,
                     float* dkey, int64_t R, int N1, int D, int Smax, int S, void* stream) {
  return as::su_prefix_bwd(demb, reinterpret_cast<const long long*>(selected_units), reinterpret_cast<const long long*>(su_num), cnt,
                           isnew, dkey, R, N1, D, Smax, S, ST(stream));
}
int as_su_mask_logits(const float* x, const int64_t* selected_units, const int64_t* entity_num, float* out, int64_t R, int N1,
                      int Smax, int S, float scale, float fill, void* stream) {
  return as::su_mask_logits(x, reinterpret_cast<const long long*>(selected_units), reinterpret_cast<const long long*>(entity_num), out,
                            R, N1, Smax, S, scale, fill, ST(stream));
}
int as_pointer_logits_fwd(const float* query, const float* key, const int64_t* key_len, float* out, int64_t R, int S, int N, int D,
                          float scale, float fill, void* stream) {
  return as::pointer_logits_fwd(query, key, reinterpret_cast<const long long*>(key_len), out, R, S, N, D, scale, fill, ST(stream));
}
int as_pointer_logits_bwd(const float* dout, const float* query, const float* key, const int64_t* key_len, float* dquery, float* dkey,
                          int64_t R, int S, int N, int D, float scale, void* stream) {
  return as::pointer_logits_bwd(dout, query, key, reinterpret_cast<const long long*>(key_len), dquery, dkey, R, S, N, D, scale,
                                ST(stream));
}
int as_conv_weight_layouts(const float* weight, void* w_fwd, void* w_dgrad, int cout, int cin, int cin_pad, int taps, void* stream) {
  return as::conv_weight_layouts(weight, w_fwd, w_dgrad, cout, cin, cin_pad, taps, ST(stream));
}
int as_spatial_fc_weight_fwd(const float* weight, void* w16, int out_features, int C, int H, int W, void* stream) {
  return as::spatial_fc_weight_fwd(weight, w16, out_features, C, H, W, ST(stream));
}
int as_spatial_fc_weight_bwd(const float* dw_padded, float* dweight, int out_features, int C, int H, int W, void* stream) {
  return as::spatial_fc_weight_bwd(dw_padded, dweight, out_features, C, H, W, ST(stream));
}
int as_spatial_front_fwd(const void* const* planes, const void* const* effects, int n_eff, const float* emb, const void* xs,
                         const void* ys, int idx_bytes, const int64_t* entity_num, int N, const float* weight, const float* bias,
                         void* cells, void* out, void* codes, int64_t R, int H, int W, void* stream) {
  return as::spatial_front_fwd(planes, effects, n_eff, emb, xs, ys, idx_bytes, reinterpret_cast<const long long*>(entity_num), N,
                               weight, bias, cells, out, codes, R, H, W, ST(stream));
}
int64_t as_spatial_front_cells_bytes(int64_t R) { return as::spatial_front_cells_bytes(R); }
int64_t as_spatial_front_bwd_scratch_floats(int64_t R, int H) { return as::spatial_front_bwd_scratch_floats(R, H); }
int as_spatial_front_bwd(const void* const* planes, const void* const* effects, int n_eff, const float* emb, const void* xs,
                         const void* ys, int idx_bytes, const int64_t* entity_num, int N, const float* weight, const float* bias,
                         const void* cells, const void* codes, const void* dpool, float* dweight, float* dbias, float* g_entity,
                         float* scratch, int64_t R, int H, int W, void* stream) {
  return as::spatial_front_bwd(planes, effects, n_eff, emb, xs, ys, idx_bytes, reinterpret_cast<const long long*>(entity_num), N,
                               weight, bias, cells, codes, dpool, dweight, dbias, g_entity, scratch, R, H, W, ST(stream));
}
int as_upconv_c32_1_fwd(const void* x, const float* w9, const float* bias, float* out, int64_t R, int h, int w, float inv_temp,
                        void* stream) {
  return as::upconv_fwd(x, w9, bias, out, R, h, w, inv_temp, ST(stream));
}
int64_t as_upconv_c32_1_scratch_floats(int64_t R, int h) { return as::upconv_scratch_floats(R, h); }
int as_upconv_c32_1_bwd(const float* dout, const void* x, const float* w9, void* dx, float* dw9, float* dbias, float* scratch,
                        int64_t R, int h, int w, float inv_temp, int relu_input, void* stream) {
  return as::upconv_bwd(dout, x, w9, dx, dw9, dbias, scratch, R, h, w, inv_temp, relu_input, ST(stream));
}
int as_ln256_bf16_fwd(const void* z, const void* res, const float* gamma, const float* beta, void* y, float* stats, int64_t rows,
                      int relu, void* stream) {
  return as::ln256_bf16_fwd(z, res, gamma, beta, y, stats, rows, relu, ST(stream));
}
int64_t as_ln256_bf16_bwd_scratch_floats(void) { return as::ln256_bf16_bwd_scratch_floats(); }
int as_ln256_bf16_bwd(const void* dy_a, const void* dy_b, const void* z, const void* res, const float* gamma, const float* beta,
                      const float* stats, void* dres, void* dz, int mask_z, float* dgamma, float* dbeta, float* dbias,
                      float* scratch, int64_t rows, int relu, void* stream) {
  return as::ln256_bf16_bwd(dy_a, dy_b, z, res, gamma, beta, stats, dres, dz, mask_z, dgamma, dbeta, dbias, scratch, rows, relu,
                            ST(stream));
}
int as_masked_mean256_bf16_fwd(const void* x, const int64_t* len, float* out, int64_t R, int N, void* stream) {
  return as::masked_mean256_bf16_fwd(x, reinterpret_cast<const long long*>(len), out, R, N, ST(stream));
}
int as_masked_mean256_bf16_bwd(const float* dout, const int64_t* len, void* dx, int64_t R, int N, void* stream) {
  return as::masked_mean256_bf16_bwd(dout, reinterpret_cast<const long long*>(len), dx, R, N, ST(stream));
}
int as_embed_relu_fwd(const void* idx, int dtype, const float* table, float* out, int64_t R, int V, int D, void* stream) {
  return as::embed_relu_fwd(idx, dtype, table, out, R, V, D, ST(stream));
}
int as_one_hot_bf16(const void* idx, int dtype, void* out, int64_t R, int V, int Vpad, void* stream) {
  return as::one_hot_bf16(idx, dtype, out, R, V, Vpad, ST(stream));
}
int as_time_encode(const void* t, int dtype, const float* position_array, float* out, int64_t R, int D, void* stream) {
  return as::time_encode(t, dtype, position_array, out, R, D, ST(stream));
}
int as_bo_features(const void* order, int order_dtype, const void* location, int location_dtype, void* out, int64_t tokens, int L,
                   int A, int NB, int map_w, int Kpad, void* stream) {
  return as::bo_features(order, order_dtype, location, location_dtype, out, tokens, L, A, NB, map_w, Kpad, ST(stream));
}
int as_cast_any_bf16(const void* src, int dtype, void* dst, int64_t rows, int K, int Kpad, void* stream) {
  return as::cast_any_bf16(src, dtype, dst, rows, K, Kpad, ST(stream));
}
int as_expand_rows_f32(const float* rows, const int32_t* row_of, float* out, int64_t total, int K, float fill, void* stream) {
  return as::expand_rows(rows, row_of, out, total, K, fill, ST(stream));
}
int as_ce_rows_fwd(const float* logits, int64_t ld, const int64_t* labels, const int64_t* excl, const int64_t* excl_len, int S, int K,
                   float smoothing, float* loss, float* lse, int32_t* argmax, int64_t R, void* stream) {
  return as::ce_rows_fwd(logits, ld, (const long long*)labels, (const long long*)excl, (const long long*)excl_len, S, K, smoothing, loss, lse,
                         argmax, R, ST(stream));
}
int as_ce_rows_bwd(const float* logits, int64_t ld, const int64_t* labels, const int64_t* excl, const int64_t* excl_len, int S, int K,
                   float smoothing, const float* lse, const float* g, float* dlogits, int64_t R, void* stream) {
  return as::ce_rows_bwd(logits, ld, (const long long*)labels, (const long long*)excl, (const long long*)excl_len, S, K, smoothing, lse, g,
                         dlogits, R, ST(stream));
}
int as_pointer_keys_fwd(const float* key, const float* end_embedding, const int64_t* entity_num, float* out, int64_t R, int N, int D,
                        void* stream) {
  return as::pointer_keys_fwd(key, end_embedding, (const long long*)entity_num, out, R, N, D, ST(stream));
}
int as_pointer_keys_bwd(const float* dout, const int64_t* entity_num, float* dkey, float* dend, int64_t R, int N, int D, void* stream) {
  return as::pointer_keys_bwd(dout, (const long long*)entity_num, dkey, dend, R, N, D, ST(stream));
}
int as_attn_fwd_ok(int N, int D) { return as::attn_fwd_ok(N, D); }
int as_attn_fwd(const void* qkv, const int64_t* key_len, void* out, void* probs, int64_t R, int N, int H, int D, float scale,
                void* stream) {
  return as::attn_fwd(qkv, reinterpret_cast<const long long*>(key_len), out, probs, R, N, H, D, scale, ST(stream));
}
int as_attn_bwd_dq(const void* qkv, const void* out, const void* dout, const void* probs, void* dscores, void* dqkv, int64_t R, int N,
                   int H, int D, float scale, void* stream) {
  return as::attn_bwd_dq(qkv, out, dout, probs, dscores, dqkv, R, N, H, D, scale, ST(stream));
}
int as_attn_small_ok(int N, int D) { return as::attn_small_ok(N, D); }
int as_attn_small_fwd(const void* qkv, const int64_t* key_len, void* out, int64_t R, int N, int H, int D, float scale, void* stream) {
  return as::attn_small_fwd(qkv, reinterpret_cast<const long long*>(key_len), out, R, N, H, D, scale, ST(stream));
}
int as_attn_small_bwd(const void* qkv, const int64_t* key_len, const void* dout, void* dqkv, int64_t R, int N, int H, int D,
                      float scale, void* stream) {
  return as::attn_small_bwd(qkv, reinterpret_cast<const long long*>(key_len), dout, dqkv, R, N, H, D, scale, ST(stream));
}
int as_su_sample(const as_su_sample_args* a, void* stream) {
  if (!a) return AS_EINVAL;
  as::SuSampleDev p;
  p.key = a->key; p.c0 = a->c0; p.c1 = a->c1; p.mt = a->mt; p.wf1t = a->wf1t; p.bf1 = a->bf1; p.wq2t = a->wq2t; p.bq2 = a->bq2;
  p.wih_t = a->wih_t; p.whh_t = a->whh_t; p.ln_i_w = a->ln_i_w; p.ln_i_b = a->ln_i_b; p.ln_h_w = a->ln_h_w; p.ln_h_b = a->ln_h_b;
  p.ln_c_w = a->ln_c_w; p.ln_c_b = a->ln_c_b; p.entity_num = reinterpret_cast<const long long*>(a->entity_num);
  p.su_mask = a->su_mask; p.noise = a->noise; p.logits = a->logits; p.units = reinterpret_cast<long long*>(a->units);
  p.num = reinterpret_cast<long long*>(a->num); p.extra = a->extra; p.emb = a->emb; p.steps_row = a->steps_row; p.ended_rows = a->ended_rows;
  p.R = a->R; p.N1 = a->N1; p.S = a->S; p.extra_units = a->extra_units; p.temperature = a->temperature;
  return as::su_sample(p, ST(stream));
}
int as_masked_mean_fwd(const float* x, const int64_t* len, float* out, int64_t R, int N, int C, void* stream) {
  return as::masked_mean_fwd(x, reinterpret_cast<const long long*>(len), out, R, N, C, ST(stream));
}
int as_masked_mean_bwd(const float* dout, const int64_t* len, float* dx, int64_t R, int N, int C, void* stream) {
  return as::masked_mean_bwd(dout, reinterpret_cast<const long long*>(len), dx, R, N, C, ST(stream));
}
int as_linear_bwd_prep(const float* dy, const float* y32, const void* y16, void* dz16, float* colsum, float* scratch,
                       int64_t rows, int C, void* stream) {
  return as::linear_bwd_prep(dy, nullptr, y32, y16, dz16, colsum, scratch, rows, C, ST(stream));
}
int as_linear_bwd_prep2(const float* dy32, const void* dy16, const float* y32, const void* y16, void* dz16, float* colsum,
                        float* scratch, int64_t rows, int C, void* stream) {
  return as::linear_bwd_prep(dy32, dy16, y32, y16, dz16, colsum, scratch, rows, C, ST(stream));
}
int as_colsum_bf16_static(const void* x, float* out, int64_t rows, int C, void* stream) {
  return as::colsum_bf16_static(x, out, rows, C, ST(stream));
}

int as_layer_norm_fwd(const float* x, const float* res, const float* gamma, const float* beta, float* y32, void* y16,
                      float* stats, int64_t rows, int C, int relu, void* stream) {
  return as::layer_norm_fwd(x, res, gamma, beta, y32, y16, stats, rows, C, relu, ST(stream));
}

int64_t as_layer_norm_bwd_scratch_floats(int C) { return as::layer_norm_bwd_scratch_floats(C); }

int64_t as_layer_norm_bwd_fused_scratch_floats(int C) { return as::layer_norm_bwd_fused_scratch_floats(C); }
int as_layer_norm_bwd_fused(const float* dy32, const void* dy16, const float* x, const float* res, const float* gamma,
                            const float* beta, const float* stats, float* dx, void* dz16, int mask_x, float* dgamma, float* dbeta,
                            float* dbias, float* scratch, int64_t rows, int C, int relu, void* stream) {
  return as::layer_norm_bwd_fused(dy32, dy16, x, res, gamma, beta, stats, dx, dz16, mask_x, dgamma, dbeta, dbias, scratch, rows, C,
                                  relu, ST(stream));
}
int as_layer_norm_bwd(const float* dy32, const void* dy16, const float* x, const float* res, const float* gamma,
                      const float* beta, const float* stats, float* dx, float* dgamma, float* dbeta, float* scratch,
                      int64_t rows, int C, int relu, void* stream) {
  return as::layer_norm_bwd(dy32, dy16, x, res, gamma, beta, stats, dx, dgamma, dbeta, scratch, rows, C, relu,
                            ST(stream));
}

int as_lnlstm_fwd(const float* gi, const float* h0, const float* c0, const void* whh, const float* ln_h_w,
                  const float* ln_h_b, const float* ln_c_w, const float* ln_c_b, float* y, float* c_all, float* u_all,
                  float* z_all, float* stats, int T, int B, int H, void* stream) {
  return as::lnlstm_fwd(gi, h0, c0, whh, ln_h_w, ln_h_b, ln_c_w, ln_c_b, y, c_all, u_all, z_all, stats, T, B, H,
                        ST(stream));
}
int as_lnlstm_bwd(const float* gi, const float* h0, const float* c0, const float* y, const float* c_all,
                  const float* u_all, const float* z_all, const float* stats, const void* whh_t, const float* ln_h_w,
                  const float* ln_h_b, const float* ln_c_w, const float* ln_c_b, const float* dy, const float* dhT,
                  const float* dcT, float* dgi, float* du, float* dh0, float* dc0, float* dparam, int T, int B, int H,
                  void* stream) {
  return as::lnlstm_bwd(gi, h0, c0, y, c_all, u_all, z_all, stats, whh_t, ln_h_w, ln_h_b, ln_c_w, ln_c_b, dy, dhT, dcT,
                        dgi, du, dh0, dc0, dparam, T, B, H, ST(stream));
}
int as_lnlstm_cluster_ok(int H) { return as::lnlstm_cluster_ok(H) ? 1 : 0; }
// diagnostics only (not part of the public header)
int as_debug_gemm_profile(void* buf) { return as::gemm_set_profile(reinterpret_cast<long long*>(buf)); }
int as_debug_lstm_max_clusters(int H) { return as::lnlstm_cluster_max_active(H); }
int as_debug_attn_profile(void* buf) { return as::attn_set_profile(reinterpret_cast<long long*>(buf)); }
int as_debug_lstm_profile(void* buf) { return as::lnlstm_set_profile(reinterpret_cast<long long*>(buf)); }

int as_softmax_fwd(const void* scores, const int64_t* key_len, void* probs, int64_t groups, int Nq, int Nk,
                   int groups_per_len, int scores_bf16, void* stream) {
  return as::softmax_fwd(scores, reinterpret_cast<const long long*>(key_len), probs, groups, Nq, Nk, groups_per_len,
                         scores_bf16, ST(stream));
}

int as_softmax_bwd(const void* probs, const void* dprobs, void* dscores, int64_t rows, int Nk, int dprobs_bf16,
                   void* stream) {
  return as::softmax_bwd(probs, dprobs, dscores, rows, Nk, dprobs_bf16, ST(stream));
}

int as_entity_features(const void* const* data, const int* kind, const int* num, const int* col, const int* dtype,
                       int n_fields, void* out, int64_t rows, int pitch, int* neg_flag, void* stream) {
  if (!data || !kind || !num || !col || !dtype) return AS_EINVAL;
  return as::entity_features(data, kind, num, col, dtype, n_fields, out, rows, pitch, neg_flag, ST(stream));
}

int as_pack_f32(const int64_t* table, int chunks, float* dst, void* stream) {
  return as::pack_f32(reinterpret_cast<const long long*>(table), chunks, dst, ST(stream));
}
int as_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream) {
  return as::cast_f32_bf16(src, dst, n, ST(stream));
}

int as_relu_bwd_f32(const float* dy, const float* y, float* dz, int64_t n, void* stream) {
  return as::relu_bwd_f32(dy, y, dz, n, ST(stream));
}

int as_gumbel_argmax(const float* logit, const float* noise, int64_t* out, int64_t rows, int K, void* stream) {
  return as::gumbel_argmax(logit, noise, reinterpret_cast<long long*>(out), rows, K, ST(stream));
}

int as_rl_loss_forward(const AsLossArgs* args, void* stream) {
  if (!args) return AS_EINVAL;
  return as::loss_forward(*args, ST(stream));
}

int as_rl_loss_backward(const AsLossArgs* args, float grad_scale, void* stream) {
  if (!args) return AS_EINVAL;
  return as::loss_backward(*args, grad_scale, ST(stream));
}

int64_t as_clip_adam_scratch_bytes(void) { return as::clip_adam_scratch_bytes(); }

int as_clip_adam(float* p, const float* g, float* m, float* v, void* shadow, int64_t n, float max_norm, float lr,
                 float beta1, float beta2, float eps, float weight_decay, int step, int* step_dev, float* scratch,
                 float* norm_out, void* stream) {
  return as::clip_adam(p, g, m, v, shadow, n, max_norm, lr, beta1, beta2, eps, weight_decay, step, step_dev, scratch,
                       norm_out, ST(stream));
}

int as_clip_adam_per_param(float* p, const float* g, float* m, float* v, void* shadow, int64_t n, float max_norm, float lr,
                           float beta1, float beta2, float eps, float weight_decay, const uint16_t* block_param,
                           const int32_t* param_active, int32_t* param_step, int n_params, int* step_dev, float* scratch,
                           float* norm_out, void* stream) {
  return as::clip_adam_per_param(p, g, m, v, shadow, n, max_norm, lr, beta1, beta2, eps, weight_decay, block_param, param_active,
                                 param_step, n_params, step_dev, scratch, norm_out, ST(stream));
}

}  // extern "C"
