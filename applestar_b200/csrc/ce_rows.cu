// Row-wise cross entropy of the supervised learner (reference: agent/default/sl_training/sl_loss.py:16-36 label smoothing,
// :127-285 the six heads): one pass over the logits for log-sum-exp / arg-max / label term (forward), one pass that writes
// the gradient (backward) -- 3 x 4 bytes per logit in total, against log_softmax + gather + argmax + their autograd
// (about 7 passes) in the reference.  The SelectedUnits head's label mask (sl_loss.py:190-204: every unit the sample
// selects is removed from the candidates of the OTHER steps) is a bit set per sample built in shared memory, so the masked
// copy of the [b, s, n] logits is never written.
#include "common.cuh"

namespace as {

constexpr int CE_THREADS = 256;
constexpr int CE_MAX_WORDS = 1024;          // label bit set: K <= 32768

struct CeAcc { float m, l, sx, best; int arg; };

__device__ __forceinline__ void ce_merge(CeAcc& a, const CeAcc& b) {
  const float M = fmaxf(a.m, b.m);
  a.l = a.l * __expf(a.m - M) + b.l * __expf(b.m - M);
  a.m = M;
  a.sx += b.sx;
  if (b.best > a.best || (b.best == a.best && b.arg < a.arg)) { a.best = b.best; a.arg = b.arg; }
}

// excluded(k) for row (b, s): k is one of labels[b, j], j < excl_len[b], and not this step's own label
__device__ __forceinline__ void ce_build_mask(unsigned* bits, int words, const long long* excl, long long len, int S, int s, int K) {
  for (int w = threadIdx.x; w < words; w += blockDim.x) bits[w] = 0u;
  __syncthreads();
  for (int j = threadIdx.x; j < len && j < S; j += blockDim.x) {
    const long long v = excl[j];
    if (v >= 0 && v < K) atomicOr(bits + (v >> 5), 1u << (v & 31));
  }
  __syncthreads();
  if (threadIdx.x == 0 && s < len) {
    const long long v = excl[s];
    if (v >= 0 && v < K) bits[v >> 5] &= ~(1u << (v & 31));
  }
  __syncthreads();
}

__global__ void __launch_bounds__(CE_THREADS) ce_rows_fwd_kernel(const float* __restrict__ logits, long long ld,
    const long long* __restrict__ labels, const long long* __restrict__ excl, const long long* __restrict__ excl_len, int S, int K,
    float smoothing, float* __restrict__ loss, float* __restrict__ lse_out, int* __restrict__ argmax_out, long long R) {
  __shared__ unsigned bits[CE_MAX_WORDS];
  __shared__ CeAcc part[CE_THREADS / 32];
  const int words = (K + 31) >> 5;
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    const long long b = r / S;
    const int s = (int)(r - b * S);
    if (excl) ce_build_mask(bits, words, excl + b * S, excl_len[b], S, s, K);
    const float* x = logits + r * ld;
    CeAcc a{-INFINITY, 0.f, 0.f, -INFINITY, 0x7fffffff};
    for (int k = threadIdx.x; k < K; k += CE_THREADS) {
      float v = __ldg(x + k);
      if (excl && (bits[k >> 5] >> (k & 31) & 1u)) v = -1e9f;
      if (v > a.best) { a.best = v; a.arg = k; }
      a.sx += v;
      if (v > a.m) { a.l = a.l * __expf(a.m - v) + 1.f; a.m = v; } else { a.l += __expf(v - a.m); }
    }
    for (int o = 16; o; o >>= 1) {
      CeAcc t{__shfl_xor_sync(~0u, a.m, o), __shfl_xor_sync(~0u, a.l, o), __shfl_xor_sync(~0u, a.sx, o),
              __shfl_xor_sync(~0u, a.best, o), __shfl_xor_sync(~0u, a.arg, o)};
      if (t.l > 0.f) { if (a.l > 0.f) ce_merge(a, t); else a = t; }
    }
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
      CeAcc t = part[0];
      for (int w = 1; w < CE_THREADS / 32; ++w) if (part[w].l > 0.f) { if (t.l > 0.f) ce_merge(t, part[w]); else t = part[w]; }
      const float lse = t.m + logf(t.l);
      long long y = labels[r];
      y = y < 0 ? 0 : (y >= K ? K - 1 : y);
      float xy = __ldg(x + y);
      if (excl && (bits[y >> 5] >> (y & 31) & 1u)) xy = -1e9f;
      const float nll = lse - xy;
      loss[r] = smoothing > 0.f ? (1.f - smoothing) * nll + smoothing * (lse - t.sx / (float)K) : nll;
      lse_out[r] = lse;
      argmax_out[r] = t.arg;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(CE_THREADS) ce_rows_bwd_kernel(const float* __restrict__ logits, long long ld,
    const long long* __restrict__ labels, const long long* __restrict__ excl, const long long* __restrict__ excl_len, int S, int K,
    float smoothing, const float* __restrict__ lse, const float* __restrict__ g, float* __restrict__ dlogits, long long R) {
  __shared__ unsigned bits[CE_MAX_WORDS];
  const int words = (K + 31) >> 5;
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    const float gr = g[r];
    float* d = dlogits + r * ld;
    if (gr == 0.f) {
      for (int k = threadIdx.x; k < K; k += CE_THREADS) d[k] = 0.f;
      continue;
    }
    const long long b = r / S;
    const int s = (int)(r - b * S);
    if (excl) ce_build_mask(bits, words, excl + b * S, excl_len[b], S, s, K);
    const float* x = logits + r * ld;
    const float L = lse[r];
    long long y = labels[r];
    y = y < 0 ? 0 : (y >= K ? K - 1 : y);
    const float uni = smoothing / (float)K;
    for (int k = threadIdx.x; k < K; k += CE_THREADS) {
      const bool masked = excl && (bits[k >> 5] >> (k & 31) & 1u);
      float v = masked ? 0.f : __expf(__ldg(x + k) - L) - uni;
      if (k == y && !masked) v -= 1.f - smoothing;
      d[k] = gr * v;
    }
    if (excl) __syncthreads();
  }
}

int ce_rows_fwd(const float* logits, long long ld, const long long* labels, const long long* excl, const long long* excl_len, int S, int K,
                float smoothing, float* loss, float* lse, int* argmax, long long R, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!logits || !labels || !loss || !lse || !argmax || K <= 0 || K > CE_MAX_WORDS * 32 || ld < K || S <= 0 || (excl && !excl_len))
    return AS_EINVAL;
  ce_rows_fwd_kernel<<<(unsigned)(R < 148 * 8 ? R : 148 * 8), CE_THREADS, 0, st>>>(logits, ld, labels, excl, excl_len, S, K, smoothing, loss,
                                                                                lse, argmax, R);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int ce_rows_bwd(const float* logits, long long ld, const long long* labels, const long long* excl, const long long* excl_len, int S, int K,
                float smoothing, const float* lse, const float* g, float* dlogits, long long R, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!logits || !labels || !lse || !g || !dlogits || K <= 0 || K > CE_MAX_WORDS * 32 || ld < K || S <= 0 || (excl && !excl_len))
    return AS_EINVAL;
  ce_rows_bwd_kernel<<<(unsigned)(R < 148 * 8 ? R : 148 * 8), CE_THREADS, 0, st>>>(logits, ld, labels, excl, excl_len, S, K, smoothing, lse, g,
                                                                                dlogits, R);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
