// Attention for short sequences and narrow heads -- the build-order Transformer of the scalar encoder
// (obs_encoder/scalar_encoder.py:19-53: 20 tokens, 2 heads x 8, module_utils.py:71-111).  The tensor-core path spends a
// 128-row tile, a TMEM allocation and five launches per direction on 20 x 20 x 8 products; here ONE WARP owns a
// (row, head) pair, lane = query (forward) / key (second half of the backward), everything in fp32 registers and a few
// KB of shared memory, probabilities recomputed in the backward pass instead of being stored.
#include <cuda_bf16.h>

#include "common.cuh"

namespace as {

constexpr int kAsWarps = 2;      // (row, head) pairs per block
constexpr int kAsMaxN = 32;

template <int D>
struct AsSmem {
  float q[kAsMaxN][D + 1], k[kAsMaxN][D + 1], v[kAsMaxN][D + 1], g[kAsMaxN][D + 1];
  float p[kAsMaxN][kAsMaxN + 1], ds[kAsMaxN][kAsMaxN + 1];
};

template <int D>
__device__ __forceinline__ void as_load_rows(float (*dst)[D + 1], const __nv_bfloat16* src, long long ld, int N, int lane) {
  for (int idx = lane; idx < N * D; idx += 32) {
    const int row = idx / D, d = idx - row * D;
    dst[row][d] = __bfloat162float(src[(long long)row * ld + d]);
  }
}

// scores of query `lane` against every key, softmax in registers; returns 1 / sum, p[] unnormalised exp
template <int D>
__device__ __forceinline__ float as_probs(const AsSmem<D>& s, int lane, int N, int kl, float scale, float (&p)[kAsMaxN]) {
  float qi[D];
#pragma unroll
  for (int d = 0; d < D; ++d) qi[d] = s.q[lane][d];
  float m = -INFINITY;
#pragma unroll
  for (int j = 0; j < kAsMaxN; ++j) {
    float a = 0.f;
    if (j < N) {
#pragma unroll
      for (int d = 0; d < D; ++d) a = fmaf(qi[d], s.k[j][d], a);
      a = j < kl ? a * scale : -1e9f;
      m = fmaxf(m, a);
    }
    p[j] = a;
  }
  float l = 0.f;
#pragma unroll
  for (int j = 0; j < kAsMaxN; ++j) {
    p[j] = j < N ? __expf(p[j] - m) : 0.f;
    l += p[j];
  }
  return 1.f / l;
}

template <int D>
__global__ void __launch_bounds__(kAsWarps * 32)
attn_small_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, const long long* __restrict__ key_len,
                      __nv_bfloat16* __restrict__ out, long long pairs, int N, int H, float scale) {
  __shared__ AsSmem<D> sm[kAsWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long pair = (long long)blockIdx.x * kAsWarps + warp;
  if (pair >= pairs) return;
  AsSmem<D>& s = sm[warp];
  const long long r = pair / H;
  const int h = (int)(pair - r * H);
  const long long ld = 3LL * H * D;
  const __nv_bfloat16* base = qkv + r * N * ld + h * D;
  as_load_rows<D>(s.q, base, ld, N, lane);
  as_load_rows<D>(s.k, base + H * D, ld, N, lane);
  as_load_rows<D>(s.v, base + 2 * H * D, ld, N, lane);
  __syncwarp();
  if (lane >= N) return;
  int kl = key_len ? (int)key_len[r] : N;
  if (kl > N) kl = N;
  float p[kAsMaxN];
  const float rinv = as_probs<D>(s, lane, N, kl, scale, p);
  float o[D];
#pragma unroll
  for (int d = 0; d < D; ++d) o[d] = 0.f;
#pragma unroll
  for (int j = 0; j < kAsMaxN; ++j) {
    if (j < N) {
#pragma unroll
      for (int d = 0; d < D; ++d) o[d] = fmaf(p[j], s.v[j][d], o[d]);
    }
  }
  __nv_bfloat16* dst = out + (r * N + lane) * (long long)(H * D) + h * D;
#pragma unroll
  for (int d = 0; d < D; ++d) dst[d] = __float2bfloat16(o[d] * rinv);
}

template <int D>
__global__ void __launch_bounds__(kAsWarps * 32)
attn_small_bwd_kernel(const __nv_bfloat16* __restrict__ qkv, const long long* __restrict__ key_len,
                      const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ dqkv, long long pairs, int N, int H,
                      float scale) {
  __shared__ AsSmem<D> sm[kAsWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long pair = (long long)blockIdx.x * kAsWarps + warp;
  if (pair >= pairs) return;
  AsSmem<D>& s = sm[warp];
  const long long r = pair / H;
  const int h = (int)(pair - r * H);
  const long long ld = 3LL * H * D;
  const __nv_bfloat16* base = qkv + r * N * ld + h * D;
  as_load_rows<D>(s.q, base, ld, N, lane);
  as_load_rows<D>(s.k, base + H * D, ld, N, lane);
  as_load_rows<D>(s.v, base + 2 * H * D, ld, N, lane);
  as_load_rows<D>(s.g, dout + r * N * (long long)(H * D) + h * D, (long long)H * D, N, lane);
  __syncwarp();
  int kl = key_len ? (int)key_len[r] : N;
  if (kl > N) kl = N;
  __nv_bfloat16* dbase = dqkv + r * N * ld + h * D;
  if (lane < N) {                                              // lane = query i: dS row, dQ_i
    float p[kAsMaxN];
    const float rinv = as_probs<D>(s, lane, N, kl, scale, p);
    float gi[D];
#pragma unroll
    for (int d = 0; d < D; ++d) gi[d] = s.g[lane][d];
    float dp[kAsMaxN];
    float delta = 0.f;
#pragma unroll
    for (int j = 0; j < kAsMaxN; ++j) {
      float a = 0.f;
      if (j < N) {
#pragma unroll
        for (int d = 0; d < D; ++d) a = fmaf(gi[d], s.v[j][d], a);
      }
      p[j] *= rinv;
      dp[j] = a;
      delta = fmaf(p[j], a, delta);
    }
    float dq[D];
#pragma unroll
    for (int d = 0; d < D; ++d) dq[d] = 0.f;
#pragma unroll
    for (int j = 0; j < kAsMaxN; ++j) {
      if (j < N) {
        const float dsj = p[j] * (dp[j] - delta);
        s.p[lane][j] = p[j];
        s.ds[lane][j] = dsj;
#pragma unroll
        for (int d = 0; d < D; ++d) dq[d] = fmaf(dsj, s.k[j][d], dq[d]);
      }
    }
    __nv_bfloat16* dst = dbase + (long long)lane * ld;
#pragma unroll
    for (int d = 0; d < D; ++d) dst[d] = __float2bfloat16(dq[d] * scale);
  }
  __syncwarp();
  if (lane < N) {                                              // lane = key j: dK_j, dV_j summed over the queries
    float dk[D], dv[D];
#pragma unroll
    for (int d = 0; d < D; ++d) { dk[d] = 0.f; dv[d] = 0.f; }
    for (int i = 0; i < N; ++i) {
      const float dsi = s.ds[i][lane], pi = s.p[i][lane];
#pragma unroll
      for (int d = 0; d < D; ++d) {
        dk[d] = fmaf(dsi, s.q[i][d], dk[d]);
        dv[d] = fmaf(pi, s.g[i][d], dv[d]);
      }
    }
    __nv_bfloat16* dst = dbase + (long long)lane * ld;
#pragma unroll
    for (int d = 0; d < D; ++d) {
      dst[H * D + d] = __float2bfloat16(dk[d] * scale);
      dst[2 * H * D + d] = __float2bfloat16(dv[d]);
    }
  }
}

int attn_small_ok(int N, int D) { return N >= 1 && N <= kAsMaxN && (D == 8 || D == 16); }

int attn_small_fwd(const void* qkv, const long long* key_len, void* out, long long R, int N, int H, int D, float scale,
                   cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!attn_small_ok(N, D) || H <= 0) return AS_EUNSUPPORTED;
  if (!qkv || !out) return AS_EINVAL;
  const long long pairs = R * H;
  const unsigned grid = (unsigned)((pairs + kAsWarps - 1) / kAsWarps);
  auto q = reinterpret_cast<const __nv_bfloat16*>(qkv);
  auto o = reinterpret_cast<__nv_bfloat16*>(out);
  if (D == 8) attn_small_fwd_kernel<8><<<grid, kAsWarps * 32, 0, st>>>(q, key_len, o, pairs, N, H, scale);
  else attn_small_fwd_kernel<16><<<grid, kAsWarps * 32, 0, st>>>(q, key_len, o, pairs, N, H, scale);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int attn_small_bwd(const void* qkv, const long long* key_len, const void* dout, void* dqkv, long long R, int N, int H, int D,
                   float scale, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!attn_small_ok(N, D) || H <= 0) return AS_EUNSUPPORTED;
  if (!qkv || !dout || !dqkv) return AS_EINVAL;
  const long long pairs = R * H;
  const unsigned grid = (unsigned)((pairs + kAsWarps - 1) / kAsWarps);
  auto q = reinterpret_cast<const __nv_bfloat16*>(qkv);
  auto g = reinterpret_cast<const __nv_bfloat16*>(dout);
  auto dq = reinterpret_cast<__nv_bfloat16*>(dqkv);
  if (D == 8) attn_small_bwd_kernel<8><<<grid, kAsWarps * 32, 0, st>>>(q, key_len, g, dq, pairs, N, H, scale);
  else attn_small_bwd_kernel<16><<<grid, kAsWarps * 32, 0, st>>>(q, key_len, g, dq, pairs, N, H, scale);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
