// Teacher-forced SelectedUnits pointer loop without the loop (reference: head/action_arg_head.py:168-216).
//
// In the learner the picks of every step are given (teacher forcing), so everything that the reference recomputes
// step by step from one-hot tensors is a function of `selected_units` alone:
//   * the autoregressive embedding of step i is the mean of the key rows of the distinct units picked up to step i
//     (before the end token)                                   -> su_prefix_fwd / bwd: one warp per row, running sum;
//   * the logits mask of step i removes padding, the end token at step 0 and every unit picked before step i
//                                                              -> su_mask_logits: first-pick step per entity in
//                                                                 shared memory, then one pass over [S, N+1].
// With these two kernels the S steps of query MLP, LayerNorm-LSTM, key dot products and embedding MLP run as a few
// batched launches over all steps instead of ~65 small launches per step.
#include "common.cuh"

namespace as {

// emb[r,i,:] = (sum over distinct live picks j <= i of key[r, sel[r,j], :]) / (has_units ? count : 1)
// cnt[r,i] = count, isnew[r,i] = 1 when step i added a unit.  D must be 32 (one lane per channel).
__global__ void __launch_bounds__(256)
su_prefix_fwd_kernel(const float* __restrict__ key, const long long* __restrict__ sel, const long long* __restrict__ entity_num,
                     const long long* __restrict__ su_num, float* __restrict__ emb, float* __restrict__ cnt,
                     unsigned char* __restrict__ isnew, long long R, int N1, int Smax, int S) {
  const int lane = threadIdx.x & 31;
  const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= R) return;
  const long long en = entity_num[r];
  const bool has_units = su_num[r] != 0;
  const long long s_lo = lane < S ? sel[r * Smax + lane] : -1, s_hi = lane + 32 < S ? sel[r * Smax + lane + 32] : -1;
  unsigned long long newmask = 0ull;
  float acc = 0.f, count = 0.f;
  bool ended = false;
  for (int i = 0; i < S; ++i) {
    const long long s = __shfl_sync(0xffffffffu, i < 32 ? s_lo : s_hi, i & 31);
    ended = ended || (s == en);
    bool add = !ended;
    if (add) {   // a unit already in the set is not counted twice (the reference takes the maximum of one-hots)
      const bool dup = (((newmask >> lane) & 1ull) && s_lo == s) || (((newmask >> (lane + 32)) & 1ull) && s_hi == s);
      if (__any_sync(0xffffffffu, dup)) add = false;
    }
    if (add) {
      const long long sc = s < 0 ? 0 : (s > N1 - 1 ? N1 - 1 : s);
      acc += __ldg(key + (r * N1 + sc) * 32 + lane);
      count += 1.f;
      newmask |= 1ull << i;
    }
    emb[(r * S + i) * 32 + lane] = has_units ? acc / count : acc;
    if (lane == 0) { cnt[r * S + i] = count; isnew[r * S + i] = add ? 1 : 0; }
  }
}

// dkey[r, sel[r,j], :] += sum_{i >= j} demb[r,i,:] / count_i for every step j that added a unit (dkey zeroed by the caller)
__global__ void __launch_bounds__(256)
su_prefix_bwd_kernel(const float* __restrict__ demb, const long long* __restrict__ sel, const long long* __restrict__ su_num,
                     const float* __restrict__ cnt, const unsigned char* __restrict__ isnew, float* __restrict__ dkey,
                     long long R, int N1, int Smax, int S) {
  const int lane = threadIdx.x & 31;
  const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= R) return;
  const bool has_units = su_num[r] != 0;
  float suffix = 0.f;
  for (int i = S - 1; i >= 0; --i) {
    float g = demb[(r * S + i) * 32 + lane];
    if (has_units) {
      const float c = cnt[r * S + i];
      g = c > 0.f ? g / c : 0.f;
    }
    suffix += g;
    if (isnew[r * S + i]) {
      long long s = sel[r * Smax + i];
      s = s < 0 ? 0 : (s > N1 - 1 ? N1 - 1 : s);
      dkey[(r * N1 + s) * 32 + lane] += suffix;
    }
  }
}

// out[r,i,n] = valid(r,i,n) ? x[r,i,n] * scale : fill, valid = n <= entity_num[r] and not (i == 0 and n == entity_num[r])
// and n not picked at any step before i (action_arg_head.py:179-187).  fill = -1e9 * scale forward, 0 backward.
__global__ void __launch_bounds__(256)
su_mask_logits_kernel(const float* __restrict__ x, const long long* __restrict__ sel, const long long* __restrict__ entity_num,
                      float* __restrict__ out, int N1, int Smax, int S, float scale, float fill) {
  extern __shared__ int first_sel[];                 // [N1] first step at which entity n was picked (S = never)
  const long long r = blockIdx.x;
  for (int n = threadIdx.x; n < N1; n += blockDim.x) first_sel[n] = S;
  __syncthreads();
  for (int j = threadIdx.x; j < S; j += blockDim.x) {
    const long long s = sel[r * Smax + j];
    if (s >= 0 && s < N1) atomicMin(&first_sel[(int)s], j);
  }
  __syncthreads();
  const long long en = entity_num[r];
  const int total = S * N1;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int i = idx / N1, n = idx - i * N1;
    const bool valid = (long long)n <= en && !(i == 0 && (long long)n == en) && first_sel[n] >= i;
    const long long o = r * total + idx;
    out[o] = valid ? x[o] * scale : fill;
  }
}

// Pointer logits (action_arg_head.py:343-363 target unit, :190-196 selected units): one CTA per row r keeps the row's
// keys [N, 32] in shared memory; thread n forms <query[r,s,:], key[r,n,:]> for every query step s (queries broadcast
// from shared memory), applies the key-length mask and the 1/temperature scale, and stores coalesced along n.
//   out[r,s,n] = n < len[r] ? scale * sum_k q[r,s,k] key[r,n,k] : fill          (len == nullptr: no mask)
constexpr int kPtrThreads = 256;
__global__ void __launch_bounds__(kPtrThreads)
pointer_logits_fwd_kernel(const float* __restrict__ query, const float* __restrict__ key, const long long* __restrict__ len,
                          float* __restrict__ out, int S, int N, float scale, float fill) {
  extern __shared__ float psm[];
  float* s_key = psm;                     // [N][33] (padded: thread n reads row n, conflict-free)
  float* s_q = psm + (size_t)N * 33;      // [S][32]
  const long long r = blockIdx.x;
  const float4* k4 = reinterpret_cast<const float4*>(key + r * (long long)N * 32);
  for (int i = threadIdx.x; i < N * 8; i += kPtrThreads) {
    const float4 v = __ldg(k4 + i);
    float* d = s_key + (i >> 3) * 33 + (i & 7) * 4;
    d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
  }
  for (int i = threadIdx.x; i < S * 32; i += kPtrThreads) s_q[i] = __ldg(query + r * (long long)S * 32 + i);
  __syncthreads();
  const long long L = len ? len[r] : (long long)N;
  for (int n = threadIdx.x; n < N; n += kPtrThreads) {
    float kr[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) kr[k] = s_key[n * 33 + k];
    const bool valid = (long long)n < L;
    for (int s = 0; s < S; ++s) {
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < 32; ++k) acc = fmaf(s_q[s * 32 + k], kr[k], acc);
      out[(r * S + s) * (long long)N + n] = valid ? acc * scale : fill;
    }
  }
}

// dquery[r,s,k] = scale * sum_{n < len} dout[r,s,n] key[r,n,k];   dkey[r,n,k] = scale * sum_s dout[r,s,n] query[r,s,k]
// (n >= len: zero).  Phase 1: thread n accumulates its dkey row over s; phase 2: warp w owns the steps s = w, w + 8, ...,
// lane = k, and walks n (keys from shared memory, dout broadcast).  Fixed summation order: deterministic.
__global__ void __launch_bounds__(kPtrThreads)
pointer_logits_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ query, const float* __restrict__ key,
                          const long long* __restrict__ len, float* __restrict__ dquery, float* __restrict__ dkey, int S, int N,
                          float scale) {
  extern __shared__ float psm[];
  float* s_key = psm;
  float* s_q = psm + (size_t)N * 33;
  const long long r = blockIdx.x;
  const float4* k4 = reinterpret_cast<const float4*>(key + r * (long long)N * 32);
  for (int i = threadIdx.x; i < N * 8; i += kPtrThreads) {
    const float4 v = __ldg(k4 + i);
    float* d = s_key + (i >> 3) * 33 + (i & 7) * 4;
    d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
  }
  for (int i = threadIdx.x; i < S * 32; i += kPtrThreads) s_q[i] = __ldg(query + r * (long long)S * 32 + i);
  __syncthreads();
  const long long Lr = len ? len[r] : (long long)N;
  const int L = (int)(Lr < 0 ? 0 : (Lr > N ? N : Lr));
  const float* dro = dout + r * (long long)S * N;
  if (dkey) {
    for (int n = threadIdx.x; n < N; n += kPtrThreads) {
      float acc[32];
#pragma unroll
      for (int k = 0; k < 32; ++k) acc[k] = 0.f;
      if (n < L) {
        for (int s = 0; s < S; ++s) {
          const float g = __ldg(dro + (long long)s * N + n) * scale;
#pragma unroll
          for (int k = 0; k < 32; ++k) acc[k] = fmaf(g, s_q[s * 32 + k], acc[k]);
        }
      }
      float4* d4 = reinterpret_cast<float4*>(dkey + (r * (long long)N + n) * 32);
#pragma unroll
      for (int q = 0; q < 8; ++q) d4[q] = make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
    }
  }
  if (dquery) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int s = warp; s < S; s += kPtrThreads / 32) {
      float acc = 0.f;
      for (int n0 = 0; n0 < L; n0 += 32) {
        const float gl = (n0 + lane < L) ? __ldg(dro + (long long)s * N + n0 + lane) : 0.f;   // coalesced, then broadcast
        const int m = min(32, L - n0);
        for (int j = 0; j < m; ++j) acc = fmaf(__shfl_sync(0xffffffffu, gl, j), s_key[(n0 + j) * 33 + lane], acc);
      }
      dquery[(r * S + s) * 32 + lane] = acc * scale;
    }
  }
}

int pointer_logits_fwd(const float* query, const float* key, const long long* len, float* out, long long R, int S, int N,
                       int D, float scale, float fill, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (D != 32 || S <= 0 || S > 64 || N <= 0 || N > 1024) return AS_ESIZE;
  if (!query || !key || !out) return AS_EINVAL;
  const size_t smem = ((size_t)N * 33 + (size_t)S * 32) * sizeof(float);
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(pointer_logits_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    cudaFuncSetAttribute(pointer_logits_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    attr = true;
  }
  pointer_logits_fwd_kernel<<<(unsigned)R, kPtrThreads, smem, st>>>(query, key, len, out, S, N, scale, fill);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int pointer_logits_bwd(const float* dout, const float* query, const float* key, const long long* len, float* dquery, float* dkey,
                       long long R, int S, int N, int D, float scale, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (D != 32 || S <= 0 || S > 64 || N <= 0 || N > 1024) return AS_ESIZE;
  if (!dout || !query || !key || (!dquery && !dkey)) return AS_EINVAL;
  const size_t smem = ((size_t)N * 33 + (size_t)S * 32) * sizeof(float);
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(pointer_logits_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    cudaFuncSetAttribute(pointer_logits_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    attr = true;
  }
  pointer_logits_bwd_kernel<<<(unsigned)R, kPtrThreads, smem, st>>>(dout, query, key, len, dquery, dkey, S, N, scale);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int su_prefix_fwd(const float* key, const long long* sel, const long long* entity_num, const long long* su_num, float* emb,
                  float* cnt, unsigned char* isnew, long long R, int N1, int D, int Smax, int S, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (D != 32 || S <= 0 || S > 64 || S > Smax) return AS_ESIZE;
  if (!key || !sel || !entity_num || !su_num || !emb || !cnt || !isnew) return AS_EINVAL;
  su_prefix_fwd_kernel<<<(unsigned)((R * 32 + 255) / 256), 256, 0, st>>>(key, sel, entity_num, su_num, emb, cnt, isnew, R, N1, Smax, S);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int su_prefix_bwd(const float* demb, const long long* sel, const long long* su_num, const float* cnt,
                  const unsigned char* isnew, float* dkey, long long R, int N1, int D, int Smax, int S, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (D != 32 || S <= 0 || S > 64 || S > Smax) return AS_ESIZE;
  if (!demb || !sel || !su_num || !cnt || !isnew || !dkey) return AS_EINVAL;
  su_prefix_bwd_kernel<<<(unsigned)((R * 32 + 255) / 256), 256, 0, st>>>(demb, sel, su_num, cnt, isnew, dkey, R, N1, Smax, S);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int su_mask_logits(const float* x, const long long* sel, const long long* entity_num, float* out, long long R, int N1,
                   int Smax, int S, float scale, float fill, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (S <= 0 || S > Smax || N1 <= 0 || N1 > 8192) return AS_ESIZE;
  if (!x || !sel || !entity_num || !out) return AS_EINVAL;
  su_mask_logits_kernel<<<(unsigned)R, 256, N1 * sizeof(int), st>>>(x, sel, entity_num, out, N1, Smax, S, scale, fill);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// Pointer keys of the selected-units head (action_arg_head.py:118-129): one row appended to the entity keys, the end-flag
// embedding placed at row entity_num.  out [R, N + 1, D] fp32; key [R, N, D]; D a multiple of 4.
__global__ void pointer_keys_fwd_kernel(const float* __restrict__ key, const float* __restrict__ end_emb, const long long* __restrict__ entity_num,
                                        float* __restrict__ out, long long R, int N, int D4) {
  const long long total = R * (N + 1) * D4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(i % D4);
    const long long rn = i / D4;
    const int n = (int)(rn % (N + 1));
    const long long r = rn / (N + 1);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (n == entity_num[r]) v = __ldg(reinterpret_cast<const float4*>(end_emb) + d);
    else if (n < N) v = __ldg(reinterpret_cast<const float4*>(key) + (r * N + n) * D4 + d);
    reinterpret_cast<float4*>(out)[i] = v;
  }
}
// dkey[r, n, :] = dout[r, n, :] except at the end-flag row (zero); dend[d] = sum_r dout[r, entity_num[r], d] in row order
__global__ void pointer_keys_bwd_kernel(const float* __restrict__ dout, const long long* __restrict__ entity_num, float* __restrict__ dkey,
                                        long long R, int N, int D4) {
  const long long total = R * N * D4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(i % D4);
    const long long rn = i / D4;
    const int n = (int)(rn % N);
    const long long r = rn / N;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (n != entity_num[r]) v = __ldg(reinterpret_cast<const float4*>(dout) + (r * (N + 1) + n) * D4 + d);
    reinterpret_cast<float4*>(dkey)[i] = v;
  }
}
__global__ void pointer_keys_dend_kernel(const float* __restrict__ dout, const long long* __restrict__ entity_num, float* __restrict__ dend,
                                         long long R, int N, int D) {
  // one block; thread = (row lane, d): rows strided over the lanes, then a fixed-order sum over the lanes
  extern __shared__ float part[];                       // [lanes][D]
  const int lanes = blockDim.x / D;
  const int d = threadIdx.x % D, l = threadIdx.x / D;
  float acc = 0.f;
  if (l < lanes)
    for (long long r = l; r < R; r += lanes) {
      const long long e = entity_num[r];
      if (e >= 0 && e <= N) acc += dout[(r * (N + 1) + e) * D + d];
    }
  if (l < lanes) part[l * D + d] = acc;
  __syncthreads();
  if (threadIdx.x < D) {
    float s = 0.f;
    for (int k = 0; k < lanes; ++k) s += part[k * D + threadIdx.x];
    dend[threadIdx.x] = s;
  }
}
int pointer_keys_fwd(const float* key, const float* end_emb, const long long* entity_num, float* out, long long R, int N, int D,
                     cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (N <= 0 || D <= 0 || (D % 4) != 0 || D > 256) return AS_ESIZE;
  if (!key || !end_emb || !entity_num || !out) return AS_EINVAL;
  const long long total = R * (N + 1) * (D / 4);
  long long blocks = (total + 255) / 256;
  pointer_keys_fwd_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, st>>>(key, end_emb, entity_num, out, R, N, D / 4);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int pointer_keys_bwd(const float* dout, const long long* entity_num, float* dkey, float* dend, long long R, int N, int D, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (N <= 0 || D <= 0 || (D % 4) != 0 || D > 256) return AS_ESIZE;
  if (!dout || !entity_num || (!dkey && !dend)) return AS_EINVAL;
  if (dkey) {
    const long long total = R * N * (D / 4);
    long long blocks = (total + 255) / 256;
    pointer_keys_bwd_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, st>>>(dout, entity_num, dkey, R, N, D / 4);
    AS_LAUNCH_CHECK();
  }
  if (dend) {
    const int threads = (1024 / D) * D;
    pointer_keys_dend_kernel<<<1, threads, (size_t)threads * sizeof(float), st>>>(dout, entity_num, dend, R, N, D);
    AS_LAUNCH_CHECK();
  }
  return AS_OK;
}

}  // namespace as
