// Scalar encoder front (reference: obs_encoder/scalar_encoder.py:91-132 and :41-53): the pieces that are index work or
// tiny element-wise math -- ReLU(Embedding(clamp(idx))) of the five one-hot fields, the sinusoidal time code, the bf16
// operand builders (one-hot rows, build-order token features [one_hot(action,174) | eye(20)[position] | bits10(x) |
// bits10(y)], integer -> bf16 casts of the fc fields) whose products and weight gradients then run on the tcgen05 GEMM.
#include "common.cuh"

namespace as {

// dtype codes shared with the entity feature kernel: 0 uint8, 1 int8, 2 int16, 3 float16, 4 int64, 5 float32
__device__ __forceinline__ long long ld_index(const void* p, int dt, long long i) {
  switch (dt) {
    case 0: return reinterpret_cast<const uint8_t*>(p)[i];
    case 1: return reinterpret_cast<const int8_t*>(p)[i];
    case 2: return reinterpret_cast<const int16_t*>(p)[i];
    case 4: return reinterpret_cast<const long long*>(p)[i];
    case 5: return (long long)reinterpret_cast<const float*>(p)[i];
    default: return (long long)__half2float(reinterpret_cast<const __half*>(p)[i]);
  }
}
__device__ __forceinline__ float ld_value(const void* p, int dt, long long i) {
  switch (dt) {
    case 0: return (float)reinterpret_cast<const uint8_t*>(p)[i];
    case 1: return (float)reinterpret_cast<const int8_t*>(p)[i];
    case 2: return (float)reinterpret_cast<const int16_t*>(p)[i];
    case 4: return (float)reinterpret_cast<const long long*>(p)[i];
    case 5: return reinterpret_cast<const float*>(p)[i];
    default: return __half2float(reinterpret_cast<const __half*>(p)[i]);
  }
}
__device__ __forceinline__ int clamp_idx(long long v, int V) { return v < 0 ? 0 : (v > V - 1 ? V - 1 : (int)v); }

// out[r, :] = relu(table[min(idx[r], V - 1), :])      (scalar_encoder.py:112-116: clamp, embedding, ReLU)
__global__ void embed_relu_fwd_kernel(const void* __restrict__ idx, int dt, const float* __restrict__ table, float* __restrict__ out,
                                      long long R, int V, int D) {
  const long long total = R * D;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / D;
    const int d = (int)(i - r * D);
    out[i] = fmaxf(__ldg(table + (long long)clamp_idx(ld_index(idx, dt, r), V) * D + d), 0.f);
  }
}
// one_hot[r, :] = e_{min(idx[r], V - 1)} as a bf16 GEMM operand [R, Vpad] (columns >= V zero): the weight gradient of an
// embedding / of a linear layer over a one-hot input is then an ordinary (deterministic, split-K) tensor-core GEMM
__global__ void one_hot_bf16_kernel(const void* __restrict__ idx, int dt, __nv_bfloat16* __restrict__ out, long long R, int V, int Vpad) {
  const long long total = R * Vpad;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / Vpad;
    const int v = (int)(i - r * Vpad);
    out[i] = __float2bfloat16_rn(v == clamp_idx(ld_index(idx, dt, r), V) ? 1.f : 0.f);
  }
}

// v[r, 2k] = sin(t[r] * pos[2k]), v[r, 2k + 1] = cos(t[r] * pos[2k + 1])            (scalar_encoder.py:91-97)
__global__ void time_encode_kernel(const void* __restrict__ t, int dt, const float* __restrict__ pos, float* __restrict__ out,
                                   long long R, int D) {
  const long long total = R * D;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / D;
    const int d = (int)(i - r * D);
    const float a = ld_value(t, dt, r) * __ldg(pos + d);
    out[i] = (d & 1) ? cosf(a) : sinf(a);
  }
}

// build-order token features [tokens, Kpad] bf16 = [one_hot(order, A) | eye(L)[position] | bits_NB(x) | bits_NB(y) | 0 pad]
// (scalar_encoder.py:41-50), x = loc % map_w, y = loc / map_w, bits MSB first (entity_encoder.py:12-17): the operand of the
// build-order Transformer's embedding GEMM
__global__ void bo_features_kernel(const void* __restrict__ order, int odt, const void* __restrict__ loc, int ldt,
                                   __nv_bfloat16* __restrict__ out, long long tokens, int L, int A, int NB, int map_w, int Kpad) {
  const long long total = tokens * Kpad;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long tk = i / Kpad;
    const int col = (int)(i - tk * Kpad);
    bool on = false;
    if (col < A) on = clamp_idx(ld_index(order, odt, tk), A) == col;
    else if (col < A + L) on = (int)(tk % L) == col - A;
    else if (col < A + L + 2 * NB) {
      const long long lc = ld_index(loc, ldt, tk);
      const int b = col - A - L;
      const int v = b < NB ? (int)(lc % map_w) : (int)(lc / map_w);
      on = (v >> (NB - 1 - (b % NB))) & 1;
    }
    out[i] = __float2bfloat16_rn(on ? 1.f : 0.f);
  }
}

// integer / half / float rows [rows, K] -> bf16 [rows, Kpad] (zero padded): GEMM operand of an fc field
__global__ void cast_any_bf16_kernel(const void* __restrict__ src, int dt, __nv_bfloat16* __restrict__ dst, long long rows, int K,
                                     int Kpad) {
  const long long total = rows * Kpad;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / Kpad;
    const int k = (int)(i - r * Kpad);
    dst[i] = __float2bfloat16_rn(k < K ? ld_value(src, dt, r * K + k) : 0.f);
  }
}

static int ew_blocks(long long n) { long long b = (n + 255) / 256; return (int)(b < 1 ? 1 : (b > 148 * 8 ? 148 * 8 : b)); }

int embed_relu_fwd(const void* idx, int dt, const float* table, float* out, long long R, int V, int D, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!idx || !table || !out || V <= 0 || D <= 0 || dt < 0 || dt > 5) return AS_EINVAL;
  embed_relu_fwd_kernel<<<ew_blocks(R * D), 256, 0, st>>>(idx, dt, table, out, R, V, D);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int one_hot_bf16(const void* idx, int dt, void* out, long long R, int V, int Vpad, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!idx || !out || V <= 0 || Vpad < V || dt < 0 || dt > 5) return AS_EINVAL;
  one_hot_bf16_kernel<<<ew_blocks(R * Vpad), 256, 0, st>>>(idx, dt, reinterpret_cast<__nv_bfloat16*>(out), R, V, Vpad);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int time_encode(const void* t, int dt, const float* pos, float* out, long long R, int D, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!t || !pos || !out || D <= 0 || dt < 0 || dt > 5) return AS_EINVAL;
  time_encode_kernel<<<ew_blocks(R * D), 256, 0, st>>>(t, dt, pos, out, R, D);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int bo_features(const void* order, int odt, const void* loc, int ldt, void* out, long long tokens, int L, int A, int NB, int map_w,
                int Kpad, cudaStream_t st) {
  if (tokens <= 0) return AS_OK;
  if (!order || !loc || !out || L <= 0 || A <= 0 || NB <= 0 || NB > 16 || map_w <= 0 || Kpad < A + L + 2 * NB) return AS_EINVAL;
  bo_features_kernel<<<ew_blocks(tokens * Kpad), 256, 0, st>>>(order, odt, loc, ldt, reinterpret_cast<__nv_bfloat16*>(out), tokens, L, A,
                                                              NB, map_w, Kpad);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int cast_any_bf16(const void* src, int dt, void* dst, long long rows, int K, int Kpad, cudaStream_t st) {
  if (rows <= 0) return AS_OK;
  if (!src || !dst || K <= 0 || Kpad < K || dt < 0 || dt > 5) return AS_EINVAL;
  cast_any_bf16_kernel<<<ew_blocks(rows * Kpad), 256, 0, st>>>(src, dt, reinterpret_cast<__nv_bfloat16*>(dst), rows, K, Kpad);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
