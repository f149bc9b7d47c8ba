// Small HBM-bound helpers: dtype casts, Gumbel-max sampling, fused global-norm clip + Adam.
#include "common.cuh"

namespace as {

// ------------------------------------------------------------------ fp32 -> bf16 (8 elements / thread / iteration)
__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n8 = n / 8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * i);
    const float4 b = __ldg(reinterpret_cast<const float4*>(src) + 2 * i + 1);
    __nv_bfloat162 h0 = __floats2bfloat162_rn(a.x, a.y), h1 = __floats2bfloat162_rn(a.z, a.w);
    __nv_bfloat162 h2 = __floats2bfloat162_rn(b.x, b.y), h3 = __floats2bfloat162_rn(b.z, b.w);
    uint4 u;
    u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
    u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
    reinterpret_cast<uint4*>(dst)[i] = u;
  }
  for (long long i = n8 * 8 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __float2bfloat16_rn(src[i]);
}

// ------------------------------------------------------------------ gradient packing
// table[c] = {source pointer (0: fill with zeros), destination offset, element count}: one CTA per chunk copies a piece of
// a parameter's gradient tensor into the flat gradient buffer.  Replaces one `grad += g` launch per parameter.
__global__ void __launch_bounds__(256) pack_f32_kernel(const long long* __restrict__ table, float* __restrict__ dst) {
  const long long* e = table + 3LL * blockIdx.x;
  const float* src = reinterpret_cast<const float*>(e[0]);
  float* d = dst + e[1];
  const int n = (int)e[2];
  if (src && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(d)) & 15) == 0) {
    const int n4 = n >> 2;
    for (int i = threadIdx.x; i < n4; i += 256) reinterpret_cast<float4*>(d)[i] = __ldg(reinterpret_cast<const float4*>(src) + i);
    for (int i = n4 * 4 + threadIdx.x; i < n; i += 256) d[i] = src[i];
  } else {
    for (int i = threadIdx.x; i < n; i += 256) d[i] = src ? src[i] : 0.f;
  }
}

int pack_f32(const long long* table, int chunks, float* dst, cudaStream_t st) {
  if (chunks <= 0) return AS_OK;
  if (!table || !dst) return AS_EINVAL;
  pack_f32_kernel<<<chunks, 256, 0, st>>>(table, dst);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int cast_f32_bf16(const float* src, void* dst, long long n, cudaStream_t st) {
  if (n <= 0) return AS_OK;
  if (!src || !dst) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(dst) & 15)) return AS_EALIGN;
  long long blocks = (n / 8 + 255) / 256;
  const int grid = (int)(blocks < 148LL * 8 ? (blocks > 0 ? blocks : 1) : 148LL * 8);
  cast_f32_bf16_kernel<<<grid, 256, 0, st>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------ ReLU backward on fp32 rows: dz = dy * (y > 0)
__global__ void relu_bwd_f32_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ dz,
                                    long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n4 = n / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 d = __ldg(reinterpret_cast<const float4*>(dy) + i);
    const float4 v = __ldg(reinterpret_cast<const float4*>(y) + i);
    reinterpret_cast<float4*>(dz)[i] = make_float4(v.x > 0.f ? d.x : 0.f, v.y > 0.f ? d.y : 0.f, v.z > 0.f ? d.z : 0.f,
                                                   v.w > 0.f ? d.w : 0.f);
  }
  for (long long i = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dz[i] = y[i] > 0.f ? dy[i] : 0.f;
}

int relu_bwd_f32(const float* dy, const float* y, float* dz, long long n, cudaStream_t st) {
  if (n <= 0) return AS_OK;
  if (!dy || !y || !dz) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(dz)) & 15) return AS_EALIGN;
  long long blocks = (n / 4 + 255) / 256;
  const int grid = (int)(blocks < 148LL * 8 ? (blocks > 0 ? blocks : 1) : 148LL * 8);
  relu_bwd_f32_kernel<<<grid, 256, 0, st>>>(dy, y, dz, n);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------ Gumbel-max sampling
// out[r] = argmax_k (logit[r,k] + noise[r,k]) in fp32, lowest index on ties (torch.argmax semantics).
template <int THREADS_PER_ROW>
__global__ void gumbel_argmax_kernel(const float* __restrict__ logit, const float* __restrict__ noise,
                                     long long* __restrict__ out, long long rows, int K) {
  constexpr int RPB = 256 / THREADS_PER_ROW;
  const int sub = threadIdx.x / THREADS_PER_ROW, tr = threadIdx.x % THREADS_PER_ROW;
  __shared__ float bv[8];
  __shared__ int bi[8];
  for (long long row0 = (long long)blockIdx.x * RPB; row0 < rows; row0 += (long long)gridDim.x * RPB) {
    const long long row = row0 + sub;
    float best = -INFINITY;
    int idx = 0x7fffffff;
    if (row < rows) {
      const float* x = logit + row * (long long)K;
      const float* g = noise + row * (long long)K;
      for (int k = tr; k < K; k += THREADS_PER_ROW) {
        const float v = __ldg(x + k) + __ldg(g + k);
        if (v > best || (v == best && k < idx)) { best = v; idx = k; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
      if (ov > best || (ov == best && oi < idx)) { best = ov; idx = oi; }
    }
    if (THREADS_PER_ROW > 32) {
      __syncthreads();
      if ((threadIdx.x & 31) == 0) { bv[threadIdx.x >> 5] = best; bi[threadIdx.x >> 5] = idx; }
      __syncthreads();
      if (threadIdx.x == 0) {
        for (int i = 1; i < 8; ++i)
          if (bv[i] > best || (bv[i] == best && bi[i] < idx)) { best = bv[i]; idx = bi[i]; }
        if (row < rows) out[row] = (idx == 0x7fffffff) ? 0 : idx;
      }
    } else if (tr == 0 && row < rows) {
      out[row] = (idx == 0x7fffffff) ? 0 : idx;
    }
  }
}

int gumbel_argmax(const float* logit, const float* noise, long long* out, long long rows, int K, cudaStream_t st) {
  if (rows <= 0) return AS_OK;
  if (K <= 0) return AS_ESIZE;
  if (!logit || !noise || !out) return AS_EINVAL;
  if (K >= 4096) {
    const int grid = (int)(rows < 148LL * 8 ? rows : 148LL * 8);
    gumbel_argmax_kernel<256><<<grid, 256, 0, st>>>(logit, noise, out, rows, K);
  } else {
    long long blocks = (rows + 7) / 8;
    const int grid = (int)(blocks < 148LL * 8 ? blocks : 148LL * 8);
    gumbel_argmax_kernel<32><<<grid, 256, 0, st>>>(logit, noise, out, rows, K);
  }
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------ fused global-norm clip + Adam
// pass 1: per-block partial sums of g^2 into scratch; pass 2: every block re-reduces the partials in a fixed order
// (deterministic), derives clip = min(1, max_norm / (norm + 1e-6)) as torch.nn.utils.clip_grad_norm_ does, then
// applies Adam (reference settings rl_learner.py:73-80) and refreshes the bf16 shadow of the weights.
__global__ void sumsq_partial_kernel(const float* __restrict__ g, long long n, float* __restrict__ partial,
                                     int* __restrict__ step_dev) {
  if (step_dev != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *step_dev += 1;   // device-side step counter (graph replay)
  float s = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n4 = n / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(g) + i);
    s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  for (long long i = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) s += g[i] * g[i];
  __shared__ float red[32];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) partial[blockIdx.x] = v;
  }
}

__global__ void clip_adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                 float* __restrict__ v, __nv_bfloat16* __restrict__ shadow, long long n,
                                 const float* __restrict__ partial, int n_partial, float max_norm, float lr,
                                 float beta1, float beta2, float eps, float wd, float bc1, float bc2,
                                 const int* __restrict__ step_dev, float* __restrict__ norm_out,
                                 const unsigned short* __restrict__ blk_param, const int* __restrict__ p_active,
                                 const int* __restrict__ p_step) {
  // blk_param != nullptr: torch.optim.Adam semantics per parameter tensor -- element i belongs to parameter
  // blk_param[i / 8] (tensors start at multiples of 8 elements); a parameter without a gradient this step
  // (p_active == 0) is left untouched, the others use their OWN step count p_step[p] + 1 for the bias corrections.
  __shared__ float s_clip;
  if (step_dev != nullptr) {                      // bias corrections from the device counter
    const float t = (float)(*step_dev);
    bc1 = 1.f - powf(beta1, t);
    bc2 = 1.f - powf(beta2, t);
  }
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int i = 0; i < n_partial; ++i) tot += partial[i];
    const float norm = sqrtf(tot);
    float c = max_norm > 0.f ? max_norm / (norm + 1e-6f) : 1.f;
    s_clip = c < 1.f ? c : 1.f;
    if (blockIdx.x == 0 && norm_out) *norm_out = norm;
  }
  __syncthreads();
  const float clip = s_clip;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  auto update = [&](float gi, float& pi, float& mi, float& vi) {
    gi *= clip;
    if (wd != 0.f) gi += wd * pi;
    mi = beta1 * mi + (1.f - beta1) * gi;
    vi = beta2 * vi + (1.f - beta2) * gi * gi;
    pi -= lr * (mi / bc1) / (sqrtf(vi / bc2) + eps);
  };
  int last_p = -1;
  bool on = true;
  auto select = [&](long long blk) {                  // per-parameter mode: activity and bias corrections of this block
    const int pp = blk_param[blk];
    if (pp != last_p) {
      last_p = pp;
      on = p_active[pp] != 0;
      const float t = (float)(p_step[pp] + 1);
      bc1 = 1.f - powf(beta1, t);
      bc2 = 1.f - powf(beta2, t);
    }
    return on;
  };
  // 16-byte accesses when the four streams allow it (the flat buffers of learner.FlatParams are 32-byte aligned)
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                     reinterpret_cast<uintptr_t>(v)) & 15) == 0 && (shadow == nullptr || (reinterpret_cast<uintptr_t>(shadow) & 7) == 0);
  const long long n4 = vec ? n / 4 : 0;
  for (long long i = tid; i < n4; i += stride) {
    if (blk_param != nullptr && !select(i >> 1)) continue;
    const float4 g4 = reinterpret_cast<const float4*>(g)[i];
    float4 p4 = reinterpret_cast<float4*>(p)[i], m4 = reinterpret_cast<float4*>(m)[i], v4 = reinterpret_cast<float4*>(v)[i];
    update(g4.x, p4.x, m4.x, v4.x); update(g4.y, p4.y, m4.y, v4.y); update(g4.z, p4.z, m4.z, v4.z); update(g4.w, p4.w, m4.w, v4.w);
    reinterpret_cast<float4*>(m)[i] = m4; reinterpret_cast<float4*>(v)[i] = v4; reinterpret_cast<float4*>(p)[i] = p4;
    if (shadow) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(p4.x, p4.y), hi = __floats2bfloat162_rn(p4.z, p4.w);
      uint2 u;
      u.x = *reinterpret_cast<uint32_t*>(&lo); u.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(shadow)[i] = u;
    }
  }
  for (long long i = n4 * 4 + tid; i < n; i += stride) {
    if (blk_param != nullptr && !select(i >> 3)) continue;
    float pi = p[i], mi = m[i], vi = v[i];
    update(g[i], pi, mi, vi);
    m[i] = mi; v[i] = vi; p[i] = pi;
    if (shadow) shadow[i] = __float2bfloat16_rn(pi);
  }
}

constexpr int kAdamBlocks = 148 * 4;

long long clip_adam_scratch_bytes() { return kAdamBlocks * sizeof(float); }

int clip_adam(float* p, const float* g, float* m, float* v, void* shadow, long long n, float max_norm, float lr,
              float beta1, float beta2, float eps, float wd, int step, int* step_dev, float* scratch, float* norm_out,
              cudaStream_t st) {
  if (n <= 0) return AS_OK;
  if (!p || !g || !m || !v || !scratch || (step <= 0 && !step_dev)) return AS_EINVAL;
  if (reinterpret_cast<uintptr_t>(g) & 15) return AS_EALIGN;
  sumsq_partial_kernel<<<kAdamBlocks, 256, 0, st>>>(g, n, scratch, step_dev);
  AS_LAUNCH_CHECK();
  const float bc1 = 1.f - powf(beta1, (float)(step > 0 ? step : 1)), bc2 = 1.f - powf(beta2, (float)(step > 0 ? step : 1));
  clip_adam_kernel<<<kAdamBlocks, 256, 0, st>>>(p, g, m, v, reinterpret_cast<__nv_bfloat16*>(shadow), n, scratch,
                                                kAdamBlocks, max_norm, lr, beta1, beta2, eps, wd, bc1, bc2, step_dev,
                                                norm_out, nullptr, nullptr, nullptr);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

__global__ void adam_bump_steps_kernel(const int* __restrict__ p_active, int* __restrict__ p_step, int n_params) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_params && p_active[i] != 0) p_step[i] += 1;
}

int clip_adam_per_param(float* p, const float* g, float* m, float* v, void* shadow, long long n, float max_norm, float lr,
                        float beta1, float beta2, float eps, float wd, const unsigned short* blk_param, const int* p_active,
                        int* p_step, int n_params, int* step_dev, float* scratch, float* norm_out, cudaStream_t st) {
  if (n <= 0) return AS_OK;
  if (!p || !g || !m || !v || !scratch || !blk_param || !p_active || !p_step || n_params <= 0 || n_params > 65535) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(g) & 15) || (n % 8) != 0) return AS_EALIGN;
  sumsq_partial_kernel<<<kAdamBlocks, 256, 0, st>>>(g, n, scratch, step_dev);
  AS_LAUNCH_CHECK();
  clip_adam_kernel<<<kAdamBlocks, 256, 0, st>>>(p, g, m, v, reinterpret_cast<__nv_bfloat16*>(shadow), n, scratch,
                                                kAdamBlocks, max_norm, lr, beta1, beta2, eps, wd, 1.f, 1.f, nullptr,
                                                norm_out, blk_param, p_active, p_step);
  AS_LAUNCH_CHECK();
  adam_bump_steps_kernel<<<(n_params + 255) / 256, 256, 0, st>>>(p_active, p_step, n_params);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// out[r, :] = row_of[r] >= 0 ? rows[row_of[r], :] : fill      (dense [total, K] from a compact list of its non-padding rows)
__global__ void expand_rows_kernel(const float* __restrict__ rows, const int* __restrict__ row_of, float* __restrict__ out,
                                   long long total, int K, float fill) {
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = warp_global; r < total; r += warps_total) {
    const int src = __ldg(row_of + r);
    float* o = out + r * K;
    if (src >= 0) {
      const float* in = rows + (long long)src * K;
      for (int k = lane; k < K; k += 32) o[k] = __ldg(in + k);
    } else {
      for (int k = lane; k < K; k += 32) o[k] = fill;
    }
  }
}
int expand_rows(const float* rows, const int* row_of, float* out, long long total, int K, float fill, cudaStream_t st) {
  if (total <= 0) return AS_OK;
  if (!row_of || !out || K <= 0) return AS_EINVAL;
  long long blocks = (total + 7) / 8;
  expand_rows_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, st>>>(rows, row_of, out, total, K, fill);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
