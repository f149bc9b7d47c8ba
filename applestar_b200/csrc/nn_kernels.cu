// Row-wise HBM-bound kernels with warp-shuffle reductions: LayerNorm (+residual, +ReLU) forward / backward,
// masked attention softmax forward / backward, and the bf16 one-hot builder that feeds the entity embedding GEMM.
#include "common.cuh"

namespace as {

constexpr int kMaxPerLane = 64;   // row width <= 32 * 64 = 2048

// ------------------------------------------------------------------------------------------ LayerNorm
// y = LN(x (+ res)) * gamma + beta (eps 1e-5, biased variance, fp32 statistics) (+ReLU); optional fp32 and bf16 copies.
template <int PER_LANE>
__global__ void __launch_bounds__(256)
layer_norm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ res, const float* __restrict__ gamma,
                      const float* __restrict__ beta, float* __restrict__ y32, __nv_bfloat16* __restrict__ y16,
                      float* __restrict__ stats, long long rows, int C, int relu) {
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp_global; row < rows; row += warps_total) {
    float v[PER_LANE];
    const float* xr = x + row * C;
    const float* rr = res ? res + row * C : nullptr;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      float t = 0.f;
      if (c < C) { t = __ldg(xr + c); if (rr) t += __ldg(rr + c); }
      v[i] = t;
      s += t;
    }
    const float mean = warp_sum(s) / (float)C;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      const float d = (c < C) ? v[i] - mean : 0.f;
      q += d * d;
    }
    const float rstd = rsqrtf(warp_sum(q) / (float)C + 1e-5f);
    if (lane == 0 && stats) { stats[row * 2] = mean; stats[row * 2 + 1] = rstd; }
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      if (c < C) {
        float o = (v[i] - mean) * rstd * __ldg(gamma + c) + __ldg(beta + c);
        if (relu) o = fmaxf(o, 0.f);
        if (y32) y32[row * C + c] = o;
        if (y16) y16[row * C + c] = __float2bfloat16_rn(o);
      }
    }
  }
}

// dx = rstd * (g*gamma - mean(g*gamma) - xhat * mean(g*gamma*xhat)),  g = dy32 + dy16 (masked by ReLU if fused)
template <int PER_LANE>
__global__ void __launch_bounds__(256)
layer_norm_bwd_kernel(const float* __restrict__ dy32, const __nv_bfloat16* __restrict__ dy16,
                      const float* __restrict__ x, const float* __restrict__ res, const float* __restrict__ gamma,
                      const float* __restrict__ beta, const float* __restrict__ stats, float* __restrict__ dx,
                      long long rows, int C, int relu) {
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp_global; row < rows; row += warps_total) {
    const float mean = stats[row * 2], rstd = stats[row * 2 + 1];
    float xh[PER_LANE], dg[PER_LANE];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      float a = 0.f, b = 0.f;
      if (c < C) {
        float t = __ldg(x + row * C + c);
        if (res) t += __ldg(res + row * C + c);
        a = (t - mean) * rstd;
        float g = 0.f;
        if (dy32) g += __ldg(dy32 + row * C + c);
        if (dy16) g += __bfloat162float(dy16[row * C + c]);
        const float gm = __ldg(gamma + c);
        if (relu && (a * gm + __ldg(beta + c)) <= 0.f) g = 0.f;
        b = g * gm;
      }
      xh[i] = a; dg[i] = b;
      s1 += b; s2 += a * b;
    }
    const float m1 = warp_sum(s1) / (float)C, m2 = warp_sum(s2) / (float)C;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      if (c < C) dx[row * C + c] = rstd * (dg[i] - m1 - xh[i] * m2);
    }
  }
}

// 128-bit variants for C % 128 == 0 (256 / 384 / 1536): each lane owns float4 columns lane + 32 i.
template <int V>
__global__ void __launch_bounds__(256)
layer_norm_fwd_v4_kernel(const float* __restrict__ x, const float* __restrict__ res, const float* __restrict__ gamma,
                         const float* __restrict__ beta, float* __restrict__ y32, __nv_bfloat16* __restrict__ y16,
                         float* __restrict__ stats, long long rows, int C, int relu) {
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp_global; row < rows; row += warps_total) {
    float4 v[V];
    const float4* xr = reinterpret_cast<const float4*>(x + row * C);
    const float4* rr = res ? reinterpret_cast<const float4*>(res + row * C) : nullptr;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      float4 t = __ldg(xr + lane + i * 32);
      if (rr) { const float4 u = __ldg(rr + lane + i * 32); t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
      v[i] = t;
      s += t.x + t.y + t.z + t.w;
    }
    const float mean = warp_sum(s) / (float)C;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += a * a + b * b + c * c + d * d;
    }
    const float rstd = rsqrtf(warp_sum(q) / (float)C + 1e-5f);
    if (lane == 0 && stats) { stats[row * 2] = mean; stats[row * 2 + 1] = rstd; }
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + lane + i * 32);
      const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + lane + i * 32);
      float4 o;
      o.x = (v[i].x - mean) * rstd * g.x + b.x; o.y = (v[i].y - mean) * rstd * g.y + b.y;
      o.z = (v[i].z - mean) * rstd * g.z + b.z; o.w = (v[i].w - mean) * rstd * g.w + b.w;
      if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
      if (y32) reinterpret_cast<float4*>(y32 + row * C)[lane + i * 32] = o;
      if (y16) {
        __nv_bfloat162 h0 = __floats2bfloat162_rn(o.x, o.y), h1 = __floats2bfloat162_rn(o.z, o.w);
        uint2 u;
        u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
        reinterpret_cast<uint2*>(y16 + row * C)[lane + i * 32] = u;
      }
    }
  }
}

template <int V, bool PG, bool DZ = false>
__global__ void __launch_bounds__(256)
layer_norm_bwd_v4_kernel(const float* __restrict__ dy32, const __nv_bfloat16* __restrict__ dy16,
                         const float* __restrict__ x, const float* __restrict__ res, const float* __restrict__ gamma,
                         const float* __restrict__ beta, const float* __restrict__ stats, float* __restrict__ dx,
                         long long rows, int C, int relu, float* __restrict__ pgrad,
                         __nv_bfloat16* __restrict__ dz16 = nullptr, int mask_x = 0) {
  // DZ: `x` is the output of a linear layer (bias, optional ReLU) whose backward follows: the kernel also writes that
  // layer's pre-activation gradient dz16 = bf16(dx (* (x > 0) when mask_x)) -- the operand of its dgrad / wgrad GEMMs --
  // and accumulates its column sums (the bias gradient) as a third partial row, so no separate pass reads dx again.
  // pgrad != NULL (PG): the parameter gradients are accumulated in the same pass -- every lane owns 4V fixed columns,
  // sums g*xhat and g over the rows of its warp, the block's 8 warps are combined in shared memory and one partial
  // row [2][C] per block goes to pgrad (reduced in fixed order by layer_norm_param_reduce_kernel).
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  constexpr int NP = DZ ? 3 : 2;
  float4 pg[PG ? V : 1], pb[PG ? V : 1], pz[DZ ? V : 1];
  if (PG) {
#pragma unroll
    for (int i = 0; i < V; ++i) { pg[i] = make_float4(0.f, 0.f, 0.f, 0.f); pb[i] = make_float4(0.f, 0.f, 0.f, 0.f); }
  }
  if (DZ) {
#pragma unroll
    for (int i = 0; i < V; ++i) pz[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  bool xpos[DZ ? 4 * V : 1];
  for (long long row = warp_global; row < rows; row += warps_total) {
    const float mean = stats[row * 2], rstd = stats[row * 2 + 1];
    float4 xh[V], dg[V];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const int idx = lane + i * 32;
      float4 t = __ldg(reinterpret_cast<const float4*>(x + row * C) + idx);
      if (DZ) { xpos[4 * i] = t.x > 0.f; xpos[4 * i + 1] = t.y > 0.f; xpos[4 * i + 2] = t.z > 0.f; xpos[4 * i + 3] = t.w > 0.f; }
      if (res) { const float4 u = __ldg(reinterpret_cast<const float4*>(res + row * C) + idx); t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
      float4 a = make_float4((t.x - mean) * rstd, (t.y - mean) * rstd, (t.z - mean) * rstd, (t.w - mean) * rstd);
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
      if (dy32) g = __ldg(reinterpret_cast<const float4*>(dy32 + row * C) + idx);
      if (dy16) {
        const uint2 u = *(reinterpret_cast<const uint2*>(dy16 + row * C) + idx);
        const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
        const float2 f1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
        g.x += f0.x; g.y += f0.y; g.z += f1.x; g.w += f1.y;
      }
      const float4 gm = __ldg(reinterpret_cast<const float4*>(gamma) + idx);
      if (relu) {
        const float4 bt = __ldg(reinterpret_cast<const float4*>(beta) + idx);
        if (a.x * gm.x + bt.x <= 0.f) g.x = 0.f;
        if (a.y * gm.y + bt.y <= 0.f) g.y = 0.f;
        if (a.z * gm.z + bt.z <= 0.f) g.z = 0.f;
        if (a.w * gm.w + bt.w <= 0.f) g.w = 0.f;
      }
      if (PG) {
        pg[i].x += g.x * a.x; pg[i].y += g.y * a.y; pg[i].z += g.z * a.z; pg[i].w += g.w * a.w;
        pb[i].x += g.x; pb[i].y += g.y; pb[i].z += g.z; pb[i].w += g.w;
      }
      const float4 b = make_float4(g.x * gm.x, g.y * gm.y, g.z * gm.z, g.w * gm.w);
      xh[i] = a; dg[i] = b;
      s1 += b.x + b.y + b.z + b.w;
      s2 += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
    }
    const float m1 = warp_sum(s1) / (float)C, m2 = warp_sum(s2) / (float)C;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      float4 o;
      o.x = rstd * (dg[i].x - m1 - xh[i].x * m2); o.y = rstd * (dg[i].y - m1 - xh[i].y * m2);
      o.z = rstd * (dg[i].z - m1 - xh[i].z * m2); o.w = rstd * (dg[i].w - m1 - xh[i].w * m2);
      if (dx) reinterpret_cast<float4*>(dx + row * C)[lane + i * 32] = o;
      if (DZ) {
        if (mask_x) {
          o.x = xpos[4 * i] ? o.x : 0.f; o.y = xpos[4 * i + 1] ? o.y : 0.f;
          o.z = xpos[4 * i + 2] ? o.z : 0.f; o.w = xpos[4 * i + 3] ? o.w : 0.f;
        }
        pz[i].x += o.x; pz[i].y += o.y; pz[i].z += o.z; pz[i].w += o.w;
        __nv_bfloat162 h0 = __floats2bfloat162_rn(o.x, o.y), h1 = __floats2bfloat162_rn(o.z, o.w);
        uint2 u;
        u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
        reinterpret_cast<uint2*>(dz16 + row * C)[lane + i * 32] = u;
      }
    }
  }
  if (PG) {
    extern __shared__ float4 s_pg[];                     // [8 warps][NP][V*32] float4
    const int w = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      s_pg[(w * NP + 0) * (V * 32) + lane + i * 32] = pg[i];
      s_pg[(w * NP + 1) * (V * 32) + lane + i * 32] = pb[i];
      if (DZ) s_pg[(w * NP + 2) * (V * 32) + lane + i * 32] = pz[i];
    }
    __syncthreads();
    for (int j = threadIdx.x; j < NP * V * 32; j += blockDim.x) {
      const int which = j / (V * 32), col4 = j - which * (V * 32);
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float4 u = s_pg[(k * NP + which) * (V * 32) + col4];
        t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
      }
      reinterpret_cast<float4*>(pgrad + ((long long)blockIdx.x * NP + which) * C)[col4] = t;
    }
  }
}

// fixed-order reduction of [chunks][3][C] partials -> dgamma, dbeta, dbias
__global__ void __launch_bounds__(256)
layer_norm_param_reduce3_kernel(const float* __restrict__ partial, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                float* __restrict__ dbias, int C, int chunks) {
  const int cx = threadIdx.x & 31, ky = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f;
  if (c < C)
    for (int k = ky; k < chunks; k += 8) {
      a0 += partial[((long long)k * 3 + 0) * C + c];
      a1 += partial[((long long)k * 3 + 1) * C + c];
      a2 += partial[((long long)k * 3 + 2) * C + c];
    }
  __shared__ float s0[8][33], s1[8][33], s2[8][33];
  s0[ky][cx] = a0; s1[ky][cx] = a1; s2[ky][cx] = a2;
  __syncthreads();
  if (ky == 0 && c < C) {
    float t0 = 0.f, t1 = 0.f, t2 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { t0 += s0[i][cx]; t1 += s1[i][cx]; t2 += s2[i][cx]; }
    dgamma[c] = t0;
    dbeta[c] = t1;
    if (dbias) dbias[c] = t2;
  }
}

// dgamma[c] = sum_r g * xhat, dbeta[c] = sum_r g.  Block = 32 columns x 8 row lanes over one row chunk (128-byte
// coalesced row segments), shared-memory reduce over the row lanes, per-chunk partials, then a fixed-order reduce
// over chunks (deterministic, no atomics).
__global__ void __launch_bounds__(256)
layer_norm_param_grad_kernel(const float* __restrict__ dy32, const __nv_bfloat16* __restrict__ dy16,
                             const float* __restrict__ x, const float* __restrict__ res,
                             const float* __restrict__ gamma, const float* __restrict__ beta,
                             const float* __restrict__ stats, float* __restrict__ partial,
                             long long rows, int C, int relu, int chunks) {
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  const int chunk = blockIdx.y;
  const long long per = (rows + chunks - 1) / chunks;
  const long long r0 = chunk * per, r1 = (r0 + per < rows) ? r0 + per : rows;
  float ag = 0.f, ab = 0.f;
  if (c < C) {
    const float gm = gamma[c], bt = beta[c];
    for (long long r = r0 + ry; r < r1; r += 8) {
      float t = __ldg(x + r * C + c);
      if (res) t += __ldg(res + r * C + c);
      const float a = (t - __ldg(stats + r * 2)) * __ldg(stats + r * 2 + 1);
      float g = 0.f;
      if (dy32) g += __ldg(dy32 + r * C + c);
      if (dy16) g += __bfloat162float(dy16[r * C + c]);
      if (relu && (a * gm + bt) <= 0.f) g = 0.f;
      ag += g * a;
      ab += g;
    }
  }
  __shared__ float sg[8][33], sb[8][33];
  sg[ry][cx] = ag; sb[ry][cx] = ab;
  __syncthreads();
  if (ry == 0 && c < C) {
    float tg = 0.f, tb = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { tg += sg[i][cx]; tb += sb[i][cx]; }
    partial[((long long)chunk * 2 + 0) * C + c] = tg;
    partial[((long long)chunk * 2 + 1) * C + c] = tb;
  }
}

// fixed-order reduction of the per-chunk partials: 32 columns x 8 chunk lanes per block
__global__ void __launch_bounds__(256)
layer_norm_param_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dgamma,
                               float* __restrict__ dbeta, int C, int chunks) {
  const int cx = threadIdx.x & 31, ky = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float ag = 0.f, ab = 0.f;
  if (c < C)
    for (int k = ky; k < chunks; k += 8) {
      ag += partial[((long long)k * 2 + 0) * C + c];
      ab += partial[((long long)k * 2 + 1) * C + c];
    }
  __shared__ float sg[8][33], sb[8][33];
  sg[ky][cx] = ag; sb[ky][cx] = ab;
  __syncthreads();
  if (ky == 0 && c < C) {
    float tg = 0.f, tb = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { tg += sg[i][cx]; tb += sb[i][cx]; }
    dgamma[c] = tg;
    dbeta[c] = tb;
  }
}

static int ln_grid(long long rows) {
  long long blocks = (rows + 7) / 8;
  const long long cap = 148LL * 8;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

#define LN_DISPATCH(KERNEL, ...)                                                         \
  do {                                                                                   \
    const int per = (C + 31) / 32;                                                       \
    if (per <= 1) KERNEL<1><<<grid, 256, 0, st>>>(__VA_ARGS__);                          \
    else if (per <= 4) KERNEL<4><<<grid, 256, 0, st>>>(__VA_ARGS__);                     \
    else if (per <= 8) KERNEL<8><<<grid, 256, 0, st>>>(__VA_ARGS__);                     \
    else if (per <= 12) KERNEL<12><<<grid, 256, 0, st>>>(__VA_ARGS__);                   \
    else if (per <= 32) KERNEL<32><<<grid, 256, 0, st>>>(__VA_ARGS__);                   \
    else KERNEL<48><<<grid, 256, 0, st>>>(__VA_ARGS__);                                  \
  } while (0)

int layer_norm_fwd(const float* x, const float* res, const float* gamma, const float* beta, float* y32, void* y16,
                   float* stats, long long rows, int C, int relu, cudaStream_t st) {
  if (rows <= 0) return AS_OK;
  if (C <= 0 || C > 32 * 48) return AS_ESIZE;
  if (!x || !gamma || !beta || (!y32 && !y16)) return AS_EINVAL;
  const int grid = ln_grid(rows);
  const bool al = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(res) | reinterpret_cast<uintptr_t>(y32) |
                    reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta)) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(y16) & 7) == 0;
  __nv_bfloat16* y16b = reinterpret_cast<__nv_bfloat16*>(y16);
  if (al && C == 256) layer_norm_fwd_v4_kernel<2><<<grid, 256, 0, st>>>(x, res, gamma, beta, y32, y16b, stats, rows, C, relu);
  else if (al && C == 384) layer_norm_fwd_v4_kernel<3><<<grid, 256, 0, st>>>(x, res, gamma, beta, y32, y16b, stats, rows, C, relu);
  else if (al && C == 1536) layer_norm_fwd_v4_kernel<12><<<grid, 256, 0, st>>>(x, res, gamma, beta, y32, y16b, stats, rows, C, relu);
  else
  LN_DISPATCH(layer_norm_fwd_kernel, x, res, gamma, beta, y32, reinterpret_cast<__nv_bfloat16*>(y16), stats, rows, C, relu);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

constexpr int kLnChunks = 148 * 8;      // = the largest grid of the fused backward (one partial row per block)
long long layer_norm_bwd_scratch_floats(int C) { return (long long)kLnChunks * 2 * C; }

int layer_norm_bwd(const float* dy32, const void* dy16, const float* x, const float* res, const float* gamma,
                   const float* beta, const float* stats, float* dx, float* dgamma, float* dbeta, float* scratch,
                   long long rows, int C, int relu, cudaStream_t st) {
  if (rows <= 0) return AS_OK;
  if (C <= 0 || C > 32 * 48) return AS_ESIZE;
  if (!x || !gamma || !beta || !stats || (!dy32 && !dy16)) return AS_EINVAL;
  const __nv_bfloat16* d16 = reinterpret_cast<const __nv_bfloat16*>(dy16);
  bool pg_done = false;
  if (dx) {
    const int grid = ln_grid(rows);
    const bool al = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(res) | reinterpret_cast<uintptr_t>(dy32) |
                      reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta) | reinterpret_cast<uintptr_t>(dx)) & 15) == 0 &&
                    (reinterpret_cast<uintptr_t>(dy16) & 7) == 0;
    const bool fuse = dgamma && dbeta && scratch && (reinterpret_cast<uintptr_t>(scratch) & 15) == 0;
    if (al && C == 256 && fuse) {
      layer_norm_bwd_v4_kernel<2, true><<<grid, 256, 8 * 2 * 64 * 16, st>>>(dy32, d16, x, res, gamma, beta, stats, dx, rows, C, relu, scratch);
      pg_done = true;
    } else if (al && C == 384 && fuse) {
      layer_norm_bwd_v4_kernel<3, true><<<grid, 256, 8 * 2 * 96 * 16, st>>>(dy32, d16, x, res, gamma, beta, stats, dx, rows, C, relu, scratch);
      pg_done = true;
    }
    else if (al && C == 256) layer_norm_bwd_v4_kernel<2, false><<<grid, 256, 0, st>>>(dy32, d16, x, res, gamma, beta, stats, dx, rows, C, relu, nullptr);
    else if (al && C == 384) layer_norm_bwd_v4_kernel<3, false><<<grid, 256, 0, st>>>(dy32, d16, x, res, gamma, beta, stats, dx, rows, C, relu, nullptr);
    else if (al && C == 1536) layer_norm_bwd_v4_kernel<12, false><<<grid, 256, 0, st>>>(dy32, d16, x, res, gamma, beta, stats, dx, rows, C, relu, nullptr);
    else
    LN_DISPATCH(layer_norm_bwd_kernel, dy32, d16, x, res, gamma, beta, stats, dx, rows, C, relu);
    AS_LAUNCH_CHECK();
    if (pg_done) {
      layer_norm_param_reduce_kernel<<<(C + 31) / 32, 256, 0, st>>>(scratch, dgamma, dbeta, C, grid);
      AS_LAUNCH_CHECK();
    }
  }
  if (dgamma && dbeta && !pg_done) {
    if (!scratch) return AS_EINVAL;
    long long want = (rows + 255) / 256;
    const int chunks = (int)(want < 1 ? 1 : (want > kLnChunks ? kLnChunks : want));
    dim3 g((C + 31) / 32, chunks);
    layer_norm_param_grad_kernel<<<g, 256, 0, st>>>(dy32, d16, x, res, gamma, beta, stats, scratch, rows, C, relu, chunks);
    AS_LAUNCH_CHECK();
    layer_norm_param_reduce_kernel<<<(C + 31) / 32, 256, 0, st>>>(scratch, dgamma, dbeta, C, chunks);
    AS_LAUNCH_CHECK();
  }
  return AS_OK;
}

int layer_norm_param_reduce3(const float* partial, float* dgamma, float* dbeta, float* dbias, int C, int chunks, cudaStream_t st) {
  layer_norm_param_reduce3_kernel<<<(C + 31) / 32, 256, 0, st>>>(partial, dgamma, dbeta, dbias, C, chunks);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// LayerNorm backward of y = LN(res + x) where x = act(linear(...)) is followed directly by that linear layer's backward:
// one pass gives dx (fp32, the residual-path gradient; may be NULL), dz16 = bf16(dx * (x > 0 if mask_x)), dgamma, dbeta
// and dbias = column sums of dz.  C = 256 or 384; scratch = as_layer_norm_bwd_fused_scratch_floats(C) floats.
long long layer_norm_bwd_fused_scratch_floats(int C) { return (long long)kLnChunks * 3 * C; }

int layer_norm_bwd_fused(const float* dy32, const void* dy16, const float* x, const float* res, const float* gamma,
                         const float* beta, const float* stats, float* dx, void* dz16, int mask_x, float* dgamma,
                         float* dbeta, float* dbias, float* scratch, long long rows, int C, int relu, cudaStream_t st) {
  if (rows <= 0) return AS_OK;
  if (C != 256 && C != 384) return AS_EUNSUPPORTED;
  if (!x || !gamma || !beta || !stats || (!dy32 && !dy16) || !dz16 || !dgamma || !dbeta || !scratch) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(res) | reinterpret_cast<uintptr_t>(dy32) |
       reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta) | reinterpret_cast<uintptr_t>(dx) |
       reinterpret_cast<uintptr_t>(scratch)) & 15) return AS_EALIGN;
  if ((reinterpret_cast<uintptr_t>(dy16) | reinterpret_cast<uintptr_t>(dz16)) & 7) return AS_EALIGN;
  const __nv_bfloat16* d16 = reinterpret_cast<const __nv_bfloat16*>(dy16);
  __nv_bfloat16* z16 = reinterpret_cast<__nv_bfloat16*>(dz16);
  const int grid = ln_grid(rows);
  if (C == 256)
    layer_norm_bwd_v4_kernel<2, true, true><<<grid, 256, 8 * 3 * 64 * 16, st>>>(dy32, d16, x, res, gamma, beta, stats, dx, rows, C,
                                                                                  relu, scratch, z16, mask_x);
  else
    layer_norm_bwd_v4_kernel<3, true, true><<<grid, 256, 8 * 3 * 96 * 16, st>>>(dy32, d16, x, res, gamma, beta, stats, dx, rows, C,
                                                                                  relu, scratch, z16, mask_x);
  AS_LAUNCH_CHECK();
  layer_norm_param_reduce3_kernel<<<(C + 31) / 32, 256, 0, st>>>(scratch, dgamma, dbeta, dbias, C, grid);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------------------------------ masked softmax
// scores [G, Nq, Nk] fp32 (already scaled), key j of group g is valid iff j < key_len[g / groups_per_len]
// (NULL = all valid) -> probabilities bf16, masked keys exactly 0 (== exp(-1e9 - max) in the reference).
__device__ __forceinline__ float ld_score(const float* p) { return __ldg(p); }
__device__ __forceinline__ float ld_score(const __nv_bfloat16* p) { return __bfloat162float(*p); }

template <int PER_LANE, typename ST>
__global__ void __launch_bounds__(256)
softmax_fwd_kernel(const ST* __restrict__ s, const long long* __restrict__ key_len, __nv_bfloat16* __restrict__ p,
                   long long rows, int Nq, int Nk, int groups_per_len) {
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp_global; row < rows; row += warps_total) {
    int len = Nk;
    if (key_len) {
      const long long l = key_len[(row / Nq) / groups_per_len];
      len = l < Nk ? (int)l : Nk;
    }
    float v[PER_LANE];
    float m = -INFINITY;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      v[i] = (c < len) ? ld_score(s + row * Nk + c) : -INFINITY;
      m = fmaxf(m, v[i]);
    }
    m = warp_max(m);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      v[i] = (c < len) ? expf(v[i] - m) : 0.f;
      sum += v[i];
    }
    const float inv = 1.f / warp_sum(sum);
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      if (c < Nk) p[row * Nk + c] = __float2bfloat16_rn(v[i] * inv);
    }
  }
}

// ds = p * (dp - sum_j dp_j p_j)   (bf16 out; the 1/sqrt(d) scale is folded into the following GEMMs)
template <int PER_LANE, typename ST>
__global__ void __launch_bounds__(256)
softmax_bwd_kernel(const __nv_bfloat16* __restrict__ p, const ST* __restrict__ dp, __nv_bfloat16* __restrict__ ds,
                   long long rows, int Nk) {
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp_global; row < rows; row += warps_total) {
    float pv[PER_LANE], dv[PER_LANE];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      pv[i] = (c < Nk) ? __bfloat162float(p[row * Nk + c]) : 0.f;
      dv[i] = (c < Nk) ? ld_score(dp + row * Nk + c) : 0.f;
      dot += pv[i] * dv[i];
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
      const int c = lane + i * 32;
      if (c < Nk) ds[row * Nk + c] = __float2bfloat16_rn(pv[i] * (dv[i] - dot));
    }
  }
}

// bf16 scores, Nk a multiple of 256: a lane owns chunks of 8 consecutive keys (16-byte loads / stores); the scalar
// kernels above move 2 bytes per lane per instruction and reach only half of the HBM rate on [*, 512] rows
template <int CH>      // chunks of 8 keys per lane: Nk = 256 * CH
__global__ void __launch_bounds__(256)
softmax_fwd_v8_kernel(const __nv_bfloat16* __restrict__ s, const long long* __restrict__ key_len,
                      __nv_bfloat16* __restrict__ p, long long rows, int Nq, int groups_per_len) {
  constexpr int Nk = 256 * CH;
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp_global; row < rows; row += warps_total) {
    int len = Nk;
    if (key_len) {
      const long long l = key_len[(row / Nq) / groups_per_len];
      len = l < Nk ? (int)l : Nk;
    }
    float v[CH][8];
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      const int c0 = (lane + 32 * j) * 8;
      uint4 q = make_uint4(0, 0, 0, 0);
      if (c0 < len) q = __ldg(reinterpret_cast<const uint4*>(s + row * Nk + c0));
      const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&q);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        v[j][e] = (c0 + e < len) ? __bfloat162float(h[e]) : -INFINITY;
        m = fmaxf(m, v[j][e]);
      }
    }
    m = warp_max(m);
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < CH; ++j)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int c = (lane + 32 * j) * 8 + e;
        v[j][e] = (c < len) ? __expf(v[j][e] - m) : 0.f;
        sum += v[j][e];
      }
    const float inv = 1.f / warp_sum(sum);
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      uint4 q;
      __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
      for (int e = 0; e < 4; ++e) h2[e] = __floats2bfloat162_rn(v[j][2 * e] * inv, v[j][2 * e + 1] * inv);
      *reinterpret_cast<uint4*>(p + row * Nk + (lane + 32 * j) * 8) = q;
    }
  }
}
template <int CH>
__global__ void __launch_bounds__(256)
softmax_bwd_v8_kernel(const __nv_bfloat16* __restrict__ p, const __nv_bfloat16* __restrict__ dp,
                      __nv_bfloat16* __restrict__ ds, long long rows) {
  constexpr int Nk = 256 * CH;
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp_global; row < rows; row += warps_total) {
    float pv[CH][8], dv[CH][8];
    float dot = 0.f;
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      const uint4 qp = __ldg(reinterpret_cast<const uint4*>(p + row * Nk + (lane + 32 * j) * 8));
      const uint4 qd = __ldg(reinterpret_cast<const uint4*>(dp + row * Nk + (lane + 32 * j) * 8));
      const __nv_bfloat16* hp = reinterpret_cast<const __nv_bfloat16*>(&qp);
      const __nv_bfloat16* hd = reinterpret_cast<const __nv_bfloat16*>(&qd);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        pv[j][e] = __bfloat162float(hp[e]);
        dv[j][e] = __bfloat162float(hd[e]);
        dot += pv[j][e] * dv[j][e];
      }
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      uint4 q;
      __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
      for (int e = 0; e < 4; ++e)
        h2[e] = __floats2bfloat162_rn(pv[j][2 * e] * (dv[j][2 * e] - dot), pv[j][2 * e + 1] * (dv[j][2 * e + 1] - dot));
      *reinterpret_cast<uint4*>(ds + row * Nk + (lane + 32 * j) * 8) = q;
    }
  }
}

#define SM_DISPATCH(KERNEL, ST, ...)                                        \
  do {                                                                      \
    const int per = (Nk + 31) / 32;                                         \
    if (per <= 1) KERNEL<1, ST><<<grid, 256, 0, st>>>(__VA_ARGS__);         \
    else if (per <= 4) KERNEL<4, ST><<<grid, 256, 0, st>>>(__VA_ARGS__);    \
    else if (per <= 8) KERNEL<8, ST><<<grid, 256, 0, st>>>(__VA_ARGS__);    \
    else if (per <= 16) KERNEL<16, ST><<<grid, 256, 0, st>>>(__VA_ARGS__);  \
    else KERNEL<32, ST><<<grid, 256, 0, st>>>(__VA_ARGS__);                 \
  } while (0)

int softmax_fwd(const void* s, const long long* key_len, void* p, long long groups, int Nq, int Nk,
                int groups_per_len, int scores_bf16, cudaStream_t st) {
  const long long rows = groups * Nq;
  if (rows <= 0) return AS_OK;
  if (Nk <= 0 || Nk > 1024 || groups_per_len <= 0) return AS_ESIZE;
  if (!s || !p) return AS_EINVAL;
  const int grid = ln_grid(rows);
  const bool al16 = ((reinterpret_cast<uintptr_t>(s) | reinterpret_cast<uintptr_t>(p)) & 15) == 0;
  if (scores_bf16 && al16 && Nk == 512) softmax_fwd_v8_kernel<2><<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(s), key_len, reinterpret_cast<__nv_bfloat16*>(p), rows, Nq, groups_per_len);
  else if (scores_bf16 && al16 && Nk == 256) softmax_fwd_v8_kernel<1><<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(s), key_len, reinterpret_cast<__nv_bfloat16*>(p), rows, Nq, groups_per_len);
  else if (scores_bf16) SM_DISPATCH(softmax_fwd_kernel, __nv_bfloat16, reinterpret_cast<const __nv_bfloat16*>(s), key_len, reinterpret_cast<__nv_bfloat16*>(p), rows, Nq, Nk, groups_per_len);
  else SM_DISPATCH(softmax_fwd_kernel, float, reinterpret_cast<const float*>(s), key_len, reinterpret_cast<__nv_bfloat16*>(p), rows, Nq, Nk, groups_per_len);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int softmax_bwd(const void* p, const void* dp, void* ds, long long rows, int Nk, int dprobs_bf16, cudaStream_t st) {
  if (rows <= 0) return AS_OK;
  if (Nk <= 0 || Nk > 1024) return AS_ESIZE;
  if (!p || !dp || !ds) return AS_EINVAL;
  const int grid = ln_grid(rows);
  const bool al16 = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(dp) | reinterpret_cast<uintptr_t>(ds)) & 15) == 0;
  if (dprobs_bf16 && al16 && Nk == 512) softmax_bwd_v8_kernel<2><<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(p), reinterpret_cast<const __nv_bfloat16*>(dp), reinterpret_cast<__nv_bfloat16*>(ds), rows);
  else if (dprobs_bf16 && al16 && Nk == 256) softmax_bwd_v8_kernel<1><<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(p), reinterpret_cast<const __nv_bfloat16*>(dp), reinterpret_cast<__nv_bfloat16*>(ds), rows);
  else if (dprobs_bf16) SM_DISPATCH(softmax_bwd_kernel, __nv_bfloat16, reinterpret_cast<const __nv_bfloat16*>(p), reinterpret_cast<const __nv_bfloat16*>(dp), reinterpret_cast<__nv_bfloat16*>(ds), rows, Nk);
  else SM_DISPATCH(softmax_bwd_kernel, float, reinterpret_cast<const __nv_bfloat16*>(p), reinterpret_cast<const float*>(dp), reinterpret_cast<__nv_bfloat16*>(ds), rows, Nk);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------------------------------ entity features
// Builds the bf16 feature matrix the embedding GEMM consumes (entity_encoder.py:60-78): row = entity, 997 real
// columns + zero padding to `pitch`.  Field table: kind 0 one-hot (clamped to num-1), 1 binary (num bits, MSB first),
// 2 value.  Each field's data is a separate [rows] array with its own element type.
struct EntityFieldTable {
  const void* data[40];
  int kind[40];
  int num[40];
  int col[40];
  int dtype[40];   // 0 u8, 1 i8, 2 i16, 3 f16, 4 i64, 5 f32
  int n_fields;
};

__device__ __forceinline__ float field_value(const void* base, int dtype, long long i) {
  switch (dtype) {
    case 0: return (float)reinterpret_cast<const uint8_t*>(base)[i];
    case 1: return (float)reinterpret_cast<const int8_t*>(base)[i];
    case 2: return (float)reinterpret_cast<const int16_t*>(base)[i];
    case 3: return __half2float(reinterpret_cast<const __half*>(base)[i]);
    case 4: return (float)reinterpret_cast<const long long*>(base)[i];
    default: return reinterpret_cast<const float*>(base)[i];
  }
}

constexpr int kEfRows = 32;     // rows per CTA
__global__ void __launch_bounds__(256)
entity_features_kernel(const EntityFieldTable tab, __nv_bfloat16* __restrict__ out, long long rows, int pitch,
                       int* __restrict__ neg_flag) {
  // A CTA assembles kEfRows consecutive rows in shared memory (zeros, then the ones / values of every field; field
  // arrays are read 32 consecutive rows at a time = coalesced) and streams them out with 16-byte stores.
  extern __shared__ __align__(16) uint8_t ef_smem[];
  __nv_bfloat16* srow = reinterpret_cast<__nv_bfloat16*>(ef_smem);       // [kEfRows][pitch]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n16 = kEfRows * pitch / 8;
  for (long long row0 = (long long)blockIdx.x * kEfRows; row0 < rows; row0 += (long long)gridDim.x * kEfRows) {
    for (int i = threadIdx.x; i < n16; i += blockDim.x) reinterpret_cast<uint4*>(srow)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const long long row = row0 + lane;
    if (row < rows) {
      __nv_bfloat16* o = srow + lane * pitch;
      for (int f = warp; f < tab.n_fields; f += 8) {
        const float v = field_value(tab.data[f], tab.dtype[f], row);
        if (tab.kind[f] == 0) {
          int idx = (int)v;
          if (idx < 0) { atomicExch(neg_flag, 1); idx = 0; }
          if (idx > tab.num[f] - 1) idx = tab.num[f] - 1;
          o[tab.col[f] + idx] = __float2bfloat16_rn(1.f);
        } else if (tab.kind[f] == 1) {
          const int iv = (int)v;
          for (int b = 0; b < tab.num[f]; ++b)
            if ((iv >> (tab.num[f] - 1 - b)) & 1) o[tab.col[f] + b] = __float2bfloat16_rn(1.f);
        } else {
          o[tab.col[f]] = __float2bfloat16_rn(v);
        }
      }
    }
    __syncthreads();
    const long long left = rows - row0;
    const int nvalid = (int)((left < kEfRows ? left : kEfRows) * (pitch / 8));
    uint4* dst = reinterpret_cast<uint4*>(out + row0 * pitch);
    for (int i = threadIdx.x; i < nvalid; i += blockDim.x) dst[i] = reinterpret_cast<const uint4*>(srow)[i];
    __syncthreads();
  }
}

int entity_features(const void* const* data, const int* kind, const int* num, const int* col, const int* dtype,
                    int n_fields, void* out, long long rows, int pitch, int* neg_flag, cudaStream_t st) {
  if (rows <= 0) return AS_OK;
  if (n_fields <= 0 || n_fields > 40 || pitch % 8 != 0) return AS_ESIZE;
  if (!out || !neg_flag) return AS_EINVAL;
  EntityFieldTable tab;
  tab.n_fields = n_fields;
  for (int i = 0; i < n_fields; ++i) {
    tab.data[i] = data[i]; tab.kind[i] = kind[i]; tab.num[i] = num[i]; tab.col[i] = col[i]; tab.dtype[i] = dtype[i];
    if (!data[i]) return AS_EINVAL;
  }
  const int smem = kEfRows * pitch * 2;
  if (smem > 200 * 1024 || (reinterpret_cast<uintptr_t>(out) & 15)) return AS_ESIZE;
  static int attr_smem = 0;
  if (smem > attr_smem) {
    cudaError_t e = cudaFuncSetAttribute(entity_features_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    attr_smem = smem;
  }
  long long blocks = (rows + kEfRows - 1) / kEfRows;
  const int grid = (int)(blocks < 148LL * 12 ? blocks : 148LL * 12);
  entity_features_kernel<<<grid, 256, smem, st>>>(tab, reinterpret_cast<__nv_bfloat16*>(out), rows, pitch, neg_flag);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------------------------------ masked mean over entities
// out[r,c] = sum_{n < len[r]} x[r,n,c] / len[r]   (entity_encoder.py:83-85).  Block = (row r, 32-entity slab) partial
// sums would need a second pass; rows are short (<= 512 x 1 KB), so one block per row with 4 entity lanes per channel
// quad keeps every load a coalesced float4 and finishes in shared memory.
__global__ void __launch_bounds__(256)
masked_mean_fwd_kernel(const float* __restrict__ x, const long long* __restrict__ len, float* __restrict__ out, int N, int C) {
  const long long r = blockIdx.x;
  const int c4n = C >> 2;                              // float4 columns
  const int cq = threadIdx.x % c4n, lane_n = threadIdx.x / c4n, n_lanes = blockDim.x / c4n;
  long long l = len[r];
  l = l < 0 ? 0 : (l > N ? N : l);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (lane_n < n_lanes) {
    const float4* xr = reinterpret_cast<const float4*>(x + r * (long long)N * C);
    for (int n = lane_n; n < (int)l; n += n_lanes) {
      const float4 v = __ldg(xr + (long long)n * c4n + cq);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  extern __shared__ float4 s_mm[];                     // [n_lanes][c4n]
  if (lane_n < n_lanes) s_mm[lane_n * c4n + cq] = acc;
  __syncthreads();
  if (threadIdx.x < c4n) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < n_lanes; ++k) { const float4 v = s_mm[k * c4n + threadIdx.x]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
    const float inv = 1.f / (float)len[r];             // the reference divides by entity_num as given
    reinterpret_cast<float4*>(out + r * C)[threadIdx.x] = make_float4(t.x * inv, t.y * inv, t.z * inv, t.w * inv);
  }
}
// dx[r,n,c] = n < len[r] ? dout[r,c] / len[r] : 0
__global__ void __launch_bounds__(256)
masked_mean_bwd_kernel(const float* __restrict__ dout, const long long* __restrict__ len, float* __restrict__ dx, long long R,
                       int N, int C) {
  const int c4n = C >> 2;
  const long long total = R * (long long)N * c4n;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long per_row = (long long)N * c4n;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long r = i / per_row;
    const int rem = (int)(i - r * per_row);
    const int n = rem / c4n, cq = rem - n * c4n;
    const long long l = len[r];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if ((long long)n < l) {
      const float inv = 1.f / (float)l;
      const float4 g = __ldg(reinterpret_cast<const float4*>(dout + r * C) + cq);
      v = make_float4(g.x * inv, g.y * inv, g.z * inv, g.w * inv);
    }
    reinterpret_cast<float4*>(dx)[i] = v;
  }
}
int masked_mean_fwd(const float* x, const long long* len, float* out, long long R, int N, int C, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (C <= 0 || (C % 4) != 0 || C / 4 > 256 || N <= 0) return AS_ESIZE;
  if (!x || !len || !out) return AS_EINVAL;
  const int c4n = C / 4, n_lanes = 256 / c4n;
  masked_mean_fwd_kernel<<<(unsigned)R, 256, n_lanes * c4n * 16, st>>>(x, len, out, N, C);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int masked_mean_bwd(const float* dout, const long long* len, float* dx, long long R, int N, int C, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (C <= 0 || (C % 4) != 0 || N <= 0) return AS_ESIZE;
  if (!dout || !len || !dx) return AS_EINVAL;
  long long blocks = (R * (long long)N * (C / 4) + 255) / 256;
  masked_mean_bwd_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, st>>>(dout, len, dx, R, N, C);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
