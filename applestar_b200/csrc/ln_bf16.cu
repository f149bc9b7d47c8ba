// bf16 entity stream of the post-LN Transformer (reference: model/module_utils.py:133-139, y = LN(x + f(x))).
// The three Transformer layers are HBM-bound around their LayerNorms; with the residual stream, the GEMM outputs and the
// gradients of both held in bf16 (what the reference's own autocast path holds) the LayerNorm kernels move 0.43x
// (forward) / 0.6x (backward) of the bytes of the fp32-stream version in nn_kernels.cu.  Statistics, the normalisation
// and all reductions stay fp32.  C = 256: a lane owns 8 contiguous channels = one 16-byte load per tensor and row.
#include "common.cuh"

namespace as {

int layer_norm_param_reduce3(const float* partial, float* dgamma, float* dbeta, float* dbias, int C, int chunks, cudaStream_t st);   // nn_kernels.cu

__device__ __forceinline__ void bf8_to_f(const uint4& u, float (&f)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) { f[2 * k] = __uint_as_float(w[k] << 16); f[2 * k + 1] = __uint_as_float(w[k] & 0xffff0000u); }
}
__device__ __forceinline__ uint4 f_to_bf8(const float (&f)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * k], f[2 * k + 1]);
    w[k] = *reinterpret_cast<uint32_t*>(&h);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

constexpr int kLn16C = 256;

// y16 = LN(z16 + res16) * gamma + beta (+ReLU); stats[row] = (mean, rstd)
__global__ void __launch_bounds__(256)
ln256_bf16_fwd_kernel(const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* __restrict__ res, const float* __restrict__ gamma,
                      const float* __restrict__ beta, __nv_bfloat16* __restrict__ y, float* __restrict__ stats, long long rows,
                      int relu) {
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  float gm[8], bt[8];
  {
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma) + lane * 2), g1 = __ldg(reinterpret_cast<const float4*>(gamma) + lane * 2 + 1);
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta) + lane * 2), b1 = __ldg(reinterpret_cast<const float4*>(beta) + lane * 2 + 1);
    gm[0] = g0.x; gm[1] = g0.y; gm[2] = g0.z; gm[3] = g0.w; gm[4] = g1.x; gm[5] = g1.y; gm[6] = g1.z; gm[7] = g1.w;
    bt[0] = b0.x; bt[1] = b0.y; bt[2] = b0.z; bt[3] = b0.w; bt[4] = b1.x; bt[5] = b1.y; bt[6] = b1.z; bt[7] = b1.w;
  }
  for (long long row = warp_global; row < rows; row += warps_total) {
    float v[8], rr[8];
    bf8_to_f(__ldg(reinterpret_cast<const uint4*>(z + row * kLn16C) + lane), v);
    if (res) {
      bf8_to_f(__ldg(reinterpret_cast<const uint4*>(res + row * kLn16C) + lane), rr);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += rr[j];
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += v[j];
    const float mean = warp_sum(s) * (1.f / kLn16C);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { const float d = v[j] - mean; q += d * d; }
    const float rstd = rsqrtf(warp_sum(q) * (1.f / kLn16C) + 1e-5f);
    if (lane == 0 && stats) { stats[row * 2] = mean; stats[row * 2 + 1] = rstd; }
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      o[j] = (v[j] - mean) * rstd * gm[j] + bt[j];
      if (relu) o[j] = fmaxf(o[j], 0.f);
    }
    reinterpret_cast<uint4*>(y + row * kLn16C)[lane] = f_to_bf8(o);
  }
}

// Backward of y = LN(res + z) with z the output of a linear layer (optional ReLU: mask_z) whose backward follows:
// dres16 (residual path, may be NULL), dz16 = dx * (z > 0 | 1), and per-block partial rows [3][256]: d gamma, d beta,
// d bias(z).  g = dyA + dyB (dyB may be NULL): the two consumers of y (residual add, next GEMM) hand their gradients over
// separately and they are summed here instead of by a separate pass.
__global__ void __launch_bounds__(256)
ln256_bf16_bwd_kernel(const __nv_bfloat16* __restrict__ dya, const __nv_bfloat16* __restrict__ dyb,
                      const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* __restrict__ res, const float* __restrict__ gamma,
                      const float* __restrict__ beta, const float* __restrict__ stats, __nv_bfloat16* __restrict__ dres,
                      __nv_bfloat16* __restrict__ dz, int mask_z, float* __restrict__ partial, long long rows, int relu) {
  const int lane = threadIdx.x & 31;
  const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
  float gm[8], bt[8], pg[8], pb[8], pz[8];
  {
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma) + lane * 2), g1 = __ldg(reinterpret_cast<const float4*>(gamma) + lane * 2 + 1);
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta) + lane * 2), b1 = __ldg(reinterpret_cast<const float4*>(beta) + lane * 2 + 1);
    gm[0] = g0.x; gm[1] = g0.y; gm[2] = g0.z; gm[3] = g0.w; gm[4] = g1.x; gm[5] = g1.y; gm[6] = g1.z; gm[7] = g1.w;
    bt[0] = b0.x; bt[1] = b0.y; bt[2] = b0.z; bt[3] = b0.w; bt[4] = b1.x; bt[5] = b1.y; bt[6] = b1.z; bt[7] = b1.w;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) { pg[j] = 0.f; pb[j] = 0.f; pz[j] = 0.f; }
  for (long long row = warp_global; row < rows; row += warps_total) {
    const float mean = stats[row * 2], rstd = stats[row * 2 + 1];
    float zv[8], t[8], g[8], xh[8], b[8];
    bf8_to_f(__ldg(reinterpret_cast<const uint4*>(z + row * kLn16C) + lane), zv);
    bf8_to_f(__ldg(reinterpret_cast<const uint4*>(dya + row * kLn16C) + lane), g);
    if (res) {
      bf8_to_f(__ldg(reinterpret_cast<const uint4*>(res + row * kLn16C) + lane), t);
#pragma unroll
      for (int j = 0; j < 8; ++j) t[j] += zv[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) t[j] = zv[j];
    }
    if (dyb) {
      float g2[8];
      bf8_to_f(__ldg(reinterpret_cast<const uint4*>(dyb + row * kLn16C) + lane), g2);
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] += g2[j];
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      xh[j] = (t[j] - mean) * rstd;
      if (relu && (xh[j] * gm[j] + bt[j]) <= 0.f) g[j] = 0.f;
      pg[j] += g[j] * xh[j];
      pb[j] += g[j];
      b[j] = g[j] * gm[j];
      s1 += b[j];
      s2 += xh[j] * b[j];
    }
    const float m1 = warp_sum(s1) * (1.f / kLn16C), m2 = warp_sum(s2) * (1.f / kLn16C);
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = rstd * (b[j] - m1 - xh[j] * m2);
    if (dres) reinterpret_cast<uint4*>(dres + row * kLn16C)[lane] = f_to_bf8(o);
    if (mask_z) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = zv[j] > 0.f ? o[j] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) pz[j] += o[j];
    reinterpret_cast<uint4*>(dz + row * kLn16C)[lane] = f_to_bf8(o);
  }
  __shared__ float sp[8][3][kLn16C];
  const int w = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sp[w][0][lane * 8 + j] = pg[j];
    sp[w][1][lane * 8 + j] = pb[j];
    sp[w][2][lane * 8 + j] = pz[j];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * kLn16C; i += blockDim.x) {
    const int which = i / kLn16C, c = i - which * kLn16C;
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sp[k][which][c];
    partial[((long long)blockIdx.x * 3 + which) * kLn16C + c] = t;
  }
}

static int ln16_grid(long long rows) {
  long long blocks = (rows + 7) / 8;
  const long long cap = 148LL * 8;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

int ln256_bf16_fwd(const void* z, const void* res, const float* gamma, const float* beta, void* y, float* stats, long long rows,
                   int relu, cudaStream_t st) {
  if (rows <= 0) return AS_OK;
  if (!z || !gamma || !beta || !y) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(res) | reinterpret_cast<uintptr_t>(y) |
       reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta)) & 15) return AS_EALIGN;
  ln256_bf16_fwd_kernel<<<ln16_grid(rows), 256, 0, st>>>((const __nv_bfloat16*)z, (const __nv_bfloat16*)res, gamma, beta,
                                                        (__nv_bfloat16*)y, stats, rows, relu);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

long long ln256_bf16_bwd_scratch_floats() { return 148LL * 8 * 3 * kLn16C; }

int ln256_bf16_bwd(const void* dya, const void* dyb, const void* z, const void* res, const float* gamma, const float* beta,
                   const float* stats, void* dres, void* dz, int mask_z, float* dgamma, float* dbeta, float* dbias, float* scratch,
                   long long rows, int relu, cudaStream_t st) {
  if (rows <= 0) return AS_OK;
  if (!dya || !z || !gamma || !beta || !stats || !dz || !dgamma || !dbeta || !scratch) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(dya) | reinterpret_cast<uintptr_t>(dyb) | reinterpret_cast<uintptr_t>(z) |
       reinterpret_cast<uintptr_t>(res) | reinterpret_cast<uintptr_t>(dres) | reinterpret_cast<uintptr_t>(dz) |
       reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta)) & 15) return AS_EALIGN;
  const int grid = ln16_grid(rows);
  ln256_bf16_bwd_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)dya, (const __nv_bfloat16*)dyb, (const __nv_bfloat16*)z,
                                              (const __nv_bfloat16*)res, gamma, beta, stats, (__nv_bfloat16*)dres, (__nv_bfloat16*)dz,
                                              mask_z, scratch, rows, relu);
  AS_LAUNCH_CHECK();
  return layer_norm_param_reduce3(scratch, dgamma, dbeta, dbias, kLn16C, grid, st);
}

// masked mean over the valid entities of a bf16 stream [R, N, 256] (entity_encoder.py:83-85) and its backward (bf16 dx)
__global__ void __launch_bounds__(256)
masked_mean256_bf16_fwd_kernel(const __nv_bfloat16* __restrict__ x, const long long* __restrict__ len, float* __restrict__ out, int N) {
  const long long r = blockIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  long long l = len[r];
  l = l < 0 ? 0 : (l > N ? N : l);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int n = w; n < (int)l; n += 8) {
    float v[8];
    bf8_to_f(__ldg(reinterpret_cast<const uint4*>(x + (r * N + n) * kLn16C) + lane), v);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += v[j];
  }
  __shared__ float sm[8][kLn16C];
#pragma unroll
  for (int j = 0; j < 8; ++j) sm[w][lane * 8 + j] = acc[j];
  __syncthreads();
  const int c = threadIdx.x;
  float t = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) t += sm[k][c];
  out[r * kLn16C + c] = t / (float)len[r];                                   // the reference divides by entity_num as given
}
__global__ void __launch_bounds__(256)
masked_mean256_bf16_bwd_kernel(const float* __restrict__ dout, const long long* __restrict__ len, __nv_bfloat16* __restrict__ dx,
                               long long R, int N) {
  const long long total = R * (long long)N * 32;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c8 = (int)(i & 31);
    const long long rn = i >> 5;
    const long long r = rn / N;
    const int n = (int)(rn - r * N);
    const long long l = len[r];
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = 0.f;
    if ((long long)n < l) {
      const float inv = 1.f / (float)l;
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(dout + r * kLn16C) + c8 * 2);
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(dout + r * kLn16C) + c8 * 2 + 1);
      o[0] = g0.x * inv; o[1] = g0.y * inv; o[2] = g0.z * inv; o[3] = g0.w * inv;
      o[4] = g1.x * inv; o[5] = g1.y * inv; o[6] = g1.z * inv; o[7] = g1.w * inv;
    }
    reinterpret_cast<uint4*>(dx)[i] = f_to_bf8(o);
  }
}
int masked_mean256_bf16_fwd(const void* x, const long long* len, float* out, long long R, int N, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!x || !len || !out || N <= 0) return AS_EINVAL;
  masked_mean256_bf16_fwd_kernel<<<(unsigned)R, 256, 0, st>>>((const __nv_bfloat16*)x, len, out, N);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int masked_mean256_bf16_bwd(const float* dout, const long long* len, void* dx, long long R, int N, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!dout || !len || !dx || N <= 0) return AS_EINVAL;
  long long blocks = (R * (long long)N * 32 + 255) / 256;
  masked_mean256_bf16_bwd_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, st>>>(dout, len, (__nv_bfloat16*)dx, R, N);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
