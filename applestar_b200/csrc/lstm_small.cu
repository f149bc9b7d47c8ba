// LayerNorm-LSTM recurrence for hidden size 32 (the SelectedUnits pointer LSTM, action_arg_head.py / lstm.py:138-167):
// one WARP per trajectory, lane = hidden unit.  W_hh (128 x 32) sits in shared memory as fp32; h is exchanged with
// shuffles, both LayerNorms are warp reductions, so a step has no block-level synchronisation at all.  The generic
// kernel in lstm.cu spends a 512-thread CTA and ~10 barriers per step on the same 4 K multiply-adds.
#include "common.cuh"

namespace as {

constexpr int kSH = 32, kSG = 128;

__device__ __forceinline__ float sgm(float x) { return 1.f / (1.f + expf(-x)); }

__global__ void __launch_bounds__(256)
lnlstm_small_fwd_kernel(const float* __restrict__ gi, const float* __restrict__ h0, const float* __restrict__ c0,
                        const __nv_bfloat16* __restrict__ whh, const float* __restrict__ gh_w, const float* __restrict__ gh_b,
                        const float* __restrict__ gc_w, const float* __restrict__ gc_b, float* __restrict__ y,
                        float* __restrict__ c_all, float* __restrict__ u_all, float* __restrict__ z_all,
                        float* __restrict__ stats, int T, int B) {
  __shared__ float sWt[kSH * kSG];                     // [k][j] = W_hh[j][k]
  for (int i = threadIdx.x; i < kSH * kSG; i += blockDim.x) {
    const int j = i / kSH, k = i - j * kSH;
    sWt[k * kSG + j] = __bfloat162float(whh[i]);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (b >= B) return;
  float ghw[4], ghb[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) { ghw[g] = gh_w[g * kSH + lane]; ghb[g] = gh_b[g * kSH + lane]; }
  const float gcw = gc_w[lane], gcb = gc_b[lane];
  float h = h0[(long long)b * kSH + lane], c = c0[(long long)b * kSH + lane];
  for (int t = 0; t < T; ++t) {
    const long long tb = (long long)t * B + b;
    float gir[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) gir[g] = __ldg(gi + tb * kSG + g * kSH + lane);
    float u[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
    for (int k = 0; k < kSH; ++k) {
      const float hk = __shfl_sync(0xffffffffu, h, k);
#pragma unroll
      for (int g = 0; g < 4; ++g) u[g] += sWt[k * kSG + g * kSH + lane] * hk;
    }
    const float mean_h = warp_sum(u[0] + u[1] + u[2] + u[3]) * (1.f / kSG);
    float q = 0.f;
#pragma unroll
    for (int g = 0; g < 4; ++g) { const float d = u[g] - mean_h; q += d * d; }
    const float rstd_h = rsqrtf(warp_sum(q) * (1.f / kSG) + 1e-5f);
    float a[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (u_all) u_all[tb * kSG + g * kSH + lane] = u[g];
      a[g] = gir[g] + (u[g] - mean_h) * rstd_h * ghw[g] + ghb[g];
    }
    const float z = sgm(a[1]) * c + sgm(a[0]) * tanhf(a[2]);
    const float mean_c = warp_sum(z) * (1.f / kSH);
    const float dz = z - mean_c;
    const float rstd_c = rsqrtf(warp_sum(dz * dz) * (1.f / kSH) + 1e-5f);
    c = dz * rstd_c * gcw + gcb;
    h = sgm(a[3]) * tanhf(c);
    y[tb * kSH + lane] = h;
    c_all[tb * kSH + lane] = c;
    if (z_all) z_all[tb * kSH + lane] = z;
    if (stats && lane == 0) { stats[tb * 4] = mean_h; stats[tb * 4 + 1] = rstd_h; stats[tb * 4 + 2] = mean_c; stats[tb * 4 + 3] = rstd_c; }
  }
}

__global__ void __launch_bounds__(256)
lnlstm_small_bwd_kernel(const float* __restrict__ gi, const float* __restrict__ c0, const float* __restrict__ c_all,
                        const float* __restrict__ u_all, const float* __restrict__ z_all, const float* __restrict__ stats,
                        const __nv_bfloat16* __restrict__ whh_t, const float* __restrict__ gh_w, const float* __restrict__ gh_b,
                        const float* __restrict__ gc_w, const float* __restrict__ gc_b, const float* __restrict__ dy,
                        const float* __restrict__ dhT, const float* __restrict__ dcT, float* __restrict__ dgi,
                        float* __restrict__ du_out, float* __restrict__ dh0, float* __restrict__ dc0,
                        float* __restrict__ dparam, int T, int B) {
  __shared__ float sW[kSG * kSH];                      // [j][k] = W_hh[j][k] = whh_t[k][j]
  for (int i = threadIdx.x; i < kSG * kSH; i += blockDim.x) {
    const int k = i / kSG, j = i - k * kSG;
    sW[j * kSH + k] = __bfloat162float(whh_t[i]);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (b >= B) return;
  float ghw[4], ghb[4], pgw[4] = {0.f, 0.f, 0.f, 0.f}, pgb[4] = {0.f, 0.f, 0.f, 0.f}, pcw = 0.f, pcb = 0.f;
#pragma unroll
  for (int g = 0; g < 4; ++g) { ghw[g] = gh_w[g * kSH + lane]; ghb[g] = gh_b[g * kSH + lane]; }
  const float gcw = gc_w[lane], gcb = gc_b[lane];
  float dh = dhT ? dhT[(long long)b * kSH + lane] : 0.f, dc = dcT ? dcT[(long long)b * kSH + lane] : 0.f;
  for (int t = T - 1; t >= 0; --t) {
    const long long tb = (long long)t * B + b;
    const float mean_h = __ldg(stats + tb * 4), rstd_h = __ldg(stats + tb * 4 + 1);
    const float mean_c = __ldg(stats + tb * 4 + 2), rstd_c = __ldg(stats + tb * 4 + 3);
    float uh[4], a[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      uh[g] = (__ldg(u_all + tb * kSG + g * kSH + lane) - mean_h) * rstd_h;
      a[g] = __ldg(gi + tb * kSG + g * kSH + lane) + uh[g] * ghw[g] + ghb[g];
    }
    const float i_ = sgm(a[0]), f_ = sgm(a[1]), g_ = tanhf(a[2]), o_ = sgm(a[3]);
    const float zhat = (__ldg(z_all + tb * kSH + lane) - mean_c) * rstd_c;
    const float tc = tanhf(zhat * gcw + gcb);
    const float dht = dh + (dy ? __ldg(dy + tb * kSH + lane) : 0.f);
    float da[4];
    da[3] = dht * tc * o_ * (1.f - o_);
    const float dct = dc + dht * o_ * (1.f - tc * tc);
    pcw += dct * zhat;
    pcb += dct;
    const float dzh = dct * gcw;
    const float m1 = warp_sum(dzh) * (1.f / kSH), m2 = warp_sum(dzh * zhat) * (1.f / kSH);
    const float dz = rstd_c * (dzh - m1 - zhat * m2);
    const float cprev = t > 0 ? __ldg(c_all + ((long long)(t - 1) * B + b) * kSH + lane) : __ldg(c0 + (long long)b * kSH + lane);
    da[0] = dz * g_ * i_ * (1.f - i_);
    da[1] = dz * cprev * f_ * (1.f - f_);
    da[2] = dz * i_ * (1.f - g_ * g_);
    dc = dz * f_;
    float a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      dgi[tb * kSG + g * kSH + lane] = da[g];
      pgw[g] += da[g] * uh[g];
      pgb[g] += da[g];
      const float d = da[g] * ghw[g];
      a1 += d; a2 += d * uh[g];
    }
    a1 = warp_sum(a1) * (1.f / kSG); a2 = warp_sum(a2) * (1.f / kSG);
    float du[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      du[g] = rstd_h * (da[g] * ghw[g] - a1 - uh[g] * a2);
      du_out[tb * kSG + g * kSH + lane] = du[g];
    }
    // dh_{t-1}[k = lane] = sum_j W[j][k] du[j]
    float acc = 0.f;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
#pragma unroll 8
      for (int jj = 0; jj < kSH; ++jj) acc += sW[(g * kSH + jj) * kSH + lane] * __shfl_sync(0xffffffffu, du[g], jj);
    }
    dh = acc;
  }
  if (dh0) dh0[(long long)b * kSH + lane] = dh;
  if (dc0) dc0[(long long)b * kSH + lane] = dc;
  float* dp = dparam + (long long)b * (2 * kSG + 2 * kSH);
#pragma unroll
  for (int g = 0; g < 4; ++g) { dp[g * kSH + lane] = pgw[g]; dp[kSG + g * kSH + lane] = pgb[g]; }
  dp[2 * kSG + lane] = pcw;
  dp[2 * kSG + kSH + lane] = pcb;
}

int lnlstm_small_fwd(const float* gi, const float* h0, const float* c0, const void* whh, const float* gh_w,
                     const float* gh_b, const float* gc_w, const float* gc_b, float* y, float* c_all, float* u_all,
                     float* z_all, float* stats, int T, int B, cudaStream_t st) {
  lnlstm_small_fwd_kernel<<<(B + 7) / 8, 256, 0, st>>>(gi, h0, c0, reinterpret_cast<const __nv_bfloat16*>(whh), gh_w, gh_b, gc_w,
                                                     gc_b, y, c_all, u_all, z_all, stats, T, B);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int lnlstm_small_bwd(const float* gi, const float* c0, const float* c_all, const float* u_all, const float* z_all,
                     const float* stats, const void* whh_t, const float* gh_w, const float* gh_b, const float* gc_w,
                     const float* gc_b, const float* dy, const float* dhT, const float* dcT, float* dgi, float* du,
                     float* dh0, float* dc0, float* dparam, int T, int B, cudaStream_t st) {
  lnlstm_small_bwd_kernel<<<(B + 7) / 8, 256, 0, st>>>(gi, c0, c_all, u_all, z_all, stats,
                                                     reinterpret_cast<const __nv_bfloat16*>(whh_t), gh_w, gh_b, gc_w, gc_b, dy,
                                                     dhT, dcT, dgi, du, dh0, dc0, dparam, T, B);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
