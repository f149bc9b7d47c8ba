// LayerNorm-LSTM recurrence (reference: distar/agent/default/model/lstm.py:138-167) as one persistent kernel per
// layer and direction of time.  The input projection LN_i(x W_ih^T) for ALL timesteps is a single tcgen05 GEMM +
// LayerNorm launch outside; this kernel runs the strictly sequential part:
//     u = W_hh h_{t-1};  a = gi_t + LN_h(u);  (i,f,g,o) = a;  z = s(f) c_{t-1} + s(i) tanh(g);
//     c_t = LN_c(z);  h_t = s(o) tanh(c_t)
// Trajectories are independent, so one CTA owns one batch row for the whole unroll: no inter-CTA synchronisation.
// W_hh (bf16, 1.2 MB for the core) streams from L2 every step; h, c and all LayerNorm statistics stay on chip.
// The backward kernel is the reverse-time twin; it emits d gi_t, d u_t (for dW_hh = sum_t du_t h_{t-1}^T, done as a
// tcgen05 GEMM outside) and per-row partial LayerNorm parameter gradients.
#include "common.cuh"

namespace as {

constexpr int kLstmThreads = 512;
constexpr int kLstmWarps = kLstmThreads / 32;

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// block-wide sum of two values (all threads get the result)
__device__ __forceinline__ void block_sum2(float& a, float& b, float* red) {
  a = warp_sum(a); b = warp_sum(b);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) { red[w] = a; red[kLstmWarps + w] = b; }
  __syncthreads();
  float sa = 0.f, sb = 0.f;
#pragma unroll
  for (int i = 0; i < kLstmWarps; ++i) { sa += red[i]; sb += red[kLstmWarps + i]; }
  a = sa; b = sb;
}

// out[j] = sum_k W[j][k] * v[k], W bf16 row-major [rows][cols] (cols % 8 == 0), v fp32 in shared memory.
// One warp per group of kRowsPerIter (2-8) output rows: lanes split each row in 16-byte chunks (coalesced); the loads of
// all rows of the group are issued before any is consumed, so ~kRowsPerIter * cols / 256 L2 requests are in flight
// per lane instead of one (the recurrence is latency-bound: W_hh streams from L2 every step).

__device__ __forceinline__ float dot8(const uint4& q, const float* vv) {
  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&q);
  float acc = 0.f;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 f = __bfloat1622float2(h2[e]);
    acc += f.x * vv[2 * e] + f.y * vv[2 * e + 1];
  }
  return acc;
}

// A lane always multiplies the same 8-element slices of v (chunk index = lane + 32 c), whatever the row: they are
// read from shared memory ONCE per mat-vec into registers (a per-row read would be an 8-way bank conflict).
template <int CHUNKS_PER_LANE, int kRowsPerIter>
__device__ __forceinline__ void matvec_rows(const __nv_bfloat16* __restrict__ Wm, const float* v, float* out, int rows,
                                            int cols) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chunks = cols >> 3;
  float vr[CHUNKS_PER_LANE][8];
#pragma unroll
  for (int c = 0; c < CHUNKS_PER_LANE; ++c) {
    const int ch = lane + c * 32;
#pragma unroll
    for (int e = 0; e < 8; ++e) vr[c][e] = (ch < chunks) ? v[ch * 8 + e] : 0.f;
  }
  for (int j0 = warp * kRowsPerIter; j0 < rows; j0 += kLstmWarps * kRowsPerIter) {
    uint4 q[kRowsPerIter][CHUNKS_PER_LANE];
#pragma unroll
    for (int r = 0; r < kRowsPerIter; ++r) {
      const uint4* wr = reinterpret_cast<const uint4*>(Wm + (long long)(j0 + r) * cols);
#pragma unroll
      for (int c = 0; c < CHUNKS_PER_LANE; ++c) {
        const int ch = lane + c * 32;
        q[r][c] = (j0 + r < rows && ch < chunks) ? __ldg(wr + ch) : make_uint4(0, 0, 0, 0);
      }
    }
#pragma unroll
    for (int r = 0; r < kRowsPerIter; ++r) {
      float acc = 0.f;
#pragma unroll
      for (int c = 0; c < CHUNKS_PER_LANE; ++c) acc += dot8(q[r][c], vr[c]);
      acc = warp_sum(acc);
      if (lane == 0 && j0 + r < rows) out[j0 + r] = acc;
    }
  }
}

__device__ __forceinline__ void matvec_bf16(const __nv_bfloat16* __restrict__ Wm, const float* v, float* out, int rows,
                                            int cols) {
  const int per_lane = ((cols >> 3) + 31) / 32;
  if (per_lane <= 1) matvec_rows<1, 8>(Wm, v, out, rows, cols);
  else if (per_lane <= 2) matvec_rows<2, 6>(Wm, v, out, rows, cols);
  else if (per_lane <= 6) matvec_rows<6, 2>(Wm, v, out, rows, cols);
  else {
    // generic (wide rows): one row at a time
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunks = cols >> 3;
    for (int j = warp; j < rows; j += kLstmWarps) {
      const uint4* wr = reinterpret_cast<const uint4*>(Wm + (long long)j * cols);
      float acc = 0.f;
      for (int ch = lane; ch < chunks; ch += 32) acc += dot8(__ldg(wr + ch), v + ch * 8);
      acc = warp_sum(acc);
      if (lane == 0) out[j] = acc;
    }
  }
}

struct LstmFwd {
  const float* gi;            // [T,B,4H]  LN_i(x W_ih^T)
  const float* h0; const float* c0;   // [B,H]
  const __nv_bfloat16* whh;   // [4H,H]
  const float* gh_w; const float* gh_b;   // LN_h affine [4H]
  const float* gc_w; const float* gc_b;   // LN_c affine [H]
  float* y;                   // [T,B,H] outputs h_t
  float* c_all;               // [T,B,H] cell states c_t
  float* u_all;               // [T,B,4H] saved W_hh h_{t-1} (pre-LN)   (may be NULL: inference)
  float* z_all;               // [T,B,H]  saved pre-LN cell            (may be NULL)
  float* stats;               // [T,B,4]  mean_h, rstd_h, mean_c, rstd_c (may be NULL)
  int T, B, H;
};

__global__ void __launch_bounds__(kLstmThreads) lnlstm_fwd_kernel(const LstmFwd p) {
  extern __shared__ float sm[];
  const int H = p.H, G = 4 * p.H, b = blockIdx.x;
  float* s_h = sm;            // [H]
  float* s_c = s_h + H;       // [H]
  float* s_u = s_c + H;       // [4H]
  float* red = s_u + G;       // [2*warps]
  for (int k = threadIdx.x; k < H; k += kLstmThreads) { s_h[k] = p.h0[(long long)b * H + k]; s_c[k] = p.c0[(long long)b * H + k]; }
  __syncthreads();
  for (int t = 0; t < p.T; ++t) {
    const long long tb = (long long)t * p.B + b;
    matvec_bf16(p.whh, s_h, s_u, G, H);
    __syncthreads();
    // LN_h statistics over 4H
    float s1 = 0.f, s2 = 0.f;
    for (int j = threadIdx.x; j < G; j += kLstmThreads) s1 += s_u[j];
    float dummy = 0.f;
    block_sum2(s1, dummy, red);
    const float mean_h = s1 / (float)G;
    for (int j = threadIdx.x; j < G; j += kLstmThreads) { const float d = s_u[j] - mean_h; s2 += d * d; }
    dummy = 0.f;
    block_sum2(s2, dummy, red);
    const float rstd_h = rsqrtf(s2 / (float)G + 1e-5f);
    // gates -> a (in place in s_u), save u
    const float* gi = p.gi + tb * G;
    for (int j = threadIdx.x; j < G; j += kLstmThreads) {
      const float u = s_u[j];
      if (p.u_all) p.u_all[tb * G + j] = u;
      s_u[j] = __ldg(gi + j) + (u - mean_h) * rstd_h * __ldg(p.gh_w + j) + __ldg(p.gh_b + j);
    }
    __syncthreads();
    // z = s(f) c + s(i) tanh(g); LN_c
    float zloc[2] = {0.f, 0.f};   // H <= 2 * kLstmThreads
    float z1 = 0.f;
    int n = 0;
    for (int k = threadIdx.x; k < H; k += kLstmThreads, ++n) {
      const float z = sigmoidf_(s_u[H + k]) * s_c[k] + sigmoidf_(s_u[k]) * tanhf(s_u[2 * H + k]);
      zloc[n] = z; z1 += z;
      if (p.z_all) p.z_all[tb * H + k] = z;
    }
    dummy = 0.f;
    block_sum2(z1, dummy, red);
    const float mean_c = z1 / (float)H;
    float z2 = 0.f;
    n = 0;
    for (int k = threadIdx.x; k < H; k += kLstmThreads, ++n) { const float d = zloc[n] - mean_c; z2 += d * d; }
    dummy = 0.f;
    block_sum2(z2, dummy, red);
    const float rstd_c = rsqrtf(z2 / (float)H + 1e-5f);
    n = 0;
    for (int k = threadIdx.x; k < H; k += kLstmThreads, ++n) {
      const float c = (zloc[n] - mean_c) * rstd_c * __ldg(p.gc_w + k) + __ldg(p.gc_b + k);
      const float h = sigmoidf_(s_u[3 * H + k]) * tanhf(c);
      s_c[k] = c; s_h[k] = h;
      p.y[tb * H + k] = h;
      p.c_all[tb * H + k] = c;
    }
    if (threadIdx.x == 0 && p.stats) {
      p.stats[tb * 4 + 0] = mean_h; p.stats[tb * 4 + 1] = rstd_h; p.stats[tb * 4 + 2] = mean_c; p.stats[tb * 4 + 3] = rstd_c;
    }
    __syncthreads();
  }
}

struct LstmBwd {
  const float* gi; const float* h0; const float* c0;
  const float* y; const float* c_all; const float* u_all; const float* z_all; const float* stats;
  const __nv_bfloat16* whh_t;   // [H,4H] = W_hh^T
  const float* gh_w; const float* gh_b; const float* gc_w; const float* gc_b;
  const float* dy;              // [T,B,H] grad of outputs (may be NULL)
  const float* dhT; const float* dcT;   // [B,H] grads of the final state (may be NULL)
  float* dgi;                   // [T,B,4H]
  float* du;                    // [T,B,4H]
  float* dh0; float* dc0;       // [B,H]
  float* dparam;                // [B, 2*4H + 2*H]: d gh_w, d gh_b, d gc_w, d gc_b per batch row
  int T, B, H;
};

__global__ void __launch_bounds__(kLstmThreads) lnlstm_bwd_kernel(const LstmBwd p) {
  extern __shared__ float sm[];
  const int H = p.H, G = 4 * p.H, b = blockIdx.x;
  float* s_dh = sm;             // [H]   recurrent dh
  float* s_dc = s_dh + H;       // [H]   recurrent dc
  float* s_da = s_dc + H;       // [4H]  d gates pre-activation, then d u
  float* s_pg = s_da + G;       // [2*4H + 2*H] param grad accumulators
  float* red = s_pg + 2 * G + 2 * H;
  for (int k = threadIdx.x; k < H; k += kLstmThreads) {
    s_dh[k] = p.dhT ? p.dhT[(long long)b * H + k] : 0.f;
    s_dc[k] = p.dcT ? p.dcT[(long long)b * H + k] : 0.f;
  }
  for (int j = threadIdx.x; j < 2 * G + 2 * H; j += kLstmThreads) s_pg[j] = 0.f;
  __syncthreads();
  for (int t = p.T - 1; t >= 0; --t) {
    const long long tb = (long long)t * p.B + b;
    const float mean_h = p.stats[tb * 4], rstd_h = p.stats[tb * 4 + 1], mean_c = p.stats[tb * 4 + 2], rstd_c = p.stats[tb * 4 + 3];
    const float* gi = p.gi + tb * G;
    const float* u = p.u_all + tb * G;
    const float* cprev = t > 0 ? p.c_all + ((long long)(t - 1) * p.B + b) * H : p.c0 + (long long)b * H;
    // ---- cell: per hidden unit k
    float dzhat[2], zh[2];
    float m1 = 0.f, m2 = 0.f;
    int n = 0;
    for (int k = threadIdx.x; k < H; k += kLstmThreads, ++n) {
      const float dh = s_dh[k] + (p.dy ? __ldg(p.dy + tb * H + k) : 0.f);
      const float ao = __ldg(gi + 3 * H + k) + (__ldg(u + 3 * H + k) - mean_h) * rstd_h * __ldg(p.gh_w + 3 * H + k) + __ldg(p.gh_b + 3 * H + k);
      const float o = sigmoidf_(ao);
      const float zhat = (__ldg(p.z_all + tb * H + k) - mean_c) * rstd_c;
      const float c = zhat * __ldg(p.gc_w + k) + __ldg(p.gc_b + k);
      const float tc = tanhf(c);
      s_da[3 * H + k] = dh * tc * o * (1.f - o);
      const float dc = s_dc[k] + dh * o * (1.f - tc * tc);
      s_pg[2 * G + k] += dc * zhat;            // d gc_w
      s_pg[2 * G + H + k] += dc;               // d gc_b
      const float dzh = dc * __ldg(p.gc_w + k);
      dzhat[n] = dzh; zh[n] = zhat;
      m1 += dzh; m2 += dzh * zhat;
    }
    block_sum2(m1, m2, red);
    m1 /= (float)H; m2 /= (float)H;
    n = 0;
    for (int k = threadIdx.x; k < H; k += kLstmThreads, ++n) {
      const float dz = rstd_c * (dzhat[n] - m1 - zh[n] * m2);
      const float ai = __ldg(gi + k) + (__ldg(u + k) - mean_h) * rstd_h * __ldg(p.gh_w + k) + __ldg(p.gh_b + k);
      const float af = __ldg(gi + H + k) + (__ldg(u + H + k) - mean_h) * rstd_h * __ldg(p.gh_w + H + k) + __ldg(p.gh_b + H + k);
      const float ag = __ldg(gi + 2 * H + k) + (__ldg(u + 2 * H + k) - mean_h) * rstd_h * __ldg(p.gh_w + 2 * H + k) + __ldg(p.gh_b + 2 * H + k);
      const float i = sigmoidf_(ai), f = sigmoidf_(af), g = tanhf(ag);
      s_da[k] = dz * g * i * (1.f - i);
      s_da[H + k] = dz * __ldg(cprev + k) * f * (1.f - f);
      s_da[2 * H + k] = dz * i * (1.f - g * g);
      s_dc[k] = dz * f;
    }
    __syncthreads();
    // ---- LN_h backward over 4H: d gi = da; d u = rstd (da*gamma - mean(.) - uhat mean(. uhat))
    float a1 = 0.f, a2 = 0.f;
    for (int j = threadIdx.x; j < G; j += kLstmThreads) {
      const float da = s_da[j];
      const float uh = (__ldg(u + j) - mean_h) * rstd_h;
      p.dgi[tb * G + j] = da;
      s_pg[j] += da * uh;                      // d gh_w
      s_pg[G + j] += da;                       // d gh_b
      const float d = da * __ldg(p.gh_w + j);
      a1 += d; a2 += d * uh;
    }
    block_sum2(a1, a2, red);
    a1 /= (float)G; a2 /= (float)G;
    for (int j = threadIdx.x; j < G; j += kLstmThreads) {
      const float uh = (__ldg(u + j) - mean_h) * rstd_h;
      const float d = rstd_h * (s_da[j] * __ldg(p.gh_w + j) - a1 - uh * a2);
      s_da[j] = d;
      p.du[tb * G + j] = d;
    }
    __syncthreads();
    // ---- dh_{t-1} = W_hh^T du
    matvec_bf16(p.whh_t, s_da, s_dh, H, G);
    __syncthreads();
  }
  for (int k = threadIdx.x; k < H; k += kLstmThreads) {
    if (p.dh0) p.dh0[(long long)b * H + k] = s_dh[k];
    if (p.dc0) p.dc0[(long long)b * H + k] = s_dc[k];
  }
  for (int j = threadIdx.x; j < 2 * G + 2 * H; j += kLstmThreads) p.dparam[(long long)b * (2 * G + 2 * H) + j] = s_pg[j];
}

bool lnlstm_cluster_ok(int H);
int lnlstm_small_fwd(const float*, const float*, const float*, const void*, const float*, const float*, const float*,
                     const float*, float*, float*, float*, float*, float*, int, int, cudaStream_t);
int lnlstm_small_bwd(const float*, const float*, const float*, const float*, const float*, const float*, const void*,
                     const float*, const float*, const float*, const float*, const float*, const float*, const float*,
                     float*, float*, float*, float*, float*, int, int, cudaStream_t);
int lnlstm_fwd_cluster(const float*, const float*, const float*, const void*, const float*, const float*, const float*,
                       const float*, float*, float*, float*, float*, float*, int, int, int, cudaStream_t);
int lnlstm_bwd_cluster(const float*, const float*, const float*, const float*, const float*, const float*, const void*,
                       const float*, const float*, const float*, const float*, const float*, const float*, const float*,
                       float*, float*, float*, float*, float*, int, int, int, cudaStream_t);

int lnlstm_fwd(const float* gi, const float* h0, const float* c0, const void* whh, const float* gh_w, const float* gh_b,
               const float* gc_w, const float* gc_b, float* y, float* c_all, float* u_all, float* z_all, float* stats,
               int T, int B, int H, cudaStream_t st) {
  if (T <= 0 || B <= 0) return AS_OK;
  if (H <= 0 || (H % 8) != 0 || H > 2 * kLstmThreads) return AS_ESIZE;
  if (!gi || !h0 || !c0 || !whh || !gh_w || !gh_b || !gc_w || !gc_b || !y || !c_all) return AS_EINVAL;
  if (H == 32)                   // warp per trajectory (lstm_small.cu)
    return lnlstm_small_fwd(gi, h0, c0, whh, gh_w, gh_b, gc_w, gc_b, y, c_all, u_all, z_all, stats, T, B, st);
  if (lnlstm_cluster_ok(H))      // W_hh slices resident in shared memory, 8 CTAs per trajectory (lstm_cluster.cu)
    return lnlstm_fwd_cluster(gi, h0, c0, whh, gh_w, gh_b, gc_w, gc_b, y, c_all, u_all, z_all, stats, T, B, H, st);
  LstmFwd p{gi, h0, c0, reinterpret_cast<const __nv_bfloat16*>(whh), gh_w, gh_b, gc_w, gc_b, y, c_all, u_all, z_all, stats, T, B, H};
  const int smem = (6 * H + 2 * kLstmWarps) * 4;
  lnlstm_fwd_kernel<<<B, kLstmThreads, smem, st>>>(p);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int lnlstm_bwd(const float* gi, const float* h0, const float* c0, const float* y, const float* c_all, const float* u_all,
               const float* z_all, const float* stats, const void* whh_t, const float* gh_w, const float* gh_b,
               const float* gc_w, const float* gc_b, const float* dy, const float* dhT, const float* dcT, float* dgi,
               float* du, float* dh0, float* dc0, float* dparam, int T, int B, int H, cudaStream_t st) {
  if (T <= 0 || B <= 0) return AS_OK;
  if (H <= 0 || (H % 8) != 0 || H > 2 * kLstmThreads) return AS_ESIZE;
  if (!gi || !c0 || !c_all || !u_all || !z_all || !stats || !whh_t || !dgi || !du || !dparam) return AS_EINVAL;
  if (H == 32)
    return lnlstm_small_bwd(gi, c0, c_all, u_all, z_all, stats, whh_t, gh_w, gh_b, gc_w, gc_b, dy, dhT, dcT, dgi, du, dh0,
                            dc0, dparam, T, B, st);
  if (lnlstm_cluster_ok(H))
    return lnlstm_bwd_cluster(gi, c0, c_all, u_all, z_all, stats, whh_t, gh_w, gh_b, gc_w, gc_b, dy, dhT, dcT, dgi, du,
                              dh0, dc0, dparam, T, B, H, st);
  LstmBwd p{gi, h0, c0, y, c_all, u_all, z_all, stats, reinterpret_cast<const __nv_bfloat16*>(whh_t), gh_w, gh_b, gc_w, gc_b,
            dy, dhT, dcT, dgi, du, dh0, dc0, dparam, T, B, H};
  const int smem = (2 * H + 4 * H + 8 * H + 2 * H + 2 * kLstmWarps) * 4;
  static bool attr = false;
  if (!attr) { cudaFuncSetAttribute(lnlstm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024); attr = true; }
  if (smem > 96 * 1024) return AS_ESIZE;
  lnlstm_bwd_kernel<<<B, kLstmThreads, smem, st>>>(p);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
