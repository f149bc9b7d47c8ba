// LayerNorm-LSTM recurrence on thread-block CLUSTERS (reference: distar/agent/default/model/lstm.py:138-167).
//
// The one-CTA-per-trajectory kernels in lstm.cu stream W_hh (1.2 MB for the core LSTM) from L2 at every time step and
// keep only 16 of 148 SMs busy at 16 trajectories.  Here a cluster of 8 CTAs owns one or two trajectories and keeps
// W_hh in shared memory for the WHOLE unroll: CTA r holds the rows that produce hidden units [r H/8, (r+1) H/8)
// (forward: the 4 x H/8 gate rows of W_hh, backward: the H/8 rows of W_hh^T; 147 KB of bf16 at H = 384).
//
// Per time step there is ONE exchange through distributed shared memory and ONE cluster barrier:
//   forward : every CTA computes its slice of u = W_hh h_{t-1}, stores it into the u buffer of all 8 CTAs (all-gather),
//             and after the barrier every CTA redundantly runs LN_h, the gates, LN_c and h_t for ALL units -- so h_t is
//             already complete in local shared memory for the next mat-vec.
//   backward: every CTA redundantly runs the cell / LayerNorm backward for all units (inputs are saved tensors that
//             do not depend on the recurrence, prefetched before the barrier), obtains the full du, computes its
//             slice of dh_{t-1} = W_hh^T du and all-gathers that slice.
// Thread t owns hidden unit t (H <= 384 threads) in the element-wise phases.  In the mat-vec, eight lanes share a
// weight row (interleaved 16-byte chunks: a quarter warp reads 128 contiguous bytes, conflict free) and combine with
// three shuffles.  B200 fits 15 such clusters at once (cudaOccupancyMaxActiveClusters), hence two trajectories per
// cluster when B > 15: the weight reads and every synchronisation are shared between the two.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace as {

constexpr int kCS = 8;        // CTAs per cluster
constexpr int kNT = 384;      // threads per CTA
constexpr int kNW = kNT / 32;

// MUFU.TANH (abs. error ~5e-4, far inside the bf16 W_hh budget): one instruction instead of ~30 on the serial path
__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sigm(float x) { return fmaf(tanh_fast(0.5f * x), 0.5f, 0.5f); }
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// block-wide sums of N values; every thread gets all results.  `red` holds N * kNW floats; consecutive calls must
// alternate between two buffers (the single barrier inside then also protects the buffer's previous use).
template <int N>
__device__ __forceinline__ void cta_sum(float (&v)[N], float* red) {
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    v[i] = warp_sum(v[i]);
    if (l == 0) red[i * kNW + w] = v[i];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < N; ++i) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kNW; ++k) s += red[i * kNW + k];
    v[i] = s;
  }
}

__device__ __forceinline__ void unpack8f(const uint4& q, float* f) {
  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 v = __bfloat1622float2(h2[e]);
    f[2 * e] = v.x; f[2 * e + 1] = v.y;
  }
}
// v0 / v1 = logical first / second half of one 8-float chunk of the vector
__device__ __forceinline__ float dot8f(const float* w, const float4& v0, const float4& v1) {
  return w[0] * v0.x + w[1] * v0.y + w[2] * v0.z + w[3] * v0.w + w[4] * v1.x + w[5] * v1.y + w[6] * v1.z + w[7] * v1.w;
}
// The mat-vec operand vectors (h, du) live in shared memory with the two 16-byte halves of every second group of four
// 32-byte chunks exchanged.  Lane pt of an octet reads chunk pt + 8c: lanes 0-3 fetch physical half 0 first, lanes
// 4-7 physical half 1 first -- both are the chunk's LOGICAL first half -- so one load instruction of the octet
// touches eight distinct bank groups instead of four.
__device__ __forceinline__ int swz(int e) { return e ^ ((e >> 3) & 4); }

__device__ __forceinline__ float sum8lanes(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}

// diagnostics: per-CTA cycle counts of the forward phases (as_debug_lstm_profile); nullptr in normal operation
__device__ long long* g_lstm_prof = nullptr;
#define LSTM_TICK(i)                                                    \
  do {                                                                  \
    if (prof && threadIdx.x == 0) { const long long now = clock64(); pacc[i] += now - tick; tick = now; } \
  } while (0)

struct LstmClFwd {
  const float* gi; const float* h0; const float* c0; const __nv_bfloat16* whh;
  const float* gh_w; const float* gh_b; const float* gc_w; const float* gc_b;
  float* y; float* c_all; float* u_all; float* z_all; float* stats;
  int T, B, H;
};
struct LstmClBwd {
  const float* gi; const float* c0; const float* c_all; const float* u_all; const float* z_all; const float* stats;
  const __nv_bfloat16* whh_t;
  const float* gh_w; const float* gh_b; const float* gc_w; const float* gc_b;
  const float* dy; const float* dhT; const float* dcT;
  float* dgi; float* du; float* dh0; float* dc0; float* dparam;
  int T, B, H;
};

// ---------------------------------------------------------------------------------------------------- forward
template <int NB, int H>
__global__ void __launch_bounds__(kNT, 1) lnlstm_fwd_cluster_kernel(const LstmClFwd p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  cg::cluster_group cluster = cg::this_cluster();
  const int r = (int)cluster.block_rank();
  const int b0 = (blockIdx.x / kCS) * NB;             // first trajectory of this cluster
  constexpr int G = 4 * H, HS = H / kCS, rows = 4 * HS;
  constexpr int chunks = H >> 3, cpl = chunks >> 3;    // 16-byte chunks per weight row, per lane
  static_assert(H <= kNT && H % 64 == 0 && HS * 8 <= kNT && HS % 4 == 0, "unsupported hidden size");
  uint8_t* sW = smem_raw;                              // [rows][H] bf16: rows g*HS + us  <->  W_hh[g*H + r*HS + us]
  float* s_h = reinterpret_cast<float*>(sW + (long long)rows * H * 2);   // [NB][H]   h_{t-1}
  float* s_uf = s_h + NB * H;                          // [2][NB][G] all-gathered u, double buffered (written by peers)
  float* s_sl = s_uf + 2 * NB * G;                     // [NB][4][HS] this CTA's u slice, staged for vector stores
  float* redA = s_sl + NB * rows;                      // [2 NB kNW]
  float* redB = redA + 2 * NB * kNW;
  const int t_id = threadIdx.x;

  for (int i = t_id; i < rows * chunks; i += kNT) {
    const int lr = i / chunks, ch = i - lr * chunks;
    const int g = lr / HS, us = lr - g * HS;
    reinterpret_cast<uint4*>(sW)[i] = __ldg(reinterpret_cast<const uint4*>(p.whh + ((long long)g * H + r * HS + us) * H) + ch);
  }
  // element-wise role: thread t = hidden unit t (all NB trajectories)
  const bool unit = t_id < H;
  const bool owner = unit && (t_id / HS == r);         // this CTA writes the saved tensors of its own units
  float ghw[4], ghb[4], gcw = 0.f, gcb = 0.f, c_reg[NB];
#pragma unroll
  for (int g = 0; g < 4; ++g) { ghw[g] = unit ? p.gh_w[g * H + t_id] : 0.f; ghb[g] = unit ? p.gh_b[g * H + t_id] : 0.f; }
  if (unit) { gcw = p.gc_w[t_id]; gcb = p.gc_b[t_id]; }
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
    const int b = min(b0 + nb, p.B - 1);               // a padding trajectory repeats the last one (never stored)
    c_reg[nb] = unit ? p.c0[(long long)b * H + t_id] : 0.f;
    if (unit) s_h[nb * H + swz(t_id)] = p.h0[(long long)b * H + t_id];
  }
  // mat-vec role: unit slot us (its four gate rows), lane part pt
  const int us = t_id >> 3, pt = t_id & 7, hs4 = pt & 4;
  long long* prof = g_lstm_prof;
  long long pacc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tick = 0;
  cluster.sync();                                      // all CTAs resident; sW / s_h ready
  if (prof) tick = clock64();

  for (int t = 0; t < p.T; ++t) {
    const int buf = t & 1;
    // prefetch this step's input projections (independent of the recurrence)
    float gi_r[NB][4];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      const int b = min(b0 + nb, p.B - 1);
#pragma unroll
      for (int g = 0; g < 4; ++g) gi_r[nb][g] = unit ? __ldg(p.gi + ((long long)t * p.B + b) * G + g * H + t_id) : 0.f;
    }
    // ---- u slice = W_slice h
    float acc[4][NB];
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) acc[g][nb] = 0.f;
    if (us < HS) {
      const uint4* w4 = reinterpret_cast<const uint4*>(sW) + us * chunks + pt;          // + g*HS*chunks + 8c
      const float4* h4 = reinterpret_cast<const float4*>(s_h) + 2 * pt + (hs4 >> 2);    // + nb*H/4 + 16c, ^1 = 2nd half
#pragma unroll
      for (int c = 0; c < cpl; ++c) {
        float w[4][8];
#pragma unroll
        for (int g = 0; g < 4; ++g) unpack8f(w4[g * HS * chunks + 8 * c], w[g]);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
          const float4* hp = h4 + nb * (H / 4) + 16 * c;
          const float4 v0 = hp[0], v1 = hp[1 - (hs4 >> 1)];
#pragma unroll
          for (int g = 0; g < 4; ++g) acc[g][nb] += dot8f(w[g], v0, v1);
        }
      }
    }
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) acc[g][nb] = sum8lanes(acc[g][nb]);
    LSTM_TICK(0);
    if (us < HS && pt == 0) {
#pragma unroll
      for (int nb = 0; nb < NB; ++nb)
#pragma unroll
        for (int g = 0; g < 4; ++g) s_sl[(nb * 4 + g) * HS + us] = acc[g][nb];
    }
    __syncthreads();
    LSTM_TICK(1);
    // all-gather: 16-byte stores of the slice into s_uf[buf][nb][g*H + r*HS ..] of every CTA of the cluster
    {
      const int q4 = HS >> 2, per_peer = NB * 4 * q4;
      for (int i = t_id; i < kCS * per_peer; i += kNT) {
        const int peer = i / per_peer, k = i - peer * per_peer;
        const int ng = k / q4, x4 = k - ng * q4;       // ng = nb * 4 + g
        const int nb = ng >> 2, g = ng & 3;
        float* dst = cluster.map_shared_rank(s_uf, peer) + (buf * NB + nb) * G + g * H + r * HS + x4 * 4;
        *reinterpret_cast<float4*>(dst) = *reinterpret_cast<const float4*>(s_sl + ng * HS + x4 * 4);
      }
    }
    LSTM_TICK(2);
    cluster_arrive();
    cluster_wait();
    LSTM_TICK(3);
    // ---- LN_h over all 4H gate rows, per trajectory: one block reduction of (sum, sum of squares).  The inputs are
    // O(1) with |mean| no larger than the spread, so the single-pass variance loses nothing visible in fp32.
    float u[NB][4], s[2 * NB];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      s[2 * nb] = 0.f; s[2 * nb + 1] = 0.f;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        u[nb][g] = unit ? s_uf[(buf * NB + nb) * G + g * H + t_id] : 0.f;
        s[2 * nb] += u[nb][g];
        s[2 * nb + 1] += u[nb][g] * u[nb][g];
      }
    }
    cta_sum(s, redA);
    LSTM_TICK(4);
    float mean_h[NB], rstd_h[NB], z[NB], a_o[NB];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      mean_h[nb] = s[2 * nb] * (1.f / G);
      rstd_h[nb] = rsqrtf(fmaxf(s[2 * nb + 1] * (1.f / G) - mean_h[nb] * mean_h[nb], 0.f) + 1e-5f);
      float a[4];
#pragma unroll
      for (int g = 0; g < 4; ++g) a[g] = gi_r[nb][g] + (u[nb][g] - mean_h[nb]) * rstd_h[nb] * ghw[g] + ghb[g];
      z[nb] = unit ? sigm(a[1]) * c_reg[nb] + sigm(a[0]) * tanh_fast(a[2]) : 0.f;
      a_o[nb] = a[3];
      s[2 * nb] = z[nb];
      s[2 * nb + 1] = z[nb] * z[nb];
    }
    cta_sum(s, redB);
    float mean_c[NB];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      mean_c[nb] = s[2 * nb] * (1.f / H);
      const float rstd_c = rsqrtf(fmaxf(s[2 * nb + 1] * (1.f / H) - mean_c[nb] * mean_c[nb], 0.f) + 1e-5f);
      const float c = (z[nb] - mean_c[nb]) * rstd_c * gcw + gcb;
      const float h = sigm(a_o[nb]) * tanh_fast(c);
      c_reg[nb] = c;
      if (unit) s_h[nb * H + swz(t_id)] = h;
      const int b = b0 + nb;
      if (owner && b < p.B) {
        const long long tb = (long long)t * p.B + b;
        p.y[tb * H + t_id] = h;
        p.c_all[tb * H + t_id] = c;
        if (p.z_all) p.z_all[tb * H + t_id] = z[nb];
        if (p.u_all) {
#pragma unroll
          for (int g = 0; g < 4; ++g) p.u_all[tb * G + g * H + t_id] = u[nb][g];
        }
      }
      if (t_id == 0 && r == 0 && p.stats && b < p.B) {
        float* st = p.stats + ((long long)t * p.B + b) * 4;
        st[0] = mean_h[nb]; st[1] = rstd_h[nb]; st[2] = mean_c[nb]; st[3] = rstd_c;
      }
    }
    __syncthreads();                                   // h_t complete in local shared memory
    LSTM_TICK(5);
  }
  if (prof && threadIdx.x == 0)
    for (int i = 0; i < 8; ++i) prof[blockIdx.x * 8 + i] = pacc[i];
  cluster.sync();                                      // nobody exits while a peer may still address its shared memory
}

// ---------------------------------------------------------------------------------------------------- backward
template <int NB, int H>
__global__ void __launch_bounds__(kNT, 1) lnlstm_bwd_cluster_kernel(const LstmClBwd p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  cg::cluster_group cluster = cg::this_cluster();
  const int r = (int)cluster.block_rank();
  const int b0 = (blockIdx.x / kCS) * NB;
  constexpr int G = 4 * H, HS = H / kCS;
  constexpr int chunks = G >> 3, cpl = chunks >> 3;    // chunks per W_hh^T row, per lane
  uint8_t* sW = smem_raw;                              // [HS][G] bf16: row us <-> W_hh^T[r*HS + us]
  float* s_du = reinterpret_cast<float*>(sW + (long long)HS * G * 2);    // [NB][G] full du of this step
  float* s_dhf = s_du + NB * G;                        // [2][NB][H] all-gathered dh_{t-1}, double buffered
  float* s_sl = s_dhf + 2 * NB * H;                    // [NB][HS] this CTA's dh slice, staged for vector stores
  float* redA = s_sl + NB * HS;                        // [2 NB kNW]
  float* redB = redA + 2 * NB * kNW;
  const int t_id = threadIdx.x;

  for (int i = t_id; i < HS * chunks; i += kNT) {
    const int us = i / chunks, ch = i - us * chunks;
    reinterpret_cast<uint4*>(sW)[i] = __ldg(reinterpret_cast<const uint4*>(p.whh_t + (long long)(r * HS + us) * G) + ch);
  }
  const bool unit = t_id < H;
  const bool owner = unit && (t_id / HS == r);
  float ghw[4], ghb[4], gcw = 0.f, gcb = 0.f;
#pragma unroll
  for (int g = 0; g < 4; ++g) { ghw[g] = unit ? p.gh_w[g * H + t_id] : 0.f; ghb[g] = unit ? p.gh_b[g * H + t_id] : 0.f; }
  if (unit) { gcw = p.gc_w[t_id]; gcb = p.gc_b[t_id]; }
  float dc_reg[NB], pg_w[NB][4], pg_b[NB][4], pg_cw[NB], pg_cb[NB];
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
    const int b = min(b0 + nb, p.B - 1);
    dc_reg[nb] = (unit && p.dcT) ? p.dcT[(long long)b * H + t_id] : 0.f;
    if (unit) s_dhf[(((p.T - 1) & 1) * NB + nb) * H + t_id] = p.dhT ? p.dhT[(long long)b * H + t_id] : 0.f;
    pg_cw[nb] = 0.f; pg_cb[nb] = 0.f;
#pragma unroll
    for (int g = 0; g < 4; ++g) { pg_w[nb][g] = 0.f; pg_b[nb][g] = 0.f; }
  }
  const int us = t_id >> 3, pt = t_id & 7, hs4 = pt & 4;
  cluster.sync();

  for (int t = p.T - 1; t >= 0; --t) {
    const int buf = t & 1;                             // dh_t lives in s_dhf[buf]; dh_{t-1} goes to s_dhf[buf ^ 1]
    // ---- saved tensors of this step (independent of the recurrence): issued before the barrier wait
    float gi_r[NB][4], u_r[NB][4], z_r[NB], cprev[NB], dy_r[NB], st[NB][4];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      const int b = min(b0 + nb, p.B - 1);
      const long long tb = (long long)t * p.B + b;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        gi_r[nb][g] = unit ? __ldg(p.gi + tb * G + g * H + t_id) : 0.f;
        u_r[nb][g] = unit ? __ldg(p.u_all + tb * G + g * H + t_id) : 0.f;
        st[nb][g] = __ldg(p.stats + tb * 4 + g);
      }
      z_r[nb] = unit ? __ldg(p.z_all + tb * H + t_id) : 0.f;
      dy_r[nb] = (unit && p.dy) ? __ldg(p.dy + tb * H + t_id) : 0.f;
      cprev[nb] = !unit ? 0.f : (t > 0 ? __ldg(p.c_all + ((long long)(t - 1) * p.B + b) * H + t_id) : __ldg(p.c0 + (long long)b * H + t_id));
    }
    if (t != p.T - 1) cluster_wait();                  // dh_t from every CTA has landed
    // ---- cell backward for all units
    float uh[NB][4], da[NB][4], dzh[NB], zhat[NB], gate_i[NB], gate_f[NB], gate_g[NB], s[2 * NB];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      float a[4];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uh[nb][g] = (u_r[nb][g] - st[nb][0]) * st[nb][1];
        a[g] = gi_r[nb][g] + uh[nb][g] * ghw[g] + ghb[g];
      }
      gate_i[nb] = sigm(a[0]); gate_f[nb] = sigm(a[1]); gate_g[nb] = tanh_fast(a[2]);
      const float o = sigm(a[3]);
      zhat[nb] = (z_r[nb] - st[nb][2]) * st[nb][3];
      const float tc = tanh_fast(zhat[nb] * gcw + gcb);
      const float dh = (unit ? s_dhf[(buf * NB + nb) * H + t_id] : 0.f) + dy_r[nb];
      da[nb][3] = dh * tc * o * (1.f - o);
      const float dc = dc_reg[nb] + dh * o * (1.f - tc * tc);
      pg_cw[nb] += dc * zhat[nb];
      pg_cb[nb] += dc;
      dzh[nb] = unit ? dc * gcw : 0.f;
      s[2 * nb] = dzh[nb];
      s[2 * nb + 1] = dzh[nb] * zhat[nb];
    }
    cta_sum(s, redA);
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      const float m1 = s[2 * nb] / (float)H, m2 = s[2 * nb + 1] / (float)H;
      const float dz = st[nb][3] * (dzh[nb] - m1 - zhat[nb] * m2);
      const float i = gate_i[nb], f = gate_f[nb], g = gate_g[nb];
      da[nb][0] = dz * g * i * (1.f - i);
      da[nb][1] = dz * cprev[nb] * f * (1.f - f);
      da[nb][2] = dz * i * (1.f - g * g);
      dc_reg[nb] = dz * f;
      float a1 = 0.f, a2 = 0.f;
#pragma unroll
      for (int gg = 0; gg < 4; ++gg) {
        if (!unit) da[nb][gg] = 0.f;
        pg_w[nb][gg] += da[nb][gg] * uh[nb][gg];
        pg_b[nb][gg] += da[nb][gg];
        const float d = da[nb][gg] * ghw[gg];
        a1 += d; a2 += d * uh[nb][gg];
      }
      s[2 * nb] = a1; s[2 * nb + 1] = a2;
    }
    cta_sum(s, redB);
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      const float a1 = s[2 * nb] / (float)G, a2 = s[2 * nb + 1] / (float)G;
      const int b = b0 + nb;
      const long long tb = (long long)t * p.B + b;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const float d = st[nb][1] * (da[nb][g] * ghw[g] - a1 - uh[nb][g] * a2);
        if (unit) s_du[nb * G + swz(g * H + t_id)] = d;
        if (owner && b < p.B) {
          p.dgi[tb * G + g * H + t_id] = da[nb][g];
          p.du[tb * G + g * H + t_id] = d;
        }
      }
    }
    __syncthreads();
    // ---- dh_{t-1} slice = W_hh^T[r*HS + us, :] . du, all-gathered into s_dhf[buf ^ 1] of every CTA
    float acc[NB];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) acc[nb] = 0.f;
    if (us < HS) {
      const uint4* w4 = reinterpret_cast<const uint4*>(sW) + us * chunks + pt;
      const float4* d4 = reinterpret_cast<const float4*>(s_du) + 2 * pt + (hs4 >> 2);
#pragma unroll 8
      for (int c = 0; c < cpl; ++c) {
        float w[8];
        unpack8f(w4[8 * c], w);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
          const float4* dp = d4 + nb * (G / 4) + 16 * c;
          acc[nb] += dot8f(w, dp[0], dp[1 - (hs4 >> 1)]);
        }
      }
    }
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) acc[nb] = sum8lanes(acc[nb]);
    if (us < HS && pt == 0) {
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) s_sl[nb * HS + us] = acc[nb];
    }
    __syncthreads();
    {
      const int q4 = HS >> 2, per_peer = NB * q4;
      for (int i = t_id; i < kCS * per_peer; i += kNT) {
        const int peer = i / per_peer, k = i - peer * per_peer;
        const int nb = k / q4, x4 = k - nb * q4;
        float* dst = cluster.map_shared_rank(s_dhf, peer) + ((buf ^ 1) * NB + nb) * H + r * HS + x4 * 4;
        *reinterpret_cast<float4*>(dst) = *reinterpret_cast<const float4*>(s_sl + nb * HS + x4 * 4);
      }
    }
    cluster_arrive();
  }
  cluster_wait();                                      // dh_{-1} (= d h0) complete
  const int fin = ((0 & 1) ^ 1);                       // buffer written at t = 0
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
    const int b = b0 + nb;
    if (b >= p.B) continue;
    if (owner) {
      if (p.dh0) p.dh0[(long long)b * H + t_id] = s_dhf[(fin * NB + nb) * H + t_id];
      if (p.dc0) p.dc0[(long long)b * H + t_id] = dc_reg[nb];
      float* dp = p.dparam + (long long)b * (2 * G + 2 * H);
#pragma unroll
      for (int g = 0; g < 4; ++g) { dp[g * H + t_id] = pg_w[nb][g]; dp[G + g * H + t_id] = pg_b[nb][g]; }
      dp[2 * G + t_id] = pg_cw[nb];
      dp[2 * G + H + t_id] = pg_cb[nb];
    }
  }
  cluster.sync();
}

static int fwd_smem_bytes(int H, int NB) { return 4 * (H / kCS) * H * 2 + (NB * H + 2 * NB * 4 * H + NB * 4 * (H / kCS) + 4 * NB * kNW) * 4; }
static int bwd_smem_bytes(int H, int NB) { return (H / kCS) * 4 * H * 2 + (NB * 4 * H + 2 * NB * H + NB * (H / kCS) + 4 * NB * kNW) * 4; }

// The kernels are written for any H <= 384 with H % 64 == 0 (one thread per hidden unit, one lane octet per unit of
// the CTA's slice, weight rows made of whole octets of 16-byte chunks) but are instantiated for the core LSTM only.
constexpr int kClusterH = 384;      // the instantiated hidden size (core LSTM); other sizes take the lstm.cu kernels
bool lnlstm_cluster_ok(int H) { return H == kClusterH; }

template <typename Params>
static int launch_cluster(void (*kern)(const Params), const Params& p, int clusters, int smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * kCS);
  cfg.blockDim = dim3(kNT);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kCS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  e = cudaLaunchKernelEx(&cfg, kern, p);
  if (e != cudaSuccess) return (int)e;
  return AS_OK;
}

// diagnostics: how many clusters of the forward kernel can be resident at once on the current device
int lnlstm_cluster_max_active(int H) {
  const int smem = fwd_smem_bytes(H, 2);
  if (cudaFuncSetAttribute(lnlstm_fwd_cluster_kernel<2, kClusterH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(64 * kCS);
  cfg.blockDim = dim3(kNT);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kCS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, lnlstm_fwd_cluster_kernel<2, kClusterH>, &cfg) != cudaSuccess) return -2;
  return n;
}

int lnlstm_set_profile(long long* buf) {
  return (int)cudaMemcpyToSymbol(g_lstm_prof, &buf, sizeof(buf));
}

constexpr int kOneWaveClusters = 15;   // measured on B200 for this footprint (as_debug_lstm_max_clusters)

int lnlstm_fwd_cluster(const float* gi, const float* h0, const float* c0, const void* whh, const float* gh_w,
                       const float* gh_b, const float* gc_w, const float* gc_b, float* y, float* c_all, float* u_all,
                       float* z_all, float* stats, int T, int B, int H, cudaStream_t st) {
  LstmClFwd p{gi, h0, c0, reinterpret_cast<const __nv_bfloat16*>(whh), gh_w, gh_b, gc_w, gc_b, y, c_all, u_all, z_all, stats, T, B, H};
  if (B <= kOneWaveClusters) return launch_cluster(lnlstm_fwd_cluster_kernel<1, kClusterH>, p, B, fwd_smem_bytes(H, 1), st);
  return launch_cluster(lnlstm_fwd_cluster_kernel<2, kClusterH>, p, (B + 1) / 2, fwd_smem_bytes(H, 2), st);
}

int lnlstm_bwd_cluster(const float* gi, const float* c0, const float* c_all, const float* u_all, const float* z_all,
                       const float* stats, const void* whh_t, const float* gh_w, const float* gh_b, const float* gc_w,
                       const float* gc_b, const float* dy, const float* dhT, const float* dcT, float* dgi, float* du,
                       float* dh0, float* dc0, float* dparam, int T, int B, int H, cudaStream_t st) {
  LstmClBwd p{gi, c0, c_all, u_all, z_all, stats, reinterpret_cast<const __nv_bfloat16*>(whh_t), gh_w, gh_b, gc_w, gc_b,
              dy, dhT, dcT, dgi, du, dh0, dc0, dparam, T, B, H};
  if (B <= kOneWaveClusters) return launch_cluster(lnlstm_bwd_cluster_kernel<1, kClusterH>, p, B, bwd_smem_bytes(H, 1), st);
  return launch_cluster(lnlstm_bwd_cluster_kernel<2, kClusterH>, p, (B + 1) / 2, bwd_smem_bytes(H, 2), st);
}

}  // namespace as
