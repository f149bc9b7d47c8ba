// LayerNorm-LSTM recurrence on a thread-block CLUSTER (reference: distar/agent/default/model/lstm.py:138-167).
//
// The one-CTA-per-trajectory kernels in lstm.cu stream W_hh (1.2 MB for the core LSTM) from L2 at every time step and
// keep only 16 of 148 SMs busy at 16 trajectories.  Here a cluster of 8 CTAs owns one trajectory: CTA r holds the
// 4 x H/8 gate rows of W_hh that belong to hidden units [r H/8, (r+1) H/8) in shared memory for the WHOLE unroll
// (147 KB of bf16 at H = 384), so a step reads weights only from shared memory.  What crosses CTAs per step goes
// through distributed shared memory: two LayerNorm statistics exchanges (eight (mean, M2) pairs each, merged with
// the parallel-variance formula) and the new h slice (forward) or the partial W_hh^T du products (backward,
// a reduce-scatter).  Three cluster barriers per step.
//
// Both directions use the same row slice of W_hh: forward computes u = W h (row dot products), backward computes
// dh = W^T du (column sums over the CTA's rows, then summed over the cluster).
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace as {

constexpr int kCS = 8;        // CTAs per cluster (= per trajectory)
constexpr int kNT = 384;      // threads per CTA
constexpr int kNW = kNT / 32;

__device__ __forceinline__ float sigm(float x) { return 1.f / (1.f + expf(-x)); }

// block-wide sums of two values; every thread gets both results.  `red` holds 2 * kNW floats.
__device__ __forceinline__ void cta_sum2(float& a, float& b, float* red) {
  a = warp_sum(a); b = warp_sum(b);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();                                 // previous use of `red` is over
  if (l == 0) { red[w] = a; red[kNW + w] = b; }
  __syncthreads();
  float sa = 0.f, sb = 0.f;
#pragma unroll
  for (int i = 0; i < kNW; ++i) { sa += red[i]; sb += red[kNW + i]; }
  a = sa; b = sb;
}

__device__ __forceinline__ float dot8s(const uint4& q, const float* vv) {
  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&q);
  const float4 v0 = *reinterpret_cast<const float4*>(vv), v1 = *reinterpret_cast<const float4*>(vv + 4);
  const float2 a = __bfloat1622float2(h2[0]), b = __bfloat1622float2(h2[1]), c = __bfloat1622float2(h2[2]),
               d = __bfloat1622float2(h2[3]);
  return a.x * v0.x + a.y * v0.y + b.x * v0.z + b.y * v0.w + c.x * v1.x + c.y * v1.y + d.x * v1.z + d.y * v1.w;
}

struct LstmClFwd {
  const float* gi; const float* h0; const float* c0; const __nv_bfloat16* whh;
  const float* gh_w; const float* gh_b; const float* gc_w; const float* gc_b;
  float* y; float* c_all; float* u_all; float* z_all; float* stats;
  int T, B, H;
};
struct LstmClBwd {
  const float* gi; const float* c0; const float* c_all; const float* u_all; const float* z_all; const float* stats;
  const __nv_bfloat16* whh;
  const float* gh_w; const float* gh_b; const float* gc_w; const float* gc_b;
  const float* dy; const float* dhT; const float* dcT;
  float* dgi; float* du; float* dh0; float* dc0; float* dparam;
  int T, B, H;
};

// shared-memory carve-up shared by both kernels: the W_hh slice first (row pitch 2H + 16 bytes keeps 16-byte row
// reads of consecutive rows on distinct banks), then fp32 scratch
__device__ __forceinline__ void load_w_slice(uint8_t* sW, const __nv_bfloat16* whh, int H, int HS, int r, int pitchB) {
  const int rows = 4 * HS, chunks = H >> 3;
  for (int i = threadIdx.x; i < rows * chunks; i += kNT) {
    const int lr = i / chunks, ch = i - lr * chunks;
    const int g = lr / HS, kk = lr - g * HS;
    const long long j = (long long)g * H + r * HS + kk;
    *reinterpret_cast<uint4*>(sW + (long long)lr * pitchB + ch * 16) =
        __ldg(reinterpret_cast<const uint4*>(whh + j * H) + ch);
  }
}

// merge kCS (mean, M2) pairs of equal count n into (mean, rstd) over kCS * n values
__device__ __forceinline__ void merge_stats(const float* st, float n, float& mean, float& rstd) {
  float m = 0.f;
#pragma unroll
  for (int q = 0; q < kCS; ++q) m += st[2 * q];
  m *= (1.f / kCS);
  float M2 = 0.f;
#pragma unroll
  for (int q = 0; q < kCS; ++q) { const float d = st[2 * q] - m; M2 += st[2 * q + 1] + n * d * d; }
  mean = m;
  rstd = rsqrtf(M2 / (n * kCS) + 1e-5f);
}

__global__ void __launch_bounds__(kNT, 1) lnlstm_fwd_cluster_kernel(const LstmClFwd p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  cg::cluster_group cluster = cg::this_cluster();
  const int r = (int)cluster.block_rank();
  const int b = blockIdx.x / kCS;
  const int H = p.H, G = 4 * H, HS = H / kCS, rows = 4 * HS;
  const int pitchB = 2 * H + 16;
  const int parts = kNT / rows, chunks = H >> 3, cpp = chunks / parts;
  uint8_t* sW = smem_raw;
  float* s_h = reinterpret_cast<float*>(sW + (long long)rows * pitchB);   // [H]   h_{t-1}, written by all CTAs
  float* s_up = s_h + H;                 // [parts][rows] partial dot products
  float* s_a = s_up + parts * rows;      // [rows] gate pre-activations
  float* s_c = s_a + rows;               // [HS]   cell state slice
  float* s_st = s_c + HS;                // [kCS][2] LN_h partial statistics (written by peers)
  float* s_stc = s_st + 2 * kCS;         // [kCS][2] LN_c partial statistics
  float* red = s_stc + 2 * kCS;          // [2 kNW]
  const int t_id = threadIdx.x;

  load_w_slice(sW, p.whh, H, HS, r, pitchB);
  for (int k = t_id; k < H; k += kNT) s_h[k] = p.h0[(long long)b * H + k];
  if (t_id < HS) s_c[t_id] = p.c0[(long long)b * H + r * HS + t_id];
  // per-thread constants: thread t < rows owns gate row t of the slice, thread t < HS owns hidden unit t
  const int lr = t_id % rows, part = t_id / rows;
  const int g_of = lr / HS, kk_of = lr - g_of * HS;
  const long long j_row = (long long)g_of * H + r * HS + kk_of;
  float ghw = 0.f, ghb = 0.f, gcw = 0.f, gcb = 0.f;
  if (t_id < rows) { ghw = p.gh_w[j_row]; ghb = p.gh_b[j_row]; }
  if (t_id < HS) { gcw = p.gc_w[r * HS + t_id]; gcb = p.gc_b[r * HS + t_id]; }
  cluster.sync();                                   // every CTA of the cluster is resident, s_h / sW are ready

  for (int t = 0; t < p.T; ++t) {
    const long long tb = (long long)t * p.B + b;
    // ---- u = W_slice h  (thread = (row, column part))
    if (part < parts) {
      const uint4* wrow = reinterpret_cast<const uint4*>(sW + (long long)lr * pitchB) + part * cpp;
      const float* hv = s_h + part * cpp * 8;
      float acc0 = 0.f, acc1 = 0.f;
      int c = 0;
      for (; c + 1 < cpp; c += 2) { acc0 += dot8s(wrow[c], hv + c * 8); acc1 += dot8s(wrow[c + 1], hv + c * 8 + 8); }
      if (c < cpp) acc0 += dot8s(wrow[c], hv + c * 8);
      s_up[part * rows + lr] = acc0 + acc1;
    }
    __syncthreads();
    float u = 0.f;
    if (t_id < rows)
      for (int pp = 0; pp < parts; ++pp) u += s_up[pp * rows + t_id];
    // ---- LN_h statistics: local (sum, sum of squares) -> (mean, M2) -> all CTAs of the cluster
    float s1 = (t_id < rows) ? u : 0.f, s2 = (t_id < rows) ? u * u : 0.f;
    cta_sum2(s1, s2, red);
    if (t_id < kCS) {
      const float ml = s1 / (float)rows;
      float* dst = cluster.map_shared_rank(s_st, t_id);
      dst[2 * r] = ml;
      dst[2 * r + 1] = fmaxf(s2 - s1 * ml, 0.f);
    }
    cluster.sync();                                                            // (1)
    float mean_h, rstd_h;
    merge_stats(s_st, (float)rows, mean_h, rstd_h);
    if (t_id < rows) {
      if (p.u_all) p.u_all[tb * G + j_row] = u;
      s_a[t_id] = __ldg(p.gi + tb * G + j_row) + (u - mean_h) * rstd_h * ghw + ghb;
    }
    __syncthreads();
    // ---- cell: z = s(f) c + s(i) tanh(g); LN_c over all H units
    float z = 0.f;
    if (t_id < HS) {
      z = sigm(s_a[HS + t_id]) * s_c[t_id] + sigm(s_a[t_id]) * tanhf(s_a[2 * HS + t_id]);
      if (p.z_all) p.z_all[tb * H + r * HS + t_id] = z;
    }
    float z1 = (t_id < HS) ? z : 0.f, z2 = (t_id < HS) ? z * z : 0.f;
    cta_sum2(z1, z2, red);
    if (t_id < kCS) {
      const float ml = z1 / (float)HS;
      float* dst = cluster.map_shared_rank(s_stc, t_id);
      dst[2 * r] = ml;
      dst[2 * r + 1] = fmaxf(z2 - z1 * ml, 0.f);
    }
    cluster.sync();                                                            // (2)
    float mean_c, rstd_c;
    merge_stats(s_stc, (float)HS, mean_c, rstd_c);
    if (t_id < HS) {
      const float c = (z - mean_c) * rstd_c * gcw + gcb;
      const float h = sigm(s_a[3 * HS + t_id]) * tanhf(c);
      s_c[t_id] = c;
      p.y[tb * H + r * HS + t_id] = h;
      p.c_all[tb * H + r * HS + t_id] = c;
#pragma unroll
      for (int q = 0; q < kCS; ++q) cluster.map_shared_rank(s_h, q)[r * HS + t_id] = h;
    }
    if (t_id == 0 && r == 0 && p.stats) {
      p.stats[tb * 4 + 0] = mean_h; p.stats[tb * 4 + 1] = rstd_h; p.stats[tb * 4 + 2] = mean_c; p.stats[tb * 4 + 3] = rstd_c;
    }
    cluster.sync();                                                            // (3) h_t complete everywhere
  }
}

__global__ void __launch_bounds__(kNT, 1) lnlstm_bwd_cluster_kernel(const LstmClBwd p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  cg::cluster_group cluster = cg::this_cluster();
  const int r = (int)cluster.block_rank();
  const int b = blockIdx.x / kCS;
  const int H = p.H, G = 4 * H, HS = H / kCS, rows = 4 * HS;
  const int pitchB = 2 * H + 16;
  const int cpairs = H >> 1, rparts = kNT / cpairs, rpp = rows / rparts;
  uint8_t* sW = smem_raw;
  float* s_dh = reinterpret_cast<float*>(sW + (long long)rows * pitchB);   // [HS] recurrent dh slice
  float* s_dc = s_dh + HS;               // [HS]
  float* s_ap = s_dc + HS;               // [rows] gate pre-activations (recomputed)
  float* s_uh = s_ap + rows;             // [rows] normalised u
  float* s_da = s_uh + rows;             // [rows] d pre-activation
  float* s_du = s_da + rows;             // [rows] d u
  float* s_tp = s_du + rows;             // [rparts][H] partial W^T du over row parts
  float* s_dhp = s_tp + rparts * H;      // [kCS][HS] partial dh slices from every CTA (written by peers)
  float* s_x1 = s_dhp + kCS * HS;        // [kCS][2] LN_c backward sums
  float* s_x2 = s_x1 + 2 * kCS;          // [kCS][2] LN_h backward sums
  float* red = s_x2 + 2 * kCS;           // [2 kNW]
  const int t_id = threadIdx.x;

  load_w_slice(sW, p.whh, H, HS, r, pitchB);
  if (t_id < HS) {
    s_dh[t_id] = p.dhT ? p.dhT[(long long)b * H + r * HS + t_id] : 0.f;
    s_dc[t_id] = p.dcT ? p.dcT[(long long)b * H + r * HS + t_id] : 0.f;
  }
  const int lr = t_id % rows;
  const int g_of = lr / HS, kk_of = lr - g_of * HS;
  const long long j_row = (long long)g_of * H + r * HS + kk_of;
  const int k_unit = r * HS + t_id;              // valid for t_id < HS
  float ghw = 0.f, ghb = 0.f, gcw = 0.f, gcb = 0.f;
  if (t_id < rows) { ghw = p.gh_w[j_row]; ghb = p.gh_b[j_row]; }
  if (t_id < HS) { gcw = p.gc_w[k_unit]; gcb = p.gc_b[k_unit]; }
  float pg_w = 0.f, pg_b = 0.f, pg_cw = 0.f, pg_cb = 0.f;    // LayerNorm parameter gradients owned by this thread
  const int cp = t_id % cpairs, rpart = t_id / cpairs;
  cluster.sync();

  for (int t = p.T - 1; t >= 0; --t) {
    const long long tb = (long long)t * p.B + b;
    const float mean_h = __ldg(p.stats + tb * 4), rstd_h = __ldg(p.stats + tb * 4 + 1);
    const float mean_c = __ldg(p.stats + tb * 4 + 2), rstd_c = __ldg(p.stats + tb * 4 + 3);
    if (t_id < rows) {
      const float uh = (__ldg(p.u_all + tb * G + j_row) - mean_h) * rstd_h;
      s_uh[t_id] = uh;
      s_ap[t_id] = __ldg(p.gi + tb * G + j_row) + uh * ghw + ghb;
    }
    __syncthreads();
    // ---- cell backward for the CTA's hidden units
    float dzh = 0.f, zhat = 0.f, da_o = 0.f;
    if (t_id < HS) {
      const float dh = s_dh[t_id] + (p.dy ? __ldg(p.dy + tb * H + k_unit) : 0.f);
      const float o = sigm(s_ap[3 * HS + t_id]);
      zhat = (__ldg(p.z_all + tb * H + k_unit) - mean_c) * rstd_c;
      const float tc = tanhf(zhat * gcw + gcb);
      da_o = dh * tc * o * (1.f - o);
      const float dc = s_dc[t_id] + dh * o * (1.f - tc * tc);
      pg_cw += dc * zhat;
      pg_cb += dc;
      dzh = dc * gcw;
    }
    float m1 = dzh, m2 = dzh * zhat;
    cta_sum2(m1, m2, red);
    if (t_id < kCS) {
      float* dst = cluster.map_shared_rank(s_x1, t_id);
      dst[2 * r] = m1; dst[2 * r + 1] = m2;
    }
    cluster.sync();                                                            // (A)
    m1 = 0.f; m2 = 0.f;
#pragma unroll
    for (int q = 0; q < kCS; ++q) { m1 += s_x1[2 * q]; m2 += s_x1[2 * q + 1]; }
    m1 /= (float)H; m2 /= (float)H;
    if (t_id < HS) {
      const float dz = rstd_c * (dzh - m1 - zhat * m2);
      const float i = sigm(s_ap[t_id]), f = sigm(s_ap[HS + t_id]), g = tanhf(s_ap[2 * HS + t_id]);
      const float cprev = t > 0 ? __ldg(p.c_all + ((long long)(t - 1) * p.B + b) * H + k_unit) : __ldg(p.c0 + (long long)b * H + k_unit);
      s_da[t_id] = dz * g * i * (1.f - i);
      s_da[HS + t_id] = dz * cprev * f * (1.f - f);
      s_da[2 * HS + t_id] = dz * i * (1.f - g * g);
      s_da[3 * HS + t_id] = da_o;
      s_dc[t_id] = dz * f;
    }
    __syncthreads();
    // ---- LN_h backward over all 4H rows
    float da = 0.f, uh = 0.f, a1 = 0.f, a2 = 0.f;
    if (t_id < rows) {
      da = s_da[t_id]; uh = s_uh[t_id];
      p.dgi[tb * G + j_row] = da;
      pg_w += da * uh;
      pg_b += da;
      const float d = da * ghw;
      a1 = d; a2 = d * uh;
    }
    cta_sum2(a1, a2, red);
    if (t_id < kCS) {
      float* dst = cluster.map_shared_rank(s_x2, t_id);
      dst[2 * r] = a1; dst[2 * r + 1] = a2;
    }
    cluster.sync();                                                            // (B)
    a1 = 0.f; a2 = 0.f;
#pragma unroll
    for (int q = 0; q < kCS; ++q) { a1 += s_x2[2 * q]; a2 += s_x2[2 * q + 1]; }
    a1 /= (float)G; a2 /= (float)G;
    if (t_id < rows) {
      const float d = rstd_h * (da * ghw - a1 - uh * a2);
      s_du[t_id] = d;
      p.du[tb * G + j_row] = d;
    }
    __syncthreads();
    // ---- partial dh_{t-1} = W_slice^T du  (thread = (column pair, row part)), then reduce-scatter over the cluster
    if (rpart < rparts) {
      const uint8_t* wcol = sW + (long long)(rpart * rpp) * pitchB + cp * 4;
      const float* dv = s_du + rpart * rpp;
      float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 8
      for (int i = 0; i < rpp; ++i) {
        const float2 w = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(wcol + (long long)i * pitchB));
        const float d = dv[i];
        acc0 += w.x * d; acc1 += w.y * d;
      }
      s_tp[rpart * H + 2 * cp] = acc0;
      s_tp[rpart * H + 2 * cp + 1] = acc1;
    }
    __syncthreads();
    for (int k = t_id; k < H; k += kNT) {
      float v = 0.f;
      for (int rp = 0; rp < rparts; ++rp) v += s_tp[rp * H + k];
      const int q = k / HS;
      cluster.map_shared_rank(s_dhp, q)[r * HS + (k - q * HS)] = v;
    }
    cluster.sync();                                                            // (C)
    if (t_id < HS) {
      float v = 0.f;
#pragma unroll
      for (int q = 0; q < kCS; ++q) v += s_dhp[q * HS + t_id];
      s_dh[t_id] = v;
    }
    // s_dh is next read by the same threads; s_dhp is rewritten only after two more cluster barriers
  }
  if (t_id < HS) {
    if (p.dh0) p.dh0[(long long)b * H + k_unit] = s_dh[t_id];
    if (p.dc0) p.dc0[(long long)b * H + k_unit] = s_dc[t_id];
    float* dp = p.dparam + (long long)b * (2 * G + 2 * H);
    dp[2 * G + k_unit] = pg_cw;
    dp[2 * G + H + k_unit] = pg_cb;
  }
  if (t_id < rows) {
    float* dp = p.dparam + (long long)b * (2 * G + 2 * H);
    dp[j_row] = pg_w;
    dp[G + j_row] = pg_b;
  }
  cluster.sync();          // no CTA may exit while a peer can still write into its shared memory
}

static int fwd_smem_bytes(int H) {
  const int HS = H / kCS, rows = 4 * HS, parts = kNT / rows;
  return rows * (2 * H + 16) + (H + parts * rows + rows + HS + 4 * kCS + 2 * kNW) * 4;
}
static int bwd_smem_bytes(int H) {
  const int HS = H / kCS, rows = 4 * HS, rparts = kNT / (H / 2);
  return rows * (2 * H + 16) + (2 * HS + 4 * rows + rparts * H + kCS * HS + 4 * kCS + 2 * kNW) * 4;
}

// shapes the cluster kernels take: H splits into 8 slices of whole 16-byte chunks, one thread per gate row, and the
// W_hh slice fits in shared memory
bool lnlstm_cluster_ok(int H) {
  if (H <= 0 || H % (8 * kCS) != 0) return false;
  const int HS = H / kCS, rows = 4 * HS;
  if (rows > kNT || kNT % rows != 0) return false;
  const int parts = kNT / rows;
  if ((H / 8) % parts != 0) return false;
  if (H / 2 > kNT || kNT % (H / 2) != 0) return false;
  const int rparts = kNT / (H / 2);
  if (rows % rparts != 0) return false;
  return fwd_smem_bytes(H) <= 220 * 1024 && bwd_smem_bytes(H) <= 220 * 1024;
}

template <typename Params>
static int launch_cluster(void (*kern)(const Params), const Params& p, int B, int smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(B * kCS);
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

int lnlstm_fwd_cluster(const float* gi, const float* h0, const float* c0, const void* whh, const float* gh_w,
                       const float* gh_b, const float* gc_w, const float* gc_b, float* y, float* c_all, float* u_all,
                       float* z_all, float* stats, int T, int B, int H, cudaStream_t st) {
  LstmClFwd p{gi, h0, c0, reinterpret_cast<const __nv_bfloat16*>(whh), gh_w, gh_b, gc_w, gc_b, y, c_all, u_all, z_all, stats, T, B, H};
  return launch_cluster(lnlstm_fwd_cluster_kernel, p, B, fwd_smem_bytes(H), st);
}

int lnlstm_bwd_cluster(const float* gi, const float* c0, const float* c_all, const float* u_all, const float* z_all,
                       const float* stats, const void* whh, const float* gh_w, const float* gh_b, const float* gc_w,
                       const float* gc_b, const float* dy, const float* dhT, const float* dcT, float* dgi, float* du,
                       float* dh0, float* dc0, float* dparam, int T, int B, int H, cudaStream_t st) {
  LstmClBwd p{gi, c0, c_all, u_all, z_all, stats, reinterpret_cast<const __nv_bfloat16*>(whh), gh_w, gh_b, gc_w, gc_b,
              dy, dhT, dcT, dgi, du, dh0, dc0, dparam, T, B, H};
  return launch_cluster(lnlstm_bwd_cluster_kernel, p, B, bwd_smem_bytes(H), st);
}

}  // namespace as
