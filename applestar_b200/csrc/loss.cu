// Fused V-trace / UPGO / TD(lambda) / entropy / teacher-KL loss, forward and backward
// (reference: distar/agent/default/rl_training/rl_loss.py:33-185, as_rl_utils.py).
//
//   K1 loss_stats_big / loss_stats_small : ONE pass (online softmax) over target + teacher logits per active row ->
//                       {lse, log pi(a), entropy, lse_teacher, KL(teacher||pi)}; the small kernel covers the five short
//                       heads in one launch and sums the selected-units sub-rows per step             (HBM-bound)
//   K2 loss_scan      : one CTA per trajectory b: stages the [T] columns in shared memory, then a handful of
//                       lanes run the reverse-time scans (V-trace per field x head, UPGO per head, TD(lambda) per
//                       field) and emit per-row backward coefficients, dV and per-b partial sums of every logged
//                       scalar into scratch                                                 (latency-bound, tiny)
//   K3 loss_finalize  : fixed-order reduction over b -> means[], total_loss                 (deterministic)
//   K4 loss_grad_big / loss_grad_small : second pass over the logits:
//                       dlogit = c_a (onehot - p) - c_e p (log p + H) + c_k (p - p_t), 16-byte loads / stores
//
// Rows whose head is masked for the taken action (actions_mask == 0) and selected-units sub-rows beyond
// selected_units_num contribute exactly zero to every term, so they are skipped (zero gradient is written).
#include "common.cuh"
#include "loss.h"

namespace as {

constexpr int kStatStride = 8;   // lse, logp_a, ent, lse_t, kl, lse_successive, kl_successive, (pad)
constexpr int kCoefStride = 4;   // c_a, c_e, c_k, c_dapo

struct Triple { float m, s, w; };

__device__ __forceinline__ bool row_active(const AsLossArgs& a, int h, long long row, long long* step_out) {
  long long step = (h == 3) ? row / a.S : row;
  *step_out = step;
  if (h == 3 && !a.su_mask[row]) return false;
  if (a.actions_mask[h] != nullptr && a.actions_mask[h][step] == 0) return false;
  return true;
}

// ---------------------------------------------------------------------------------------------- K1
// rho needs log pi(a) of every step of every head, masked or not (a masked step's rho still enters the V-trace
// recurrence of earlier steps), so rows with actions_mask == 0 still read their target logits; only the teacher
// row, entropy and KL are skipped there.  Selected-units sub-rows beyond selected_units_num are skipped entirely.
//
// Row statistics, one pass: running maximum with rescaling (online softmax), so a row is read from HBM once.
//   x: (mx, s = sum e^(x-mx), w = sum e^(x-mx) x)      teacher: (mt, s2, w2 = sum e^(t-mt) t, u = sum e^(t-mt) x)
struct RowAcc {
  float mx, s, w, mt, s2, w2, u;
  __device__ __forceinline__ void init() { mx = mt = -INFINITY; s = w = s2 = w2 = u = 0.f; }
  __device__ __forceinline__ void raise_x(float m) {
    if (m > mx) { const float c = expf(mx - m); s *= c; w *= c; mx = m; }
  }
  __device__ __forceinline__ void raise_t(float m) {
    if (m > mt) { const float c = expf(mt - m); s2 *= c; w2 *= c; u *= c; mt = m; }
  }
  __device__ __forceinline__ void add_x(float xv) { const float e = expf(xv - mx); s += e; w += e * xv; }
  __device__ __forceinline__ void add_xt(float xv, float tv) {
    const float e = expf(xv - mx), et = expf(tv - mt);
    s += e; w += e * xv; s2 += et; w2 += et * tv; u += et * xv;
  }
  // bring the partial sums to a common maximum (the group's) before they are added up
  __device__ __forceinline__ void rebase(float gx, float gt) {
    const float cx = (mx == -INFINITY) ? 0.f : expf(mx - gx), ct = (mt == -INFINITY) ? 0.f : expf(mt - gt);
    s *= cx; w *= cx; s2 *= ct; w2 *= ct; u *= ct; mx = gx; mt = gt;
  }
};

// one thread's share of a row: elements tr, tr + nthr, ... (16-byte loads when the row allows)
__device__ __forceinline__ void row_accumulate(RowAcc& r, const float* __restrict__ x, const float* __restrict__ t, int K, int tr, int nthr,
                                               bool full) {
  const bool vec = (K % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(t)) & 15) == 0;
  if (vec) {
    const float4* x4 = reinterpret_cast<const float4*>(x);
    const float4* t4 = reinterpret_cast<const float4*>(t);
    auto one = [&](const float4& xv, const float4& tv) {
      r.raise_x(fmaxf(fmaxf(xv.x, xv.y), fmaxf(xv.z, xv.w)));
      if (full) {
        r.raise_t(fmaxf(fmaxf(tv.x, tv.y), fmaxf(tv.z, tv.w)));
        r.add_xt(xv.x, tv.x); r.add_xt(xv.y, tv.y); r.add_xt(xv.z, tv.z); r.add_xt(xv.w, tv.w);
      } else {
        r.add_x(xv.x); r.add_x(xv.y); r.add_x(xv.z); r.add_x(xv.w);
      }
    };
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    int i = tr;
    for (; i + 3 * nthr < K / 4; i += 4 * nthr) {        // four 16-byte loads per tensor in flight per thread
      float4 xv[4], tv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) { xv[q] = __ldg(x4 + i + q * nthr); tv[q] = full ? __ldg(t4 + i + q * nthr) : z4; }
#pragma unroll
      for (int q = 0; q < 4; ++q) one(xv[q], tv[q]);
    }
    for (; i < K / 4; i += nthr) one(__ldg(x4 + i), full ? __ldg(t4 + i) : z4);
  } else {
    for (int k = tr; k < K; k += nthr) {
      const float xv = __ldg(x + k);
      r.raise_x(xv);
      if (full) { const float tv = __ldg(t + k); r.raise_t(tv); r.add_xt(xv, tv); } else r.add_x(xv);
    }
  }
}

__device__ __forceinline__ void warp_combine(RowAcc& r) {
  const float gx = warp_max(r.mx), gt = warp_max(r.mt);
  r.rebase(gx, gt);
  r.s = warp_sum(r.s); r.w = warp_sum(r.w); r.s2 = warp_sum(r.s2); r.w2 = warp_sum(r.w2); r.u = warp_sum(r.u);
}

// DAPO (as_rl_utils.py:105-127): KL(successive || pi) needs ms = max c, s3 = sum e^(c-ms), w3 = sum e^(c-ms) c,
// u3 = sum e^(c-ms) x (two passes; the successive row is read while use_dapo is on only).  Returns kl via st[5], st[6].
__device__ __forceinline__ void warp_dapo(const float* __restrict__ x, const float* __restrict__ sc, int K, int lane, float lse, float* st) {
  float ms = -INFINITY;
  for (int k = lane; k < K; k += 32) ms = fmaxf(ms, __ldg(sc + k));
  ms = warp_max(ms);
  float s3 = 0.f, w3 = 0.f, u3 = 0.f;
  for (int k = lane; k < K; k += 32) {
    const float cv = __ldg(sc + k);
    const float ec = expf(cv - ms);
    s3 += ec; w3 += ec * cv; u3 += ec * __ldg(x + k);
  }
  s3 = warp_sum(s3); w3 = warp_sum(w3); u3 = warp_sum(u3);
  if (lane == 0) {
    const float lse_s = ms + logf(s3);
    st[5] = lse_s;
    st[6] = (w3 / s3 - lse_s) - (u3 / s3 - lse);            // KL  sum p_s (log p_s - log p)
  }
}

__device__ __forceinline__ void write_row_stats(const AsLossArgs& a, int h, long long row, const RowAcc& r, bool full, float* st,
                                                const float* x) {
  const float lse = r.mx + logf(r.s);
  st[0] = lse;
  st[1] = __ldg(x + a.action[h][row]) - lse;                  // log pi(a)
  if (full) {
    const float lse_t = r.mt + logf(r.s2);
    st[2] = lse - r.w / r.s;                                  // entropy  -sum p log p
    st[3] = lse_t;
    st[4] = (r.w2 / r.s2 - lse_t) - (r.u / r.s2 - lse);       // KL  sum p_t (log p_t - log p)
  } else {
    st[2] = 0.f; st[3] = 0.f; st[4] = 0.f;
  }
  st[5] = 0.f; st[6] = 0.f;
}

// warp per row (K < 4096); returns lse to every lane.  Rows of up to 17 x 32 logits (every short head: 327, 128, 2, 513,
// 512) are held in registers: all loads of the row are issued back to back, then maximum and sums come from registers.
constexpr int kRegRow = 17;
__device__ __forceinline__ float warp_row_stats(const AsLossArgs& a, int h, long long row, int K, int lane, bool full) {
  float* st = a.row_stats[h] + row * kStatStride;
  const float* x = a.logit[h] + row * (long long)K;
  const float* t = a.teacher[h] + row * (long long)K;
  RowAcc r; r.init();
  if (K <= kRegRow * 32) {
    float xr[kRegRow], tr[kRegRow];
#pragma unroll
    for (int i = 0; i < kRegRow; ++i) {
      const int k = lane + 32 * i;
      xr[i] = (k < K) ? __ldg(x + k) : -INFINITY;
      tr[i] = (full && k < K) ? __ldg(t + k) : -INFINITY;
    }
#pragma unroll
    for (int i = 0; i < kRegRow; ++i) { r.mx = fmaxf(r.mx, xr[i]); r.mt = fmaxf(r.mt, tr[i]); }
    r.mx = warp_max(r.mx); r.mt = warp_max(r.mt);
#pragma unroll
    for (int i = 0; i < kRegRow; ++i) {
      if (lane + 32 * i < K) {
        if (full) r.add_xt(xr[i], tr[i]); else r.add_x(xr[i]);
      }
    }
    r.s = warp_sum(r.s); r.w = warp_sum(r.w); r.s2 = warp_sum(r.s2); r.w2 = warp_sum(r.w2); r.u = warp_sum(r.u);
  } else {
    row_accumulate(r, x, t, K, lane, 32, full);
    warp_combine(r);
  }
  if (lane == 0) write_row_stats(a, h, row, r, full, st, x);
  const float lse = r.mx + logf(r.s);
  if (full && a.successive[h] != nullptr) {
    __syncwarp();
    warp_dapo(x, a.successive[h] + row * (long long)K, K, lane, lse, st);
  }
  return lse;
}

// K >= 4096 (target_location): one CTA of 256 threads per row
__global__ void __launch_bounds__(256) loss_stats_big_kernel(const AsLossArgs a, int h, long long rows, int K) {
  __shared__ float red[8][8];
  const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    long long step = 0;
    const bool full = row_active(a, h, row, &step);   // false: only lse / log pi(a) are needed
    float* st = a.row_stats[h] + row * kStatStride;
    const float* x = a.logit[h] + row * (long long)K;
    const float* t = a.teacher[h] + row * (long long)K;
    RowAcc r; r.init();
    row_accumulate(r, x, t, K, threadIdx.x, 256, full);
    warp_combine(r);
    __syncthreads();                                  // red[] of the previous row has been consumed
    if (lane == 0) { red[0][wi] = r.mx; red[1][wi] = r.mt; }
    __syncthreads();
    float gx = red[0][0], gt = red[1][0];
#pragma unroll
    for (int i = 1; i < 8; ++i) { gx = fmaxf(gx, red[0][i]); gt = fmaxf(gt, red[1][i]); }
    r.rebase(gx, gt);
    if (lane == 0) { red[2][wi] = r.s; red[3][wi] = r.w; red[4][wi] = r.s2; red[5][wi] = r.w2; red[6][wi] = r.u; }
    __syncthreads();
    if (threadIdx.x == 0) {
      r.s = r.w = r.s2 = r.w2 = r.u = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { r.s += red[2][i]; r.w += red[3][i]; r.s2 += red[4][i]; r.w2 += red[5][i]; r.u += red[6][i]; }
      write_row_stats(a, h, row, r, full, st, x);
      red[7][0] = r.mx + logf(r.s);
    }
    if (full && a.successive[h] != nullptr) {          // DAPO: rare path, one warp walks the row
      __syncthreads();
      if (wi == 0) warp_dapo(x, a.successive[h] + row * (long long)K, K, lane, red[7][0], st);
    }
  }
}

// The four short heads (action_type, delay, queued, target_unit: warp per row, 8 rows per CTA) and selected_units (CTA per
// step: its valid sub-rows are spread over the 8 warps, then summed in row order into the per-step aggregates the scan
// needs) in ONE launch.  grid = 4 * ceil(TB / 8) + TB.
// su_step[step] = {sum log pi, sum (log pi - log mu), sum entropy stat, sum KL stat, sum DAPO stat, valid rows, valid logits, -}
__global__ void __launch_bounds__(256) loss_stats_small_kernel(const AsLossArgs a, int nb) {
  const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
  const long long TB = (long long)a.T * a.B;
  if ((int)blockIdx.x < 4 * nb) {
    const int hsel = blockIdx.x / nb;
    const int h = hsel == 3 ? 4 : hsel;
    const long long row = (long long)(blockIdx.x - hsel * nb) * 8 + wi;
    if (row >= TB) return;
    long long step = 0;
    const bool full = row_active(a, h, row, &step);
    warp_row_stats(a, h, row, a.head_k[h], lane, full);
    if (h == 4) {                                     // valid target-unit logits of the step (entropy normaliser)
      float n = 0.f;
      for (int j = lane; j < a.NE; j += 32) n += a.tu_logits_mask[row * a.NE + j] ? 1.f : 0.f;
      n = warp_sum(n);
      if (lane == 0) a.row_stats[4][row * kStatStride + 7] = n;
    }
    return;
  }
  __shared__ float rowv[64][5];
  __shared__ unsigned char rowon[64];
  const long long step = blockIdx.x - 4 * nb;
  const int S = a.S;                                  // <= 64 (host-checked)
  for (int j = wi; j < S; j += 8) {
    const long long row = step * S + j;
    const bool on = a.su_mask[row] != 0;
    if (lane == 0) rowon[j] = on;
    if (!on) continue;
    long long st_ = 0;
    const bool full = row_active(a, 3, row, &st_);
    warp_row_stats(a, 3, row, a.head_k[3], lane, full);
    __syncwarp();
    if (lane == 0) {
      const float* st = a.row_stats[3] + row * kStatStride;
      rowv[j][0] = st[1]; rowv[j][1] = st[1] - a.mu_logp[3][row]; rowv[j][2] = st[2]; rowv[j][3] = st[4]; rowv[j][4] = st[6];
    }
  }
  float n = 0.f;
  if (wi == 7) {
    for (int j = lane; j < a.N1; j += 32) n += a.su_logits_mask[step * a.N1 + j] ? 1.f : 0.f;
    n = warp_sum(n);
  }
  __syncthreads();
  float* out = a.su_step + step * 8;
  if (threadIdx.x < 5) {
    float acc = 0.f;
    for (int j = 0; j < S; ++j) if (rowon[j]) acc += rowv[j][threadIdx.x];
    out[threadIdx.x] = acc;
  } else if (threadIdx.x == 5) {
    float c = 0.f;
    for (int j = 0; j < S; ++j) c += rowon[j] ? 1.f : 0.f;
    out[5] = c;
  } else if (wi == 7 && lane == 0) {
    out[6] = n; out[7] = 0.f;
  }
}

// ---------------------------------------------------------------------------------------------- K2
// Stat slots (per-b partial sums, later divided by T*B):
//   [f*10 + h]      pg loss of field f, head h (h<6);  [f*10+6] td;  [f*10+7] reward;  [f*10+8] value (mean over T+1)
//   base = F*10:    [base + h] upgo h;  [base+6+h] entropy h;  [base+12+h] kl h;  [base+18] kl extra_at;
//                   [base+19] battle reward;  [base+20+h] DAPO kl h
__device__ __forceinline__ float head_mask_at(const AsLossArgs& a, int h, long long step) {
  return a.actions_mask[h] ? (float)a.actions_mask[h][step] : 1.f;
}

// One CTA per trajectory, three stages:
//   A  (all 256 threads) gather everything the scans need into shared memory with coalesced / warp-parallel reads:
//      per-step log-probs, importance ratios, head and field masks, values, rewards, the per-step sums over the
//      selected-units rows and the valid-logit counts of the two pointer heads (513 / 512 mask bytes per step);
//   B  (one thread per (head | field | term) job) the time recurrences, reading shared memory only;
//   C  (all threads) coalesced write-out of the backward coefficients.
// The earlier version ran every loop of A and C inside the single-thread jobs (4096 dependent global accesses per
// selected-units job): 1.43 ms; this one is bound by the 64-step recurrences themselves.
constexpr int kScanThreads = 256;

__global__ void __launch_bounds__(kScanThreads) loss_scan_kernel(const AsLossArgs a) {
  const int b = blockIdx.x;
  const int T = a.T, B = a.B, F = a.F;
  extern __shared__ float sm[];
  float* s_lp = sm;                       // [6][T]  log pi(a) per head (selected units: sum over valid rows)
  float* s_rho = s_lp + 6 * T;            // [6][T]  min(1, pi / mu)
  float* s_hm = s_rho + 6 * T;            // [6][T]  head masks
  float* s_st2 = s_hm + 6 * T;            // [6][T]  entropy statistic (selected units: sum over valid rows)
  float* s_st4 = s_st2 + 6 * T;           // [6][T]  KL statistic
  float* s_c0 = s_st4 + 6 * T;            // [6][T]  out: policy-gradient + UPGO coefficient
  float* s_c1 = s_c0 + 6 * T;             // [6][T]  out: entropy coefficient (per valid row for selected units)
  float* s_c2 = s_c1 + 6 * T;             // [6][T]  out: KL coefficient
  float* s_v = s_c2 + 6 * T;              // [F][T+1]
  float* s_r = s_v + F * (T + 1);         // [F][T]
  float* s_fm = s_r + F * T;              // [F][T]  field masks
  float* s_dv = s_fm + F * T;             // [F][T]  out: d value
  float* s_cnt = s_dv + F * T;            // [T] valid selected-units rows
  float* s_nlsu = s_cnt + T;              // [T] valid selected-units logits
  float* s_nltu = s_nlsu + T;             // [T] valid target-unit logits
  float* s_atm = s_nltu + T;              // [T] action-type extra-KL mask
  float* s_br = s_atm + T;                // [T] battle reward
  float* s_st6 = s_br + T;                // [6][T]  DAPO KL statistic
  float* s_c3 = s_st6 + 6 * T;            // [6][T]  out: DAPO KL coefficient
  float* s_dpm = s_c3 + 6 * T;            // [T] DAPO step flag
  float* s_cu = s_dpm + T;                // [6][T]  out: UPGO part of the policy-gradient coefficient
  float* s_ent = s_cu + 6 * T;            // [6][T]  normalised, masked entropy per step
  const bool dapo = a.successive[0] != nullptr;
  const float inv_tb = 1.f / (float)(T * B);
  const int tid = threadIdx.x;

  // ---- A1: per-step scalars, thread = (t, item)
  for (int i = tid; i < T * 6; i += kScanThreads) {
    const int h = i / T, t = i - h * T;
    const long long step = (long long)t * B + b;
    s_hm[i] = head_mask_at(a, h, step);
    s_c0[i] = 0.f;
    if (h != 3) {
      const float lp = a.row_stats[h][step * kStatStride + 1];
      s_lp[i] = lp;
      s_rho[i] = fminf(expf(lp - a.mu_logp[h][step]), 1.f);
      s_st2[i] = a.row_stats[h][step * kStatStride + 2];
      s_st4[i] = a.row_stats[h][step * kStatStride + 4];
      s_st6[i] = a.row_stats[h][step * kStatStride + 6];
    }
    s_c3[i] = 0.f;
  }
  for (int i = tid; i < T * F; i += kScanThreads) {
    const int f = i / T, t = i - f * T;
    const long long step = (long long)t * B + b;
    s_r[i] = a.reward[f][step];
    s_fm[i] = a.field_mask[f] ? a.field_mask[f][step] : 1.f;
  }
  for (int i = tid; i < (T + 1) * F; i += kScanThreads) {
    const int f = i / (T + 1), t = i - f * (T + 1);
    float v = a.value[f][(long long)t * B + b];
    if (t == T) v *= (a.winloss_reward[(long long)(T - 1) * B + b] == 0.f) ? 1.f : 0.f;   // rl_loss.py:47-49
    s_v[i] = v;
  }
  for (int t = tid; t < T; t += kScanThreads) {
    const long long step = (long long)t * B + b;
    s_atm[t] = ((a.step[step] < a.at_kl_steps) ? 1.f : 0.f) * a.cum_action_mask[step];
    s_br[t] = a.battle_reward[step];
    s_dpm[t] = (a.step[step] < a.dapo_steps) ? 1.f : 0.f;
  }
  // ---- A2: per-step aggregates of the selected-units rows and the pointer-head logit counts (written by K1)
  for (int t = tid; t < T; t += kScanThreads) {
    const long long step = (long long)t * B + b;
    const float* g = a.su_step + step * 8;
    s_lp[3 * T + t] = g[0];
    s_rho[3 * T + t] = fminf(expf(g[1]), 1.f);
    s_st2[3 * T + t] = g[2];
    s_st4[3 * T + t] = g[3];
    s_st6[3 * T + t] = g[4];
    s_cnt[t] = g[5]; s_nlsu[t] = g[6];
    s_nltu[t] = a.row_stats[4][step * kStatStride + 7];
  }
  __syncthreads();

  // ---- A3: entropy per (head, step) and its backward coefficient (as_rl_utils.py:52-72): element-wise, all threads
  for (int i = tid; i < T * 6; i += kScanThreads) {
    const int h = i / T, t = i - h * T;
    const float hm = s_hm[i];
    float norm, denom = 1.f;
    if (h == 3) { norm = 1e-9f + logf(s_nlsu[t] + 1.f); denom = s_cnt[t] + 1e-9f; }
    else norm = (h == 4) ? 1e-9f + logf(s_nltu[t] + 1.f) : logf((float)a.head_k[h]);
    s_ent[i] = s_st2[i] / norm / denom * hm;
    s_c1[i] = -a.w_ent * a.hw_ent[h] * hm * inv_tb / (norm * denom);
    s_cu[i] = 0.f;
  }
  __syncthreads();

  // ---- B: recurrences (shared memory only)
  float* part = a.partial + (long long)b * AS_LOSS_MAX_STATS;
  const int base = F * 10;
  if (tid < 6) {
    // head job: V-trace per field (gamma = lambda = 1, rl_loss.py:102-103) + UPGO on winloss
    const int h = tid;
    const float* rho = s_rho + h * T;
    const float* lp = s_lp + h * T;
    const float* hm = s_hm + h * T;
    float* c0 = s_c0 + h * T;
    for (int f = 0; f < F; ++f) {
      const float* v = s_v + f * (T + 1);
      const float* r = s_r + f * T;
      const float* fm = s_fm + f * T;
      const float wv = a.w_pg[f] * a.hw_pg[h] * inv_tb;
      float acc = 0.f;
      float vs_next = v[T];
      for (int t = T - 1; t >= 0; --t) {
        const float delta = rho[t] * (r[t] + v[t + 1] - v[t]);
        const float vs = v[t] + delta + rho[t] * (vs_next - v[t + 1]);
        const float adv = rho[t] * (r[t] + vs_next - v[t]);
        vs_next = vs;
        const float m = hm[t] * fm[t];
        acc += -adv * lp[t] * m;
        c0[t] += -adv * m * wv;
      }
      part[f * 10 + h] = acc;
    }
  } else if (tid >= 192 && tid < 198) {
    // UPGO on winloss (as_rl_utils.py:265-281, :31-49), one thread per head
    const int h = tid - 192;
    const float* rho = s_rho + h * T;
    const float* lp = s_lp + h * T;
    const float* hm = s_hm + h * T;
    float* cu = s_cu + h * T;
    const int f = a.upgo_field;
    const float* v = s_v + f * (T + 1);
    const float* r = s_r + f * T;
    const float wu = a.w_upgo * a.hw_upgo[h] * inv_tb;
    float acc = 0.f;
    float g_next = 0.f;
    for (int t = T - 1; t >= 0; --t) {
      float g;
      if (t == T - 1) {
        g = r[t] + v[t + 1];
      } else {
        const float lam = (r[t + 1] + v[t + 2] >= v[t + 1]) ? 1.f : 0.f;    // as_rl_utils.py:279-280
        g = r[t] + lam * g_next + (1.f - lam) * v[t + 1];
      }
      g_next = g;
      const float adv = rho[t] * (g - v[t]);
      acc += -adv * lp[t] * hm[t];
      cu[t] = -adv * hm[t] * wu;
    }
    part[base + h] = acc;
  } else if (tid >= 32 && tid < 32 + F) {
    // TD(lambda = 0.8) per field (as_rl_utils.py:221-243)
    const int f = tid - 32;
    const float* v = s_v + f * (T + 1);
    const float* r = s_r + f * T;
    const float gamma = a.gamma_td[f], lam = 0.8f;
    float acc = 0.f, racc = 0.f, vacc = v[T];
    float g_next = 0.f;
    for (int t = T - 1; t >= 0; --t) {
      float g;
      if (t == T - 1) g = r[t] + gamma * v[t + 1];
      else {
        const float d = gamma * lam;
        g = r[t] + d * g_next + (gamma - d) * v[t + 1];
      }
      g_next = g;
      const float m = s_fm[f * T + t];
      const float diff = g - v[t];
      acc += 0.5f * diff * diff * m;
      racc += r[t];
      vacc += v[t];
      s_dv[f * T + t] = -diff * m * a.w_td[f] * inv_tb;
    }
    part[f * 10 + 6] = acc;
    part[f * 10 + 7] = racc;
    part[f * 10 + 8] = vacc * ((float)T / (float)(T + 1));      // finalize multiplies by 1/(T*B); mean is over (T+1)*B
    part[f * 10 + 9] = 0.f;
  } else if (tid >= 64 && tid < 70) {
    // entropy per head: sum of the per-step values of A3
    const int h = tid - 64;
    float acc = 0.f;
    for (int t = 0; t < T; ++t) acc += s_ent[h * T + t];
    part[base + 6 + h] = acc;
  } else if (tid >= 96 && tid < 102) {
    // KL(teacher || pi) per head (as_rl_utils.py:75-103) and its backward coefficient
    const int h = tid - 96;
    float acc = 0.f, acc_at = 0.f;
    for (int t = 0; t < T; ++t) {
      const float hm = s_hm[h * T + t];
      float ck = a.w_kl * a.hw_kl[h] * hm * inv_tb;
      const float kl = s_st4[h * T + t];
      if (h == 0) {
        const float m = s_atm[t];
        acc_at += kl * m;
        ck += a.w_atkl * m * inv_tb;
      }
      s_c2[h * T + t] = ck;
      acc += kl * hm;
    }
    part[base + 12 + h] = acc;
    if (h == 0) part[base + 18] = acc_at;
  } else if (tid == 128) {
    float acc = 0.f;
    for (int t = 0; t < T; ++t) acc += s_br[t];
    part[base + 19] = acc;
  } else if (tid >= 160 && tid < 166) {
    // DAPO: KL(successive || pi) per head while step < dapo_steps (as_rl_utils.py:105-127)
    const int h = tid - 160;
    float acc = 0.f;
    if (dapo) {
      for (int t = 0; t < T; ++t) {
        const float m = s_hm[h * T + t] * s_dpm[t];
        s_c3[h * T + t] = a.w_dapo * a.hw_dapo[h] * m * inv_tb;
        acc += s_st6[h * T + t] * m;
      }
    }
    part[base + 20 + h] = acc;
  }
  __syncthreads();

  // ---- C: coefficients and value gradients, coalesced
  for (int i = tid; i < T * 6; i += kScanThreads) {
    const int h = i / T, t = i - h * T;
    const long long step = (long long)t * B + b;
    // selected units: one coefficient set per STEP; the gradient kernel applies it to the step's valid sub-rows
    *reinterpret_cast<float4*>(a.coef[h] + step * kCoefStride) = make_float4(s_c0[i] + s_cu[i], s_c1[i], s_c2[i], s_c3[i]);
  }
  for (int i = tid; i < T * F; i += kScanThreads) {
    const int f = i / T, t = i - f * T;
    a.dvalue[f][(long long)t * B + b] = s_dv[i];
  }
  if (tid < F) a.dvalue[tid][(long long)T * B + b] = 0.f;
}

// ---------------------------------------------------------------------------------------------- K3
__global__ void loss_finalize_kernel(const AsLossArgs a) {
  const int n = a.F * 10 + 26;
  __shared__ float tot[64];
  const float inv_tb = 1.f / (float)(a.T * a.B);
  float contrib = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float s = 0.f;
    for (int b = 0; b < a.B; ++b) s += a.partial[(long long)b * AS_LOSS_MAX_STATS + i];   // fixed order
    s *= inv_tb;
    a.out[i] = s;
    contrib += s * a.stat_weight[i];
  }
  contrib = warp_sum(contrib);
  if ((threadIdx.x & 31) == 0) tot[threadIdx.x >> 5] = contrib;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += tot[i];
    a.out[AS_LOSS_MAX_STATS - 1] = s;                    // total_loss
  }
}

// ---------------------------------------------------------------------------------------------- K4
struct GradRow {                 // everything one active row needs
  float lse, ent, lse_t, lse_s, c_a, c_e, c_k, c_d;
  int act_idx;
};
__device__ __forceinline__ GradRow grad_row_params(const AsLossArgs& a, int h, long long row, long long step, float gscale) {
  GradRow g;
  const float* st = a.row_stats[h] + row * kStatStride;
  g.lse = st[0]; g.ent = st[2]; g.lse_t = st[3]; g.lse_s = st[5];
  const float4 cf = *reinterpret_cast<const float4*>(a.coef[h] + step * kCoefStride);    // per step (also for selected units)
  g.c_a = cf.x * gscale; g.c_e = cf.y * gscale; g.c_k = cf.z * gscale; g.c_d = cf.w * gscale;
  g.act_idx = (int)a.action[h][row];
  return g;
}
__device__ __forceinline__ float grad_elem(const GradRow& g, float xv, float tv, int k) {
  const float lp = xv - g.lse;
  const float p = expf(lp);
  float v = -g.c_a * p - g.c_e * p * (lp + g.ent) + g.c_k * (p - expf(tv - g.lse_t));
  if (k == g.act_idx) v += g.c_a;
  return v;
}

// K >= 4096: one CTA per row, 16-byte loads and stores
__global__ void __launch_bounds__(256) loss_grad_big_kernel(const AsLossArgs a, int h, long long rows, int K, float gscale) {
  if (a.grad_scale != nullptr) gscale *= __ldg(a.grad_scale);
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    long long step = 0;
    const bool act = row_active(a, h, row, &step);
    float* dx = a.dlogit[h] + row * (long long)K;
    const float* x = a.logit[h] + row * (long long)K;
    const float* t = a.teacher[h] + row * (long long)K;
    const bool vec = (K % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(t) | reinterpret_cast<uintptr_t>(dx)) & 15) == 0;
    if (!act) {
      if (vec) for (int i = threadIdx.x; i < K / 4; i += 256) reinterpret_cast<float4*>(dx)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      else for (int k = threadIdx.x; k < K; k += 256) dx[k] = 0.f;
      continue;
    }
    const GradRow g = grad_row_params(a, h, row, step, gscale);
    const float* sc = (a.successive[h] != nullptr && g.c_d != 0.f) ? a.successive[h] + row * (long long)K : nullptr;
    if (vec && sc == nullptr) {
      const float4* x4 = reinterpret_cast<const float4*>(x);
      const float4* t4 = reinterpret_cast<const float4*>(t);
      auto one = [&](int i, const float4& xv, const float4& tv) {
        reinterpret_cast<float4*>(dx)[i] = make_float4(grad_elem(g, xv.x, tv.x, 4 * i), grad_elem(g, xv.y, tv.y, 4 * i + 1),
                                                       grad_elem(g, xv.z, tv.z, 4 * i + 2), grad_elem(g, xv.w, tv.w, 4 * i + 3));
      };
      int i = threadIdx.x;
      for (; i + 3 * 256 < K / 4; i += 4 * 256) {
        float4 xv[4], tv[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) { xv[q] = __ldg(x4 + i + q * 256); tv[q] = __ldg(t4 + i + q * 256); }
#pragma unroll
        for (int q = 0; q < 4; ++q) one(i + q * 256, xv[q], tv[q]);
      }
      for (; i < K / 4; i += 256) one(i, __ldg(x4 + i), __ldg(t4 + i));
    } else {
      for (int k = threadIdx.x; k < K; k += 256) {
        const float xv = __ldg(x + k);
        float v = grad_elem(g, xv, __ldg(t + k), k);
        if (sc) v += g.c_d * (expf(xv - g.lse) - expf(__ldg(sc + k) - g.lse_s));
        dx[k] = v;
      }
    }
  }
}

// The four short heads (warp per row) and selected_units in ONE launch.  The selected-units gradient [T*B*S, N1] is walked
// flat (its rows are 513 floats, so a row-wise walk cannot use vector stores): 95 % of it is the zero gradient of
// sub-rows beyond selected_units_num, and this pass is bound by writing those zeros.
// grid = 4 * ceil(TB / 8) + TB.
__global__ void __launch_bounds__(256) loss_grad_small_kernel(const AsLossArgs a, int nb, float gscale) {
  if (a.grad_scale != nullptr) gscale *= __ldg(a.grad_scale);
  const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
  const long long TB = (long long)a.T * a.B;
  if ((int)blockIdx.x < 4 * nb) {
    const int hsel = blockIdx.x / nb;
    const int h = hsel == 3 ? 4 : hsel;
    const long long row = (long long)(blockIdx.x - hsel * nb) * 8 + wi;
    if (row >= TB) return;
    const int K = a.head_k[h];
    long long step = 0;
    const bool act = row_active(a, h, row, &step);
    float* dx = a.dlogit[h] + row * (long long)K;
    if (!act) {
      for (int k = lane; k < K; k += 32) dx[k] = 0.f;
      return;
    }
    const GradRow g = grad_row_params(a, h, row, step, gscale);
    const float* x = a.logit[h] + row * (long long)K;
    const float* t = a.teacher[h] + row * (long long)K;
    const float* sc = (a.successive[h] != nullptr && g.c_d != 0.f) ? a.successive[h] + row * (long long)K : nullptr;
    for (int k = lane; k < K; k += 32) {
      const float xv = __ldg(x + k);
      float v = grad_elem(g, xv, __ldg(t + k), k);
      if (sc) v += g.c_d * (expf(xv - g.lse) - expf(__ldg(sc + k) - g.lse_s));
      dx[k] = v;
    }
    return;
  }
  // selected units: CTA per step.  The step's [S, N1] gradient block is contiguous; the parameters of its valid sub-rows
  // go to shared memory once, then the block is walked flat in 16-byte pieces (most of it is the zero gradient of the
  // sub-rows beyond selected_units_num).
  __shared__ GradRow sg[64];
  __shared__ unsigned char son[64];
  const long long step = blockIdx.x - 4 * nb;
  const int K = a.head_k[3], S = a.S;
  if (threadIdx.x < S) {
    const long long row = step * S + threadIdx.x;
    long long st_ = 0;
    const bool on = a.su_mask[row] != 0 && row_active(a, 3, row, &st_);
    son[threadIdx.x] = on;
    if (on) sg[threadIdx.x] = grad_row_params(a, 3, row, step, gscale);
  }
  __syncthreads();
  const long long base = step * S * (long long)K;
  const float* x = a.logit[3] + base;
  const float* t = a.teacher[3] + base;
  const float* sc = a.successive[3] ? a.successive[3] + base : nullptr;
  float* dx = a.dlogit[3] + base;
  const int total = S * K;
  auto elem = [&](int e) -> float {
    const int j = e / K;
    if (!son[j]) return 0.f;
    const GradRow& g = sg[j];
    const float xv = __ldg(x + e);
    float v = grad_elem(g, xv, __ldg(t + e), e - j * K);
    if (sc != nullptr && g.c_d != 0.f) v += g.c_d * (expf(xv - g.lse) - expf(__ldg(sc + e) - g.lse_s));
    return v;
  };
  const int vec_end = ((reinterpret_cast<uintptr_t>(dx) & 15) == 0) ? total / 4 : 0;
  for (int i = threadIdx.x; i < vec_end; i += 256) {
    const int e = 4 * i;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (son[e / K] | son[(e + 3) / K]) v = make_float4(elem(e), elem(e + 1), elem(e + 2), elem(e + 3));
    reinterpret_cast<float4*>(dx)[i] = v;
  }
  for (int e = vec_end * 4 + threadIdx.x; e < total; e += 256) dx[e] = elem(e);
}

// ---------------------------------------------------------------------------------------------- host
static int stats_grid(long long rows, int rpb) {
  long long blocks = (rows + rpb - 1) / rpb;
  const long long cap = 148LL * 8;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

int loss_forward(const AsLossArgs& a, cudaStream_t st) {
  if (a.T <= 0 || a.B <= 0 || a.F <= 0 || a.F > AS_LOSS_MAX_FIELDS || a.S <= 0 || a.S > 64) return AS_ESIZE;
  if (a.F * 10 + 26 >= AS_LOSS_MAX_STATS) return AS_ESIZE;
  for (int h = 1; h < 6; ++h)
    if ((a.successive[h] != nullptr) != (a.successive[0] != nullptr)) return AS_EINVAL;   // DAPO: all heads or none
  if (a.upgo_field < 0 || a.upgo_field >= a.F) return AS_EINVAL;
  if (!a.su_step) return AS_EINVAL;
  const long long TB = (long long)a.T * a.B;
  for (int h = 0; h < 6; ++h) {
    if (!a.logit[h] || !a.teacher[h] || !a.action[h] || !a.mu_logp[h] || !a.row_stats[h] || !a.coef[h]) return AS_EINVAL;
    if (h != 5 && a.head_k[h] >= 4096) return AS_ESIZE;     // short heads: one warp per row
  }
  const int nb = (int)((TB + 7) / 8);
  loss_stats_small_kernel<<<4 * nb + (int)TB, 256, 0, st>>>(a, nb);
  AS_LAUNCH_CHECK();
  loss_stats_big_kernel<<<stats_grid(TB, 1), 256, 0, st>>>(a, 5, TB, a.head_k[5]);
  AS_LAUNCH_CHECK();
  const int smem = (72 * a.T + a.F * (4 * a.T + 1) + 6 * a.T) * 4;
  if (smem > 48 * 1024) return AS_ESIZE;
  loss_scan_kernel<<<a.B, kScanThreads, smem, st>>>(a);
  AS_LAUNCH_CHECK();
  loss_finalize_kernel<<<1, 64, 0, st>>>(a);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int loss_backward(const AsLossArgs& a, float gscale, cudaStream_t st) {
  const long long TB = (long long)a.T * a.B;
  for (int h = 0; h < 6; ++h)
    if (!a.dlogit[h]) return AS_EINVAL;
  const int nb = (int)((TB + 7) / 8);
  loss_grad_small_kernel<<<4 * nb + (int)TB, 256, 0, st>>>(a, nb, gscale);
  AS_LAUNCH_CHECK();
  loss_grad_big_kernel<<<stats_grid(TB, 1), 256, 0, st>>>(a, 5, TB, a.head_k[5], gscale);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
