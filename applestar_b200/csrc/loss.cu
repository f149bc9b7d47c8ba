// Fused V-trace / UPGO / TD(lambda) / entropy / teacher-KL loss, forward and backward
// (reference: distar/agent/default/rl_training/rl_loss.py:33-185, as_rl_utils.py).
//
//   K1 loss_row_stats : one pass over target + teacher logits per active row ->
//                       {lse, log pi(a), entropy, lse_teacher, KL(teacher||pi)}            (HBM-bound)
//   K2 loss_scan      : one CTA per trajectory b: stages the [T] columns in shared memory, then a handful of
//                       lanes run the reverse-time scans (V-trace per field x head, UPGO per head, TD(lambda) per
//                       field) and emit per-row backward coefficients, dV and per-b partial sums of every logged
//                       scalar into scratch                                                 (latency-bound, tiny)
//   K3 loss_finalize  : fixed-order reduction over b -> means[], total_loss                 (deterministic)
//   K4 loss_row_grad  : second pass over the logits: dlogit = c_a (onehot - p) - c_e p (log p + H) + c_k (p - p_t)
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
template <int THREADS_PER_ROW>
__global__ void loss_row_stats_kernel(const AsLossArgs a, int h, long long rows, int K) {
  constexpr int RPB = 256 / THREADS_PER_ROW;
  const int sub = threadIdx.x / THREADS_PER_ROW;
  const int tr = threadIdx.x % THREADS_PER_ROW;
  __shared__ float red[8][8];
  for (long long row0 = (long long)blockIdx.x * RPB; row0 < rows; row0 += (long long)gridDim.x * RPB) {
    const long long row = row0 + sub;
    if (THREADS_PER_ROW > 32) __syncthreads();        // protects red[] reuse; row is block-uniform here
    if (row >= rows) continue;
    float* st = a.row_stats[h] + row * kStatStride;
    if (h == 3 && !a.su_mask[row]) {
      if (tr < kStatStride) st[tr] = 0.f;
      continue;
    }
    long long step = 0;
    const bool full = row_active(a, h, row, &step);   // false: only lse / log pi(a) are needed
    const float* x = a.logit[h] + row * (long long)K;
    const float* t = a.teacher[h] + row * (long long)K;
    const float* sc = (full && a.successive[h] != nullptr) ? a.successive[h] + row * (long long)K : nullptr;   // DAPO
    float mx = -INFINITY, mt = -INFINITY, ms = -INFINITY;
    for (int k = tr; k < K; k += THREADS_PER_ROW) {
      mx = fmaxf(mx, __ldg(x + k));
      if (full) mt = fmaxf(mt, __ldg(t + k));
      if (sc) ms = fmaxf(ms, __ldg(sc + k));
    }
    mx = warp_max(mx);
    mt = warp_max(mt);
    ms = warp_max(ms);
    if (THREADS_PER_ROW > 32) {
      const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
      if (l == 0) { red[0][w] = mx; red[1][w] = mt; red[7][w] = ms; }
      __syncthreads();
      mx = red[0][0]; mt = red[1][0]; ms = red[7][0];
#pragma unroll
      for (int i = 1; i < 8; ++i) { mx = fmaxf(mx, red[0][i]); mt = fmaxf(mt, red[1][i]); ms = fmaxf(ms, red[7][i]); }
      __syncthreads();                                  // red[7] is reused below
    }
    // pass 2 (rows come back from L1/L2): s = sum e^(x-mx), w = sum e^(x-mx) x, s2, w2 (teacher), u = sum e^(t-mt) x
    float s = 0.f, w = 0.f, s2 = 0.f, w2 = 0.f, u = 0.f;
    if (full) {
      for (int k = tr; k < K; k += THREADS_PER_ROW) {
        const float xv = __ldg(x + k), tv = __ldg(t + k);
        const float ex = expf(xv - mx), et = expf(tv - mt);
        s += ex; w += ex * xv;
        s2 += et; w2 += et * tv; u += et * xv;
      }
    } else {
      for (int k = tr; k < K; k += THREADS_PER_ROW) s += expf(__ldg(x + k) - mx);
    }
    s = warp_sum(s); w = warp_sum(w); s2 = warp_sum(s2); w2 = warp_sum(w2); u = warp_sum(u);
    if (THREADS_PER_ROW > 32) {
      const int wi = threadIdx.x >> 5, l = threadIdx.x & 31;
      if (l == 0) { red[2][wi] = s; red[3][wi] = w; red[4][wi] = s2; red[5][wi] = w2; red[6][wi] = u; }
      __syncthreads();
      s = w = s2 = w2 = u = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { s += red[2][i]; w += red[3][i]; s2 += red[4][i]; w2 += red[5][i]; u += red[6][i]; }
    }
    // DAPO (as_rl_utils.py:105-127): KL(successive || pi) needs s3 = sum e^(c-ms), w3 = sum e^(c-ms) c, u3 = sum e^(c-ms) x
    float s3 = 0.f, w3 = 0.f, u3 = 0.f;
    if (sc) {
      for (int k = tr; k < K; k += THREADS_PER_ROW) {
        const float cv = __ldg(sc + k);
        const float ec = expf(cv - ms);
        s3 += ec; w3 += ec * cv; u3 += ec * __ldg(x + k);
      }
      s3 = warp_sum(s3); w3 = warp_sum(w3); u3 = warp_sum(u3);
      if (THREADS_PER_ROW > 32) {
        const int wi = threadIdx.x >> 5, l = threadIdx.x & 31;
        __syncthreads();
        if (l == 0) { red[2][wi] = s3; red[3][wi] = w3; red[4][wi] = u3; }
        __syncthreads();
        s3 = w3 = u3 = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) { s3 += red[2][i]; w3 += red[3][i]; u3 += red[4][i]; }
      }
    }
    if (tr == 0) {
      const float lse = mx + logf(s);
      const long long act_idx = a.action[h][row];
      st[0] = lse;
      st[1] = __ldg(x + act_idx) - lse;                         // log pi(a)
      if (full) {
        const float lse_t = mt + logf(s2);
        st[2] = lse - w / s;                                    // entropy  -sum p log p
        st[3] = lse_t;
        st[4] = (w2 / s2 - lse_t) - (u / s2 - lse);             // KL  sum p_t (log p_t - log p)
      } else {
        st[2] = 0.f; st[3] = 0.f; st[4] = 0.f;
      }
      if (sc) {
        const float lse_s = ms + logf(s3);
        st[5] = lse_s;
        st[6] = (w3 / s3 - lse_s) - (u3 / s3 - lse);            // KL  sum p_s (log p_s - log p)
      } else {
        st[5] = 0.f; st[6] = 0.f;
      }
      st[7] = 0.f;
    }
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
  const int T = a.T, B = a.B, F = a.F, S = a.S;
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
  const bool dapo = a.successive[0] != nullptr;
  const float inv_tb = 1.f / (float)(T * B);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = kScanThreads / 32;

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
  // ---- A2: warp per step: sums over the selected-units rows and the pointer-head logit masks
  for (int t = warp; t < T; t += nwarp) {
    const long long step = (long long)t * B + b;
    float lp = 0.f, lr = 0.f, e2 = 0.f, k4 = 0.f, k6 = 0.f, cnt = 0.f, nsu = 0.f, ntu = 0.f;
    for (int j = lane; j < S; j += 32) {
      const long long row = step * S + j;
      if (a.su_mask[row]) {
        const float l = a.row_stats[3][row * kStatStride + 1];
        lp += l;
        lr += l - a.mu_logp[3][row];
        e2 += a.row_stats[3][row * kStatStride + 2];
        k4 += a.row_stats[3][row * kStatStride + 4];
        k6 += a.row_stats[3][row * kStatStride + 6];
        cnt += 1.f;
      }
    }
    for (int j = lane; j < a.N1; j += 32) nsu += a.su_logits_mask[step * a.N1 + j] ? 1.f : 0.f;
    for (int j = lane; j < a.NE; j += 32) ntu += a.tu_logits_mask[step * a.NE + j] ? 1.f : 0.f;
    lp = warp_sum(lp); lr = warp_sum(lr); e2 = warp_sum(e2); k4 = warp_sum(k4); k6 = warp_sum(k6);
    cnt = warp_sum(cnt); nsu = warp_sum(nsu); ntu = warp_sum(ntu);
    if (lane == 0) {
      s_lp[3 * T + t] = lp;
      s_rho[3 * T + t] = fminf(expf(lr), 1.f);
      s_st2[3 * T + t] = e2;
      s_st4[3 * T + t] = k4;
      s_st6[3 * T + t] = k6;
      s_cnt[t] = cnt; s_nlsu[t] = nsu; s_nltu[t] = ntu;
    }
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
      float acc = 0.f;
      float vs_next = v[T];
      for (int t = T - 1; t >= 0; --t) {
        const float delta = rho[t] * (r[t] + v[t + 1] - v[t]);
        const float vs = v[t] + delta + rho[t] * (vs_next - v[t + 1]);
        const float adv = rho[t] * (r[t] + vs_next - v[t]);
        vs_next = vs;
        const float m = hm[t] * fm[t];
        acc += -adv * lp[t] * m;
        c0[t] += -adv * m * a.w_pg[f] * a.hw_pg[h] * inv_tb;
      }
      part[f * 10 + h] = acc;
    }
    {
      const int f = a.upgo_field;
      const float* v = s_v + f * (T + 1);
      const float* r = s_r + f * T;
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
        c0[t] += -adv * hm[t] * a.w_upgo * a.hw_upgo[h] * inv_tb;
      }
      part[base + h] = acc;
    }
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
    // entropy per head (as_rl_utils.py:52-72) and its backward coefficient
    const int h = tid - 64;
    float acc = 0.f;
    for (int t = 0; t < T; ++t) {
      const float hm = s_hm[h * T + t];
      float e;
      if (h == 3) {
        const float norm = 1e-9f + logf(s_nlsu[t] + 1.f);
        const float denom = s_cnt[t] + 1e-9f;
        e = s_st2[3 * T + t] / norm / denom;
        s_c1[3 * T + t] = -a.w_ent * a.hw_ent[3] * hm * inv_tb / (norm * denom);
      } else {
        const float norm = (h == 4) ? 1e-9f + logf(s_nltu[t] + 1.f) : logf((float)a.head_k[h]);
        e = s_st2[h * T + t] / norm;
        s_c1[h * T + t] = -a.w_ent * a.hw_ent[h] * hm * inv_tb / norm;
      }
      acc += e * hm;
    }
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
    a.coef[h][step * kCoefStride + 0] = s_c0[i];
    if (h != 3) {
      a.coef[h][step * kCoefStride + 1] = s_c1[i];
      a.coef[h][step * kCoefStride + 2] = s_c2[i];
      a.coef[h][step * kCoefStride + 3] = s_c3[i];
    }
  }
  for (int i = tid; i < T * S; i += kScanThreads) {
    const int t = i / S, j = i - t * S;
    const long long row = ((long long)t * B + b) * S + j;
    const bool on = a.su_mask[row];
    a.coef[3][row * kCoefStride + 1] = on ? s_c1[3 * T + t] : 0.f;
    a.coef[3][row * kCoefStride + 2] = on ? s_c2[3 * T + t] : 0.f;
    a.coef[3][row * kCoefStride + 3] = on ? s_c3[3 * T + t] : 0.f;
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
template <int THREADS_PER_ROW>
__global__ void loss_row_grad_kernel(const AsLossArgs a, int h, long long rows, int K, float gscale) {
  constexpr int RPB = 256 / THREADS_PER_ROW;
  if (a.grad_scale != nullptr) gscale *= __ldg(a.grad_scale);
  const int sub = threadIdx.x / THREADS_PER_ROW;
  const int tr = threadIdx.x % THREADS_PER_ROW;
  for (long long row0 = (long long)blockIdx.x * RPB; row0 < rows; row0 += (long long)gridDim.x * RPB) {
    const long long row = row0 + sub;
    if (row >= rows) continue;
    long long step = 0;
    const bool act = row_active(a, h, row, &step);
    float* dx = a.dlogit[h] + row * (long long)K;
    if (!act) {
      for (int k = tr; k < K; k += THREADS_PER_ROW) dx[k] = 0.f;
      continue;
    }
    const float* st = a.row_stats[h] + row * kStatStride;
    const float lse = st[0], ent = st[2], lse_t = st[3], lse_s = st[5];
    const float c_a = a.coef[h][step * kCoefStride + 0] * gscale;
    const float* cf = a.coef[h] + row * kCoefStride;
    const float c_e = cf[1] * gscale, c_k = cf[2] * gscale, c_d = cf[3] * gscale;
    const float* x = a.logit[h] + row * (long long)K;
    const float* t = a.teacher[h] + row * (long long)K;
    const float* sc = (a.successive[h] != nullptr && c_d != 0.f) ? a.successive[h] + row * (long long)K : nullptr;
    const int act_idx = (int)a.action[h][row];
    for (int k = tr; k < K; k += THREADS_PER_ROW) {
      const float lp = __ldg(x + k) - lse;
      const float p = expf(lp);
      const float pt = expf(__ldg(t + k) - lse_t);
      float g = -c_a * p - c_e * p * (lp + ent) + c_k * (p - pt);
      if (sc) g += c_d * (p - expf(__ldg(sc + k) - lse_s));
      if (k == act_idx) g += c_a;
      dx[k] = g;
    }
  }
}

// ---------------------------------------------------------------------------------------------- host
static int stats_grid(long long rows, int rpb) {
  long long blocks = (rows + rpb - 1) / rpb;
  const long long cap = 148LL * 8;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

int loss_forward(const AsLossArgs& a, cudaStream_t st) {
  if (a.T <= 0 || a.B <= 0 || a.F <= 0 || a.F > AS_LOSS_MAX_FIELDS || a.S <= 0) return AS_ESIZE;
  if (a.F * 10 + 26 >= AS_LOSS_MAX_STATS) return AS_ESIZE;
  for (int h = 1; h < 6; ++h)
    if ((a.successive[h] != nullptr) != (a.successive[0] != nullptr)) return AS_EINVAL;   // DAPO: all heads or none
  if (a.upgo_field < 0 || a.upgo_field >= a.F) return AS_EINVAL;
  const long long TB = (long long)a.T * a.B;
  for (int h = 0; h < 6; ++h) {
    if (!a.logit[h] || !a.teacher[h] || !a.action[h] || !a.mu_logp[h] || !a.row_stats[h] || !a.coef[h]) return AS_EINVAL;
    const long long rows = (h == 3) ? TB * a.S : TB;
    const int K = a.head_k[h];
    if (K >= 4096) loss_row_stats_kernel<256><<<stats_grid(rows, 1), 256, 0, st>>>(a, h, rows, K);
    else loss_row_stats_kernel<32><<<stats_grid(rows, 8), 256, 0, st>>>(a, h, rows, K);
    AS_LAUNCH_CHECK();
  }
  const int smem = (60 * a.T + a.F * (4 * a.T + 1) + 6 * a.T) * 4;
  if (smem > 48 * 1024) return AS_ESIZE;
  loss_scan_kernel<<<a.B, kScanThreads, smem, st>>>(a);
  AS_LAUNCH_CHECK();
  loss_finalize_kernel<<<1, 64, 0, st>>>(a);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int loss_backward(const AsLossArgs& a, float gscale, cudaStream_t st) {
  const long long TB = (long long)a.T * a.B;
  for (int h = 0; h < 6; ++h) {
    if (!a.dlogit[h]) return AS_EINVAL;
    const long long rows = (h == 3) ? TB * a.S : TB;
    const int K = a.head_k[h];
    if (K >= 4096) loss_row_grad_kernel<256><<<stats_grid(rows, 1), 256, 0, st>>>(a, h, rows, K, gscale);
    else loss_row_grad_kernel<32><<<stats_grid(rows, 8), 256, 0, st>>>(a, h, rows, K, gscale);
    AS_LAUNCH_CHECK();
  }
  return AS_OK;
}

}  // namespace as
