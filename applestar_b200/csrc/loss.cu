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

constexpr int kStatStride = 8;   // lse, logp_a, ent, lse_t, kl, (pad)
constexpr int kCoefStride = 4;   // c_a, c_e, c_k, (pad)

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
    float mx = -INFINITY, mt = -INFINITY;
    for (int k = tr; k < K; k += THREADS_PER_ROW) {
      mx = fmaxf(mx, __ldg(x + k));
      if (full) mt = fmaxf(mt, __ldg(t + k));
    }
    mx = warp_max(mx);
    mt = warp_max(mt);
    if (THREADS_PER_ROW > 32) {
      const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
      if (l == 0) { red[0][w] = mx; red[1][w] = mt; }
      __syncthreads();
      mx = red[0][0]; mt = red[1][0];
#pragma unroll
      for (int i = 1; i < 8; ++i) { mx = fmaxf(mx, red[0][i]); mt = fmaxf(mt, red[1][i]); }
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
      st[5] = 0.f; st[6] = 0.f; st[7] = 0.f;
    }
  }
}

// ---------------------------------------------------------------------------------------------- K2
// Stat slots (per-b partial sums, later divided by T*B):
//   [f*10 + h]      pg loss of field f, head h (h<6);  [f*10+6] td;  [f*10+7] reward;  [f*10+8] value (mean over T+1)
//   base = F*10:    [base + h] upgo h;  [base+6+h] entropy h;  [base+12+h] kl h;  [base+18] kl extra_at;
//                   [base+19] battle reward
__device__ __forceinline__ float head_mask_at(const AsLossArgs& a, int h, long long step) {
  return a.actions_mask[h] ? (float)a.actions_mask[h][step] : 1.f;
}

__global__ void __launch_bounds__(64) loss_scan_kernel(const AsLossArgs a) {
  const int b = blockIdx.x;
  const int T = a.T, B = a.B, F = a.F;
  extern __shared__ float sm[];
  // shared columns: logp_a[6][T], mu[6][T], value[F][T+1], reward[F][T]
  float* s_lp = sm;
  float* s_rho = s_lp + 6 * T;
  float* s_v = s_rho + 6 * T;
  float* s_r = s_v + F * (T + 1);
  const float inv_tb = 1.f / (float)(T * B);

  for (int t = threadIdx.x; t < T; t += blockDim.x) {
    const long long step = (long long)t * B + b;
    for (int h = 0; h < 6; ++h) {
      float lp, lr;
      if (h == 3) {
        lp = 0.f; lr = 0.f;
        for (int j = 0; j < a.S; ++j) {
          const long long row = step * a.S + j;
          if (a.su_mask[row]) {
            const float l = a.row_stats[3][row * kStatStride + 1];
            lp += l;
            lr += l - a.mu_logp[3][row];
          }
        }
      } else {
        lp = a.row_stats[h][step * kStatStride + 1];
        lr = lp - a.mu_logp[h][step];
      }
      s_lp[h * T + t] = lp;
      s_rho[h * T + t] = fminf(expf(lr), 1.f);
    }
    for (int f = 0; f < F; ++f) s_r[f * T + t] = a.reward[f][step];
  }
  for (int t = threadIdx.x; t <= T; t += blockDim.x) {
    for (int f = 0; f < F; ++f) {
      float v = a.value[f][(long long)t * B + b];
      if (t == T) v *= (a.winloss_reward[(long long)(T - 1) * B + b] == 0.f) ? 1.f : 0.f;   // rl_loss.py:47-49
      s_v[f * (T + 1) + t] = v;
    }
  }
  __syncthreads();

  float* part = a.partial + (long long)b * AS_LOSS_MAX_STATS;
  const int base = F * 10;
  const int lane = threadIdx.x;
  if (lane < 6) {
    // ---- head job: V-trace per field (gamma = lambda = 1, rl_loss.py:102-103) + UPGO on winloss
    const int h = lane;
    const float* rho = s_rho + h * T;
    const float* lp = s_lp + h * T;
    for (int t = 0; t < T; ++t) a.coef[h][((long long)t * B + b) * kCoefStride + 0] = 0.f;
    for (int f = 0; f < F; ++f) {
      const float* v = s_v + f * (T + 1);
      const float* r = s_r + f * T;
      float acc = 0.f;
      float vs_next = v[T];
      for (int t = T - 1; t >= 0; --t) {
        const float delta = rho[t] * (r[t] + v[t + 1] - v[t]);
        const float vs = v[t] + delta + rho[t] * (vs_next - v[t + 1]);
        const float adv = rho[t] * (r[t] + vs_next - v[t]);
        vs_next = vs;
        const long long step = (long long)t * B + b;
        float m = head_mask_at(a, h, step);
        if (a.field_mask[f]) m *= a.field_mask[f][step];
        acc += -adv * lp[t] * m;
        a.coef[h][step * kCoefStride + 0] += -adv * m * a.w_pg[f] * a.hw_pg[h] * inv_tb;
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
        const long long step = (long long)t * B + b;
        const float m = head_mask_at(a, h, step);
        acc += -adv * lp[t] * m;
        a.coef[h][step * kCoefStride + 0] += -adv * m * a.w_upgo * a.hw_upgo[h] * inv_tb;
      }
      part[base + h] = acc;
    }
  } else if (lane < 6 + F) {
    // ---- TD(lambda = 0.8) per field (as_rl_utils.py:221-243)
    const int f = lane - 6;
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
      const long long step = (long long)t * B + b;
      const float m = a.field_mask[f] ? a.field_mask[f][step] : 1.f;
      const float diff = g - v[t];
      acc += 0.5f * diff * diff * m;
      racc += r[t];
      vacc += v[t];
      a.dvalue[f][step] = -diff * m * a.w_td[f] * inv_tb;
    }
    a.dvalue[f][(long long)T * B + b] = 0.f;
    part[f * 10 + 6] = acc;
    part[f * 10 + 7] = racc;
    part[f * 10 + 8] = vacc * ((float)T / (float)(T + 1));      // finalize multiplies by 1/(T*B); mean is over (T+1)*B
    part[f * 10 + 9] = 0.f;
  } else if (lane >= 16 && lane < 22) {
    // ---- entropy per head (as_rl_utils.py:52-72) and its backward coefficient
    const int h = lane - 16;
    float acc = 0.f;
    for (int t = 0; t < T; ++t) {
      const long long step = (long long)t * B + b;
      const float hm = head_mask_at(a, h, step);
      float e = 0.f;
      if (h == 3) {
        float cnt = 0.f, nl = 0.f;
        for (int j = 0; j < a.N1; ++j) nl += a.su_logits_mask[step * a.N1 + j] ? 1.f : 0.f;
        for (int j = 0; j < a.S; ++j) cnt += a.su_mask[step * a.S + j] ? 1.f : 0.f;
        const float norm = 1e-9f + logf(nl + 1.f);
        const float denom = cnt + 1e-9f;
        for (int j = 0; j < a.S; ++j) {
          const long long row = step * a.S + j;
          const bool on = a.su_mask[row];
          if (on) e += a.row_stats[3][row * kStatStride + 2] / norm;
          a.coef[3][row * kCoefStride + 1] = on ? (-a.w_ent * a.hw_ent[3] * hm * inv_tb / (norm * denom)) : 0.f;
        }
        e /= denom;
      } else {
        float norm;
        if (h == 4) {
          float nl = 0.f;
          for (int j = 0; j < a.NE; ++j) nl += a.tu_logits_mask[step * a.NE + j] ? 1.f : 0.f;
          norm = 1e-9f + logf(nl + 1.f);
        } else {
          norm = logf((float)a.head_k[h]);
        }
        e = a.row_stats[h][step * kStatStride + 2] / norm;
        a.coef[h][step * kCoefStride + 1] = -a.w_ent * a.hw_ent[h] * hm * inv_tb / norm;
      }
      acc += e * hm;
    }
    part[base + 6 + h] = acc;
  } else if (lane >= 24 && lane < 30) {
    // ---- KL(teacher || pi) per head (as_rl_utils.py:75-103) and its backward coefficient
    const int h = lane - 24;
    float acc = 0.f, acc_at = 0.f;
    for (int t = 0; t < T; ++t) {
      const long long step = (long long)t * B + b;
      const float hm = head_mask_at(a, h, step);
      float ck = a.w_kl * a.hw_kl[h] * hm * inv_tb;
      float kl = 0.f;
      if (h == 3) {
        for (int j = 0; j < a.S; ++j) {
          const long long row = step * a.S + j;
          const bool on = a.su_mask[row];
          if (on) kl += a.row_stats[3][row * kStatStride + 4];
          a.coef[3][row * kCoefStride + 2] = on ? ck : 0.f;
        }
      } else {
        kl = a.row_stats[h][step * kStatStride + 4];
        if (h == 0) {
          const float flag = (a.step[step] < a.at_kl_steps) ? 1.f : 0.f;
          const float m = flag * a.cum_action_mask[step];
          acc_at += kl * m;
          ck += a.w_atkl * m * inv_tb;
        }
        a.coef[h][step * kCoefStride + 2] = ck;
      }
      acc += kl * hm;
    }
    part[base + 12 + h] = acc;
    if (h == 0) part[base + 18] = acc_at;
  } else if (lane == 30) {
    float acc = 0.f;
    for (int t = 0; t < T; ++t) acc += a.battle_reward[(long long)t * B + b];
    part[base + 19] = acc;
  }
}

// ---------------------------------------------------------------------------------------------- K3
__global__ void loss_finalize_kernel(const AsLossArgs a) {
  const int n = a.F * 10 + 20;
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
    const float lse = st[0], ent = st[2], lse_t = st[3];
    const float c_a = a.coef[h][step * kCoefStride + 0] * gscale;
    const float* cf = a.coef[h] + row * kCoefStride;
    const float c_e = cf[1] * gscale, c_k = cf[2] * gscale;
    const float* x = a.logit[h] + row * (long long)K;
    const float* t = a.teacher[h] + row * (long long)K;
    const int act_idx = (int)a.action[h][row];
    for (int k = tr; k < K; k += THREADS_PER_ROW) {
      const float lp = __ldg(x + k) - lse;
      const float p = expf(lp);
      const float pt = expf(__ldg(t + k) - lse_t);
      float g = -c_a * p - c_e * p * (lp + ent) + c_k * (p - pt);
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
  if (a.F * 10 + 20 >= AS_LOSS_MAX_STATS) return AS_ESIZE;
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
  const int smem = (12 * a.T + a.F * (2 * a.T + 1)) * 4;
  if (smem > 48 * 1024) return AS_ESIZE;
  loss_scan_kernel<<<a.B, 64, smem, st>>>(a);
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
