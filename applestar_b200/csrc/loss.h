/* Argument block of the fused RL-loss kernels (plain C; also included by include/applestar_b200.h).
 * Head order: 0 action_type, 1 delay, 2 queued, 3 selected_units, 4 target_unit, 5 target_location.
 * All tensors are contiguous; "step" = t*B + b; selected_units rows = step*S + j (S = 64). */
#ifndef APPLESTAR_B200_LOSS_H
#define APPLESTAR_B200_LOSS_H
#include <stdint.h>

#define AS_LOSS_MAX_FIELDS 6
#define AS_LOSS_MAX_STATS 96   /* out[0 .. F*10+26) = means, out[AS_LOSS_MAX_STATS-1] = total_loss */

typedef struct AsLossArgs {
  int32_t T, B, F, S;            /* unroll length, trajectories, enabled value fields, selected-units rows (64) */
  int32_t N1, NE;                /* widths of the selected_units (N+1) and target_unit (N) logits */
  int32_t head_k[6];             /* logits per row of each head */
  int32_t upgo_field;            /* index of 'winloss' among the F fields */
  float at_kl_steps;             /* extra action-type KL while step < at_kl_steps */
  float dapo_steps;              /* DAPO KL while step < dapo_steps */
  const float* logit[6];         /* target policy logits  [rows, K] fp32 */
  const float* teacher[6];       /* teacher logits        [rows, K] fp32 */
  const float* successive[6];    /* DAPO: logits of the successive policy [rows, K] fp32, or all NULL (use_dapo off) */
  float* dlogit[6];              /* out: d total_loss / d logit */
  const int64_t* action[6];      /* taken action per row */
  const float* mu_logp[6];       /* behaviour log-prob per row */
  const int64_t* actions_mask[6];/* per step, NULL for action_type / delay */
  const uint8_t* su_mask;        /* bool [T*B, S] */
  const uint8_t* su_logits_mask; /* bool [T*B, N1] */
  const uint8_t* tu_logits_mask; /* bool [T*B, NE] */
  const float* cum_action_mask;  /* [T*B] */
  const float* step;             /* game step [T*B] */
  const float* value[AS_LOSS_MAX_FIELDS];      /* [T+1, B] */
  float* dvalue[AS_LOSS_MAX_FIELDS];           /* out [T+1, B] */
  const float* reward[AS_LOSS_MAX_FIELDS];     /* [T, B] */
  const float* field_mask[AS_LOSS_MAX_FIELDS]; /* [T*B] or NULL */
  const float* winloss_reward;   /* [T, B]: value[T] is zeroed where winloss_reward[T-1] != 0 */
  const float* battle_reward;    /* [T, B] (logged only) */
  float gamma_td[AS_LOSS_MAX_FIELDS];
  float w_pg[AS_LOSS_MAX_FIELDS], w_td[AS_LOSS_MAX_FIELDS];
  float hw_pg[6], hw_upgo[6], hw_ent[6], hw_kl[6], hw_dapo[6];
  float w_upgo, w_ent, w_kl, w_atkl, w_dapo;
  const float* stat_weight;      /* [AS_LOSS_MAX_STATS]: total_loss = sum_i stat_weight[i] * means[i] */
  const float* grad_scale;       /* backward only: device scalar d(objective)/d(total_loss) multiplied into dlogit, or NULL (= 1) */
  /* scratch (caller-provided device memory) */
  float* row_stats[6];           /* [rows, 8]; rows = T*B, selected units T*B*S */
  float* coef[6];                /* [T*B, 4] per head (selected units: per step as well) */
  float* su_step;                /* [T*B, 8] per-step sums over the selected-units sub-rows */
  float* partial;                /* [B, AS_LOSS_MAX_STATS] */
  float* out;                    /* [AS_LOSS_MAX_STATS] */
} AsLossArgs;

#endif
