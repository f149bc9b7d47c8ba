// SelectedUnits sampling loop (head/action_arg_head.py:262-313 of the reference) as ONE kernel: all 64 autoregressive
// pointer steps of a row run inside a CTA, with no launch between steps.  The eager loop is ~40 launches per step.
//
// Per step the reference computes, for every row,
//     q   = query_fc2(relu(query_fc1(ae)))            ae = ae0 + embed_fc2(relu(embed_fc1(emb)))   (ae = ae0 at step 0)
//     h,c = LN-LSTM(32)(q, h, c)
//     l_n = <h, key_n>  masked to -1e9, / temperature ; pick = argmax(l + gumbel) ; bookkeeping ; emb = mean picked keys
// embed_fc2 and query_fc1 are both affine with no activation in between, so the 1024-wide autoregressive embedding
// never has to be formed inside the loop:
//     query_fc1(ae) = [Wq1 ae0 + bq1] + [Wq1 bf2] + (Wq1 Wf2) relu(embed_fc1(emb)) = c0 + c1 + M e1
// with c0 [R,256] one GEMM before the loop and M [256,256], c1 [256] functions of the weights only.  A step is then
// 32->256, 256->256, 256->32 mat-vecs, the H=32 cell and a 513 x 32 pointer product: all fp32, M read from L2,
// the small matrices resident in shared memory.  Rows are independent (the reference's early `break` only trims the
// step axis, done by the caller from steps_row), so there is no inter-CTA communication.
#include "common.cuh"
#include "su_sample.h"

namespace as {

constexpr int kSuQ = 256, kSuD = 32, kSuG = 128, kSuMaxWords = 64;

__device__ __forceinline__ float su_sgm(float x) { return 1.f / (1.f + expf(-x)); }

template <int NR>
__global__ void __launch_bounds__(256, 1) su_sample_kernel(const SuSampleDev p) {
  extern __shared__ __align__(16) float smem[];
  float* sWf1t = smem;                         // 32*256
  float* sWq2t = sWf1t + kSuD * kSuQ;          // 256*32
  float* sWih = sWq2t + kSuQ * kSuD;           // 32*128
  float* sWhh = sWih + kSuD * kSuG;            // 32*128
  float* sPart = sWhh + kSuD * kSuG;           // 4*NR*256
  float* sE1 = sPart + 4 * NR * kSuQ;          // NR*256
  float* sQ1 = sE1 + NR * kSuQ;                // NR*256
  float* sP2 = sQ1 + NR * kSuQ;                // 8*NR*32
  float* sEmb = sP2 + 8 * NR * kSuD;           // NR*32
  float* sH = sEmb + NR * kSuD;                // NR*32
  __shared__ unsigned sLm[NR][kSuMaxWords], sOh[NR][kSuMaxWords];
  __shared__ float sBv[NR][8];
  __shared__ int sBi[NR][8];
  __shared__ float sEndLogit[NR];
  __shared__ int sStop;
  __shared__ int sEnded[NR];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int row0 = blockIdx.x * NR;
  const int N1 = p.N1, S = p.S;
  const int words = (N1 + 31) >> 5;

  for (int i = tid; i < kSuD * kSuQ; i += 256) { sWf1t[i] = p.wf1t[i]; sWq2t[i] = p.wq2t[i]; }
  for (int i = tid; i < kSuD * kSuG; i += 256) { sWih[i] = p.wih_t[i]; sWhh[i] = p.whh_t[i]; }
  for (int i = tid; i < NR * kSuD; i += 256) sEmb[i] = 0.f;
  for (int i = tid; i < NR * kSuMaxWords; i += 256) {
    const int r = i / kSuMaxWords, w = i - r * kSuMaxWords;
    const int row = row0 + r;
    unsigned m = 0;
    if (row < p.R && w < words) {
      const long long en = p.entity_num[row];            // valid keys n < entity_num; the end token joins at step 1
      const long long lo = (long long)w * 32;
      if (en >= lo + 32) m = 0xffffffffu;
      else if (en > lo) m = (1u << (int)(en - lo)) - 1u;
    }
    sLm[r][w] = m;
    sOh[r][w] = 0u;
  }
  float c0r[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) c0r[r] = (row0 + r < p.R) ? p.c0[(long long)(row0 + r) * kSuQ + tid] : 0.f;
  const float c1r = p.c1[tid], bf1r = p.bf1[tid];
  // per-row state owned by warp r (lane = hidden unit / key column)
  float h = 0.f, c = 0.f, ksum = 0.f;
  int cnt = 0, ended = 0, num = 0, steps_row = 0;
  long long en_w = 0;
  if (warp < NR && row0 + warp < p.R) {
    en_w = p.entity_num[row0 + warp];
    ended = p.su_mask[row0 + warp] ? 0 : 1;
    num = ended ? 0 : S;
    steps_row = ended ? 0 : S;
    if (ended && lane == 0 && p.ended_rows) atomicAdd(p.ended_rows, 1);
  }
  if (tid == 0) sStop = 0;
  float lniw[4], lnib[4], lnhw[4], lnhb[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    lniw[g] = p.ln_i_w[g * kSuD + lane]; lnib[g] = p.ln_i_b[g * kSuD + lane];
    lnhw[g] = p.ln_h_w[g * kSuD + lane]; lnhb[g] = p.ln_h_b[g * kSuD + lane];
  }
  const float lncw = p.ln_c_w[lane], lncb = p.ln_c_b[lane], bq2r = p.bq2[lane];
  const bool use_temp = p.temperature != 1.0f;
  __syncthreads();

  for (int step = 0; step < S; ++step) {
    // every row of the batch has raised its end flag and this CTA is past the latest of them (the reference's early
    // break, :309-310): the remaining steps would be trimmed by the caller anyway.  sStop was written before the barrier that closed the previous step.
    if (sStop) break;
    // ---- query_fc1 pre-activation: c0 (+ c1 + M e1 from step 1 on)
    float pre[NR];
#pragma unroll
    for (int r = 0; r < NR; ++r) pre[r] = c0r[r];
    if (step > 0) {
      float e[NR];
#pragma unroll
      for (int r = 0; r < NR; ++r) e[r] = bf1r;
#pragma unroll 8
      for (int k = 0; k < kSuD; ++k) {
        const float w = sWf1t[k * kSuQ + tid];
#pragma unroll
        for (int r = 0; r < NR; ++r) e[r] = fmaf(w, sEmb[r * kSuD + k], e[r]);
      }
#pragma unroll
      for (int r = 0; r < NR; ++r) sE1[r * kSuQ + tid] = fmaxf(e[r], 0.f);
      __syncthreads();
      const int jg = tid & 63, ks = tid >> 6;
      float4 acc[NR];
#pragma unroll
      for (int r = 0; r < NR; ++r) acc[r] = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4* mrow = reinterpret_cast<const float4*>(p.mt) + (long long)(ks * 64) * (kSuQ / 4) + jg;
#pragma unroll 8
      for (int kk = 0; kk < 64; ++kk) {
        const float4 m = __ldg(mrow + kk * (kSuQ / 4));
#pragma unroll
        for (int r = 0; r < NR; ++r) {
          const float ev = sE1[r * kSuQ + ks * 64 + kk];
          acc[r].x = fmaf(m.x, ev, acc[r].x); acc[r].y = fmaf(m.y, ev, acc[r].y);
          acc[r].z = fmaf(m.z, ev, acc[r].z); acc[r].w = fmaf(m.w, ev, acc[r].w);
        }
      }
#pragma unroll
      for (int r = 0; r < NR; ++r) reinterpret_cast<float4*>(sPart + (ks * NR + r) * kSuQ)[jg] = acc[r];
      __syncthreads();
#pragma unroll
      for (int r = 0; r < NR; ++r)
        pre[r] += c1r + ((sPart[(0 * NR + r) * kSuQ + tid] + sPart[(1 * NR + r) * kSuQ + tid]) +
                         (sPart[(2 * NR + r) * kSuQ + tid] + sPart[(3 * NR + r) * kSuQ + tid]));
    }
#pragma unroll
    for (int r = 0; r < NR; ++r) sQ1[r * kSuQ + tid] = fmaxf(pre[r], 0.f);
    __syncthreads();
    // ---- query_fc2: 8 slices of 32 inputs, output column = lane
    {
      float a2[NR];
#pragma unroll
      for (int r = 0; r < NR; ++r) a2[r] = 0.f;
#pragma unroll 8
      for (int jj = 0; jj < 32; ++jj) {
        const int j = warp * 32 + jj;
        const float w = sWq2t[j * kSuD + lane];
#pragma unroll
        for (int r = 0; r < NR; ++r) a2[r] = fmaf(w, sQ1[r * kSuQ + j], a2[r]);
      }
#pragma unroll
      for (int r = 0; r < NR; ++r) sP2[(warp * NR + r) * kSuD + lane] = a2[r];
    }
    __syncthreads();
    // ---- LN-LSTM cell (lstm.py:138-153), one warp per row
    if (warp < NR) {
      float x = bq2r;
#pragma unroll
      for (int s = 0; s < 8; ++s) x += sP2[(s * NR + warp) * kSuD + lane];
      float gi[4] = {0.f, 0.f, 0.f, 0.f}, gh[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
      for (int k = 0; k < kSuD; ++k) {
        const float xk = __shfl_sync(0xffffffffu, x, k), hk = __shfl_sync(0xffffffffu, h, k);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          gi[g] = fmaf(sWih[k * kSuG + g * kSuD + lane], xk, gi[g]);
          gh[g] = fmaf(sWhh[k * kSuG + g * kSuD + lane], hk, gh[g]);
        }
      }
      const float mi = warp_sum(gi[0] + gi[1] + gi[2] + gi[3]) * (1.f / kSuG);
      const float mh = warp_sum(gh[0] + gh[1] + gh[2] + gh[3]) * (1.f / kSuG);
      float qi = 0.f, qh = 0.f;
#pragma unroll
      for (int g = 0; g < 4; ++g) { const float di = gi[g] - mi, dh = gh[g] - mh; qi += di * di; qh += dh * dh; }
      const float ri = rsqrtf(warp_sum(qi) * (1.f / kSuG) + 1e-5f), rh = rsqrtf(warp_sum(qh) * (1.f / kSuG) + 1e-5f);
      float a[4];
#pragma unroll
      for (int g = 0; g < 4; ++g)
        a[g] = ((gi[g] - mi) * ri * lniw[g] + lnib[g]) + ((gh[g] - mh) * rh * lnhw[g] + lnhb[g]);
      const float z = su_sgm(a[1]) * c + su_sgm(a[0]) * tanhf(a[2]);
      const float mc = warp_sum(z) * (1.f / kSuD);
      const float dz = z - mc;
      const float rc = rsqrtf(warp_sum(dz * dz) * (1.f / kSuD) + 1e-5f);
      c = dz * rc * lncw + lncb;
      h = su_sgm(a[3]) * tanhf(c);
      sH[warp * kSuD + lane] = h;
    }
    __syncthreads();
    // ---- pointer logits, mask, store, Gumbel arg-max (lowest index on ties)
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      const int row = row0 + r;
      float best = -INFINITY;
      int idx = 0x7fffffff;
      if (row < p.R) {
        float q[kSuD];
#pragma unroll
        for (int k = 0; k < kSuD; ++k) q[k] = sH[r * kSuD + k];
        const float* keyr = p.key + (long long)row * N1 * kSuD;
        float* out = p.logits + ((long long)row * S + step) * N1;
        const float* nz = p.noise + ((long long)step * p.R + row) * N1;
        for (int n = tid; n < N1; n += 256) {
          const float4* kp = reinterpret_cast<const float4*>(keyr + (long long)n * kSuD);
          float d = 0.f;
#pragma unroll
          for (int k4 = 0; k4 < kSuD / 4; ++k4) {
            const float4 kv = __ldg(kp + k4);
            d = fmaf(q[4 * k4], kv.x, d); d = fmaf(q[4 * k4 + 1], kv.y, d);
            d = fmaf(q[4 * k4 + 2], kv.z, d); d = fmaf(q[4 * k4 + 3], kv.w, d);
          }
          float v = ((sLm[r][n >> 5] >> (n & 31)) & 1u) ? d : -1e9f;
          if (use_temp) v = v / p.temperature;
          out[n] = v;
          if (step == S - 1 && n == (int)p.entity_num[row]) sEndLogit[r] = v;
          const float s = v + __ldg(nz + n);
          if (s > best || (s == best && n < idx)) { best = s; idx = n; }
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov > best || (ov == best && oi < idx)) { best = ov; idx = oi; }
      }
      if (lane == 0) { sBv[r][warp] = best; sBi[r][warp] = idx; }
    }
    __syncthreads();
    // ---- bookkeeping (:279-301), warp r owns row r
    if (warp < NR && row0 + warp < p.R) {
      const int r = warp, row = row0 + r;
      float best = sBv[r][0];
      int idx = sBi[r][0];
#pragma unroll
      for (int w = 1; w < 8; ++w)
        if (sBv[r][w] > best || (sBv[r][w] == best && sBi[r][w] < idx)) { best = sBv[r][w]; idx = sBi[r][w]; }
      if (idx == 0x7fffffff) idx = 0;
      const bool hit = (long long)idx == en_w;
      if (hit && !ended) {
        num = step + 1; steps_row = step + 1;
        if (lane == 0 && p.ended_rows) {                 // latest end step first, then the count that publishes it
          atomicMax(p.ended_rows + 1, step + 1);
          __threadfence();
          atomicAdd(p.ended_rows, 1);
        }
      }
      ended |= hit ? 1 : 0;
      if (!ended) {
        const unsigned bit = 1u << (idx & 31);
        const unsigned old = sOh[r][idx >> 5];                        // same value in every lane (read before the write)
        __syncwarp();
        if (!(old & bit)) {
          if (lane == 0) sOh[r][idx >> 5] = old | bit;
          cnt += 1;
          ksum += __ldg(p.key + ((long long)row * N1 + idx) * kSuD + lane);
        }
      }
      sEmb[r * kSuD + lane] = cnt ? ksum / (float)cnt : ksum;
      if (lane == 0) {
        p.units[(long long)row * S + step] = idx;
        if (step == 0 && en_w < N1) sLm[r][en_w >> 5] |= 1u << (en_w & 31);     // end token selectable from step 1
        sLm[r][idx >> 5] &= ~(1u << (idx & 31));                            // a picked unit is masked afterwards
      }
    }
    if (tid == 0 && p.ended_rows) {
      // CTAs are not in step with each other: stop only when every row has ended AND this CTA has computed as many
      // steps as the latest of them needed (the caller keeps steps [0, max steps_row) of EVERY row)
      const int cnt = *reinterpret_cast<volatile int*>(p.ended_rows);
      __threadfence();
      const int last = *reinterpret_cast<volatile int*>(p.ended_rows + 1);
      sStop = cnt >= p.R && step + 1 >= last;
    }
    __syncthreads();
  }

  // ---- outputs after the loop
  if (warp < NR && row0 + warp < p.R) {
    const int row = row0 + warp;
    p.emb[(long long)row * kSuD + lane] = sEmb[warp * kSuD + lane];
    if (lane == 0) { p.num[row] = num; p.steps_row[row] = steps_row; sEnded[warp] = ended; }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    const int row = row0 + r;
    if (row >= p.R) continue;
    const float* last = p.logits + ((long long)row * S + (S - 1)) * N1;    // each thread re-reads its own stores
    const bool on = p.extra_units && !sEnded[r];
    const float endl = sEndLogit[r];
    for (int n = tid; n < N1; n += 256) p.extra[(long long)row * N1 + n] = (on && last[n] > endl) ? 1.f : 0.f;
  }
}

template <int NR>
static int su_sample_launch(const SuSampleDev& p, cudaStream_t st) {
  const size_t smem = sizeof(float) * (2 * kSuD * kSuQ + 2 * kSuD * kSuG + 4 * NR * kSuQ + 2 * NR * kSuQ + 8 * NR * kSuD +
                                       2 * NR * kSuD);
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(su_sample_kernel<NR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    attr_done = true;
  }
  const int grid = (p.R + NR - 1) / NR;
  su_sample_kernel<NR><<<grid, 256, smem, st>>>(p);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int su_sample(const SuSampleDev& p, cudaStream_t st) {
  if (p.R <= 0) return AS_OK;
  if (p.N1 <= 0 || p.N1 > 32 * kSuMaxWords || p.S <= 0) return AS_ESIZE;
  if (!p.key || !p.c0 || !p.c1 || !p.mt || !p.wf1t || !p.bf1 || !p.wq2t || !p.bq2 || !p.wih_t || !p.whh_t || !p.ln_i_w ||
      !p.ln_i_b || !p.ln_h_w || !p.ln_h_b || !p.ln_c_w || !p.ln_c_b || !p.entity_num || !p.su_mask || !p.noise || !p.logits ||
      !p.units || !p.num || !p.extra || !p.emb || !p.steps_row)
    return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(p.key) | reinterpret_cast<uintptr_t>(p.mt)) & 15) return AS_EALIGN;
  return p.R > 148 ? su_sample_launch<2>(p, st) : su_sample_launch<1>(p, st);
}

}  // namespace as
