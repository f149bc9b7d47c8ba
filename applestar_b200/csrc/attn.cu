// Fused attention forward for the entity Transformer (module_utils.py:71-111 of the reference: softmax(Q K^T / sqrt(d),
// padded keys filled with -1e9) V, 2 heads x 128, up to 512 keys): scores never leave the SM.
//
// One work item = 128 queries of one (observation, head).  All keys of the row fit the tensor memory at once
// (4 blocks x 128 fp32 columns = the full 512 TMEM columns), so the softmax is EXACT two-pass rather than the online
// form with accumulator rescaling:
//   TMA   : Q tile + the K blocks -> shared memory (128-byte swizzle), V blocks behind them
//   MMA   : S_kb = Q K_kb^T                  tcgen05.mma 128x128x16, accumulators in TMEM columns [128 kb, 128 kb+128)
//   pass 1: 8 warps read S from TMEM (thread = query row, two threads share a row), scale, mask, running max / sum
//   pass 2: read S again, P = exp2(s - m) / l as bf16 into the shared-memory tiles that held K (K is dead by then),
//           laid out as the K-major A operand of the second product; the same tiles are sent to global memory by TMA
//           when the backward pass needs P
//   MMA   : O += P_kb V_kb                   V straight from the packed qkv buffer as an MN-major B operand; O reuses
//           the TMEM columns of S_0
//   out   : O from TMEM -> bf16 -> global
// The CTA is persistent over work items; every mbarrier completes exactly once per item, so one parity bit tracks them.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "gemm.h"

namespace as {

constexpr int kAtParts = 4;                // softmax threads per query row (each owns 128 / kAtParts columns of a block)
constexpr int kAtCols = 128 / kAtParts;
constexpr int kAtSoft = 128 * kAtParts;    // softmax / epilogue threads
constexpr int kAtThreads = 128 + kAtSoft;  // warp 0: TMA, warp 1: MMA, warps 4..: softmax / epilogue
constexpr int kAtSub = 128 * 64 * 2;       // one [128 rows][64 bf16] swizzled sub-tile: 16 KB
constexpr int kAtKP = 0;                   // 4 key blocks x 2 sub-tiles (K by d-half, later P by key-half)
constexpr int kAtQ = 8 * kAtSub;           // 2 sub-tiles (d halves); V block 2 lands here once Q is dead
constexpr int kAtV = 10 * kAtSub;          // 2 stages x 2 sub-tiles (d chunks)
constexpr int kAtBars = 14 * kAtSub;       // 229376
constexpr int kAtRed = kAtBars + 256;      // exchange slots of the backward (float [kAtParts][128]); the forward uses a dead K tile
constexpr int kAtTotal = kAtRed + 2 * 128 * 8;
constexpr int kAtSmemMax = 232448;
constexpr float kLog2e = 1.4426950408889634f;

struct AttnDev {
  const long long* key_len;
  __nv_bfloat16* out;                 // forward: O [R,N,H*128]; backward: the dq column block of dqkv [R,N,3*H*128]
  const __nv_bfloat16* O;             // backward only: forward output, dO, stored probabilities
  const __nv_bfloat16* dO;
  const __nv_bfloat16* P;
  int R, N, H, NB, NQ, store_p;
  int out_ld;                         // row pitch of `out` in elements
  int b1_col, b2_col;                 // qkv column of the first / second B operand (forward: K, V; backward: V, K)
  float scale_log2, out_scale;
  long long tiles;
  long long* prof;                    // diagnostics: cycle sums of the softmax warps' phases (CTA 0)
};

static long long* g_attn_prof = nullptr;
int attn_set_profile(long long* buf) { g_attn_prof = buf; return AS_OK; }

// barrier indices
enum { kBarK = 0, kBarS = 4, kBarP = 8, kBarV = 12, kBarPv = 16, kBarO = 20, kBarDone = 21, kBarPf = 22, kBarCount = 26 };

// second-operand block at position o of the item's block order -> buffer: 0 and 3 share stage 0, 1 -> stage 1,
// 2 -> the dead Q tile
__device__ __forceinline__ int at_v_offset(int o) { return o == 2 ? kAtQ : kAtV + (o == 1 ? 2 * kAtSub : 0); }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// BWD = false: forward as described above (A = Q, first B = K, second B = V).
// BWD = true : the query-side half of the backward pass with the same pipeline: dP = dO V^T into TMEM (A = dO tile,
//              first B = V), dS = P o (dP - delta) with delta = rowsum(dO o O) and P read from global memory, dS -> bf16
//              -> shared-memory A tiles (and by TMA to global for the dK product), dQ = scale dS K (second B = K).
template <bool BWD>
__global__ void __launch_bounds__(kAtThreads, 1)
attn_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmQKV,
            const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmPin, const AttnDev p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t pad = (1024u - (raw & 1023u)) & 1023u;
  if (pad + kAtTotal > kAtSmemMax) __trap();
  uint8_t* smem = smem_raw + pad;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kAtBars);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kAtBars + kBarCount * 8);
  float2* red = reinterpret_cast<float2*>(smem + kAtRed);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NB = p.NB, HD = p.H * 128;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kBarCount; ++i) mbar_init(&bars[i], (i >= kBarP && i < kBarP + 4) ? (uint32_t)kAtSoft : (i == kBarDone ? (uint32_t)kAtSoft : 1u));
    mbar_fence_init();
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmA);
    if (p.store_p) tma_prefetch_desc(&tmP);
    if (BWD) tma_prefetch_desc(&tmPin);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (+ TMA stores of P / dS)
    // Buffers are handed back block by block: K/P tile kb of the next item is loaded as soon as P_kb V_kb of this item
    // has retired (kBarPv + kb), so the next item's first products run under this item's epilogue.
    if (lane == 0) {
      uint32_t it = 0;
      for (long long tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        const uint32_t ph = it & 1u;
        const int qt = (int)(tile % p.NQ);
        const int bh = (int)(tile / p.NQ);
        const int h = bh % p.H, b = bh / p.H;
        if (it > 0) {
          if (p.store_p) tma_store_wait_read();                    // the stores of the previous item have read their tiles
          for (int kb = 0; kb < NB; ++kb) mbar_wait(&bars[kBarPv + kb], ph ^ 1u);
        }
        const int rot = (int)(it % (uint32_t)NB);                  // block order of this item: rot, rot+1, .. (MMA warp)
        for (int o = 0; o < NB; ++o) {
          const int kb = rot + o < NB ? rot + o : rot + o - NB;
          mbar_expect_tx(&bars[kBarK + kb], (o == 0 ? 4u : 2u) * kAtSub);
          if (o == 0) {
            tma_load_4d(smem + kAtQ, &tmA, &bars[kBarK + kb], h * 128, qt * 128, b, 0);
            tma_load_4d(smem + kAtQ + kAtSub, &tmA, &bars[kBarK + kb], h * 128 + 64, qt * 128, b, 0);
          }
          tma_load_4d(smem + kAtKP + (2 * kb) * kAtSub, &tmQKV, &bars[kBarK + kb], p.b1_col + h * 128, kb * 128, b, 0);
          tma_load_4d(smem + kAtKP + (2 * kb + 1) * kAtSub, &tmQKV, &bars[kBarK + kb], p.b1_col + h * 128 + 64, kb * 128, b, 0);
        }
        // second-operand blocks by POSITION o in the order: 0 -> stage 0, 1 -> stage 1, 2 -> the Q tile (once every
        // first product has retired), 3 -> stage 0 again (once the position-0 second product has retired)
        auto load_v = [&](int o) {
          const int kb = rot + o < NB ? rot + o : rot + o - NB;
          uint8_t* dst = smem + at_v_offset(o);
          mbar_expect_tx(&bars[kBarV + o], 2u * kAtSub);
          tma_load_4d(dst, &tmQKV, &bars[kBarV + o], p.b2_col + h * 128, kb * 128, b, 0);
          tma_load_4d(dst + kAtSub, &tmQKV, &bars[kBarV + o], p.b2_col + h * 128 + 64, kb * 128, b, 0);
        };
        load_v(0);
        if (NB > 1) load_v(1);
        if constexpr (BWD) {
          // stored probabilities: block kb lands in the tile that held the first-product operand of block kb as soon as
          // that product has retired; the softmax warps turn it into dS in place
          for (int o = 0; o < NB; ++o) {
            const int kb = rot + o < NB ? rot + o : rot + o - NB;
            mbar_wait(&bars[kBarS + kb], ph);
            mbar_expect_tx(&bars[kBarPf + kb], 2u * kAtSub);
            tma_load_4d(smem + kAtKP + (2 * kb) * kAtSub, &tmPin, &bars[kBarPf + kb], kb * 128, qt * 128, bh, 0);
            tma_load_4d(smem + kAtKP + (2 * kb + 1) * kAtSub, &tmPin, &bars[kBarPf + kb], kb * 128 + 64, qt * 128, bh, 0);
          }
        }
        if (NB > 2) {
          const int last = rot + NB - 1 < NB ? rot + NB - 1 : rot - 1;
          mbar_wait(&bars[kBarS + last], ph);
          load_v(2);
        }
        for (int o = 0; o < NB; ++o) {
          const int kb = rot + o < NB ? rot + o : rot + o - NB;
          if (p.store_p) {
            mbar_wait(&bars[kBarP + kb], ph);                      // tile kb written (and proxy-fenced) by all softmax threads
            tma_store_4d(&tmP, smem + kAtKP + (2 * kb) * kAtSub, kb * 128, qt * 128, bh, 0);
            tma_store_4d(&tmP, smem + kAtKP + (2 * kb + 1) * kAtSub, kb * 128 + 64, qt * 128, bh, 0);
            tma_store_commit();
          }
          if (o == 0 && NB > 3) {
            mbar_wait(&bars[kBarPv + kb], ph);
            load_v(3);
          }
        }
      }
      if (p.store_p) tma_store_wait_all();
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (warp-uniform, one lane issues)
    // Item `it` walks its key blocks in the order rot, rot+1, .. (rot = it mod NB) and accumulates the second product
    // in the TMEM columns of block rot, the first one the softmax warps consume.  The next item starts at rot+1: the
    // columns of all its blocks but the last are free once this item's softmax has consumed them (this warp has waited
    // for that before issuing this item's second products), and its last block -- this item's accumulator -- waits for
    // the epilogue.  So the next item's first products run under this item's epilogue.
    constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t idesc_qk = (1u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint32_t idesc_pv = idesc_qk | (1u << 16);           // second B operand is MN-major
    const uint32_t sbase = smem_u32(smem);
    uint32_t it = 0;
    for (long long tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1u;
      const int rot = (int)(it % (uint32_t)NB);
      for (int o = 0; o < NB; ++o) {
        const int kb = rot + o < NB ? rot + o : rot + o - NB;
        if (o == NB - 1 && it > 0) mbar_wait(&bars[kBarDone], ph ^ 1u);  // accumulator drained by the previous epilogue
        mbar_wait(&bars[kBarK + kb], ph);
        tc_fence_after();
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const uint32_t a_lo = (1u << 16) | ((sbase + kAtQ + j * kAtSub) >> 4);
          const uint32_t b_lo = (1u << 16) | ((sbase + kAtKP + (2 * kb + j) * kAtSub) >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t da = ((uint64_t)kDescHi << 32) | (a_lo + 2u * k);
            const uint64_t db = ((uint64_t)kDescHi << 32) | (b_lo + 2u * k);
            umma_bf16_lead(0, tmem_base + (uint32_t)(kb * 128), da, db, idesc_qk, (j | k) ? 1u : 0u);
          }
        }
        umma_commit_lead(0, &bars[kBarS + kb]);
      }
      for (int o = 0; o < NB; ++o) {
        const int kb = rot + o < NB ? rot + o : rot + o - NB;
        mbar_wait(&bars[kBarP + kb], ph);
        mbar_wait(&bars[kBarV + o], ph);
        tc_fence_after();
        const uint32_t vaddr = sbase + (uint32_t)at_v_offset(o);
        const uint32_t b_lo = ((uint32_t)(kAtSub >> 4) << 16) | (vaddr >> 4);
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          const uint32_t a_lo = (1u << 16) | ((sbase + kAtKP + (2 * kb + (s >> 2)) * kAtSub) >> 4);
          const uint64_t da = ((uint64_t)kDescHi << 32) | (a_lo + 2u * (s & 3));
          const uint64_t db = ((uint64_t)kDescHi << 32) | (b_lo + (uint32_t)(s * (16 * 128 / 16)));
          umma_bf16_lead(0, tmem_base + (uint32_t)(rot * 128), da, db, idesc_pv, (o | s) ? 1u : 0u);
        }
        umma_commit_lead(0, &bars[kBarPv + kb]);
      }
      umma_commit_lead(0, &bars[kBarO]);
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ softmax + output (4 * kAtParts warps)
    const int e = warp - 4;
    const int quarter = warp & 3;                                  // TMEM lane quarter this warp may read
    const int half = e >> 2;                                       // which kAtCols of a block's 128 columns
    const int row = quarter * 32 + lane;
    const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * kAtCols);
    uint32_t it = 0;
    for (long long tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1u;
      const int qt = (int)(tile % p.NQ);
      const int bh = (int)(tile / p.NQ);
      const int h = bh % p.H, b = bh / p.H;
      long long kl = (!BWD && p.key_len) ? p.key_len[b] : (long long)p.N;
      if (kl > p.N) kl = p.N;
      uint32_t v[kAtCols];
      float out_scale = 1.f;
      const int rot = (int)(it % (uint32_t)NB);                    // block order rot, rot+1, ..; accumulator in block rot
      const bool prof = p.prof && blockIdx.x == 0 && threadIdx.x == 128;
      long long tp0 = 0, tp1 = 0, tp2 = 0, tp3 = 0;
      if (prof) tp0 = clock64();
      if constexpr (!BWD) {
      float m = -INFINITY, l = 0.f;
      // pass 1: running max / sum over this thread's 64 columns of every key block
      for (int o = 0; o < NB; ++o) {
        const int kb = rot + o < NB ? rot + o : rot + o - NB;
        mbar_wait(&bars[kBarS + kb], ph);
        tc_fence_after();
        tmem_ld32(t_row + kb * 128, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
        if constexpr (kAtCols > 32) tmem_ld32(t_row + kb * 128 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[kAtCols - 32]));
        tmem_ld_wait();
        const int lim = (int)kl - (kb * 128 + half * kAtCols);
        float cm = -INFINITY;
#pragma unroll
        for (int i = 0; i < kAtCols; ++i) {
          const float t = i < lim ? __uint_as_float(v[i]) * p.scale_log2 : -1e9f * kLog2e;
          v[i] = __float_as_uint(t);
          cm = fmaxf(cm, t);
        }
        if (cm > m) { l *= fast_exp2(m - cm); m = cm; }
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int i = 0; i < kAtCols; i += 2) {
          s0 += fast_exp2(__uint_as_float(v[i]) - m);
          s1 += fast_exp2(__uint_as_float(v[i + 1]) - m);
        }
        l += s0 + s1;
      }
      if (prof) tp1 = clock64();
      // the row's kAtParts threads combine (max, sum) through the first K tile: every thread has waited for all first
      // products, so K is dead; the second barrier keeps pass 2 from overwriting the slots while they are still read
      {
        float2* xch = reinterpret_cast<float2*>(smem + kAtKP);
        xch[half * 128 + row] = make_float2(m, l);
        asm volatile("bar.sync 1, %0;" ::"n"(kAtSoft) : "memory");
        float mm = m;
#pragma unroll
        for (int q = 0; q < kAtParts; ++q) mm = fmaxf(mm, xch[q * 128 + row].x);
        float ll = 0.f;
#pragma unroll
        for (int q = 0; q < kAtParts; ++q) { const float2 o = xch[q * 128 + row]; ll += o.y * fast_exp2(o.x - mm); }
        m = mm; l = ll;
        asm volatile("bar.sync 1, %0;" ::"n"(kAtSoft) : "memory");
      }
      const float rinv = 1.f / l;
      // pass 2: normalised probabilities -> bf16 -> the A-operand tiles (128-byte swizzle: chunk c of row r at c ^ (r & 7))
      for (int o = 0; o < NB; ++o) {
        const int kb = rot + o < NB ? rot + o : rot + o - NB;
        tmem_ld32(t_row + kb * 128, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
        if constexpr (kAtCols > 32) tmem_ld32(t_row + kb * 128 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[kAtCols - 32]));
        tmem_ld_wait();
        const int lim = (int)kl - (kb * 128 + half * kAtCols);
        uint8_t* dst = smem + kAtKP + (2 * kb + (half * kAtCols) / 64) * kAtSub + row * 128;
        const int c0 = ((half * kAtCols) % 64) / 8;              // first 16-byte chunk of this thread in the 128-byte row
#pragma unroll
        for (int c = 0; c < kAtCols / 8; ++c) {
          uint32_t w[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int i = c * 8 + 2 * j;
            const float t0 = i < lim ? __uint_as_float(v[i]) * p.scale_log2 : -1e9f * kLog2e;
            const float t1 = i + 1 < lim ? __uint_as_float(v[i + 1]) * p.scale_log2 : -1e9f * kLog2e;
            w[j] = pack_bf16(fast_exp2(t0 - m) * rinv, fast_exp2(t1 - m) * rinv);
          }
          *reinterpret_cast<uint4*>(dst + (((c0 + c) ^ (row & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
        }
        tc_fence_before();
        fence_proxy_async_smem();
        mbar_arrive(&bars[kBarP + kb]);
      }
      } else {
        // delta = rowsum(dO o O): each of the row's kAtParts threads sums its kAtCols columns
        const long long orow = ((long long)b * p.N + qt * 128 + row) * HD + h * 128 + half * kAtCols;
        const uint4* o4 = reinterpret_cast<const uint4*>(p.O + orow);
        const uint4* g4 = reinterpret_cast<const uint4*>(p.dO + orow);
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < kAtCols / 8; ++c) {
          const uint4 ov = __ldg(o4 + c), gv = __ldg(g4 + c);
          const uint32_t ow[4] = {ov.x, ov.y, ov.z, ov.w}, gw[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            d = fmaf(__uint_as_float(ow[j] << 16), __uint_as_float(gw[j] << 16), d);
            d = fmaf(__uint_as_float(ow[j] & 0xffff0000u), __uint_as_float(gw[j] & 0xffff0000u), d);
          }
        }
        float* xch = reinterpret_cast<float*>(red);                // [kAtParts][128] floats: 2 KB
        if (it > 0) asm volatile("bar.sync 1, %0;" ::"n"(kAtSoft) : "memory");   // previous item's reads are done
        xch[half * 128 + row] = d;
        asm volatile("bar.sync 1, %0;" ::"n"(kAtSoft) : "memory");
        float delta = 0.f;
#pragma unroll
        for (int q = 0; q < kAtParts; ++q) delta += xch[q * 128 + row];
        out_scale = p.out_scale;
        for (int o = 0; o < NB; ++o) {
          const int kb = rot + o < NB ? rot + o : rot + o - NB;
          mbar_wait(&bars[kBarS + kb], ph);
          tc_fence_after();
          tmem_ld32(t_row + kb * 128, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
          if constexpr (kAtCols > 32) tmem_ld32(t_row + kb * 128 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[kAtCols - 32]));
          tmem_ld_wait();
          uint8_t* dst = smem + kAtKP + (2 * kb + (half * kAtCols) / 64) * kAtSub + row * 128;
          const int c0 = ((half * kAtCols) % 64) / 8;               // first 16-byte chunk of this thread in the 128-byte row
          mbar_wait(&bars[kBarPf + kb], ph);                       // P block in place (TMA, same swizzle as the dS tile)
#pragma unroll
          for (int c = 0; c < kAtCols / 8; ++c) {
            const uint4 pvc = *reinterpret_cast<const uint4*>(dst + (((c0 + c) ^ (row & 7)) << 4));
            const uint32_t pw[4] = {pvc.x, pvc.y, pvc.z, pvc.w};
            uint32_t w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int i = c * 8 + 2 * j;
              const float p0 = __uint_as_float(pw[j] << 16), p1 = __uint_as_float(pw[j] & 0xffff0000u);
              w[j] = pack_bf16(p0 * (__uint_as_float(v[i]) - delta), p1 * (__uint_as_float(v[i + 1]) - delta));
            }
            *reinterpret_cast<uint4*>(dst + (((c0 + c) ^ (row & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
          }
          tc_fence_before();
          fence_proxy_async_smem();
          mbar_arrive(&bars[kBarP + kb]);
        }
      }
      // output: accumulator (fp32, TMEM columns [0,128)) -> bf16 rows of O / dq
      if (prof) tp2 = clock64();
      mbar_wait(&bars[kBarO], ph);
      if (prof) tp3 = clock64();
      tc_fence_after();
      tmem_ld32(t_row + rot * 128, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      if constexpr (kAtCols > 32) tmem_ld32(t_row + rot * 128 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[kAtCols - 32]));
      tmem_ld_wait();
      {
        uint4* out = reinterpret_cast<uint4*>(p.out + ((long long)b * p.N + qt * 128 + row) * p.out_ld + h * 128 + half * kAtCols);
#pragma unroll
        for (int c = 0; c < kAtCols / 8; ++c)
          out[c] = make_uint4(pack_bf16(__uint_as_float(v[8 * c]) * out_scale, __uint_as_float(v[8 * c + 1]) * out_scale),
                              pack_bf16(__uint_as_float(v[8 * c + 2]) * out_scale, __uint_as_float(v[8 * c + 3]) * out_scale),
                              pack_bf16(__uint_as_float(v[8 * c + 4]) * out_scale, __uint_as_float(v[8 * c + 5]) * out_scale),
                              pack_bf16(__uint_as_float(v[8 * c + 6]) * out_scale, __uint_as_float(v[8 * c + 7]) * out_scale));
      }
      tc_fence_before();
      mbar_arrive(&bars[kBarDone]);
      if (prof) {
        const long long tp4 = clock64();
        p.prof[0] += 1; p.prof[1] += tp1 - tp0; p.prof[2] += tp2 - tp1; p.prof[3] += tp3 - tp2; p.prof[4] += tp4 - tp3;
        if (it > 0) p.prof[5] += tp0 - p.prof[7];
        p.prof[7] = tp4;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

int make_map_bf16(CUtensorMap* tm, const void* ptr, long long inner, long long rows, long long ld, int nb0, long long sb0,
                  int nb1, long long sb1, int box_inner, int box_rows);
int gemm_num_sms();

int attn_fwd_ok(int N, int D) { return D == 128 && N >= 128 && N <= 512 && N % 128 == 0; }

template <bool BWD>
static int attn_launch(const CUtensorMap& ta, const CUtensorMap& tq, const CUtensorMap& tp, const CUtensorMap& tpin,
                       const AttnDev& p, cudaStream_t st) {
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(attn_kernel<BWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAtSmemMax);
    if (e != cudaSuccess) return (int)e;
    attr_done = true;
  }
  const int sms = gemm_num_sms();
  const int grid = (int)(p.tiles < sms ? p.tiles : sms);
  attn_kernel<BWD><<<grid, kAtThreads, kAtSmemMax, st>>>(ta, tq, tp, tpin, p);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int attn_fwd(const void* qkv, const long long* key_len, void* O, void* P, long long R, int N, int H, int D, float scale,
             cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!attn_fwd_ok(N, D) || H <= 0) return AS_EUNSUPPORTED;
  if (!qkv || !O) return AS_EINVAL;
  CUtensorMap tq, tp;
  int rc = make_map_bf16(&tq, qkv, 3LL * H * D, N, 3LL * H * D, (int)R, (long long)N * 3 * H * D, 1, 0, 64, 128);
  if (rc != AS_OK) return rc;
  if (P) {
    rc = make_map_bf16(&tp, P, N, N, N, (int)(R * H), (long long)N * N, 1, 0, 64, 128);
    if (rc != AS_OK) return rc;
  } else {
    tp = tq;
  }
  AttnDev p = {};
  p.key_len = key_len;
  p.out = reinterpret_cast<__nv_bfloat16*>(O);
  p.R = (int)R; p.N = N; p.H = H; p.NB = N / 128; p.NQ = N / 128; p.store_p = P ? 1 : 0;
  p.out_ld = H * D; p.b1_col = H * D; p.b2_col = 2 * H * D;
  p.scale_log2 = scale * kLog2e; p.out_scale = 1.f;
  p.tiles = R * H * p.NQ;
  p.prof = g_attn_prof;
  return attn_launch<false>(tq, tq, tp, tp, p, st);
}

// dS [R,H,N,N] and dq (column block 0 of dqkv [R,N,3*H*D]) from P, O, dO, qkv; dK and dV remain GEMMs over dS / P.
int attn_bwd_dq(const void* qkv, const void* O, const void* dO, const void* P, void* dS, void* dqkv, long long R, int N, int H,
                int D, float scale, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!attn_fwd_ok(N, D) || H <= 0) return AS_EUNSUPPORTED;
  if (!qkv || !O || !dO || !P || !dS || !dqkv) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(O) | reinterpret_cast<uintptr_t>(P)) & 15) return AS_EALIGN;
  CUtensorMap tq, ta, tp, tpin;
  int rc = make_map_bf16(&tpin, P, N, N, N, (int)(R * H), (long long)N * N, 1, 0, 64, 128);
  if (rc != AS_OK) return rc;
  rc = make_map_bf16(&tq, qkv, 3LL * H * D, N, 3LL * H * D, (int)R, (long long)N * 3 * H * D, 1, 0, 64, 128);
  if (rc != AS_OK) return rc;
  rc = make_map_bf16(&ta, dO, (long long)H * D, N, (long long)H * D, (int)R, (long long)N * H * D, 1, 0, 64, 128);
  if (rc != AS_OK) return rc;
  rc = make_map_bf16(&tp, dS, N, N, N, (int)(R * H), (long long)N * N, 1, 0, 64, 128);
  if (rc != AS_OK) return rc;
  AttnDev p = {};
  p.out = reinterpret_cast<__nv_bfloat16*>(dqkv);
  p.O = reinterpret_cast<const __nv_bfloat16*>(O);
  p.dO = reinterpret_cast<const __nv_bfloat16*>(dO);
  p.P = reinterpret_cast<const __nv_bfloat16*>(P);
  p.R = (int)R; p.N = N; p.H = H; p.NB = N / 128; p.NQ = N / 128; p.store_p = 1;
  p.out_ld = 3 * H * D; p.b1_col = 2 * H * D; p.b2_col = H * D;
  p.scale_log2 = 0.f; p.out_scale = scale;
  p.tiles = R * H * p.NQ;
  p.prof = g_attn_prof;
  return attn_launch<true>(ta, tq, tp, tpin, p, st);
}

}  // namespace as
