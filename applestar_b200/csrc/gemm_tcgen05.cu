// Persistent warp-specialised bf16 GEMM for sm_100a: TMA (4-D tensor maps, 128B swizzle) -> shared memory ring
// -> tcgen05.mma (cta_group::1, M=128, N=BN) with fp32 accumulators double-buffered in TMEM -> tcgen05.ld
// epilogue (bias / ReLU / border mask / dtype cast) written straight to global memory.
//
//   C[b][m][n] = epi( sum_k A[b][m][k] * B[b][n][k] )
//
// Either operand may be "K-major" (contraction index contiguous in global memory) or "MN-major" (the m / n index
// contiguous), so X*W^T (forward), dY*W (dgrad) and dY^T*X (wgrad) all run without materialising a transpose.
// Extra modes: split-K (fp32 partials into a caller workspace, reduced by as_splitk_reduce) and an implicit-GEMM
// 3x3 convolution over zero-padded NHWC pixels (per-tap row shift of the A operand, see as_conv3x3_*).
// Multi-tap convolutions normally run in SLAB mode: one CTA step covers 256 consecutive pixels (two accumulators);
// per 64-channel block the pixels plus a halo of +-(row pitch + 1) are loaded ONCE, and each tap's A operand is a
// shared-memory descriptor whose start address is offset by whole 128-byte rows into that slab.  That cuts the L2 -> shared-memory traffic of A by ~9x and of the weights by 2x; the per-tap
// mode re-reads A for every tap and is L2-bandwidth bound (measured 190-730 TFLOP/s).
//
// Warp roles (12 warps): w0 TMA producer, w1 MMA issuer, w2 TMEM allocator, w4-11 epilogue.
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "gemm.h"

// Register budget by role (setmaxnreg): warps 0-3 (TMA producer, MMA issuer, TMEM allocator) give registers up, the two
// epilogue warpgroups take them.  0 = one budget for all warps (168 at 384 threads).
#ifndef AS_WG0_REGS
#define AS_WG0_REGS 104
#endif
#ifndef AS_EPI_REGS
#define AS_EPI_REGS 200
#endif

namespace as {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle row
constexpr int kEpiWarps = 8;
constexpr int kGemmThreads = 128 + 32 * kEpiWarps;
constexpr int kStgPitch = 20;   // floats per staged row (16 columns + pad): 16-byte aligned, conflict-free float4 access

struct GemmDev {
  int M, N, K;
  int nb0, nb1;
  int m_blocks, n_blocks;
  int splits, kb_per_split, num_kb;
  // conv mode
  int taps, cin_blocks, cin;
  int shift[9];
  int img_w, img_h;  // padded image dims for border masking (0 = off)
  int b_k_shift;     // MN-major B only: row offset added to the contraction coordinate
  int wgrad_taps;    // >0: batch index b0 = conv tap; A and B are shared by all taps, B rows shifted by shift[b0]
  // wg_pair: batch index b0 = PAIR of taps.  The M side has only 64 channels, so the two 64-column chunks of the A tile
  // are loaded at different contraction rows (pair_a[b0][j]) and one 128-row MMA produces both taps: accumulator rows
  // 0-63 = first tap, 64-127 = second tap.  B rows are shifted by pair_b[b0].
  int wg_pair, pair_a[5][2], pair_b[5];
  // slab mode (multi-tap conv): 256-row steps, A slab = 2 TMA boxes of slab_box rows starting at row m0 - halo
  int slab, halo, slab_box, slab_bytes, slab_slots, b_slots, tpb, b_resident;
  int half_k;        // slab mode: only K steps 0-1 of a block hold real input channels (Cin <= 32)
  int mc;            // slab ring mode on CTA pairs: each CTA loads half of every weight tile and multicasts it to both
  // epilogue
  void* C;
  void* C2;   // optional bf16 copy of the output (same ldc / batch strides)
  int c_bf16;
  long long ldc, c_sb0, c_sb1;
  const float* bias;
  const __nv_bfloat16* resid;   // bf16 [.., M, ldc]: added to the accumulator row before ReLU / border masking (or nullptr)
  uint32_t* bits_out;           // ReLU sign bits of the bf16 output, [M][N / 32] words (or nullptr)
  const uint32_t* bits_in;      // output zeroed where the bit is clear (or nullptr)
  const __nv_bfloat16* mask;    // bf16 [.., M, ldc]: output zeroed where mask <= 0 (ReLU backward of the producing layer)
  int relu;
  float alpha;
  float* ws;  // split-K partials
  int tma_store;   // 1: C (and C2) are written with TMA bulk stores from a swizzled staging tile
  int plain;  // host: 1 = no addend / mask / sign bits / border / diagnostics and the TMA-store epilogue (EPI 1), 2 = sign bits only (EPI 2)
  int debug;  // diagnostics: 1 = epilogue skips global stores, 2 = epilogue only drains TMEM
};

// diagnostics: cycles the slab-mode MMA issuer spends in each wait (as_debug_gemm_profile); nullptr normally
__device__ long long* g_gemm_prof = nullptr;

template <int BN>
struct SmemLayout {
  static constexpr int kABytes = BM * 128;
  static constexpr int kBBytes = BN * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128 ? 6 : 8);
  static constexpr int kRingBytes = 192 * 1024;                              // operand ring (>= kStages * kStageBytes); slab mode re-carves it
  static_assert(kStages * kStageBytes <= kRingBytes, "ring too small");
  static constexpr int kBarBytes = 1024;
  static constexpr int kStgBytes = kEpiWarps * 4096;                         // per-warp epilogue staging tiles (32 rows x 128 B)
  static constexpr int kTotal = kRingBytes + kBarBytes + kStgBytes + 1024;   // +1024 alignment slack
};

// EPI 1 (plain): the epilogue of the TMA-store path is compiled without the side inputs (addend, ReLU masks / sign bits,
// diagnostics); EPI 2: the same, but with the sign-bit input / output.  The K = 256 GEMMs of the
// Transformer are bound by their epilogue's instruction stream, and every runtime test of an option that is off costs
// them issue slots (measured: -10 % on the qkv / FFN forward GEMMs).  EPI 0: everything, decided at run time.
template <int BN, bool A_MN, bool B_MN, int EPI = 0>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmC2,
            const __grid_constant__ CUtensorMap tmB2, const GemmDev p) {
  using L = SmemLayout<BN>;
  constexpr bool PLAIN = EPI != 0;          // no addend / bf16 mask / diagnostics
  constexpr bool NOBITS = EPI == 1;         // ... and no sign bits either
  constexpr int kStages = L::kStages;
  // accumulator ring in TMEM: narrow tiles have few k-blocks per tile, so the MMA issuer may run up to 8 tiles ahead
  // of the epilogue; for BN < 128 the two epilogue warp sets also alternate tiles instead of splitting columns
  constexpr int kAcc = (BN >= 256) ? 2 : (BN >= 128 ? 4 : 8);
  constexpr bool kSplitCols = BN >= 128;
  constexpr uint32_t kTmemCols = kAcc * BN;
  static_assert(kTmemCols == 512 || kTmemCols == 256, "TMEM allocation must be a power of two <= 512");

  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle; pointer arithmetic (not an integer round trip) keeps the shared
  // address space visible to the compiler
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* bar_base = smem + L::kRingBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bar_base);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + kAcc;
  uint64_t* slab_full = tmem_empty + kAcc;            // slab mode: 3 A slabs
  uint64_t* slab_empty = slab_full + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(slab_empty + 3);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], p.mc ? 2 : 1);       // CTA pair: a weight slot is free when BOTH issuers are done with it
    }
    for (int i = 0; i < kAcc; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], kSplitCols ? kEpiWarps : kEpiWarps / 2);
    }
    for (int i = 0; i < 3; ++i) {
      mbar_init(&slab_full[i], 1);
      mbar_init(&slab_empty[i], 1);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  if (p.mc) cluster_sync_all();                     // the peer's barriers are initialised before anything targets them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int batch = p.nb0 * p.nb1;
  const long long tiles_per_split = (long long)batch * p.m_blocks * p.n_blocks;
  const long long num_tiles = tiles_per_split * p.splits;

  // slab mode: work item = (n block, pair of m blocks); shared memory = the A slabs followed by the weight tiles
  // (a ring of per-tap tiles, or -- when they all fit -- every tile of the convolution, loaded once)
  long long num_work = p.slab ? (long long)((p.m_blocks + 1) / 2) * p.n_blocks : num_tiles;
  if (p.mc) num_work = (num_work + 1) & ~1LL;       // both CTAs of a pair run the same number of steps (a phantom one is all zeros)
  const uint32_t crank = p.mc ? cluster_ctarank() : 0u;
  uint8_t* b_ring = smem + p.slab_slots * p.slab_bytes;

  if (warp < 4) {
#if AS_WG0_REGS
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(AS_WG0_REGS));
#endif
  if (warp == 0 && p.slab) {
    // ------------------------------------------------------------------ TMA producer, slab mode
    if (lane == 0 && !A_MN && !B_MN) {
      // chunk = (work item, 64-channel block).  The slab of chunk i+1 is requested before the weight tiles of chunk i.
      const long long my_work = (num_work - blockIdx.x + gridDim.x - 1) / gridDim.x;
      const long long chunks = my_work * p.cin_blocks;
      const int nsl = p.slab_slots;
      int bslot = 0; uint32_t bphase = 0;
      auto load_slab = [&](long long ci) {
        const long long w = blockIdx.x + (ci / p.cin_blocks) * gridDim.x;
        const int cb = (int)(ci % p.cin_blocks);
        const int m0 = (int)(w / p.n_blocks) * (2 * BM);
        const int sl = (int)(ci % nsl);
        const uint32_t ph = (uint32_t)((ci / nsl) & 1);
        mbar_wait(&slab_empty[sl], ph ^ 1);
        uint8_t* dst = smem + sl * p.slab_bytes;
        mbar_expect_tx(&slab_full[sl], 2 * p.slab_box * 128);
        tma_load_4d(dst, &tmA, &slab_full[sl], cb * BK, m0 - p.halo, 0, 0);
        tma_load_4d(dst + p.slab_box * 128, &tmA, &slab_full[sl], cb * BK, m0 - p.halo + p.slab_box, 0, 0);
      };
      if (p.b_resident) {
        // all weight tiles of the convolution (n_blocks == 1): one barrier, loaded once per CTA
        mbar_expect_tx(&full_bar[0], p.cin_blocks * p.taps * L::kBBytes);
        for (int cb = 0; cb < p.cin_blocks; ++cb)
          for (int tap = 0; tap < p.taps; ++tap)
            tma_load_4d(b_ring + (cb * p.taps + tap) * L::kBBytes, &tmB, &full_bar[0], tap * p.cin + cb * BK, 0, 0, 0);
        for (long long ci = 0; ci < chunks; ++ci) load_slab(ci);
      } else {
        if (chunks > 0) load_slab(0);
        for (long long ci = 0; ci < chunks; ++ci) {
          if (ci + 1 < chunks) load_slab(ci + 1);
          const long long w = blockIdx.x + (ci / p.cin_blocks) * gridDim.x;
          const int cb = (int)(ci % p.cin_blocks);
          const int n0 = (int)(w % p.n_blocks) * BN;
          for (int tap = 0; tap < p.taps; tap += p.tpb) {
            mbar_wait(&empty_bar[bslot], bphase ^ 1);
            mbar_expect_tx(&full_bar[bslot], p.tpb * L::kBBytes);
            for (int j = 0; j < p.tpb; ++j) {
              uint8_t* dstb = b_ring + (bslot * p.tpb + j) * L::kBBytes;
              const int kc = (tap + j) * p.cin + cb * BK;
              if (p.mc)   // this CTA's half of the tile (rows crank*BN/2 ..) goes to both CTAs of the pair
                tma_load_4d_mc(dstb + crank * (L::kBBytes / 2), &tmB2, &full_bar[bslot], kc, n0 + (int)crank * (BN / 2), 0, 0, (uint16_t)3);
              else
                tma_load_4d(dstb, &tmB, &full_bar[bslot], kc, n0, 0, 0);
            }
            if (++bslot == p.b_slots) { bslot = 0; bphase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1 && p.slab) {
    // ------------------------------------------------------------------ MMA issuer, slab mode (warp-uniform loop,
    // lane 0 issues: see umma_bf16_lead)
    if (!A_MN && !B_MN) {
      const uint32_t leader = lane == 0 ? 1u : 0u;
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      // descriptor = constant high word | (lbo << 16) | (address >> 4): per MMA only the low word changes, by an
      // immediate (the issuing thread is the critical path when N is small)
      constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);
      constexpr uint32_t kDescLbo = (16u >> 4) << 16;
      const int nsl = p.slab_slots;
      const bool half_k = p.half_k != 0;
      int bslot = 0; uint32_t bphase = 0;
      long long ci = 0;
      uint32_t it2 = 0;
      if (p.b_resident) { mbar_wait(&full_bar[0], 0); tc_fence_after(); }
      long long* prof = g_gemm_prof;
      long long t_acc = 0, t_slab = 0, t_b = 0, t_all = prof ? clock64() : 0, tk;
      for (long long w = blockIdx.x; w < num_work; w += gridDim.x, ++it2) {
        const int acc0 = (int)((2 * it2) % kAcc);
        const uint32_t acc_phase = ((2 * it2) / kAcc) & 1u;
        if (prof) tk = clock64();
        mbar_wait(&tmem_empty[acc0], acc_phase ^ 1);
        mbar_wait(&tmem_empty[acc0 + 1], acc_phase ^ 1);
        if (prof) t_acc += clock64() - tk;
        tc_fence_after();
        const uint32_t d0 = __shfl_sync(0xffffffffu, tmem_base + (uint32_t)(acc0 * BN), 0);
        for (int cb = 0; cb < p.cin_blocks; ++cb, ++ci) {
          const int sl = (int)(ci % nsl);
          if (prof) tk = clock64();
          mbar_wait(&slab_full[sl], (uint32_t)((ci / nsl) & 1));
          if (prof) t_slab += clock64() - tk;
          tc_fence_after();
          const uint32_t slab_lo = kDescLbo | ((smem_u32(smem + sl * p.slab_bytes) >> 4) + (uint32_t)(p.halo * 8));
          for (int tap = 0; tap < p.taps; tap += p.tpb) {
            uint32_t b_lo;
            if (p.b_resident) {
              b_lo = kDescLbo | (smem_u32(b_ring + (cb * p.taps + tap) * L::kBBytes) >> 4);
            } else {
              if (prof) tk = clock64();
              mbar_wait(&full_bar[bslot], bphase);
              if (prof) t_b += clock64() - tk;
              tc_fence_after();
              b_lo = kDescLbo | (smem_u32(b_ring + bslot * p.tpb * L::kBBytes) >> 4);
            }
            for (int j = 0; j < p.tpb; ++j) {
              // the 128-byte swizzle is a function of the absolute shared-memory address, so a start address that
              // is only 128-byte aligned needs no descriptor base offset (verified against the per-tap mode)
              // shuffles from lane 0 mark the operands as warp-uniform for the compiler (uniform-register operands)
              const uint32_t a_lo = __shfl_sync(0xffffffffu, slab_lo + (uint32_t)(p.shift[tap + j] * 8), 0);
              const uint32_t bj_lo = __shfl_sync(0xffffffffu, b_lo + (uint32_t)(j * (L::kBBytes >> 4)), 0);
              const uint32_t first = (cb > 0 || tap + j > 0) ? 1u : 0u;
              // consecutive MMAs alternate between the two accumulators (no back-to-back dependent accumulation)
              auto mma_k = [&](int k) {
#pragma unroll
                for (int sub = 0; sub < 2; ++sub) {
                  const uint64_t da = ((uint64_t)kDescHi << 32) | (a_lo + (uint32_t)(sub * (BM * 128 / 16) + k * 2));
                  const uint64_t db = ((uint64_t)kDescHi << 32) | (bj_lo + (uint32_t)(k * 2));
                  umma_bf16_lead(leader, d0 + (uint32_t)(sub * BN), da, db, idesc, k > 0 ? 1u : first);
                }
              };
              // a layer with <= 32 input channels fills only half of the 64-wide K block (the rest is TMA zero fill):
              // two fully unrolled variants, chosen by a warp-uniform flag, instead of a data-dependent trip count
              if (half_k) {
                mma_k(0); mma_k(1);
              } else {
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) mma_k(k);
              }
            }
            if (!p.b_resident) {
              if (p.mc) umma_commit_mc_lead(leader, &empty_bar[bslot], (uint16_t)3);
              else umma_commit_lead(leader, &empty_bar[bslot]);
              if (++bslot == p.b_slots) { bslot = 0; bphase ^= 1; }
            }
          }
          umma_commit_lead(leader, &slab_empty[sl]);
        }
        umma_commit_lead(leader, &tmem_full[acc0]);
        umma_commit_lead(leader, &tmem_full[acc0 + 1]);
      }
      if (prof && lane == 0) {
        prof[blockIdx.x * 4 + 0] = clock64() - t_all; prof[blockIdx.x * 4 + 1] = t_acc;
        prof[blockIdx.x * 4 + 2] = t_slab; prof[blockIdx.x * 4 + 3] = t_b;
      }
    }
  } else if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        long long t = tile;
        const int n_blk = (int)(t % p.n_blocks); t /= p.n_blocks;
        const int m_blk = (int)(t % p.m_blocks); t /= p.m_blocks;
        const int b = (int)(t % batch);
        const int split = (int)(t / batch);
        const int b0 = b % p.nb0, b1 = b / p.nb0;
        const int m0 = m_blk * BM, n0 = n_blk * BN;
        const int kb_begin = split * p.kb_per_split;
        const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * L::kStageBytes;
          uint8_t* sb = sa + L::kABytes;
          mbar_expect_tx(&full_bar[stage], L::kStageBytes);
          int ka, kbcol, mshift = 0;
          if (p.taps > 0) {
            const int tap = kb / p.cin_blocks, cb = kb - tap * p.cin_blocks;
            ka = cb * BK;
            kbcol = tap * p.cin + cb * BK;
            mshift = p.shift[tap];
          } else {
            ka = kbcol = kb * BK;
          }
          const int ob0 = p.wgrad_taps > 0 ? 0 : b0;                       // operands shared by all taps
          const int kshift = p.wg_pair ? p.pair_b[b0] : (p.wgrad_taps > 0 ? p.shift[b0] : p.b_k_shift);
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_4d(sa + j * (BK * 128), &tmA, &full_bar[stage], p.wg_pair ? 0 : m0 + j * 64,
                          p.wg_pair ? ka + p.pair_a[b0][j] : ka, ob0, b1);
          } else {
            tma_load_4d(sa, &tmA, &full_bar[stage], ka, m0 + mshift, ob0, b1);
          }
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_4d(sb + j * (BK * 128), &tmB, &full_bar[stage], n0 + j * 64, kbcol + kshift, ob0, b1);
          } else {
            tma_load_4d(sb, &tmB, &full_bar[stage], kbcol, n0, ob0, b1);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (warp-uniform loop, lane 0 issues)
    {
      const uint32_t leader = lane == 0 ? 1u : 0u;
      // instruction descriptor: fp32 accum, bf16 A/B, majors, N>>3, M>>4
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                             ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;
      for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int acc = (int)(it % kAcc);
        const uint32_t acc_phase = (it / kAcc) & 1u;
        const int split = (int)(tile / tiles_per_split);
        const int kb_begin = split * p.kb_per_split;
        const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = __shfl_sync(0xffffffffu, tmem_base + (uint32_t)(acc * BN), 0);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * L::kStageBytes);
          const uint32_t sb = sa + L::kABytes;
          // descriptor low word = (lbo << 16) | (address >> 4); the high word is constant, so per MMA the low words
          // only advance by an immediate
          constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);
          const uint32_t a_lo = __shfl_sync(0xffffffffu, ((A_MN ? (uint32_t)(BK * 128) : 16u) >> 4 << 16) | (sa >> 4), 0);
          const uint32_t b_lo = __shfl_sync(0xffffffffu, ((B_MN ? (uint32_t)(BK * 128) : 16u) >> 4 << 16) | (sb >> 4), 0);
          const uint32_t first = kb > kb_begin ? 1u : 0u;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = ((uint64_t)kDescHi << 32) | (a_lo + (uint32_t)(k * (A_MN ? 16 * 128 / 16 : 2)));
            const uint64_t db = ((uint64_t)kDescHi << 32) | (b_lo + (uint32_t)(k * (B_MN ? 16 * 128 / 16 : 2)));
            umma_bf16_lead(leader, d_tmem, da, db, idesc, k > 0 ? 1u : first);
          }
          umma_commit_lead(leader, &empty_bar[stage]);  // frees the smem slot when these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit_lead(leader, &tmem_full[acc]);      // accumulator complete -> epilogue
      }
    }
  }
  } else {
#if AS_WG0_REGS
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(AS_EPI_REGS));
#endif
    // ------------------------------------------------------------------ epilogue (8 warps)
    // Warp e reads TMEM lanes 32*(e%4).. (its hardware quarter) and the column half e/4 of the tile, so two warps per
    // scheduler hide each other's fixed-latency stalls.  tcgen05.ld gives each lane one accumulator row (32 columns);
    // the chunk goes through a per-warp shared-memory tile 16 columns at a time so that global stores are
    // row-contiguous full sectors; bias / ReLU / border mask / dtype conversion happen on the way out.
    const int e = warp - 4;
    const int ew = e & 3;
    const int half = e >> 2;
    float* stg = reinterpret_cast<float*>(bar_base + L::kBarBytes) + e * (32 * kStgPitch);
    constexpr int kChunks = BN / 32;
    const int c_begin = kSplitCols ? half * (kChunks / 2) : 0;
    const int c_end = kSplitCols ? c_begin + kChunks / 2 : kChunks;
    uint32_t it = 0;
    const int subs = p.slab ? 2 : 1;
    for (long long tile = blockIdx.x; tile < num_work; tile += gridDim.x)
    for (int sub = 0; sub < subs; ++sub, ++it) {
      if (!kSplitCols && (int)(it & 1u) != half) continue;     // narrow tiles: this tile belongs to the other warp set
      const int acc = (int)(it % kAcc);
      const uint32_t acc_phase = (it / kAcc) & 1u;
      long long t = tile;
      const int n_blk = (int)(t % p.n_blocks); t /= p.n_blocks;
      int m_blk, b, split;
      if (p.slab) { m_blk = (int)t * 2 + sub; b = 0; split = 0; }
      else { m_blk = (int)(t % p.m_blocks); t /= p.m_blocks; b = (int)(t % batch); split = (int)(t / batch); }
      const int b0 = b % p.nb0, b1 = b / p.nb0;
      const int row_base = m_blk * BM + ew * 32;
      const int n0 = n_blk * BN;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(acc * BN);
      const bool to_ws = p.splits > 1;
      const long long out_off = to_ws ? (((long long)split * batch + b) * p.M) * (long long)p.N
                                      : (long long)b0 * p.c_sb0 + (long long)b1 * p.c_sb1;
      const long long ldo = to_ws ? (long long)p.N : p.ldc;
      const bool f32_out = to_ws || !p.c_bf16;
      if (p.tma_store && !to_ws) {
        // ---- TMA-store path: lane = accumulator row.  Each 128-byte row segment (32 fp32 or 64 bf16 columns) goes to
        // the warp's 128B-swizzled staging tile with eight conflict-free 16-byte stores; one bulk tensor store per
        // warp then sends the 32-row x 128-byte box (tile / matrix edges are clipped by the TMA unit).
        uint8_t* sbuf = bar_base + L::kBarBytes + e * 4096;
        const int grow = row_base + lane;
        bool border = false;
        if (p.img_w > 0) {
          const int q = grow % (p.img_w * p.img_h);
          const int yy = q / p.img_w, xx = q - yy * p.img_w;
          border = (yy == 0) | (yy == p.img_h - 1) | (xx == 0) | (xx == p.img_w - 1);
        }
        constexpr int kHalfCols = kSplitCols ? BN / 2 : BN;           // narrow tiles: one warp set takes all columns
        const int col_begin = kSplitCols ? half * kHalfCols : 0;
        const int col_end = col_begin + kHalfCols;
        const bool need_bf16 = !f32_out || p.C2 != nullptr;
        const CUtensorMap* map16 = f32_out ? &tmC2 : &tmC;
        // Software pipeline over 64-column steps: the tcgen05.ld pair of the NEXT step is issued as soon as the
        // current step's accumulators have been consumed, so its latency hides behind the staging writes and stores.
        uint32_t v0[32], v1[32];
        auto issue = [&](int col) {
          if (col < col_end && n0 + col < p.N) tmem_ld32(t_row + col, v0);
          if (col + 32 < col_end && n0 + col + 32 < p.N) tmem_ld32(t_row + col + 32, v1);
        };
        // alpha / bias / ReLU / border on one 32-column chunk.  Every runtime option is tested once per chunk (warp
        // uniform), never per element.
        auto finish = [&](const uint32_t (&v)[32], int c, float (&f)[32], uint32_t bw) {
          const int gc = n0 + c;
          if (p.bias == nullptr) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * p.alpha;
          } else if (gc + 32 <= p.N && (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0) {
            const float4* b4 = reinterpret_cast<const float4*>(p.bias + gc);    // same address in every lane: broadcast
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 bq = __ldg(b4 + q);
              f[4 * q] = fmaf(__uint_as_float(v[4 * q]), p.alpha, bq.x);
              f[4 * q + 1] = fmaf(__uint_as_float(v[4 * q + 1]), p.alpha, bq.y);
              f[4 * q + 2] = fmaf(__uint_as_float(v[4 * q + 2]), p.alpha, bq.z);
              f[4 * q + 3] = fmaf(__uint_as_float(v[4 * q + 3]), p.alpha, bq.w);
            }
          } else {
            const float bl = (gc + lane < p.N) ? __ldg(p.bias + gc + lane) : 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaf(__uint_as_float(v[j]), p.alpha, __shfl_sync(0xffffffffu, bl, j));
          }
          if (!PLAIN && p.resid != nullptr && grow < p.M && gc + 32 <= p.N) {
            // residual / second gradient contribution with C's layout: this lane's 32 columns are 64 contiguous bytes
            const uint4* r4 = reinterpret_cast<const uint4*>(p.resid + (long long)b0 * p.c_sb0 + (long long)b1 * p.c_sb1 +
                                                             (long long)grow * p.ldc + gc);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint4 rv = __ldg(r4 + q);
              const uint32_t rw[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                f[8 * q + 2 * t] += __uint_as_float(rw[t] << 16);
                f[8 * q + 2 * t + 1] += __uint_as_float(rw[t] & 0xffff0000u);
              }
            }
          }
          if (!PLAIN && p.mask != nullptr && grow < p.M && gc + 32 <= p.N) {
            // ReLU backward of the layer whose output `mask` is: keep the gradient where that output was positive
            const uint4* m4 = reinterpret_cast<const uint4*>(p.mask + (long long)b0 * p.c_sb0 + (long long)b1 * p.c_sb1 +
                                                             (long long)grow * p.ldc + gc);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint4 mv = __ldg(m4 + q);
              const uint32_t mw[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                // bf16 > 0  <=>  sign bit clear and magnitude non-zero
                if (!((mw[t] & 0x7fffu) != 0u && (mw[t] & 0x8000u) == 0u)) f[8 * q + 2 * t] = 0.f;
                if (!((mw[t] & 0x7fff0000u) != 0u && (mw[t] & 0x80000000u) == 0u)) f[8 * q + 2 * t + 1] = 0.f;
              }
            }
          }
          if (!NOBITS && p.bits_in != nullptr) {
            // sign bits of the layer whose ReLU this gradient passes: columns 2k / 2k+1 of the chunk are bits 15-k / 31-k
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = (bw >> (((j & 1) ? 31 : 15) - (j >> 1)) & 1u) ? f[j] : 0.f;
          }
          if (p.relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
          }
          if (p.img_w > 0) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = border ? 0.f : f[j];
          }
        };
        auto stage32 = [&](const float (&f)[32], int c) {
          if (!PLAIN && p.debug >= 4) return;
          if (lane == 0) tma_store_wait_read();                       // previous box has left the staging tile
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(sbuf + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
          if (PLAIN || p.debug < 3) fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0 && (PLAIN || p.debug == 0) && row_base < p.M) {
            tma_store_4d(&tmC, sbuf, n0 + c, row_base, b0, b1);
            tma_store_commit();
          }
        };
        issue(col_begin);
#pragma unroll 1
        for (int col = col_begin; col < col_end; col += 64) {
          const bool has0 = n0 + col < p.N, has1 = (col + 32 < col_end) && (n0 + col + 32 < p.N);   // warp-uniform
          uint32_t bw0 = 0xffffffffu, bw1 = 0xffffffffu;          // requested before the wait for the accumulator columns
          if (!NOBITS && p.bits_in != nullptr && grow < p.M) {
            const uint32_t* brow = p.bits_in + (long long)grow * (p.N >> 5) + ((n0 + col) >> 5);
            if (has0) bw0 = __ldg(brow);
            if (has1) bw1 = __ldg(brow + 1);
          }
          tmem_ld_wait();
          if ((!PLAIN && p.debug == 2) || !has0) { issue(col + 64); continue; }
          uint32_t packed[32];
          if (!f32_out) {
            float f[32];
            finish(v0, col, f, bw0);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              __nv_bfloat162 hv = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
              packed[j] = *reinterpret_cast<uint32_t*>(&hv);
            }
            if (has1) {
              finish(v1, col + 32, f, bw1);
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                __nv_bfloat162 hv = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
                packed[16 + j] = *reinterpret_cast<uint32_t*>(&hv);
              }
            }
            issue(col + 64);
            if (!NOBITS && p.bits_out != nullptr && grow < p.M) {
              // a relu output half-word is positive iff adding 0x7fff carries into its top bit; pair k of a 32-column
              // chunk lands on bits 15-k (even column) and 31-k (odd column)
              uint32_t w0 = 0u, w1 = 0u;
#pragma unroll
              for (int k = 0; k < 16; ++k) {
                w0 |= ((packed[k] + 0x7fff7fffu) & 0x80008000u) >> k;
                w1 |= ((packed[16 + k] + 0x7fff7fffu) & 0x80008000u) >> k;
              }
              uint32_t* brow = p.bits_out + (long long)grow * (p.N >> 5) + ((n0 + col) >> 5);
              if (has1 && ((p.N >> 5) & 1) == 0) *reinterpret_cast<uint2*>(brow) = make_uint2(w0, w1);
              else { brow[0] = w0; if (has1) brow[1] = w1; }
            }
          } else {
            float f[32];
            finish(v0, col, f, bw0);
            if (need_bf16) {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                __nv_bfloat162 hv = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
                packed[j] = *reinterpret_cast<uint32_t*>(&hv);
              }
            }
            stage32(f, col);
            if (has1) {
              finish(v1, col + 32, f, bw1);
              if (need_bf16) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  __nv_bfloat162 hv = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
                  packed[16 + j] = *reinterpret_cast<uint32_t*>(&hv);
                }
              }
              stage32(f, col + 32);
            }
            issue(col + 64);
          }
          if (need_bf16 && (PLAIN || p.debug < 4)) {
            if (lane == 0) tma_store_wait_read();
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (j < 4 || has1)
                *reinterpret_cast<uint4*>(sbuf + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                    make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
            }
            if (PLAIN || p.debug < 3) fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && (PLAIN || p.debug == 0) && row_base < p.M) {
              tma_store_4d(map16, sbuf, n0 + col, row_base, b0, b1);
              tma_store_commit();
            }
          }
        }
      } else
#pragma unroll 1
      for (int c = c_begin; c < c_end; ++c) {
        uint32_t v[32];
        tmem_ld32(t_row + c * 32, v);
        tmem_ld_wait();
        if (n0 + c * 32 >= p.N || p.debug == 2) continue; // warp-uniform
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
          const int col0 = n0 + c * 32 + sub * 16;
#pragma unroll
          for (int q = 0; q < 4; ++q)
            *reinterpret_cast<float4*>(stg + lane * kStgPitch + q * 4) =
                make_float4(__uint_as_float(v[sub * 16 + 4 * q]), __uint_as_float(v[sub * 16 + 4 * q + 1]),
                            __uint_as_float(v[sub * 16 + 4 * q + 2]), __uint_as_float(v[sub * 16 + 4 * q + 3]));
          __syncwarp();
          if (col0 < p.N) {
            if (f32_out) {
              // fp32: 4 lanes cover one 64-byte row segment, 8 rows per instruction
              const int cq = (lane & 3) * 4;
              const int gcol = col0 + cq;
              float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
              if (!to_ws && p.bias != nullptr) {
                if (gcol < p.N) bv.x = __ldg(p.bias + gcol);
                if (gcol + 1 < p.N) bv.y = __ldg(p.bias + gcol + 1);
                if (gcol + 2 < p.N) bv.z = __ldg(p.bias + gcol + 2);
                if (gcol + 3 < p.N) bv.w = __ldg(p.bias + gcol + 3);
              }
              float* obase = (to_ws ? p.ws : reinterpret_cast<float*>(p.C)) + out_off + gcol;
#pragma unroll
              for (int it = 0; it < 4; ++it) {
                const int r = it * 8 + (lane >> 2);
                const int grow = row_base + r;
                float4 a = *reinterpret_cast<const float4*>(stg + r * kStgPitch + cq);
                if (!to_ws) {
                  a.x = a.x * p.alpha + bv.x; a.y = a.y * p.alpha + bv.y; a.z = a.z * p.alpha + bv.z; a.w = a.w * p.alpha + bv.w;
                  if (p.relu) { a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); a.z = fmaxf(a.z, 0.f); a.w = fmaxf(a.w, 0.f); }
                  if (p.img_w > 0) {
                    const int q = grow % (p.img_w * p.img_h);
                    const int yy = q / p.img_w, xx = q - yy * p.img_w;
                    if ((yy == 0) | (yy == p.img_h - 1) | (xx == 0) | (xx == p.img_w - 1)) a = make_float4(0.f, 0.f, 0.f, 0.f);
                  }
                }
                if (grow < p.M && gcol < p.N && p.debug == 0) {
                  float* dst = obase + (long long)grow * ldo;
                  if (gcol + 4 <= p.N && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
                    *reinterpret_cast<float4*>(dst) = a;
                  } else {
                    dst[0] = a.x;
                    if (gcol + 1 < p.N) dst[1] = a.y;
                    if (gcol + 2 < p.N) dst[2] = a.z;
                    if (gcol + 3 < p.N) dst[3] = a.w;
                  }
                  if (!to_ws && p.C2 != nullptr) {
                    __nv_bfloat16* d2 = reinterpret_cast<__nv_bfloat16*>(p.C2) + out_off + (long long)grow * ldo + gcol;
                    if (gcol + 4 <= p.N && (reinterpret_cast<uintptr_t>(d2) & 7) == 0) {
                      __nv_bfloat162 h0 = __floats2bfloat162_rn(a.x, a.y), h1 = __floats2bfloat162_rn(a.z, a.w);
                      uint2 u;
                      u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
                      *reinterpret_cast<uint2*>(d2) = u;
                    } else {
                      d2[0] = __float2bfloat16_rn(a.x);
                      if (gcol + 1 < p.N) d2[1] = __float2bfloat16_rn(a.y);
                      if (gcol + 2 < p.N) d2[2] = __float2bfloat16_rn(a.z);
                      if (gcol + 3 < p.N) d2[3] = __float2bfloat16_rn(a.w);
                    }
                  }
                }
              }
            } else {
              // bf16: 2 lanes cover one 32-byte row segment, 16 rows per instruction
              const int cq = (lane & 1) * 8;
              const int gcol = col0 + cq;
              float bv[8];
#pragma unroll
              for (int k = 0; k < 8; ++k) bv[k] = (p.bias != nullptr && gcol + k < p.N) ? __ldg(p.bias + gcol + k) : 0.f;
              __nv_bfloat16* obase = reinterpret_cast<__nv_bfloat16*>(p.C) + out_off + gcol;
#pragma unroll
              for (int it = 0; it < 2; ++it) {
                const int r = it * 16 + (lane >> 1);
                const int grow = row_base + r;
                const float4 a0 = *reinterpret_cast<const float4*>(stg + r * kStgPitch + cq);
                const float4 a1 = *reinterpret_cast<const float4*>(stg + r * kStgPitch + cq + 4);
                float f[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                bool border = false;
                if (p.img_w > 0) {
                  const int q = grow % (p.img_w * p.img_h);
                  const int yy = q / p.img_w, xx = q - yy * p.img_w;
                  border = (yy == 0) | (yy == p.img_h - 1) | (xx == 0) | (xx == p.img_w - 1);
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                  float x = f[k] * p.alpha + bv[k];
                  if (p.relu) x = fmaxf(x, 0.f);
                  f[k] = border ? 0.f : x;
                }
                if (grow < p.M && gcol < p.N && p.debug == 0) {
                  __nv_bfloat16* dst = obase + (long long)grow * ldo;
                  if (gcol + 8 <= p.N && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
                    __nv_bfloat162 h0 = __floats2bfloat162_rn(f[0], f[1]), h1 = __floats2bfloat162_rn(f[2], f[3]);
                    __nv_bfloat162 h2 = __floats2bfloat162_rn(f[4], f[5]), h3 = __floats2bfloat162_rn(f[6], f[7]);
                    uint4 u;
                    u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
                    u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
                    *reinterpret_cast<uint4*>(dst) = u;
                  } else {
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                      if (gcol + k < p.N) dst[k] = __float2bfloat16_rn(f[k]);
                  }
                }
              }
            }
          }
          __syncwarp();                                  // staging tile is reused by the next half chunk
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
  }

  if (warp >= 4 && lane == 0) tma_store_wait_all();     // bulk stores must have drained before shared memory goes away
  tc_fence_before();
  __syncthreads();
  if (p.mc) cluster_sync_all();                          // the peer may still multicast into / arrive on this CTA
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// Operand = up to 4-D bf16 tensor: dims (inner contiguous, rows, b0, b1) with element strides.
static int make_map(CUtensorMap* tm, const void* ptr, long long inner, long long rows, long long ld, int nb0,
                    long long sb0, int nb1, long long sb1, int box_inner, int box_rows, int es = 2) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return AS_EDRIVER;
  // the encode is a driver-API call: a thread that has not touched the runtime yet (e.g. autograd's backward thread
  // when its first kernel is one of ours) has no current context and gets CUDA_ERROR_INVALID_CONTEXT
  static thread_local bool ctx_ready = false;
  if (!ctx_ready) { cudaFree(nullptr); ctx_ready = true; }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || ((ld * es) & 15)) return AS_EALIGN;
  cuuint64_t dims[4] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)nb0, (cuuint64_t)nb1};
  long long s0 = nb0 > 1 ? sb0 : ld * rows, s1 = nb1 > 1 ? sb1 : s0 * nb0;
  if (((s0 * es) & 15) || ((s1 * es) & 15)) return AS_EALIGN;
  cuuint64_t strides[3] = {(cuuint64_t)(ld * es), (cuuint64_t)(s0 * es), (cuuint64_t)(s1 * es)};
  cuuint32_t box[4] = {(cuuint32_t)box_inner, (cuuint32_t)box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(tm, es == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                   const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS && knobs().debug_tmap)
    fprintf(stderr, "tensor map encode failed: CUresult %d ptr %p dims %llu %llu %llu %llu strides %llu %llu %llu box %u %u es %d\n",
            (int)r, ptr, (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)dims[2],
            (unsigned long long)dims[3], (unsigned long long)strides[0], (unsigned long long)strides[1],
            (unsigned long long)strides[2], box[0], box[1], es);
  return r == CUDA_SUCCESS ? AS_OK : AS_EDRIVER;
}

static int g_num_sms = 0;
static int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = kNumSMsB200;
  }
  return g_num_sms;
}

// used by attn.cu
int make_map_bf16(CUtensorMap* tm, const void* ptr, long long inner, long long rows, long long ld, int nb0, long long sb0,
                  int nb1, long long sb1, int box_inner, int box_rows) {
  return make_map(tm, ptr, inner, rows, ld, nb0, sb0, nb1, sb1, box_inner, box_rows);
}
int gemm_num_sms() { return num_sms(); }

template <int BN, bool A_MN, bool B_MN, int EPI = 0>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tc2,
                  const CUtensorMap& tb2, const GemmDev& p, cudaStream_t st) {
  using L = SmemLayout<BN>;
  static bool attr_set = false;
  auto kern = gemm_kernel<BN, A_MN, B_MN, EPI>;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  long long tiles = (long long)p.splits * p.nb0 * p.nb1 * p.m_blocks * p.n_blocks;
  if (p.slab) tiles = (long long)((p.m_blocks + 1) / 2) * p.n_blocks;
  int grid = (int)(tiles < num_sms() ? tiles : num_sms());
  if (grid <= 0) return AS_OK;
  if (p.mc) {
    // CTA pairs (thread-block clusters of 2): blockIdx.x = 2 * pair + rank, both ranks walk the step list in lockstep
    grid = (int)(((tiles + 1) & ~1LL) < num_sms() ? ((tiles + 1) & ~1LL) : (num_sms() & ~1));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = L::kTotal;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, tc2, tb2, p);
    if (e != cudaSuccess) return (int)e;
    return AS_OK;
  }
  kern<<<grid, kGemmThreads, L::kTotal, st>>>(ta, tb, tc, tc2, tb2, p);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

template <int BN>
static int dispatch_major(bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                          const CUtensorMap& tc2, const CUtensorMap& tb2, const GemmDev& p, cudaStream_t st) {
  if constexpr (BN >= 128) {
    // row-major A with a plain epilogue: the Transformer's forward / data-gradient GEMMs
    if (p.plain == 1 && !a_mn && !b_mn) return launch<BN, false, false, 1>(ta, tb, tc, tc2, tb2, p, st);
    if (p.plain == 1 && !a_mn && b_mn) return launch<BN, false, true, 1>(ta, tb, tc, tc2, tb2, p, st);
  }
  if constexpr (BN >= 128) {
    // sign bits: the FFN pair (forward GEMM that writes them, data-gradient GEMM that masks from them) and the conv chains
    if (p.plain == 2 && !a_mn && !b_mn) return launch<BN, false, false, 2>(ta, tb, tc, tc2, tb2, p, st);
  }
  if constexpr (BN == 256) {
    if (p.plain == 2 && !a_mn && b_mn) return launch<BN, false, true, 2>(ta, tb, tc, tc2, tb2, p, st);
  }
  if (!a_mn && !b_mn) return launch<BN, false, false>(ta, tb, tc, tc2, tb2, p, st);
  if (!a_mn && b_mn) return launch<BN, false, true>(ta, tb, tc, tc2, tb2, p, st);
  if (a_mn && !b_mn) return launch<BN, true, false>(ta, tb, tc, tc2, tb2, p, st);
  return launch<BN, true, true>(ta, tb, tc, tc2, tb2, p, st);
}

static int pick_bn(int N, bool b_mn) {
  if (b_mn) return N > 128 ? 256 : (N > 64 ? 128 : 64);
  if (N > 128) return 256;
  if (N > 64) return 128;
  if (N > 32) return 64;
  return 32;
}

int gemm_set_profile(long long* buf) { return (int)cudaMemcpyToSymbol(g_gemm_prof, &buf, sizeof(buf)); }

int g_gemm_debug = 0;
int g_gemm_no_tma_store = 0;
int g_gemm_no_slab = 0;       // diagnostics: force the per-tap convolution mode

int gemm_run(const GemmCall& g, cudaStream_t st) {
  if (g.M <= 0 || g.N <= 0 || g.K <= 0 || g.nb0 <= 0 || g.nb1 <= 0) return AS_ESIZE;
  if (!g.A || !g.B || (!g.C && g.splits <= 1)) return AS_EINVAL;
  if (g.splits > 1 && !g.ws) return AS_EINVAL;
  int BN = g.bn_override ? g.bn_override : pick_bn(g.N, g.b_mn);
  if (g.b_mn && BN < 64) return AS_EUNSUPPORTED;
  GemmDev p{};
  p.M = g.wg_pair ? 128 : g.M; p.N = g.N; p.K = g.K; p.nb0 = g.nb0; p.nb1 = g.nb1;
  p.wg_pair = g.wg_pair;
  if (g.wg_pair) {
    if (g.M != 64 || !g.a_mn || !g.b_mn || g.wgrad_taps != 5 || g.nb0 != 5 || g.splits <= 1) return AS_EUNSUPPORTED;
    for (int i = 0; i < 5; ++i) { p.pair_a[i][0] = g.pair_a[i][0]; p.pair_a[i][1] = g.pair_a[i][1]; p.pair_b[i] = g.pair_b[i]; }
  }
  p.m_blocks = (g.M + BM - 1) / BM;
  p.n_blocks = (g.N + BN - 1) / BN;
  p.taps = g.taps; p.cin = g.K;
  if (g.taps > 0) {
    // Cin need not be a multiple of 64: the A box is zero-filled past Cin, which also cancels whatever the B box
    // picks up from the next tap's columns.
    if (g.taps > 9 || (g.K % 8) != 0 || g.a_mn || g.nb0 * g.nb1 != 1) return AS_EUNSUPPORTED;
    p.cin_blocks = (g.K + BK - 1) / BK;
    p.num_kb = g.taps * p.cin_blocks;
    for (int i = 0; i < g.taps; ++i) p.shift[i] = g.shift[i];
  } else {
    p.cin_blocks = 1;
    p.num_kb = (g.K + BK - 1) / BK;
  }
  p.img_w = g.img_w; p.img_h = g.img_h;
  p.b_k_shift = g.b_mn ? g.b_k_shift : 0;
  p.wgrad_taps = g.wgrad_taps;
  if (g.wgrad_taps > 0) {
    if (g.wgrad_taps > 9 || g.nb0 != g.wgrad_taps || !g.b_mn || g.taps > 0) return AS_EUNSUPPORTED;
    for (int i = 0; i < g.wgrad_taps; ++i) p.shift[i] = g.shift[i];
  }
  p.splits = g.splits > 1 ? g.splits : 1;
  p.kb_per_split = (p.num_kb + p.splits - 1) / p.splits;
  p.splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;   // drop empty splits
  if (g.splits > 1 && p.splits != g.splits) return AS_ESIZE;     // caller sizes ws from as_gemm_splits()
  p.C = g.C; p.C2 = g.C2; p.c_bf16 = g.c_bf16; p.ldc = g.ldc; p.c_sb0 = g.c_sb0; p.c_sb1 = g.c_sb1;
  p.resid = reinterpret_cast<const __nv_bfloat16*>(g.resid);
  p.mask = reinterpret_cast<const __nv_bfloat16*>(g.mask);
  p.bits_out = g.bits_out;
  p.bits_in = g.bits_in;
  p.bias = g.bias; p.relu = g.relu; p.alpha = g.alpha; p.ws = g.ws;
  p.debug = g_gemm_debug;
  // slab mode for multi-tap convolutions (see the header comment): needs the halo'd slab of a 256-row step to fit two
  // TMA boxes of <= 256 rows and three slabs plus >= 3 weight tiles to fit the operand ring
  p.slab = 0;
  int a_box_rows = BM;
  if (g.taps > 1 && !g.a_mn && !g.b_mn && BN <= 128 && !g_gemm_no_slab) {
    int halo = 0;
    for (int i = 0; i < g.taps; ++i) halo = halo > abs(g.shift[i]) ? halo : abs(g.shift[i]);
    const int rows = 2 * BM + 2 * halo;
    const int box = (((rows + 1) / 2) + 7) & ~7;
    const int slab_bytes = 2 * box * 128;                 // multiple of 2048
    const int ring = 192 * 1024;
    const int b_bytes = BN * 128;
    const int max_slots = (BN >= 128) ? 6 : 8;            // = kStages of the instantiation (barrier array size)
    const int cin_blocks = (g.K + BK - 1) / BK;
    const long long all_b = (long long)cin_blocks * g.taps * b_bytes;
    if (box <= 256) {
      if (p.n_blocks == 1 && 2 * slab_bytes + all_b <= ring && !knobs().slab_no_resident) {
        // every weight tile stays in shared memory for the whole kernel; two slabs double-buffer the pixels
        p.slab = 1; p.slab_slots = 2; p.b_resident = 1; p.tpb = 1; p.b_slots = 1;
      } else {
        // ring of weight tiles, `tpb` taps per slot (one barrier round trip per slot) and three slabs
        int tpb = BN >= 128 ? 1 : 3;
        if (g.taps % tpb != 0) tpb = 1;
        int slots = (ring - 3 * slab_bytes) / (tpb * b_bytes);
        if (slots < 3 && tpb > 1) { tpb = 1; slots = (ring - 3 * slab_bytes) / b_bytes; }
        if (slots > max_slots) slots = max_slots;
        if (knobs().slab_bslots >= 2 && knobs().slab_bslots < slots) slots = knobs().slab_bslots;
        if (slots >= 2) { p.slab = 1; p.slab_slots = 3; p.b_resident = 0; p.tpb = tpb; p.b_slots = slots; }
        // weight multicast across a CTA pair halves the L2 -> SM weight stream, but measured 4 % SLOWER than independent
        // CTAs on the 128-channel convs (the issuer, not the weight stream, is the limit there): opt-in only
        if (p.slab && p.n_blocks == 1 && BN >= 64 && (p.m_blocks + 1) / 2 >= 2 && knobs().slab_mc) p.mc = 1;
      }
      if (p.slab) { p.halo = halo; p.slab_box = box; p.slab_bytes = slab_bytes; a_box_rows = box; }
      if (p.slab && g.K <= 32 && !knobs().slab_no_halfk) p.half_k = 1;
    }
  }

  CUtensorMap ta, tb;
  int rc;
  const long long k_total_b = g.taps > 0 ? (long long)g.taps * g.K : g.K;
  const int onb0 = g.wgrad_taps > 0 ? 1 : g.nb0;      // tensor-map extent of the first batch dim
  if (g.a_mn) rc = make_map(&ta, g.A, g.M, g.K, g.lda, onb0, g.a_sb0, g.nb1, g.a_sb1, 64, BK);
  else rc = make_map(&ta, g.A, g.K, g.a_rows > 0 ? g.a_rows : g.M, g.lda, onb0, g.a_sb0, g.nb1, g.a_sb1, BK, a_box_rows);
  if (rc) return rc;
  if (g.b_mn) rc = make_map(&tb, g.B, g.N, k_total_b, g.ldb, onb0, g.b_sb0, g.nb1, g.b_sb1, 64, BK);
  else rc = make_map(&tb, g.B, k_total_b, g.N, g.ldb, onb0, g.b_sb0, g.nb1, g.b_sb1, BK, BN);
  if (rc) return rc;
  CUtensorMap tb2 = tb;                              // half-tile box for the multicast loads of a CTA pair
  if (p.mc) {
    rc = make_map(&tb2, g.B, k_total_b, g.N, g.ldb, onb0, g.b_sb0, g.nb1, g.b_sb1, BK, BN / 2);
    if (rc) return rc;
  }
  // output tensor maps for the TMA-store epilogue (fall back to register stores when C is not 16-byte pitched)
  CUtensorMap tc = ta, tc2 = ta;
  p.tma_store = 0;
  if (p.splits == 1 && g.C != nullptr && !g_gemm_no_tma_store) {
    const int es = g.c_bf16 ? 2 : 4;
    const long long c_rows = (long long)p.m_blocks * BM;      // rows beyond M are clipped through the map's extent
    (void)c_rows;
    int ok = make_map(&tc, g.C, g.N, g.M, g.ldc, g.nb0, g.c_sb0, g.nb1, g.c_sb1, 128 / es, 32, es) == AS_OK;
    if (ok && g.C2 != nullptr) ok = make_map(&tc2, g.C2, g.N, g.M, g.ldc, g.nb0, g.c_sb0, g.nb1, g.c_sb1, 64, 32, 2) == AS_OK;
    // a bf16 box is 64 columns wide: with 32-wide tiles it would spill into the neighbouring column tile
    if ((g.c_bf16 || g.C2 != nullptr) && BN < 64 && p.n_blocks > 1) ok = 0;
    p.tma_store = ok;
  }
  // the residual add lives in the TMA-store epilogue only; whole 32-column chunks, 16-byte aligned rows
  if (g.resid != nullptr && (!p.tma_store || (g.N % 32) != 0 || (g.ldc % 8) != 0 ||
                             (reinterpret_cast<uintptr_t>(g.resid) & 15) != 0))
    return AS_EUNSUPPORTED;
  if ((g.bits_in != nullptr || g.bits_out != nullptr) &&
      (!p.tma_store || (g.N % 32) != 0 || g.nb0 * g.nb1 != 1 || (g.bits_out != nullptr && (!g.relu || !g.c_bf16 || g.C2 != nullptr))))
    return AS_EUNSUPPORTED;
  if (g.mask != nullptr && (!p.tma_store || (g.N % 32) != 0 || (g.ldc % 8) != 0 ||
                            (reinterpret_cast<uintptr_t>(g.mask) & 15) != 0))
    return AS_EUNSUPPORTED;
  p.plain = 0;
  if (p.tma_store && g.resid == nullptr && g.mask == nullptr && g_gemm_debug == 0 && !knobs().no_plain_epilogue)
    p.plain = (g.bits_in == nullptr && g.bits_out == nullptr) ? 1 : 2;
  switch (BN) {
    case 32: return dispatch_major<32>(g.a_mn, g.b_mn, ta, tb, tc, tc2, tb2, p, st);
    case 64: return dispatch_major<64>(g.a_mn, g.b_mn, ta, tb, tc, tc2, tb2, p, st);
    case 128: return dispatch_major<128>(g.a_mn, g.b_mn, ta, tb, tc, tc2, tb2, p, st);
    case 256: return dispatch_major<256>(g.a_mn, g.b_mn, ta, tb, tc, tc2, tb2, p, st);
  }
  return AS_EUNSUPPORTED;
}

// ---- split-K reduction: out[i] = (accumulate ? out[i] : 0) + sum_s ws[s][i], fixed order (deterministic)
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, float* __restrict__ out, long long n, int splits,
                                     int accumulate) {
  // 16-byte loads, four splits in flight per thread; summation order over splits is fixed
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 s = accumulate ? reinterpret_cast<const float4*>(out)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    int k = 0;
    for (; k + 8 <= splits; k += 8) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldg(reinterpret_cast<const float4*>(ws + (long long)(k + u) * n) + i);
#pragma unroll
      for (int u = 0; u < 8; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
    }
    for (; k < splits; ++k) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(ws + (long long)k * n) + i);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    reinterpret_cast<float4*>(out)[i] = s;
  }
  for (long long i = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float s = accumulate ? out[i] : 0.f;
    for (int k = 0; k < splits; ++k) s += ws[(long long)k * n + i];
    out[i] = s;
  }
}
__global__ void splitk_reduce_scalar_kernel(const float* __restrict__ ws, float* __restrict__ out, long long n, int splits,
                                            int accumulate) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    float s = accumulate ? out[i] : 0.f;
    for (int k = 0; k < splits; ++k) s += ws[(long long)k * n + i];
    out[i] = s;
  }
}

// out[t][co][ci] (t < taps) from the tap-pair partials ws[split][5][128][N]: row t * 64 + m of a split's [640][N] matrix;
// plain (m = co, N = cin) or transposed (m = ci, N = cout: the operands were swapped to get a 64-channel M side)
__global__ void wgrad_pair_reduce_kernel(const float* __restrict__ ws, float* __restrict__ out, int splits, long long stride,
                                         int total, int cout, int cin, int transposed, int swap_halves) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int ci = i % cin, co = (i / cin) % cout, t = i / (cin * cout);
  // row of tap t inside a split's [5 pairs][128][N] block; swap_halves: the later tap of a pair sits in rows 0-63
  const int half = swap_halves ? ((t & 1) ^ 1) : (t & 1);
  const long long row0 = (long long)(t >> 1) * 128 + half * 64;
  const long long src = transposed ? (row0 + ci) * cout + co : (row0 + co) * cin + ci;
  float a = 0.f;
  for (int s = 0; s < splits; ++s) a += ws[(long long)s * stride + src];
  out[i] = a;
}

int wgrad_pair_reduce(const float* ws, float* out, int splits, long long stride, int taps, int cout, int cin, int transposed,
                      int swap_halves, cudaStream_t st) {
  const int total = taps * cout * cin;
  wgrad_pair_reduce_kernel<<<(total + 127) / 128, 128, 0, st>>>(ws, out, splits, stride, total, cout, cin, transposed,
                                                                swap_halves);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int splitk_reduce(const float* ws, float* out, long long n, int splits, int accumulate, cudaStream_t st) {
  if (n <= 0) return AS_OK;
  const bool vec = (n % 4 == 0) && ((reinterpret_cast<uintptr_t>(ws) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  long long blocks = ((vec ? n / 4 : n) + 63) / 64;
  int grid = (int)(blocks < 148 * 16 ? (blocks > 0 ? blocks : 1) : 148 * 16);
  if (vec) splitk_reduce_kernel<<<grid, 64, 0, st>>>(ws, out, n, splits, accumulate);
  else splitk_reduce_scalar_kernel<<<grid, 64, 0, st>>>(ws, out, n, splits, accumulate);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
