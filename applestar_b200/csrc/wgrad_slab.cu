// 3x3 convolution weight gradient for layers whose M side has 64 channels (the 76x80 maps: 32 -> 64 and 64 -> 32), all
// nine taps from ONE pass over the pixels.
//
//   dW[tap][co][ci] = sum_p dy[p][co] * x[p + shift(tap)][ci]          (padded NHWC bf16 maps, p = padded pixel index)
//
// The tap-per-tile GEMM (gemm_tcgen05.cu, wgrad_taps) streams dy and x through shared memory once per tap (pair), which
// makes these launches L2-bandwidth bound.  Here a CTA owns a contiguous range of pixels (split-K over the grid) and per
// 64-pixel block loads ONE slab of each operand with its halo; every tap pair is then an MMA whose operand descriptors
// merely start at different rows of the slabs:
//   * A is MN-major [pixel rows][64 channels = one 128-byte row].  Its two 64-row halves (M = 128) are two "chunks" of
//     the descriptor, LBO bytes apart -- so chunk 0 / chunk 1 can be the SAME slab at two different pixel offsets, i.e.
//     two taps per instruction (accumulator rows 0-63 / 64-127).
//   * B is MN-major [pixel rows][<= 64 channels], one chunk, started at the tap's pixel offset.
//   The 128-byte swizzle is a function of the absolute shared-memory address, so row-granular start offsets are legal
//   (same argument as the forward slab convolution).
// Five accumulators (tap pairs) of 64 fp32 columns live in TMEM for the whole range; the epilogue writes the partials
// [split][5][128][N], reduced by wgrad_pair_reduce.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "gemm.h"

namespace as {

constexpr int kWsThreads = 192;           // warp 0: TMA, warp 1: MMA, warps 2-5: epilogue
constexpr int kWsBK = 64;                 // pixels per block
constexpr int kWsStages = 4;

struct WgradSlabDev {
  int num_kb, kb_per_split;
  int rows_a, rows_b;                     // slab heights (box rows)
  int halo_a, halo_b;                     // slab row 0 = pixel k0 - halo
  int bytes_a, bytes_b;                   // slab sizes rounded up to 1024
  int a_row[5], a_lbo[5], b_row[5];       // per tap pair: chunk-0 row of A, chunk distance in bytes, row of B
  int N;                                  // real columns of B (<= 128: blockIdx.y selects a 64-column half)
  float* ws;                              // [gridDim][5][128][N]
};

__global__ void __launch_bounds__(kWsThreads, 1)
wgrad_slab_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const WgradSlabDev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int stage_bytes = p.bytes_a + p.bytes_b;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kWsStages * stage_bytes);
  uint64_t* empty_bar = full_bar + kWsStages;
  uint64_t* done_bar = empty_bar + kWsStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.y * 64;
  const int kb_begin = blockIdx.x * p.kb_per_split;
  const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);

  if (threadIdx.x == 0) {
    for (int i = 0; i < kWsStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(done_bar, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * stage_bytes;
        mbar_expect_tx(&full_bar[stage], (uint32_t)((p.rows_a + p.rows_b) * 128));
        tma_load_4d(sa, &tmA, &full_bar[stage], 0, kb * kWsBK - p.halo_a, 0, 0);
        tma_load_4d(sa + p.bytes_a, &tmB, &full_bar[stage], n0, kb * kWsBK - p.halo_b, 0, 0);
        if (++stage == kWsStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // instruction descriptor: fp32 accumulate, bf16 operands, both MN-major, N = 64, M = 128
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb_begin; kb < kb_end; ++kb) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      const uint32_t sa = smem_u32(smem + stage * stage_bytes);
      const uint32_t sb = sa + (uint32_t)p.bytes_a;
#pragma unroll
      for (int b = 0; b < 5; ++b) {
        const uint32_t a_lo = __shfl_sync(0xffffffffu, ((uint32_t)(p.a_lbo[b] >> 4) << 16) | ((sa + (uint32_t)(p.a_row[b] * 128)) >> 4), 0);
        const uint32_t b_lo = __shfl_sync(0xffffffffu, ((uint32_t)(kWsBK * 128 >> 4) << 16) | ((sb + (uint32_t)(p.b_row[b] * 128)) >> 4), 0);
#pragma unroll
        for (int k = 0; k < kWsBK / 16; ++k) {
          const uint64_t da = ((uint64_t)kDescHi << 32) | (a_lo + (uint32_t)(k * 128));
          const uint64_t db = ((uint64_t)kDescHi << 32) | (b_lo + (uint32_t)(k * 128));
          umma_bf16_lead(0, tmem_base + (uint32_t)(b * 64), da, db, idesc, (kb > kb_begin || k > 0) ? 1u : 0u);
        }
      }
      umma_commit_lead(0, &empty_bar[stage]);
      if (++stage == kWsStages) { stage = 0; phase ^= 1; }
    }
    umma_commit_lead(0, done_bar);
  } else {
    // epilogue: warp w reads TMEM lanes 32 (w % 4) .. +31 (its hardware quarter); thread = accumulator row
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    if (kb_end > kb_begin) {
      mbar_wait(done_bar, 0);
      tc_fence_after();
    }
    for (int b = 0; b < 5; ++b) {
      float* dst = p.ws + (((long long)blockIdx.x * 5 + b) * 128 + row) * p.N + n0;
      const int ncols = min(64, p.N - n0);
      for (int c0 = 0; c0 < ncols; c0 += 32) {
        uint32_t v[32];
        if (kb_end > kb_begin) {
          tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(b * 64 + c0), v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          if (c0 + j < ncols)
            *reinterpret_cast<float4*>(dst + c0 + j) = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                                   __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
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

// mode 1: A = dy [P][64], B = x [P][Cin <= 64]     accumulator rows 0-63 = tap 2b+1, 64-127 = tap 2b
// mode 2: A = x  [P][64], B = dy [P][Cout <= 64]   accumulator rows 0-63 = tap 2b,   64-127 = tap 2b+1 (transposed dW)
int wgrad_slab_ok(int mode, int Cin, int Cout, int pw) {
  if (mode == 1 && !(Cout == 64 && Cin <= 128 && Cin % 8 == 0)) return 0;
  if (mode == 2 && !(Cin == 64 && Cout <= 128 && Cout % 8 == 0)) return 0;
  if (mode != 1 && mode != 2) return 0;
  return kWsBK + 2 * (pw + 1) <= 256;           // slab box height limit
}

int wgrad_slab(int mode, const void* dy, const void* x, float* ws, long long P, int pw, int Cin, int Cout, int splits,
               cudaStream_t st) {
  if (!wgrad_slab_ok(mode, Cin, Cout, pw)) return AS_EUNSUPPORTED;
  if (!dy || !x || !ws || splits <= 0) return AS_EINVAL;
  int shift[10];
  for (int ky = 0; ky < 3; ++ky)
    for (int kx = 0; kx < 3; ++kx) shift[ky * 3 + kx] = (ky - 1) * pw + (kx - 1);
  shift[9] = shift[8];
  WgradSlabDev p{};
  p.num_kb = (int)((P + kWsBK - 1) / kWsBK);
  p.kb_per_split = (p.num_kb + splits - 1) / splits;
  p.ws = ws;
  const void* A;
  const void* B;
  int ca, cb;
  if (mode == 1) {
    A = dy; B = x; ca = Cout; cb = Cin;
    int dmax = 0;
    for (int b = 0; b < 5; ++b) dmax = max(dmax, shift[2 * b + 1] - shift[2 * b]);
    p.halo_a = (dmax + 7) & ~7;
    p.rows_a = kWsBK + p.halo_a;
    p.halo_b = pw + 1;
    p.rows_b = kWsBK + 2 * (pw + 1);
    for (int b = 0; b < 5; ++b) {
      const int d = shift[2 * b + 1] - shift[2 * b];
      p.a_row[b] = p.halo_a - d;            // chunk 0: dy rows read d pixels earlier -> the LATER tap of the pair
      p.a_lbo[b] = d * 128;                 // chunk 1: dy rows at the block itself   -> the earlier tap
      p.b_row[b] = p.halo_b + shift[2 * b];
    }
  } else {
    A = x; B = dy; ca = Cin; cb = Cout;
    p.halo_a = pw + 1;
    p.rows_a = kWsBK + 2 * (pw + 1);
    p.halo_b = 0;
    p.rows_b = kWsBK;
    for (int b = 0; b < 5; ++b) {
      p.a_row[b] = p.halo_a + shift[2 * b];
      p.a_lbo[b] = (shift[2 * b + 1] - shift[2 * b]) * 128;
      p.b_row[b] = 0;
    }
  }
  p.N = cb;
  p.bytes_a = (p.rows_a * 128 + 1023) & ~1023;
  p.bytes_b = (p.rows_b * 128 + 1023) & ~1023;
  CUtensorMap ta, tb;
  int rc = make_map_bf16(&ta, A, ca, P, ca, 1, 0, 1, 0, 64, p.rows_a);
  if (rc) return rc;
  rc = make_map_bf16(&tb, B, cb, P, cb, 1, 0, 1, 0, 64, p.rows_b);
  if (rc) return rc;
  const int smem = kWsStages * (p.bytes_a + p.bytes_b) + 256 + 1024;
  if (smem > 232448) return AS_ESIZE;
  static int attr_smem = 0;
  if (smem > attr_smem) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_slab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    attr_smem = smem;
  }
  wgrad_slab_kernel<<<dim3((unsigned)splits, (unsigned)((cb + 63) / 64)), kWsThreads, smem, st>>>(ta, tb, p);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
