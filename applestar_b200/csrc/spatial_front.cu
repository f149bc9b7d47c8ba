// Front of the spatial encoder as ONE kernel per direction (reference: model/encoder.py:37-42 ScatterConnection 'add',
// obs_encoder/spatial_encoder.py:52-72 one-hot planes + effect planes + 1x1 project conv 56 -> 32 + ReLU, :79 first
// 2x2 max-pool).  The 56-channel input of the project convolution is almost entirely one-hot, so the convolution is
// evaluated EXACTLY and sparsely in fp32 on the CUDA cores instead of as a dense GEMM over a materialised map:
//
//   pre[o] = b[o] + W[o,0] h/256 + T1[vis,pr][o] + T2[creep,alerts,pathable,buildable][o]        (two table rows)
//          + sum_{effect e at this pixel} W[o,18+e] + contrib[cell][o]                            (rare)
//   contrib[cell] = W[:,24:56] . (sum of the projected embeddings of the entities on that cell, ascending entity order)
//
// and the kernel writes only max over the 2x2 window of relu(pre) as a zero-padded NHWC bf16 map [R, H/2+2, W/2+2, 32]:
// the 64-channel bf16 input map (3.3 GB per learner step), the full-resolution project output (1.7 GB) and the separate
// pool pass are never formed.  HBM traffic per observation: 7 planes (170 KB) read, 6080 x 64 B written.
//
// The backward recomputes pre for the window of every pooled pixel (same arithmetic, so ReLU mask and arg-max agree
// bit for bit with the forward), routes d pooled to the arg-max pixel, and accumulates
//   d T1 / d T2 / d(height weight) / d(effect weights)  in per-thread private shared-memory tables (no atomics: deterministic),
//   g_entity[r,e,:] = routed gradient at the cell of entity e  (global; feeds d proj = g W_s and d W_s = g^T emb).
// Three kernels per direction: the tables (weights only, one CTA), the per-observation cell list (sort of the <= 512
// live entities by cell, segment sums in ascending entity order, contribution vectors), and the dense kernel, one CTA
// per band of kBandRows input rows of one observation, whose set-up is a few coalesced copies.  The cell list and the
// tables written by the forward are reused by the backward.
#include "common.cuh"

namespace as {

constexpr int kFrontC = 32;            // output channels of the project conv
constexpr int kFrontCin = 56;
constexpr int kBandRows = 38;          // input rows per CTA (H = 152 -> 4 bands); even
constexpr int kFrontMaxN = 512;
constexpr int kT1Rows = 20, kT2Rows = 16;
constexpr int kPitch = 36;             // table row pitch (floats): 16-byte aligned rows (128-bit shared loads), rows spread over the banks

struct FrontArgs {
  const uint8_t* plane[7];             // height, visibility, creep, player_relative, alerts, pathable, buildable [R,H,W]
  const int16_t* effect[6];            // [R, n_eff] flat pixel ids, zero padded (pixel 0 quirk, spatial_encoder.py:61-69)
  int n_eff;
  const float* emb;                    // [R, N, 32] projected entity embeddings (fp32)
  const void* xs; const void* ys;      // [R, N] uint8 or int64
  int idx_bytes;
  const long long* entity_num;         // [R] or nullptr
  int N;
  const float* weight;                 // [32, 56] fp32 (Conv2d weight [32,56,1,1])
  const float* bias;                   // [32]
  int H, W;
};

struct FrontTables {                   // global memory, written by front_tables_kernel
  float t1[kT1Rows * kPitch];          // bias + visibility x player_relative
  float t2[kT2Rows * kPitch];          // creep x alerts x pathable x buildable
  float wh[kFrontC];                   // height weight (already / 256)
  float weff[6 * kPitch];
  float ws_t[32 * kPitch];             // ws_t[c][o] = W[o][24 + c]
};
constexpr int kTabFloats = (kT1Rows + kT2Rows + 6) * kPitch + kFrontC;      // t1, t2, wh, weff: what the dense kernels copy

// layout of the `cells` buffer shared by forward and backward
struct FrontCells {
  const FrontTables* tabs;
  const int* ncell;                    // [R] live entities
  const uint32_t* skey;                // [R][512] (cell << 9 | entity), ascending
  const __nv_bfloat16* contrib;        // [R][512][32], valid at segment heads
};
static size_t cells_tab_bytes() { return (sizeof(FrontTables) + 255) & ~size_t(255); }
static size_t cells_ncell_bytes(long long R) { return ((size_t)R * 4 + 255) & ~size_t(255); }
long long spatial_front_cells_bytes(long long R) {
  return (long long)(cells_tab_bytes() + cells_ncell_bytes(R) + (size_t)R * kFrontMaxN * 4 + (size_t)R * kFrontMaxN * kFrontC * 2);
}
static FrontCells cells_view(void* buf, long long R) {
  uint8_t* p = reinterpret_cast<uint8_t*>(buf);
  FrontCells c;
  c.tabs = reinterpret_cast<const FrontTables*>(p); p += cells_tab_bytes();
  c.ncell = reinterpret_cast<const int*>(p); p += cells_ncell_bytes(R);
  c.skey = reinterpret_cast<const uint32_t*>(p); p += (size_t)R * kFrontMaxN * 4;
  c.contrib = reinterpret_cast<const __nv_bfloat16*>(p);
  return c;
}

__global__ void front_tables_kernel(const float* __restrict__ weight, const float* __restrict__ bias, FrontTables* __restrict__ t) {
  const int tid = threadIdx.x, nth = blockDim.x;
  for (int i = tid; i < kT1Rows * kFrontC; i += nth) {
    const int row = i / kFrontC, o = i - row * kFrontC;
    const int vis = row / 5, pr = row - vis * 5;
    const float* w = weight + o * kFrontCin;
    t->t1[row * kPitch + o] = bias[o] + w[1 + vis] + w[7 + pr];
  }
  for (int i = tid; i < kT2Rows * kFrontC; i += nth) {
    const int row = i / kFrontC, o = i - row * kFrontC;
    const float* w = weight + o * kFrontCin;
    t->t2[row * kPitch + o] = w[5 + ((row >> 3) & 1)] + w[12 + ((row >> 2) & 1)] + w[14 + ((row >> 1) & 1)] + w[16 + (row & 1)];
  }
  for (int i = tid; i < kFrontC; i += nth) t->wh[i] = weight[i * kFrontCin] * (1.f / 256.f);
  for (int i = tid; i < 6 * kFrontC; i += nth) {
    const int e = i / kFrontC, o = i - e * kFrontC;
    t->weff[e * kPitch + o] = weight[o * kFrontCin + 18 + e];
  }
  for (int i = tid; i < 32 * kFrontC; i += nth) {
    const int c = i / kFrontC, o = i - c * kFrontC;
    t->ws_t[c * kPitch + o] = weight[o * kFrontCin + 24 + c];
  }
}

// Per observation: live entities sorted by (clamped cell, entity index); at every segment head the contribution vector
// W_s . (sum of the segment's embeddings in ascending entity order).  Embedding rows are staged in shared memory first,
// so the per-head loop has no global-memory latency in it.
constexpr int kCellThreads = 512;
__global__ void __launch_bounds__(kCellThreads)
front_cells_kernel(const FrontArgs a, const FrontTables* __restrict__ tabs, int* __restrict__ ncell, uint32_t* __restrict__ skey,
                   __nv_bfloat16* __restrict__ contrib) {
  extern __shared__ __align__(16) uint8_t cells_smem[];
  float* s_emb = reinterpret_cast<float*>(cells_smem);                         // [512][32]
  float* s_ws = s_emb + kFrontMaxN * 32;                                        // [32][33]
  uint32_t* keys = reinterpret_cast<uint32_t*>(s_ws + 32 * kPitch);            // [512]
  const int r = blockIdx.x, tid = threadIdx.x;
  const int H = a.H, W = a.W, N = a.N;
  const long long en = a.entity_num ? a.entity_num[r] : (long long)N;
  for (int i = tid; i < N * 8; i += kCellThreads)
    reinterpret_cast<float4*>(s_emb)[i] = __ldg(reinterpret_cast<const float4*>(a.emb + (long long)r * N * 32) + i);
  for (int i = tid; i < 32 * kPitch; i += kCellThreads) s_ws[i] = tabs->ws_t[i];
  uint32_t key = 0xFFFFFFFFu;
  if (tid < N && (long long)tid < en) {                                        // encoder.py:38: masked entities add nothing
    long long x, y;
    if (a.idx_bytes == 1) {
      x = reinterpret_cast<const uint8_t*>(a.xs)[(long long)r * N + tid];
      y = reinterpret_cast<const uint8_t*>(a.ys)[(long long)r * N + tid];
    } else {
      x = reinterpret_cast<const long long*>(a.xs)[(long long)r * N + tid];
      y = reinterpret_cast<const long long*>(a.ys)[(long long)r * N + tid];
    }
    x = x < 0 ? 0 : (x > W - 1 ? W - 1 : x);                                   // module_utils.py:18-19
    y = y < 0 ? 0 : (y > H - 1 ? H - 1 : y);
    key = ((uint32_t)(y * W + x) << 9) | (uint32_t)tid;
  }
  keys[tid] = key;
  __syncthreads();
  for (int k = 2; k <= kFrontMaxN; k <<= 1) {                                   // bitonic sort, ascending; dead keys sink to the end
    for (int j = k >> 1; j > 0; j >>= 1) {
      const int ixj = tid ^ j;
      if (ixj > tid) {
        const uint32_t ka = keys[tid], kb = keys[ixj];
        const bool up = (tid & k) == 0;
        if ((ka > kb) == up) { keys[tid] = kb; keys[ixj] = ka; }
      }
      __syncthreads();
    }
  }
  const int nlive = (int)(en < 0 ? 0 : (en > N ? N : en));
  if (tid == 0) ncell[r] = nlive;
  skey[(long long)r * kFrontMaxN + tid] = keys[tid];
  const int lane = tid & 31, warp = tid >> 5;
  for (int i = warp; i < nlive; i += kCellThreads / 32) {
    const uint32_t cell = keys[i] >> 9;
    if (i > 0 && (keys[i - 1] >> 9) == cell) continue;                         // not a head (warp-uniform)
    float seg = 0.f;
    for (int j = i; j < nlive && (keys[j] >> 9) == cell; ++j) seg += s_emb[(keys[j] & 511u) * 32 + lane];
    float out = 0.f;
#pragma unroll
    for (int c = 0; c < 32; ++c) out = fmaf(s_ws[c * kPitch + lane], __shfl_sync(0xffffffffu, seg, c), out);
    contrib[((long long)r * kFrontMaxN + i) * kFrontC + lane] = __float2bfloat16_rn(out);
  }
}

// ------------------------------------------------------------------------------------------------ dense kernels: set-up
struct alignas(16) FrontSmem {
  float tab[kTabFloats];               // t1 | t2 | wh | weff (layout of FrontTables)
  uint32_t pixinfo[kBandRows * 160];   // row1 | row2 << 5 | height << 9
  uint16_t pixmap[kBandRows * 160];    // bits 0-9: 1 + sorted index of the cell's segment head (0 = none); bits 10-15: effects
};
constexpr int kOffT2 = kT1Rows * kPitch, kOffWh = kOffT2 + kT2Rows * kPitch, kOffEff = kOffWh + kFrontC;

__device__ __forceinline__ uint32_t front_info(int h, int vis, int crp, int pr, int alr, int pth, int bld) {
  vis = vis > 3 ? 3 : vis; pr = pr > 4 ? 4 : pr;                               // one-hot indices are clamped
  crp = crp > 1 ? 1 : crp; alr = alr > 1 ? 1 : alr; pth = pth > 1 ? 1 : pth; bld = bld > 1 ? 1 : bld;
  return (uint32_t)(vis * 5 + pr) | ((uint32_t)((crp << 3) | (alr << 2) | (pth << 1) | bld) << 5) | ((uint32_t)h << 9);
}

// tables, per-pixel table rows, effect bits and segment-head slots of one band.  Ends with a __syncthreads().
__device__ __forceinline__ void front_setup(const FrontArgs& a, const FrontCells& c, FrontSmem& s, int r, int band) {
  const int tid = threadIdx.x, nth = blockDim.x;
  const int H = a.H, W = a.W, HW = H * W;
  const int y_lo = band * kBandRows;
  const int n_pix = kBandRows * W;
  const float* tg = reinterpret_cast<const float*>(c.tabs);
  for (int i = tid; i < kTabFloats; i += nth) s.tab[i] = __ldg(tg + i);
  {
    uint32_t* pm32 = reinterpret_cast<uint32_t*>(s.pixmap);
    for (int i = tid; i < n_pix / 2; i += nth) pm32[i] = 0u;
  }
  // plane bytes -> table rows (two pixels per thread-iteration, coalesced 2-byte loads)
  const long long g0 = ((long long)r * H + y_lo) * W;
  for (int i = tid; i < n_pix / 2; i += nth) {
    uchar2 pv[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) pv[k] = __ldg(reinterpret_cast<const uchar2*>(a.plane[k] + g0) + i);
    s.pixinfo[2 * i] = front_info(pv[0].x, pv[1].x, pv[2].x, pv[3].x, pv[4].x, pv[5].x, pv[6].x);
    s.pixinfo[2 * i + 1] = front_info(pv[0].y, pv[1].y, pv[2].y, pv[3].y, pv[4].y, pv[5].y, pv[6].y);
  }
  __syncthreads();
  for (int e = 0; e < 6; ++e) {            // effect bits (zero padding of the lists sets pixel 0: the reference's quirk, kept)
    const int16_t* ids = a.effect[e] + (long long)r * a.n_eff;
    for (int i = tid; i < a.n_eff; i += nth) {
      const int id = (int)ids[i];
      if (id < 0 || id >= HW) continue;
      const int yl = id / W - y_lo;
      if (yl >= 0 && yl < kBandRows) {
        const int local = yl * W + id % W;
        atomicOr(reinterpret_cast<uint32_t*>(s.pixmap) + (local >> 1), (0x400u << e) << ((local & 1) * 16));
      }
    }
  }
  const int nlive = c.ncell[r];
  const uint32_t* sk = c.skey + (long long)r * kFrontMaxN;
  for (int j = tid; j < nlive; j += nth) {
    const uint32_t cell = __ldg(sk + j) >> 9;
    if (j > 0 && (__ldg(sk + j - 1) >> 9) == cell) continue;
    const int yl = (int)(cell / (uint32_t)W) - y_lo;
    if (yl >= 0 && yl < kBandRows) {
      const int local = yl * W + (int)(cell % (uint32_t)W);
      atomicOr(reinterpret_cast<uint32_t*>(s.pixmap) + (local >> 1), (uint32_t)(j + 1) << ((local & 1) * 16));
    }
  }
  __syncthreads();
}

// pre-activation of channel o at a pixel with table info `pi` and map bits `pm`
__device__ __forceinline__ float front_pre(const FrontSmem& s, const __nv_bfloat16* __restrict__ contrib_r, uint32_t pi, uint32_t pm, int o) {
  float v = s.tab[(pi & 31u) * kPitch + o] + s.tab[kOffT2 + ((pi >> 5) & 15u) * kPitch + o];
  v = fmaf((float)(pi >> 9), s.tab[kOffWh + o], v);
  if (pm) {
    const uint32_t slot = pm & 0x3FFu;
    if (slot) v += __bfloat162float(contrib_r[(slot - 1) * kFrontC + o]);
#pragma unroll
    for (int e = 0; e < 6; ++e)
      if (pm & (0x400u << e)) v += s.tab[kOffEff + e * kPitch + o];
  }
  return v;
}

// ------------------------------------------------------------------------------------------------ forward
// thread = (pooled pixel, 8-channel quarter)
constexpr int kFrontThreads = 256;

__global__ void __launch_bounds__(kFrontThreads, 4)
spatial_front_fwd_kernel(const FrontArgs a, const FrontCells c, __nv_bfloat16* __restrict__ out, uint32_t* __restrict__ codes) {
  __shared__ FrontSmem s;
  const int r = blockIdx.x, band = blockIdx.y;
  front_setup(a, c, s, r, band);
  const __nv_bfloat16* contrib_r = c.contrib + (long long)r * kFrontMaxN * kFrontC;
  const int H = a.H, W = a.W, PW = W / 2, pw2 = PW + 2, PH = H / 2;
  const int o0 = (threadIdx.x & 3) * 8;
  const int prow0 = band * (kBandRows / 2);
  const int n_pool = (kBandRows / 2) * PW;
  float wh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) wh[j] = s.tab[kOffWh + o0 + j];
  // one pixel: two 8-float table rows (four 128-bit shared loads), height FMA, the rare extras, running maximum
  auto pixel = [&](uint32_t pi, uint32_t pm, float (&best)[8], uint32_t& code, const uint32_t q) {
    const float4* t1 = reinterpret_cast<const float4*>(s.tab + (pi & 31u) * kPitch + o0);
    const float4* t2 = reinterpret_cast<const float4*>(s.tab + kOffT2 + ((pi >> 5) & 15u) * kPitch + o0);
    const float4 a0 = t1[0], a1 = t1[1], b0 = t2[0], b1 = t2[1];
    const float h = (float)(pi >> 9);
    float v[8] = {a0.x + b0.x, a0.y + b0.y, a0.z + b0.z, a0.w + b0.w, a1.x + b1.x, a1.y + b1.y, a1.z + b1.z, a1.w + b1.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fmaf(h, wh[j], v[j]);
    if (pm) {
      const uint32_t slot = pm & 0x3FFu;
      if (slot) {
        const uint4 cv = *reinterpret_cast<const uint4*>(contrib_r + (slot - 1) * kFrontC + o0);
        const uint32_t cw[4] = {cv.x, cv.y, cv.z, cv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v[2 * j] += __uint_as_float(cw[j] << 16);
          v[2 * j + 1] += __uint_as_float(cw[j] & 0xffff0000u);
        }
      }
#pragma unroll
      for (int e = 0; e < 6; ++e)
        if (pm & (0x400u << e)) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] += s.tab[kOffEff + e * kPitch + o0 + j];
        }
    }
    // strictly greater: the FIRST maximum in scan order wins (torch max-pool), and nothing wins unless it is > 0 (ReLU);
    // nibble j of `code` = winning pixel of channel o0 + j (15 = none): the backward routes by it without recomputing
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (v[j] > best[j]) { best[j] = v[j]; code = (code & ~(15u << (4 * j))) | (q << (4 * j)); }
  };
  for (int pp = threadIdx.x >> 2; pp < n_pool; pp += kFrontThreads / 4) {
    const int py = pp / PW, px = pp - py * PW;
    float best[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) best[j] = 0.f;                                // max(relu(.)) == max(0, pre ...)
    uint32_t code = 0xFFFFFFFFu;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const int l0 = (2 * py + dy) * W + 2 * px;
      const uint2 pi2 = *reinterpret_cast<const uint2*>(s.pixinfo + l0);
      const uint32_t pm2 = *reinterpret_cast<const uint32_t*>(s.pixmap + l0);
      pixel(pi2.x, pm2 & 0xFFFFu, best, code, (uint32_t)(2 * dy));
      pixel(pi2.y, pm2 >> 16, best, code, (uint32_t)(2 * dy + 1));
    }
    if (codes) codes[(((long long)r * PH + prow0 + py) * PW + px) * 4 + (threadIdx.x & 3)] = code;
    uint32_t pk[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      __nv_bfloat162 hv = __floats2bfloat162_rn(best[2 * j], best[2 * j + 1]);
      pk[j] = *reinterpret_cast<uint32_t*>(&hv);
    }
    *reinterpret_cast<uint4*>(out + ((((long long)r * (PH + 2)) + prow0 + py + 1) * pw2 + px + 1) * kFrontC + o0) =
        make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
  // zero border of the padded map: left / right pixel of the band's rows, top / bottom row from the first / last band
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < (kBandRows / 2) * 2 * 4; i += kFrontThreads) {
    const int q = i & 3, side = (i >> 2) & 1, py = i >> 3;
    uint4* dst = reinterpret_cast<uint4*>(out + ((((long long)r * (PH + 2)) + prow0 + py + 1) * pw2 + (side ? PW + 1 : 0)) * kFrontC);
    dst[q] = z;
  }
  if (band == 0) {
    uint4* dst = reinterpret_cast<uint4*>(out + ((long long)r * (PH + 2)) * pw2 * kFrontC);
    for (int i = threadIdx.x; i < pw2 * 4; i += kFrontThreads) dst[i] = z;
  }
  if (band == (int)gridDim.y - 1) {
    uint4* dst = reinterpret_cast<uint4*>(out + (((long long)r * (PH + 2)) + PH + 1) * pw2 * kFrontC);
    for (int i = threadIdx.x; i < pw2 * 4; i += kFrontThreads) dst[i] = z;
  }
}

// ------------------------------------------------------------------------------------------------ backward
// warp = a stream of pooled pixels, lane = output channel.  Private accumulator rows (per thread, shared memory
// [row][thread]: bank = lane, conflict-free, no atomics): 0-19 d T1, 20-35 d T2, 36-41 d effect weights, 42 sum g * h.
constexpr int kFrontBwdThreads = 256;
constexpr int kPrivRows = 43;

__global__ void __launch_bounds__(kFrontBwdThreads, 2)
spatial_front_bwd_kernel(const FrontArgs a, const FrontCells c, const __nv_bfloat16* __restrict__ dpool,
                         const uint32_t* __restrict__ codes, float* __restrict__ partial, float* __restrict__ g_entity) {
  extern __shared__ __align__(16) uint8_t front_smem_raw[];
  FrontSmem& s = *reinterpret_cast<FrontSmem*>(front_smem_raw);
  float* priv = reinterpret_cast<float*>(front_smem_raw + ((sizeof(FrontSmem) + 15) & ~size_t(15)));   // [kPrivRows][threads]
  const int r = blockIdx.x, band = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = 0; i < kPrivRows; ++i) priv[i * kFrontBwdThreads + tid] = 0.f;
  front_setup(a, c, s, r, band);
  const uint32_t* sk = c.skey + (long long)r * kFrontMaxN;
  const int nlive = c.ncell[r];
  const int H = a.H, W = a.W, PW = W / 2, pw2 = PW + 2, PH = H / 2;
  const int prow0 = band * (kBandRows / 2);
  const int n_pool = (kBandRows / 2) * PW;
  if (band == 0) {                                                           // masked entities: zero gradient rows
    for (long long i = (long long)nlive * 32 + tid; i < (long long)a.N * 32; i += kFrontBwdThreads)
      g_entity[(long long)r * a.N * 32 + i] = 0.f;
  }
  constexpr int kWarps = kFrontBwdThreads / 32;
  // one pooled pixel: window (pi[4], pm[4]) in scan order, g = d pooled of this lane's channel, code = the forward's
  // winning pixel of this channel (15: the ReLU killed the window)
  auto window = [&](const uint32_t (&pi)[4], const uint32_t (&pm)[4], float g, uint32_t code) {
    const int arg = code > 3u ? -1 : (int)code;
    if (arg < 0) g = 0.f;
    const int aq = arg < 0 ? 0 : arg;
    const uint32_t pia = aq == 0 ? pi[0] : (aq == 1 ? pi[1] : (aq == 2 ? pi[2] : pi[3]));
    const uint32_t pma = aq == 0 ? pm[0] : (aq == 1 ? pm[1] : (aq == 2 ? pm[2] : pm[3]));
    priv[(pia & 31u) * kFrontBwdThreads + tid] += g;
    priv[(kT1Rows + ((pia >> 5) & 15u)) * kFrontBwdThreads + tid] += g;
    priv[42 * kFrontBwdThreads + tid] += g * (float)(pia >> 9);
    if (pma & 0xFC00u) {
#pragma unroll
      for (int e = 0; e < 6; ++e)
        if (pma & (0x400u << e)) priv[(36 + e) * kFrontBwdThreads + tid] += g;
    }
    // entities on the window's cells: their projection receives this channel's gradient only if their pixel won the max
    if ((pm[0] | pm[1] | pm[2] | pm[3]) & 0x3FFu) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t slot = pm[q] & 0x3FFu;                                // warp-uniform
        if (slot) {
          const uint32_t cell = __ldg(sk + slot - 1) >> 9;
          const float gq = (arg == q) ? g : 0.f;
          for (int j = (int)slot - 1; j < nlive; ++j) {
            const uint32_t kj = __ldg(sk + j);
            if ((kj >> 9) != cell) break;
            g_entity[((long long)r * a.N + (kj & 511u)) * 32 + lane] = gq;
          }
        }
      }
    }
  };
  // warp w takes the pooled-pixel pairs (px, px + 1), px = 2 w, 2 w + 16, ... of every pooled row: the two windows are
  // independent chains (instruction-level parallelism), their table info is one 16-byte shared load per input row
  for (int py = 0; py < kBandRows / 2; ++py) {
    const __nv_bfloat16* grow = dpool + ((((long long)r * (PH + 2)) + prow0 + py + 1) * pw2 + 1) * kFrontC + lane;
    const uint32_t* crow = codes + (((long long)r * PH + prow0 + py) * PW) * 4 + (lane >> 3);
    const int nsh = (lane & 7) * 4;
    // all of this warp's d-pooled values and codes of the row are requested before the first window is processed
    constexpr int kMaxIt = 5;                                                  // PW <= 80 -> at most 5 pairs per warp and row
    float gv[kMaxIt][2];
    uint32_t cv[kMaxIt][2];
#pragma unroll
    for (int it = 0; it < kMaxIt; ++it) {
      const int px = 2 * warp + it * 2 * kWarps;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const bool ok = px + u < PW;
        gv[it][u] = ok ? __bfloat162float(grow[(long long)(px + u) * kFrontC]) : 0.f;
        cv[it][u] = ok ? __ldg(crow + (px + u) * 4) : 0xFFFFFFFFu;
      }
    }
#pragma unroll
    for (int it = 0; it < kMaxIt; ++it) {
      const int px = 2 * warp + it * 2 * kWarps;
      if (px >= PW) break;
      const float g0 = gv[it][0], g1 = gv[it][1];
      const uint32_t c0 = (cv[it][0] >> nsh) & 15u, c1 = (cv[it][1] >> nsh) & 15u;
      const int l0 = (2 * py) * W + 2 * px;
      const uint4 ia = *reinterpret_cast<const uint4*>(s.pixinfo + l0);      // same address in every lane: broadcast
      const uint4 ib = *reinterpret_cast<const uint4*>(s.pixinfo + l0 + W);
      const uint2 ma = *reinterpret_cast<const uint2*>(s.pixmap + l0);
      const uint2 mb = *reinterpret_cast<const uint2*>(s.pixmap + l0 + W);
      const uint32_t pi0[4] = {ia.x, ia.y, ib.x, ib.y}, pi1[4] = {ia.z, ia.w, ib.z, ib.w};
      const uint32_t pm0[4] = {ma.x & 0xFFFFu, ma.x >> 16, mb.x & 0xFFFFu, mb.x >> 16};
      const uint32_t pm1[4] = {ma.y & 0xFFFFu, ma.y >> 16, mb.y & 0xFFFFu, mb.y >> 16};
      window(pi0, pm0, g0, c0);
      if (px + 1 < PW) window(pi1, pm1, g1, c1);
    }
  }
  __syncthreads();
  // per-CTA partial sums, fixed order over the 8 warps
  float* dst = partial + ((long long)blockIdx.x * gridDim.y + blockIdx.y) * (kPrivRows * 32);
  for (int i = tid; i < kPrivRows * 32; i += kFrontBwdThreads) {
    const int row = i >> 5, l = i & 31;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) t += priv[row * kFrontBwdThreads + w * 32 + l];
    dst[i] = t;
  }
}

// sum of the per-CTA partials [n_cta][43][32] over n_cta in fixed order: block = (row), 8 x 32 threads
__global__ void __launch_bounds__(256)
spatial_front_reduce_kernel(const float* __restrict__ partial, float* __restrict__ tables, int n_cta) {
  const int row = blockIdx.x, lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
  float acc = 0.f;
  for (int i = grp; i < n_cta; i += 8) acc += partial[((long long)i * kPrivRows + row) * 32 + lane];
  __shared__ float sm[8][32];
  sm[grp][lane] = acc;
  __syncthreads();
  if (grp == 0) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sm[k][lane];
    tables[row * 32 + lane] = t;
  }
}

// tables [43][32] -> d weight[:, 0:24] and d bias (the scatter block 24:56 comes from g_entity^T emb)
__global__ void spatial_front_finish_kernel(const float* __restrict__ t, float* __restrict__ dw, float* __restrict__ db) {
  const int o = threadIdx.x;
  if (o >= kFrontC) return;
  auto T1 = [&](int row) { return t[row * 32 + o]; };
  auto T2 = [&](int row) { return t[(kT1Rows + row) * 32 + o]; };
  float* w = dw + o * kFrontCin;
  w[0] = t[42 * 32 + o] * (1.f / 256.f);
  float all = 0.f;
  for (int vis = 0; vis < 4; ++vis) {
    float sv = 0.f;
    for (int pr = 0; pr < 5; ++pr) sv += T1(vis * 5 + pr);
    w[1 + vis] = sv;
    all += sv;
  }
  for (int pr = 0; pr < 5; ++pr) {
    float sp = 0.f;
    for (int vis = 0; vis < 4; ++vis) sp += T1(vis * 5 + pr);
    w[7 + pr] = sp;
  }
  for (int bit = 0; bit < 4; ++bit) {                                        // bit 3 creep, 2 alerts, 1 pathable, 0 buildable
    float s0 = 0.f, s1 = 0.f;
    for (int row = 0; row < kT2Rows; ++row) {
      if ((row >> bit) & 1) s1 += T2(row); else s0 += T2(row);
    }
    const int base = bit == 3 ? 5 : (bit == 2 ? 12 : (bit == 1 ? 14 : 16));
    w[base] = s0;
    w[base + 1] = s1;
  }
  for (int e = 0; e < 6; ++e) w[18 + e] = t[(36 + e) * 32 + o];
  db[o] = all;
}

// ------------------------------------------------------------------------------------------------ host
static int front_check(const FrontArgs& a, long long R) {
  if (R <= 0) return AS_ESIZE;
  if (a.W <= 0 || a.W > 160 || (a.W & 3) || a.H <= 0 || (a.H % kBandRows) != 0) return AS_EUNSUPPORTED;
  if (a.N <= 0 || a.N > kFrontMaxN || (a.idx_bytes != 1 && a.idx_bytes != 8)) return AS_EUNSUPPORTED;
  if ((long long)a.H * a.W >= (1 << 23)) return AS_ESIZE;
  if (!a.emb || !a.xs || !a.ys || !a.weight || !a.bias) return AS_EINVAL;
  if (reinterpret_cast<uintptr_t>(a.emb) & 15) return AS_EALIGN;
  for (int i = 0; i < 7; ++i) if (!a.plane[i]) return AS_EINVAL;
  for (int i = 0; i < 6; ++i) if (!a.effect[i]) return AS_EINVAL;
  return AS_OK;
}

static FrontArgs front_args(const void* const* planes, const void* const* effects, int n_eff, const float* emb, const void* xs,
                            const void* ys, int idx_bytes, const long long* en, int N, const float* weight, const float* bias,
                            int H, int W) {
  FrontArgs a{};
  for (int i = 0; i < 7; ++i) a.plane[i] = reinterpret_cast<const uint8_t*>(planes[i]);
  for (int i = 0; i < 6; ++i) a.effect[i] = reinterpret_cast<const int16_t*>(effects[i]);
  a.n_eff = n_eff; a.emb = emb; a.xs = xs; a.ys = ys; a.idx_bytes = idx_bytes; a.entity_num = en; a.N = N;
  a.weight = weight; a.bias = bias; a.H = H; a.W = W;
  return a;
}

int spatial_front_fwd(const void* const* planes, const void* const* effects, int n_eff, const float* emb, const void* xs,
                      const void* ys, int idx_bytes, const long long* en, int N, const float* weight, const float* bias,
                      void* cells, void* out, void* codes, long long R, int H, int W, cudaStream_t st) {
  if (!planes || !effects || !out || !cells) return AS_EINVAL;
  if (reinterpret_cast<uintptr_t>(cells) & 255) return AS_EALIGN;
  const FrontArgs a = front_args(planes, effects, n_eff, emb, xs, ys, idx_bytes, en, N, weight, bias, H, W);
  int rc = front_check(a, R);
  if (rc) return rc;
  const FrontCells c = cells_view(cells, R);
  front_tables_kernel<<<1, 256, 0, st>>>(weight, bias, const_cast<FrontTables*>(c.tabs));
  AS_LAUNCH_CHECK();
  static bool attr = false;
  const int csmem = (kFrontMaxN * 32 + 32 * kPitch + kFrontMaxN) * 4;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(front_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, csmem);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  front_cells_kernel<<<(unsigned)R, kCellThreads, csmem, st>>>(a, c.tabs, const_cast<int*>(c.ncell), const_cast<uint32_t*>(c.skey),
                                                               const_cast<__nv_bfloat16*>(c.contrib));
  AS_LAUNCH_CHECK();
  dim3 grid((unsigned)R, (unsigned)(H / kBandRows));
  spatial_front_fwd_kernel<<<grid, kFrontThreads, 0, st>>>(a, c, reinterpret_cast<__nv_bfloat16*>(out), reinterpret_cast<uint32_t*>(codes));
  AS_LAUNCH_CHECK();
  return AS_OK;
}

long long spatial_front_bwd_scratch_floats(long long R, int H) {
  return (R * (H / kBandRows) + 1) * (long long)(kPrivRows * 32);
}

int spatial_front_bwd(const void* const* planes, const void* const* effects, int n_eff, const float* emb, const void* xs,
                      const void* ys, int idx_bytes, const long long* en, int N, const float* weight, const float* bias,
                      const void* cells, const void* codes, const void* dpool, float* dweight, float* dbias, float* g_entity,
                      float* scratch, long long R, int H, int W, cudaStream_t st) {
  if (!planes || !effects || !cells || !codes || !dpool || !dweight || !dbias || !g_entity || !scratch) return AS_EINVAL;
  const FrontArgs a = front_args(planes, effects, n_eff, emb, xs, ys, idx_bytes, en, N, weight, bias, H, W);
  int rc = front_check(a, R);
  if (rc) return rc;
  const FrontCells c = cells_view(const_cast<void*>(cells), R);
  static bool attr = false;
  const int smem = (((int)sizeof(FrontSmem) + 15) & ~15) + kPrivRows * kFrontBwdThreads * 4;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(spatial_front_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  const int bands = H / kBandRows;
  dim3 grid((unsigned)R, (unsigned)bands);
  spatial_front_bwd_kernel<<<grid, kFrontBwdThreads, smem, st>>>(a, c, reinterpret_cast<const __nv_bfloat16*>(dpool),
                                                                 reinterpret_cast<const uint32_t*>(codes), scratch, g_entity);
  AS_LAUNCH_CHECK();
  float* tables = scratch + R * bands * (long long)(kPrivRows * 32);
  spatial_front_reduce_kernel<<<kPrivRows, 256, 0, st>>>(scratch, tables, (int)(R * bands));
  AS_LAUNCH_CHECK();
  spatial_front_finish_kernel<<<1, 32, 0, st>>>(tables, dweight, dbias);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
