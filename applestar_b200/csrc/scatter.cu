// ScatterConnection (reference: distar/agent/default/model/module_utils.py:11-34) as a deterministic,
// atomic-free sort-then-segment scatter.
//
// One CTA per (observation, group of DG channels):
//   1. keys[i] = (clamped cell id << 9) | entity index   (unique, so the sort is trivially stable)
//   2. bitonic sort of <=512 keys in shared memory
//   3. each segment head sums its segment's rows in ascending entity order (== torch CPU scatter_add_ order,
//      bit-exact) or keeps the last row ('cover'); the head also stamps cellmap[cell] = segment id + 1
//   4. the CTA streams its DG full H*W planes to HBM with 16-byte stores: zero where cellmap is empty, the
//      segment sum otherwise.  Zeros are written by the same pass, so no separate memset of the 3.1 MB/obs map.
// HBM traffic per observation: N*D*4 (embeddings, read once per channel group slice) + D*H*W*4 written.
#include "common.cuh"

namespace as {

constexpr int kScatterMaxN = 512;
constexpr int kScatterThreads = 512;

template <typename IdxT, int DG>
__global__ void __launch_bounds__(kScatterThreads)
scatter_fwd_kernel(const float* __restrict__ emb, const IdxT* __restrict__ xs, const IdxT* __restrict__ ys,
                   const long long* __restrict__ entity_num, float* __restrict__ out, int N, int D, int H, int W,
                   int cover) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int HW = H * W;
  uint32_t* keys = reinterpret_cast<uint32_t*>(smem);                       // [512]
  float* segsum = reinterpret_cast<float*>(smem + kScatterMaxN * 4);        // [512][DG]
  uint16_t* cellmap = reinterpret_cast<uint16_t*>(smem + kScatterMaxN * 4 + kScatterMaxN * DG * 4);  // [HW]

  const int r = blockIdx.x;
  const int d0 = blockIdx.y * DG;
  const int tid = threadIdx.x;
  const long long en = entity_num ? entity_num[r] : (long long)N;

  // 1. clear the cell map, build keys
  {
    uint4* cm4 = reinterpret_cast<uint4*>(cellmap);
    const int n16 = (HW * 2 + 15) / 16;
    for (int i = tid; i < n16; i += kScatterThreads) cm4[i] = make_uint4(0, 0, 0, 0);
  }
  if (tid < N) {
    long long x = (long long)xs[(long long)r * N + tid];
    long long y = (long long)ys[(long long)r * N + tid];
    x = x < 0 ? 0 : (x > W - 1 ? W - 1 : x);                                // module_utils.py:18-19
    y = y < 0 ? 0 : (y > H - 1 ? H - 1 : y);
    keys[tid] = ((uint32_t)(y * W + x) << 9) | (uint32_t)tid;
  } else {
    keys[tid] = 0xFFFFFFFFu;
  }
  __syncthreads();
  // 2. bitonic sort (ascending)
  for (int k = 2; k <= kScatterMaxN; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      const int ixj = tid ^ j;
      if (ixj > tid) {
        const uint32_t a = keys[tid], b = keys[ixj];
        const bool up = (tid & k) == 0;
        if ((a > b) == up) { keys[tid] = b; keys[ixj] = a; }
      }
      __syncthreads();
    }
  }
  // 3. segment heads reduce in entity order
  if (tid < N) {
    const uint32_t key = keys[tid];
    const uint32_t cell = key >> 9;
    const bool head = (tid == 0) || ((keys[tid - 1] >> 9) != cell);
    if (head) {
      float acc[DG];
#pragma unroll
      for (int c = 0; c < DG; ++c) acc[c] = 0.f;
      for (int j = tid; j < N; ++j) {
        const uint32_t kj = keys[j];
        if ((kj >> 9) != cell) break;
        const int e = (int)(kj & 511u);
        const float* src = emb + ((long long)r * N + e) * D + d0;
        const bool live = (long long)e < en;                                 // encoder.py:38 mask
#pragma unroll
        for (int c = 0; c < DG; ++c) {
          const float v = live ? __ldg(src + c) : 0.f;
          acc[c] = cover ? v : (acc[c] + v);
        }
      }
#pragma unroll
      for (int c = 0; c < DG; ++c) segsum[tid * DG + c] = acc[c];
      cellmap[cell] = (uint16_t)(tid + 1);
    }
  }
  __syncthreads();
  // 4. stream the planes, ONE CHANNEL AT A TIME: the CTA writes a plane (H*W*4 bytes, contiguous) front to back with
  //    16-byte stores before it starts the next one, so that the few hundred CTAs in flight are a few hundred linear
  //    write streams (DRAM page locality) instead of DG interleaved streams each.  The cell map (shared memory) is
  //    re-read per channel; zeros are written by the same pass.
  const int q4 = HW / 4;                                                     // HW % 4 == 0 checked on host
  float* obase = out + ((long long)r * D + d0) * HW;
#pragma unroll 1
  for (int c = 0; c < DG; ++c) {
    float4* plane = reinterpret_cast<float4*>(obase + (long long)c * HW);
#pragma unroll 4
    for (int p4 = tid; p4 < q4; p4 += kScatterThreads) {
      const uint2 m = *reinterpret_cast<const uint2*>(cellmap + p4 * 4);
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if ((m.x | m.y) != 0u) {
        const uint32_t m0 = m.x & 0xFFFFu, m1 = m.x >> 16, m2 = m.y & 0xFFFFu, m3 = m.y >> 16;
        if (m0) v.x = segsum[(m0 - 1) * DG + c];
        if (m1) v.y = segsum[(m1 - 1) * DG + c];
        if (m2) v.z = segsum[(m2 - 1) * DG + c];
        if (m3) v.w = segsum[(m3 - 1) * DG + c];
      }
      plane[p4] = v;
    }
  }
}

// backward of 'add': d_emb[r,e,:] = live ? d_map[r,:,cell(e)] : 0     (pure gather)
template <typename IdxT>
__global__ void scatter_bwd_kernel(const float* __restrict__ dmap, const IdxT* __restrict__ xs,
                                   const IdxT* __restrict__ ys, const long long* __restrict__ entity_num,
                                   float* __restrict__ demb, long long R, int N, int D, int H, int W) {
  const long long total = R * N * (long long)D;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int d = (int)(i % D);
    const long long re = i / D;
    const int e = (int)(re % N);
    const long long r = re / N;
    const long long en = entity_num ? entity_num[r] : (long long)N;
    float g = 0.f;
    if ((long long)e < en) {
      long long x = (long long)xs[re], y = (long long)ys[re];
      x = x < 0 ? 0 : (x > W - 1 ? W - 1 : x);
      y = y < 0 ? 0 : (y > H - 1 ? H - 1 : y);
      g = __ldg(dmap + ((r * D + d) * H + y) * W + x);
    }
    demb[i] = g;
  }
}

template <typename IdxT>
static int scatter_fwd_launch(const float* emb, const void* xs, const void* ys, const long long* en, float* out,
                              long long R, int N, int D, int H, int W, int cover, cudaStream_t st) {
  constexpr int DG = 8;
  const int smem = kScatterMaxN * 4 + kScatterMaxN * DG * 4 + ((H * W * 2 + 15) / 16) * 16;
  auto kern = scatter_fwd_kernel<IdxT, DG>;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  if (smem > 160 * 1024) return AS_ESIZE;
  dim3 grid((unsigned)R, (unsigned)(D / DG));
  kern<<<grid, kScatterThreads, smem, st>>>(emb, (const IdxT*)xs, (const IdxT*)ys, en, out, N, D, H, W, cover);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int scatter_fwd(const float* emb, const void* xs, const void* ys, int idx_bytes, const long long* en, float* out,
                long long R, int N, int D, int H, int W, int cover, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (N <= 0 || N > kScatterMaxN || D <= 0 || (D % 8) != 0 || H <= 0 || W <= 0) return AS_ESIZE;
  if (((long long)H * W) % 4 != 0 || (long long)H * W > 65535 * 4 || R > 2147483647LL) return AS_ESIZE;
  if ((long long)H * W >= (1 << 23)) return AS_ESIZE;
  if (!emb || !xs || !ys || !out) return AS_EINVAL;
  if (reinterpret_cast<uintptr_t>(out) & 15) return AS_EALIGN;
  if (idx_bytes == 1) return scatter_fwd_launch<uint8_t>(emb, xs, ys, en, out, R, N, D, H, W, cover, st);
  if (idx_bytes == 8) return scatter_fwd_launch<long long>(emb, xs, ys, en, out, R, N, D, H, W, cover, st);
  return AS_EUNSUPPORTED;
}

int scatter_bwd(const float* dmap, const void* xs, const void* ys, int idx_bytes, const long long* en, float* demb,
                long long R, int N, int D, int H, int W, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!dmap || !xs || !ys || !demb) return AS_EINVAL;
  const long long total = R * N * (long long)D;
  long long blocks = (total + 255) / 256;
  const int grid = (int)(blocks < 148LL * 16 ? blocks : 148LL * 16);
  if (idx_bytes == 1)
    scatter_bwd_kernel<uint8_t><<<grid, 256, 0, st>>>(dmap, (const uint8_t*)xs, (const uint8_t*)ys, en, demb, R, N, D, H, W);
  else if (idx_bytes == 8)
    scatter_bwd_kernel<long long><<<grid, 256, 0, st>>>(dmap, (const long long*)xs, (const long long*)ys, en, demb, R, N, D, H, W);
  else
    return AS_EUNSUPPORTED;
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
