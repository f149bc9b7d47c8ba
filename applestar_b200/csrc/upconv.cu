// Tail of the location head as one kernel per direction (reference: head/action_arg_head.py:440-446: the third
// F.interpolate(scale_factor=2, mode='bilinear') followed by the 3x3 convolution 32 -> 1, flatten, / temperature).
//
// Both steps are linear and the convolution has ONE output channel, so they commute into something much smaller:
//   logit[p] = b + sum_t sum_c w[t][c] U[p + t][c],   U = bilinear x2 of X (zero outside the image: the conv's padding)
//            = b + sum_t (bilinear x2 of Y_t)(p + t),  Y_t[n] = sum_c w[t][c] X[n][c]      (9 numbers per LOW-res pixel)
// The 152x160x32 upsampled map (1.6 GB per learner step, written once and read twice, plus its gradient) is never
// formed: the forward reads the 76x80x32 map and writes the logits, the backward reads the logit gradients and writes
// the gradient of the 76x80x32 map (already masked by X > 0 when X is a ReLU output), the weight and bias gradients.
// All arithmetic fp32 from the bf16 map (the unfused path rounded U to bf16).
#include "common.cuh"

namespace as {

constexpr int kUcC = 32;
constexpr int kUcThreads = 320;                // forward: two row groups x 160 output columns
constexpr int kUcBandHi = 8;                 // output rows per forward CTA
constexpr int kUcBandLo = 8;                 // low-res rows per backward CTA
constexpr int kUcMaxWlo = 80;                // low-res width limit (shared-memory tiles)

// F.interpolate(scale 2, bilinear, align_corners=False): source position of output index o, neighbours clamped
__device__ __forceinline__ void uc_src(int o, int n_in, int* i0, int* i1, float* w1) {
  float s = (o + 0.5f) * 0.5f - 0.5f;
  if (s < 0.f) s = 0.f;
  const int a = (int)s;
  *i0 = a;
  *i1 = a + 1 < n_in ? a + 1 : n_in - 1;
  *w1 = s - (float)a;
}
// transposed: low index y receives from outputs 2y-1+k (k = 0..3) with these weights
__device__ __forceinline__ void uc_tw(int y, int n, float* w) {
  w[0] = y > 0 ? 0.25f : 0.f; w[1] = y > 0 ? 0.75f : 1.f;
  w[2] = y < n - 1 ? 0.75f : 1.f; w[3] = y < n - 1 ? 0.25f : 0.f;
}

// ------------------------------------------------------------------------------------------------ forward
// CTA = (image, band of kUcBandHi output rows).  Phase 1: Y_t for the low-res rows the band touches; phase 2: one
// thread per output pixel gathers 9 taps x 4 bilinear neighbours from shared memory.
__global__ void __launch_bounds__(kUcThreads)
upconv_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w9, const float* __restrict__ bias,
                  float* __restrict__ out, int h, int w, float inv_temp) {
  constexpr int kRows = kUcBandHi / 2 + 3;                                   // low rows needed by a band (incl. halo)
  __shared__ float sY[kRows][kUcMaxWlo][9];
  __shared__ __align__(16) float sW[9 * kUcC];
  const int r = blockIdx.x, band = blockIdx.y;
  const int H = 2 * h, W = 2 * w;
  const int y0 = band * kUcBandHi;
  const int lo0 = max(0, (y0 - 1) / 2 - 1);                                  // first low row that can be a neighbour
  for (int i = threadIdx.x; i < 9 * kUcC; i += kUcThreads) sW[i] = w9[i];
  __syncthreads();
  // two low-res pixels per thread-iteration: every 16-byte weight load serves eight multiply-adds
  const int n_lo = kRows * w;
  for (int i = threadIdx.x; i < (n_lo + 1) / 2; i += kUcThreads) {
    float acc[2][9];
    uint4 xv[2][4];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int ii = 2 * i + u;
      const int lr = ii / w, lx = ii - lr * w;
      const int ly = lo0 + lr;
#pragma unroll
      for (int t = 0; t < 9; ++t) acc[u][t] = 0.f;
#pragma unroll
      for (int q = 0; q < 4; ++q) xv[u][q] = make_uint4(0, 0, 0, 0);
      if (ii < n_lo && ly < h) {
        const uint4* px = reinterpret_cast<const uint4*>(x + (((long long)r * (h + 2) + ly + 1) * (w + 2) + lx + 1) * kUcC);
#pragma unroll
        for (int q = 0; q < 4; ++q) xv[u][q] = __ldg(px + q);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
#pragma unroll
      for (int hq = 0; hq < 2; ++hq) {                                         // channels 8 q + 4 hq .. + 3
        float xa[2][4];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const uint32_t w0 = hq ? xv[u][q].z : xv[u][q].x, w1 = hq ? xv[u][q].w : xv[u][q].y;
          xa[u][0] = __uint_as_float(w0 << 16); xa[u][1] = __uint_as_float(w0 & 0xffff0000u);
          xa[u][2] = __uint_as_float(w1 << 16); xa[u][3] = __uint_as_float(w1 & 0xffff0000u);
        }
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          const float4 wv = *reinterpret_cast<const float4*>(&sW[t * kUcC + q * 8 + hq * 4]);
#pragma unroll
          for (int u = 0; u < 2; ++u)
            acc[u][t] = fmaf(xa[u][0], wv.x, fmaf(xa[u][1], wv.y, fmaf(xa[u][2], wv.z, fmaf(xa[u][3], wv.w, acc[u][t]))));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int ii = 2 * i + u;
      if (ii < n_lo) {
        const int lr = ii / w, lx = ii - lr * w;
#pragma unroll
        for (int t = 0; t < 9; ++t) sY[lr][lx][t] = acc[u][t];
      }
    }
  }
  __syncthreads();
  const float b0 = bias ? bias[0] : 0.f;
  // thread = output column (and a group of rows): the three column taps' neighbours / weights are computed once
  const int ncol_groups = kUcThreads / W > 0 ? kUcThreads / W : 1;             // threads beyond ncol_groups * W idle here
  const int xx = threadIdx.x % W, rg = threadIdx.x / W;
  if (rg < ncol_groups) {
    int j0[3], j1[3]; float wx[3]; bool okx[3];
#pragma unroll
    for (int tx = 0; tx < 3; ++tx) {
      const int qx = xx + tx - 1;
      okx[tx] = qx >= 0 && qx < W;
      uc_src(okx[tx] ? qx : 0, w, &j0[tx], &j1[tx], &wx[tx]);
    }
    for (int yy = rg; yy < kUcBandHi; yy += ncol_groups) {
      const int y = y0 + yy;
      if (y >= H) break;
      float acc = b0;
#pragma unroll
      for (int ty = 0; ty < 3; ++ty) {
        const int qy = y + ty - 1;
        if (qy < 0 || qy >= H) continue;
        int i0, i1; float wy;
        uc_src(qy, h, &i0, &i1, &wy);
        const float (*r0)[9] = sY[i0 - lo0];
        const float (*r1)[9] = sY[i1 - lo0];
#pragma unroll
        for (int tx = 0; tx < 3; ++tx) {
          if (!okx[tx]) continue;
          const int t = ty * 3 + tx;
          const float top = (1.f - wx[tx]) * r0[j0[tx]][t] + wx[tx] * r0[j1[tx]][t];
          const float bot = (1.f - wx[tx]) * r1[j0[tx]][t] + wx[tx] * r1[j1[tx]][t];
          acc += (1.f - wy) * top + wy * bot;
        }
      }
      out[(long long)r * H * W + (long long)y * W + xx] = acc * inv_temp;
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward
// CTA = (image, band of kUcBandLo low-res rows).  A: logit gradients of the 2 * band + 5 output rows that reach the band (times
// 1/temperature, zero outside the image) -> shared memory; B: dY_t[n] = sum over the 4x4 outputs q around n of
// bilinear weight x dlogit[q - t]; C: thread = (low pixel, 8 channels): dX = sum_t w[t][c] dY_t (masked by X > 0 when
// relu_input), dW partials in registers (fixed channel group per thread), reduced over the CTA in fixed order.
constexpr int kUcBwdThreads = 256;
constexpr int kUcDRows = 2 * kUcBandLo + 5;                                  // output rows 2 n0 - 2 .. 2 n0 + 2 * band + 2

__global__ void __launch_bounds__(kUcBwdThreads)
upconv_bwd_kernel(const float* __restrict__ dout, const __nv_bfloat16* __restrict__ x, const float* __restrict__ w9,
                  __nv_bfloat16* __restrict__ dx, float* __restrict__ partial, int h, int w, float inv_temp, int relu_input) {
  __shared__ float sD[kUcDRows][2 * kUcMaxWlo + 4];                          // column index = output column + 2
  __shared__ float sdY[kUcBandLo * kUcMaxWlo][9];
  __shared__ __align__(16) float sW[9 * kUcC];
  __shared__ float sRed[kUcBwdThreads / 32][4][72];
  __shared__ float sDb[kUcBwdThreads / 32];
  const int r = blockIdx.x, band = blockIdx.y;
  const int H = 2 * h, W = 2 * w;
  const int n0 = band * kUcBandLo;
  const int prow0 = 2 * n0 - 2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 9 * kUcC; i += kUcBwdThreads) sW[i] = w9[i];
  float dbl = 0.f;
  for (int i = tid; i < kUcDRows * (W + 4); i += kUcBwdThreads) {
    const int rr = i / (W + 4), cc = i - rr * (W + 4);
    const int py = prow0 + rr, px = cc - 2;
    float v = 0.f;
    if (py >= 0 && py < H && px >= 0 && px < W) {
      v = __ldg(dout + (long long)r * H * W + (long long)py * W + px) * inv_temp;
      if (py >= 2 * n0 && py < 2 * (n0 + kUcBandLo)) dbl += v;              // every output row belongs to exactly one band
    }
    sD[rr][cc] = v;
  }
  dbl = warp_sum(dbl);
  if (lane == 0) sDb[warp] = dbl;
  __syncthreads();
  // B
  for (int i = tid; i < kUcBandLo * w * 9; i += kUcBwdThreads) {
    const int t = i % 9, n = i / 9;
    const int ny = n / w, nx = n - ny * w;
    const int gy = n0 + ny;
    float acc = 0.f;
    if (gy < h) {
      float wy[4], wx[4];
      uc_tw(gy, h, wy);
      uc_tw(nx, w, wx);
      const int ty = t / 3, tx = t - ty * 3;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int qy = 2 * gy - 1 + k;                                       // output row whose neighbour is gy
        if (wy[k] == 0.f || qy < 0 || qy >= H) continue;
        const int prow = qy - (ty - 1) - prow0;                              // p = q - tap offset; zero rows outside the image are staged
        float rs = 0.f;
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          const int qx = 2 * nx - 1 + l;
          if (wx[l] == 0.f || qx < 0 || qx >= W) continue;
          rs = fmaf(wx[l], sD[prow][qx - (tx - 1) + 2], rs);
        }
        acc = fmaf(wy[k], rs, acc);
      }
    }
    sdY[n][t] = acc;
  }
  __syncthreads();
  // C
  const int cg = tid & 3;
  float dw[72];
#pragma unroll
  for (int j = 0; j < 72; ++j) dw[j] = 0.f;
  for (int i = tid; i < kUcBandLo * w * 4; i += kUcBwdThreads) {
    const int n = i >> 2;
    const int ny = n / w, nx = n - ny * w;
    const int gy = n0 + ny;
    if (gy >= h) break;
    const long long off = (((long long)r * (h + 2) + gy + 1) * (w + 2) + nx + 1) * kUcC + cg * 8;
    const uint4 xv = __ldg(reinterpret_cast<const uint4*>(x + off));
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
    float xf[8], g[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) { xf[2 * k] = __uint_as_float(xw[k] << 16); xf[2 * k + 1] = __uint_as_float(xw[k] & 0xffff0000u); }
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = 0.f;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float dy = sdY[n][t];
      const float4 wa = *reinterpret_cast<const float4*>(&sW[t * kUcC + cg * 8]);
      const float4 wb = *reinterpret_cast<const float4*>(&sW[t * kUcC + cg * 8 + 4]);
      const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        g[j] = fmaf(dy, wv[j], g[j]);
        dw[t * 8 + j] = fmaf(dy, xf[j], dw[t * 8 + j]);
      }
    }
    if (relu_input) {
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = xf[j] > 0.f ? g[j] : 0.f;
    }
    uint32_t pk[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      __nv_bfloat162 hv = __floats2bfloat162_rn(g[2 * k], g[2 * k + 1]);
      pk[k] = *reinterpret_cast<uint32_t*>(&hv);
    }
    *reinterpret_cast<uint4*>(dx + off) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
  // zero border of dx: left / right pixel of the band's rows, top / bottom row from the first / last band
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < kUcBandLo * 2 * 4; i += kUcBwdThreads) {
    const int q = i & 3, side = (i >> 2) & 1, ny = i >> 3;
    if (n0 + ny < h)
      reinterpret_cast<uint4*>(dx + (((long long)r * (h + 2) + n0 + ny + 1) * (w + 2) + (side ? w + 1 : 0)) * kUcC)[q] = z;
  }
  if (band == 0)
    for (int i = tid; i < (w + 2) * 4; i += kUcBwdThreads) reinterpret_cast<uint4*>(dx + ((long long)r * (h + 2)) * (w + 2) * kUcC)[i] = z;
  if (band == (int)gridDim.y - 1)
    for (int i = tid; i < (w + 2) * 4; i += kUcBwdThreads)
      reinterpret_cast<uint4*>(dx + (((long long)r * (h + 2)) + h + 1) * (w + 2) * kUcC)[i] = z;
  // dW: lanes with the same channel group (lane & 3) -> butterfly over lane bits 2..4, then the 8 warps in fixed order
#pragma unroll
  for (int j = 0; j < 72; ++j) {
    float v = dw[j];
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    if (lane < 4) sRed[warp][lane][j] = v;
  }
  __syncthreads();
  float* dst = partial + ((long long)blockIdx.x * gridDim.y + blockIdx.y) * 289;
  for (int i = tid; i < 288; i += kUcBwdThreads) {                              // i = t * 32 + c
    const int t = i >> 5, c = i & 31;
    float s = 0.f;
#pragma unroll
    for (int wq = 0; wq < kUcBwdThreads / 32; ++wq) s += sRed[wq][c >> 3][t * 8 + (c & 7)];
    dst[i] = s;
  }
  if (tid == 0) {
    float s = 0.f;
    for (int wq = 0; wq < kUcBwdThreads / 32; ++wq) s += sDb[wq];
    dst[288] = s;
  }
}

// fixed-order sum of the per-CTA partials [n][289] -> dw9 [9][32], db [1], in two levels (kUcChunks partial sums first)
constexpr int kUcChunks = 64;
__global__ void __launch_bounds__(256)
upconv_reduce1_kernel(const float* __restrict__ partial, float* __restrict__ part2, int n) {
  const int col = blockIdx.x * 32 + (threadIdx.x & 31), grp = threadIdx.x >> 5;
  const int per = (n + kUcChunks - 1) / kUcChunks;
  const int i0 = blockIdx.y * per, i1 = min(n, i0 + per);
  float acc = 0.f;
  if (col < 289)
    for (int i = i0 + grp; i < i1; i += 8) acc += partial[(long long)i * 289 + col];
  __shared__ float sm[8][33];
  sm[grp][threadIdx.x & 31] = acc;
  __syncthreads();
  if (grp == 0 && col < 289) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sm[k][threadIdx.x & 31];
    part2[blockIdx.y * 289 + col] = t;
  }
}
__global__ void upconv_reduce2_kernel(const float* __restrict__ part2, float* __restrict__ dw9, float* __restrict__ db) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= 289) return;
  float t = 0.f;
  for (int k = 0; k < kUcChunks; ++k) t += part2[k * 289 + col];
  if (col < 288) dw9[col] = t; else db[0] = t;
}

static int uc_check(long long R, int h, int w) {
  if (R <= 0 || h <= 0 || w <= 0) return AS_ESIZE;
  if (w > kUcMaxWlo || R > 2147483647LL) return AS_EUNSUPPORTED;
  return AS_OK;
}

int upconv_fwd(const void* x, const float* w9, const float* bias, float* out, long long R, int h, int w, float inv_temp,
               cudaStream_t st) {
  int rc = uc_check(R, h, w);
  if (rc) return rc;
  if (!x || !w9 || !out) return AS_EINVAL;
  if (reinterpret_cast<uintptr_t>(x) & 15) return AS_EALIGN;
  dim3 grid((unsigned)R, (unsigned)((2 * h + kUcBandHi - 1) / kUcBandHi));
  upconv_fwd_kernel<<<grid, kUcThreads, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(x), w9, bias, out, h, w, inv_temp);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

long long upconv_scratch_floats(long long R, int h) { return (R * ((h + kUcBandLo - 1) / kUcBandLo) + kUcChunks) * 289LL; }

int upconv_bwd(const float* dout, const void* x, const float* w9, void* dx, float* dw9, float* db, float* scratch, long long R,
               int h, int w, float inv_temp, int relu_input, cudaStream_t st) {
  int rc = uc_check(R, h, w);
  if (rc) return rc;
  if (!dout || !x || !w9 || !dx || !dw9 || !db || !scratch) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dx)) & 15) return AS_EALIGN;
  const int bands = (h + kUcBandLo - 1) / kUcBandLo;
  dim3 grid((unsigned)R, (unsigned)bands);
  upconv_bwd_kernel<<<grid, kUcBwdThreads, 0, st>>>(dout, reinterpret_cast<const __nv_bfloat16*>(x), w9,
                                                 reinterpret_cast<__nv_bfloat16*>(dx), scratch, h, w, inv_temp, relu_input);
  AS_LAUNCH_CHECK();
  float* part2 = scratch + R * bands * 289LL;
  upconv_reduce1_kernel<<<dim3((289 + 31) / 32, kUcChunks), 256, 0, st>>>(scratch, part2, (int)(R * bands));
  AS_LAUNCH_CHECK();
  upconv_reduce2_kernel<<<3, 128, 0, st>>>(part2, dw9, db);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
