// HBM-bound kernels of the spatial path on zero-padded NHWC bf16 feature maps [R, H+2, W+2, C]:
// layout conversion, residual / gated-residual tails, 2x bilinear upsampling, the 132-channel location-head input,
// the final 32->1 convolution, and column sums (bias gradients).  The convolutions themselves are implicit GEMMs on
// the tcgen05 kernel (gemm_tcgen05.cu); every kernel here keeps the one-pixel border at exactly zero so that the
// next convolution sees correct zero padding.
#include "common.cuh"

namespace as {

__device__ __forceinline__ float bf2f(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ __nv_bfloat16 f2bf(float v) { return __float2bfloat16_rn(v); }

static int ew_grid(long long n, int per_thread) {
  long long blocks = (n / per_thread + 255) / 256;
  const long long cap = 148LL * 16;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

// ------------------------------------------------------------------------------------------ layout conversion
// NCHW fp32 [R,C,H,W] -> padded NHWC bf16 [R,H+2,W+2,Cp] (channels >= C and the border are zero).  One block per
// (r, y); a [Cp][W] tile goes through shared memory so both sides are coalesced.
__global__ void nchw_to_pnhwc_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int C, int Cp,
                                     int H, int W, int relu) {
  extern __shared__ float tile[];                 // [C][W+1]
  const int r = blockIdx.x / (H + 2), yy = blockIdx.x % (H + 2);
  __nv_bfloat16* drow = dst + ((long long)r * (H + 2) + yy) * (W + 2) * Cp;
  const int n_out = (W + 2) * Cp;
  if (yy == 0 || yy == H + 1) {
    for (int i = threadIdx.x; i < n_out; i += blockDim.x) drow[i] = f2bf(0.f);
    return;
  }
  const int y = yy - 1;
  for (int i = threadIdx.x; i < C * W; i += blockDim.x) {
    const int c = i / W, x = i - c * W;
    float v = src[(((long long)r * C + c) * H + y) * W + x];
    tile[c * (W + 1) + x] = relu ? fmaxf(v, 0.f) : v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n_out; i += blockDim.x) {
    const int xx = i / Cp, c = i - xx * Cp;
    float v = 0.f;
    if (xx >= 1 && xx <= W && c < C) v = tile[c * (W + 1) + xx - 1];
    drow[i] = f2bf(v);
  }
}

// padded NHWC bf16 -> NCHW fp32 (interior only, first C channels)
__global__ void pnhwc_to_nchw_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int C, int Cp,
                                     int H, int W) {
  extern __shared__ float tile[];                 // [C][W+1]
  const int r = blockIdx.x / H, y = blockIdx.x % H;
  const __nv_bfloat16* srow = src + (((long long)r * (H + 2) + y + 1) * (W + 2) + 1) * Cp;
  for (int i = threadIdx.x; i < W * C; i += blockDim.x) {
    const int x = i / C, c = i - x * C;
    tile[c * (W + 1) + x] = bf2f(srow[(long long)x * Cp + c]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C * W; i += blockDim.x) {
    const int c = i / W, x = i - c * W;
    dst[(((long long)r * C + c) * H + y) * W + x] = tile[c * (W + 1) + x];
  }
}

int nchw_to_pnhwc(const float* src, void* dst, long long R, int C, int Cp, int H, int W, int relu, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (C <= 0 || Cp < C || H <= 0 || W <= 0) return AS_ESIZE;
  const int smem = C * (W + 1) * 4;
  if (smem > 96 * 1024) return AS_ESIZE;
  static bool attr = false;
  if (!attr) { cudaFuncSetAttribute(nchw_to_pnhwc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024); attr = true; }
  nchw_to_pnhwc_kernel<<<(unsigned)(R * (H + 2)), 256, smem, st>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), C, Cp, H, W, relu);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int pnhwc_to_nchw(const void* src, float* dst, long long R, int C, int Cp, int H, int W, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (C <= 0 || Cp < C || H <= 0 || W <= 0) return AS_ESIZE;
  const int smem = C * (W + 1) * 4;
  if (smem > 96 * 1024) return AS_ESIZE;
  static bool attr = false;
  if (!attr) { cudaFuncSetAttribute(pnhwc_to_nchw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024); attr = true; }
  pnhwc_to_nchw_kernel<<<(unsigned)(R * H), 256, smem, st>>>(reinterpret_cast<const __nv_bfloat16*>(src), dst, C, Cp, H, W);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------------------------------ residual tails
union BF8 { uint4 u; __nv_bfloat16 h[8]; };

// mode 0: y = a + b;  mode 1: y = relu(a + b)
__global__ void add_act_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ y,
                               long long n8, int mode) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    BF8 va, vb, vy;
    va.u = __ldg(a + i); vb.u = __ldg(b + i);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = bf2f(va.h[j]) + bf2f(vb.h[j]);
      if (mode == 1) s = fmaxf(s, 0.f);
      vy.h[j] = f2bf(s);
    }
    y[i] = vy.u;
  }
}
// dz = dy * (y > 0)
__global__ void relu_bwd_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y, uint4* __restrict__ dz,
                                long long n8) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    BF8 vd, vy, vo;
    vd.u = __ldg(dy + i); vy.u = __ldg(y + i);
#pragma unroll
    for (int j = 0; j < 8; ++j) vo.h[j] = bf2f(vy.h[j]) > 0.f ? vd.h[j] : f2bf(0.f);
    dz[i] = vo.u;
  }
}

int add_act(const void* a, const void* b, void* y, long long n, int mode, cudaStream_t st) {
  if (n <= 0) return AS_OK;
  if (n % 8) return AS_ESIZE;
  if (!a || !b || !y) return AS_EINVAL;
  add_act_kernel<<<ew_grid(n, 8), 256, 0, st>>>((const uint4*)a, (const uint4*)b, (uint4*)y, n / 8, mode);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int relu_bwd(const void* dy, const void* y, void* dz, long long n, cudaStream_t st) {
  if (n <= 0) return AS_OK;
  if (n % 8) return AS_ESIZE;
  if (!dy || !y || !dz) return AS_EINVAL;
  relu_bwd_kernel<<<ew_grid(n, 8), 256, 0, st>>>((const uint4*)dy, (const uint4*)y, (uint4*)dz, n / 8);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// MUFU versions for the gate math: the results are rounded to bf16 (2^-9) right away, the approximations are ~2^-11
__device__ __forceinline__ float gate_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float gate_sigmoid(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }

// GatedResBlock tail (module_utils.py:227-229): y = relu(tanh(r * sigmoid(g)) * sp + x)
__global__ void gated_fwd_kernel(const uint4* __restrict__ r, const uint4* __restrict__ g, const uint4* __restrict__ x,
                                 const float* __restrict__ sp, uint4* __restrict__ y, long long n8) {
  const float s = *sp;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    BF8 vr, vg, vx, vy;
    vr.u = __ldg(r + i); vg.u = __ldg(g + i); vx.u = __ldg(x + i);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sg = gate_sigmoid(bf2f(vg.h[j]));
      const float t = gate_tanh(bf2f(vr.h[j]) * sg);
      vy.h[j] = f2bf(fmaxf(t * s + bf2f(vx.h[j]), 0.f));
    }
    y[i] = vy.u;
  }
}
// dz = dy (y>0); dx = dz; t = tanh(r s(g)); dsp = sum dz t; du = dz sp (1-t^2); dr = du s(g); dg = du r s(g)(1-s(g))
__global__ void gated_bwd_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y, const uint4* __restrict__ r,
                                 const uint4* __restrict__ g, const float* __restrict__ sp, uint4* __restrict__ dr,
                                 uint4* __restrict__ dg, uint4* __restrict__ dx, float* __restrict__ dsp_partial,
                                 long long n8) {
  const float s = *sp;
  float acc = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    BF8 vd, vy, vr, vg, odr, odg, odx;
    vd.u = __ldg(dy + i); vy.u = __ldg(y + i); vr.u = __ldg(r + i); vg.u = __ldg(g + i);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float dz = bf2f(vy.h[j]) > 0.f ? bf2f(vd.h[j]) : 0.f;
      const float rv = bf2f(vr.h[j]);
      const float sg = gate_sigmoid(bf2f(vg.h[j]));
      const float t = gate_tanh(rv * sg);
      acc += dz * t;
      const float du = dz * s * (1.f - t * t);
      odr.h[j] = f2bf(du * sg);
      odg.h[j] = f2bf(du * rv * sg * (1.f - sg));
      odx.h[j] = f2bf(dz);
    }
    dr[i] = odr.u; dg[i] = odg.u; dx[i] = odx.u;
  }
  __shared__ float red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    dsp_partial[blockIdx.x] = t;
  }
}
__global__ void sum_partials_kernel(const float* __restrict__ partial, int n, float* __restrict__ out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < n; ++i) s += partial[i];
    *out = s;
  }
}

constexpr int kGatedBlocks = 148 * 8;
long long gated_scratch_floats() { return kGatedBlocks; }

int gated_fwd(const void* r, const void* g, const void* x, const float* sp, void* y, long long n, cudaStream_t st) {
  if (n <= 0) return AS_OK;
  if (n % 8) return AS_ESIZE;
  if (!r || !g || !x || !sp || !y) return AS_EINVAL;
  gated_fwd_kernel<<<ew_grid(n, 8), 256, 0, st>>>((const uint4*)r, (const uint4*)g, (const uint4*)x, sp, (uint4*)y, n / 8);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int gated_bwd(const void* dy, const void* y, const void* r, const void* g, const float* sp, void* dr, void* dg, void* dx,
              float* dsp, float* scratch, long long n, cudaStream_t st) {
  if (n <= 0) return AS_OK;
  if (n % 8) return AS_ESIZE;
  if (!dy || !y || !r || !g || !sp || !dr || !dg || !dx || !dsp || !scratch) return AS_EINVAL;
  gated_bwd_kernel<<<kGatedBlocks, 256, 0, st>>>((const uint4*)dy, (const uint4*)y, (const uint4*)r, (const uint4*)g, sp,
                                                 (uint4*)dr, (uint4*)dg, (uint4*)dx, scratch, n / 8);
  AS_LAUNCH_CHECK();
  sum_partials_kernel<<<1, 32, 0, st>>>(scratch, kGatedBlocks, dsp);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------------------------------ bilinear x2 (align_corners=False)
// out[2i] = .25 in[i-1] + .75 in[i], out[2i+1] = .75 in[i] + .25 in[i+1], indices clamped (== F.interpolate).
__device__ __forceinline__ void up2_src(int o, int n_in, int* i0, int* i1, float* w1) {
  float s = (o + 0.5f) * 0.5f - 0.5f;
  if (s < 0.f) s = 0.f;
  const int a = (int)s;
  *i0 = a;
  *i1 = a + 1 < n_in ? a + 1 : n_in - 1;
  *w1 = s - (float)a;
}

// Forward: one thread = one SOURCE pixel (8 channels) -> its 2x2 block of outputs.  With scale 2 and half-pixel
// centres the even / odd outputs are 0.25 * left + 0.75 * centre and 0.75 * centre + 0.25 * right (indices clamped at
// the edges, which reproduces F.interpolate's clamping), so a 3x3 neighbourhood is loaded once for four outputs.
__global__ void __launch_bounds__(256)
up2_fwd_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, long long R, int H, int W, int C,
               const Dec3 dec) {
  const int C8 = C / 8, W2 = 2 * W, H2 = 2 * H;
  const uint32_t total = (uint32_t)(R * (long long)H * W * C8);
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    int c8, xs, ys, ri;
    dec3(i, dec, c8, xs, ys, ri);
    const long long r = ri;
    const int ym = ys > 0 ? ys - 1 : 0, yp = ys < H - 1 ? ys + 1 : H - 1;
    const int xm = xs > 0 ? xs - 1 : 0, xp = xs < W - 1 ? xs + 1 : W - 1;
    const __nv_bfloat16* base = in + (r * (H + 2)) * (long long)(W + 2) * C + c8 * 8;
    const int rows[3] = {ym + 1, ys + 1, yp + 1};
    float he[3][8], ho[3][8];                       // horizontally interpolated even / odd output columns per source row
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const __nv_bfloat16* rowp = base + (long long)rows[q] * (W + 2) * C;
      BF8 l, m, n;
      l.u = __ldg(reinterpret_cast<const uint4*>(rowp + (long long)(xm + 1) * C));
      m.u = __ldg(reinterpret_cast<const uint4*>(rowp + (long long)(xs + 1) * C));
      n.u = __ldg(reinterpret_cast<const uint4*>(rowp + (long long)(xp + 1) * C));
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float c = 0.75f * bf2f(m.h[j]);
        he[q][j] = 0.25f * bf2f(l.h[j]) + c;
        ho[q][j] = c + 0.25f * bf2f(n.h[j]);
      }
    }
    BF8 o00, o01, o10, o11;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      o00.h[j] = f2bf(0.25f * he[0][j] + 0.75f * he[1][j]);
      o01.h[j] = f2bf(0.25f * ho[0][j] + 0.75f * ho[1][j]);
      o10.h[j] = f2bf(0.75f * he[1][j] + 0.25f * he[2][j]);
      o11.h[j] = f2bf(0.75f * ho[1][j] + 0.25f * ho[2][j]);
    }
    __nv_bfloat16* ob = out + ((r * (H2 + 2) + 2 * ys + 1) * (long long)(W2 + 2) + 2 * xs + 1) * C + c8 * 8;
    *reinterpret_cast<uint4*>(ob) = o00.u;
    *reinterpret_cast<uint4*>(ob + C) = o01.u;
    *reinterpret_cast<uint4*>(ob + (long long)(W2 + 2) * C) = o10.u;
    *reinterpret_cast<uint4*>(ob + (long long)(W2 + 2) * C + C) = o11.u;
  }
}
// Backward (exact transpose): one thread = a 2x2 block of INPUT pixels (8 channels), gathering the 6x6 window of
// output gradients that touch it.  Input index y receives from output rows 2y-1 .. 2y+2 with weights
// {0.25, 0.75, 0.75, 0.25}; at the first / last index the missing outer row's weight folds into its neighbour (1.0).
__device__ __forceinline__ void up2_tw(int y, int n, float* w) {
  w[0] = y > 0 ? 0.25f : 0.f; w[1] = y > 0 ? 0.75f : 1.f;
  w[2] = y < n - 1 ? 0.75f : 1.f; w[3] = y < n - 1 ? 0.25f : 0.f;
}
constexpr int kUp2Ty = 4;     // input rows per CTA in the backward
// Separable in shared memory: a CTA owns kUp2Ty input rows of one image.  Phase 1 combines the four output rows that
// touch each input row (vertical taps) for every OUTPUT column into fp32 V[ty][ox][c]; phase 2 combines four columns
// of V per input pixel.  Each output gradient is read ~1.25 times (rows shared by neighbouring tiles) instead of
// nine times from L1/L2 in a per-pixel gather.
__global__ void __launch_bounds__(512, 2)
up2_bwd_kernel(const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ din, int H, int W, int C, int nty,
               const Dec3 d1, const Dec3 d2) {
  extern __shared__ __align__(16) float sV[];        // [kUp2Ty][2W][C + 4]  (+4: column stride off the bank period)
  const int Cp = C + 4;
  const int C8 = C / 8, W2 = 2 * W, H2 = 2 * H;
  const long long r = blockIdx.x / nty;
  const int y0 = (blockIdx.x % nty) * kUp2Ty;
  const __nv_bfloat16* base = dout + (r * (H2 + 2)) * (long long)(W2 + 2) * C;
  const int n1 = kUp2Ty * W2 * C8;
  // phase 1 in batches of kB items per thread: all 4 * kB loads are issued before any is consumed (the kernel is
  // bound by the number of bytes in flight, not by arithmetic)
  constexpr int kB = 2;
  for (int i0 = threadIdx.x; i0 < n1; i0 += kB * blockDim.x) {
    uint4 v[kB][4];
    float wy[kB][4];
    int so[kB];
#pragma unroll
    for (int u = 0; u < kB; ++u) {
      const int i = i0 + u * blockDim.x;
      so[u] = -1;
#pragma unroll
      for (int k = 0; k < 4; ++k) { v[u][k] = make_uint4(0, 0, 0, 0); wy[u][k] = 0.f; }
      if (i < n1) {
        int c8, ox, ty, unused;
        dec3((uint32_t)i, d1, c8, ox, ty, unused);
        const int y = y0 + ty;
        if (y < H) {
          up2_tw(y, H, wy[u]);
          so[u] = (ty * W2 + ox) * Cp + c8 * 4;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (wy[u][k] != 0.f)                     // also excludes the rows outside the image
              v[u][k] = __ldg(reinterpret_cast<const uint4*>(base + ((long long)(2 * y + k) * (W2 + 2) + ox + 1) * C + c8 * 8));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kB; ++u) {
      if (so[u] < 0) continue;
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        BF8 t;
        t.u = v[u][k];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += wy[u][k] * bf2f(t.h[j]);
      }
      // channels 8 c8 .. +3 at column 4 c8, channels +4 .. +7 at column C / 2 + 4 c8: consecutive threads are 16 bytes
      // apart in both stores (a 32-byte stride made every float4 store a 2-way bank conflict)
      *reinterpret_cast<float4*>(sV + so[u]) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(sV + so[u] + C / 2) = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
  }
  __syncthreads();
  const int n2 = kUp2Ty * W * C8;
  for (int i = threadIdx.x; i < n2; i += blockDim.x) {
    int c8, x, ty, unused;
    dec3((uint32_t)i, d2, c8, x, ty, unused);
    const int y = y0 + ty;
    if (y >= H) continue;
    float wx[4];
    up2_tw(x, W, wx);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (wx[k] == 0.f) continue;
      const float* src = sV + ((long long)ty * W2 + 2 * x - 1 + k) * Cp + c8 * 4;
      const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + C / 2);
      acc[0] += wx[k] * a.x; acc[1] += wx[k] * a.y; acc[2] += wx[k] * a.z; acc[3] += wx[k] * a.w;
      acc[4] += wx[k] * b.x; acc[5] += wx[k] * b.y; acc[6] += wx[k] * b.z; acc[7] += wx[k] * b.w;
    }
    BF8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.h[j] = f2bf(acc[j]);
    *reinterpret_cast<uint4*>(din + ((r * (H + 2) + y + 1) * (long long)(W + 2) + x + 1) * C + c8 * 8) = o.u;
  }
}

int zero_border(void* x, long long R, int H, int W, int C, cudaStream_t st);
int up2_fwd(const void* in, void* out, long long R, int H, int W, int C, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (C % 8 || H <= 0 || W <= 0) return AS_ESIZE;
  if (!in || !out) return AS_EINVAL;
  const long long total = R * (long long)H * W * (C / 8);
  if (total >= (1LL << 31)) return AS_ESIZE;
  up2_fwd_kernel<<<ew_grid(total, 1), 256, 0, st>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, R, H, W, C,
                                                    make_dec3(C / 8, W, H));
  AS_LAUNCH_CHECK();
  return zero_border(out, R, 2 * H, 2 * W, C, st);
}
int up2_bwd(const void* dout, void* din, long long R, int H, int W, int C, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (C % 8 || H <= 0 || W <= 0) return AS_ESIZE;
  if (!dout || !din) return AS_EINVAL;
  const int nty = (H + kUp2Ty - 1) / kUp2Ty;
  const int smem = kUp2Ty * 2 * W * (C + 4) * 4;
  if (smem > 200 * 1024 || R * nty >= (1LL << 31)) return AS_ESIZE;
  static int attr_smem = 0;
  if (smem > attr_smem) {
    cudaError_t e = cudaFuncSetAttribute(up2_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    attr_smem = smem;
  }
  up2_bwd_kernel<<<(unsigned)(R * nty), 512, smem, st>>>((const __nv_bfloat16*)dout, (__nv_bfloat16*)din, H, W, C, nty,
                                                       make_dec3(C / 8, 2 * W, kUp2Ty), make_dec3(C / 8, W, kUp2Ty));
  AS_LAUNCH_CHECK();
  return zero_border(din, R, H, W, C, st);
}

// ------------------------------------------------------------------------------------------ location-head input
// x1 = relu(cat[reshape(emb,[4,H,W]), skip]) (action_arg_head.py:419-422) as padded NHWC bf16 with Cp = 192 channels
// (0-3 embedding, 4-131 skip, rest zero).  emb fp32 [R, 4*H*W] (post-ReLU), skip padded NHWC bf16 [R,H+2,W+2,128].
// thread = (padded pixel, group of 8 output channels): one 16-byte store; the skip channels arrive shifted by four
// channels, i.e. as two 8-byte loads
__global__ void __launch_bounds__(256)
loc_cat_fwd_kernel(const float* __restrict__ emb, const __nv_bfloat16* __restrict__ skip, __nv_bfloat16* __restrict__ out,
                   long long R, int H, int W, int Cs, int Cp, const Dec3 dec) {
  const int G = Cp / 8;
  const uint32_t total = (uint32_t)(R * (H + 2) * (long long)(W + 2) * G);
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    int g, xx, yy, ri;
    dec3(i, dec, g, xx, yy, ri);
    const long long r = ri;
    const long long p = (r * (H + 2) + yy) * (long long)(W + 2) + xx;
    BF8 o;
    o.u = make_uint4(0, 0, 0, 0);
    if (xx >= 1 && xx <= W && yy >= 1 && yy <= H) {
      const int c0 = g * 8;                                        // output channels c0 .. c0+7
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int c = c0 + half * 4;                               // four channels at a time
        if (c == 0) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            o.h[j] = f2bf(fmaxf(__ldg(emb + (r * 4 + j) * (long long)(H * W) + (yy - 1) * W + xx - 1), 0.f));
        } else if (c - 4 + 4 <= Cs) {
          const uint2 v = __ldg(reinterpret_cast<const uint2*>(skip + p * Cs + c - 4));
          const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const float2 f = __bfloat1622float2(h2[j]);
            o.h[half * 4 + 2 * j] = f2bf(fmaxf(f.x, 0.f));
            o.h[half * 4 + 2 * j + 1] = f2bf(fmaxf(f.y, 0.f));
          }
        }
      }
    }
    *reinterpret_cast<uint4*>(out + p * Cp + g * 8) = o.u;
  }
}
// thread = (padded pixel, group of 8 skip channels); the group-0 thread also scatters the four embedding gradients
__global__ void __launch_bounds__(256)
loc_cat_bwd_kernel(const __nv_bfloat16* __restrict__ dcat, const float* __restrict__ emb,
                   const __nv_bfloat16* __restrict__ skip, float* __restrict__ demb, __nv_bfloat16* __restrict__ dskip,
                   long long R, int H, int W, int Cs, int Cp, const Dec3 dec) {
  const int G = Cs / 8;
  const uint32_t total = (uint32_t)(R * (H + 2) * (long long)(W + 2) * G);
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    int g, xx, yy, ri;
    dec3(i, dec, g, xx, yy, ri);
    const long long r = ri;
    const long long p = (r * (H + 2) + yy) * (long long)(W + 2) + xx;
    const bool inner = xx >= 1 && xx <= W && yy >= 1 && yy <= H;
    BF8 o;
    o.u = make_uint4(0, 0, 0, 0);
    if (inner) {
      BF8 sk;
      sk.u = __ldg(reinterpret_cast<const uint4*>(skip + p * Cs + g * 8));
      const uint2 d0 = __ldg(reinterpret_cast<const uint2*>(dcat + p * Cp + 4 + g * 8));
      const uint2 d1 = __ldg(reinterpret_cast<const uint2*>(dcat + p * Cp + 8 + g * 8));
      const __nv_bfloat16* dh0 = reinterpret_cast<const __nv_bfloat16*>(&d0);
      const __nv_bfloat16* dh1 = reinterpret_cast<const __nv_bfloat16*>(&d1);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        o.h[j] = bf2f(sk.h[j]) > 0.f ? dh0[j] : f2bf(0.f);
        o.h[4 + j] = bf2f(sk.h[4 + j]) > 0.f ? dh1[j] : f2bf(0.f);
      }
      if (g == 0) {
        const uint2 de = __ldg(reinterpret_cast<const uint2*>(dcat + p * Cp));
        const __nv_bfloat16* eh = reinterpret_cast<const __nv_bfloat16*>(&de);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const long long e = (r * 4 + j) * (long long)(H * W) + (yy - 1) * W + xx - 1;
          demb[e] = __ldg(emb + e) > 0.f ? bf2f(eh[j]) : 0.f;
        }
      }
    }
    *reinterpret_cast<uint4*>(dskip + p * Cs + g * 8) = o.u;
  }
}

int loc_cat_fwd(const float* emb, const void* skip, void* out, long long R, int H, int W, int Cs, int Cp, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (Cp < 4 + Cs || (Cp % 8) || (Cs % 8)) return AS_ESIZE;
  if (!emb || !skip || !out) return AS_EINVAL;
  const long long total = R * (H + 2) * (long long)(W + 2) * (Cp / 8);
  if (total >= (1LL << 31)) return AS_ESIZE;
  loc_cat_fwd_kernel<<<ew_grid(total, 1), 256, 0, st>>>(emb, (const __nv_bfloat16*)skip, (__nv_bfloat16*)out, R, H, W, Cs, Cp,
                                                        make_dec3(Cp / 8, W + 2, H + 2));
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int loc_cat_bwd(const void* dcat, const float* emb, const void* skip, float* demb, void* dskip, long long R, int H, int W,
                int Cs, int Cp, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!dcat || !emb || !skip || !demb || !dskip) return AS_EINVAL;
  if (Cp < 4 + Cs || (Cp % 8) || (Cs % 8)) return AS_ESIZE;
  const long long total = R * (H + 2) * (long long)(W + 2) * (Cs / 8);
  if (total >= (1LL << 31)) return AS_ESIZE;
  loc_cat_bwd_kernel<<<ew_grid(total, 1), 256, 0, st>>>((const __nv_bfloat16*)dcat, emb, (const __nv_bfloat16*)skip, demb,
                                                        (__nv_bfloat16*)dskip, R, H, W, Cs, Cp, make_dec3(Cs / 8, W + 2, H + 2));
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------------------------------ final 3x3 conv, 32 -> 1
// y[r,y,x] = (b + sum_{tap,c} w[tap][c] * in[r, y+ky, x+kx, c]) / T   (in padded NHWC bf16, C = 32; y fp32 unpadded =
// the location logits).  Four threads per output pixel, eight channels each: every tap is ONE 16-byte load per thread
// and a warp's load covers 512 contiguous bytes (8 pixels); two xor-shuffles finish the channel sum.
__device__ __forceinline__ void unpack8(const uint4& q, float* f) {
  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 v = __bfloat1622float2(h2[e]);
    f[2 * e] = v.x; f[2 * e + 1] = v.y;
  }
}

// These three kernels are bound by the L1 / LSU data pipe, not by HBM, when every output re-loads its nine taps
// (ncu: 14.3 GB through L1 for 1.66 GB of input, lsu wavefronts at 84 % of peak).  So: four lanes (8 channels each)
// own one pixel COLUMN and walk down a segment of rows with a sliding 3x3 window in registers -- three 16-byte loads
// per output instead of nine, and a warp's load is still 512 contiguous bytes.
constexpr int kC32Seg = 38;      // rows per thread segment (152 = 4 * 38)

__device__ __forceinline__ float dot8w(const uint4& q, const float (&wt)[8]) {
  float f[8];
  unpack8(q, f);
  return f[0] * wt[0] + f[1] * wt[1] + f[2] * wt[2] + f[3] * wt[3] + f[4] * wt[4] + f[5] * wt[5] + f[6] * wt[6] + f[7] * wt[7];
}

__global__ void __launch_bounds__(256, 2)
conv_c32_1_fwd_kernel(const __nv_bfloat16* __restrict__ in, const float* __restrict__ w, const float* __restrict__ bias,
                      float* __restrict__ out, long long R, int H, int W, float inv_temp, int nseg, const Dec3 dec) {
  const int cg = threadIdx.x & 3;
  float sw[9][8];                                   // this lane's 8 channels of every tap, in registers (shared-memory
#pragma unroll                                      // reads would share the LSU pipe with the pixel loads)
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int c = 0; c < 8; ++c) sw[t][c] = __ldg(w + t * 32 + cg * 8 + c);
  const float b = bias ? bias[0] : 0.f;
  const long long total = R * (long long)nseg * W;                 // pixel columns x row segments
  const long long total_pad = (total + 7) & ~7LL;                  // whole warps stay in the loop for the shuffles
  const long long stride = ((long long)gridDim.x * blockDim.x) >> 2;
  for (long long col = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 2; col < total_pad; col += stride) {
    const bool ok = col < total;
    int x, seg, ri, unused;
    dec3((uint32_t)(ok ? col : 0), dec, x, seg, ri, unused);
    const long long r = ri;
    const int y0 = seg * kC32Seg, y1 = min(H, y0 + kC32Seg);
    const long long pitch = (long long)(W + 2) * 32;
    const __nv_bfloat16* p = in + ((r * (H + 2) + y0) * (long long)(W + 2) + x) * 32 + cg * 8;
    // Every padded input row is loaded and unpacked ONCE and feeds the three output rows it belongs to (as tap row 0 of
    // output j, 1 of j - 1, 2 of j - 2): the bf16 -> fp32 unpacking runs on the half-rate ALU pipe, which was the limit
    // (73 % busy) when each row was unpacked again in each of the three iterations that used it.
    const int nrow = y1 - y0 + 2;
    uint4 rn[3];
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) rn[kx] = __ldg(reinterpret_cast<const uint4*>(p + kx * 32));
    float o1 = 0.f, o2 = 0.f;                          // partial sums of output rows j - 1 and j - 2
    for (int j = 0; j < nrow; ++j) {
      uint4 rr[3];
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) rr[kx] = rn[kx];
      if (j + 1 < nrow) {
        const __nv_bfloat16* pr = p + (long long)(j + 1) * pitch;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) rn[kx] = __ldg(reinterpret_cast<const uint4*>(pr + kx * 32));
      }
      float d0 = 0.f, d1 = 0.f, d2 = 0.f;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        float f[8];
        unpack8(rr[kx], f);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          d0 = fmaf(f[c], sw[0 * 3 + kx][c], d0);
          d1 = fmaf(f[c], sw[1 * 3 + kx][c], d1);
          d2 = fmaf(f[c], sw[2 * 3 + kx][c], d2);
        }
      }
      if (j >= 2) {
        float acc = o2 + d2;
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        if (ok && cg == 0) out[(r * H + (y0 + j - 2)) * (long long)W + x] = (acc + b) * inv_temp;
      }
      o2 = o1 + d1;
      o1 = d0;
    }
  }
}

// din[r,yy,xx,c] = sum_tap w[tap][c] * dy[r, yy-ky, xx-kx] / T  (interior; border 0), dy fp32 unpadded [R,H,W].
// Thread per input pixel, 32 channel accumulators in registers, 64 contiguous bytes written per thread.
__global__ void __launch_bounds__(256)
conv_c32_1_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w, __nv_bfloat16* __restrict__ din,
                        long long R, int H, int W, float inv_temp, const Dec3 dec) {
  __shared__ __align__(16) float sw[288];
  for (int i = threadIdx.x; i < 288; i += blockDim.x) sw[i] = w[i] * inv_temp;
  __syncthreads();
  // Two pixels per thread (p and p + 256 of a 512-pixel group): every 16-byte weight read from shared memory feeds
  // both, which halves the traffic of the L1 / shared-memory pipe (82 % busy with one pixel per thread).
  const long long total = R * (long long)(H + 2) * (W + 2);
  const long long stride = (long long)gridDim.x * 512;
  for (long long base = (long long)blockIdx.x * 512 + threadIdx.x; base < total; base += stride) {
    float acc[2][32];
    float g[2][9];
    long long pp[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long p = base + u * 256;
      pp[u] = p;
#pragma unroll
      for (int c = 0; c < 32; ++c) acc[u][c] = 0.f;
#pragma unroll
      for (int t = 0; t < 9; ++t) g[u][t] = 0.f;
      if (p < total) {
        int xx, yy, ri, unused;
        dec3((uint32_t)p, dec, xx, yy, ri, unused);
        const long long r = ri;
        if (xx >= 1 && xx <= W && yy >= 1 && yy <= H) {
#pragma unroll
          for (int ky = 0; ky < 3; ++ky) {
            const int oy = yy - ky;
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
              const int ox = xx - kx;
              if (oy >= 0 && oy < H && ox >= 0 && ox < W) g[u][ky * 3 + kx] = __ldg(dy + (r * H + oy) * (long long)W + ox);
            }
          }
        }
      }
    }
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float* wt = sw + t * 32;
#pragma unroll
      for (int c4 = 0; c4 < 8; ++c4) {
        const float4 wv = *reinterpret_cast<const float4*>(wt + c4 * 4);
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          acc[u][c4 * 4] = fmaf(g[u][t], wv.x, acc[u][c4 * 4]);
          acc[u][c4 * 4 + 1] = fmaf(g[u][t], wv.y, acc[u][c4 * 4 + 1]);
          acc[u][c4 * 4 + 2] = fmaf(g[u][t], wv.z, acc[u][c4 * 4 + 2]);
          acc[u][c4 * 4 + 3] = fmaf(g[u][t], wv.w, acc[u][c4 * 4 + 3]);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (pp[u] >= total) continue;
      uint4* dst = reinterpret_cast<uint4*>(din + pp[u] * 32);
#pragma unroll
      for (int c4 = 0; c4 < 4; ++c4) {
        __nv_bfloat162 h0 = __floats2bfloat162_rn(acc[u][c4 * 8], acc[u][c4 * 8 + 1]), h1 = __floats2bfloat162_rn(acc[u][c4 * 8 + 2], acc[u][c4 * 8 + 3]);
        __nv_bfloat162 h2 = __floats2bfloat162_rn(acc[u][c4 * 8 + 4], acc[u][c4 * 8 + 5]), h3 = __floats2bfloat162_rn(acc[u][c4 * 8 + 6], acc[u][c4 * 8 + 7]);
        uint4 q;
        q.x = *reinterpret_cast<uint32_t*>(&h0); q.y = *reinterpret_cast<uint32_t*>(&h1);
        q.z = *reinterpret_cast<uint32_t*>(&h2); q.w = *reinterpret_cast<uint32_t*>(&h3);
        dst[c4] = q;
      }
    }
  }
}
// dw[tap][c] = sum_p dy[p] in[p+tap][c];  db = sum dy  -> per-block partials [blocks][9*32+1].  Same column-walking
// thread mapping as the forward: 72 fp32 accumulators per thread, three 16-byte loads per pixel, then a shuffle
// reduction over the warp's eight pixel columns and a shared-memory reduction over the block's warps.
__global__ void __launch_bounds__(256, 2)
conv_c32_1_wgrad_kernel(const float* __restrict__ dy, const __nv_bfloat16* __restrict__ in,
                        float* __restrict__ partial, long long R, int H, int W, float inv_temp, int nseg, const Dec3 dec) {
  const int cg = threadIdx.x & 3, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float acc[9][8], accb = 0.f;
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int c = 0; c < 8; ++c) acc[t][c] = 0.f;
  const long long total = R * (long long)nseg * W;
  const long long stride = ((long long)gridDim.x * blockDim.x) >> 2;
  for (long long col = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 2; col < total; col += stride) {
    int x, seg, ri, unused;
    dec3((uint32_t)col, dec, x, seg, ri, unused);
    const long long r = ri;
    const int y0 = seg * kC32Seg, y1 = min(H, y0 + kC32Seg);
    const long long pitch = (long long)(W + 2) * 32;
    const __nv_bfloat16* p = in + ((r * (H + 2) + y0) * (long long)(W + 2) + x) * 32 + cg * 8;
    // each padded row is loaded and unpacked once and multiplied by the gradients of the three output rows it belongs
    // to (tap row 0 of output j, 1 of j - 1, 2 of j - 2)
    const int nout = y1 - y0, nrow = nout + 2;
    uint4 rn[3];
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) rn[kx] = __ldg(reinterpret_cast<const uint4*>(p + kx * 32));
    float gn = __ldg(dy + (r * H + y0) * (long long)W + x);
    float g1 = 0.f, g2 = 0.f;
    for (int j = 0; j < nrow; ++j) {
      uint4 rr[3];
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) rr[kx] = rn[kx];
      const float g0 = j < nout ? gn * inv_temp : 0.f;
      if (j + 1 < nrow) {
        const __nv_bfloat16* pr = p + (long long)(j + 1) * pitch;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) rn[kx] = __ldg(reinterpret_cast<const uint4*>(pr + kx * 32));
        if (j + 1 < nout) gn = __ldg(dy + (r * H + y0 + j + 1) * (long long)W + x);
      }
      accb += g0;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        float f[8];
        unpack8(rr[kx], f);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          acc[kx][c] = fmaf(g0, f[c], acc[kx][c]);
          acc[3 + kx][c] = fmaf(g1, f[c], acc[3 + kx][c]);
          acc[6 + kx][c] = fmaf(g2, f[c], acc[6 + kx][c]);
        }
      }
      g2 = g1;
      g1 = g0;
    }
  }
  __shared__ float red[8][9 * 32 + 1];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      float a = acc[t][c];
      a += __shfl_xor_sync(0xffffffffu, a, 4);
      a += __shfl_xor_sync(0xffffffffu, a, 8);
      a += __shfl_xor_sync(0xffffffffu, a, 16);
      if (lane < 4) red[wid][t * 32 + cg * 8 + c] = a;
    }
  accb += __shfl_xor_sync(0xffffffffu, accb, 4);
  accb += __shfl_xor_sync(0xffffffffu, accb, 8);
  accb += __shfl_xor_sync(0xffffffffu, accb, 16);
  if (lane == 0) red[wid][288] = accb;
  __syncthreads();
  for (int i = threadIdx.x; i < 289; i += blockDim.x) {
    float s = 0.f;
    for (int k = 0; k < 8; ++k) s += red[k][i];
    partial[(long long)blockIdx.x * 289 + i] = s;
  }
}
__global__ void conv_c32_1_wreduce_kernel(const float* __restrict__ partial, int blocks, float* __restrict__ dw,
                                          float* __restrict__ db) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 289) return;
  float s = 0.f;
  for (int k = 0; k < blocks; ++k) s += partial[(long long)k * 289 + i];
  if (i < 288) dw[i] = s; else if (db) db[0] = s;
}

constexpr int kC321Blocks = 148 * 8;
long long conv_c32_1_scratch_floats() { return (long long)kC321Blocks * 289; }

int conv_c32_1_fwd(const void* in, const float* w, const float* bias, float* out, long long R, int H, int W, float inv_temp,
                   cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!in || !w || !out) return AS_EINVAL;
  const int nseg = (H + kC32Seg - 1) / kC32Seg;
  const long long cols = R * (long long)nseg * W;
  if (R * (long long)(H + 2) * (W + 2) >= (1LL << 31)) return AS_ESIZE;
  conv_c32_1_fwd_kernel<<<ew_grid(cols * 4, 1), 256, 0, st>>>((const __nv_bfloat16*)in, w, bias, out, R, H, W, inv_temp, nseg,
                                                             make_dec3(W, nseg, 1u << 30));
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int conv_c32_1_bwd(const float* dy, const void* in, const float* w, void* din, float* dw, float* db, float* scratch,
                   long long R, int H, int W, float inv_temp, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!dy || !in || !w || !din || !dw || !scratch) return AS_EINVAL;
  if (R * (long long)(H + 2) * (W + 2) >= (1LL << 31)) return AS_ESIZE;
  conv_c32_1_dgrad_kernel<<<ew_grid((R * (long long)(H + 2) * (W + 2) + 1) / 2, 1), 256, 0, st>>>(dy, w, (__nv_bfloat16*)din, R, H, W, inv_temp,
                                                                                            make_dec3(W + 2, H + 2, 1u << 30));
  AS_LAUNCH_CHECK();
  const int nseg = (H + kC32Seg - 1) / kC32Seg;
  conv_c32_1_wgrad_kernel<<<kC321Blocks, 256, 0, st>>>(dy, (const __nv_bfloat16*)in, scratch, R, H, W, inv_temp, nseg,
                                                       make_dec3(W, nseg, 1u << 30));
  AS_LAUNCH_CHECK();
  conv_c32_1_wreduce_kernel<<<3, 128, 0, st>>>(scratch, kC321Blocks, dw, db);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------------------------------ column sums (bias grads)
// out[c] = sum_r x[r][c] for bf16 x [rows, C]; two-stage, fixed order.  C % 8 == 0: a thread owns 8 columns (one
// 16-byte load per row), 16 threads span 128 columns, 16 row lanes per block with four loads in flight each.
__global__ void __launch_bounds__(256)
colsum_partial_v8_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ partial, long long rows, int C,
                         int chunks, int txn) {
  // txn (4, 8 or 16) threads span the columns of a block (8 columns each), 256 / txn row lanes
  const int tx = threadIdx.x & (txn - 1), ty = threadIdx.x / txn, tyn = 256 / txn;
  const int c = (blockIdx.x * txn + tx) * 8;
  const long long per = (rows + chunks - 1) / chunks;
  const long long r0 = blockIdx.y * per, r1 = (r0 + per < rows) ? r0 + per : rows;
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  if (c < C) {
    const __nv_bfloat16* col = x + c;
    long long r = r0 + ty;
    for (; r + 3 * tyn < r1; r += 4 * tyn) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(reinterpret_cast<const uint4*>(col + (r + (long long)tyn * u) * C));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = __bfloat1622float2(h2[e]);
          s[2 * e] += f.x; s[2 * e + 1] += f.y;
        }
      }
    }
    for (; r < r1; r += tyn) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(col + r * C));
      const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h2[e]);
        s[2 * e] += f.x; s[2 * e + 1] += f.y;
      }
    }
  }
  __shared__ float sm[64 * 36];                    // [tyn][txn * 8 + 4]
  const int pitch = txn * 8 + 4;
#pragma unroll
  for (int j = 0; j < 8; ++j) sm[ty * pitch + tx * 8 + j] = s[j];
  __syncthreads();
  if (threadIdx.x < txn * 8) {
    const int cc = blockIdx.x * txn * 8 + threadIdx.x;
    if (cc < C) {
      float t = 0.f;
      for (int i = 0; i < tyn; ++i) t += sm[i * pitch + threadIdx.x];
      partial[(long long)blockIdx.y * C + cc] = t;
    }
  }
}
// any even C: 32 column pairs x 8 row lanes
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ partial, long long rows, int C,
                      int chunks) {
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;      // 32 column pairs x 8 row lanes
  const int c = (blockIdx.x * 32 + cx) * 2;
  const long long per = (rows + chunks - 1) / chunks;
  const long long r0 = blockIdx.y * per, r1 = (r0 + per < rows) ? r0 + per : rows;
  float s0 = 0.f, s1 = 0.f;
  if (c < C) {
    for (long long r = r0 + ry; r < r1; r += 8) {
      const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(x + r * C + c));
      s0 += v.x; s1 += v.y;
    }
  }
  __shared__ float sa[8][33], sb[8][33];
  sa[ry][cx] = s0; sb[ry][cx] = s1;
  __syncthreads();
  if (ry == 0 && c < C) {
    float t0 = 0.f, t1 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { t0 += sa[i][cx]; t1 += sb[i][cx]; }
    partial[(long long)blockIdx.y * C + c] = t0;
    partial[(long long)blockIdx.y * C + c + 1] = t1;
  }
}
__global__ void __launch_bounds__(256)
colsum_partial_f32_kernel(const float* __restrict__ x, float* __restrict__ partial, long long rows, int C, int chunks) {
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  const long long per = (rows + chunks - 1) / chunks;
  const long long r0 = blockIdx.y * per, r1 = (r0 + per < rows) ? r0 + per : rows;
  float s0 = 0.f;
  if (c < C)
    for (long long r = r0 + ry; r < r1; r += 8) s0 += __ldg(x + r * C + c);
  __shared__ float sa[8][33];
  sa[ry][cx] = s0;
  __syncthreads();
  if (ry == 0 && c < C) {
    float t0 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t0 += sa[i][cx];
    partial[(long long)blockIdx.y * C + c] = t0;
  }
}

// second stage: 32 columns x 8 chunk lanes per block, fixed summation order
__global__ void __launch_bounds__(256)
colsum_reduce_kernel(const float* __restrict__ partial, float* __restrict__ out, int C, int chunks) {
  const int cx = threadIdx.x & 31, ky = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float s = 0.f;
  if (c < C)
    for (int k = ky; k < chunks; k += 8) s += partial[(long long)k * C + c];
  __shared__ float sa[8][33];
  sa[ky][cx] = s;
  __syncthreads();
  if (ky == 0 && c < C) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sa[i][cx];
    out[c] = t;
  }
}
constexpr int kColsumChunks = 1024;          // scratch rows the callers provide (as_colsum_scratch_floats)
constexpr int kColsumMaxUsed = 148 * 4;      // chunks actually launched: four row chunks per SM
long long colsum_scratch_floats(int C) { return (long long)kColsumChunks * C; }
int colsum_bf16(const void* x, float* out, float* scratch, long long rows, int C, cudaStream_t st) {
  if (rows <= 0 || C <= 0 || (C % 2) != 0) return AS_ESIZE;
  if (!x || !out || !scratch) return AS_EINVAL;
  long long want = (rows + 255) / 256;
  const int chunks = (int)(want < 1 ? 1 : (want > kColsumMaxUsed ? kColsumMaxUsed : want));
  if (C % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    const int txn = C >= 128 ? 16 : (C >= 64 ? 8 : 4);
    dim3 g((C + txn * 8 - 1) / (txn * 8), chunks);
    colsum_partial_v8_kernel<<<g, 256, 0, st>>>((const __nv_bfloat16*)x, scratch, rows, C, chunks, txn);
  } else {
    dim3 g((C + 63) / 64, chunks);
    colsum_partial_kernel<<<g, 256, 0, st>>>((const __nv_bfloat16*)x, scratch, rows, C, chunks);
  }
  AS_LAUNCH_CHECK();
  colsum_reduce_kernel<<<(C + 31) / 32, 256, 0, st>>>(scratch, out, C, chunks);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as

namespace as {

// ------------------------------------------------------------------------------------------ spatial input features
// The 56-channel spatial input of the reference (spatial_encoder.py:52-71, channel order App. A8: height/256,
// visibility(4), creep(2), player_relative(5), alerts(2), pathable(2), buildable(2), 6 effect planes, 32 scatter
// channels) written once as a zero-padded NHWC bf16 map with Cp = 64 channels, ready for the 1x1 project GEMM.
// CTA = (observation, band of rows); the effect index lists become a per-band bitmap in shared memory (a zero
// padded list sets pixel 0, App. A7).  Thread = pixel: 128 contiguous bytes written per thread.
struct SpatialPlanes {
  const uint8_t* plane[7];     // height_map, visibility_map, creep, player_relative, alerts, pathable, buildable [R,H,W]
  const int16_t* effect[6];    // [R, n_eff] flat pixel ids
  int n_eff;
};

constexpr int kFeatBand = 8;   // rows per CTA

// ScatterConnection operands for the fused mode (scatter_map == nullptr): the 32 scatter channels of a pixel are summed
// straight from the entity embeddings, in ascending entity order like as_scatter_connection_fwd (bit-identical fp32
// sums), so the dense fp32 map [R,32,H,W] (3.2 GB written and read back at the benchmark size) never exists.
struct ScatterSrc {
  const float* emb;            // [R, N, 32]
  const void* xs;              // [R, N] uint8 or int64
  const void* ys;
  const long long* entity_num; // [R] or nullptr
  int idx_bytes, N;
};

__global__ void __launch_bounds__(256, 4)
spatial_features_kernel(const SpatialPlanes sp, const float* __restrict__ scatter_map, const ScatterSrc sc,
                        __nv_bfloat16* __restrict__ out, int H, int W, int Cs) {
  __shared__ unsigned int effbits[kFeatBand * 256];      // one byte-sized bit set per pixel, W <= 1024/… (W*band <= 1024*… checked on host)
  __shared__ __align__(16) uint4 stile[8][32 * 8];       // per-warp output staging (below); before that: unsorted keys
  __shared__ uint32_t sorted[512];                       // fused scatter: (band-local cell << 9 | entity) of the band's entities
  uint32_t* skeys = reinterpret_cast<uint32_t*>(&stile[0][0]);
  __shared__ uint16_t cmap[kFeatBand * 256];             // band-local cell -> 1 + index of its first sorted key
  __shared__ int s_count;
  const int r = blockIdx.x;
  const int band = blockIdx.y;                            // padded row band: rows [band*kFeatBand, +kFeatBand) of H+2
  const int HW = H * W;
  const int n_pix = kFeatBand * W;
  for (int i = threadIdx.x; i < (n_pix + 3) / 4; i += blockDim.x) effbits[i] = 0u;
  __syncthreads();
  const int y_lo = band * kFeatBand - 1;                  // first interior row (unpadded coords) of the band
  for (int e = 0; e < 6; ++e) {
    const int16_t* ids = sp.effect[e] + (long long)r * sp.n_eff;
    for (int i = threadIdx.x; i < sp.n_eff; i += blockDim.x) {
      const int id = (int)ids[i];
      if (id < 0 || id >= HW) continue;
      const int y = id / W - y_lo;
      if (y >= 0 && y < kFeatBand) {
        const int local = y * W + id % W;
        atomicOr(&effbits[local >> 2], (1u << e) << ((local & 3) * 8));
      }
    }
  }
  __syncthreads();
  int n_b = 0;
  long long en_r = 0;
  if (!scatter_map) {
    en_r = sc.entity_num ? sc.entity_num[r] : (long long)sc.N;
    if (threadIdx.x == 0) s_count = 0;
    for (int i = threadIdx.x; i < n_pix; i += blockDim.x) cmap[i] = 0;
    __syncthreads();
    for (int e = threadIdx.x; e < sc.N; e += blockDim.x) {
      if ((long long)e >= en_r) continue;                                   // encoder.py:38 mask: contributes 0
      long long x, y;
      if (sc.idx_bytes == 1) {
        x = reinterpret_cast<const uint8_t*>(sc.xs)[(long long)r * sc.N + e];
        y = reinterpret_cast<const uint8_t*>(sc.ys)[(long long)r * sc.N + e];
      } else {
        x = reinterpret_cast<const long long*>(sc.xs)[(long long)r * sc.N + e];
        y = reinterpret_cast<const long long*>(sc.ys)[(long long)r * sc.N + e];
      }
      x = x < 0 ? 0 : (x > W - 1 ? W - 1 : x);                              // module_utils.py:18-19
      y = y < 0 ? 0 : (y > H - 1 ? H - 1 : y);
      const int yl = (int)y - y_lo;
      if (yl >= 0 && yl < kFeatBand) skeys[atomicAdd(&s_count, 1)] = ((uint32_t)(yl * W + (int)x) << 9) | (uint32_t)e;
    }
    __syncthreads();
    n_b = s_count;
    for (int i = threadIdx.x; i < n_b; i += blockDim.x) {                    // rank sort (keys are unique)
      const uint32_t k = skeys[i];
      int rank = 0;
      for (int j = 0; j < n_b; ++j) rank += skeys[j] < k ? 1 : 0;
      sorted[rank] = k;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n_b; i += blockDim.x) {
      const uint32_t cell = sorted[i] >> 9;
      if (i == 0 || (sorted[i - 1] >> 9) != cell) cmap[cell] = (uint16_t)(i + 1);
    }
    __syncthreads();
  }
  const int Cp = 64;
  // every lane builds the 128 bytes of its pixel (zeros for the border ring and for lanes past the band); the warp's 32
  // pixels are contiguous in the output, so they are staged in a swizzled 4 KB tile and written as 512-byte rows
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n_band = kFeatBand * (W + 2);
  for (int i0 = 0; i0 < n_band; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const int yy = band * kFeatBand + i / (W + 2);
    const int xx = i % (W + 2);
    const bool valid = i < n_band && yy < H + 2;
    // the pixel's eight 16-byte chunks are packed as bf16 bit patterns directly (one-hot = 0x3F80 or 0) and stored into
    // the staging tile one by one: no 64-float image of the pixel is ever live (it cost 128 registers = 2 CTAs per SM)
    uint4* srow = &stile[warp][lane * 8];
    const int sw = lane & 7;
    const bool inside = valid && !(yy == 0 || yy == H + 1 || xx == 0 || xx == W + 1);
    if (!inside) {
#pragma unroll
      for (int q = 0; q < 8; ++q) srow[q ^ sw] = make_uint4(0, 0, 0, 0);
    } else {
      const int y = yy - 1, x = xx - 1;
      const long long pix = ((long long)r * H + y) * W + x;
      const uint32_t kOne = 0x3F80u;
      auto oh = [&](int v, int k) -> uint32_t { return v == k ? kOne : 0u; };
      const int hgt = sp.plane[0][pix];
      int vis = sp.plane[1][pix], crp = sp.plane[2][pix], pr = sp.plane[3][pix], alr = sp.plane[4][pix];
      int pth = sp.plane[5][pix], bld = sp.plane[6][pix];
      vis = vis > 3 ? 3 : vis; crp = crp > 1 ? 1 : crp; pr = pr > 4 ? 4 : pr; alr = alr > 1 ? 1 : alr;
      pth = pth > 1 ? 1 : pth; bld = bld > 1 ? 1 : bld;
      const int local = (y - y_lo) * W + x;
      const unsigned int bits = (effbits[local >> 2] >> ((local & 3) * 8)) & 0xFFu;
      const uint32_t hb = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn((float)hgt * (1.f / 256.f)));
      srow[0 ^ sw] = make_uint4(hb | (oh(vis, 0) << 16), oh(vis, 1) | (oh(vis, 2) << 16), oh(vis, 3) | (oh(crp, 0) << 16),
                                oh(crp, 1) | (oh(pr, 0) << 16));
      srow[1 ^ sw] = make_uint4(oh(pr, 1) | (oh(pr, 2) << 16), oh(pr, 3) | (oh(pr, 4) << 16), oh(alr, 0) | (oh(alr, 1) << 16),
                                oh(pth, 0) | (oh(pth, 1) << 16));
      auto eb = [&](int e) -> uint32_t { return ((bits >> e) & 1u) ? kOne : 0u; };
      srow[2 ^ sw] = make_uint4(oh(bld, 0) | (oh(bld, 1) << 16), eb(0) | (eb(1) << 16), eb(2) | (eb(3) << 16), eb(4) | (eb(5) << 16));
      int m = 0;
      uint32_t cell = 0;
      if (!scatter_map) { cell = (uint32_t)local; m = cmap[cell]; }
#pragma unroll
      for (int q = 0; q < 4; ++q) {                                // scatter channels 8 q .. 8 q + 7 -> chunk 3 + q
        float f[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) f[c] = 0.f;
        if (scatter_map) {
          const float* smp = scatter_map + (long long)r * Cs * HW + (long long)y * W + x;
#pragma unroll
          for (int c = 0; c < 8; ++c) f[c] = (8 * q + c < Cs) ? __ldg(smp + (long long)(8 * q + c) * HW) : 0.f;
        } else if (m) {
          for (int j = m - 1; j < n_b && (sorted[j] >> 9) == cell; ++j) {       // ascending entity index: exact sum order
            const float4* src = reinterpret_cast<const float4*>(sc.emb + ((long long)r * sc.N + (sorted[j] & 511u)) * 32) + 2 * q;
            const float4 v0 = __ldg(src), v1 = __ldg(src + 1);
            f[0] += v0.x; f[1] += v0.y; f[2] += v0.z; f[3] += v0.w;
            f[4] += v1.x; f[5] += v1.y; f[6] += v1.z; f[7] += v1.w;
          }
        }
        __nv_bfloat162 h0 = __floats2bfloat162_rn(f[0], f[1]), h1 = __floats2bfloat162_rn(f[2], f[3]);
        __nv_bfloat162 h2 = __floats2bfloat162_rn(f[4], f[5]), h3 = __floats2bfloat162_rn(f[6], f[7]);
        srow[(3 + q) ^ sw] = make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1),
                                        *reinterpret_cast<uint32_t*>(&h2), *reinterpret_cast<uint32_t*>(&h3));
      }
      srow[7 ^ sw] = make_uint4(0, 0, 0, 0);
    }
    __syncwarp();
    // first pixel of this warp in the band, and how many of its 32 pixels exist
    const int iw = i0 + warp * 32;
    int nvalid = n_band - iw;
    const int rows_left = (H + 2 - band * kFeatBand) * (W + 2) - iw;         // pixels of the image below this point
    nvalid = nvalid < rows_left ? nvalid : rows_left;
    nvalid = nvalid > 32 ? 32 : nvalid;
    if (nvalid > 0) {
      uint4* gbase = reinterpret_cast<uint4*>(out + (((long long)r * (H + 2) + band * kFeatBand) * (W + 2) + iw) * Cp);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int idx = k * 32 + lane;
        const int px = idx >> 3, ch = idx & 7;
        if (px < nvalid) gbase[idx] = stile[warp][px * 8 + (ch ^ (px & 7))];
      }
    }
    __syncwarp();
  }
}


int spatial_features(const void* const* planes, const void* const* effects, int n_eff, const float* scatter_map,
                     const float* emb, const void* xs, const void* ys, int idx_bytes, const long long* entity_num, int N,
                     void* out, long long R, int H, int W, int Cs, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (H <= 0 || W <= 0 || W > 256 || Cs != 32 || n_eff <= 0) return AS_ESIZE;
  if (!planes || !effects || !out) return AS_EINVAL;
  ScatterSrc sc = {emb, xs, ys, entity_num, idx_bytes, N};
  if (!scatter_map) {
    if (!emb || !xs || !ys) return AS_EINVAL;
    if (N <= 0 || N > 512) return AS_ESIZE;
    if (idx_bytes != 1 && idx_bytes != 8) return AS_EUNSUPPORTED;
    if (reinterpret_cast<uintptr_t>(emb) & 15) return AS_EALIGN;
  }
  SpatialPlanes sp;
  for (int i = 0; i < 7; ++i) { sp.plane[i] = reinterpret_cast<const uint8_t*>(planes[i]); if (!planes[i]) return AS_EINVAL; }
  for (int i = 0; i < 6; ++i) { sp.effect[i] = reinterpret_cast<const int16_t*>(effects[i]); if (!effects[i]) return AS_EINVAL; }
  sp.n_eff = n_eff;
  dim3 grid((unsigned)R, (unsigned)((H + 2 + kFeatBand - 1) / kFeatBand));
  spatial_features_kernel<<<grid, 256, 0, st>>>(sp, scatter_map, sc, reinterpret_cast<__nv_bfloat16*>(out), H, W, Cs);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ScatterConnection backward straight from the gradient of the padded NHWC feature map:
// d_emb[r,e,d] = e < entity_num[r] ? d_feat[r, y(e)+1, x(e)+1, ch0 + d] : 0
template <typename IdxT>
__global__ void scatter_bwd_pnhwc_kernel(const __nv_bfloat16* __restrict__ dfeat, const IdxT* __restrict__ xs,
                                         const IdxT* __restrict__ ys, const long long* __restrict__ entity_num,
                                         float* __restrict__ demb, long long R, int N, int D, int H, int W, int Cp,
                                         int ch0) {
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
      g = bf2f(dfeat[((r * (H + 2) + y + 1) * (long long)(W + 2) + x + 1) * Cp + ch0 + d]);
    }
    demb[i] = g;
  }
}

int scatter_bwd_pnhwc(const void* dfeat, const void* xs, const void* ys, int idx_bytes, const long long* en, float* demb,
                      long long R, int N, int D, int H, int W, int Cp, int ch0, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if (!dfeat || !xs || !ys || !demb) return AS_EINVAL;
  const long long total = R * N * (long long)D;
  const int grid = ew_grid(total, 1);
  if (idx_bytes == 1)
    scatter_bwd_pnhwc_kernel<uint8_t><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dfeat, (const uint8_t*)xs, (const uint8_t*)ys, en, demb, R, N, D, H, W, Cp, ch0);
  else if (idx_bytes == 8)
    scatter_bwd_pnhwc_kernel<long long><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dfeat, (const long long*)xs, (const long long*)ys, en, demb, R, N, D, H, W, Cp, ch0);
  else
    return AS_EUNSUPPORTED;
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// ------------------------------------------------------------------------------------------ 2x2 max-pool, padded NHWC bf16
// forward: out[r,py,px,c] = max over the 2x2 window; backward: gradient to the FIRST maximum in row-major window
// order (torch.nn.functional.max_pool2d semantics), zero elsewhere and on the border.
__global__ void maxpool2_fwd_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, long long R,
                                    int H, int W, int C, const Dec3 dec) {
  const int Ho = H / 2, Wo = W / 2, C8 = C / 8;
  const uint32_t total = (uint32_t)(R * (long long)(Ho + 2) * (Wo + 2) * C8);
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    int c8, xx, yy, ri;
    dec3(i, dec, c8, xx, yy, ri);
    const long long r = ri;
    BF8 o;
    if (xx == 0 || xx == Wo + 1 || yy == 0 || yy == Ho + 1) {
      o.u = make_uint4(0, 0, 0, 0);
    } else {
      const __nv_bfloat16* base = in + ((r * (H + 2) + 2 * (yy - 1) + 1) * (long long)(W + 2) + 2 * (xx - 1) + 1) * C + c8 * 8;
      BF8 a, b, c, d;
      a.u = *reinterpret_cast<const uint4*>(base);
      b.u = *reinterpret_cast<const uint4*>(base + C);
      c.u = *reinterpret_cast<const uint4*>(base + (long long)(W + 2) * C);
      d.u = *reinterpret_cast<const uint4*>(base + (long long)(W + 2) * C + C);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        o.h[j] = f2bf(fmaxf(fmaxf(bf2f(a.h[j]), bf2f(b.h[j])), fmaxf(bf2f(c.h[j]), bf2f(d.h[j]))));
    }
    *reinterpret_cast<uint4*>(out + (long long)i * 8) = o.u;
  }
}

__global__ void maxpool2_bwd_kernel(const __nv_bfloat16* __restrict__ in, const __nv_bfloat16* __restrict__ dout,
                                    __nv_bfloat16* __restrict__ din, long long R, int H, int W, int C, const Dec3 dec,
                                    int relu_input) {
  // thread per (pooled pixel incl. border ring handled separately, 8 channels): writes the four input pixels.
  // relu_input: `in` is the output of a ReLU layer -- a window whose maximum is 0 routes nothing, which IS that layer's
  // ReLU backward mask (every routed element is positive), so its backward can skip the mask pass
  const int Ho = H / 2, Wo = W / 2, C8 = C / 8;
  const uint32_t total = (uint32_t)(R * (long long)Ho * Wo * C8);
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    int c8, px, py, ri;
    dec3(i, dec, c8, px, py, ri);
    const long long r = ri;
    const long long ibase = ((r * (H + 2) + 2 * py + 1) * (long long)(W + 2) + 2 * px + 1) * C + c8 * 8;
    BF8 a, b, c, d, g, oa, ob, oc, od;
    a.u = *reinterpret_cast<const uint4*>(in + ibase);
    b.u = *reinterpret_cast<const uint4*>(in + ibase + C);
    c.u = *reinterpret_cast<const uint4*>(in + ibase + (long long)(W + 2) * C);
    d.u = *reinterpret_cast<const uint4*>(in + ibase + (long long)(W + 2) * C + C);
    g.u = *reinterpret_cast<const uint4*>(dout + ((r * (Ho + 2) + py + 1) * (long long)(Wo + 2) + px + 1) * C + c8 * 8);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float va = bf2f(a.h[j]), vb = bf2f(b.h[j]), vc = bf2f(c.h[j]), vd = bf2f(d.h[j]);
      int best = 0; float m = va;
      if (vb > m) { m = vb; best = 1; }
      if (vc > m) { m = vc; best = 2; }
      if (vd > m) { m = vd; best = 3; }
      if (relu_input && !(m > 0.f)) best = -1;
      const __nv_bfloat16 z = f2bf(0.f);
      oa.h[j] = best == 0 ? g.h[j] : z; ob.h[j] = best == 1 ? g.h[j] : z;
      oc.h[j] = best == 2 ? g.h[j] : z; od.h[j] = best == 3 ? g.h[j] : z;
    }
    *reinterpret_cast<uint4*>(din + ibase) = oa.u;
    *reinterpret_cast<uint4*>(din + ibase + C) = ob.u;
    *reinterpret_cast<uint4*>(din + ibase + (long long)(W + 2) * C) = oc.u;
    *reinterpret_cast<uint4*>(din + ibase + (long long)(W + 2) * C + C) = od.u;
  }
}

// zero the one-pixel border ring of a padded NHWC map
__global__ void zero_border_kernel(__nv_bfloat16* __restrict__ x, long long R, int H, int W, int C) {
  const int C8 = C / 8;
  const int ring = 2 * (W + 2) + 2 * H;
  const long long total = R * (long long)ring * C8;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c8 = (int)(i % C8);
    long long p = i / C8;
    const int k = (int)(p % ring);
    const long long r = p / ring;
    int yy, xx;
    if (k < W + 2) { yy = 0; xx = k; }
    else if (k < 2 * (W + 2)) { yy = H + 1; xx = k - (W + 2); }
    else { const int q = k - 2 * (W + 2); yy = 1 + q / 2; xx = (q & 1) ? W + 1 : 0; }
    *reinterpret_cast<uint4*>(x + ((r * (H + 2) + yy) * (long long)(W + 2) + xx) * C + c8 * 8) = make_uint4(0, 0, 0, 0);
  }
}

int zero_border(void* x, long long R, int H, int W, int C, cudaStream_t st) {
  const long long ring = R * (long long)(2 * (W + 2) + 2 * H) * (C / 8);
  zero_border_kernel<<<ew_grid(ring, 1), 256, 0, st>>>((__nv_bfloat16*)x, R, H, W, C);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int maxpool2_fwd(const void* in, void* out, long long R, int H, int W, int C, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if ((C % 8) || (H % 2) || (W % 2) || H <= 0 || W <= 0) return AS_ESIZE;
  if (!in || !out) return AS_EINVAL;
  const long long total = R * (long long)(H / 2 + 2) * (W / 2 + 2) * (C / 8);
  if (total >= (1LL << 31)) return AS_ESIZE;
  maxpool2_fwd_kernel<<<ew_grid(total, 1), 256, 0, st>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, R, H, W, C,
                                                         make_dec3(C / 8, W / 2 + 2, H / 2 + 2));
  AS_LAUNCH_CHECK();
  return AS_OK;
}

int maxpool2_bwd(const void* in, const void* dout, void* din, long long R, int H, int W, int C, int relu_input, cudaStream_t st) {
  if (R <= 0) return AS_OK;
  if ((C % 8) || (H % 2) || (W % 2) || H <= 0 || W <= 0) return AS_ESIZE;
  if (!in || !dout || !din) return AS_EINVAL;
  const long long total = R * (long long)(H / 2) * (W / 2) * (C / 8);
  if (total >= (1LL << 31)) return AS_ESIZE;
  maxpool2_bwd_kernel<<<ew_grid(total, 1), 256, 0, st>>>((const __nv_bfloat16*)in, (const __nv_bfloat16*)dout, (__nv_bfloat16*)din, R, H, W, C,
                                                         make_dec3(C / 8, W / 2, H / 2), relu_input);
  AS_LAUNCH_CHECK();
  const long long ring = R * (long long)(2 * (W + 2) + 2 * H) * (C / 8);
  zero_border_kernel<<<ew_grid(ring, 1), 256, 0, st>>>((__nv_bfloat16*)din, R, H, W, C);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// dz = dy * (y > 0) AND out[c] = sum_r dz[r][c] in one pass (the ReLU backward and the bias gradient of a
// convolution read the same tensor; fusing them saves one full read of dz).  Same thread mapping as
// colsum_partial_v8_kernel: C % 8 == 0, 16-byte loads / stores.
__global__ void __launch_bounds__(256)
relu_bwd_colsum_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                       __nv_bfloat16* __restrict__ dz, float* __restrict__ partial, long long rows, int C, int chunks,
                       int txn) {
  const int tx = threadIdx.x & (txn - 1), ty = threadIdx.x / txn, tyn = 256 / txn;
  const int c = (blockIdx.x * txn + tx) * 8;
  const long long per = (rows + chunks - 1) / chunks;
  const long long r0 = blockIdx.y * per, r1 = (r0 + per < rows) ? r0 + per : rows;
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  if (c < C) {
    long long r = r0 + ty;
    for (; r + 3 * tyn < r1; r += 4 * tyn) {
      BF8 vd[4], vy[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        vd[u].u = __ldg(reinterpret_cast<const uint4*>(dy + (r + (long long)tyn * u) * C + c));
        vy[u].u = __ldg(reinterpret_cast<const uint4*>(y + (r + (long long)tyn * u) * C + c));
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        BF8 vo;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const bool on = bf2f(vy[u].h[j]) > 0.f;
          vo.h[j] = on ? vd[u].h[j] : f2bf(0.f);
          s[j] += on ? bf2f(vd[u].h[j]) : 0.f;
        }
        *reinterpret_cast<uint4*>(dz + (r + (long long)tyn * u) * C + c) = vo.u;
      }
    }
    for (; r < r1; r += tyn) {
      BF8 vd, vy, vo;
      vd.u = __ldg(reinterpret_cast<const uint4*>(dy + r * C + c));
      vy.u = __ldg(reinterpret_cast<const uint4*>(y + r * C + c));
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const bool on = bf2f(vy.h[j]) > 0.f;
        vo.h[j] = on ? vd.h[j] : f2bf(0.f);
        s[j] += on ? bf2f(vd.h[j]) : 0.f;
      }
      *reinterpret_cast<uint4*>(dz + r * C + c) = vo.u;
    }
  }
  __shared__ float sm[64 * 36];
  const int pitch = txn * 8 + 4;
#pragma unroll
  for (int j = 0; j < 8; ++j) sm[ty * pitch + tx * 8 + j] = s[j];
  __syncthreads();
  if (threadIdx.x < txn * 8) {
    const int cc = blockIdx.x * txn * 8 + threadIdx.x;
    if (cc < C) {
      float t = 0.f;
      for (int i = 0; i < tyn; ++i) t += sm[i * pitch + threadIdx.x];
      partial[(long long)blockIdx.y * C + cc] = t;
    }
  }
}

int relu_bwd_colsum(const void* dy, const void* y, void* dz, float* out, float* scratch, long long rows, int C,
                    cudaStream_t st) {
  if (rows <= 0 || C <= 0 || (C % 8) != 0) return AS_ESIZE;
  if (!dy || !y || !dz || !out || !scratch) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(dz)) & 15) return AS_EALIGN;
  long long want = (rows + 127) / 128;
  const int chunks = (int)(want < 1 ? 1 : (want > 1024 ? 1024 : want));   // scratch: as_colsum_scratch_floats(C) = 1024 * C floats
  const int txn = C >= 128 ? 16 : (C >= 64 ? 8 : 4);
  dim3 g((C + txn * 8 - 1) / (txn * 8), chunks);
  relu_bwd_colsum_kernel<<<g, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (__nv_bfloat16*)dz, scratch,
                                            rows, C, chunks, txn);
  AS_LAUNCH_CHECK();
  colsum_reduce_kernel<<<(C + 31) / 32, 256, 0, st>>>(scratch, out, C, chunks);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// Backward "prologue" of a linear layer in ONE pass over the fp32 upstream gradient [rows, C] (C % 8 == 0):
//   dz = dy * (y > 0)   (y fp32 or bf16, optional)   ->  bf16 copy for the dgrad / wgrad GEMMs
//   colsum[c] = sum_r dz[r][c]                         ->  bias gradient (optional)
// replaces three passes (ReLU mask, fp32 -> bf16 cast, column sum).  Same thread mapping as colsum_partial_v8_kernel.
__global__ void __launch_bounds__(256)
linear_bwd_prep_kernel(const float* __restrict__ dy, const __nv_bfloat16* __restrict__ dy_extra, const float* __restrict__ y32,
                       const __nv_bfloat16* __restrict__ y16,
                       __nv_bfloat16* __restrict__ dz, float* __restrict__ partial, long long rows, int C, int chunks,
                       int txn) {
  const int tx = threadIdx.x & (txn - 1), ty = threadIdx.x / txn, tyn = 256 / txn;
  const int c = (blockIdx.x * txn + tx) * 8;
  const long long per = (rows + chunks - 1) / chunks;
  const long long r0 = blockIdx.y * per, r1 = (r0 + per < rows) ? r0 + per : rows;
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  if (c < C) {
    for (long long r = r0 + ty; r < r1; r += 2 * tyn) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const long long rr = r + (long long)u * tyn;
        if (rr >= r1) break;
        const float4 a = __ldg(reinterpret_cast<const float4*>(dy + rr * C + c));
        const float4 b = __ldg(reinterpret_cast<const float4*>(dy + rr * C + c) + 1);
        float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        if (dy_extra) {                                  // second gradient of the same output (its bf16 copy had a consumer too)
          BF8 ev;
          ev.u = __ldg(reinterpret_cast<const uint4*>(dy_extra + rr * C + c));
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] += bf2f(ev.h[j]);
        }
        if (y32) {
          const float4 ya = __ldg(reinterpret_cast<const float4*>(y32 + rr * C + c));
          const float4 yb = __ldg(reinterpret_cast<const float4*>(y32 + rr * C + c) + 1);
          const float yy[8] = {ya.x, ya.y, ya.z, ya.w, yb.x, yb.y, yb.z, yb.w};
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = yy[j] > 0.f ? f[j] : 0.f;
        } else if (y16) {
          BF8 yv;
          yv.u = __ldg(reinterpret_cast<const uint4*>(y16 + rr * C + c));
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = bf2f(yv.h[j]) > 0.f ? f[j] : 0.f;
        }
        BF8 o;
#pragma unroll
        for (int j = 0; j < 8; ++j) { o.h[j] = f2bf(f[j]); s[j] += f[j]; }
        *reinterpret_cast<uint4*>(dz + rr * C + c) = o.u;
      }
    }
  }
  if (partial == nullptr) return;
  __shared__ float sm[64 * 36];
  const int pitch = txn * 8 + 4;
#pragma unroll
  for (int j = 0; j < 8; ++j) sm[ty * pitch + tx * 8 + j] = s[j];
  __syncthreads();
  if (threadIdx.x < txn * 8) {
    const int cc = blockIdx.x * txn * 8 + threadIdx.x;
    if (cc < C) {
      float t = 0.f;
      for (int i = 0; i < tyn; ++i) t += sm[i * pitch + threadIdx.x];
      partial[(long long)blockIdx.y * C + cc] = t;
    }
  }
}

int linear_bwd_prep(const float* dy, const void* dy_extra, const float* y32, const void* y16, void* dz, float* colsum,
                    float* scratch, long long rows, int C, cudaStream_t st) {
  if (rows <= 0 || C <= 0 || (C % 8) != 0) return AS_ESIZE;
  if (!dy || !dz || (colsum && !scratch)) return AS_EINVAL;
  if ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dy_extra) | reinterpret_cast<uintptr_t>(y32) |
       reinterpret_cast<uintptr_t>(y16) |
       reinterpret_cast<uintptr_t>(dz)) & 15) return AS_EALIGN;
  long long want = (rows + 63) / 64;
  const int chunks = (int)(want < 1 ? 1 : (want > 1024 ? 1024 : want));   // scratch: as_colsum_scratch_floats(C) = 1024 * C floats
  const int txn = C >= 128 ? 16 : (C >= 64 ? 8 : 4);
  dim3 g((C + txn * 8 - 1) / (txn * 8), chunks);
  linear_bwd_prep_kernel<<<g, 256, 0, st>>>(dy, (const __nv_bfloat16*)dy_extra, y32, (const __nv_bfloat16*)y16, (__nv_bfloat16*)dz, colsum ? scratch : nullptr,
                                            rows, C, chunks, txn);
  AS_LAUNCH_CHECK();
  if (colsum) {
    colsum_reduce_kernel<<<(C + 31) / 32, 256, 0, st>>>(scratch, colsum, C, chunks);
    AS_LAUNCH_CHECK();
  }
  return AS_OK;
}

int colsum_f32(const float* x, float* out, float* scratch, long long rows, int C, cudaStream_t st) {
  if (rows <= 0 || C <= 0) return AS_ESIZE;
  if (!x || !out || !scratch) return AS_EINVAL;
  long long want = (rows + 255) / 256;
  const int chunks = (int)(want < 1 ? 1 : (want > kColsumChunks ? kColsumChunks : want));
  dim3 g((C + 31) / 32, chunks);
  colsum_partial_f32_kernel<<<g, 256, 0, st>>>(x, scratch, rows, C, chunks);
  AS_LAUNCH_CHECK();
  colsum_reduce_kernel<<<(C + 31) / 32, 256, 0, st>>>(scratch, out, C, chunks);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as

// ------------------------------------------------------------------------------------------ column sums without any allocation
// as_colsum_bf16 needs a caller-provided scratch buffer for its per-chunk partial sums.  Some hosts cannot provide
// one: processes that hand their whole device-memory budget to an arena at start-up and forbid any later cudaMalloc,
// and step loops in which the device-wide synchronisation that cudaMalloc / cudaFree imply would stall every other
// stream.  For them the same two kernels run on a fixed scratch area that lives in the library image itself.
//
// The area is a file-scope __device__ array, so every GPU that loads the library carries it (1 MiB) until the
// process exits.  In host code the array's NAME is only a host-side handle, not a device address: the address is
// looked up once per device with cudaGetSymbolAddress (with that device current) and that pointer is what the
// kernels receive.
namespace as {

constexpr int kStaticColsumMaxC = 256;                                   // widest row the static scratch can serve
static __device__ __align__(256) float g_colsum_static_scratch[kColsumChunks * kStaticColsumMaxC];

static float* g_colsum_static_ptr[64] = {nullptr};                       // per-device address of the array above

int colsum_bf16_static(const void* x, float* out, long long rows, int C, cudaStream_t st) {
  if (rows <= 0 || C <= 0 || (C % 2) != 0) return AS_ESIZE;
  if (C > kStaticColsumMaxC) return AS_ESIZE;                            // beyond the built-in scratch: use as_colsum_bf16
  if (!x || !out) return AS_EINVAL;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return AS_EUNSUPPORTED;
  if (g_colsum_static_ptr[dev] == nullptr) {                             // set-up: first call on this device
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_static_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return AS_EDRIVER;
    g_colsum_static_ptr[dev] = p;
  }
  // every partial the reduce kernel reads is written by the partial kernel of the same call, so nothing in the
  // shared area has to be cleared between calls
  return colsum_bf16(x, out, g_colsum_static_ptr[dev], rows, C, st);
}

// bf16 operand layouts of a Conv2d weight [Cout, Cin, k, k] fp32 for the implicit-GEMM kernels, one launch:
//   w_fwd[co][tap][ci]          (ci < cin_pad, zero beyond Cin)            forward:  A = pixels x (tap, ci)
//   w_dg [ci][flipped tap][co]  (ci < cin_pad, rows beyond Cin zero)       data gradient: contraction over (tap, co)
__global__ void conv_weight_layouts_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ w_fwd,
                                           __nv_bfloat16* __restrict__ w_dg, int cout, int cin, int cin_pad, int taps) {
  const long long total = (long long)cout * taps * cin_pad;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    {   // forward layout: i = (co * taps + tap) * cin_pad + ci
      const int ci = (int)(i % cin_pad);
      const int tap = (int)((i / cin_pad) % taps);
      const int co = (int)(i / ((long long)cin_pad * taps));
      const float v = ci < cin ? __ldg(w + ((long long)co * cin + ci) * taps + tap) : 0.f;
      w_fwd[i] = f2bf(v);
    }
    {   // data-gradient layout: i = (ci * taps + tf) * cout + co, tf = flipped tap
      const int co = (int)(i % cout);
      const int tf = (int)((i / cout) % taps);
      const int ci = (int)(i / ((long long)cout * taps));
      const float v = ci < cin ? __ldg(w + ((long long)co * cin + ci) * taps + (taps - 1 - tf)) : 0.f;
      w_dg[i] = f2bf(v);
    }
  }
}
int conv_weight_layouts(const float* w, void* w_fwd, void* w_dg, int cout, int cin, int cin_pad, int taps, cudaStream_t st) {
  if (cout <= 0 || cin <= 0 || cin_pad < cin || taps <= 0) return AS_ESIZE;
  if (!w || !w_fwd || !w_dg) return AS_EINVAL;
  const long long total = (long long)cout * taps * cin_pad;
  const int grid = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  conv_weight_layouts_kernel<<<grid, 256, 0, st>>>(w, (__nv_bfloat16*)w_fwd, (__nv_bfloat16*)w_dg, cout, cin, cin_pad, taps);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

// Spatial-encoder FC weight (spatial_encoder.py:88-89): the reference flattens NCHW (column c*H*W + y*W + x); the
// activations here are padded NHWC, column ((y+1)*(W+2) + (x+1))*C + c, border columns zero.
//   fwd: w16[o][padded column] = bf16(w[o][nchw column])            bwd: dw[o][nchw column] = dwp[o][padded column]
__global__ void spatial_fc_weight_fwd_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ w16, int O, int C, int H,
                                             int W) {
  const int pw = W + 2, ph = H + 2;
  const long long total = (long long)O * ph * pw * C;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = (int)(i % C);
    long long t = i / C;
    const int px = (int)(t % pw); t /= pw;
    const int py = (int)(t % ph);
    const long long o = t / ph;
    float v = 0.f;
    if (px >= 1 && px <= W && py >= 1 && py <= H) v = __ldg(w + o * ((long long)C * H * W) + ((long long)c * H + (py - 1)) * W + (px - 1));
    w16[i] = f2bf(v);
  }
}
__global__ void spatial_fc_weight_bwd_kernel(const float* __restrict__ dwp, float* __restrict__ dw, int O, int C, int H, int W) {
  const int pw = W + 2, ph = H + 2;
  const long long total = (long long)O * C * H * W;
  const long long stride = (long long)gridDim.x * blockDim.x;
  // thread index follows the PADDED (source) layout so that the big read is coalesced; the scattered 4-byte writes of one
  // warp land in C different 128-byte lines that neighbouring warps complete
  const long long total_p = (long long)O * ph * pw * C;
  (void)total;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total_p; i += stride) {
    const int c = (int)(i % C);
    long long t = i / C;
    const int px = (int)(t % pw); t /= pw;
    const int py = (int)(t % ph);
    const long long o = t / ph;
    if (px >= 1 && px <= W && py >= 1 && py <= H)
      dw[o * ((long long)C * H * W) + ((long long)c * H + (py - 1)) * W + (px - 1)] = __ldg(dwp + i);
  }
}
int spatial_fc_weight_fwd(const float* w, void* w16, int O, int C, int H, int W, cudaStream_t st) {
  if (O <= 0 || C <= 0 || H <= 0 || W <= 0) return AS_ESIZE;
  if (!w || !w16) return AS_EINVAL;
  spatial_fc_weight_fwd_kernel<<<148 * 16, 256, 0, st>>>(w, (__nv_bfloat16*)w16, O, C, H, W);
  AS_LAUNCH_CHECK();
  return AS_OK;
}
int spatial_fc_weight_bwd(const float* dwp, float* dw, int O, int C, int H, int W, cudaStream_t st) {
  if (O <= 0 || C <= 0 || H <= 0 || W <= 0) return AS_ESIZE;
  if (!dwp || !dw) return AS_EINVAL;
  spatial_fc_weight_bwd_kernel<<<148 * 16, 256, 0, st>>>(dwp, dw, O, C, H, W);
  AS_LAUNCH_CHECK();
  return AS_OK;
}

}  // namespace as
