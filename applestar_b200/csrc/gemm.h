// Internal (C++) description of one tcgen05 GEMM launch; the extern "C" wrappers live in abi.cu.
#pragma once
#include <cuda_runtime.h>

namespace as {

struct GemmCall {
  // C[b][m][n] = epi(alpha * sum_k A[b][m][k] * B[b][n][k] + bias[n])
  const void* A = nullptr;   // bf16
  const void* B = nullptr;   // bf16
  int M = 0, N = 0, K = 0;
  bool a_mn = false;         // false: A stored [M rows][K contiguous] (lda = row pitch); true: [K rows][M contiguous]
  bool b_mn = false;         // false: B stored [N rows][K contiguous];                   true: [K rows][N contiguous]
  long long lda = 0, ldb = 0;
  long long a_rows = 0;      // K-major A only: number of valid rows in global memory (defaults to M)
  int nb0 = 1, nb1 = 1;      // batch = nb0 * nb1
  long long a_sb0 = 0, a_sb1 = 0, b_sb0 = 0, b_sb1 = 0;   // batch strides (elements)
  void* C = nullptr;
  void* C2 = nullptr;        // optional bf16 copy of C (only when C is fp32)
  int c_bf16 = 0;
  long long ldc = 0, c_sb0 = 0, c_sb1 = 0;
  const float* bias = nullptr;
  const void* resid = nullptr;   // optional bf16 tensor with C's layout (ldc, batch strides), added before ReLU (TMA-store path)
  // ReLU sign bits, one uint32 per 32 output columns, row pitch N / 32 words (batch 1, N % 32 == 0, TMA-store epilogue):
  // bits_out is written by a relu GEMM (bit j of word [m][n / 32] = output [m][n + j] > 0), bits_in zeroes the output where
  // the bit is clear -- the same ReLU backward as `mask` at 1/16 of its bytes
  uint32_t* bits_out = nullptr;
  const uint32_t* bits_in = nullptr;
  const void* mask = nullptr;    // optional bf16 tensor with C's layout: the output is zeroed where mask <= 0 (the ReLU
                                 // backward of the layer that produced `mask`, applied to the gradient this GEMM computes)
  int relu = 0;
  float alpha = 1.f;
  int splits = 1;            // >1: fp32 partials to ws[splits][batch][M][N]
  float* ws = nullptr;
  int taps = 0;              // >0: implicit-GEMM conv; K = Cin, B = [N][taps*Cin], A rows shifted by shift[tap]
  int shift[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  int img_w = 0, img_h = 0;  // padded image dims: border rows are written as zero
  int b_k_shift = 0;         // MN-major B: offset (rows) added to the contraction coordinate
  int wgrad_taps = 0;        // >0: conv weight-gradient mode, batch dim 0 = tap (see gemm_tcgen05.cu)
  int bn_override = 0;
  // conv weight gradients with a 64-channel M side: two taps share one 128-row MMA (see gemm_tcgen05.cu, wg_pair)
  int wg_pair = 0;
  int pair_a[5][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}};   // contraction-row offset of A's two 64-column chunks
  int pair_b[5] = {0, 0, 0, 0, 0};                                 // contraction-row offset of B
};

int gemm_run(const GemmCall& g, cudaStream_t st);
int splitk_reduce(const float* ws, float* out, long long n, int splits, int accumulate, cudaStream_t st);
int wgrad_pair_reduce(const float* ws, float* out, int splits, long long stride, int taps, int cout, int cin, int transposed,
                      int swap_halves, cudaStream_t st);
int wgrad_slab_ok(int mode, int Cin, int Cout, int pw);
int wgrad_slab(int mode, const void* dy, const void* x, float* ws, long long P, int pw, int Cin, int Cout, int splits,
               cudaStream_t st);

}  // namespace as
