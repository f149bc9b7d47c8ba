"""Diagnostics: where does a small-K GEMM spend its time?  Times one shape with the epilogue progressively disabled."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import _lib, kernels as K
lib = _lib.lib()
lib.as_debug_gemm_mode.argtypes = [ctypes.c_int]
def t(M, N, Kd, bf16, mode, iters=5):
    A = torch.randn(M, Kd, device='cuda').to(torch.bfloat16); B = torch.randn(N, Kd, device='cuda').to(torch.bfloat16)
    C = torch.empty((M, N), device='cuda', dtype=torch.bfloat16 if bf16 else torch.float32)
    lib.as_debug_gemm_mode(mode)
    for _ in range(2): K.gemm(A, B, C, M, N, Kd)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): K.gemm(A, B, C, M, N, Kd)
    e1.record(); torch.cuda.synchronize(); lib.as_debug_gemm_mode(0)
    return e0.elapsed_time(e1) / iters
for (M, N, Kd, bf) in [(532480, 768, 256, True), (532480, 256, 256, False), (532480, 1024, 256, True), (532480, 768, 64, True), (532480, 768, 1024, True),
                      (8192, 8192, 8192, True)]:
    print(M, N, Kd, 'bf16' if bf else 'f32', ' full %.3f  no-store %.3f  no-fence %.3f  legacy-store %.3f  drain-only %.3f ms' % tuple(t(M, N, Kd, bf, m) for m in (0, 1, 3, 8, 2)), flush=True)


def tm(M, N, Kd, mode, masked, iters=5):
    """FFN dgrad shape: dz1[M, N] = (dz2[M, Kd] @ W2[Kd, N]) * (h > 0), B read MN-major."""
    A = torch.randn(M, Kd, device='cuda').to(torch.bfloat16); B = torch.randn(Kd, N, device='cuda').to(torch.bfloat16)
    C = torch.empty((M, N), device='cuda', dtype=torch.bfloat16)
    H = torch.randn(M, N, device='cuda').to(torch.bfloat16) if masked else None
    lib.as_debug_gemm_mode(mode)
    def run():
        _lib.call('as_gemm_bf16_masked', A.data_ptr(), B.data_ptr(), C.data_ptr(), H.data_ptr() if masked else None, M, N, Kd, 0, 1,
                  Kd, N, N, 1, 1.0, _lib.stream_ptr())
    for _ in range(2): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): run()
    e1.record(); torch.cuda.synchronize(); lib.as_debug_gemm_mode(0)
    return e0.elapsed_time(e1) / iters
for masked in (False, True):
    print('dgrad 532480x1024x256 B mn-major, masked=%s' % masked, ' full %.3f  no-store %.3f  no-fence %.3f  drain-only %.3f ms' % tuple(tm(532480, 1024, 256, m, masked) for m in (0, 1, 3, 2)), flush=True)


def tk(M, N, Kd, mode, masked, iters=5):
    """same product with the weight held K-major (W2^T [N, Kd])"""
    A = torch.randn(M, Kd, device='cuda').to(torch.bfloat16); B = torch.randn(N, Kd, device='cuda').to(torch.bfloat16)
    C = torch.empty((M, N), device='cuda', dtype=torch.bfloat16)
    H = torch.randn(M, N, device='cuda').to(torch.bfloat16) if masked else None
    lib.as_debug_gemm_mode(mode)
    def run():
        _lib.call('as_gemm_bf16_masked', A.data_ptr(), B.data_ptr(), C.data_ptr(), H.data_ptr() if masked else None, M, N, Kd, 0, 0,
                  Kd, Kd, N, 1, 1.0, _lib.stream_ptr())
    for _ in range(2): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): run()
    e1.record(); torch.cuda.synchronize(); lib.as_debug_gemm_mode(0)
    return e0.elapsed_time(e1) / iters
for masked in (False, True):
    print('dgrad 532480x1024x256 B k-major, masked=%s' % masked, ' full %.3f  no-store %.3f  no-fence %.3f  drain-only %.3f ms' % tuple(tk(532480, 1024, 256, m, masked) for m in (0, 1, 3, 2)), flush=True)
