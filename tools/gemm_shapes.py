"""Diagnostics: plain tcgen05 GEMM timings for a list of shapes M,N,K[,bf16out]."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import kernels as K
for spec in sys.argv[1:]:
    v = [int(x) for x in spec.split(',')]
    M, N, Kd = v[:3]; bf = len(v) > 3 and v[3] == 1
    A = torch.randn(M, Kd, device='cuda').to(torch.bfloat16); B = torch.randn(N, Kd, device='cuda').to(torch.bfloat16)
    C = torch.empty((M, N), device='cuda', dtype=torch.bfloat16 if bf else torch.float32)
    for _ in range(3): K.gemm(A, B, C, M, N, Kd)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): K.gemm(A, B, C, M, N, Kd)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print('%s  %.3f ms  %.0f TF/s' % (spec, ms, 2.0 * M * N * Kd / ms / 1e9), flush=True)
