"""Runs one tcgen05 GEMM shape a few times (target for `ncu --set full`).  usage: python tools/one_gemm.py M N K [bf16out]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from applestar_b200 import kernels as K  # noqa: E402

M, N, Kd = (int(v) for v in sys.argv[1:4])
out_bf16 = len(sys.argv) > 4 and sys.argv[4] == '1'
A = torch.randn(M, Kd, device='cuda').to(torch.bfloat16)
B = torch.randn(N, Kd, device='cuda').to(torch.bfloat16)
C = torch.empty((M, N), device='cuda', dtype=torch.bfloat16 if out_bf16 else torch.float32)
for _ in range(4):
    K.gemm(A, B, C, M, N, Kd)
torch.cuda.synchronize()
print('ok')
