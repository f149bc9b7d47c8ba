"""One Transformer-shaped GEMM launch for an ncu capture: python tools/gemm_one.py [fwd|bits|bitmasked] [M N K]
  ncu --set full --import-source on --clock-control none -k regex:gemm_kernel -s 2 -c 1 -o gpurun_out/gemm_one python tools/gemm_one.py fwd"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import _lib, kernels as K

kind = sys.argv[1] if len(sys.argv) > 1 else 'fwd'
M, N, Kd = (int(v) for v in sys.argv[2:5]) if len(sys.argv) > 4 else (532480, 1024, 256)
dev = torch.device('cuda:0')
A = torch.randn(M, Kd, device=dev).to(torch.bfloat16)
C = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
bias = torch.randn(N, device=dev)
bits = torch.full((M, N // 32), 0x55555555, dtype=torch.int32, device=dev)
for _ in range(3):
    if kind == 'fwd':
        B = torch.randn(N, Kd, device=dev).to(torch.bfloat16)
        K.gemm(A, B, C, M, N, Kd, bias=bias, relu=True)
    elif kind == 'bits':
        B = torch.randn(N, Kd, device=dev).to(torch.bfloat16)
        _lib.call('as_gemm_bf16_relu_bits', A.data_ptr(), B.data_ptr(), C.data_ptr(), bias.data_ptr(), bits.data_ptr(), M, N, Kd, Kd, Kd, N,
                  _lib.stream_ptr())
    else:
        B = torch.randn(Kd, N, device=dev).to(torch.bfloat16)
        _lib.call('as_gemm_bf16_bitmasked', A.data_ptr(), B.data_ptr(), C.data_ptr(), bits.data_ptr(), M, N, Kd, 0, 1, Kd, N, N, 1, 1.0,
                  _lib.stream_ptr())
torch.cuda.synchronize()
print('ok')
