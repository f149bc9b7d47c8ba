"""Times the Transformer-shaped GEMMs (K = 256 epilogue-bound shapes, masked / sign-bit variants, 8192^3, one slab conv) stand-alone."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import _lib, kernels as K
which = sys.argv[1] if len(sys.argv) > 1 else 'new'
dev = torch.device('cuda:0')
M = 532480
def timeit(fn, iters=6):
    for _ in range(2): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
A256 = torch.randn(M, 256, device=dev).to(torch.bfloat16)
A1024 = torch.randn(M, 1024, device=dev).to(torch.bfloat16)
bias = torch.randn(1024, device=dev)
out = []
for N in (768, 1024):
    B = torch.randn(N, 256, device=dev).to(torch.bfloat16); C = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    out.append('fwd N=%d K=256 bias relu: %.3f' % (N, timeit(lambda: K.gemm(A256, B, C, M, N, 256, bias=bias[:N], relu=True))))
B = torch.randn(256, 1024, device=dev).to(torch.bfloat16); C32 = torch.empty(M, 256, dtype=torch.float32, device=dev); C16 = torch.empty(M, 256, dtype=torch.bfloat16, device=dev)
out.append('fwd N=256 K=1024 f32+bf16 out: %.3f' % timeit(lambda: K.gemm(A1024, B, C32, M, 256, 1024, bias=bias[:256], relu=True, C2=C16)))
Bm = torch.randn(256, 1024, device=dev).to(torch.bfloat16); C = torch.empty(M, 1024, dtype=torch.bfloat16, device=dev)
H = torch.randn(M, 1024, device=dev).to(torch.bfloat16)
def masked(mask):
    _lib.call('as_gemm_bf16_masked', A256.data_ptr(), Bm.data_ptr(), C.data_ptr(), mask.data_ptr() if mask is not None else None, M, 1024, 256,
              0, 1, 256, 1024, 1024, 1, 1.0, _lib.stream_ptr())
out.append('dgrad N=1024 K=256 mn-major plain: %.3f  bf16-masked: %.3f' % (timeit(lambda: masked(None)), timeit(lambda: masked(H))))
if which == 'new':
    bits = torch.full((M, 32), 0x55555555, dtype=torch.int32, device=dev)
    Bk = torch.randn(1024, 256, device=dev).to(torch.bfloat16)
    t1 = timeit(lambda: _lib.call('as_gemm_bf16_relu_bits', A256.data_ptr(), Bk.data_ptr(), C.data_ptr(), bias.data_ptr(), bits.data_ptr(), M, 1024, 256, 256, 256, 1024, _lib.stream_ptr()))
    t2 = timeit(lambda: _lib.call('as_gemm_bf16_bitmasked', A256.data_ptr(), Bm.data_ptr(), C.data_ptr(), bits.data_ptr(), M, 1024, 256, 0, 1, 256, 1024, 1024, 1, 1.0, _lib.stream_ptr()))
    out.append('relu_bits fwd: %.3f  bitmasked dgrad: %.3f' % (t1, t2))
A = torch.randn(8192, 8192, device=dev).to(torch.bfloat16); Cb = torch.empty(8192, 8192, dtype=torch.bfloat16, device=dev)
out.append('8192^3: %.3f' % timeit(lambda: K.gemm(A, A, Cb, 8192, 8192, 8192)))
# a 19x20 slab conv (border mask path) through the spatial module
from applestar_b200 import spatial
x = torch.randn(1040, 21, 22, 128, device=dev).to(torch.bfloat16)
w = torch.randn(128, 128, 3, 3, device=dev) * 0.05
try:
    y = torch.empty_like(x)
    wl = spatial._conv_weights(w, 128)[0]
    out.append('conv 19x20 128->128 3x3: %.3f' % timeit(lambda: _lib.call('as_conv_nhwc_bf16', x.data_ptr(), wl.data_ptr(), y.data_ptr(), None, 1040, 19, 20, 128, 128, 9, 0, 1, 1, _lib.stream_ptr())))
except Exception as exc:
    out.append('conv: skipped (%r)' % (exc,))
print('\n'.join(out))
