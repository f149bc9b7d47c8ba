"""Diagnostics: bilinear x2 upsample forward / backward kernels at the step's three sizes."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import _lib
def t(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
st = _lib.stream_ptr()
for (R, H, W, C) in [(1024, 19, 20, 128), (1024, 38, 40, 64), (1024, 76, 80, 32)]:
    x = torch.randn(R, H + 2, W + 2, C, device='cuda').to(torch.bfloat16)
    y = torch.empty(R, 2 * H + 2, 2 * W + 2, C, device='cuda', dtype=torch.bfloat16)
    g = torch.randn_like(y); dx = torch.empty_like(x)
    f = t(lambda: _lib.call('as_upsample2_fwd_bf16', x.data_ptr(), y.data_ptr(), R, H, W, C, st))
    b = t(lambda: _lib.call('as_upsample2_bwd_bf16', g.data_ptr(), dx.data_ptr(), R, H, W, C, st))
    gb = (x.numel() + y.numel()) * 2 / 1e9
    print((R, H, W, C), 'fwd %.3f ms (%.0f GB/s)  bwd %.3f ms (%.0f GB/s)' % (f, gb / f * 1e3, b, gb / b * 1e3), flush=True)
