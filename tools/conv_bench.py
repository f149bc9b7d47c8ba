"""Diagnostics: implicit-GEMM convolution modes on the step's shapes (same tensors, cached bf16 weights)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import _lib, spatial
lib = _lib.lib()
lib.as_debug_gemm_mode.argtypes = [ctypes.c_int]
def bench(x, w, b, iters=10):
    with torch.no_grad():
        for _ in range(3):
            spatial.conv(x, w, b, relu=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            spatial.conv(x, w, b, relu=True)
        e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
variants = [('slab', 0, {}), ('slab no-multicast', 0, {'AS_SLAB_NO_MC': '1'}),
            ('per-tap', 16, {})]
for (R, H, W, Cin, Cout, k) in [(1040, 19, 20, 128, 128, 3), (1024, 38, 40, 128, 64, 3), (1024, 38, 40, 64, 128, 3), (1040, 76, 80, 64, 32, 3), (1040, 76, 80, 32, 64, 3)]:
    x = torch.randn(R, H + 2, W + 2, Cin, device='cuda').to(torch.bfloat16)
    x[:, 0] = 0; x[:, -1] = 0; x[:, :, 0] = 0; x[:, :, -1] = 0
    w = torch.nn.Parameter(torch.randn(Cout, Cin, k, k, device='cuda') * 0.05)
    b = torch.nn.Parameter(torch.randn(Cout, device='cuda'))
    fl = 2.0 * R * (H + 2) * (W + 2) * Cin * Cout * k * k
    out = []
    for name, mode, env in variants:
        for kk in ('AS_SLAB_NO_RESIDENT', 'AS_SLAB_BSLOTS', 'AS_SLAB_NO_MC'):
            os.environ.pop(kk, None)
        os.environ.update(env)
        lib.as_debug_gemm_mode(mode)
        ms = bench(x, w, b)
        out.append('%s %.3f ms (%.0f TF/s)' % (name, ms, fl / ms / 1e9))
    lib.as_debug_gemm_mode(0)
    print((R, H, W, Cin, Cout), ' | '.join(out), flush=True)
