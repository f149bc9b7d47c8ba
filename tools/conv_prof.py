"""Diagnostics: where the slab-mode MMA issuer waits (cycles per CTA)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import _lib, spatial
lib = _lib.lib()
lib.as_debug_gemm_profile.argtypes = [ctypes.c_void_p]
R, H, W, Cin, Cout = [int(v) for v in (sys.argv[1:6] if len(sys.argv) > 5 else (1040, 19, 20, 128, 128))]
x = torch.randn(R, H + 2, W + 2, Cin, device='cuda').to(torch.bfloat16)
w = torch.nn.Parameter(torch.randn(Cout, Cin, 3, 3, device='cuda') * 0.05)
b = torch.nn.Parameter(torch.randn(Cout, device='cuda'))
with torch.no_grad():
    for _ in range(3): spatial.conv(x, w, b, relu=True)
    buf = torch.zeros(148 * 4, dtype=torch.int64, device='cuda')
    lib.as_debug_gemm_profile(ctypes.c_void_p(buf.data_ptr()))
    spatial.conv(x, w, b, relu=True)
    torch.cuda.synchronize()
    lib.as_debug_gemm_profile(ctypes.c_void_p(0))
v = buf.view(148, 4).double().cpu()
v = v[v[:, 0] > 0]
print('CTAs %d: total %.0f cycles | wait accumulators %.0f | wait slab %.0f | wait weights %.0f  (means)' % (len(v), v[:, 0].mean(), v[:, 1].mean(), v[:, 2].mean(), v[:, 3].mean()))
