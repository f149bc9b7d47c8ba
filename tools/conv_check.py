"""Diagnostics: slab-mode convolution vs per-tap mode (same kernel, debug switch) and vs cuDNN fp32."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from applestar_b200 import _lib, spatial
lib = _lib.lib()
lib.as_debug_gemm_mode.argtypes = [ctypes.c_int]
torch.backends.cudnn.allow_tf32 = False
def run(R, H, W, Cin, Cout, k, mode):
    g = torch.Generator().manual_seed(1)
    xn = torch.randn(R, Cin, H, W, generator=g).cuda()
    x = spatial.to_pnhwc(xn)
    w = (torch.randn(Cout, Cin, k, k, generator=g) * 0.05).cuda()
    b = torch.randn(Cout, generator=g).cuda()
    lib.as_debug_gemm_mode(mode)
    with torch.no_grad():
        y = spatial.conv(x, w, b, relu=False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            spatial.conv(x, w, b, relu=False)
        e1.record(); torch.cuda.synchronize()
    lib.as_debug_gemm_mode(0)
    ref = F.conv2d(xn.to(torch.bfloat16).float(), w.to(torch.bfloat16).float(), b, padding=k // 2)
    got = y[:, 1:-1, 1:-1].permute(0, 3, 1, 2).float()
    ms = e0.elapsed_time(e1) / 5
    return (got - ref).abs().max().item() / ref.abs().max().item(), ms, 2.0 * R * (H + 2) * (W + 2) * Cin * Cout * k * k / ms / 1e9
for shape in [(1040, 19, 20, 128, 128, 3), (1040, 38, 40, 128, 64, 3), (1040, 76, 80, 64, 32, 3), (1040, 76, 80, 32, 64, 3), (3, 19, 20, 192, 128, 3)]:
    for name, mode in (('warm', 0), ('slab', 0), ('per-tap', 16)):
        err, ms, tf = run(*shape, mode)
        print('%-28s %-20s rel.err %.4f  %.3f ms  %.0f TF/s' % (shape, name, err, ms, tf), flush=True)
