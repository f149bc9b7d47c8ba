"""Runs one implicit-GEMM convolution shape a few times (target for `ncu --set full`).
usage: python tools/one_conv.py R H W Cin Cout k"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from applestar_b200 import spatial  # noqa: E402

R, H, W, Cin, Cout, k = (int(v) for v in sys.argv[1:7])
x = torch.randn(R, H + 2, W + 2, Cin, device='cuda').to(torch.bfloat16)
x[:, 0] = 0; x[:, -1] = 0; x[:, :, 0] = 0; x[:, :, -1] = 0
w = torch.randn(Cout, Cin, k, k, device='cuda') * 0.05
b = torch.randn(Cout, device='cuda')
with torch.no_grad():
    for _ in range(3):
        y = spatial.conv(x, w, b, relu=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        y = spatial.conv(x, w, b, relu=True)
    e1.record()
    torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print('conv %s: %.3f ms  %.0f TF/s' % (sys.argv[1:7], ms, 2.0 * R * (H + 2) * (W + 2) * Cin * Cout * k * k / ms / 1e9))
