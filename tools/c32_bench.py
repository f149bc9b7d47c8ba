"""Diagnostics: final 3x3 32->1 convolution kernels (location logits) at the step's size."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import spatial
R, H, W = int(os.environ.get('R', 1024)), 152, 160
x = (torch.randn(R, H + 2, W + 2, 32, device='cuda') * 0.5).to(torch.bfloat16).requires_grad_(True)
w = (torch.randn(1, 32, 3, 3, device='cuda') * 0.1).requires_grad_(True)
b = torch.randn(1, device='cuda').requires_grad_(True)
g = torch.randn(R, H * W, device='cuda')
def step():
    y = spatial.conv_c32_to_1(x, w, b, 1.25)
    y.backward(g)
for _ in range(2): step()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(); torch.cuda.synchronize()
for e in prof.key_averages():
    if 'c32' in e.key: print('%-50s %8.1f us' % (e.key[:50], e.device_time_total))
