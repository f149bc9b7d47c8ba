"""Fused location-head tail (csrc/upconv.cu) at the learner's row count."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import spatial
R = 1024
x = torch.relu(torch.randn(R, 78, 82, 32, device='cuda')).to(torch.bfloat16).requires_grad_(True)
w = (torch.randn(1, 32, 3, 3, device='cuda') * 0.2).requires_grad_(True)
b = torch.zeros(1, device='cuda', requires_grad=True)
out = spatial.upconv_c32_to_1(x, w, b, 1.0, True)
go = torch.randn_like(out)
for _ in range(3):
    out = spatial.upconv_c32_to_1(x, w, b, 1.0, True)
    out.backward(go)
torch.cuda.synchronize()
print('ok')
