"""Fused spatial front (csrc/spatial_front.cu) at the learner's row count: forward / backward kernel times."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import spatial, synth
R = int(sys.argv[1]) if len(sys.argv) > 1 else 1040
obs = synth.tree_map(lambda t: t.cuda(), synth.gen_observations(R, 512, 3, hidden=False))
proj = torch.relu(torch.randn(R, 512, 32, device='cuda')).requires_grad_(True)
w = (torch.randn(32, 56, 1, 1, device='cuda') * 0.2).requires_grad_(True)
b = torch.zeros(32, device='cuda', requires_grad=True)
def fwd():
    return spatial.spatial_front(proj, obs['entity_info']['x'], obs['entity_info']['y'], obs['entity_num'], obs['spatial_info'], w, b)
out = fwd()
gy = torch.randn_like(out)
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
with torch.no_grad():
    print('front forward  R=%d: %.3f ms' % (R, t(fwd)))
def fb():
    o = fwd()
    o.backward(gy)
print('front fwd+bwd  R=%d: %.3f ms' % (R, t(fb)))
