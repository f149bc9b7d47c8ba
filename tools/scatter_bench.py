"""Stand-alone ScatterConnection kernel: achieved HBM GB/s against its algorithmic bytes (3.179 MB per observation)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import ops, synth
R = int(sys.argv[1]) if len(sys.argv) > 1 else 1040
obs = synth.gen_observations(R, 512, 3, hidden=False)
x, y, en = (obs['entity_info']['x'].cuda(), obs['entity_info']['y'].cuda(), obs['entity_num'].cuda())
proj = torch.randn(R, 512, 32, device='cuda')
for _ in range(3):
    out = ops.scatter_connection(proj, x, y, en, 152, 160)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    out = ops.scatter_connection(proj, x, y, en, 152, 160)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
nbytes = R * (512 * 32 * 4 + 512 * 2 + 32 * 152 * 160 * 4)
print('scatter_connection R=%d: %.3f ms, %.0f GB/s' % (R, ms, nbytes / ms / 1e6))
