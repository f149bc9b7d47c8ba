import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from applestar_b200 import spatial, synth
R = 256
obs = synth.gen_observations(R, 512, 3)
sp = {k: v.cuda() for k, v in obs['spatial_info'].items()}
proj = torch.relu(torch.randn(R, 512, 32)).cuda()
x = obs['entity_info']['x'].cuda(); y = obs['entity_info']['y'].cuda(); en = obs['entity_num'].cuda()
for _ in range(3):
    f, H, W = spatial._build_features(proj, x, y, en, sp)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    f, H, W = spatial._build_features(proj, x, y, en, sp)
e1.record(); torch.cuda.synchronize()
print('features R=%d: %.3f ms, %.0f GB/s written' % (R, e0.elapsed_time(e1) / 10, f.numel() * 2 / (e0.elapsed_time(e1) / 10) / 1e6))
