"""Per-kernel time table of one learner step (torch.profiler / CUPTI) -- run on the GPU box.
usage: python tools/profile_step.py [--traj 16 --unroll 64] > gpurun_out/step_profile.txt"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import profile, ProfilerActivity  # noqa: E402

import bench  # noqa: E402
from applestar_b200 import Model, ReinforcementLoss, synth  # noqa: E402
from applestar_b200.learner import LearnerStep  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--traj', type=int, default=16)
ap.add_argument('--unroll', type=int, default=64)
ap.add_argument('--rows', type=int, default=60)
a = ap.parse_args()
dev = torch.device('cuda:0')
model = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
model.load_state_dict(synth.full_state_dict(bench.load_spec(), 0), strict=False)
model = model.to(dev)
step = LearnerStep(model, ReinforcementLoss({}, 'MP0'))
data = synth.tree_map(lambda t: t.to(dev), synth.gen_trajectories(a.unroll, a.traj, 512, 1234))
for _ in range(3):
    step(data)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step(data)
    torch.cuda.synchronize()
ka = prof.key_averages()
n_kernels = sum(e.count for e in ka if e.device_type == torch.autograd.DeviceType.CUDA)
print('CUDA kernel launches in one step: %d' % n_kernels)
print(ka.table(sort_by='cuda_time_total', row_limit=a.rows, max_name_column_width=90))
print(ka.table(sort_by='self_cpu_time_total', row_limit=40, max_name_column_width=70))
if os.environ.get('EACH'):
    pats = os.environ['EACH'].split(',')
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA and any(p in e.name for p in pats):
            print('EACH %-60s %10.1f us' % (e.name[:60], e.device_time_total))
