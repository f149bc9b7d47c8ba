"""One eager (no CUDA graph) learner step at the benchmark size -- target for
  ncu --kernel-name-base demangled -k regex:as:: --metrics <list> --clock-control none python tools/ncu_step.py
Warm-up step first (allocator, weight caches), then one step between cudaProfilerStart/Stop
(use ncu --profile-from-start off)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from applestar_b200 import Model, ReinforcementLoss, synth  # noqa: E402
from applestar_b200.learner import LearnerStep  # noqa: E402

T = int(os.environ.get('T', 64)); B = int(os.environ.get('B', 16))
dev = torch.device('cuda:0')
model = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
model.load_state_dict(synth.full_state_dict(bench.load_spec(), 0), strict=False)
model = model.to(dev)
step = LearnerStep(model, ReinforcementLoss({}, 'MP0'), use_graph=False)
data = synth.tree_map(lambda t: t.to(dev), synth.gen_trajectories(T, B, 512, 1234))
step(data)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
step(data)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print('ok')
