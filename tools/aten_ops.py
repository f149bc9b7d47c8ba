"""Which ATen operators still launch kernels inside one learner step (eager), by device time: torch.profiler over one
step at the benchmark size.  usage: python tools/aten_ops.py [rows]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import bench  # noqa: E402
from applestar_b200 import Model, ReinforcementLoss, synth  # noqa: E402
from applestar_b200.learner import LearnerStep  # noqa: E402

dev = torch.device('cuda:0')
model = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
model.load_state_dict(synth.full_state_dict(bench.load_spec(), 0), strict=False)
model = model.to(dev)
step = LearnerStep(model, ReinforcementLoss({}, 'MP0'), use_graph=False)
data = synth.tree_map(lambda t: t.to(dev), synth.gen_trajectories(64, 16, 512, 1234))
step(data)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
    step(data)
    torch.cuda.synchronize()
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 45
print(prof.key_averages().table(sort_by='self_cuda_time_total', row_limit=rows, max_name_column_width=60))
print(prof.key_averages(group_by_stack_n=4).table(sort_by='self_cuda_time_total', row_limit=30, max_name_column_width=50,
                                                  max_src_column_width=90))
