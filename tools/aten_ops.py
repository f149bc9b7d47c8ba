"""Which library (ATen) kernels still run inside one eager learner step, and from which line of this repo: torch profiler with
Python stacks, CUDA time per (op, input shapes, innermost applestar_b200 frame when the profiler has one).  python tools/aten_ops.py > gpurun_out/aten_ops.txt"""
import collections
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
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True, record_shapes=True) as prof:
    step(data)
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
for ev in prof.events():
    if ev.device_type != torch.autograd.DeviceType.CPU or not ev.name.startswith('aten::'):
        continue
    cuda_us = sum(k.duration for k in ev.kernels) if ev.kernels else 0.0
    if cuda_us <= 0:
        continue
    frame = next((s for s in (ev.stack or []) if 'applestar_b200' in s or 'bench.py' in s), 'autograd engine / other')
    shapes = str([tuple(s) for s in (ev.input_shapes or []) if s])[:70]
    key = (ev.name, (frame.strip()[-60:] if frame != 'autograd engine / other' else '') + ' ' + shapes)
    agg[key][0] += 1
    agg[key][1] += cuda_us
tot = sum(v[1] for v in agg.values())
print('library ops with GPU time inside one eager step: %.1f us total' % tot)
for (name, frame), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:60]:
    print('%8.1f us %4d  %-28s %s' % (us, n, name, frame))
