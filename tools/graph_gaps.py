"""GPU idle time inside one CUDA-graph replay of the learner step: kernel intervals from the profiler, union of busy time,
largest gaps and what sits around them.  python tools/graph_gaps.py"""
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
step = LearnerStep(model, ReinforcementLoss({}, 'MP0'))
data = synth.tree_map(lambda t: t.to(dev), synth.gen_trajectories(64, 16, 512, 1234))
step(data)
static = step.static_inputs(data)
for _ in range(3):
    step(static)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(static)
    torch.cuda.synchronize()
ks = sorted(((e.time_range.start, e.time_range.end, e.name) for e in prof.events()
             if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start), key=lambda t: t[0])
ks = [k for k in ks if 'Memcpy' not in k[2] and 'Memset' not in k[2]] or ks
t0, t1 = ks[0][0], max(k[1] for k in ks)
busy, cur_s, cur_e = 0.0, ks[0][0], ks[0][1]
gaps = []
last_name = ks[0][2]
for s, e, n in ks[1:]:
    if s > cur_e:
        busy += cur_e - cur_s
        gaps.append((s - cur_e, last_name, n))
        cur_s, cur_e = s, e
    else:
        cur_e = max(cur_e, e)
    if e >= cur_e:
        last_name = n
busy += cur_e - cur_s
print('kernels %d  span %.2f ms  busy (union) %.2f ms  idle %.2f ms in %d gaps' % (len(ks), (t1 - t0) / 1e3, busy / 1e3, (t1 - t0 - busy) / 1e3, len(gaps)))
hist = [0, 0, 0, 0]
for g, _, _ in gaps:
    hist[0 if g < 2 else 1 if g < 5 else 2 if g < 20 else 3] += 1
print('gaps < 2 us: %d, 2-5 us: %d, 5-20 us: %d, > 20 us: %d' % tuple(hist))
tot_by_after = {}
for g, a, b in gaps:
    key = (a[:40], b[:40])
    tot_by_after[key] = tot_by_after.get(key, 0.0) + g
for (a, b), g in sorted(tot_by_after.items(), key=lambda kv: -kv[1])[:25]:
    print('%8.1f us  %-42s -> %s' % (g, a, b))
import collections
import re
agg = collections.defaultdict(lambda: [0, 0.0])
for s, e, n in ks:
    n = re.sub(r'\(.*', '', n)
    n = re.sub(r'^void ', '', n)
    agg[n[:70]][0] += 1
    agg[n[:70]][1] += e - s
tot = sum(v[1] for v in agg.values())
print('\nkernel time inside the replay (warm, concurrent streams): %.2f ms summed' % (tot / 1e3))
for n, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print('%9.1f us %5.1f%% %5d  %s' % (us, 100 * us / tot, c, n))
lib = sum(v[1] for n, v in agg.items() if not n.startswith('as::'))
print('library kernels: %.2f ms in %d launches' % (lib / 1e3, sum(v[0] for n, v in agg.items() if not n.startswith('as::'))))
# context of the largest gap: the kernels (all streams) that end / start around it
big = max(gaps, key=lambda g: g[0])
allk = sorted(((e.time_range.start, e.time_range.end, e.name, getattr(e, 'device_index', 0)) for e in prof.events()
               if e.device_type == torch.autograd.DeviceType.CUDA), key=lambda t: t[0])
edge = None
cur_e = allk[0][1]
for s, e, n, _ in allk[1:]:
    if s - cur_e >= big[0] - 1:
        edge = (cur_e, s)
        break
    cur_e = max(cur_e, e)
if edge:
    print('\nlargest gap: %.1f us, %.2f ms after the first kernel' % (edge[1] - edge[0], (edge[0] - allk[0][0]) / 1e3))
    before = [k for k in allk if k[1] <= edge[0]][-6:]
    after = [k for k in allk if k[0] >= edge[1]][:6]
    for s, e, n, _ in before:
        print('  before: ends %8.1f us before the gap, %7.1f us long  %s' % (edge[0] - e, e - s, n[:90]))
    for s, e, n, _ in after:
        print('  after : starts %7.1f us after the gap, %7.1f us long  %s' % (s - edge[1], e - s, n[:90]))
    inside = [k for k in sorted(((e.time_range.start, e.time_range.end, e.name) for e in prof.events()
                                 if e.device_type == torch.autograd.DeviceType.CUDA), key=lambda t: t[0])
              if k[0] < edge[1] and k[1] > edge[0]]
    for s, e, n in inside[:10]:
        print('  overlapping the gap: %8.1f us .. %8.1f us  %s' % (s - edge[0], e - edge[0], n[:90]))
