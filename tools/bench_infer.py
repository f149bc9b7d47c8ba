"""Batched actor inference (BASELINE.json configs[2]): Model.compute_logp_action on B observations x 512 entities,
inputs resident on the device, Gumbel noise supplied.  Prints obs/s.  usage: python tools/bench_infer.py [B] [iters] [graph=1]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from applestar_b200 import Model, synth  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
use_graph = (int(sys.argv[3]) if len(sys.argv) > 3 else 1) != 0
dev = torch.device('cuda:0')
model = Model({}, use_value_network=False)
model.load_state_dict(synth.full_state_dict(bench.load_spec(), 0), strict=False)
model = model.to(dev).eval()
obs = synth.tree_map(lambda t: t.to(dev), synth.gen_observations(B, 512, 7))
hidden = [(torch.zeros(B, 384, device=dev), torch.zeros(B, 384, device=dev)) for _ in range(3)]
noise = synth.tree_map(lambda t: t.to(dev), synth.gen_gumbel(B, 11))
from applestar_b200.learner import ActorStep  # noqa: E402
actor = ActorStep(model, use_graph=use_graph)


def call():
    return actor(obs['spatial_info'], obs['entity_info'], obs['scalar_info'], obs['entity_num'], hidden, gumbel_noise=noise)


with torch.no_grad():
    for _ in range(2):
        out = call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()                       # `ncu --profile-from-start off` captures only the timed calls
    e0.record()
    for _ in range(iters):
        out = call()
    e1.record()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
ms = e0.elapsed_time(e1) / iters
print(json.dumps({'metric': 'inference_obs_per_sec_per_gpu', 'value': B / ms * 1e3, 'unit': 'obs/s', 'ms_per_call': ms, 'batch': B,
                  'entities': 512, 'cuda_graph': use_graph, 'launches_per_call': actor.launches_per_call, 'dtype': 'bf16', 'data': 'synthetic', 'selected_units_steps': int(out['logit']['selected_units'].shape[1])}))
