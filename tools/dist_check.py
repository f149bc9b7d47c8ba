"""Multi-GPU gradient check (run under torchrun, one rank per GPU): the bucketed, hook-driven all-reduce of
learner.DistModule (gradient packing + NCCL AVG, overlapped with backward) must leave in the flat gradient buffer exactly
the mean over ranks of the gradients a plain autograd backward produces on each rank's own batch.
usage: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dist_check.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from applestar_b200 import Model, ReinforcementLoss, synth  # noqa: E402
from applestar_b200.learner import DistModule  # noqa: E402

rank, local = int(os.environ['RANK']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl')
dev = torch.device('cuda', local)
sd = synth.full_state_dict(bench.load_spec(), 0)
data = synth.tree_map(lambda t: t.to(dev), synth.gen_trajectories(4, 2, 64, 100 + rank))     # a different batch per rank


def build():
    m = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
    m.load_state_dict(sd, strict=False)
    return m.to(dev)


loss = ReinforcementLoss({}, 'MP0')
# reference: plain autograd on this rank, then one manual all-reduce of the flattened gradients
plain = build()
loss.compute_loss(plain.rl_learner_forward(**data))['total_loss'].backward()
named = {k: p for k, p in plain.named_parameters() if p.requires_grad}
# DistModule path
dm = DistModule(build(), bucket_mb=4.0)
dm.flat.zero_grad()
dm.prepare_backward()
loss.compute_loss(dm.module.rl_learner_forward(**data))['total_loss'].backward()
dm.sync_gradients()
torch.cuda.synchronize()
worst, n_buckets = 0.0, len(dm._buckets)
for name, p, o in zip(dm.flat.names, dm.flat.params, dm.flat.offsets):
    g = named[name].grad
    ref = torch.zeros_like(p) if g is None else g.detach().float().clone()
    dist.all_reduce(ref, op=dist.ReduceOp.AVG)
    got = dm.flat.grad[o:o + p.numel()].view(p.shape)
    err = (got - ref).abs().max().item() / (ref.abs().max().item() + 1e-12)
    worst = max(worst, err)
    assert p.grad.data_ptr() == got.data_ptr()
print('rank %d: %d parameters in %d buckets, pack_mode=%s, worst relative gradient mismatch %.3g' %
      (rank, len(dm.flat.params), n_buckets, dm.flat.pack_mode, worst))
assert worst < 1e-5, worst
dist.barrier()
dist.destroy_process_group()
