"""CPU suite, world_size 2 over gloo: the data-parallel plumbing of applestar_b200/learner.py (flat parameter
buffer, one-shot broadcast, bucketed gradient all-reduce launched from autograd hooks, sharding of trajectories).
The kernels themselves are GPU-only; this covers the N > 1 host logic the reference keeps in
ctools/utils/dist_helper.py:372-439."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


class Tiny(nn.Module):
    def __init__(self):
        super().__init__()
        self.encoder = nn.Sequential(nn.Linear(12, 33), nn.ReLU(), nn.Linear(33, 17))
        self.policy = nn.Linear(17, 5)
        self.value_networks = nn.Linear(17, 1)
        self.unused = nn.Linear(3, 3)            # never receives a gradient: sync must still complete

    def forward(self, x):
        h = torch.relu(self.encoder(x))
        return self.policy(h).logsumexp(-1).mean() + self.value_networks(h).pow(2).mean()


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from applestar_b200.learner import DistModule
    torch.manual_seed(100 + rank)                # different init per rank: broadcast must make them equal
    model = Tiny()
    dm = DistModule(model, bucket_mb=0.0002)     # ~52 floats per bucket -> several buckets
    assert len(dm._buckets) >= 3
    gathered = [torch.zeros_like(dm.flat.flat) for _ in range(world)]
    dist.all_gather(gathered, dm.flat.flat)
    assert torch.equal(gathered[0], gathered[1])
    # each rank gets its own shard of a global batch of 8 "trajectories"
    g = torch.Generator().manual_seed(7)
    # the reference learner's sequence, nothing else (rl_learner.py:114-132): zero_grad / backward / sync_gradients
    # / optimizer.step, for THREE steps with a plain torch optimizer -- from the second step on the buckets must
    # have re-armed themselves
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    ref = Tiny()
    ref.load_state_dict(model.state_dict())
    ref_opt = torch.optim.SGD(ref.parameters(), lr=0.1)
    worst = 0.0
    for step in range(3):
        full = torch.randn(8, 12, generator=g)
        shard = full[rank * 4:(rank + 1) * 4]
        opt.zero_grad()
        model(shard).backward()
        stats = torch.full((6,), float(rank + 1 + step))          # the step's packed log scalars: one collective, rank mean
        dm.reduce_stats(stats)
        dm.sync_gradients()
        assert torch.allclose(stats, torch.full((6,), 1.5 + step)), stats
        # single-process gradient of the mean over both shards
        ref_opt.zero_grad()
        (0.5 * (ref(full[:4]) + ref(full[4:]))).backward()
        for (k, p), (_, pr) in zip(model.named_parameters(), ref.named_parameters()):
            gr = pr.grad if pr.grad is not None else torch.zeros_like(pr)
            mine = p.grad if p.grad is not None else torch.zeros_like(p)
            worst = max(worst, (mine - gr).abs().max().item())
        opt.step()
        ref_opt.step()
    for p, pr in zip(model.parameters(), ref.parameters()):
        worst = max(worst, (p - pr).abs().max().item())
    q.put((rank, worst))
    dist.destroy_process_group()


def test_bucketed_allreduce_world2_gloo():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, worst in res:
        assert worst < 1e-6, (rank, worst)


def test_flat_params_views_and_buckets():
    from applestar_b200.learner import FlatParams
    m = Tiny()
    before = {k: v.clone() for k, v in m.state_dict().items()}
    fp = FlatParams(m)
    for k, v in m.state_dict().items():
        assert torch.equal(v, before[k])
    for p, o in zip(fp.params, fp.offsets):
        assert o % FlatParams.ALIGN == 0
        assert p.data_ptr() == fp.flat.data_ptr() + 4 * o
    b = fp.buckets(100)
    assert b[0][1] == fp.numel and b[-1][0] == 0
    assert sorted(i for _, _, idx in b for i in idx) == list(range(len(fp.params)))
    for (s0, e0, _), (s1, e1, _) in zip(b, b[1:]):
        assert e1 == s0
