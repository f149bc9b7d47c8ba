"""Shared helpers for the parity tests."""
import json
import os

import torch

from applestar_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load_spec(name='winloss'):
    with open(os.path.join(GOLD, 'state_dict_spec.json')) as f:
        return [(k, tuple(s), t) for k, s, t in json.load(f)[name]]


def golden(name):
    return torch.load(os.path.join(GOLD, name), map_location='cpu', weights_only=False)


def valid_logits(a, b):
    """Masked logits (-1e9 fill) are compared as a predicate, never numerically."""
    ma, mb = a <= -1e8, b <= -1e8
    assert torch.equal(ma, mb), 'mask predicate differs'
    return a[~ma], b[~mb]


def max_abs(a, b):
    return (a.double() - b.double()).abs().max().item() if a.numel() else 0.0


def clone_tree(x):
    return synth.tree_map(lambda t: t.clone(), x)


def grad_sample(g, cap=65536):
    """The stored sample of a gradient tensor (oracle/make_golden.py::grad_sample_index): everything up to `cap`
    elements, else a fixed odd stride."""
    g = g.detach().float().reshape(-1).cpu()
    n = g.numel()
    if n <= cap:
        return g
    stride = (n // cap) | 1
    return g[torch.arange(0, n, stride)[:cap]]


def grad_budgets(fixture):
    """-> (per-tensor budgets, typical relative error): the reference's own bf16-autocast gradient error per tensor
    [relative L2, cosine] and the median relative error over all tensors."""
    gerr = fixture['bf16_autocast_grad_err']
    rels = sorted(float(v[0]) for v in gerr.values() if float(v[0]) < 1.0)
    return gerr, rels[len(rels) // 2]


def grad_check(mine, ref, budget, typical, slack=3.0):
    """One gradient tensor of a bf16 implementation against the fp32 reference gradient.  `budget` = [relative L2,
    cosine] of the REFERENCE'S OWN bf16-autocast gradient against its fp32 gradient for this tensor, `typical` = the
    median of that relative error over all tensors of the model (grad_budgets).  Pass:
        relative L2 <= max(slack x budget, typical)  and  cosine >= min(1 - slack^2 x (1 - budget cosine), 0.99)
    (1 - cosine is quadratic in the error, hence slack^2; where the reference's own bf16 gradient is more than 30 % off
    -- sums that cancel, e.g. the two-element bias of the queued head -- the direction carries no information and
    only the distance is checked).  Per tensor the slack is 3: the network is chaotic at random
    init (unscaled randn LSTM weights, few trajectories), a single tensor of the reference's own bf16 run is a noisy
    estimate of its error level; the 2x rule of SURVEY §8(c) is enforced on module-level aggregates (grad_group_check).
    Tensors whose reference gradient is numerically zero (budget relative error > 1: e.g. a bias in front of a
    softmax) must be numerically zero here too.  -> None when fine, else a description."""
    a, b = mine.double().reshape(-1), ref.double().reshape(-1)
    rel_b, cos_b = float(budget[0]), float(budget[1])
    if rel_b > 1.0 or b.norm().item() < 1e-8:
        return None if a.norm().item() <= max(1e-4, 100 * b.norm().item()) else 'expected ~0, |g| = %.3e' % a.norm().item()
    rel = ((a - b).norm() / b.norm()).item()
    cos = (torch.dot(a, b) / (a.norm() * b.norm()).clamp(min=1e-30)).item()
    rel_max = max(slack * rel_b, typical)
    cos_min = -1.0 if rel_b > 0.3 else min(1.0 - slack * slack * (1.0 - cos_b), 0.99)
    if rel <= rel_max and cos >= cos_min:
        return None
    return 'rel-L2 %.3f (max %.3f), cosine %.4f (min %.4f)' % (rel, rel_max, cos, cos_min)


GRAD_GROUPS = ('encoder.entity_encoder', 'encoder.spatial_encoder', 'encoder.scalar_encoder', 'encoder.scatter_project',
               'core_lstm', 'policy.action_type_head', 'policy.delay_head', 'policy.queued_head',
               'policy.selected_units_head', 'policy.target_unit_head', 'policy.location_head', 'value_networks')


def grad_group_check(mine, ref, gerr, typical, slack=2.0):
    """Module-level aggregate: the concatenated gradient of every tensor of a module group (mine / ref: dicts of flat
    tensors) must be within `slack` x the reference's own bf16-autocast error of the same aggregate, which follows
    from the per-tensor budgets: err_group^2 = sum_i (rel_i |g_i|)^2 -- or within half the reference's typical
    per-tensor error where its aggregate happens to be very accurate.  -> (failures, table)."""
    bad, table = [], []
    for grp in GRAD_GROUPS:
        num = den = bud = 0.0
        for k, r in ref.items():
            if not k.startswith(grp) or float(gerr[k][0]) > 1.0:
                continue
            a, b = mine[k].double().reshape(-1), r.double().reshape(-1)
            num += float((a - b).pow(2).sum())
            den += float(b.pow(2).sum())
            bud += (float(gerr[k][0]) ** 2) * float(b.pow(2).sum())
        if den == 0.0:
            continue
        rel, rel_b = (num / den) ** 0.5, (bud / den) ** 0.5
        table.append((grp, rel, rel_b))
        if rel > max(slack * rel_b, 0.5 * typical):
            bad.append((grp, 'aggregate rel-L2 %.4f, reference bf16 %.4f' % (rel, rel_b)))
    return bad, table
