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


def grad_check(mine, ref, budget, slack=2.0, rel_floor=0.02, cos_floor=0.005):
    """One gradient tensor of a bf16 implementation against the fp32 reference gradient.  `budget` = [relative L2,
    cosine] of the REFERENCE'S OWN bf16-autocast gradient against its fp32 gradient (fixture `bf16_autocast_grad_err`):
    the implementation may be `slack` x as far off, plus small floors.  Tensors whose reference gradient is numerically
    zero (budget relative error > 1: e.g. a bias in front of a softmax) must be numerically zero here too.
    -> None when fine, else a description."""
    a, b = mine.double().reshape(-1), ref.double().reshape(-1)
    rel_b, cos_b = float(budget[0]), float(budget[1])
    if rel_b > 1.0 or b.norm().item() < 1e-8:
        return None if a.norm().item() <= max(1e-4, 100 * b.norm().item()) else 'expected ~0, |g| = %.3e' % a.norm().item()
    rel = ((a - b).norm() / b.norm()).item()
    cos = (torch.dot(a, b) / (a.norm() * b.norm()).clamp(min=1e-30)).item()
    rel_max = slack * rel_b + rel_floor
    cos_min = 1.0 - slack * (1.0 - cos_b) - cos_floor
    if rel <= rel_max and cos >= cos_min:
        return None
    return 'rel-L2 %.3f (max %.3f), cosine %.4f (min %.4f)' % (rel, rel_max, cos, cos_min)
