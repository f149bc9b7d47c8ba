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
