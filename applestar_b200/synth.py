"""Seeded synthetic observations / trajectories / weights of the named shapes.

Schema = the learner/actor input layout of the reference (field dtypes from
distar/agent/default/lib/features.py:41-82, learner layout from
distar/agent/default/rl_training/rl_dataloader.py:45-76,206-245). Distributions follow SURVEY.md §8(d).
Everything is generated on the CPU with a `torch.Generator`, so the oracle (CPU), the golden
fixtures (container) and the CUDA path (GPU box) see bit-identical inputs.
"""
import math

import torch

from . import config as C
from . import tables

_DT = {'uint8': torch.uint8, 'int8': torch.int8, 'int16': torch.int16, 'float16': torch.float16,
       'float32': torch.float32, 'int64': torch.int64}


def _gen(seed):
    g = torch.Generator(device='cpu')
    g.manual_seed(int(seed))
    return g


def _randint(g, lo, hi, size, dtype=torch.int64):
    return torch.randint(lo, hi, size, generator=g, dtype=torch.int64).to(dtype)


def gen_observations(rows, n_entity=C.MAX_ENTITY_NUM, seed=1234, entity_num_range=None, special=False,
                     hidden=True):
    """`rows` observations: spatial_info / entity_info / scalar_info / entity_num (/ hidden_state)."""
    g = _gen(seed)
    H, W = C.SPATIAL_Y, C.SPATIAL_X
    lo, hi = entity_num_range if entity_num_range is not None else (max(16, n_entity // 2), n_entity)
    entity_num = _randint(g, lo, hi + 1, (rows,))
    if special and rows >= 2:
        entity_num[0] = n_entity
        entity_num[1] = 1
    valid = torch.arange(n_entity).unsqueeze(0) < entity_num.unsqueeze(1)

    entity_info = {}
    for name, dt, arc, num in C.ENTITY_FIELDS:
        if arc == 'unsqueeze':
            v = torch.rand((rows, n_entity), generator=g).to(torch.float16)
        elif name == 'x':
            v = _randint(g, 0, W, (rows, n_entity), _DT[dt])
        elif name == 'y':
            v = _randint(g, 0, H, (rows, n_entity), _DT[dt])
        else:
            v = _randint(g, 0, num, (rows, n_entity), _DT[dt])
        entity_info[name] = v * valid.to(v.dtype)          # padded entities are zero (rl_dataloader.py:213-215)
    if special and rows >= 1:                               # heavy collision: every entity on one cell
        entity_info['x'][0] = 17
        entity_info['y'][0] = 33

    spatial_info = {}
    for name, arc, num in C.SPATIAL_FIELDS:
        if arc == 'other':
            spatial_info[name] = _randint(g, 0, 256, (rows, H, W), torch.uint8)
        elif arc == 'one_hot':
            spatial_info[name] = _randint(g, 0, num, (rows, H, W), torch.uint8)
        else:                                               # effect lists: k ids then zero padding (features.py:486-493)
            k = _randint(g, 0, 11, (rows,))
            ids = _randint(g, 0, H * W, (rows, C.EFFECT_LEN), torch.int16)
            keep = torch.arange(C.EFFECT_LEN).unsqueeze(0) < k.unsqueeze(1)
            spatial_info[name] = ids * keep.to(torch.int16)

    def bern(shape, p=0.1, dtype=torch.uint8):
        return (torch.rand(shape, generator=g) < p).to(dtype)

    scalar_info = {
        'home_race': _randint(g, 0, 5, (rows,), torch.uint8),
        'away_race': _randint(g, 0, 5, (rows,), torch.uint8),
        'upgrades': bern((rows, C.NUM_UPGRADES), dtype=torch.int16),
        'time': torch.rand((rows,), generator=g) * 30000.0,
        'unit_counts_bow': bern((rows, C.NUM_UNIT_TYPES)) * _randint(g, 1, 6, (rows, C.NUM_UNIT_TYPES), torch.uint8),
        'agent_statistics': torch.rand((rows, 10), generator=g) * 8.0,
        'cumulative_stat': bern((rows, C.NUM_CUMULATIVE_STAT_ACTIONS)),
        'beginning_order': _randint(g, 0, C.NUM_BEGINNING_ORDER_ACTIONS, (rows, C.BEGINNING_ORDER_LENGTH), torch.int16),
        'last_queued': _randint(g, 0, 2, (rows,), torch.int16),
        'last_delay': _randint(g, 0, 128, (rows,), torch.int16),
        'last_action_type': _randint(g, 0, C.NUM_ACTIONS, (rows,), torch.int16),
        'bo_location': _randint(g, 0, H * W, (rows, C.BEGINNING_ORDER_LENGTH), torch.int16),
        'unit_order_type': bern((rows, C.NUM_UNIT_MIX_ABILITIES)),
        'unit_type_bool': bern((rows, C.NUM_UNIT_TYPES)),
        'enemy_unit_type_bool': bern((rows, C.NUM_UNIT_TYPES)),
    }
    out = {'spatial_info': spatial_info, 'entity_info': entity_info, 'scalar_info': scalar_info,
           'entity_num': entity_num}
    if hidden:
        out['hidden_state'] = [(torch.randn((rows, 384), generator=g) * 0.1,
                                torch.randn((rows, 384), generator=g) * 0.1) for _ in range(3)]
    return out


def gen_trajectories(unroll_len, batch_size, n_entity=C.MAX_ENTITY_NUM, seed=1234,
                     fields=('winloss', 'build_order', 'built_unit', 'battle'), special=False):
    """Learner batch: `(T+1)*B` flattened observations (row = t*B + b) + the `[T,B,...]` learner-only
    tensors, as `rl_dataloader.collate_fn` would deliver them (padded to `n_entity`).  `special` puts the
    all-entities-on-one-cell observation in row 0 and the `entity_num == 1` observation in row 1."""
    T, B, N = unroll_len, batch_size, n_entity
    S = C.MAX_SELECTED_UNITS_NUM
    data = gen_observations((T + 1) * B, N, seed, hidden=True, special=special)
    g = _gen(seed + 7919)
    entity_num = data['entity_num'][:T * B].view(T, B)

    action_type = _randint(g, 0, C.NUM_ACTIONS, (T, B))
    has = {k: torch.tensor(tables.flags(getattr(tables, 'ACTION_HAS_' + k.upper())))
           for k in ('queued', 'selected_units', 'target_unit', 'target_location')}
    actions_mask = {k: has[k][action_type].long() for k in has}
    sun = _randint(g, 2, 13, (T, B)) * actions_mask['selected_units']         # {0} U [2,12]
    sun = torch.minimum(sun, entity_num + 1)                                  # distinct valid units + the end flag
    # distinct random unit ids < entity_num, last valid slot = entity_num (the end flag)
    selected_units = torch.zeros((T, B, S), dtype=torch.int64)
    scores = torch.rand((T, B, N), generator=g)
    scores = scores.masked_fill(torch.arange(N).view(1, 1, N) >= entity_num.unsqueeze(-1), 2.0)
    order = scores.argsort(dim=-1)[..., :S]
    if N < S:
        order = torch.nn.functional.pad(order, (0, S - N))
    slot = torch.arange(S).view(1, 1, S)
    selected_units = torch.where(slot < (sun - 1).unsqueeze(-1), order, selected_units)
    selected_units = torch.where((slot == (sun - 1).unsqueeze(-1)) & (sun > 0).unsqueeze(-1),
                                 entity_num.unsqueeze(-1).expand(T, B, S), selected_units)
    su_mask = slot < sun.unsqueeze(-1)

    action_info = {
        'action_type': action_type,
        'delay': _randint(g, 0, 128, (T, B)),
        'queued': _randint(g, 0, 2, (T, B)),
        'selected_units': selected_units,
        'target_unit': (torch.rand((T, B), generator=g) * entity_num).long().clamp_(max=N - 1),
        'target_location': _randint(g, 0, C.SPATIAL_Y * C.SPATIAL_X, (T, B)),
    }

    def neg_exp(shape):
        return torch.log(torch.rand(shape, generator=g).clamp_(min=1e-6))     # -Exp(1)

    behaviour_logp = {k: neg_exp((T, B)) for k in C.HEADS if k != 'selected_units'}
    behaviour_logp['selected_units'] = torch.where(su_mask, neg_exp((T, B, S)), torch.full((T, B, S), -1e9))

    su_logits_mask = torch.arange(N + 1).view(1, 1, N + 1) < (entity_num + 1).unsqueeze(-1)
    tu_logits_mask = torch.arange(N).view(1, 1, N) < entity_num.unsqueeze(-1)
    # teacher selected-units logits carry the same evolving mask the policy applies
    # (action_arg_head.py:163,179-187) and -1e9 padding (rl_dataloader.py:229-232)
    idx_tb = torch.arange(T * B)
    en_flat = entity_num.reshape(-1)
    su_flat = selected_units.view(T * B, S)
    cur = su_logits_mask.reshape(T * B, N + 1).clone()
    cur[idx_tb, en_flat] = False                                              # step 0: end flag disabled
    steps = [cur]
    for s in range(1, S):
        cur = cur.clone()
        if s == 1:
            cur[idx_tb, en_flat] = True
        cur[idx_tb, su_flat[:, s - 1]] = False
        steps.append(cur)
    step_mask = torch.stack(steps, dim=1).view(T, B, S, N + 1)
    rows_valid = slot < torch.clamp(sun, min=1).unsqueeze(-1)                 # teacher emits max(sun,1) rows
    step_mask = step_mask & rows_valid.unsqueeze(-1)
    teacher_logit = {
        'action_type': torch.randn((T, B, C.NUM_ACTIONS), generator=g),
        'delay': torch.randn((T, B, 128), generator=g),
        'queued': torch.randn((T, B, 2), generator=g),
        'selected_units': torch.where(step_mask, torch.randn((T, B, S, N + 1), generator=g),
                                      torch.full((1,), -1e9)),
        'target_unit': torch.where(tu_logits_mask, torch.randn((T, B, N), generator=g), torch.full((1,), -1e9)),
        'target_location': torch.randn((T, B, C.SPATIAL_Y * C.SPATIAL_X), generator=g),
    }
    reward = {}
    for f in ('winloss', 'build_order', 'built_unit', 'battle'):
        if f == 'winloss':
            r = torch.zeros((T, B))
            win = torch.rand((B,), generator=g) < 0.1
            sign = torch.where(torch.rand((B,), generator=g) < 0.5, -1.0, 1.0)
            r[-1] = win.float() * sign
        else:
            r = torch.randn((T, B), generator=g) * 0.1
        reward[f] = r
    mask = {
        'actions_mask': {k: actions_mask[k] for k in ('queued', 'selected_units', 'target_unit', 'target_location')},
        'cum_action_mask': torch.ones((T, B)), 'build_order_mask': torch.ones((T, B)),
        'built_unit_mask': torch.ones((T, B)),
        'selected_units_mask': su_mask, 'selected_units_logits_mask': su_logits_mask,
        'target_units_logits_mask': tu_logits_mask,
    }
    data.update({
        'action_info': action_info, 'selected_units_num': sun, 'behaviour_logp': behaviour_logp,
        'teacher_logit': teacher_logit, 'mask': mask, 'reward': reward,
        'step': torch.rand((T, B), generator=g) * 20000.0,
        'batch_size': B, 'unroll_len': T, 'aux_type': B,
        # host-side max(selected_units_num): the collate step has it on the CPU for free; lets the learner step run
        # without a device->host sync (optional key, ignored by the reference through **kwargs)
        'selected_units_max': max(int(sun.max()), 1),
    })
    return data


def gen_value_feature(rows, seed=77, n_units=C.MAX_ENTITY_NUM):
    """`value_feature` of the self-play learner (lib/features.py:690-765 + behaviour Z, agent.py:628-630): opponent
    statistics, own + enemy unit lists padded to `n_units`, two unit-presence planes, build order / cumulative stat."""
    g = _gen(seed)
    H, W = C.SPATIAL_Y, C.SPATIAL_X
    total = _randint(g, max(2, n_units // 4), n_units + 1, (rows,))
    valid = torch.arange(n_units).unsqueeze(0) < total.unsqueeze(1)

    def bern(shape, p=0.1):
        return (torch.rand(shape, generator=g) < p).to(torch.uint8)
    return {
        'unit_type': _randint(g, 0, C.NUM_UNIT_TYPES, (rows, n_units), torch.int16) * valid.to(torch.int16),
        'unit_alliance': _randint(g, 0, 2, (rows, n_units), torch.uint8) * valid.to(torch.uint8),
        'unit_x': _randint(g, 0, W, (rows, n_units), torch.uint8) * valid.to(torch.uint8),
        'unit_y': _randint(g, 0, H, (rows, n_units), torch.uint8) * valid.to(torch.uint8),
        'total_unit_count': total,
        'enemy_unit_counts_bow': bern((rows, C.NUM_UNIT_TYPES)) * _randint(g, 1, 6, (rows, C.NUM_UNIT_TYPES), torch.uint8),
        'enemy_unit_type_bool': bern((rows, C.NUM_UNIT_TYPES)),
        'enemy_agent_statistics': torch.rand((rows, 10), generator=g) * 8.0,
        'enemy_upgrades': bern((rows, C.NUM_UPGRADES)),
        'own_units_spatial': torch.rand((rows, 1, H, W), generator=g) < 0.02,
        'enemy_units_spatial': torch.rand((rows, 1, H, W), generator=g) < 0.02,
        'beginning_order': _randint(g, 0, C.NUM_BEGINNING_ORDER_ACTIONS, (rows, C.BEGINNING_ORDER_LENGTH), torch.int16),
        'bo_location': _randint(g, 0, H * W, (rows, C.BEGINNING_ORDER_LENGTH), torch.int16),
        'cumulative_stat': bern((rows, C.NUM_CUMULATIVE_STAT_ACTIONS)),
    }


def gen_gumbel(rows, seed=99, n_entity=C.MAX_ENTITY_NUM):
    """Fixed Gumbel noise for the six sampling sites (SURVEY.md App. A24 noise contract)."""
    g = _gen(seed)

    def gum(shape):
        u = torch.rand(shape, generator=g).clamp_(1e-10, 1.0 - 1e-7)
        return -torch.log(-torch.log(u))
    return {'action_type': gum((rows, C.NUM_ACTIONS)), 'delay': gum((rows, 128)), 'queued': gum((rows, 2)),
            'selected_units': gum((C.MAX_SELECTED_UNITS_NUM, rows, n_entity + 1)),
            'target_unit': gum((rows, n_entity)), 'target_location': gum((rows, C.SPATIAL_Y * C.SPATIAL_X))}


def synth_state_dict(spec, seed=0):
    """Deterministic weights for every *trainable* tensor of `spec` = [(key, shape), ...].

    Distributions track the reference initialisers (xavier-normal Linear/Conv weights,
    nn_module.py:17-46; unscaled randn LSTM matrices, lstm.py:125-126; LayerNorm affine near identity)
    but every tensor gets its own seeded stream so tensors are reproducible key by key.
    """
    out = {}
    for i, (key, shape) in enumerate(spec):
        g = _gen(seed * 1000003 + i * 7919 + 13)
        shape = tuple(shape)
        leaf = key.rsplit('.', 1)[-1]
        if 'weight_ih' in leaf or 'weight_hh' in leaf:
            t = torch.randn(shape, generator=g)
        elif 'layernorm' in key or key.endswith('.norm.weight') or key.endswith('.norm.bias') or \
                (len(shape) == 1 and leaf == 'weight'):
            t = (1.0 + 0.1 * torch.randn(shape, generator=g)) if leaf == 'weight' else 0.1 * torch.randn(shape, generator=g)
        elif leaf == 'UpdateSP':
            t = torch.full(shape, 0.1)
        elif len(shape) >= 2:
            rf = 1
            for s in shape[2:]:
                rf *= s
            fan_in, fan_out = shape[1] * rf, shape[0] * rf
            t = torch.randn(shape, generator=g) * math.sqrt(2.0 / (fan_in + fan_out))
        else:
            t = (torch.rand(shape, generator=g) - 0.5) * 0.1
        out[key] = t
    return out


def tree_map(fn, x):
    """Apply fn to every tensor of a nested dict/list/tuple structure."""
    if isinstance(x, torch.Tensor):
        return fn(x)
    if isinstance(x, dict):
        return {k: tree_map(fn, v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(tree_map(fn, v) for v in x)
    return x


def time_position_array(dim=32):
    """Sinusoidal time-code denominators, scalar_encoder.py:11-16,86."""
    x = torch.arange(0, dim, dtype=torch.float)
    x = x // 2 * 2
    x = torch.div(x, dim)
    x = torch.pow(10000., x)
    return torch.div(1., x)


def full_state_dict(spec, seed=0):
    """`spec` = [(key, shape, trainable)]: seeded trainables + the frozen position array.
    (Frozen one-hot / binary embedding tables are identities both sides never read.)"""
    sd = synth_state_dict([(k, s) for k, s, t in spec if t], seed)
    sd['encoder.scalar_encoder.position_array'] = time_position_array(32)
    return sd


def gen_sl_loss_inputs(b, s, N, seed):
    """Seeded inputs of SupervisedLoss.compute_loss (sl_learner.py:50-53): six logit tensors as `Model.sl_train` returns
    them (selected_units [b, s, N+1], -1e9 beyond each sample's entities), labels, per-head action masks, lengths.
    Sample 0 selects nothing (length 0, masked out), sample 1 has the maximum length s, one sample repeats a unit."""
    g = torch.Generator().manual_seed(seed)
    entity_num = torch.randint(max(s + 1, N // 2), N + 1, (b,), generator=g)
    sun = torch.randint(2, s + 1, (b,), generator=g)
    sun[0], sun[1 % b] = 0, s
    dims = {'action_type': 327, 'delay': 128, 'queued': 2, 'target_unit': N, 'target_location': 152 * 160}
    logits = {k: torch.randn(b, d, generator=g) * 2.0 for k, d in dims.items()}
    su = torch.randn(b, s, N + 1, generator=g) * 2.0
    ent = torch.arange(N + 1).unsqueeze(0)
    su = su.masked_fill((ent > entity_num.unsqueeze(1)).unsqueeze(1), -1e9)
    logits['selected_units'] = su
    logits['target_unit'] = logits['target_unit'].masked_fill(torch.arange(N).unsqueeze(0) >= entity_num.unsqueeze(1), -1e9)
    actions = {k: torch.randint(0, d, (b,), generator=g) for k, d in dims.items()}
    actions['target_unit'] = (torch.rand(b, generator=g) * entity_num).long()
    units = torch.zeros(b, 64, dtype=torch.long)
    infer = torch.zeros(b, s, dtype=torch.long)
    for i in range(b):
        n_i, L = int(entity_num[i]), int(sun[i])
        if L > 0:
            perm = torch.randperm(n_i, generator=g)[:L - 1]
            units[i, :L - 1] = perm
            units[i, L - 1] = n_i                                      # end flag
        guess = torch.randperm(n_i, generator=g)[:s]
        infer[i] = guess
        if L > 1:
            infer[i, :(L + 1) // 2] = units[i, :(L + 1) // 2]          # partly right
        if i % 2 == 0 and s > 2:
            infer[i, int(torch.randint(1, s, (1,), generator=g))] = n_i
    if b > 2 and int(sun[2]) > 2:
        units[2, 1] = units[2, 0]                                      # a repeated pick
    actions['selected_units'] = units
    mask = {k: (torch.rand(b, generator=g) > 0.3).float() for k in list(dims) + ['selected_units']}
    mask['action_type'] = torch.ones(b)
    mask['selected_units'][0] = 0.
    mask['queued'] = torch.zeros(b)                                   # a head with no valid row at all
    return {'logits': logits, 'actions': actions, 'actions_mask': mask, 'selected_units_num': sun, 'entity_num': entity_num,
            'infer_selected_units': infer}
