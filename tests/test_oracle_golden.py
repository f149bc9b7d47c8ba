"""CPU suite: pins the oracle restatement (oracle/restate.py) to the fixtures produced by the executed
reference (oracle/make_golden.py). Tolerances: fp32 restatement vs fp32 reference -- atol 1e-5 on valid
logits / values / losses (SURVEY §8c policy), masked logits as a predicate, index outputs bit-exact."""
import pytest
import torch

from applestar_b200 import config as C, synth, tables
from oracle import restate as R
from util import clone_tree, golden, load_spec, max_abs, valid_logits


@pytest.fixture(scope='module')
def learner():
    g = golden('learner_small.pt')
    sd = synth.full_state_dict(load_spec('four'), 0)
    sd = {k: v.requires_grad_(True) for k, v in sd.items()}
    data = synth.gen_trajectories(g['T'], g['B'], g['N'], g['seed'])
    out = R.rl_learner_forward(sd, g['baselines'], **clone_tree(data))
    logits = {k: v.detach().clone() for k, v in out['target_logit'].items()}
    values = {k: v.detach().clone() for k, v in out['value'].items()}
    info = R.compute_loss(C.DEFAULT_LOSS_CONFIG, out)
    info['total_loss'].backward()
    return g, sd, logits, values, info


def test_learner_logits_and_values(learner):
    g, sd, logits, values, info = learner
    for k in C.HEADS:
        a, b = valid_logits(logits[k], g['logit'][k])
        assert max_abs(a, b) < 1e-5, k
    for k, v in values.items():
        assert max_abs(v, g['value'][k]) < 1e-5, k


def test_learner_loss_dict(learner):
    g, sd, logits, values, info = learner
    assert set(info) == set(g['loss_info'])
    for k, v in g['loss_info'].items():
        assert abs(float(info[k]) - v) <= 1e-5 * max(1.0, abs(v)), k


def test_learner_gradients(learner):
    g, sd, logits, values, info = learner
    for k, (norm, total) in g['grad_summary'].items():
        mine = sd[k].grad
        mn = 0.0 if mine is None else mine.double().norm().item()
        assert abs(mn - float(norm)) <= 1e-3 * float(norm) + 1e-8, k
    for k, v in g['grads'].items():
        assert max_abs(sd[k].grad, v) <= 1e-3 * v.abs().max().item() + 1e-8, k


def test_inference_sampling():
    g = golden('infer_small.pt')
    sd = synth.full_state_dict(load_spec('winloss'), 0)
    data = synth.gen_observations(g['B'], g['N'], g['seed'], entity_num_range=(g['N'] // 2, g['N']), special=True)
    noise = synth.gen_gumbel(g['B'], g['seed'] + 1, g['N'])
    su_table = torch.tensor(tables.flags(tables.ACTION_HAS_SELECTED_UNITS))
    with torch.no_grad():
        out = R.compute_logp_action(sd, noise=noise, su_table=su_table, **clone_tree(data))
    # heads before the first sampled pointer step depend only on teacher-free quantities
    for k in ('action_type', 'delay', 'queued'):
        assert torch.equal(out['action_info'][k], g['action_info'][k]), k
        assert max_abs(*valid_logits(out['logit'][k], g['logit'][k])) < 1e-5, k
    assert torch.equal(out['selected_units_num'], g['selected_units_num'])
    steps = g['logit']['selected_units'].shape[1]
    assert out['logit']['selected_units'].shape[1] == steps
    # picks after a row's end flag are don't-care (all logits masked); compare the live prefix
    live = torch.arange(steps).unsqueeze(0) < g['selected_units_num'].unsqueeze(1)
    assert torch.equal(out['action_info']['selected_units'][live], g['action_info']['selected_units'][live])
    assert max_abs(*valid_logits(out['logit']['selected_units'][live], g['logit']['selected_units'][live])) < 2e-5
    for k in ('target_unit', 'target_location'):
        assert max_abs(*valid_logits(out['logit'][k], g['logit'][k])) < 2e-5, k
        assert torch.equal(out['action_info'][k], g['action_info'][k]), k
    assert torch.equal(out['extra_units'], g['extra_units'])
    for (h, c), (gh, gc) in zip(out['hidden_state'], g['hidden_state']):
        assert max_abs(h, gh) < 1e-5 and max_abs(c, gc) < 1e-5
    for k in C.HEADS:
        a, b = out['action_logp'][k], g['action_logp'][k]
        keep = (b > -1e8) & (live if k == 'selected_units' else torch.ones_like(b, dtype=torch.bool))
        assert max_abs(a[keep], b[keep]) < 2e-5, k


def test_scatter_connection_bit_exact():
    g = golden('scatter_small.pt')
    gen = torch.Generator().manual_seed(g['seed'])
    Rr, N, D = g['R'], g['N'], g['D']
    emb = torch.randn(Rr, N, D, generator=gen)
    x = torch.randint(0, 8, (Rr, N), generator=gen)
    y = torch.randint(0, 8, (Rr, N), generator=gen)
    x[1, :50] = 250
    y[1, :50] = 200
    for typ in ('add', 'cover'):
        m = R.scatter_connection(emb, x, y, typ).permute(0, 2, 3, 1).reshape(Rr, -1, D)
        cells = (m.abs().sum(-1) > 0).nonzero()
        assert torch.equal(cells, g['out'][typ]['cells'])
        assert torch.equal(m[cells[:, 0], cells[:, 1]], g['out'][typ]['vals'])   # bit-exact


def test_scan_matches_python_loops():
    """V-trace / UPGO / TD(lambda) scans against a scalar, pure-Python restatement (small case)."""
    gen = torch.Generator().manual_seed(3)
    T, B = 7, 3
    rho = torch.rand(T, B, generator=gen)
    r = torch.randn(T, B, generator=gen)
    v = torch.randn(T + 1, B, generator=gen)
    adv = R.vtrace_adv(rho, r, v)
    ret = R.lambda_returns(r, 0.997, v, 0.8)
    for b in range(B):
        vs = [0.0] * (T + 1)
        vs[T] = v[T, b].item()
        for t in range(T - 1, -1, -1):
            delta = rho[t, b].item() * (r[t, b].item() + v[t + 1, b].item() - v[t, b].item())
            vs[t] = v[t, b].item() + delta + rho[t, b].item() * (vs[t + 1] - v[t + 1, b].item())
        g_next = None
        for t in range(T - 1, -1, -1):
            a = rho[t, b].item() * (r[t, b].item() + vs[t + 1] - v[t, b].item())
            assert abs(a - adv[t, b].item()) < 1e-5
            if t == T - 1:
                g_next = r[t, b].item() + 0.997 * v[t + 1, b].item()
            else:
                g_next = r[t, b].item() + 0.997 * 0.8 * g_next + 0.997 * 0.2 * v[t + 1, b].item()
            assert abs(g_next - ret[t, b].item()) < 1e-5


def test_learner_n128_fixture():
    """The N = 128 fixture (fused-attention size) pins the restatement the same way: logits, values, total loss."""
    g = golden('learner_n128.pt')
    sd = synth.full_state_dict(load_spec('winloss'), 0)
    data = synth.gen_trajectories(g['T'], g['B'], g['N'], g['seed'])
    with torch.no_grad():
        out = R.rl_learner_forward(sd, g['baselines'], **clone_tree(data))
        info = R.compute_loss(C.DEFAULT_LOSS_CONFIG, out)
    for k in C.HEADS:
        a, b = valid_logits(out['target_logit'][k], g['logit'][k])
        assert max_abs(a, b) < 1e-5, k
    for k, v in out['value'].items():
        assert max_abs(v, g['value'][k]) < 1e-5, k
    v = g['loss_info']['total_loss']
    assert abs(float(info['total_loss']) - v) <= 1e-5 * max(1.0, abs(v))


def test_learner_n512_fixture():
    """The benchmarked entity count (N = 512, special observations): restatement vs executed reference -- logits, values,
    every loss scalar, all 394 gradient norms and the ~300 stored gradient tensors."""
    from util import grad_sample
    g = golden('learner_n512.pt')
    sd = synth.full_state_dict(load_spec('winloss'), 0)
    sd = {k: v.requires_grad_(True) for k, v in sd.items()}
    data = synth.gen_trajectories(g['T'], g['B'], g['N'], g['seed'], special=True)
    out = R.rl_learner_forward(sd, g['baselines'], **clone_tree(data))
    for k in C.HEADS:
        a, b = valid_logits(out['target_logit'][k].detach(), g['logit'][k])
        assert max_abs(a, b) < 2e-5, k
    for k, v in out['value'].items():
        assert max_abs(v.detach(), g['value'][k]) < 1e-5, k
    info = R.compute_loss(C.DEFAULT_LOSS_CONFIG, out)
    assert set(info) == set(g['loss_info'])
    for k, v in g['loss_info'].items():
        assert abs(float(info[k]) - v) <= 2e-5 * max(1.0, abs(v)), k
    info['total_loss'].backward()
    for k, (norm, _) in g['grad_summary'].items():
        mn = 0.0 if sd[k].grad is None else sd[k].grad.double().norm().item()
        assert abs(mn - float(norm)) <= 2e-3 * float(norm) + 1e-7, k
    assert len(g['grads']) >= 300
    for k, v in g['grads'].items():
        mine = grad_sample(sd[k].grad)
        assert max_abs(mine, v) <= 2e-3 * v.abs().max().item() + 1e-7, k


@pytest.mark.parametrize('tag,kw', [('default', {}), ('label_smooth', {'label_smooth': True}), ('no_su_mask', {'su_mask': False})])
def test_supervised_loss_restatement_vs_reference_fixture(tag, kw):
    """oracle.restate.supervised_loss against SupervisedLoss.compute_loss of the executed reference (sl_loss.py:102-285):
    all 15 log entries and d total_loss / d logits of every head."""
    import torch
    from applestar_b200 import synth
    from oracle import restate
    fx = golden('sl_loss_small.pt')
    m = fx['meta']
    d = synth.gen_sl_loss_inputs(m['b'], m['s'], m['N'], m['seed'])
    logits = {k: v.clone().requires_grad_(True) for k, v in d['logits'].items()}
    info = restate.supervised_loss(logits, d['actions'], d['actions_mask'], d['selected_units_num'], d['entity_num'],
                                   d['infer_selected_units'], **kw)
    ref = fx[tag]
    assert set(info) == set(ref['info'])
    for k, v in ref['info'].items():
        assert abs(float(info[k]) - v) <= 1e-6 + 1e-5 * abs(v), (k, float(info[k]), v)
    info['total_loss'].backward()
    for k, g in ref['dlogits'].items():
        assert torch.allclose(logits[k].grad, g, rtol=1e-5, atol=1e-7), k
    assert torch.allclose(logits['target_location'].grad[0], ref['dlogits_location_row0'], rtol=1e-5, atol=1e-8)
