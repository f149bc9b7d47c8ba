"""GPU parity at the BENCHMARKED entity count (N = 512: four key blocks in the fused attention, the 512-key scatter
sort, the cluster LSTM, up to 12 pointer steps) against the executed reference (tests/golden/learner_n512.pt) and
against the oracle run on the box's CPU.

Tolerance policy: a bf16 implementation may be 2x as far from the truth as the REFERENCE'S OWN
bf16-autocast run -- per logit head in max-abs AND RMS, per loss scalar, per module-level gradient aggregate in
relative L2 (util.grad_group_check); every gradient TENSOR is checked in relative L2 and cosine at 3x its own budget
(util.grad_check says why 3), and the median relative error over all tensors must not exceed the reference's.
Fixture keys: `bf16_autocast_err`, `bf16_autocast_loss_info`, `bf16_autocast_grad_err`.  No tensor may fail.  tools/parity_report.py prints the full per-tensor table (profiles/r02_parity_n512*.txt).
Masks and indices are compared bit-exactly."""
import pytest
import torch

from applestar_b200 import config as C, synth, tables
from oracle import restate as R
from util import clone_tree, golden, grad_budgets, grad_check, grad_group_check, grad_sample, load_spec

gpu = pytest.mark.gpu


def _model(baselines=('winloss',), spec='winloss', **cfg):
    from applestar_b200 import Model
    sd = synth.full_state_dict(load_spec(spec), 0)
    full = {'model': {'enable_baselines': list(baselines)}, 'agent': {'extra_units': True}}
    full['model'].update(cfg)
    m = Model(full, use_value_network=True)
    res = m.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys
    return m.cuda(), sd


def _cuda(tree):
    return synth.tree_map(lambda t: t.cuda(), tree)


def _check_logits(out_logit, g, slack=2.0):
    bad = []
    for k in C.HEADS:
        a = out_logit[k].detach().double().cpu()
        ref64 = g['logit64'][k]
        masked = ref64 <= -1e8
        assert torch.equal(a <= -1e8, masked), 'mask predicate differs: ' + k          # bit-exact
        d = (a - ref64)[~masked]
        mx, rms = d.abs().max().item(), d.pow(2).mean().sqrt().item()
        bmx, brms = (slack * v.item() for v in g['bf16_autocast_err'][k])
        if mx > bmx or rms > brms:
            bad.append((k, 'max-abs %.4f (budget %.4f), rms %.5f (budget %.5f)' % (mx, bmx, rms, brms)))
    return bad


@gpu
def test_learner_n512_vs_golden():
    from applestar_b200 import ReinforcementLoss
    g = golden('learner_n512.pt')
    model, sd = _model()
    data = synth.gen_trajectories(g['T'], g['B'], g['N'], g['seed'], special=True)
    out = model.rl_learner_forward(**_cuda(data))
    bad = _check_logits(out['target_logit'], g)
    for f, v in out['value'].items():
        d = v.detach().double().cpu() - g['value64'][f]
        bmx, brms = (2.0 * t.item() for t in g['bf16_autocast_err']['value/' + f])
        if d.abs().max().item() > bmx or d.pow(2).mean().sqrt().item() > brms:
            bad.append(('value/' + f, d.abs().max().item(), bmx))
    assert not bad, bad
    info = ReinforcementLoss({}, 'MP0').compute_loss(out)
    assert set(info) == set(g['loss_info'])
    for k, ref in g['loss_info'].items():                        # every scalar of the reference's log dict
        budget = 2.0 * abs(g['bf16_autocast_loss_info'][k] - ref) + 2e-3 * max(1.0, abs(ref))
        if abs(float(info[k]) - ref) > budget:
            bad.append((k, float(info[k]), ref, budget))
    assert not bad, bad
    info['total_loss'].backward()
    params = dict(model.named_parameters())
    gerr, typical = grad_budgets(g)
    assert len(g['grads']) >= 300
    mine = {k: grad_sample(params[k].grad) for k in g['grads']}
    rels = []
    for k, ref in g['grads'].items():                            # ~300 gradient TENSORS (full, or a fixed sample)
        assert params[k].grad is not None, k
        why = grad_check(mine[k], ref, gerr[k], typical)
        if why:
            bad.append((k, why))
        if float(gerr[k][0]) < 1.0:
            rels.append(((mine[k].double() - ref.double()).norm() / ref.double().norm()).item())
    bad += grad_group_check(mine, g['grads'], gerr, typical)[0]
    assert sorted(rels)[len(rels) // 2] <= typical, 'median relative gradient error above the reference bf16 run'
    for k, (norm, _) in g['grad_summary'].items():               # all 394: gradient norm within the same budget
        gn = params[k].grad.double().norm().item()
        rel_b = float(gerr[k][0])
        if rel_b > 1.0:
            continue
        if abs(gn - norm.item()) > max(3.0 * rel_b, typical) * norm.item() + 1e-6:
            bad.append((k, 'norm %.4e vs %.4e' % (gn, norm.item())))
    assert not bad, (len(bad), bad[:12])


@gpu
def test_learner_n512_all_gradients_vs_oracle():
    """T = 8, B = 2, N = 512: the oracle restatement runs on the box's CPU (a few seconds) and EVERY one of the 394
    gradient tensors is compared in full (cosine / relative L2).  The budgets are the reference's bf16-autocast
    errors on exactly this batch (tests/golden/learner_n512_t8_budget.pt, `make_golden.py n512t8`)."""
    from applestar_b200 import ReinforcementLoss
    g = golden('learner_n512_t8_budget.pt')
    model, sd = _model()
    data = synth.gen_trajectories(8, 2, 512, 8512, special=True)
    out = model.rl_learner_forward(**_cuda(data))
    info = ReinforcementLoss({}, 'MP0').compute_loss(out)
    info['total_loss'].backward()
    sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ref = R.rl_learner_forward(sdr, ['winloss'], **clone_tree(data))
    ref_info = R.compute_loss(C.DEFAULT_LOSS_CONFIG, ref)
    ref_info['total_loss'].backward()
    for k in C.HEADS:
        a, b = out['target_logit'][k].detach().cpu(), ref['target_logit'][k].detach()
        keep = b > -1e8
        assert torch.equal(a <= -1e8, ~keep), k
        d = (a - b)[keep].double()
        bmx, brms = (2.0 * v.item() for v in g['bf16_autocast_err'][k])
        assert d.abs().max().item() <= bmx and d.pow(2).mean().sqrt().item() <= brms, (k, d.abs().max().item(), bmx)
    bad = []
    gerr, typical = grad_budgets(g)
    params = dict(model.named_parameters())
    mine, refs = {}, {}
    for k, p in params.items():
        if not p.requires_grad:
            continue
        refs[k] = (sdr[k].grad if sdr[k].grad is not None else torch.zeros_like(sdr[k])).reshape(-1)
        mine[k] = p.grad.detach().float().cpu().reshape(-1)
        why = grad_check(mine[k], refs[k], gerr[k], typical)
        if why:
            bad.append((k, why))
    assert len(refs) == 394
    bad += grad_group_check(mine, refs, gerr, typical)[0]
    assert not bad, (len(bad), bad[:12])


@gpu
def test_compute_teacher_logit_vs_oracle():
    """Model.compute_teacher_logit (model.py:76-93): the frozen-teacher path of every learner iteration."""
    g = golden('learner_n512.pt')
    model, sd = _model()
    model.eval()
    B, N = 6, 512
    data = synth.gen_trajectories(1, B, N, 31, special=True)
    obs = {k: synth.tree_map(lambda t: t[:B], data[k]) for k in ('spatial_info', 'entity_info', 'scalar_info')}
    en = data['entity_num'][:B]
    hs = [(h[:B], c[:B]) for h, c in data['hidden_state']]
    act = {k: v[0] for k, v in data['action_info'].items()}
    sun = data['selected_units_num'][0]
    with torch.no_grad():
        out = model.compute_teacher_logit(**_cuda(obs), entity_num=en.cuda(), hidden_state=_cuda(hs),
                                          selected_units_num=sun.cuda(), action_info=_cuda(act))
        ref = R.compute_teacher_logit(sd, **clone_tree(obs), entity_num=en.clone(), hidden_state=clone_tree(hs),
                                      selected_units_num=sun.clone(), action_info=clone_tree(act))
    assert set(out) == {'logit', 'hidden_state', 'entity_num', 'selected_units_num'}
    assert torch.equal(out['selected_units_num'].cpu(), sun) and torch.equal(out['entity_num'].cpu(), en)
    for k in C.HEADS:
        a, b = out['logit'][k].float().cpu(), ref['logit'][k]
        assert a.shape == b.shape, (k, a.shape, b.shape)
        keep = b > -1e8
        assert torch.equal(a <= -1e8, ~keep), k
        d = (a - b)[keep].double()
        bmx, brms = (2.0 * v.item() for v in g['bf16_autocast_err'][k])
        assert d.abs().max().item() <= bmx and d.pow(2).mean().sqrt().item() <= brms, (k, d.abs().max().item(), bmx)
    for (h, c), (hr, cr) in zip(out['hidden_state'], ref['hidden_state']):
        assert (h.cpu() - hr).abs().max().item() < 0.05 and (c.cpu() - cr).abs().max().item() < 0.1


@gpu
def test_inference_all_heads_teacher_forced_replay():
    """compute_logp_action at N = 512: the device's own sampled actions are replayed teacher-forced through the oracle
    (compute_teacher_logit), which pins the logits of ALL six heads of the sampling path at the bf16 budget; the
    Gumbel arg-max of the device logits is bit-exact; log-probs match the device logits."""
    g = golden('learner_n512.pt')
    model, sd = _model()
    model.eval()
    B, N = 5, 512
    data = synth.gen_observations(B, N, 77, entity_num_range=(N // 2, N), special=True)
    noise = synth.gen_gumbel(B, 78, N)
    with torch.no_grad():
        out = model.compute_logp_action(gumbel_noise=_cuda(noise), **_cuda(data))
    act = {k: v.cpu() for k, v in out['action_info'].items()}
    sun = out['selected_units_num'].cpu()
    steps = act['selected_units'].shape[1]
    su_table = torch.tensor(tables.flags(tables.ACTION_HAS_SELECTED_UNITS))
    assert torch.equal(sun > 0, su_table[act['action_type']])
    # bit-exact index work given the device logits
    for k, nz in (('action_type', noise['action_type']), ('delay', noise['delay']), ('queued', noise['queued']),
                  ('target_unit', noise['target_unit']), ('target_location', noise['target_location'])):
        lg = out['logit'][k].float().cpu()
        assert torch.equal((lg + nz).argmax(-1), act[k]), k
    su_lg = out['logit']['selected_units'].float().cpu()
    for s in range(steps):
        live = s < sun
        pick = (su_lg[:, s] + noise['selected_units'][s]).argmax(-1)
        assert torch.equal(pick[live], act['selected_units'][live, s]), s
    # teacher-forced oracle replay of the sampled actions
    sel = torch.zeros(B, C.MAX_SELECTED_UNITS_NUM, dtype=torch.long)
    for b in range(B):
        n = int(sun[b])
        sel[b, :n] = act['selected_units'][b, :n]
    tf_act = dict(act)
    tf_act['selected_units'] = sel
    with torch.no_grad():
        ref = R.compute_teacher_logit(sd, **clone_tree({k: data[k] for k in ('spatial_info', 'entity_info', 'scalar_info')}),
                                      entity_num=data['entity_num'].clone(), hidden_state=clone_tree(data['hidden_state']),
                                      selected_units_num=sun.clone(), action_info=tf_act)
    for k in C.HEADS:
        a, b = out['logit'][k].float().cpu(), ref['logit'][k]
        if k == 'selected_units':
            S = min(a.shape[1], b.shape[1])
            live = (torch.arange(S).unsqueeze(0) < sun.unsqueeze(1)).unsqueeze(-1)
            a, b = a[:, :S], b[:, :S]
            keep = (b > -1e8) & live
            assert torch.equal((a <= -1e8) & live, (b <= -1e8) & live), k
        else:
            keep = b > -1e8
            assert torch.equal(a <= -1e8, ~keep), k
        d = (a - b)[keep].double()
        if d.numel() == 0:
            continue
        bmx, brms = (2.0 * v.item() for v in g['bf16_autocast_err'][k])
        assert d.abs().max().item() <= bmx and d.pow(2).mean().sqrt().item() <= brms, \
            (k, d.abs().max().item(), bmx, d.pow(2).mean().sqrt().item(), brms)
    for k in ('action_type', 'delay', 'queued', 'target_unit', 'target_location'):
        lp = torch.log_softmax(out['logit'][k].float().cpu(), -1).gather(-1, act[k].unsqueeze(-1)).squeeze(-1)
        assert (lp - out['action_logp'][k].cpu()).abs().max().item() < 1e-4, k


@gpu
def test_model_forward_entry_point():
    """Model.forward (model.py:46-54): (action_info, selected_units_num, hidden_state) -- same sampled actions as
    compute_logp_action under the same Gumbel noise."""
    model, sd = _model()
    model.eval()
    B, N = 4, 512
    data = _cuda(synth.gen_observations(B, N, 5, entity_num_range=(N // 2, N)))
    noise = _cuda(synth.gen_gumbel(B, 6, N))
    with torch.no_grad():
        action, sun, state = model(gumbel_noise=noise, **data)
        full = model.compute_logp_action(gumbel_noise=noise, **data)
    for k in C.HEADS:
        assert torch.equal(action[k], full['action_info'][k]), k
    assert torch.equal(sun, full['selected_units_num'])
    assert len(state) == 3 and all(torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
                                   for a, b in zip(state, full['hidden_state']))


@gpu
def test_sl_train_entry_point_vs_oracle():
    """Model.sl_train (model.py:170-189): batch-major trajectories [B*T], LSTM from the given state, teacher-forced
    policy on every row.  Oracle: the same restated blocks with the batch-major reshape."""
    g = golden('learner_n512.pt')
    model, sd = _model()
    Bt, T, N = 2, 3, 128
    rows = Bt * T
    data = synth.gen_trajectories(1, rows, N, 91)
    obs = {k: synth.tree_map(lambda t: t[:rows], data[k]) for k in ('spatial_info', 'entity_info', 'scalar_info')}
    en = data['entity_num'][:rows]
    act = {k: v[0] for k, v in data['action_info'].items()}
    sun = data['selected_units_num'][0]
    hs = [(h[:Bt], c[:Bt]) for h, c in data['hidden_state']]
    with torch.no_grad():
        logits, action, state = model.sl_train(**_cuda(obs), entity_num=en.cuda(), selected_units_num=sun.cuda(),
                                               traj_lens=[T] * Bt, hidden_state=_cuda(hs), action_info=_cuda(act))
        lstm_input, ctx, _, ent, skip = R.encoder(sd, **clone_tree(obs), entity_num=en.clone())
        x = lstm_input.view(Bt, T, -1).permute(1, 0, 2)                     # model.py:178-181
        o, rstate = R.lnlstm(sd, 'core_lstm', x, clone_tree(hs), 3)
        o = o.permute(1, 0, 2).reshape(rows, -1)
        ref = R.policy_train(sd, o, ent, skip, ctx, en, clone_tree(act), sun.clone())
    for k in C.HEADS:
        a, b = logits[k].float().cpu(), ref[k]
        keep = b > -1e8
        assert torch.equal(a <= -1e8, ~keep), k
        d = (a - b)[keep].double()
        bmx = 3.0 * g['bf16_autocast_err'][k][0].item()
        assert d.abs().max().item() <= bmx, (k, d.abs().max().item(), bmx)
        assert torch.equal(action[k].cpu(), act[k]), k
    for (h, c), (hr, cr) in zip(state, rstate):
        assert (h.cpu() - hr).abs().max().item() < 0.05


@gpu
def test_only_update_baseline_and_only_update_value():
    """`Model.only_update_baseline` (model.py:138-140) detaches the critic input: value-tower gradients unchanged, and
    with `ReinforcementLoss.only_update_value` (rl_loss.py:170-171, value pre-training) nothing outside the value
    towers receives a gradient; total_loss == weighted TD loss.  `reset` (rl_loss.py:187-199) re-reads the weights
    and clears the flag."""
    from applestar_b200 import ReinforcementLoss
    model, sd = _model()
    data = synth.gen_trajectories(3, 2, 64, 17)
    loss = ReinforcementLoss({}, 'MP0')
    out = model.rl_learner_forward(**_cuda(data))
    info = loss.compute_loss(out)
    info['total_loss'].backward()
    full = {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
    model.zero_grad(set_to_none=True)

    model.only_update_baseline = True
    loss.only_update_value = True
    out = model.rl_learner_forward(**_cuda(data))
    info2 = loss.compute_loss(out)
    info2['total_loss'].backward()
    w_td = float(loss.loss_weights['baseline']['winloss'])
    assert abs(float(info2['total_loss']) - w_td * float(info2['winloss/td'])) < 1e-5
    assert abs(float(info2['winloss/td']) - float(info['winloss/td'])) < 1e-5        # same forward
    sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ref = R.rl_learner_forward(sdr, ['winloss'], **clone_tree(data))
    rinfo = R.compute_loss(C.DEFAULT_LOSS_CONFIG, ref, only_update_value=True)
    rinfo['total_loss'].backward()
    for k, p in model.named_parameters():
        if not p.requires_grad:
            continue
        if k.startswith('value_networks.'):
            assert p.grad is not None and float(p.grad.abs().sum()) > 0, k
            rg = sdr[k].grad
            cos = torch.dot(p.grad.cpu().double().flatten(), rg.double().flatten()) / (
                p.grad.double().norm().cpu() * rg.double().norm()).clamp(min=1e-30)
            assert cos.item() > 0.98, (k, cos.item())
        else:
            assert p.grad is None or float(p.grad.abs().sum()) == 0.0, k
    loss.reset({'loss_weights': {'entropy': 0.0}})
    assert loss.only_update_value is False and float(loss.loss_weights['entropy']) == 0.0
    model.only_update_baseline = False
    model.zero_grad(set_to_none=True)
    out = model.rl_learner_forward(**_cuda(data))
    info3 = loss.compute_loss(out)
    ent_term = float(C.DEFAULT_LOSS_CONFIG['loss_weights']['entropy']) * float(info['entropy/total'])
    want = float(info['total_loss']) - ent_term                           # the entropy term is gone, the rest stays
    assert abs(float(info3['total_loss']) - want) < 2e-4 * max(1.0, abs(want))


@gpu
def test_value_encoder_path_vs_oracle():
    """`learner.use_value_feature` (obs_encoder/value_encoder.py:12-74, model.py:141-144): second ScatterConnection call
    site (8 channels), 10-channel spatial tower, widened critics.  Values against the oracle (pinned to the executed
    reference by tests/test_oracle_vs_reference.py) and the gradient of every value-encoder / critic tensor in direction."""
    from applestar_b200 import Model
    g = golden('learner_n512.pt')
    spec = load_spec('winloss_value_feature')
    sd = synth.full_state_dict(spec, 0)
    model = Model({'model': {'enable_baselines': ['winloss']}, 'learner': {'use_value_feature': True}}, use_value_network=True)
    res = model.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys
    model = model.cuda()
    T, B, N = 3, 2, 128
    data = synth.gen_trajectories(T, B, N, 21)
    vf = synth.gen_value_feature((T + 1) * B, 22, N)
    out = model.rl_learner_forward(value_feature=_cuda(vf), **_cuda(data))
    v = out['value']['winloss']
    sdr = {k: t.clone().requires_grad_(True) for k, t in sd.items()}
    ref = R.rl_learner_forward(sdr, ['winloss'], value_feature=clone_tree(vf), **clone_tree(data))['value']['winloss']
    d = (v.detach().cpu() - ref.detach()).double()
    bmx, brms = (2.0 * t.item() for t in g['bf16_autocast_err']['value/winloss'])
    assert d.abs().max().item() <= bmx and d.pow(2).mean().sqrt().item() <= brms, (d.abs().max().item(), bmx)
    w = torch.randn(T + 1, B, generator=torch.Generator().manual_seed(3))
    (v * w.cuda()).sum().backward()
    (ref * w).sum().backward()
    bad = []
    for k, p in model.named_parameters():
        if not (k.startswith('value_encoder.') or k.startswith('value_networks.')) or not p.requires_grad:
            continue
        a, b = p.grad.detach().cpu().double().flatten(), sdr[k].grad.double().flatten()
        if b.norm().item() < 1e-9:
            continue
        cos = (torch.dot(a, b) / (a.norm() * b.norm()).clamp(min=1e-30)).item()
        rel = ((a - b).norm() / b.norm()).item()
        if cos < 0.95 or rel > 0.35:        # gradients pass the 34-layer critic tower in bf16: the reference's own bf16
            bad.append((k, cos, rel))       # run is ~0.24 off on these tensors (fixture, value_networks aggregate)
    assert not bad, bad[:8]


@gpu
def test_actor_shared_batch_inference_protocol():
    """actor_batch.SharedBatchInference (reference: Agent.gpu_batch_inference, agent.py:781-805, with copy_input_data /
    copy_output_data, :37-89): only rows whose flag is raised are served and written back, flags are cleared, the rest of
    the shared output stays untouched; served results equal the direct model call on the same rows (teacher path:
    deterministic logits; sampling path: same Gumbel noise -> same actions)."""
    from applestar_b200 import actor_batch
    model, sd = _model()
    model.eval()
    E, N = 4, 512                                            # environments = rows of the shared buffers
    data = synth.gen_trajectories(1, E, N, 41, special=True)
    obs = {k: synth.tree_map(lambda t: t[:E], data[k]) for k in ('spatial_info', 'entity_info', 'scalar_info')}
    en = data['entity_num'][:E]
    hs = [(h[:E], c[:E]) for h, c in data['hidden_state']]
    act = {k: v[0] for k, v in data['action_info'].items()}
    sun = data['selected_units_num'][0]
    full = dict(obs, entity_num=en, hidden_state=hs, action_info=act, selected_units_num=sun)
    shared_in = synth.tree_map(lambda t: torch.zeros_like(t).share_memory_(), full)
    signals = torch.zeros(E).share_memory_()
    # environment side: rows 0 and 2 publish their step, 1 and 3 stay idle
    for idx in (0, 2):
        step = {'entity_num': en[idx], 'selected_units_num': sun[idx],
                'hidden_state': [(h[idx], c[idx]) for h, c in hs],
                'spatial_info': {k: v[idx] for k, v in obs['spatial_info'].items()},
                'scalar_info': {k: v[idx] for k, v in obs['scalar_info'].items()},
                'entity_info': {k: v[idx, :int(en[idx])] for k, v in obs['entity_info'].items()},
                'action_info': {k: (v[idx, :int(sun[idx])] if k == 'selected_units' else v[idx]) for k, v in act.items()}}
        actor_batch.copy_input_data(shared_in, step, idx)
        signals[idx] = 1
    with torch.no_grad():
        want = model.compute_teacher_logit(**_cuda(synth.tree_map(lambda t: t.clone(), {k: shared_in[k] for k in (
            'spatial_info', 'entity_info', 'scalar_info')})), entity_num=shared_in['entity_num'].cuda(),
            hidden_state=_cuda(shared_in['hidden_state']), selected_units_num=shared_in['selected_units_num'].cuda(),
            action_info=_cuda(shared_in['action_info']))
    shared_out = {'logit': {k: torch.zeros((E,) + tuple(v.shape[1:])).share_memory_() for k, v in want['logit'].items()},
                  'hidden_state': [(torch.zeros(E, 384).share_memory_(), torch.zeros(E, 384).share_memory_()) for _ in range(3)]}
    srv = actor_batch.SharedBatchInference(model, shared_in, shared_out, signals, teacher=True)
    assert srv.poll() == 2 and float(signals.sum()) == 0 and srv.poll() == 0
    for k, v in want['logit'].items():
        got = shared_out['logit'][k]
        for idx in (0, 2):
            assert torch.equal(got[idx], v[idx].float().cpu()), k
        assert float(got[1].abs().sum()) == 0 and float(got[3].abs().sum()) == 0
    for layer in range(3):
        assert torch.equal(shared_out['hidden_state'][layer][0][2], want['hidden_state'][layer][0][2].float().cpu())
    # sampling path through ActorStep with fixed noise
    from applestar_b200 import ops
    noise = ops.fresh_gumbel(E, N, torch.device('cuda:0'))
    with torch.no_grad():
        ref = model.compute_logp_action(**_cuda({k: shared_in[k] for k in ('spatial_info', 'entity_info', 'scalar_info')}),
                                        entity_num=shared_in['entity_num'].cuda(), hidden_state=_cuda(shared_in['hidden_state']),
                                        gumbel_noise=noise)
    out2 = {'action_info': {k: torch.full((E,) + ((64,) if k == 'selected_units' else ()), -1, dtype=torch.int64).share_memory_()
                            for k in ref['action_info']},
            'selected_units_num': torch.zeros(E, dtype=torch.int64).share_memory_()}
    signals[1] = 1
    srv2 = actor_batch.SharedBatchInference(model, shared_in, out2, signals, teacher=False)
    assert srv2.poll(gumbel_noise=noise) == 1
    for k, v in ref['action_info'].items():
        got = out2['action_info'][k][1]
        want_row = v[1].cpu()
        assert torch.equal(got[:want_row.numel()] if got.dim() else got, want_row), k
        assert int(out2['action_info'][k][0].reshape(-1)[0]) == -1           # rows that did not ask keep their content
    assert int(out2['selected_units_num'][1]) == int(ref['selected_units_num'][1])


@gpu
def test_learner_step_at_benchmark_size_properties():
    """BASELINE.json's configuration (T = 64, B = 16, N = 512) is too large for the CPU oracle, so it is covered by
    size-independent properties: (1) run-to-run determinism -- every kernel sums in a fixed order, so two steps from
    the same weights give bit-identical statistics and gradients; (2) everything is finite; (3) trajectories are
    independent: the first 8 trajectories alone give the same policy logits as inside the batch of 16;
    (4) the weight gradient of the batch is the mean of the two half-batch gradients up to bf16 / summation-order
    effects (loss means are over T * B: as_rl_utils.py:22,45,68,99,242)."""
    from applestar_b200 import Model, ReinforcementLoss
    T, B, N = 64, 16, 512
    sd = synth.full_state_dict(load_spec('winloss'), 0)
    model = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
    model.load_state_dict(sd, strict=False)
    model = model.cuda()
    loss_fn = ReinforcementLoss({}, 'MP0')
    host = synth.gen_trajectories(T, B, N, 1234)

    def take(batch, sl):
        """trajectories `sl` of a [T(+1) * B, ...] / [T, B, ...] batch"""
        Bn = len(range(*sl.indices(B)))

        def cut(t, rows):
            if t.dim() >= 2 and t.shape[0] == T and t.shape[1] == B:
                return t[:, sl].contiguous()
            if t.dim() >= 1 and t.shape[0] == rows * B:
                return t.view(rows, B, *t.shape[1:])[:, sl].reshape(rows * Bn, *t.shape[1:]).contiguous()
            return t
        out = {}
        for k, v in batch.items():
            if k in ('batch_size',):
                out[k] = Bn
            elif isinstance(v, (dict, list, tuple, torch.Tensor)):
                out[k] = synth.tree_map(lambda t: cut(t, T + 1 if t.shape[0] == (T + 1) * B else T), v)
            else:
                out[k] = v
        return out

    def run(batch):
        model.zero_grad(set_to_none=True)
        dev_batch = synth.tree_map(lambda t: t.cuda(), batch)
        out = model.rl_learner_forward(**dev_batch)
        info = loss_fn.compute_loss(out)
        info['total_loss'].backward()
        grads = {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.grad is not None}
        stats = {k: float(v) for k, v in info.items()}
        return stats, grads, {k: v.detach().float().clone() for k, v in out['target_logit'].items()}

    s1, g1, l1 = run(host)
    s2, g2, l2 = run(host)
    assert s1 == s2                                                       # (1) bit-identical statistics ...
    assert set(g1) == set(g2) and all(torch.equal(g1[k], g2[k]) for k in g1)   # ... and gradients
    assert all(torch.equal(l1[k], l2[k]) for k in l1)
    assert all(abs(v) < 1e6 for v in s1.values())                         # (2)
    assert all(torch.isfinite(g).all() for g in g1.values())
    sa, ga, la = run(take(host, slice(0, 8)))
    sb, gb, _ = run(take(host, slice(8, 16)))
    for k in ('action_type', 'delay', 'queued', 'target_unit', 'target_location'):      # (3)
        full = l1[k][:, :8]
        keep = full > -1e8
        assert torch.equal(la[k] <= -1e8, ~keep), k
        d = torch.where(keep, (la[k] - full).abs(), torch.zeros_like(full))
        # same rows through the same kernels: no coupling across trajectories.  What may differ is the fp32 summation
        # order of the split-K GEMMs (the number of K splits follows the row count), i.e. last-bit differences that now
        # and then flip a bf16 rounding of an activation; with the randn recurrent weights of a fresh model such a flip
        # grows along the unroll (tests/test_nn_gpu.py pins the recurrence step by step instead)
        assert d[:2].max().item() <= 0.05 and d[:2].mean().item() <= 1e-3, (k, d[:2].max().item(), d[:2].mean().item())
        assert d.max().item() <= 0.5 and d.mean().item() <= 0.02, (k, d.max().item(), d.mean().item())
    num = den = 0.0
    for k in g1:                                                          # (4)
        if k in ga and k in gb:
            mean = 0.5 * (ga[k].double() + gb[k].double())
            num += (g1[k].double() - mean).pow(2).sum().item()
            den += mean.pow(2).sum().item()
    assert (num / den) ** 0.5 < 0.25, (num / den) ** 0.5
    assert abs(s1['total_loss'] - 0.5 * (sa['total_loss'] + sb['total_loss'])) <= 2e-2 * max(1.0, abs(s1['total_loss'])), (
        s1['total_loss'], sa['total_loss'], sb['total_loss'])
