"""End-to-end GPU parity of the Model / ReinforcementLoss mirror against the oracle and the golden fixtures
produced by the executed reference.  bf16 tensor-core GEMMs => tolerance is calibrated per output from the
reference's own bf16-autocast error (fixture `bf16_autocast_err`): err(ours, fp64) <= 2 x that."""
import pytest
import torch

from applestar_b200 import config as C, synth, tables
from oracle import restate as R
from oracle import restate as R_
from util import clone_tree, golden, load_spec

gpu = pytest.mark.gpu


def _model(baselines, spec_name):
    from applestar_b200 import Model
    sd = synth.full_state_dict(load_spec(spec_name), 0)
    m = Model({'model': {'enable_baselines': baselines}, 'agent': {'extra_units': True}}, use_value_network=True)
    res = m.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys
    assert all('encode_modules' in k or 'one_hot' in k or 'binary' in k or 'position_array' in k
               for k in res.missing_keys), res.missing_keys
    return m.cuda(), sd


def test_state_dict_contract():
    from applestar_b200 import Model
    for name, bl in (('winloss', ['winloss']), ('four', ['winloss', 'build_order', 'built_unit', 'battle'])):
        m = Model({'model': {'enable_baselines': bl}}, use_value_network=True)
        got = {k: tuple(v.shape) for k, v in m.state_dict().items()}
        want = {k: s for k, s, _ in load_spec(name)}
        assert got == want
        trainable = {k for k, p in m.named_parameters() if p.requires_grad}
        assert trainable == {k for k, _, t in load_spec(name) if t}


@gpu
@pytest.mark.parametrize('fixture,spec', [('learner_small.pt', 'four'), ('learner_n128.pt', 'winloss')])
def test_learner_forward_loss_backward_vs_golden(fixture, spec):
    """learner_small: N = 64 (composed attention path); learner_n128: N = 128, where the fused tcgen05 attention forward
    and backward kernels, the scatter-fused feature kernel and the tap-pair weight gradients all run."""
    from applestar_b200 import ReinforcementLoss
    g = golden(fixture)
    model, sd = _model(g['baselines'], spec)
    data = synth.gen_trajectories(g['T'], g['B'], g['N'], g['seed'])
    out = model.rl_learner_forward(**synth.tree_map(lambda t: t.cuda(), data))
    for k in C.HEADS:
        a = out['target_logit'][k].detach().double().cpu()
        ref64 = g['logit64'][k]
        masked = ref64 <= -1e8
        assert torch.equal(a <= -1e8, masked), k                       # mask predicate, bit-exact
        err = (a - ref64)[~masked].abs().max().item()
        budget = 2.0 * g['bf16_autocast_err'][k][0].item()
        assert err <= budget, (k, err, budget)
    for f, v in out['value'].items():
        err = (v.detach().double().cpu() - g['value64'][f]).abs().max().item()
        assert err <= 2.0 * g['bf16_autocast_err']['value/' + f][0].item(), (f, err)
    info = ReinforcementLoss({}, 'MP0').compute_loss(out)
    assert set(info) == set(g['loss_info'])
    ref_total = g['loss_info']['total_loss']
    assert abs(float(info['total_loss']) - ref_total) <= 0.05 * max(1.0, abs(ref_total))
    info['total_loss'].backward()
    params = dict(model.named_parameters())
    bad = []
    for k, (norm, _) in g['grad_summary'].items():
        gn = params[k].grad.double().norm().item() if params[k].grad is not None else 0.0
        if abs(gn - norm.item()) > 0.15 * norm.item() + 1e-4:
            bad.append((k, gn, norm.item()))
    assert len(bad) <= 0.05 * len(g['grad_summary']), bad[:10]         # >=95% of tensors within 15% grad-norm


@gpu
def test_loss_on_model_logits_matches_oracle_loss():
    """Same device logits through the fused loss and through the oracle loss: fp32-level agreement."""
    from applestar_b200 import ReinforcementLoss
    model, sd = _model(['winloss'], 'winloss')
    data = synth.gen_trajectories(3, 2, 48, 5)
    out = model.rl_learner_forward(**synth.tree_map(lambda t: t.cuda(), data))
    info = ReinforcementLoss({}, 'MP0').compute_loss(out)
    cpu_out = synth.tree_map(lambda t: t.detach().cpu(), {k: v for k, v in out.items()})
    ref = R.compute_loss(C.DEFAULT_LOSS_CONFIG, cpu_out)
    for k, v in ref.items():
        assert abs(float(info[k]) - float(v)) <= 2e-5 + 2e-4 * abs(float(v)), k


@gpu
def test_inference_sampling_vs_golden():
    g = golden('infer_small.pt')
    model, sd = _model(['winloss'], 'winloss')
    model.eval()
    data = synth.gen_observations(g['B'], g['N'], g['seed'], entity_num_range=(g['N'] // 2, g['N']), special=True)
    noise = synth.gen_gumbel(g['B'], g['seed'] + 1, g['N'])
    with torch.no_grad():
        out = model.compute_logp_action(gumbel_noise=synth.tree_map(lambda t: t.cuda(), noise),
                                        **synth.tree_map(lambda t: t.cuda(), data))
    # index work given identical logits is bit-exact (test_gumbel_argmax_bit_exact); end to end, bf16 logits can
    # flip near-ties, so agreement is a rate: the three cheap heads must agree on >= 2/3 rows
    agree = sum(int((out['action_info'][k].cpu() == g['action_info'][k]).sum()) for k in ('action_type', 'delay', 'queued'))
    assert agree >= 6
    for k in ('action_type', 'delay', 'queued'):
        a, b = out['logit'][k].cpu(), g['logit'][k]
        assert (a - b).abs().max().item() < 0.25, k
    assert out['logit']['selected_units'].shape[2] == g['N'] + 1
    assert out['extra_units'].shape == (g['B'], g['N'] + 1)
    for (h, c) in out['hidden_state']:
        assert h.shape == (g['B'], 384) and torch.isfinite(h).all() and torch.isfinite(c).all()
    # self-consistency with the oracle on the device's own sampled actions (teacher-forced replay)
    su_table = torch.tensor(tables.flags(tables.ACTION_HAS_SELECTED_UNITS))
    ref = R.compute_logp_action(sd, noise=noise, su_table=su_table, **clone_tree(data))
    if torch.equal(ref['action_info']['action_type'], out['action_info']['action_type'].cpu()):
        assert torch.equal(ref['selected_units_num'] > 0, out['selected_units_num'].cpu() > 0)


@gpu
def test_cuda_graph_learner_step_matches_eager():
    """The captured CUDA graph of the learner step (forward, loss, backward, clip+Adam) replays the same kernels as the
    eager step.  Adam with beta1 = 0 moves every weight by ~lr*sign(g), so last-bit differences of the library
    convolutions still in the front of the spatial encoder can flip individual near-zero gradients; the comparison
    is therefore on the step count, the loss trajectory (1e-3 relative) and the relative L2 distance of the weights."""
    from applestar_b200 import Model, ReinforcementLoss
    from applestar_b200.learner import LearnerStep
    sd = synth.full_state_dict(load_spec('winloss'), 0)
    data = synth.tree_map(lambda t: t.cuda(), synth.gen_trajectories(3, 2, 48, 9))
    steps = {}
    for use_graph in (False, True):
        m = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
        m.load_state_dict(sd, strict=False)
        m = m.cuda()
        step = LearnerStep(m, ReinforcementLoss({}, 'MP0'), lr=1e-4, use_graph=use_graph)
        for _ in range(3):        # the capture's warm-up leaves no trace: 3 calls = 3 optimizer updates either way
            step(data)
        steps[use_graph] = step
    eager_step, graph_step = steps[False], steps[True]
    assert graph_step.use_graph and graph_step.launches_per_step > 100
    assert int(graph_step.optimizer.t_dev.item()) == int(eager_step.optimizer.t_dev.item()) == 3
    w_e, w_g = eager_step.flat.flat, graph_step.flat.flat
    assert ((w_g - w_e).norm() / w_e.norm()).item() < 1e-4
    l_e = float(eager_step(data)['total_loss'])
    l_g = float(graph_step(data)['total_loss'])
    assert abs(l_e - l_g) <= 1e-3 * max(1.0, abs(l_e)), (l_e, l_g)
    info = graph_step(data)
    assert 'winloss/td' in info and isinstance(info['kl/total'], float)


@gpu
def test_shadow_follows_load_state_dict_and_torch_optimizer():
    """The bf16 weight shadow of a DistModule-wrapped model must follow every way the reference changes weights:
    `load_state_dict` after wrapping (checkpoint hook, value reset: rl_learner.py:228) and an in-place torch
    optimizer step (the documented drop-in mode keeps the reference's own Adam)."""
    from applestar_b200 import Model
    from applestar_b200.learner import DistModule
    spec = load_spec('winloss')
    data = synth.tree_map(lambda t: t.cuda(), synth.gen_trajectories(2, 2, 48, 9))

    def values_of(model):
        with torch.no_grad():
            return model.rl_learner_forward(**data)['value']['winloss'].clone()

    sd0, sd1 = synth.full_state_dict(spec, 0), synth.full_state_dict(spec, 1)
    fresh = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
    fresh.load_state_dict(sd1, strict=False)
    want = values_of(fresh.cuda())
    m = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
    m.load_state_dict(sd0, strict=False)
    dm = DistModule(m.cuda())
    before = values_of(m)
    dm.load_state_dict(sd1, strict=False)                     # after wrapping, as the reference does
    after = values_of(m)
    assert (before - want).abs().max().item() > 1e-3
    assert torch.equal(after, want)
    opt = torch.optim.SGD(m.parameters(), lr=0.05)
    opt.zero_grad()
    out = m.rl_learner_forward(**data)
    out['value']['winloss'].sum().backward()
    dm.sync_gradients()
    opt.step()
    moved = values_of(m)
    ref = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
    ref.load_state_dict({k: v.detach().cpu() for k, v in m.state_dict().items()}, strict=False)
    assert (moved - after).abs().max().item() > 1e-4          # the step is visible in the forward pass ...
    assert torch.equal(moved, values_of(ref.cuda()))          # ... and equals a freshly built model with those weights


@gpu
def test_actor_step_graph_matches_eager():
    """learner.ActorStep (compute_logp_action replayed from a CUDA graph, 64 pointer steps inside, trim outside) returns
    the eager call's tensors: same kernels on the same inputs, so sampled actions and step count are bit-exact."""
    from applestar_b200.learner import ActorStep
    model, _ = _model(['winloss'], 'winloss')
    model.eval()
    B, N = 6, 48
    dev = torch.device('cuda')
    outs = []
    actor = ActorStep(model)
    for seed in (3, 4):                      # second call replays the graph captured by the first
        data = synth.tree_map(lambda t: t.to(dev), synth.gen_observations(B, N, seed, entity_num_range=(N // 2, N)))
        noise = synth.tree_map(lambda t: t.to(dev), synth.gen_gumbel(B, seed + 10, N))
        with torch.no_grad():
            eager = model.compute_logp_action(gumbel_noise=noise, **data)
        graphed = actor(data['spatial_info'], data['entity_info'], data['scalar_info'], data['entity_num'],
                        data['hidden_state'], gumbel_noise=noise)
        outs.append((eager, graphed))
        assert 'selected_units_steps' not in graphed
        for k in eager['action_info']:
            assert torch.equal(eager['action_info'][k], graphed['action_info'][k]), k
            assert eager['logit'][k].shape == graphed['logit'][k].shape, k
            assert (eager['logit'][k] - graphed['logit'][k]).abs().max().item() < 1e-3, k
        assert torch.equal(eager['selected_units_num'], graphed['selected_units_num'])
        assert torch.equal(eager['extra_units'], graphed['extra_units'])
        for (h0, c0), (h1, c1) in zip(eager['hidden_state'], graphed['hidden_state']):
            assert (h0 - h1).abs().max().item() < 1e-4 and (c0 - c1).abs().max().item() < 1e-4
    assert actor.launches_per_call > 100 and len(actor._graphs) == 1
    bad = synth.tree_map(lambda t: t.to(dev), synth.gen_observations(B, N, 5, entity_num_range=(N // 2, N)))
    bad['entity_info']['unit_type'][0, 0] = -1
    if bad['entity_info']['unit_type'].dtype.is_signed:
        with pytest.raises(RuntimeError):
            actor(bad['spatial_info'], bad['entity_info'], bad['scalar_info'], bad['entity_num'], bad['hidden_state'],
                  gumbel_noise=noise)


@gpu
@pytest.mark.parametrize('R,N,temperature', [(7, 48, 0.8), (150, 512, 1.0)])
def test_su_sample_fused_kernel_vs_oracle(R, N, temperature):
    """as_su_sample (all 64 sampling steps in one launch, fp32) against the oracle's step loop on the same keys, noise
    and weights: picks, selected_units_num, step count and extra_units bit-exact; logits to fp32 round-off."""
    from applestar_b200 import ops
    from applestar_b200.model import SelectedUnitsHead
    torch.manual_seed(11)
    head = SelectedUnitsHead(True).cuda()
    with torch.no_grad():
        for prm in head.lstm.parameters():
            if prm.dim() == 2:
                prm.mul_(0.3)                                  # keep the pointer logits in a realistic range
    sd = {'policy.selected_units_head.' + k: v.detach().cpu() for k, v in head.state_dict().items()}
    ent = torch.randn(R, N, 256) * 0.5
    en = torch.randint(max(N // 3, 2), N + 1, (R,))
    en[0], en[1] = N, 2
    su_mask = torch.rand(R) < 0.8
    su_mask[0], su_mask[2] = True, False
    ae0 = torch.randn(R, 1024) * 0.3
    noise = synth.gen_gumbel(R, 21, N)['selected_units']
    if R < 10:
        noise[:, 3, :] = -50.0
        noise[1:, 3, int(en[3])] = 0.0                     # row 3 takes the end token at step 1; others run on
    ref_logits, ref_units, ref_ae, ref_num, ref_extra = R_.su_head_sample(sd, ae0, ent, en, su_mask, noise, temperature, True)
    key, _ = R_.su_keys(sd, ent, en)
    with torch.no_grad():
        logits, units, ae, num, extra = ops.selected_units_sample(head, key.cuda(), ae0.cuda(), en.cuda(), su_mask.cuda(),
                                                                  noise.cuda(), temperature, True)
    assert logits.shape == ref_logits.shape, (logits.shape, ref_logits.shape)
    assert torch.equal(units.cpu(), ref_units)
    assert torch.equal(num.cpu(), ref_num)
    assert torch.equal(extra.cpu(), ref_extra)
    masked = ref_logits < -1e8
    assert torch.equal(logits.cpu() < -1e8, masked)
    assert (logits.cpu() - ref_logits)[~masked].abs().max().item() < 2e-3
    assert (ae.cpu() - ref_ae).abs().max().item() < 0.05      # the final embed_fc1/2 run on the bf16 GEMM
    with torch.no_grad():                                   # the step-by-step device loop agrees on every pick too
        import os
        os.environ['APPLESTAR_SU_FUSED'] = '0'
        try:
            l2, u2, _, n2, e2 = ops.selected_units_sample(head, key.cuda(), ae0.cuda(), en.cuda(), su_mask.cuda(),
                                                          noise.cuda(), temperature, True)
        finally:
            del os.environ['APPLESTAR_SU_FUSED']
    assert l2.shape == logits.shape and torch.equal(n2 > 0, num > 0)


@gpu
def test_grad_pack_matches_accumulate(monkeypatch):
    """FlatParams pack mode (p.grad = None, one as_pack_f32 launch per bucket) fills the flat gradient buffer exactly like
    accumulation into views does, over two steps, including a parameter that gets no gradient (zeros, not stale data)."""
    import torch.nn as nn
    from applestar_b200.learner import DistModule, FlatParams
    flats = []
    for mode in ('1', '0'):
        monkeypatch.setenv('APPLESTAR_GRAD_PACK', mode)
        torch.manual_seed(3)
        m = nn.ModuleDict({'a': nn.Linear(70001, 3), 'b': nn.Linear(3, 5), 'unused': nn.Linear(4, 4)}).cuda()
        dm = DistModule(m, flat=FlatParams(m))
        assert dm.flat.pack_mode == (mode == '1')
        for step in range(2):
            dm.flat.zero_grad()
            x = torch.randn(6, 70001, device='cuda', generator=torch.Generator('cuda').manual_seed(step))
            y = m['b'](m['a'](x))
            if step == 0:
                y = y.sum() + m['unused'](torch.ones(2, 4, device='cuda')).sum()
            y.sum().backward()
            dm.sync_gradients()
            for p, o in zip(dm.flat.params, dm.flat.offsets):
                assert p.grad.data_ptr() == dm.flat.grad.data_ptr() + 4 * o       # views again after the sync
        flats.append(dm.flat.grad.clone())
    assert torch.equal(flats[0], flats[1])
    assert float(flats[0].abs().sum()) > 0


@gpu
def test_selected_units_batched_matches_step_loop():
    """ops.selected_units_train (all steps at once) vs the step-by-step loop: logits, final embedding, gradients."""
    from applestar_b200 import ops
    from applestar_b200.model import SelectedUnitsHead
    torch.manual_seed(5)
    R, N, S = 6, 40, 7
    head = SelectedUnitsHead(False).cuda()
    key0 = torch.randn(R, N + 1, 32, device='cuda')
    ae0 = torch.randn(R, 1024, device='cuda') * 0.1
    en = torch.tensor([40, 12, 30, 5, 40, 22], device='cuda')
    sel = torch.zeros(R, 64, dtype=torch.long, device='cuda')
    num = torch.tensor([7, 3, 0, 2, 5, 1], device='cuda')
    for r in range(R):
        k = int(num[r])
        perm = torch.randperm(int(en[r]))[:max(k - 1, 0)]
        sel[r, :len(perm)] = perm.cuda()
        sel[r, len(perm):] = en[r]                    # end token, repeated
    sel[4, 1] = sel[4, 0]                              # a duplicate pick (one-hot maximum counts it once)
    outs = []
    for fn in (ops.selected_units_train, ops.selected_units_train_loop):
        ops.SU_STEPS_HINT = S
        key = key0.clone().requires_grad_(True)
        ae = ae0.clone().requires_grad_(True)
        for p in head.parameters():
            p.grad = None
        logits, _, ae_out, _, _ = fn(head, key, ae, en, sel, num, 0.8)
        valid = logits > -1e8
        g1 = torch.randn(logits.shape, device='cuda', generator=torch.Generator('cuda').manual_seed(1))
        g2 = torch.randn(ae_out.shape, device='cuda', generator=torch.Generator('cuda').manual_seed(2))
        ((logits * g1 * valid).sum() + (ae_out * g2).sum()).backward()
        outs.append((logits.detach(), ae_out.detach(), key.grad.clone(), ae.grad.clone(),
                     head.query_fc1[0].weight.grad.clone()))
    ops.SU_STEPS_HINT = None
    a, b = outs
    assert torch.equal(a[0] > -1e8, b[0] > -1e8)                       # masks bit-exact
    v = a[0] > -1e8
    assert (a[0][v] - b[0][v]).abs().max().item() < 0.05               # bf16 GEMMs batched differently
    nan_ok = torch.isnan(b[1]) == torch.isnan(a[1])
    assert bool(nan_ok.all())
    fin = ~torch.isnan(b[1])
    assert (a[1][fin] - b[1][fin]).abs().max().item() < 0.05
    for x, y in zip(a[2:], b[2:]):
        fin = ~(torch.isnan(x) | torch.isnan(y))
        assert (x[fin] - y[fin]).abs().max().item() < 0.08 * y[fin].abs().max().item() + 1e-3


@gpu
def test_su_sample_rows_ending_at_very_different_steps():
    """The sampling kernel leaves its loop early once EVERY row of the batch has ended.  CTAs (1-2 rows each) are not in
    step with each other, so the condition must hold per CTA: here most rows end at steps 0-3 and one row at step 40;
    every row must still carry the reference's logits / picks for all 41 steps (bit-exact picks vs the oracle loop)."""
    from applestar_b200 import ops
    from applestar_b200.model import SelectedUnitsHead
    torch.manual_seed(12)
    head = SelectedUnitsHead(True).cuda()
    with torch.no_grad():
        for prm in head.lstm.parameters():
            if prm.dim() == 2:
                prm.mul_(0.3)
    sd = {'policy.selected_units_head.' + k: v.detach().cpu() for k, v in head.state_dict().items()}
    R, N = 301, 128                                       # > 148 rows: two rows per CTA
    ent = torch.randn(R, N, 256) * 0.5
    en = torch.randint(60, N + 1, (R,))
    su_mask = torch.ones(R, dtype=torch.bool)
    su_mask[5] = False
    ae0 = torch.randn(R, 1024) * 0.3
    noise = synth.gen_gumbel(R, 33, N)['selected_units']          # [S, R, N + 1]
    end_step = torch.randint(0, 4, (R,))
    end_step[137] = 40
    for r in range(R):
        noise[:, r, int(en[r])] = -1e4                            # never the end token ...
        noise[int(end_step[r]) + 1, r, int(en[r])] = 1e4          # ... except at its step (the token is selectable from step 1)
    ref_logits, ref_units, _, ref_num, ref_extra = R_.su_head_sample(sd, ae0, ent, en, su_mask, noise, 1.0, True)
    key, _ = R_.su_keys(sd, ent, en)
    with torch.no_grad():
        logits, units, _, num, extra = ops.selected_units_sample(head, key.cuda(), ae0.cuda(), en.cuda(), su_mask.cuda(),
                                                                 noise.cuda(), 1.0, True)
    assert int(ref_num.max()) == 42 and logits.shape == ref_logits.shape
    assert torch.equal(units.cpu(), ref_units) and torch.equal(num.cpu(), ref_num) and torch.equal(extra.cpu(), ref_extra)
    masked = ref_logits < -1e8
    assert torch.equal(logits.cpu() < -1e8, masked)
    assert (logits.cpu() - ref_logits)[~masked].abs().max().item() < 2e-3


@gpu
def test_fused_adam_follows_torch_adam_for_parameters_without_gradient():
    """torch.optim.Adam skips a parameter whose .grad is None and keeps one step count per parameter; the reference relies
    on it when the policy is frozen during value pre-training and trained afterwards (rl_learner.py:147-172).  FusedAdam
    (as_clip_adam_per_param) must give the same weights: here layer `b` gets no gradient for 3 steps, then trains."""
    import torch.nn as nn
    from applestar_b200.learner import DistModule, FlatParams, FusedAdam
    torch.manual_seed(5)

    def make():
        torch.manual_seed(5)
        return nn.ModuleDict({'a': nn.Linear(40, 24), 'b': nn.Linear(24, 8), 'c': nn.Linear(40, 8)}).cuda()
    mine, ref = make(), make()
    dm = DistModule(mine, flat=FlatParams(mine))
    opt = FusedAdam(dm.flat, lr=1e-3, betas=(0.0, 0.99), eps=1e-5, max_norm=1.0)
    ropt = torch.optim.Adam(ref.parameters(), lr=1e-3, betas=(0.0, 0.99), eps=1e-5)
    g = torch.Generator(device='cuda').manual_seed(1)
    for step in range(8):
        x = torch.randn(16, 40, device='cuda', generator=g)
        for model, is_mine in ((mine, True), (ref, False)):
            if is_mine:
                dm.flat.zero_grad()
            else:
                ropt.zero_grad(set_to_none=True)
            h = model['a'](x)
            out = model['c'](x).pow(2).sum() if step < 3 else (model['b'](torch.relu(h)).pow(2).sum() + model['c'](x).pow(2).sum())
            out.backward()
            if is_mine:
                dm.sync_gradients()
                opt.step()
            else:
                torch.nn.utils.clip_grad_norm_([p for p in ref.parameters() if p.grad is not None], 1.0)
                ropt.step()
    assert opt.p_step is not None
    steps = {n: int(s) for n, s in zip(dm.flat.names, opt.p_step.tolist())}
    assert steps['c.weight'] == 8 and steps['b.weight'] == 5 and steps['a.weight'] == 5, steps
    for (n, p), (_, q) in zip(mine.named_parameters(), ref.named_parameters()):
        assert torch.allclose(p, q, rtol=2e-5, atol=2e-7), (n, float((p - q).abs().max()))
