"""TEST INFRASTRUCTURE (container only): runs the UNMODIFIED reference on seeded synthetic inputs and
writes small golden fixtures to tests/golden/. Run: python oracle/make_golden.py

The fixtures pin oracle/restate.py (and through it the CUDA path) to the executed reference:
  state_dict_spec.json   keys / shapes / trainable flags of the reference Model (1 and 4 baselines)
  learner_small.pt       rl_learner_forward + compute_loss + backward, T=2 B=2 N=64, fp32 (+fp64 truth)
  infer_small.pt         compute_logp_action under fixed Gumbel noise, B=3 N=512 (incl. collision / 1-entity obs)
  learner_n128.pt        the learner case at N=128 entities, one baseline (the size from which the fused tcgen05 attention
                         forward / backward kernels run); written by `python oracle/make_golden.py n128` alone
  learner_n512.pt        the learner case at the benchmarked N=512 (T=4, B=2, special observations) with ~300 reference
                         gradient tensors and the reference's own bf16-autocast gradient error; `... make_golden.py n512`
  scatter_small.pt       scatter_connection on a heavy-collision case
  sl_loss_small.pt       SupervisedLoss.compute_loss (log entries + logit gradients, three configurations); `... sl_loss`
Inputs and weights are NOT stored: tests regenerate them from the same seeds (applestar_b200/synth.py).
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref_loader import load_reference  # noqa: E402
from applestar_b200 import synth, config as C, tables  # noqa: E402

GOLD = os.path.join(ROOT, 'tests', 'golden')


def ref_cfg(baselines, use_value_feature=False):
    from easydict import EasyDict
    learner = dict(C.DEFAULT_LOSS_CONFIG)
    learner.update({'use_value_feature': use_value_feature, 'player_id': 'MP0'})
    return EasyDict({'common': {'type': 'rl'}, 'agent': {'extra_units': True},
                     'model': {'enable_baselines': list(baselines)}, 'learner': learner})


def build_ref_model(baselines, seed=0, use_value_feature=False):
    from distar.agent.default.model.model import Model
    cfg = ref_cfg(baselines, use_value_feature)
    m = Model(cfg, use_value_network=True)
    trainable = {k for k, p in m.named_parameters() if p.requires_grad}
    spec = [(k, list(v.shape), k in trainable) for k, v in m.state_dict().items()]
    sd = synth.synth_state_dict([(k, s) for k, s, t in spec if t], seed)
    missing = m.load_state_dict(sd, strict=False)
    assert not missing.unexpected_keys
    return m, cfg, spec


def grad_summary(model):
    out = {}
    for k, p in model.named_parameters():
        if p.requires_grad:
            g = p.grad if p.grad is not None else torch.zeros_like(p)
            out[k] = torch.tensor([g.double().norm().item(), g.double().sum().item()])
    return out


SMALL_GRADS = ['policy.selected_units_head.end_embedding', 'policy.location_head.res.0.UpdateSP',
               'core_lstm.layers.2.cell.layernorm_c.weight', 'encoder.scatter_project.0.bias',
               'policy.target_unit_head.query_fc2.0.weight', 'value_networks.winloss.value_fc.0.weight',
               'encoder.entity_encoder.transformer.layers.0.layernorm1.bias',
               'encoder.spatial_encoder.project.0.bias', 'policy.action_type_head.action_fc.layer2.0.bias',
               'encoder.scalar_encoder.encode_modules.home_race.weight']


def learner_case(T, B, N, baselines, seed):
    from distar.agent.default.rl_training.rl_loss import ReinforcementLoss
    model, cfg, _ = build_ref_model(baselines)
    data = synth.gen_trajectories(T, B, N, seed)
    loss_fn = ReinforcementLoss(cfg.learner, 'MP0')
    out = model.rl_learner_forward(**synth.tree_map(lambda t: t.clone(), data))
    logits = {k: v.detach().clone() for k, v in out['target_logit'].items()}
    values = {k: v.detach().clone() for k, v in out['value'].items()}
    info = loss_fn.compute_loss(out)
    model.zero_grad()
    info['total_loss'].backward()
    res = {'T': T, 'B': B, 'N': N, 'seed': seed, 'baselines': list(baselines),
           'logit': logits, 'value': values,
           'loss_info': {k: (float(v) if not isinstance(v, torch.Tensor) else float(v.detach())) for k, v in info.items()},
           'grad_summary': grad_summary(model),
           'grads': {k: dict(model.named_parameters())[k].grad.clone() for k in SMALL_GRADS}}
    # fp64 ground truth of the same forward -> calibrates fp32/bf16 tolerances
    old_float = torch.Tensor.float
    torch.set_default_dtype(torch.float64)
    torch.Tensor.float = lambda self, *a, **k: self.double()
    try:
        m64 = model.double()
        d64 = synth.tree_map(lambda t: t.double() if t.dtype == torch.float32 else t.clone(), data)
        with torch.no_grad():
            o64 = m64.rl_learner_forward(**d64)
        res['logit64'] = {k: v.clone() for k, v in o64['target_logit'].items()}
        res['value64'] = {k: v.clone() for k, v in o64['value'].items()}
    finally:
        torch.Tensor.float = old_float
        torch.set_default_dtype(torch.float32)
    # bf16-autocast reference error level: what "bf16 tolerance" means per head
    model = model.float()
    import distar.agent.default.model.encoder as enc_mod
    orig_sc = enc_mod.scatter_connection
    enc_mod.scatter_connection = lambda shape, pe, loc, dim, typ='add': orig_sc(shape, pe.float(), loc, dim, typ)
    try:
        with torch.no_grad(), torch.autocast('cpu', dtype=torch.bfloat16):
            ob = model.rl_learner_forward(**synth.tree_map(lambda t: t.clone(), data))
        err = {}
        for k in logits:
            valid = res['logit64'][k] > -1e8
            d = (ob['target_logit'][k].double() - res['logit64'][k])[valid]
            err[k] = torch.tensor([d.abs().max().item(), d.pow(2).mean().sqrt().item()])
        for k in values:
            d = ob['value'][k].double() - res['value64'][k]
            err['value/' + k] = torch.tensor([d.abs().max().item(), d.pow(2).mean().sqrt().item()])
        res['bf16_autocast_err'] = err
    finally:
        enc_mod.scatter_connection = orig_sc
    return res


BIG_GRADS = [   # tensors above FULL_GRAD_NUMEL whose reference gradient is kept (strided sample when above BIG_GRAD_CAP)
    'encoder.entity_encoder.transformer.embedding.0.weight',
    'encoder.entity_encoder.transformer.layers.0.attention.attention_pre.0.weight',
    'encoder.entity_encoder.transformer.layers.1.attention.project.0.weight',
    'encoder.entity_encoder.transformer.layers.2.mlp.0.0.weight',
    'encoder.entity_encoder.transformer.layers.2.mlp.1.0.weight',
    'encoder.entity_encoder.entity_fc.0.weight', 'encoder.entity_encoder.embed_fc.0.weight',
    'encoder.spatial_encoder.downsample.0.0.weight', 'encoder.spatial_encoder.downsample.1.0.weight',
    'encoder.spatial_encoder.downsample.2.0.weight', 'encoder.spatial_encoder.res.0.conv1.0.weight',
    'encoder.spatial_encoder.res.3.conv2.0.weight', 'encoder.spatial_encoder.fc.0.weight',
    'encoder.scalar_encoder.encode_modules.unit_counts_bow.0.weight',
    'encoder.scalar_encoder.encode_modules.beginning_order.transformer.embedding.0.weight',
    'core_lstm.layers.0.cell.weight_ih', 'core_lstm.layers.0.cell.weight_hh', 'core_lstm.layers.1.cell.weight_ih',
    'core_lstm.layers.2.cell.weight_hh',
    'policy.action_type_head.project.0.weight', 'policy.action_type_head.action_fc.layer2.0.weight',
    'policy.action_type_head.action_map_fc1.0.weight', 'policy.action_type_head.glu1.layer2.0.weight',
    'policy.action_type_head.glu2.layer1.0.weight',
    'policy.delay_head.fc1.0.weight', 'policy.delay_head.fc3.0.weight', 'policy.delay_head.embed_fc2.0.weight',
    'policy.queued_head.fc1.0.weight', 'policy.queued_head.embed_fc2.0.weight',
    'policy.selected_units_head.query_fc1.0.weight', 'policy.selected_units_head.embed_fc2.0.weight',
    'policy.target_unit_head.query_fc1.0.weight',
    'policy.location_head.project_embed.0.weight', 'policy.location_head.conv1.0.weight',
    'policy.location_head.res.0.conv1.0.weight', 'policy.location_head.res.1.GateWeightG.0.0.weight',
    'policy.location_head.res.3.conv2.0.weight', 'policy.location_head.res.3.GateWeightG.3.0.weight',
    'policy.location_head.upsample.0.0.weight', 'policy.location_head.upsample.1.0.weight',
    'value_networks.winloss.project.0.weight', 'value_networks.winloss.res.0.fc1.0.weight',
    'value_networks.winloss.res.15.fc2.0.weight']
FULL_GRAD_NUMEL = 8192      # every trainable tensor up to this size keeps its full reference gradient
BIG_GRAD_CAP = 65536


def grad_sample_index(numel):
    """Flat indices of the stored gradient sample: everything up to BIG_GRAD_CAP elements, else a fixed odd stride."""
    if numel <= BIG_GRAD_CAP:
        return None
    stride = (numel // BIG_GRAD_CAP) | 1
    return torch.arange(0, numel, stride)[:BIG_GRAD_CAP]


def _kept_grads(model, everything=False):
    out = {}
    for k, p in model.named_parameters():
        if not p.requires_grad or (not everything and p.numel() > FULL_GRAD_NUMEL and k not in BIG_GRADS):
            continue
        if everything:
            out[k] = (p.grad if p.grad is not None else torch.zeros_like(p)).detach().float().reshape(-1).clone()
            continue
        g = (p.grad if p.grad is not None else torch.zeros_like(p)).detach().float().reshape(-1)
        idx = grad_sample_index(g.numel())
        out[k] = (g if idx is None else g[idx]).clone()
    return out


def learner_n512_case(T=4, B=2, N=512, seed=5120, budgets_only=False):
    """The benchmarked entity count (N = 512: four key blocks in the fused attention, 512-key scatter sort, cluster LSTM)
    with the all-on-one-cell and entity_num == 1 observations: reference fp32 logits / values / 63 loss scalars /
    gradients of ~300 tensors (every tensor <= 8192 elements in full + BIG_GRADS), the fp64 forward, and the
    reference's OWN bf16-autocast error for every logit head (max-abs, RMS) and every kept gradient (relative L2,
    cosine) as the tolerance budget of a bf16 implementation."""
    from distar.agent.default.rl_training.rl_loss import ReinforcementLoss
    baselines = ['winloss']
    model, cfg, _ = build_ref_model(baselines)
    data = synth.gen_trajectories(T, B, N, seed, special=True)
    loss_fn = ReinforcementLoss(cfg.learner, 'MP0')
    out = model.rl_learner_forward(**synth.tree_map(lambda t: t.clone(), data))
    logits = {k: v.detach().clone() for k, v in out['target_logit'].items()}
    values = {k: v.detach().clone() for k, v in out['value'].items()}
    info = loss_fn.compute_loss(out)
    model.zero_grad()
    info['total_loss'].backward()
    grads = _kept_grads(model)
    all_grads = _kept_grads(model, everything=True)
    res = {'T': T, 'B': B, 'N': N, 'seed': seed, 'baselines': baselines, 'special': True,
           'logit': logits, 'value': values,
           'loss_info': {k: (float(v) if not isinstance(v, torch.Tensor) else float(v.detach())) for k, v in info.items()},
           'grad_summary': grad_summary(model), 'grads': grads,
           'grad_rule': {'full_numel': FULL_GRAD_NUMEL, 'cap': BIG_GRAD_CAP, 'big': list(BIG_GRADS)}}
    old_float = torch.Tensor.float
    torch.set_default_dtype(torch.float64)
    torch.Tensor.float = lambda self, *a, **k: self.double()
    try:
        m64 = model.double()
        d64 = synth.tree_map(lambda t: t.double() if t.dtype == torch.float32 else t.clone(), data)
        with torch.no_grad():
            o64 = m64.rl_learner_forward(**d64)
        res['logit64'] = {k: v.clone() for k, v in o64['target_logit'].items()}
        res['value64'] = {k: v.clone() for k, v in o64['value'].items()}
    finally:
        torch.Tensor.float = old_float
        torch.set_default_dtype(torch.float32)
    model = model.float()
    import distar.agent.default.model.encoder as enc_mod
    orig_sc = enc_mod.scatter_connection
    enc_mod.scatter_connection = lambda shape, pe, loc, dim, typ='add': orig_sc(shape, pe.float(), loc, dim, typ)
    try:
        model.zero_grad()
        with torch.autocast('cpu', dtype=torch.bfloat16):
            ob = model.rl_learner_forward(**synth.tree_map(lambda t: t.clone(), data))
            ob = dict(ob)
            ob['target_logit'] = {k: v.float() for k, v in ob['target_logit'].items()}
            ob['value'] = {k: v.float() for k, v in ob['value'].items()}
        ib = loss_fn.compute_loss(ob)
        ib['total_loss'].backward()
        err = {}
        for k in logits:
            valid = res['logit64'][k] > -1e8
            d = (ob['target_logit'][k].detach().double() - res['logit64'][k])[valid]
            err[k] = torch.tensor([d.abs().max().item(), d.pow(2).mean().sqrt().item()])
        for k in values:
            d = ob['value'][k].detach().double() - res['value64'][k]
            err['value/' + k] = torch.tensor([d.abs().max().item(), d.pow(2).mean().sqrt().item()])
        res['bf16_autocast_err'] = err
        gb = _kept_grads(model, everything=True)
        gerr = {}                           # every trainable tensor: [relative L2, cosine] of autocast-bf16 vs fp32
        for k, g in all_grads.items():
            a, b = gb[k].double(), g.double()
            rel = ((a - b).norm() / b.norm().clamp(min=1e-30)).item()
            cos = (torch.dot(a, b) / (a.norm() * b.norm()).clamp(min=1e-30)).item()
            gerr[k] = torch.tensor([rel, cos])
        res['bf16_autocast_grad_err'] = gerr
        res['bf16_autocast_loss_info'] = {k: (float(v) if not isinstance(v, torch.Tensor) else float(v.detach()))
                                          for k, v in ib.items()}
    finally:
        enc_mod.scatter_connection = orig_sc
    if budgets_only:        # the GPU test recomputes the fp32 side with the oracle on the box's CPU
        for k in ('logit', 'value', 'logit64', 'value64', 'grads'):
            del res[k]
    return res


def sl_loss_case(b=6, s=5, N=24, seed=77):
    """SupervisedLoss.compute_loss of the executed reference (sl_loss.py:102-121) on synth.gen_sl_loss_inputs: every log
    entry for label_smooth off / on and su_mask on / off, and d total_loss / d logits for the default configuration."""
    from easydict import EasyDict
    from distar.agent.default.sl_training.sl_loss import SupervisedLoss
    out = {'meta': dict(b=b, s=s, N=N, seed=seed)}
    for tag, extra in (('default', {}), ('label_smooth', {'label_smooth': True}), ('no_su_mask', {'su_mask': False})):
        d = synth.gen_sl_loss_inputs(b, s, N, seed)
        logits = {k: v.clone().requires_grad_(True) for k, v in d['logits'].items()}
        loss = SupervisedLoss(EasyDict({'learner': dict(extra)}))
        info = loss.compute_loss(logits, d['actions'], d['actions_mask'], d['selected_units_num'], d['entity_num'],
                                 {'selected_units': d['infer_selected_units']})
        info['total_loss'].backward()
        out[tag] = {'info': {k: float(v) for k, v in info.items()},
                    'dlogits': {k: v.grad.clone() for k, v in logits.items() if k != 'target_location'},
                    'dlogits_location_sum_abs': float(logits['target_location'].grad.abs().sum()),
                    'dlogits_location_row0': logits['target_location'].grad[0].clone()}
    return out


def infer_case(B, N, seed):
    model, cfg, _ = build_ref_model(['winloss'])
    model.eval()
    data = synth.gen_observations(B, N, seed, entity_num_range=(N // 2, N), special=True)
    noise = synth.gen_gumbel(B, seed + 1, N)
    order = [noise['action_type'], noise['delay'], noise['queued']]
    state = {'i': 0, 'su': 0}

    def fake_multinomial(p, n, *a, **k):
        K = p.shape[1]
        if K == N + 1:
            g = noise['selected_units'][state['su']]
            state['su'] += 1
        elif K == N:
            g = noise['target_unit']
        elif K == C.SPATIAL_X * C.SPATIAL_Y:
            g = noise['target_location']
        else:
            g = order[state['i']]
            state['i'] += 1
        return (torch.log(p) + g).argmax(1, keepdim=True)
    real = torch.multinomial
    torch.multinomial = fake_multinomial
    try:
        with torch.no_grad():
            out = model.compute_logp_action(**synth.tree_map(lambda t: t.clone(), data))
    finally:
        torch.multinomial = real
    return {'B': B, 'N': N, 'seed': seed, 'logit': out['logit'], 'action_info': out['action_info'],
            'action_logp': out['action_logp'], 'selected_units_num': out['selected_units_num'],
            'extra_units': out['extra_units'], 'hidden_state': out['hidden_state']}


def scatter_case():
    from distar.agent.default.model.module_utils import scatter_connection
    g = torch.Generator().manual_seed(5)
    R, N, D = 4, 512, 32
    emb = torch.randn(R, N, D, generator=g)
    x = torch.randint(0, 8, (R, N), generator=g)
    y = torch.randint(0, 8, (R, N), generator=g)
    x[1, :50] = 250          # out-of-range -> clamps to 159
    y[1, :50] = 200          # -> 151
    loc = torch.stack([x, y], -1)
    out = {}
    for typ in ('add', 'cover'):
        m = scatter_connection((R, 152, 160), emb.clone(), loc.clone(), D, typ)
        nz = m.permute(0, 2, 3, 1).reshape(R, -1, D)
        cells = (nz.abs().sum(-1) > 0).nonzero()
        out[typ] = {'cells': cells, 'vals': nz[cells[:, 0], cells[:, 1]].clone()}
    return {'seed': 5, 'R': R, 'N': N, 'D': D, 'out': out}


def main():
    load_reference()
    os.makedirs(GOLD, exist_ok=True)
    torch.manual_seed(0)
    torch.set_num_threads(8)
    if len(sys.argv) > 1 and sys.argv[1] == 'n128':     # additional fixture only; the others stay byte-identical
        torch.save(learner_case(2, 2, 128, ['winloss'], 4242), os.path.join(GOLD, 'learner_n128.pt'))
        print('learner_n128.pt', os.path.getsize(os.path.join(GOLD, 'learner_n128.pt')))
        return
    if len(sys.argv) > 1 and sys.argv[1] == 'n512':     # the benchmarked entity count; written alone like n128
        torch.save(learner_n512_case(), os.path.join(GOLD, 'learner_n512.pt'))
        print('learner_n512.pt', os.path.getsize(os.path.join(GOLD, 'learner_n512.pt')))
        return
    if len(sys.argv) > 1 and sys.argv[1] == 'spec_vf':   # state_dict contract with learner.use_value_feature (ValueEncoder); the
        path = os.path.join(GOLD, 'state_dict_spec.json')  # existing entries of the file are kept as they are
        with open(path) as f:
            specs = json.load(f)
        _, _, spec = build_ref_model(['winloss'], use_value_feature=True)
        specs['winloss_value_feature'] = spec
        with open(path, 'w') as f:
            json.dump(specs, f)
        print('state_dict_spec.json: winloss_value_feature', len(spec), 'entries')
        return
    if len(sys.argv) > 1 and sys.argv[1] == 'sl_loss':  # SupervisedLoss
        torch.save(sl_loss_case(), os.path.join(GOLD, 'sl_loss_small.pt'))
        print('sl_loss_small.pt', os.path.getsize(os.path.join(GOLD, 'sl_loss_small.pt')))
        return
    if len(sys.argv) > 1 and sys.argv[1] == 'n512t8':   # tolerance budgets (only) of the T = 8 all-gradients test
        torch.save(learner_n512_case(8, 2, 512, 8512, budgets_only=True), os.path.join(GOLD, 'learner_n512_t8_budget.pt'))
        print('learner_n512_t8_budget.pt', os.path.getsize(os.path.join(GOLD, 'learner_n512_t8_budget.pt')))
        return
    specs = {}
    for name, bl in (('winloss', ['winloss']), ('four', ['winloss', 'build_order', 'built_unit', 'battle'])):
        _, _, spec = build_ref_model(bl)
        specs[name] = spec
    with open(os.path.join(GOLD, 'state_dict_spec.json'), 'w') as f:
        json.dump(specs, f)
    torch.save(scatter_case(), os.path.join(GOLD, 'scatter_small.pt'))
    torch.save(learner_case(2, 2, 64, ['winloss', 'build_order', 'built_unit', 'battle'], 1234),
               os.path.join(GOLD, 'learner_small.pt'))
    torch.save(infer_case(3, 512, 4321), os.path.join(GOLD, 'infer_small.pt'))
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == '__main__':
    main()
