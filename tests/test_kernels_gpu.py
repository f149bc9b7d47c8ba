"""GPU parity of the HBM-bound kernels against the oracle (bit-exact for index / scatter work)."""
import pytest
import torch

from applestar_b200 import config as C, synth
from oracle import restate as R
from util import golden

pytestmark = pytest.mark.gpu


def _scatter_inputs(g):
    gen = torch.Generator().manual_seed(g['seed'])
    Rr, N, D = g['R'], g['N'], g['D']
    emb = torch.randn(Rr, N, D, generator=gen)
    x = torch.randint(0, 8, (Rr, N), generator=gen)
    y = torch.randint(0, 8, (Rr, N), generator=gen)
    x[1, :50] = 250
    y[1, :50] = 200
    return emb, x, y


@pytest.mark.parametrize('typ', ['add', 'cover'])
def test_scatter_connection_bit_exact_vs_golden_and_oracle(typ):
    from applestar_b200 import ops
    g = golden('scatter_small.pt')
    emb, x, y = _scatter_inputs(g)
    out = ops.scatter_connection(emb.cuda(), x.cuda(), y.cuda(), None, 152, 160, typ).cpu()
    ref = R.scatter_connection(emb, x, y, typ)
    assert torch.equal(out, ref)                                    # bit-exact incl. every zero
    m = out.permute(0, 2, 3, 1).reshape(g['R'], -1, g['D'])
    cells = g['out'][typ]['cells']
    assert torch.equal(m[cells[:, 0], cells[:, 1]], g['out'][typ]['vals'])


def test_scatter_connection_full_size_uint8_masked_and_backward():
    from applestar_b200 import ops
    obs = synth.gen_observations(6, 512, 11, special=True)
    gen = torch.Generator().manual_seed(12)
    emb = torch.relu(torch.randn(6, 512, 32, generator=gen))
    x, y, en = obs['entity_info']['x'], obs['entity_info']['y'], obs['entity_num']
    mask = (torch.arange(512).unsqueeze(0) < en.unsqueeze(1)).unsqueeze(2)
    ref = R.scatter_connection(emb * mask, x, y, 'add')
    e = emb.cuda().requires_grad_(True)
    out = ops.scatter_connection(e, x.cuda(), y.cuda(), en.cuda(), 152, 160, 'add')
    assert torch.equal(out.detach().cpu(), ref)                     # all-on-one-cell + entity_num=1 rows included
    gout = torch.randn(out.shape, generator=gen)
    out.backward(gout.cuda())
    er = emb.clone().requires_grad_(True)
    R.scatter_connection(er * mask, x, y, 'add').backward(gout)
    assert torch.equal(e.grad.cpu(), er.grad)                       # pure gather -> bit-exact


def test_scatter_round_trip_property_at_bench_size():
    """Size-independent property at the op-microbench shape (B=1024 is sharded in 4 chunks of 256 here):
    sum over the map == sum over valid embeddings, and gathering the map at entity cells upper-bounds rows."""
    from applestar_b200 import ops
    obs = synth.gen_observations(256, 512, 13)
    gen = torch.Generator().manual_seed(14)
    emb = torch.relu(torch.randn(256, 512, 32, generator=gen)).cuda()
    x, y, en = obs['entity_info']['x'].cuda(), obs['entity_info']['y'].cuda(), obs['entity_num'].cuda()
    out = ops.scatter_connection(emb, x, y, en, 152, 160, 'add')
    mask = (torch.arange(512, device='cuda').unsqueeze(0) < en.unsqueeze(1)).unsqueeze(2)
    assert torch.allclose(out.double().sum((2, 3)), (emb * mask).double().sum(1), rtol=1e-6, atol=1e-4)
    cell = y.long() * 160 + x.long()
    picked = out.view(256, 32, -1).gather(2, cell.unsqueeze(1).expand(256, 32, 512)).permute(0, 2, 1)
    assert (picked >= (emb * mask) - 1e-6).all()                    # non-negative rows can only add up
    assert (out != 0).sum().item() <= 256 * 512 * 32


def test_gumbel_argmax_bit_exact():
    from applestar_b200 import ops
    gen = torch.Generator().manual_seed(15)
    for rows, K in ((5, 327), (7, 2), (3, 24320), (9, 513)):
        logits = torch.randn(rows, K, generator=gen)
        logits[0, : K // 2] = -1e9
        logits[rows - 1] = 0.25                                      # all ties -> lowest index after noise ties
        noise = synth.gen_gumbel(rows, 16)['action_type'][:, :1].expand(rows, K).contiguous() if K == 2 else \
            torch.randn(rows, K, generator=gen)
        if K == 513:
            noise[rows - 1] = 0.0
        got = ops.gumbel_argmax(logits.cuda(), noise.cuda()).cpu()
        assert torch.equal(got, R.gumbel_pick(logits, noise))


def test_clip_adam_matches_torch():
    from applestar_b200 import _lib
    gen = torch.Generator().manual_seed(17)
    n = 100003
    p = torch.randn(n, generator=gen)
    p_ref = p.clone().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], lr=1e-3, betas=(0.0, 0.99), eps=1e-5)
    pc, m, v = p.cuda(), torch.zeros(n, device='cuda'), torch.zeros(n, device='cuda')
    shadow = torch.empty(n, device='cuda', dtype=torch.bfloat16)
    scratch = torch.empty(_lib.lib().as_clip_adam_scratch_bytes() // 4, device='cuda')
    norm = torch.zeros(1, device='cuda')
    step_dev = torch.full((1,), 2, dtype=torch.int32, device='cuda')     # third step uses the device-side counter
    for step in range(1, 4):
        g = torch.randn(n, generator=gen) * 0.05
        p_ref.grad = g.clone()
        tn = torch.nn.utils.clip_grad_norm_([p_ref], 1.0)
        opt.step()
        _lib.call('as_clip_adam', pc.data_ptr(), g.cuda().data_ptr(), m.data_ptr(), v.data_ptr(), shadow.data_ptr(), n,
                  1.0, 1e-3, 0.0, 0.99, 1e-5, 0.0, step if step < 3 else 0, 0 if step < 3 else step_dev.data_ptr(),
                  scratch.data_ptr(), norm.data_ptr(), _lib.stream_ptr())
        assert abs(norm.item() - tn.item()) < 1e-4 * tn.item()
        assert (pc.cpu() - p_ref.detach()).abs().max().item() < 2e-6
    assert torch.equal(shadow.cpu(), pc.cpu().to(torch.bfloat16))


def _loss_inputs(T, B, N, seed, fields):
    data = synth.gen_trajectories(T, B, N, seed)
    gen = torch.Generator().manual_seed(seed + 1)
    logits = {}
    for k in C.HEADS:
        t = data['teacher_logit'][k]
        logits[k] = torch.where(t > -1e8, torch.randn(t.shape, generator=gen), t)   # same -1e9 mask pattern as a model
    values = {f: torch.randn(T + 1, B, generator=gen) * 0.5 for f in fields}
    inputs = {'target_logit': logits, 'value': values, 'action_log_prob': data['behaviour_logp'],
              'teacher_logit': data['teacher_logit'], 'mask': data['mask'], 'action': data['action_info'],
              'reward': data['reward'], 'step': data['step'] * 0.5}
    return inputs


@pytest.mark.parametrize('fields,dapo', [(['winloss'], False), (['winloss', 'build_order', 'built_unit', 'battle'], False),
                                         (['winloss'], True), (['winloss', 'build_order', 'built_unit', 'battle'], True)])
def test_fused_rl_loss_matches_oracle(fields, dapo):
    """fp32 vs fp32: atol 2e-5 on every logged scalar, 1e-3 relative (of the tensor max) on gradients.  `dapo`: with the
    DAPO term KL(successive || pi) (as_rl_utils.py:105-127, `use_dapo`), half of the steps below `dapo_steps`."""
    from applestar_b200 import ReinforcementLoss
    T, B, N = 6, 3, 40
    inputs = _loss_inputs(T, B, N, 21, fields)
    cfg = C.DEFAULT_LOSS_CONFIG
    if dapo:
        gen = torch.Generator().manual_seed(5)
        inputs['successive_logit'] = {k: torch.where(t > -1e8, torch.randn(t.shape, generator=gen), t)
                                      for k, t in inputs['teacher_logit'].items()}
        cfg = C.deep_merge(C.DEFAULT_LOSS_CONFIG, {'use_dapo': True, 'dapo': {'dapo_steps': 5000},
                                                   'dapo_head_weights': {'selected_units': 0.5, 'target_location': 2.0}})
    ref_in = synth.tree_map(lambda t: t.clone(), inputs)
    for k in C.HEADS:
        ref_in['target_logit'][k].requires_grad_(True)
    for f in fields:
        ref_in['value'][f].requires_grad_(True)
    ref = R.compute_loss(cfg, ref_in, use_dapo=dapo)
    ref['total_loss'].backward()

    dev_in = synth.tree_map(lambda t: t.cuda(), inputs)
    for k in C.HEADS:
        dev_in['target_logit'][k].requires_grad_(True)
    for f in fields:
        dev_in['value'][f].requires_grad_(True)
    info = ReinforcementLoss(cfg if dapo else {}, 'MP0').compute_loss(dev_in)
    if dapo:
        assert abs(float(ref['battle/total'])) > 1e-3 and 'battle/target_location' in info
        assert not ReinforcementLoss(cfg, 'SP0').use_dapo            # rl_loss.py:22-24: main players only
    info['total_loss'].backward()
    assert set(info) == set(ref)
    for k, v in ref.items():
        assert abs(float(info[k]) - float(v)) <= 2e-5 + 1e-4 * abs(float(v)), (k, float(info[k]), float(v))
    for k in C.HEADS:
        gr, gd = ref_in['target_logit'][k].grad, dev_in['target_logit'][k].grad.cpu()
        assert (gd - gr).abs().max().item() <= 1e-3 * gr.abs().max().item() + 1e-9, k
    for f in fields:
        gr, gd = ref_in['value'][f].grad, dev_in['value'][f].grad.cpu()
        assert (gd - gr).abs().max().item() <= 1e-3 * gr.abs().max().item() + 1e-9, f


@pytest.mark.parametrize('R,S,N,masked', [(5, 1, 512, True), (7, 12, 513, False), (3, 64, 40, True)])
def test_pointer_logits_fwd_bwd(R, S, N, masked):
    """as_pointer_logits_fwd/_bwd (TargetUnit logits, head/action_arg_head.py:343-363, and the SelectedUnits key product)
    against fp32 torch: logits and both gradients to fp32 round-off, mask predicate bit-exact."""
    from applestar_b200 import ops
    g = torch.Generator().manual_seed(R * 100 + S)
    q = torch.randn(R, S, 32, generator=g).cuda().requires_grad_(True)
    k = torch.randn(R, N, 32, generator=g).cuda().requires_grad_(True)
    ln = torch.randint(1, N + 1, (R,), generator=g).cuda()
    ln[0] = N
    scale = 1.0 / 0.8
    out = ops._PointerLogits.apply(q, k, ln if masked else None, scale, -1e9 * scale)
    go = torch.randn(R, S, N, generator=g).cuda()
    out.backward(go)
    qr, kr = q.detach().clone().requires_grad_(True), k.detach().clone().requires_grad_(True)
    ref = torch.einsum('rsk,rnk->rsn', qr, kr)
    if masked:
        m = torch.arange(N, device='cuda').view(1, 1, N) < ln.view(R, 1, 1)
        ref = ref.masked_fill(~m, -1e9)
    ref = ref * scale
    ref.backward(go)
    assert torch.equal(out <= -1e8, ref <= -1e8)
    v = ref > -1e8
    assert (out[v] - ref[v]).abs().max().item() < 1e-4
    assert (q.grad - qr.grad).abs().max().item() < 1e-3 * qr.grad.abs().max().item() + 1e-5
    assert (k.grad - kr.grad).abs().max().item() < 1e-3 * kr.grad.abs().max().item() + 1e-5


def test_scalar_encoder_front_kernels_vs_torch():
    """as_embed_relu_fwd / as_one_hot_bf16 / as_time_encode / as_bo_features / as_cast_any_bf16 (scalar_encoder.py:91-132,
    :41-53) against plain torch on the narrow storage dtypes: the embedding gather is exact, the operand builders are
    bit-exact, products and weight gradients carry the bf16 rounding of every other linear layer."""
    import torch.nn.functional as F
    from applestar_b200 import ops
    g = torch.Generator().manual_seed(8)
    Rn = 300
    idx = torch.randint(0, 140, (Rn,), generator=g).to(torch.int16)            # values above V - 1 are clamped
    table = torch.randn(128, 64, generator=g)
    td = table.cuda().requires_grad_(True)
    out = ops.embedding_relu(idx.cuda(), td)
    tr = table.clone().requires_grad_(True)
    ref = F.relu(F.embedding(idx.long().clamp(max=127), tr))
    assert torch.equal(out.cpu(), ref.detach())
    go = torch.randn(Rn, 64, generator=g)
    out.backward(go.cuda()); ref.backward(go)
    assert (td.grad.cpu() - tr.grad).abs().max().item() < 1e-2 * tr.grad.abs().max().item()
    assert torch.equal(ops.one_hot_bf16(idx.cuda(), 128)[:, :128].float().cpu(), F.one_hot(idx.long().clamp(max=127), 128).float())
    t = torch.rand(Rn, generator=g) * 30000
    pos = synth.time_position_array(32)
    v = ops.time_encode(t.cuda(), pos.cuda()).cpu()
    want = torch.empty(Rn, 32)
    want[:, 0::2] = torch.sin(t.unsqueeze(1) * pos[0::2]); want[:, 1::2] = torch.cos(t.unsqueeze(1) * pos[1::2])
    assert (v - want).abs().max().item() < 2e-3                                # fp32 sin / cos of arguments up to 3e4
    order = torch.randint(0, 174, (Rn, 20), generator=g).to(torch.int16)
    loc = torch.randint(0, 152 * 160, (Rn, 20), generator=g).to(torch.int16)
    W = torch.randn(64, 214, generator=g) * 0.2
    b = torch.randn(64, generator=g) * 0.1
    Wd, bd = W.cuda().requires_grad_(True), b.cuda().requires_grad_(True)
    y = ops.build_order_embed(order.cuda(), loc.cuda(), Wd, bd)
    Wr, br = W.clone().requires_grad_(True), b.clone().requires_grad_(True)
    l = loc.long()
    feats = torch.cat([F.one_hot(order.long(), 174).float(), torch.eye(20).unsqueeze(0).expand(Rn, 20, 20),
                       R.bits_msb_first(l % 160, 10), R.bits_msb_first(l // 160, 10)], dim=2)
    yr = F.relu(F.linear(feats, Wr, br))
    assert (y.cpu() - yr.detach()).abs().max().item() < 2e-2
    gy = torch.randn(Rn, 20, 64, generator=g)
    y.backward(gy.cuda()); yr.backward(gy)
    assert ((Wd.grad.cpu() - Wr.grad).norm() / Wr.grad.norm()).item() < 5e-2
    assert ((bd.grad.cpu() - br.grad).norm() / br.grad.norm()).item() < 5e-2
    x8 = torch.randint(0, 6, (Rn, 261), generator=g).to(torch.uint8)
    from applestar_b200 import kernels as K
    x16 = K.to_bf16(x8.cuda())
    assert x16.shape == (Rn, 264) and torch.equal(x16[:, :261].float().cpu(), x8.float()) and float(x16[:, 261:].abs().sum()) == 0


@pytest.mark.gpu
def test_trajectory_stager_rebuilds_the_dense_batch_bit_exactly():
    """staging.pack -> Stager.load: every tensor of the batch arrives as the reference learner would see it, the teacher's
    selected-units logits rebuilt from their emitted rows (rl_dataloader.py:206-245 padding = -1e9)."""
    from applestar_b200 import staging, synth
    dev = torch.device('cuda:0')
    for (T, B, N, seed) in ((4, 3, 64, 5), (2, 2, 512, 6)):
        host = synth.gen_trajectories(T, B, N, seed)
        packed = staging.pack(host)
        dst = synth.tree_map(lambda t: torch.zeros_like(t, device=dev), host)
        staging.Stager(packed, dev).load(packed, dst)
        torch.cuda.synchronize()
        flat_ref, flat_out = [], []
        synth.tree_map(lambda t: flat_ref.append(t) or t, host)
        synth.tree_map(lambda t: flat_out.append(t) or t, dst)
        assert len(flat_ref) == len(flat_out)
        for a, b in zip(flat_ref, flat_out):
            assert torch.equal(a, b.cpu())
    # a batch whose teacher emitted nothing at all (all padding) and one where every row is data
    host = synth.gen_trajectories(2, 2, 64, 7)
    for fill_all in (True, False):
        h = dict(host); h['teacher_logit'] = dict(host['teacher_logit'])
        su = torch.full_like(host['teacher_logit']['selected_units'], staging.FILL) if fill_all else \
            torch.randn_like(host['teacher_logit']['selected_units'])
        h['teacher_logit']['selected_units'] = su
        packed = staging.pack(h)
        dst = synth.tree_map(lambda t: torch.zeros_like(t, device=dev), h)
        staging.Stager(packed, dev).load(packed, dst)
        assert torch.equal(dst['teacher_logit']['selected_units'].cpu(), su)


@pytest.mark.gpu
@pytest.mark.parametrize('tag,cfg,kw', [('default', {}, {}), ('label_smooth', {'label_smooth': True}, {'label_smooth': True}),
                                        ('no_su_mask', {'su_mask': False}, {'su_mask': False})])
def test_supervised_loss_vs_oracle_and_reference_fixture(tag, cfg, kw):
    """applestar_b200.SupervisedLoss (as_ce_rows_fwd/bwd) against the reference's own output (fixture) and the oracle:
    15 log entries and the gradient of every logit tensor; fp32 arithmetic, tolerance 2e-5 abs + 1e-4 rel."""
    from applestar_b200 import SupervisedLoss, synth
    from oracle import restate
    from util import golden
    fx = golden('sl_loss_small.pt')
    m = fx['meta']
    d = synth.gen_sl_loss_inputs(m['b'], m['s'], m['N'], m['seed'])
    dev = torch.device('cuda:0')
    cu = synth.tree_map(lambda t: t.to(dev), d)
    logits = {k: v.clone().requires_grad_(True) for k, v in cu['logits'].items()}
    info = SupervisedLoss({'learner': cfg}).compute_loss(logits, cu['actions'], cu['actions_mask'], cu['selected_units_num'],
                                                        cu['entity_num'], {'selected_units': cu['infer_selected_units']})
    info['total_loss'].backward()
    ologits = {k: v.clone().requires_grad_(True) for k, v in d['logits'].items()}
    oinfo = restate.supervised_loss(ologits, d['actions'], d['actions_mask'], d['selected_units_num'], d['entity_num'],
                                    d['infer_selected_units'], **kw)
    oinfo['total_loss'].backward()
    ref = fx[tag]
    assert set(info) == set(ref['info'])
    for k, v in ref['info'].items():
        assert abs(float(info[k]) - v) <= 2e-5 + 1e-4 * abs(v), (k, float(info[k]), v)
        assert abs(float(info[k]) - float(oinfo[k])) <= 2e-5 + 1e-4 * abs(v), (k, float(info[k]), float(oinfo[k]))
    for k in logits:
        g, og = logits[k].grad.cpu(), ologits[k].grad
        assert torch.allclose(g, og, rtol=1e-4, atol=2e-6), (k, float((g - og).abs().max()))
        if k in ref['dlogits']:
            assert torch.allclose(g, ref['dlogits'][k], rtol=1e-4, atol=2e-6), k


@pytest.mark.gpu
def test_ce_rows_at_learner_size_vs_torch():
    """The row kernels at the SL learner's sizes (1024 rows x 24 320 location logits; 1024 x 64 x 513 selected-units logits
    with the label mask) against torch's log_softmax on the same device; arg-max bit-equal."""
    from applestar_b200.sl_loss import ce_rows
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(3)
    x = (torch.randn(1024, 24320, device=dev, generator=g) * 3).requires_grad_(True)
    y = torch.randint(0, 24320, (1024,), device=dev, generator=g)
    w = torch.rand(1024, device=dev, generator=g)
    w[::7] = 0
    row, arg = ce_rows(x, y, 0.1)
    (row * w).sum().backward()
    x2 = x.detach().clone().requires_grad_(True)
    lp = torch.log_softmax(x2, -1)
    ref = 0.9 * (-lp.gather(1, y[:, None]).squeeze(1)) + 0.1 * (-lp.mean(-1))
    (ref * w).sum().backward()
    assert torch.equal(arg.long(), x.argmax(-1))
    assert torch.allclose(row, ref, rtol=1e-5, atol=1e-5)
    assert torch.allclose(x.grad, x2.grad, rtol=1e-4, atol=1e-7)
    b, s, n = 1024, 64, 513
    su = torch.randn(b, s, n, device=dev, generator=g).requires_grad_(True)
    lens = torch.randint(0, s + 1, (b,), device=dev, generator=g)
    lab = torch.stack([torch.randperm(n, device=dev, generator=g)[:s] for _ in range(64)]).repeat(b // 64, 1)
    row, _ = ce_rows(su.view(b * s, n), lab.reshape(-1), 0.0, lab, (lens - 1).clamp(min=0), s)
    row.sum().backward()
    picked = torch.arange(s, device=dev)[None] < (lens - 1).clamp(min=0)[:, None]
    hide = torch.zeros(b, n + 1, dtype=torch.bool, device=dev).scatter_(1, torch.where(picked, lab, torch.full_like(lab, n)), True)
    hide = hide[:, None, :n].repeat(1, s, 1)
    own = torch.zeros(b, s, n + 1, dtype=torch.bool, device=dev).scatter_(2, torch.where(picked, lab, torch.full_like(lab, n))[..., None], True)
    su2 = su.detach().clone().requires_grad_(True)
    lp = torch.log_softmax(su2.masked_fill(hide & ~own[..., :n], -1e9), -1)
    ref = -lp.gather(2, lab[..., None]).squeeze(2)
    ref.sum().backward()
    assert torch.allclose(row.view(b, s), ref, rtol=1e-5, atol=1e-5)
    assert torch.allclose(su.grad, su2.grad, rtol=1e-4, atol=1e-7)


@pytest.mark.gpu
@pytest.mark.parametrize('R,N,D', [(37, 512, 32), (5, 64, 32), (3, 7, 8)])
def test_pointer_keys_fwd_bwd(R, N, D):
    """ops.pointer_keys (as_pointer_keys_fwd/bwd) against the reference's cat + where (action_arg_head.py:118-129): forward
    bit-exact, key gradient bit-exact, end-embedding gradient a fixed-order fp32 sum."""
    from applestar_b200 import ops
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(R * N)
    key = torch.randn(R, N, D, device=dev, generator=g).requires_grad_(True)
    end = torch.randn(1, D, device=dev, generator=g).requires_grad_(True)
    en = torch.randint(1, N + 1, (R,), device=dev, generator=g)
    en[0] = N                                                   # end flag on the appended row
    out = ops.pointer_keys(key, end, en)
    go = torch.randn(R, N + 1, D, device=dev, generator=g)
    out.backward(go)
    k2, e2 = key.detach().clone().requires_grad_(True), end.detach().clone().requires_grad_(True)
    cat = torch.cat([k2, k2.new_zeros(R, 1, D)], dim=1)
    is_end = torch.arange(N + 1, device=dev).unsqueeze(0) == en.unsqueeze(1)
    ref = torch.where(is_end.unsqueeze(2), e2.view(1, 1, D), cat)
    ref.backward(go)
    assert torch.equal(out, ref)
    assert torch.equal(key.grad, k2.grad)
    assert torch.allclose(end.grad, e2.grad, rtol=1e-5, atol=1e-5)
