"""GPU parity of the padded-NHWC bf16 spatial kernels against plain PyTorch fp32 references (same bf16-rounded
inputs).  Tolerances are stated per assertion; they are bf16 storage rounding (2^-8 relative) of the outputs."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

# the PyTorch references here must be true fp32 (cuDNN would otherwise use TF32 for convolutions)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def _rand(shape, seed, scale=1.0):
    return (torch.randn(shape, generator=torch.Generator().manual_seed(seed)) * scale).cuda()


def _to_p(x):
    from applestar_b200 import spatial
    return spatial.to_pnhwc(x)


def _from_p(p):
    return p[:, 1:-1, 1:-1].permute(0, 3, 1, 2).float()


def _border_zero(p):
    return (p[:, 0].abs().max() == 0 and p[:, -1].abs().max() == 0 and p[:, :, 0].abs().max() == 0
            and p[:, :, -1].abs().max() == 0)


def test_layout_round_trip_and_channel_pad():
    from applestar_b200 import spatial
    x = _rand((3, 20, 19, 20), 1).requires_grad_(True)
    p = spatial.to_pnhwc(x, 64)
    assert p.shape == (3, 21, 22, 64) and _border_zero(p) and p[..., 20:].abs().max() == 0
    assert torch.equal(_from_p(p)[:, :20], x.detach().to(torch.bfloat16).float())
    g = _rand(p.shape, 2).to(torch.bfloat16)
    p.backward(g)
    assert torch.equal(x.grad, _from_p(g)[:, :20])


@pytest.mark.parametrize('cin,cout,k,H,W', [(128, 128, 3, 19, 20), (128, 64, 3, 38, 40), (64, 32, 3, 76, 80),
                                             (128, 128, 1, 19, 20), (32, 64, 3, 76, 80), (64, 128, 3, 38, 40),
                                             (192, 128, 3, 19, 20), (32, 32, 3, 11, 14)])
def test_conv_nhwc_forward_backward(cin, cout, k, H, W):
    from applestar_b200 import spatial
    R = 3
    x = _rand((R, cin, H, W), 3).requires_grad_(True)
    w = _rand((cout, cin, k, k), 4, 0.05).requires_grad_(True)
    b = _rand((cout,), 5).requires_grad_(True)
    y = spatial.conv(_to_p(x), w, b, relu=True)
    assert _border_zero(y)
    gy = torch.zeros_like(y)
    gy[:, 1:-1, 1:-1] = _rand((R, H, W, cout), 6).to(torch.bfloat16)
    y.backward(gy)
    xr = x.detach().to(torch.bfloat16).float().requires_grad_(True)
    wr = w.detach().to(torch.bfloat16).float().requires_grad_(True)
    br = b.detach().clone().requires_grad_(True)
    yr = F.relu(F.conv2d(xr, wr, br, padding=k // 2))
    yr.backward(_from_p(gy))
    scale = yr.abs().max().item()
    assert (_from_p(y) - yr).abs().max().item() < 0.01 * scale + 1e-2
    assert (x.grad - xr.grad).abs().max().item() < 0.02 * xr.grad.abs().max().item() + 1e-3
    assert (w.grad - wr.grad).abs().max().item() < 0.02 * wr.grad.abs().max().item() + 1e-2
    assert (b.grad - br.grad).abs().max().item() < 0.01 * br.grad.abs().max().item() + 1e-2


def test_conv1_padded_input_channels():
    """location-head conv1: 132 real channels carried in a 192-channel map."""
    from applestar_b200 import spatial
    R, H, W = 2, 19, 20
    emb = torch.relu(_rand((R, 4 * H * W), 7)).requires_grad_(True)
    skip_nchw = _rand((R, 128, H, W), 8).requires_grad_(True)
    w = _rand((128, 132, 1, 1), 9, 0.05).requires_grad_(True)
    cat = spatial.loc_cat(emb, _to_p(skip_nchw), 192)
    assert _border_zero(cat) and cat[..., 132:].abs().max() == 0
    y = spatial.conv(cat, w, None, relu=False)
    gy = torch.zeros_like(y)
    gy[:, 1:-1, 1:-1] = _rand((R, H, W, 128), 10).to(torch.bfloat16)
    y.backward(gy)
    er = emb.detach().clone().requires_grad_(True)
    sr = skip_nchw.detach().to(torch.bfloat16).float().requires_grad_(True)
    wr = w.detach().to(torch.bfloat16).float().requires_grad_(True)
    catr = F.relu(torch.cat([er.view(R, 4, H, W), sr], dim=1))
    yr = F.conv2d(catr, wr)
    yr.backward(_from_p(gy))
    assert (_from_p(y) - yr).abs().max().item() < 0.02 * yr.abs().max().item() + 1e-2
    assert (emb.grad - er.grad).abs().max().item() < 0.02 * er.grad.abs().max().item() + 1e-3
    assert (skip_nchw.grad - sr.grad).abs().max().item() < 0.02 * sr.grad.abs().max().item() + 1e-3
    assert (w.grad - wr.grad).abs().max().item() < 0.02 * wr.grad.abs().max().item() + 1e-2


def test_residual_tails():
    from applestar_b200 import spatial
    a = _rand((2, 21, 22, 128), 11).to(torch.bfloat16).requires_grad_(True)
    b = _rand((2, 21, 22, 128), 12).to(torch.bfloat16).requires_grad_(True)
    g = _rand((2, 21, 22, 128), 13).to(torch.bfloat16)
    y = spatial.add_relu(a, b)
    y.backward(g)
    ref = F.relu(a.detach().float() + b.detach().float())
    assert torch.equal(y.float(), ref.to(torch.bfloat16).float())
    assert torch.equal(a.grad.float(), (g.float() * (ref > 0)).to(torch.bfloat16).float()) and torch.equal(a.grad, b.grad)
    # gated residual
    r, gg, x = (_rand((2, 21, 22, 128), s).to(torch.bfloat16).requires_grad_(True) for s in (14, 15, 16))
    sp = torch.full((1,), 0.1, device='cuda', requires_grad=True)
    y = spatial.gated_residual(r, gg, sp, x)
    y.backward(g)
    rr, gr, xr = (t.detach().float().requires_grad_(True) for t in (r, gg, x))
    spr = sp.detach().clone().requires_grad_(True)
    yr = F.relu(torch.tanh(rr * torch.sigmoid(gr)) * spr + xr)
    yr.backward(g.float())
    assert (y.float() - yr).abs().max().item() < 0.02
    for got, want in ((r.grad, rr.grad), (gg.grad, gr.grad), (x.grad, xr.grad)):
        assert (got.float() - want).abs().max().item() < 0.01 * want.abs().max().item() + 1e-3
    assert abs(sp.grad.item() - spr.grad.item()) < 0.01 * abs(spr.grad.item()) + 1e-2


@pytest.mark.parametrize('H,W,Cc', [(19, 20, 128), (38, 40, 64)])
def test_bilinear_upsample_fwd_bwd(H, W, Cc):
    from applestar_b200 import spatial
    x = _rand((2, Cc, H, W), 17).requires_grad_(True)
    y = spatial.upsample2(_to_p(x))
    assert y.shape == (2, 2 * H + 2, 2 * W + 2, Cc) and _border_zero(y)
    gy = torch.zeros_like(y)
    gy[:, 1:-1, 1:-1] = _rand((2, 2 * H, 2 * W, Cc), 18).to(torch.bfloat16)
    y.backward(gy)
    xr = x.detach().to(torch.bfloat16).float().requires_grad_(True)
    yr = F.interpolate(xr, scale_factor=2., mode='bilinear')
    yr.backward(_from_p(gy))
    assert (_from_p(y) - yr).abs().max().item() < 0.02
    assert (x.grad - xr.grad).abs().max().item() < 0.02 * xr.grad.abs().max().item() + 1e-3


def test_final_conv_32_to_1():
    from applestar_b200 import spatial
    R, H, W = 2, 152, 160
    x = _rand((R, 32, H, W), 19).requires_grad_(True)
    w = _rand((1, 32, 3, 3), 20, 0.1).requires_grad_(True)
    b = _rand((1,), 21).requires_grad_(True)
    logits = spatial.conv_c32_to_1(_to_p(x), w, b, 1.0 / 0.8)
    g = _rand((R, H * W), 22)
    logits.backward(g)
    xr = x.detach().to(torch.bfloat16).float().requires_grad_(True)
    wr, br = w.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    lr = (F.conv2d(xr, wr, br, padding=1) / 0.8).reshape(R, -1)
    lr.backward(g)
    assert (logits - lr).abs().max().item() < 1e-4
    assert (x.grad - xr.grad).abs().max().item() < 0.01 * xr.grad.abs().max().item() + 1e-3
    assert (w.grad - wr.grad).abs().max().item() < 1e-3 * wr.grad.abs().max().item() + 1e-2
    assert abs(b.grad.item() - br.grad.item()) < 1e-3 * abs(br.grad.item()) + 1e-2


def test_maxpool_fwd_bwd():
    from applestar_b200 import spatial
    x = torch.relu(_rand((2, 32, 76, 80), 23)).requires_grad_(True)      # post-ReLU: many exact ties at 0
    y = spatial.max_pool2(_to_p(x))
    assert y.shape == (2, 40, 42, 32) and _border_zero(y)
    gy = torch.zeros_like(y)
    gy[:, 1:-1, 1:-1] = _rand((2, 38, 40, 32), 24).to(torch.bfloat16)
    y.backward(gy)
    xr = x.detach().to(torch.bfloat16).float().requires_grad_(True)
    yr = F.max_pool2d(xr, 2, 2)
    yr.backward(_from_p(gy))
    assert torch.equal(_from_p(y), yr)
    assert torch.equal(x.grad, xr.grad)                                   # first-max routing == torch


def test_fused_scatter_features_bit_exact(monkeypatch):
    """as_spatial_features_scatter (scatter sums inside the feature kernel) == as_scatter_connection_fwd +
    as_spatial_features, bit for bit: heavy collisions (512 entities on one cell), entity_num = 1, uint8 and int64
    coordinates, and the oracle's scatter channels."""
    from applestar_b200 import spatial, synth
    from oracle import restate as R
    obs = synth.gen_observations(5, 512, 77, special=True)
    g = torch.Generator().manual_seed(5)
    proj = torch.relu(torch.randn(5, 512, 32, generator=g)).cuda()
    sp = {k: v.cuda() for k, v in obs['spatial_info'].items()}
    en = obs['entity_num'].cuda()
    outs = {}
    for idx_dtype in (torch.uint8, torch.int64):
        x = obs['entity_info']['x'].to(idx_dtype).cuda()
        y = obs['entity_info']['y'].to(idx_dtype).cuda()
        for mode in ('1', '0'):
            monkeypatch.setenv('APPLESTAR_SCATTER_FUSED', mode)
            feats, H, W = spatial._build_features(proj, x, y, en, sp)
            outs[(idx_dtype, mode)] = feats
        assert torch.equal(outs[(idx_dtype, '1')], outs[(idx_dtype, '0')])
    assert torch.equal(outs[(torch.uint8, '1')], outs[(torch.int64, '1')])
    mask = (torch.arange(512).unsqueeze(0) < obs['entity_num'].unsqueeze(1)).unsqueeze(2)
    smap = R.scatter_connection(proj.cpu() * mask, obs['entity_info']['x'], obs['entity_info']['y'], 'add')
    got = outs[(torch.uint8, '1')][:, 1:-1, 1:-1, 24:56].permute(0, 3, 1, 2).float().cpu()
    assert torch.equal(got, smap.to(torch.bfloat16).float())


def test_spatial_project_matches_reference_front():
    """scatter + 56-channel input + 1x1 project conv + ReLU (+ its backward) vs the oracle's fp32 front end."""
    from applestar_b200 import spatial, synth
    from oracle import restate as R
    obs = synth.gen_observations(3, 512, 31, special=True)
    g = torch.Generator().manual_seed(32)
    proj = torch.relu(torch.randn(3, 512, 32, generator=g))
    w = torch.randn(32, 56, 1, 1, generator=g) * 0.2
    b = torch.randn(32, generator=g) * 0.1
    en = obs['entity_num']
    mask = (torch.arange(512).unsqueeze(0) < en.unsqueeze(1)).unsqueeze(2)
    pr = proj.clone().requires_grad_(True)
    wr, br = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    smap = R.scatter_connection(pr * mask, obs['entity_info']['x'], obs['entity_info']['y'], 'add')
    ref = torch.relu(F.conv2d(R.spatial_input(obs['spatial_info'], smap), wr, br))
    gout = torch.randn(ref.shape, generator=g)
    ref.backward(gout)

    pd = proj.cuda().requires_grad_(True)
    wd, bd = w.cuda().requires_grad_(True), b.cuda().requires_grad_(True)
    sp = {k: v.cuda() for k, v in obs['spatial_info'].items()}
    out = spatial.spatial_project(pd, obs['entity_info']['x'].cuda(), obs['entity_info']['y'].cuda(), en.cuda(), sp, wd, bd)
    assert out.shape == (3, 154, 162, 32) and _border_zero(out)
    gy = torch.zeros_like(out)
    gy[:, 1:-1, 1:-1] = gout.permute(0, 2, 3, 1).cuda().to(torch.bfloat16)
    out.backward(gy)
    assert (_from_p(out).cpu() - ref).abs().max().item() < 0.02 * ref.abs().max().item() + 2e-2
    # bf16 features can flip a ReLU whose pre-activation is ~0 (obs 0 piles 512 entities on one cell), which toggles that
    # pixel's whole contribution: compare in relative L2, not max
    assert ((pd.grad.cpu() - pr.grad).norm() / pr.grad.norm()).item() < 0.05
    assert ((wd.grad.cpu() - wr.grad).norm() / wr.grad.norm()).item() < 0.05
    assert ((bd.grad.cpu() - br.grad).norm() / br.grad.norm()).item() < 0.05


def test_spatial_front_fused_vs_oracle():
    """csrc/spatial_front.cu: scatter + 56-channel input + 1x1 project + ReLU + first 2x2 max-pool in one kernel, evaluated
    in fp32 from the one-hot planes -> the pooled map equals the fp32 oracle up to the bf16 store (and the bf16 copy of
    the per-cell contribution); pool routing / ReLU masks / effect-plane pixel-0 quirk / clamped coordinates are index
    work and agree exactly wherever the window maximum is unambiguous.  Gradients (weight incl. the scatter block, bias,
    entity projections) against the oracle's autograd."""
    from applestar_b200 import spatial, synth
    from oracle import restate as R
    Rn = 4
    obs = synth.gen_observations(Rn, 512, 31, special=True)
    obs['entity_info']['x'][2, :40] = 250                       # out of range: clamps to 159 / 151
    obs['entity_info']['y'][2, :40] = 200
    g = torch.Generator().manual_seed(32)
    proj = torch.relu(torch.randn(Rn, 512, 32, generator=g)) * 0.5
    w = torch.randn(32, 56, 1, 1, generator=g) * 0.3
    b = torch.randn(32, generator=g) * 0.1
    en = obs['entity_num']
    mask = (torch.arange(512).unsqueeze(0) < en.unsqueeze(1)).unsqueeze(2)
    pr = proj.clone().requires_grad_(True)
    wr, br = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    smap = R.scatter_connection(pr * mask, obs['entity_info']['x'], obs['entity_info']['y'], 'add')
    ref = F.max_pool2d(torch.relu(F.conv2d(R.spatial_input(obs['spatial_info'], smap), wr, br)), 2, 2)
    gout = torch.randn(ref.shape, generator=g).to(torch.bfloat16).float()
    ref.backward(gout)

    pd = proj.cuda().requires_grad_(True)
    wd, bd = w.cuda().requires_grad_(True), b.cuda().requires_grad_(True)
    sp = {k: v.cuda() for k, v in obs['spatial_info'].items()}
    out = spatial.spatial_front(pd, obs['entity_info']['x'].cuda(), obs['entity_info']['y'].cuda(), en.cuda(), sp, wd, bd)
    assert out.shape == (Rn, 78, 82, 32) and _border_zero(out)
    got = _from_p(out).cpu()
    refd = ref.detach()
    assert torch.equal(got > 0, refd.to(torch.bfloat16).float() > 0) or ((got > 0) != (refd > 0)).float().mean().item() < 1e-4
    assert (got - refd).abs().max().item() < 0.01 * refd.abs().max().item() + 1e-2
    gy = torch.zeros_like(out)
    gy[:, 1:-1, 1:-1] = gout.permute(0, 2, 3, 1).cuda().to(torch.bfloat16)
    out.backward(gy)
    for name, mine, want, tol in (('proj', pd.grad, pr.grad, 0.05), ('weight', wd.grad, wr.grad, 0.01), ('bias', bd.grad, br.grad, 0.01)):
        rel = ((mine.cpu() - want).norm() / want.norm()).item()
        assert rel < tol, (name, rel)
    assert float(pd.grad[1, 1:].abs().sum()) == 0.0              # entity_num == 1: masked entities get exactly zero


@pytest.mark.parametrize('rows,C', [(1000, 256), (4097, 40), (77, 128), (25000, 32)])
@pytest.mark.parametrize('ykind', ['none', 'f32', 'bf16'])
def test_linear_bwd_prep(rows, C, ykind):
    """as_linear_bwd_prep: bf16(dy * (y > 0)) bit-exact, bias gradient = fp32 column sums of the masked gradient."""
    from applestar_b200 import _lib
    dy = _rand((rows, C), 41)
    y = torch.relu(_rand((rows, C), 42))
    y32 = y if ykind == 'f32' else None
    y16 = y.to(torch.bfloat16) if ykind == 'bf16' else None
    dz = torch.empty((rows, C), dtype=torch.bfloat16, device='cuda')
    db = torch.empty(C, dtype=torch.float32, device='cuda')
    scratch = torch.empty(_lib.lib().as_colsum_scratch_floats(C), dtype=torch.float32, device='cuda')
    _lib.call('as_linear_bwd_prep', dy.data_ptr(), 0 if y32 is None else y32.data_ptr(), 0 if y16 is None else y16.data_ptr(),
              dz.data_ptr(), db.data_ptr(), scratch.data_ptr(), rows, C, _lib.stream_ptr())
    mask = torch.ones_like(dy, dtype=torch.bool) if ykind == 'none' else ((y32 if y32 is not None else y16.float()) > 0)
    want = dy * mask
    assert torch.equal(dz, want.to(torch.bfloat16))
    ref = want.double().sum(0)
    assert (db.double() - ref).abs().max().item() <= 1e-5 * want.abs().sum(0).max().item() + 1e-5


@pytest.mark.parametrize('rows,C', [(480480 // 8, 128), (12345, 32), (999, 64)])
def test_relu_bwd_colsum_bf16(rows, C):
    from applestar_b200 import _lib
    dy = _rand((rows, C), 43).to(torch.bfloat16)
    y = torch.relu(_rand((rows, C), 44)).to(torch.bfloat16)
    dz = torch.empty_like(dy)
    db = torch.empty(C, dtype=torch.float32, device='cuda')
    scratch = torch.empty(_lib.lib().as_colsum_scratch_floats(C), dtype=torch.float32, device='cuda')
    _lib.call('as_relu_bwd_colsum_bf16', dy.data_ptr(), y.data_ptr(), dz.data_ptr(), db.data_ptr(), scratch.data_ptr(),
              rows, C, _lib.stream_ptr())
    want = torch.where(y.float() > 0, dy, torch.zeros_like(dy))
    assert torch.equal(dz, want)
    ref = want.double().sum(0)
    assert (db.double() - ref).abs().max().item() <= 1e-5 * want.float().abs().sum(0).max().item() + 1e-5


@pytest.mark.parametrize('relu_input', [False, True])
def test_upconv_fused_tail_vs_torch(relu_input):
    """csrc/upconv.cu: bilinear x2 + 3x3 conv 32 -> 1 + flatten + 1/temperature without the upsampled map, against
    F.interpolate + F.conv2d in fp32 (forward 1e-4 relative; dx rounds to bf16; weight / bias gradients fp32)."""
    from applestar_b200 import spatial
    g = torch.Generator().manual_seed(5)
    R, h, w = 3, 76, 80
    x = torch.relu(torch.randn(R, 32, h, w, generator=g)).to(torch.bfloat16).float()
    wt = torch.randn(1, 32, 3, 3, generator=g) * 0.2
    b = torch.randn(1, generator=g)
    inv_t = 1.0 / 0.8
    xr, wr, br = x.clone().requires_grad_(True), wt.clone().requires_grad_(True), b.clone().requires_grad_(True)
    ref = F.conv2d(F.interpolate(torch.relu(xr) if relu_input else xr, scale_factor=2., mode='bilinear'), wr, br, padding=1)
    ref = ref.reshape(R, -1) * inv_t
    go = torch.randn(ref.shape, generator=g)
    ref.backward(go)
    xp = spatial.to_pnhwc(x.cuda()).detach().requires_grad_(True)
    wd, bd = wt.cuda().requires_grad_(True), b.cuda().requires_grad_(True)
    out = spatial.upconv_c32_to_1(xp, wd, bd, inv_t, relu_input=relu_input)
    assert out.shape == (R, 4 * h * w)
    assert (out.cpu() - ref.detach()).abs().max().item() < 1e-4 * ref.abs().max().item() + 1e-4
    out.backward(go.cuda())
    assert _border_zero(xp.grad)
    gx = _from_p(xp.grad).cpu()
    assert (gx - xr.grad).abs().max().item() < 0.01 * xr.grad.abs().max().item() + 1e-3
    assert (wd.grad.cpu() - wr.grad).abs().max().item() < 1e-4 * wr.grad.abs().max().item() + 1e-3
    assert abs(float(bd.grad) - float(br.grad)) < 1e-3 * abs(float(br.grad)) + 1e-3


@pytest.mark.parametrize('k1,k2,C', [(1, 1, 128), (3, 3, 128), (3, 1, 64)])
def test_conv_chain_relu_sign_bits(monkeypatch, k1, k2, C):
    """conv -> ReLU -> conv with the ReLU backward applied in the second convolution's data-gradient epilogue: from the sign
    bits the first convolution wrote (as_conv_nhwc_bf16_relu_bits / _bitmasked) and from the bf16 map itself
    (as_conv_nhwc_bf16_masked) the gradients are bit-identical, and both match torch."""
    from applestar_b200 import _lib, spatial
    R, H, W = 3, 19, 20
    x0 = _rand((R, C, H, W), 31)
    w1_0, w2_0 = _rand((C, C, k1, k1), 32, 0.05), _rand((C, C, k2, k2), 33, 0.05)
    b1_0 = _rand((C,), 34)
    gy = None
    grads, calls = {}, {}
    for bits_on in (True, False):
        monkeypatch.setattr(spatial, 'RELU_BITS', bits_on)
        seen = []
        orig = _lib.call
        monkeypatch.setattr(_lib, 'call', lambda name, *a, _o=orig, _s=seen: (_s.append(name), _o(name, *a))[1])
        x = x0.clone().requires_grad_(True)
        w1, w2, b1 = w1_0.clone().requires_grad_(True), w2_0.clone().requires_grad_(True), b1_0.clone().requires_grad_(True)
        h = spatial.conv(_to_p(x), w1, b1, relu=True, grad_premasked=True)
        y = spatial.conv(h, w2, None, relu=False, mask_input_grad=True)
        if gy is None:
            gy = torch.zeros_like(y)
            gy[:, 1:-1, 1:-1] = _rand((R, H, W, C), 35).to(torch.bfloat16)
        y.backward(gy)
        monkeypatch.setattr(_lib, 'call', orig)
        grads[bits_on] = (x.grad.clone(), w1.grad.clone(), w2.grad.clone(), b1.grad.clone(), _from_p(y).clone())
        calls[bits_on] = seen
    assert 'as_conv_nhwc_bf16_relu_bits' in calls[True] and 'as_conv_nhwc_bf16_bitmasked' in calls[True]
    assert 'as_conv_nhwc_bf16_masked' in calls[False] and 'as_conv_nhwc_bf16_bitmasked' not in calls[False]
    for a, b in zip(grads[True], grads[False]):
        assert torch.equal(a, b)
    xr = x0.to(torch.bfloat16).float().requires_grad_(True)
    w1r, w2r = w1_0.to(torch.bfloat16).float().requires_grad_(True), w2_0.to(torch.bfloat16).float().requires_grad_(True)
    hr = F.relu(F.conv2d(xr, w1r, b1_0, padding=k1 // 2)).to(torch.bfloat16).float()
    yr = F.conv2d(hr, w2r, None, padding=k2 // 2)
    hr_full = F.relu(F.conv2d(xr, w1r, b1_0, padding=k1 // 2))
    yr = F.conv2d(hr_full, w2r, None, padding=k2 // 2)
    yr.backward(_from_p(gy))
    gx = grads[True][0]
    assert (gx - xr.grad).abs().max().item() < 0.03 * xr.grad.abs().max().item() + 1e-3
    assert (grads[True][1] - w1r.grad).abs().max().item() < 0.03 * w1r.grad.abs().max().item() + 1e-2
