"""Spatial path on zero-padded NHWC bf16 feature maps ([R, H+2, W+2, C], border == 0).

Every 3x3 / 1x1 convolution of the 19x20 ResBlocks and of the location head is an implicit GEMM on the tcgen05
kernel (forward, data-gradient and weight-gradient); the tails (residual ReLU, gated residual, bilinear x2, the
132-channel concat, the final 32->1 convolution) are bf16 HBM-bound kernels that preserve the zero border.
Reference modules: ResBlock (ctools/torch_utils/network/res_block.py:13-65), GatedResBlock
(model/module_utils.py:204-231), LocationHead (model/head/action_arg_head.py:366-450).
"""
import ctypes
import os

import torch
import torch.nn.functional as F

from . import _lib
from . import kernels as K

_ptr = K._ptr


def _conv_weights(weight, cin_pad):
    """bf16 operand layouts of a Conv2d weight [Cout, Cin, k, k]: forward [Cout][tap][ci] and data-gradient
    [ci][flipped tap][Cout]; cached on the Parameter per optimizer generation."""
    tag = (K._SHADOW_GEN[0], weight._version, weight.data_ptr(), cin_pad)
    hit = getattr(weight, '_as_conv16', None)
    if hit is not None and hit[0] == tag:
        return hit[1], hit[2]
    w = weight.detach().contiguous()
    cout, cin, kh, kw = w.shape
    w_fwd = torch.empty((cout, kh * kw * cin_pad), dtype=torch.bfloat16, device=w.device)
    w_dg = torch.empty((cin_pad, kh * kw * cout), dtype=torch.bfloat16, device=w.device)
    _lib.call('as_conv_weight_layouts', w.data_ptr(), w_fwd.data_ptr(), w_dg.data_ptr(), cout, cin, cin_pad, kh * kw,
              _lib.stream_ptr())
    try:
        weight._as_conv16 = (tag, w_fwd, w_dg)
    except AttributeError:
        pass
    return w_fwd, w_dg


def _conv_call(x, w2, y, bias, R, H, W, cin, cout, taps, relu):
    _lib.call('as_conv_nhwc_bf16', x.data_ptr(), w2.data_ptr(), y.data_ptr(), _ptr(bias), R, H, W, cin, cout, taps, 0,
              1, int(relu), _lib.stream_ptr())


FORK_CONV = os.environ.get('APPLESTAR_CONV_FORK', '1') != '0'
RELU_BITS = os.environ.get('APPLESTAR_CONV_RELU_BITS', '1') != '0'


class ConvNHWC(torch.autograd.Function):
    """3x3 / 1x1 convolution on a padded NHWC bf16 map.  With `fork` the input is handed back as a second output (an
    alias): the caller gives THAT tensor to the input's other consumers (residual add, gate branch), so in backward their
    gradient arrives here as `dxa` and is added to this convolution's data gradient inside the GEMM epilogue
    (as_conv_nhwc_bf16_add) instead of by autograd's separate add over the two maps."""

    @staticmethod
    def forward(ctx, x, weight, bias, relu, fork=False, grad_premasked=False, mask_input_grad=False):
        """grad_premasked: every consumer of this (ReLU) layer's output already applies the ReLU backward mask to the
        gradient it sends back (a following convolution called with mask_input_grad, a max-pool with relu_input), so
        the backward skips its own mask pass.  mask_input_grad: the input is the ReLU output of the previous layer and
        this convolution is its only consumer: the data-gradient GEMM zeroes the gradient where the input is <= 0."""
        ctx.set_materialize_grads(False)
        ctx.grad_premasked, ctx.mask_input_grad = bool(grad_premasked and relu), bool(mask_input_grad)
        R, ph, pw, cin_p = x.shape
        H, W = ph - 2, pw - 2
        cout, cin, kh, _ = weight.shape
        taps = kh * kh
        w_fwd, w_dg = _conv_weights(weight, cin_p)
        y = torch.empty((R, ph, pw, cout), dtype=torch.bfloat16, device=x.device)
        in_bits = getattr(x, '_as_relu_bits', None) if mask_input_grad else None
        x = x.contiguous()
        if ctx.grad_premasked and cout % 32 == 0 and any(ctx.needs_input_grad[:3]) and RELU_BITS:
            # the consumers will apply this layer's ReLU mask to the gradient they send back: hand them 1 bit per element
            # (written by this launch from its packed output) instead of making them re-read the bf16 map
            bits = torch.empty((R * ph * pw, cout // 32), dtype=torch.int32, device=x.device)
            _lib.call('as_conv_nhwc_bf16_relu_bits', x.data_ptr(), w_fwd.data_ptr(), y.data_ptr(), _ptr(bias), bits.data_ptr(),
                      R, H, W, cin_p, cout, taps, _lib.stream_ptr())
            y._as_relu_bits = bits
        else:
            _conv_call(x, w_fwd, y, bias, R, H, W, cin_p, cout, taps, relu)
        ctx.dims = (R, H, W, cin, cin_p, cout, taps, kh)
        ctx.relu = relu
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x, w_dg, y if relu else None, in_bits)
        if fork:
            return y, x.view_as(x)
        return y

    @staticmethod
    def backward(ctx, dy, dxa=None):
        if dy is None:                                   # only the pass-through alias was used downstream
            return dxa, None, None, None, None, None, None
        x, w_dg, y, in_bits = ctx.saved_tensors
        R, H, W, cin, cin_p, cout, taps, kh = ctx.dims
        dy = dy.contiguous()
        st = _lib.stream_ptr()
        dx = dw = db = None
        want_db = ctx.has_bias and ctx.needs_input_grad[2]
        if ctx.relu and not ctx.grad_premasked:
            dz = torch.empty_like(dy)
            if want_db and cout % 8 == 0:       # ReLU mask and bias gradient in one pass
                db = torch.empty(cout, dtype=torch.float32, device=x.device)
                scratch = torch.empty(_lib.lib().as_colsum_scratch_floats(cout), dtype=torch.float32, device=x.device)
                _lib.call('as_relu_bwd_colsum_bf16', dy.data_ptr(), y.data_ptr(), dz.data_ptr(), db.data_ptr(),
                          scratch.data_ptr(), R * (H + 2) * (W + 2), cout, st)
                want_db = False
            else:
                _lib.call('as_relu_bwd_bf16', dy.data_ptr(), y.data_ptr(), dz.data_ptr(), dy.numel(), st)
            dy = dz
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            if ctx.mask_input_grad and cin_p % 32 == 0 and (dxa is None or (dxa.dtype == torch.bfloat16 and dxa.shape == x.shape)):
                dxa = None if dxa is None else dxa.contiguous()
                if in_bits is not None:
                    _lib.call('as_conv_nhwc_bf16_bitmasked', dy.data_ptr(), w_dg.data_ptr(), dx.data_ptr(), _ptr(dxa),
                              in_bits.data_ptr(), R, H, W, cout, cin_p, taps, st)
                else:
                    _lib.call('as_conv_nhwc_bf16_masked', dy.data_ptr(), w_dg.data_ptr(), dx.data_ptr(), _ptr(dxa), x.data_ptr(),
                              R, H, W, cout, cin_p, taps, st)
            elif ctx.mask_input_grad:
                raise RuntimeError('mask_input_grad needs a channel count that is a multiple of 32')
            elif dxa is not None and cin_p % 32 == 0 and dxa.dtype == torch.bfloat16 and dxa.shape == x.shape:
                dxa = dxa.contiguous()
                _lib.call('as_conv_nhwc_bf16_add', dy.data_ptr(), w_dg.data_ptr(), dx.data_ptr(), 0, dxa.data_ptr(), R, H, W,
                          cout, cin_p, taps, 0, st)
            else:
                _conv_call(dy, w_dg, dx, None, R, H, W, cout, cin_p, taps, False)
                if dxa is not None:
                    dx = dx + dxa
        if ctx.needs_input_grad[1]:
            P = R * (H + 2) * (W + 2)
            tiles = taps * ((cout + 127) // 128) * max(1, (cin_p + 255) // 256)
            pmode = _lib.lib().as_conv_wgrad_pair_mode(cin_p, cout, taps, 2, W)
            if pmode >= 10:                     # one CTA per split (and 64-column half) walks all nine taps from a slab
                tiles = ((cin_p if pmode == 11 else cout) + 63) // 64
            elif pmode:
                tiles = 5                       # two taps per 128-row tile (64-channel M side)
            want = max(1, min((P + 63) // 64, K.NUM_SMS // tiles))
            splits = _lib.lib().as_gemm_splits(P, want)
            dw9 = torch.empty((taps, cout, cin_p), dtype=torch.float32, device=x.device)
            ws_n = _lib.lib().as_conv_wgrad_ws_floats(cin_p, cout, taps, splits, W)
            ws = torch.empty(ws_n, dtype=torch.float32, device=x.device) if splits > 1 else None
            _lib.call('as_conv_wgrad_nhwc_bf16', dy.data_ptr(), x.data_ptr(), dw9.data_ptr(), _ptr(ws), R, H, W, cin_p,
                      cout, taps, splits, st)
            dw = dw9[:, :, :cin].permute(1, 2, 0).reshape(cout, cin, kh, kh)
        if want_db:
            db = torch.empty(cout, dtype=torch.float32, device=x.device)
            scratch = torch.empty(_lib.lib().as_colsum_scratch_floats(cout), dtype=torch.float32, device=x.device)
            _lib.call('as_colsum_bf16', dy.data_ptr(), db.data_ptr(), scratch.data_ptr(), R * (H + 2) * (W + 2), cout, st)
        return dx, dw, db, None, None, None, None


def conv(x, weight, bias, relu=False, grad_premasked=False, mask_input_grad=False):
    return ConvNHWC.apply(x, weight, bias, relu, False, grad_premasked, mask_input_grad)


def conv_fork(x, weight, bias, relu=False, grad_premasked=False):
    """-> (conv(x), alias of x for x's other consumers); see ConvNHWC."""
    if not FORK_CONV or not (torch.is_grad_enabled() and x.requires_grad):
        return ConvNHWC.apply(x, weight, bias, relu, False, grad_premasked, False), x
    return ConvNHWC.apply(x, weight, bias, relu, True, grad_premasked, False)


class AddAct(torch.autograd.Function):
    """mode 0: a + b; mode 1: relu(a + b) on bf16 maps."""

    @staticmethod
    def forward(ctx, a, b, mode):
        a, b = a.contiguous(), b.contiguous()
        y = torch.empty_like(a)
        _lib.call('as_add_act_bf16', a.data_ptr(), b.data_ptr(), y.data_ptr(), a.numel(), mode, _lib.stream_ptr())
        ctx.mode = mode
        ctx.save_for_backward(y if mode == 1 else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        if ctx.mode == 1:
            dy = dy.contiguous()
            dz = torch.empty_like(dy)
            _lib.call('as_relu_bwd_bf16', dy.data_ptr(), y.data_ptr(), dz.data_ptr(), dy.numel(), _lib.stream_ptr())
            dy = dz
        return dy, dy, None


def add(a, b):
    return AddAct.apply(a, b, 0)


def add_relu(a, b):
    return AddAct.apply(a, b, 1)


class Gated(torch.autograd.Function):
    @staticmethod
    def forward(ctx, r, g, x, sp):
        r, g, x = r.contiguous(), g.contiguous(), x.contiguous()
        y = torch.empty_like(x)
        _lib.call('as_gated_fwd_bf16', r.data_ptr(), g.data_ptr(), x.data_ptr(), sp.data_ptr(), y.data_ptr(), x.numel(),
                  _lib.stream_ptr())
        ctx.save_for_backward(r, g, y, sp)
        return y

    @staticmethod
    def backward(ctx, dy):
        r, g, y, sp = ctx.saved_tensors
        dy = dy.contiguous()
        dr, dg, dx = torch.empty_like(r), torch.empty_like(g), torch.empty_like(y)
        dsp = torch.empty(1, dtype=torch.float32, device=y.device)
        scratch = torch.empty(_lib.lib().as_gated_scratch_floats(), dtype=torch.float32, device=y.device)
        _lib.call('as_gated_bwd_bf16', dy.data_ptr(), y.data_ptr(), r.data_ptr(), g.data_ptr(), sp.data_ptr(),
                  dr.data_ptr(), dg.data_ptr(), dx.data_ptr(), dsp.data_ptr(), scratch.data_ptr(), y.numel(),
                  _lib.stream_ptr())
        return dr, dg, dx, dsp


def gated_residual(r, g, update_sp, x):
    return Gated.apply(r, g, x, update_sp)


class Up2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        R, ph, pw, Cc = x.shape
        H, W = ph - 2, pw - 2
        x = x.contiguous()
        y = torch.empty((R, 2 * H + 2, 2 * W + 2, Cc), dtype=torch.bfloat16, device=x.device)
        _lib.call('as_upsample2_fwd_bf16', x.data_ptr(), y.data_ptr(), R, H, W, Cc, _lib.stream_ptr())
        ctx.dims = (R, H, W, Cc)
        return y

    @staticmethod
    def backward(ctx, dy):
        R, H, W, Cc = ctx.dims
        dy = dy.contiguous()
        dx = torch.empty((R, H + 2, W + 2, Cc), dtype=torch.bfloat16, device=dy.device)
        _lib.call('as_upsample2_bwd_bf16', dy.data_ptr(), dx.data_ptr(), R, H, W, Cc, _lib.stream_ptr())
        return dx


def upsample2(x):
    return Up2.apply(x)


class ToPNHWC(torch.autograd.Function):
    """NCHW fp32 -> padded NHWC bf16 (optionally zero-padding the channel count)."""

    @staticmethod
    def forward(ctx, x, cpad):
        R, Cc, H, W = x.shape
        x = x.contiguous().float()
        y = torch.empty((R, H + 2, W + 2, cpad), dtype=torch.bfloat16, device=x.device)
        _lib.call('as_nchw_to_pnhwc', x.data_ptr(), y.data_ptr(), R, Cc, cpad, H, W, 0, _lib.stream_ptr())
        ctx.dims = (R, Cc, H, W, cpad)
        return y

    @staticmethod
    def backward(ctx, dy):
        R, Cc, H, W, cpad = ctx.dims
        dy = dy.contiguous()
        dx = torch.empty((R, Cc, H, W), dtype=torch.float32, device=dy.device)
        _lib.call('as_pnhwc_to_nchw', dy.data_ptr(), dx.data_ptr(), R, Cc, cpad, H, W, _lib.stream_ptr())
        return dx, None


def to_pnhwc(x, cpad=None):
    return ToPNHWC.apply(x, cpad or x.shape[1])


class LocCat(torch.autograd.Function):
    @staticmethod
    def forward(ctx, emb, skip, cpad):
        R, ph, pw, Cs = skip.shape
        H, W = ph - 2, pw - 2
        emb = emb.contiguous().float()
        skip = skip.contiguous()
        out = torch.empty((R, ph, pw, cpad), dtype=torch.bfloat16, device=skip.device)
        _lib.call('as_loc_cat_fwd', emb.data_ptr(), skip.data_ptr(), out.data_ptr(), R, H, W, Cs, cpad, _lib.stream_ptr())
        ctx.dims = (R, H, W, Cs, cpad)
        ctx.save_for_backward(emb, skip)
        return out

    @staticmethod
    def backward(ctx, dcat):
        emb, skip = ctx.saved_tensors
        R, H, W, Cs, cpad = ctx.dims
        dcat = dcat.contiguous()
        demb = torch.empty_like(emb)
        dskip = torch.empty_like(skip)
        _lib.call('as_loc_cat_bwd', dcat.data_ptr(), emb.data_ptr(), skip.data_ptr(), demb.data_ptr(), dskip.data_ptr(),
                  R, H, W, Cs, cpad, _lib.stream_ptr())
        return demb, dskip, None


def loc_cat(emb, skip, cpad=192):
    return LocCat.apply(emb, skip, cpad)


class ConvC32To1(torch.autograd.Function):
    """3x3 conv 32 -> 1 + flatten + 1/temperature -> fp32 logits [R, H*W]."""

    @staticmethod
    def forward(ctx, x, weight, bias, inv_temp):
        R, ph, pw, Cc = x.shape
        H, W = ph - 2, pw - 2
        assert Cc == 32 and tuple(weight.shape) == (1, 32, 3, 3)
        x = x.contiguous()
        w9 = weight.detach().permute(0, 2, 3, 1).reshape(9, 32).contiguous().float()
        out = torch.empty((R, H * W), dtype=torch.float32, device=x.device)
        _lib.call('as_conv_c32_1_fwd', x.data_ptr(), w9.data_ptr(), _ptr(bias), out.data_ptr(), R, H, W, float(inv_temp),
                  _lib.stream_ptr())
        ctx.dims = (R, H, W)
        ctx.inv_temp = float(inv_temp)
        ctx.save_for_backward(x, w9)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, w9 = ctx.saved_tensors
        R, H, W = ctx.dims
        dy = dy.contiguous().float()
        dx = torch.empty_like(x)
        dw = torch.empty((9, 32), dtype=torch.float32, device=x.device)
        db = torch.empty(1, dtype=torch.float32, device=x.device)
        scratch = torch.empty(_lib.lib().as_conv_c32_1_scratch_floats(), dtype=torch.float32, device=x.device)
        _lib.call('as_conv_c32_1_bwd', dy.data_ptr(), x.data_ptr(), w9.data_ptr(), dx.data_ptr(), dw.data_ptr(),
                  db.data_ptr(), scratch.data_ptr(), R, H, W, ctx.inv_temp, _lib.stream_ptr())
        return dx, dw.view(3, 3, 32).permute(2, 0, 1).reshape(1, 32, 3, 3), db, None


def conv_c32_to_1(x, weight, bias, inv_temp=1.0):
    return ConvC32To1.apply(x, weight, bias, inv_temp)


class UpConvC32To1(torch.autograd.Function):
    """bilinear x2 + 3x3 conv 32 -> 1 + flatten + 1/temperature (action_arg_head.py:440-446) as one kernel per
    direction that never forms the upsampled map (csrc/upconv.cu).  relu_input: x is a ReLU output consumed only here,
    so the returned gradient already carries x's ReLU mask."""

    @staticmethod
    def forward(ctx, x, weight, bias, inv_temp, relu_input):
        R, ph, pw, Cc = x.shape
        h, w = ph - 2, pw - 2
        assert Cc == 32 and tuple(weight.shape) == (1, 32, 3, 3)
        x = x.contiguous()
        w9 = weight.detach().permute(0, 2, 3, 1).reshape(9, 32).contiguous().float()
        out = torch.empty((R, 4 * h * w), dtype=torch.float32, device=x.device)
        _lib.call('as_upconv_c32_1_fwd', x.data_ptr(), w9.data_ptr(), _ptr(bias), out.data_ptr(), R, h, w, float(inv_temp),
                  _lib.stream_ptr())
        ctx.dims = (R, h, w)
        ctx.inv_temp, ctx.relu_input = float(inv_temp), bool(relu_input)
        ctx.save_for_backward(x, w9)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, w9 = ctx.saved_tensors
        R, h, w = ctx.dims
        dy = dy.contiguous().float()
        dx = torch.empty_like(x)
        dw = torch.empty((9, 32), dtype=torch.float32, device=x.device)
        db = torch.empty(1, dtype=torch.float32, device=x.device)
        scratch = torch.empty(_lib.lib().as_upconv_c32_1_scratch_floats(R, h), dtype=torch.float32, device=x.device)
        _lib.call('as_upconv_c32_1_bwd', dy.data_ptr(), x.data_ptr(), w9.data_ptr(), dx.data_ptr(), dw.data_ptr(),
                  db.data_ptr(), scratch.data_ptr(), R, h, w, ctx.inv_temp, int(ctx.relu_input), _lib.stream_ptr())
        return dx, dw.view(3, 3, 32).permute(2, 0, 1).reshape(1, 32, 3, 3), db, None, None


def upconv_c32_to_1(x, weight, bias, inv_temp=1.0, relu_input=False):
    return UpConvC32To1.apply(x, weight, bias, inv_temp, relu_input)


def spatial_fc(x, weight, bias, ch, H, W):
    """ReLU(fc) of the spatial encoder (spatial_encoder.py:88-89) over a padded NHWC bf16 map x [R, H+2, W+2, ch].  The
    reference's weight is laid out for the NCHW flatten (column c*H*W + y*W + x); the GEMM reads a bf16 re-layout for
    the padded NHWC flatten, column ((y+1)*(W+2) + (x+1))*ch + c, border columns zero (as_spatial_fc_weight_fwd, cached
    until the weight changes), and the weight gradient of that layout is mapped back by as_spatial_fc_weight_bwd."""
    out = weight.shape[0]
    tag = (K._SHADOW_GEN[0], weight._version, weight.data_ptr())
    hit = getattr(weight, '_as_fc16', None)
    if hit is None or hit[0] != tag:
        w = weight.detach().contiguous()
        w16 = torch.empty((out, (H + 2) * (W + 2) * ch), dtype=torch.bfloat16, device=w.device)
        _lib.call('as_spatial_fc_weight_fwd', w.data_ptr(), w16.data_ptr(), out, ch, H, W, _lib.stream_ptr())
        hit = (tag, w16)
        try:
            weight._as_fc16 = hit
        except AttributeError:
            pass

    def unmap(dwp):
        dwp = dwp.contiguous()
        dw = torch.empty((out, ch * H * W), dtype=torch.float32, device=dwp.device)
        _lib.call('as_spatial_fc_weight_bwd', dwp.data_ptr(), dw.data_ptr(), out, ch, H, W, _lib.stream_ptr())
        return dw

    return K.linear(x.reshape(x.shape[0], -1), weight, bias, True, relayout=(hit[1], unmap))


class MaxPool2(torch.autograd.Function):
    """2x2 max-pool.  relu_input: x is a ReLU layer's output and this pool is its only consumer -- the backward then routes
    nothing through windows whose maximum is 0, i.e. the gradient it returns already carries that layer's ReLU mask."""

    @staticmethod
    def forward(ctx, x, relu_input=False):
        ctx.relu_input = bool(relu_input)
        R, ph, pw, Cc = x.shape
        H, W = ph - 2, pw - 2
        x = x.contiguous()
        y = torch.empty((R, H // 2 + 2, W // 2 + 2, Cc), dtype=torch.bfloat16, device=x.device)
        _lib.call('as_maxpool2_fwd_bf16', x.data_ptr(), y.data_ptr(), R, H, W, Cc, _lib.stream_ptr())
        ctx.dims = (R, H, W, Cc)
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        R, H, W, Cc = ctx.dims
        dy = dy.contiguous()
        dx = torch.empty_like(x)
        _lib.call('as_maxpool2_bwd_bf16', x.data_ptr(), dy.data_ptr(), dx.data_ptr(), R, H, W, Cc, int(ctx.relu_input),
                  _lib.stream_ptr())
        return dx, None


def max_pool2(x, relu_input=False):
    return MaxPool2.apply(x, relu_input)


_PLANES = ('height_map', 'visibility_map', 'creep', 'player_relative', 'alerts', 'pathable', 'buildable')
_EFFECTS = ('effect_PsiStorm', 'effect_NukeDot', 'effect_LiberatorDefenderZone', 'effect_BlindingCloud',
            'effect_CorrosiveBile', 'effect_LurkerSpines')


def _build_features(proj, x, y, entity_num, spatial_info):
    import ctypes
    R, N, D = proj.shape
    hm = spatial_info['height_map']
    H, W = hm.shape[1], hm.shape[2]
    fused = os.environ.get('APPLESTAR_SCATTER_FUSED', '1') != '0' and D == 32 and N <= 512
    if not fused:
        smap = K.ScatterConnection.apply(proj.detach(), x, y, entity_num, H, W, 'add')     # dense map: temporary
    planes = (ctypes.c_void_p * 7)()
    effects = (ctypes.c_void_p * 6)()
    keep = []
    for i, k in enumerate(_PLANES):
        t = spatial_info[k].contiguous()
        if t.dtype != torch.uint8:
            t = t.to(torch.uint8)
        keep.append(t)
        planes[i] = t.data_ptr()
    n_eff = spatial_info[_EFFECTS[0]].shape[1]
    for i, k in enumerate(_EFFECTS):
        t = spatial_info[k].contiguous()
        if t.dtype != torch.int16:
            t = t.to(torch.int16)
        keep.append(t)
        effects[i] = t.data_ptr()
    out = torch.empty((R, H + 2, W + 2, 64), dtype=torch.bfloat16, device=proj.device)
    if fused:      # scatter channels summed inside the feature kernel: the dense fp32 map is never written
        xs, ys = x.contiguous(), y.contiguous()
        if xs.dtype not in (torch.uint8, torch.int64):
            xs, ys = xs.long(), ys.long()
        en = None if entity_num is None else entity_num.contiguous().long()
        emb = proj.detach().float().contiguous()
        _lib.call('as_spatial_features_scatter', planes, effects, n_eff, emb.data_ptr(), xs.data_ptr(), ys.data_ptr(),
                  xs.element_size(), 0 if en is None else en.data_ptr(), N, out.data_ptr(), R, H, W, _lib.stream_ptr())
    else:
        _lib.call('as_spatial_features', planes, effects, n_eff, smap.data_ptr(), out.data_ptr(), R, H, W, D,
                  _lib.stream_ptr())
    return out, H, W


class SpatialProject(torch.autograd.Function):
    """ScatterConnection ('add') + 56-channel spatial input + 1x1 project conv (56 -> 32) + ReLU
    (encoder.py:37-42, spatial_encoder.py:52-72) -> zero-padded NHWC bf16 [R,H+2,W+2,32].
    Backward: weight / bias gradients by the tap-batched wgrad GEMM; the only data gradient that exists (towards the
    entity projections) is gathered at the <=512 entity cells and multiplied by the 32x32 scatter block of the
    weight, so the 3.3 GB full-resolution data gradient of the conv is never formed."""

    @staticmethod
    def forward(ctx, proj, x, y, entity_num, spatial_info, weight, bias):
        feats, H, W = _build_features(proj, x, y, entity_num, spatial_info)
        R, N, D = proj.shape
        cout = weight.shape[0]
        w_fwd, _ = _conv_weights(weight, 64)
        out = torch.empty((R, H + 2, W + 2, cout), dtype=torch.bfloat16, device=proj.device)
        _conv_call(feats, w_fwd, out, bias, R, H, W, 64, cout, 1, True)
        xs, ys = x.contiguous(), y.contiguous()
        if xs.dtype not in (torch.uint8, torch.int64):
            xs, ys = xs.long(), ys.long()
        en = None if entity_num is None else entity_num.contiguous().long()
        ctx.dims = (R, N, D, H, W, cout, weight.shape[1])
        ctx.save_for_backward(feats, out, xs, ys, en, weight)
        return out

    @staticmethod
    def backward(ctx, dy):
        feats, out, xs, ys, en, weight = ctx.saved_tensors
        R, N, D, H, W, cout, cin = ctx.dims
        st = _lib.stream_ptr()
        dy = dy.contiguous()
        dz = torch.empty_like(dy)
        P = R * (H + 2) * (W + 2)
        db = torch.empty(cout, dtype=torch.float32, device=dy.device)
        scratch = torch.empty(_lib.lib().as_colsum_scratch_floats(cout), dtype=torch.float32, device=dy.device)
        _lib.call('as_relu_bwd_colsum_bf16', dy.data_ptr(), out.data_ptr(), dz.data_ptr(), db.data_ptr(),
                  scratch.data_ptr(), P, cout, st)
        splits = _lib.lib().as_gemm_splits(P, K.NUM_SMS)
        dw1 = torch.empty((1, cout, 64), dtype=torch.float32, device=dy.device)
        ws = torch.empty((splits, 1, cout, 64), dtype=torch.float32, device=dy.device) if splits > 1 else None
        _lib.call('as_conv_wgrad_nhwc_bf16', dz.data_ptr(), feats.data_ptr(), dw1.data_ptr(), _ptr(ws), R, H, W, 64, cout,
                  1, splits, st)
        dw = dw1[0, :, :cin].reshape(cout, cin, 1, 1)
        g = torch.empty((R, N, cout), dtype=torch.float32, device=dy.device)       # dz rows at the entity cells
        _lib.call('as_scatter_connection_bwd_pnhwc', dz.data_ptr(), xs.data_ptr(), ys.data_ptr(), xs.element_size(),
                  _ptr(en), g.data_ptr(), R, N, cout, H, W, cout, 0, st)
        w_s = weight.detach()[:, 24:24 + D, 0, 0]                                  # [cout, D] scatter block
        dproj = (g.view(R * N, cout) @ w_s).view(R, N, D)
        return dproj, None, None, None, None, dw, db


def spatial_project(proj, x, y, entity_num, spatial_info, weight, bias):
    return SpatialProject.apply(proj, x, y, entity_num, spatial_info, weight, bias)


def _plane_ptrs(spatial_info):
    planes = (ctypes.c_void_p * 7)()
    effects = (ctypes.c_void_p * 6)()
    keep = []
    for i, k in enumerate(_PLANES):
        t = spatial_info[k].contiguous()
        if t.dtype != torch.uint8:
            t = t.to(torch.uint8)
        keep.append(t)
        planes[i] = t.data_ptr()
    for i, k in enumerate(_EFFECTS):
        t = spatial_info[k].contiguous()
        if t.dtype != torch.int16:
            t = t.to(torch.int16)
        keep.append(t)
        effects[i] = t.data_ptr()
    return planes, effects, keep, spatial_info[_EFFECTS[0]].shape[1]


class SpatialFront(torch.autograd.Function):
    """ScatterConnection ('add') + 56-channel spatial input + 1x1 project conv (56 -> 32) + ReLU + the first 2x2 max-pool
    (encoder.py:37-42, spatial_encoder.py:52-72,79) as ONE kernel per direction (csrc/spatial_front.cu): the one-hot
    input makes the convolution two table look-ups per pixel, evaluated exactly in fp32; only the pooled map
    [R, H/2+2, W/2+2, 32] (zero-padded NHWC bf16) touches HBM.  Backward: the kernel recomputes the windows, routes the
    pooled gradient, and returns the weight / bias gradients and the per-entity gradient rows; d proj and the scatter
    block of d weight are two small GEMMs on them."""

    @staticmethod
    def forward(ctx, proj, x, y, entity_num, spatial_info, weight, bias):
        R, N, D = proj.shape
        planes, effects, keep, n_eff = _plane_ptrs(spatial_info)
        H, W = keep[0].shape[1], keep[0].shape[2]
        xs, ys = x.contiguous(), y.contiguous()
        if xs.dtype not in (torch.uint8, torch.int64):
            xs, ys = xs.long(), ys.long()
        en = None if entity_num is None else entity_num.contiguous().long()
        emb = proj.detach().float().contiguous()
        w2 = weight.detach().reshape(weight.shape[0], -1).float().contiguous()
        b1 = bias.detach().float().contiguous()
        out = torch.empty((R, H // 2 + 2, W // 2 + 2, 32), dtype=torch.bfloat16, device=proj.device)
        cells = torch.empty(_lib.lib().as_spatial_front_cells_bytes(R), dtype=torch.uint8, device=proj.device)
        need_bwd = any(ctx.needs_input_grad)
        codes = torch.empty((R, H // 2, W // 2, 4), dtype=torch.int32, device=proj.device) if need_bwd else None
        _lib.call('as_spatial_front_fwd', planes, effects, n_eff, emb.data_ptr(), xs.data_ptr(), ys.data_ptr(),
                  xs.element_size(), _ptr(en), N, w2.data_ptr(), b1.data_ptr(), cells.data_ptr(), out.data_ptr(), _ptr(codes),
                  R, H, W, _lib.stream_ptr())
        _lib.LAUNCHES += 2
        ctx.dims = (R, N, D, H, W, n_eff)
        ctx.wshape = weight.shape
        ctx.save_for_backward(emb, xs, ys, en, w2, b1, cells, codes, *keep)
        return out

    @staticmethod
    def backward(ctx, dpool):
        emb, xs, ys, en, w2, b1, cells, codes = ctx.saved_tensors[:8]
        keep = ctx.saved_tensors[8:]
        R, N, D, H, W, n_eff = ctx.dims
        dev = dpool.device
        planes = (ctypes.c_void_p * 7)(*[t.data_ptr() for t in keep[:7]])
        effects = (ctypes.c_void_p * 6)(*[t.data_ptr() for t in keep[7:13]])
        dpool = dpool.contiguous()
        dw = torch.empty((32, 56), dtype=torch.float32, device=dev)
        db = torch.empty(32, dtype=torch.float32, device=dev)
        g = torch.empty((R, N, 32), dtype=torch.float32, device=dev)
        scratch = torch.empty(_lib.lib().as_spatial_front_bwd_scratch_floats(R, H), dtype=torch.float32, device=dev)
        _lib.call('as_spatial_front_bwd', planes, effects, n_eff, emb.data_ptr(), xs.data_ptr(), ys.data_ptr(),
                  xs.element_size(), _ptr(en), N, w2.data_ptr(), b1.data_ptr(), cells.data_ptr(), codes.data_ptr(), dpool.data_ptr(),
                  dw.data_ptr(), db.data_ptr(), g.data_ptr(), scratch.data_ptr(), R, H, W, _lib.stream_ptr())
        _lib.LAUNCHES += 2
        g16 = K.to_bf16(g.view(R * N, 32))                                   # exact: the rows are bf16 gradients
        dproj = None
        if ctx.needs_input_grad[0]:
            ws16 = K.to_bf16(w2[:, 24:56].contiguous())                      # [o, c]
            dproj = torch.empty((R * N, 32), dtype=torch.float32, device=dev)
            K.gemm(g16, ws16, dproj, R * N, 32, 32, lda=32, b_mn=True, ldb=32)
            dproj = dproj.view(R, N, D)
        dw[:, 24:56] = K._wgrad(g16, K.to_bf16(emb.view(R * N, 32)), 32, 32)
        return dproj, None, None, None, None, dw.view(ctx.wshape), db


def spatial_front(proj, x, y, entity_num, spatial_info, weight, bias):
    return SpatialFront.apply(proj, x, y, entity_num, spatial_info, weight, bias)
