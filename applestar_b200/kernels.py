"""torch.autograd bindings of the C-ABI kernels (applestar_b200/_lib.py).

Tensors stay torch tensors only so that torch owns device memory, streams and the autograd tape; every
function here hands raw `data_ptr()`s to `libapplestar_b200.so` on `torch.cuda.current_stream()`.
"""
import math
import os

import torch
import torch.nn.functional as F

from . import _lib

NUM_SMS = 148
_DT_CODE = {torch.uint8: 0, torch.int8: 1, torch.int16: 2, torch.float16: 3, torch.int64: 4, torch.float32: 5}


def _ptr(t):
    return 0 if t is None else t.data_ptr()


# ---------------------------------------------------------------------------------------------
# bf16 staging
def to_bf16(x):
    """fp32 (or any narrow storage dtype) [rows, K] -> bf16 [rows, K8] (K padded to a multiple of 8 with zeros so TMA
    pitches are 16-byte)."""
    if x.dtype in _DT_CODE and x.dtype != torch.float32 and x.dim() == 2:      # uint8 / int16 / ... straight to the operand
        x = x.contiguous()
        k8 = (x.shape[1] + 7) // 8 * 8
        out = torch.empty((x.shape[0], k8), dtype=torch.bfloat16, device=x.device)
        _lib.call('as_cast_any_bf16', x.data_ptr(), _DT_CODE[x.dtype], out.data_ptr(), x.shape[0], x.shape[1], k8,
                  _lib.stream_ptr())
        return out
    if x.dtype == torch.bfloat16:
        return x if x.shape[-1] % 8 == 0 and x.is_contiguous() else _pad8(x.contiguous())
    x = x.contiguous()
    if x.shape[-1] % 8 != 0:
        x = _pad8(x)
    out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    _lib.call('as_cast_f32_bf16', x.data_ptr(), out.data_ptr(), x.numel(), _lib.stream_ptr())
    return out


def _pad8(x):
    k = x.shape[-1]
    return F.pad(x, (0, (-k) % 8))


_SHADOW_GEN = [0]          # bumped by the optimizer whenever weights change through raw pointers


def weights_changed():
    _SHADOW_GEN[0] += 1


def weight_bf16(w):
    """bf16 copy of a weight matrix [N, K] (K padded to 8).  Parameters that live in a FlatParams buffer carry a
    view of the flat bf16 shadow the fused Adam kernel keeps fresh; the view is re-cast here whenever the parameter
    was changed behind its back (torch optimizers, `load_state_dict`, `copy_`: anything that bumps `_version`).
    Otherwise the copy is cached on the Parameter object itself (never keyed by address: freed addresses get reused
    by other models)."""
    sh = getattr(w, '_as_shadow', None)
    if sh is not None and w.shape[-1] % 8 == 0 and w.dim() == 2:
        if w._version != getattr(w, '_as_shadow_ver', None):
            _lib.call('as_cast_f32_bf16', w.data_ptr(), sh.data_ptr(), w.numel(), _lib.stream_ptr())
            w._as_shadow_ver = w._version
        return sh
    tag = (_SHADOW_GEN[0], w._version, w.data_ptr())
    hit = getattr(w, '_as_w16', None)
    if hit is not None and hit[0] == tag:
        return hit[1]
    w16 = to_bf16(w.detach().reshape(w.shape[0], -1))
    try:
        w._as_w16 = (tag, w16)
    except AttributeError:
        pass
    return w16


# ---------------------------------------------------------------------------------------------
def gemm(A, B, C, M, N, K, a_mn=False, b_mn=False, lda=None, ldb=None, ldc=None, bias=None, relu=False,
         alpha=1.0, batch=(1, 1), a_strides=(0, 0), b_strides=(0, 0), c_strides=(0, 0), splits=1, ws=None, C2=None,
         b_k_shift=0):
    """Raw tcgen05 GEMM launch: C[b][m][n] = act(alpha * sum_k A[b][m][k] B[b][n][k] + bias[n])."""
    lda = lda if lda is not None else (M if a_mn else K)
    ldb = ldb if ldb is not None else (N if b_mn else K)
    ldc = ldc if ldc is not None else N
    _lib.call('as_gemm_bf16', A.data_ptr(), B.data_ptr(), _ptr(C), _ptr(C2), _ptr(bias), M, N, K, int(a_mn), int(b_mn),
              lda, ldb, ldc, batch[0], batch[1], a_strides[0], a_strides[1], b_strides[0], b_strides[1],
              c_strides[0], c_strides[1], int(C is not None and C.dtype == torch.bfloat16), int(relu), float(alpha),
              splits, _ptr(ws), b_k_shift, _lib.stream_ptr())


def _wgrad(dy16, x16, n_out, k_in):
    """dW[n_out, k_in] = dY^T X with both operands read MN-major (no transposes); split-K over the rows."""
    M = dy16.shape[0]
    tiles = ((n_out + 127) // 128) * ((k_in + 255) // 256 if k_in > 128 else 1)
    want = max(1, min((M + 63) // 64, NUM_SMS // max(tiles, 1)))
    splits = _lib.lib().as_gemm_splits(M, want)
    dw = torch.empty((n_out, k_in), dtype=torch.float32, device=dy16.device)
    if splits > 1:
        ws = torch.empty((splits, n_out, k_in), dtype=torch.float32, device=dy16.device)
        gemm(dy16, x16, None, n_out, k_in, M, a_mn=True, b_mn=True, lda=dy16.shape[1], ldb=x16.shape[1],
             splits=splits, ws=ws)
        _lib.call('as_splitk_reduce', ws.data_ptr(), dw.data_ptr(), dw.numel(), splits, 0, _lib.stream_ptr())
    else:
        gemm(dy16, x16, dw, n_out, k_in, M, a_mn=True, b_mn=True, lda=dy16.shape[1], ldb=x16.shape[1])
    return dw


class LinearFn(torch.autograd.Function):
    """y = act(x W^T + b): forward, dgrad and wgrad all on the tcgen05 GEMM.  x may be fp32 or bf16.  Returns
    (y fp32 | None, y bf16 | None): the bf16 copy (written by the same epilogue) feeds a following GEMM directly."""

    @staticmethod
    def forward(ctx, x, weight, bias, relu, want32, want16, relayout=None):
        """`relayout` = (w16, unmap): the GEMM reads `w16` (a bf16 re-layout of `weight` for x's column order) and the
        weight gradient of that layout goes through `unmap` to become d `weight`."""
        ctx.set_materialize_grads(False)        # an unused output copy must not cost a zero fill + an add in backward
        K = x.shape[-1]
        N = weight.shape[0]
        x2 = x.reshape(-1, K)
        M = x2.shape[0]
        x16 = to_bf16(x2)
        w16 = weight_bf16(weight) if relayout is None else relayout[0]
        ctx.unmap = None if relayout is None else relayout[1]
        y32 = torch.empty((M, N), dtype=torch.float32, device=x.device) if want32 else None
        y16 = torch.empty((M, N), dtype=torch.bfloat16, device=x.device) if want16 else None
        if want32:
            gemm(x16, w16, y32, M, N, x16.shape[1], lda=x16.shape[1], ldb=w16.shape[1], bias=bias, relu=relu, C2=y16)
        else:
            gemm(x16, w16, y16, M, N, x16.shape[1], lda=x16.shape[1], ldb=w16.shape[1], bias=bias, relu=relu)
        ctx.relu = relu
        ctx.wk = weight.shape[1] if weight.dim() == 2 else x.shape[-1]
        ctx.xk = K
        ctx.xshape = x.shape
        ctx.x_bf16 = x.dtype == torch.bfloat16
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x16, w16, (y32 if want32 else y16) if relu else None)
        shp = (*x.shape[:-1], N)
        return (None if y32 is None else y32.view(shp)), (None if y16 is None else y16.view(shp))

    @staticmethod
    def backward(ctx, dy32, dy16):
        x16, w16, y = ctx.saved_tensors
        N = w16.shape[0]
        K8 = w16.shape[1]
        if dy32 is None and dy16 is None:
            return None, None, None, None, None, None, None
        extra = None
        if dy32 is not None and dy16 is not None:
            if (dy32.dtype == torch.float32 and dy16.dtype == torch.bfloat16 and N % 8 == 0 and dy32.is_contiguous()
                    and dy16.is_contiguous() and (y is None or y.is_contiguous())):
                dy, extra = dy32, dy16.reshape(-1, N)         # summed inside as_linear_bwd_prep2
            else:
                dy = dy32 + dy16
        else:
            dy = dy32 if dy32 is not None else dy16
        dy2 = dy.reshape(-1, N)
        M = dy2.shape[0]
        dx = dw = db = None
        want_db = ctx.has_bias and ctx.needs_input_grad[2]
        if dy2.dtype == torch.float32 and N % 8 == 0 and dy2.is_contiguous() and (y is None or y.is_contiguous()):
            # ReLU mask + bf16 cast + bias gradient in one pass over dy
            dy16b = torch.empty((M, N), dtype=torch.bfloat16, device=dy.device)
            y32 = y if (y is not None and y.dtype == torch.float32) else None
            y16 = y if (y is not None and y.dtype == torch.bfloat16) else None
            if want_db:
                db = torch.empty(N, dtype=torch.float32, device=dy.device)
                scratch = torch.empty(_lib.lib().as_colsum_scratch_floats(N), dtype=torch.float32, device=dy.device)
            _lib.call('as_linear_bwd_prep2', dy2.data_ptr(), _ptr(extra), _ptr(y32), _ptr(y16), dy16b.data_ptr(), _ptr(db),
                      _ptr(scratch) if want_db else 0, M, N, _lib.stream_ptr())
            want_db = False
        elif (ctx.relu and want_db and dy2.dtype == torch.bfloat16 and y is not None and y.dtype == torch.bfloat16
              and N % 8 == 0 and dy2.is_contiguous() and y.is_contiguous()):
            # bf16 gradient of a ReLU layer with bias (FFN hidden): mask and bias gradient in one pass
            dy16b = torch.empty((M, N), dtype=torch.bfloat16, device=dy.device)
            db = torch.empty(N, dtype=torch.float32, device=dy.device)
            scratch = torch.empty(_lib.lib().as_colsum_scratch_floats(N), dtype=torch.float32, device=dy.device)
            _lib.call('as_relu_bwd_colsum_bf16', dy2.data_ptr(), y.data_ptr(), dy16b.data_ptr(), db.data_ptr(),
                      scratch.data_ptr(), M, N, _lib.stream_ptr())
            dy2 = dy16b
            want_db = False
        else:
            if ctx.relu:
                dy2 = relu_backward(dy2, y)
            dy16b = to_bf16(dy2)                      # [M, N8]
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K8), dtype=torch.bfloat16 if ctx.x_bf16 else torch.float32, device=dy.device)
            # dX[m,k] = sum_n dY[m,n] W[n,k]: A = dY (K-major over n), B = W read MN-major (k contiguous)
            gemm(dy16b, w16, dx, M, K8, N, lda=dy16b.shape[1], b_mn=True, ldb=K8)
            if K8 != ctx.xk:
                dx = dx[:, :ctx.xk]
            dx = dx.reshape(ctx.xshape)
        if ctx.needs_input_grad[1]:
            dw = _wgrad(dy16b, x16, N, K8)
            if ctx.unmap is not None:
                dw = ctx.unmap(dw)
            elif K8 != ctx.wk:
                dw = dw[:, :ctx.wk]
        if want_db:
            db = column_sum(dy2)
        return dx, dw, db, None, None, None, None


def column_sum(x):
    """fp32 column sums of a [rows, C] fp32 / bf16 matrix (bias gradients), deterministic two-stage reduction."""
    x = x.contiguous()
    rows, Cc = x.shape
    if x.dtype == torch.bfloat16 and Cc % 2:
        return x.sum(dim=0, dtype=torch.float32)
    out = torch.empty(Cc, dtype=torch.float32, device=x.device)
    scratch = torch.empty(_lib.lib().as_colsum_scratch_floats(Cc), dtype=torch.float32, device=x.device)
    name = 'as_colsum_bf16' if x.dtype == torch.bfloat16 else 'as_colsum_f32'
    if x.dtype not in (torch.bfloat16, torch.float32):
        x = x.float()
    _lib.call(name, x.data_ptr(), out.data_ptr(), scratch.data_ptr(), rows, Cc, _lib.stream_ptr())
    return out


def relu_backward(dy, y):
    """dy * (y > 0) with one kernel (bf16 or fp32 pair); mixed dtypes fall back to promoting dy."""
    dy, y = dy.contiguous(), y.contiguous()
    if dy.dtype != y.dtype:
        dy = dy.to(y.dtype) if y.dtype == torch.float32 else dy
        y = y if dy.dtype == y.dtype else y.float()
    n = dy.numel()
    dz = torch.empty_like(dy)
    if dy.dtype == torch.bfloat16 and n % 8 == 0:
        _lib.call('as_relu_bwd_bf16', dy.data_ptr(), y.data_ptr(), dz.data_ptr(), n, _lib.stream_ptr())
    elif dy.dtype == torch.float32:
        _lib.call('as_relu_bwd_f32', dy.data_ptr(), y.data_ptr(), dz.data_ptr(), n, _lib.stream_ptr())
    else:
        return dy * (y > 0)
    return dz


def linear(x, weight, bias, relu=False, out_bf16=False, relayout=None):
    y32, y16 = LinearFn.apply(x, weight, bias, relu, not out_bf16, out_bf16, relayout)
    return y16 if out_bf16 else y32


def linear_dual(x, weight, bias, relu=False):
    """-> (fp32, bf16) copies of the same output."""
    return LinearFn.apply(x, weight, bias, relu, True, True, None)


# ---------------------------------------------------------------------------------------------
class ScatterConnection(torch.autograd.Function):
    @staticmethod
    def forward(ctx, proj, x, y, entity_num, H, W, scatter_type):
        R, N, D = proj.shape
        proj = proj.contiguous().float()
        x = x.contiguous()
        y = y.contiguous()
        if x.dtype not in (torch.uint8, torch.int64):
            x, y = x.long(), y.long()
        en = None if entity_num is None else entity_num.contiguous().long()
        out = torch.empty((R, D, H, W), dtype=torch.float32, device=proj.device)
        _lib.call('as_scatter_connection_fwd', proj.data_ptr(), x.data_ptr(), y.data_ptr(), x.element_size(),
                  _ptr(en), out.data_ptr(), R, N, D, H, W, int(scatter_type == 'cover'), _lib.stream_ptr())
        if scatter_type not in ('add', 'cover'):
            raise NotImplementedError(scatter_type)
        ctx.dims = (R, N, D, H, W)
        ctx.scatter_type = scatter_type
        ctx.save_for_backward(x, y, en)
        return out

    @staticmethod
    def backward(ctx, dout):
        if ctx.scatter_type != 'add':
            raise NotImplementedError("backward is defined for scatter_type='add' (the model's setting)")
        x, y, en = ctx.saved_tensors
        R, N, D, H, W = ctx.dims
        dout = dout.contiguous()
        dproj = torch.empty((R, N, D), dtype=torch.float32, device=dout.device)
        _lib.call('as_scatter_connection_bwd', dout.data_ptr(), x.data_ptr(), y.data_ptr(), x.element_size(),
                  _ptr(en), dproj.data_ptr(), R, N, D, H, W, _lib.stream_ptr())
        return dproj, None, None, None, None, None, None


def gumbel_argmax(logits, noise):
    logits = logits.detach().contiguous().float()
    noise = noise.contiguous().float()
    rows, K = logits.shape
    out = torch.empty((rows,), dtype=torch.int64, device=logits.device)
    _lib.call('as_gumbel_argmax', logits.data_ptr(), noise.data_ptr(), out.data_ptr(), rows, K, _lib.stream_ptr())
    return out


# ---------------------------------------------------------------------------------------------
class LayerNormFn(torch.autograd.Function):
    """LN(x (+ residual)) (+ReLU) with warp-shuffle statistics.  Returns (y fp32, y bf16): the fp32 copy carries the
    residual stream, the bf16 copy feeds the next GEMM without a separate cast pass.  Either can be skipped."""

    @staticmethod
    def forward(ctx, x, residual, weight, bias, relu, want32, want16):
        ctx.set_materialize_grads(False)
        C = x.shape[-1]
        x = x.contiguous().float()
        res = None if residual is None else residual.contiguous().float()
        rows = x.numel() // C
        y32 = torch.empty_like(x) if want32 else None
        y16 = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device) if want16 else None
        stats = torch.empty((rows, 2), dtype=torch.float32, device=x.device)
        _lib.call('as_layer_norm_fwd', x.data_ptr(), _ptr(res), weight.data_ptr(), bias.data_ptr(), _ptr(y32),
                  _ptr(y16), stats.data_ptr(), rows, C, int(relu), _lib.stream_ptr())
        ctx.relu, ctx.C, ctx.rows = relu, C, rows
        ctx.has_res = res is not None
        ctx.save_for_backward(x, res, weight, bias, stats)
        return y32, y16

    @staticmethod
    def backward(ctx, dy32, dy16):
        x, res, weight, bias, stats = ctx.saved_tensors
        C, rows = ctx.C, ctx.rows
        if dy32 is None and dy16 is None:
            return None, None, None, None, None, None, None
        dy32 = None if dy32 is None else dy32.contiguous()
        dy16 = None if dy16 is None else dy16.contiguous()
        dx = torch.empty_like(x) if (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]) else None
        need_w = ctx.needs_input_grad[2] or ctx.needs_input_grad[3]
        dg = torch.empty_like(weight) if need_w else None
        db = torch.empty_like(bias) if need_w else None
        scratch = torch.empty(_lib.lib().as_layer_norm_bwd_scratch_floats(C), dtype=torch.float32,
                              device=x.device) if need_w else None
        _lib.call('as_layer_norm_bwd', _ptr(dy32), _ptr(dy16), x.data_ptr(), _ptr(res), weight.data_ptr(),
                  bias.data_ptr(), stats.data_ptr(), _ptr(dx), _ptr(dg), _ptr(db), _ptr(scratch), rows, C,
                  int(ctx.relu), _lib.stream_ptr())
        return dx, (dx if ctx.has_res else None), dg, db, None, None, None


def layer_norm(x, weight, bias, residual=None, relu=False):
    return LayerNormFn.apply(x, residual, weight, bias, relu, True, False)[0]


def layer_norm_dual(x, weight, bias, residual=None, relu=False):
    """-> (fp32, bf16) copies of the same normalised rows."""
    return LayerNormFn.apply(x, residual, weight, bias, relu, True, True)


def _ln_fwd(z32, res, gamma, beta, relu):
    rows, C = z32.shape
    y32 = torch.empty_like(z32)
    y16 = torch.empty(z32.shape, dtype=torch.bfloat16, device=z32.device)
    stats = torch.empty((rows, 2), dtype=torch.float32, device=z32.device)
    _lib.call('as_layer_norm_fwd', z32.data_ptr(), _ptr(res), gamma.data_ptr(), beta.data_ptr(), y32.data_ptr(),
              y16.data_ptr(), stats.data_ptr(), rows, C, int(relu), _lib.stream_ptr())
    return y32, y16, stats


def _ln_bwd_fused(dy32, dy16, z32, res, gamma, beta, stats, relu_ln, mask_z, want_dres):
    """-> (dres fp32 | None, dz16, dgamma, dbeta, dbias): as_layer_norm_bwd_fused."""
    rows, C = z32.shape
    dev = z32.device
    dres = torch.empty_like(z32) if want_dres else None
    dz16 = torch.empty(z32.shape, dtype=torch.bfloat16, device=dev)
    dgamma, dbeta, dbias = (torch.empty(C, dtype=torch.float32, device=dev) for _ in range(3))
    scratch = torch.empty(_lib.lib().as_layer_norm_bwd_fused_scratch_floats(C), dtype=torch.float32, device=dev)
    dy32 = None if dy32 is None else dy32.contiguous().reshape(rows, C)
    dy16 = None if dy16 is None else dy16.contiguous().reshape(rows, C)
    _lib.call('as_layer_norm_bwd_fused', _ptr(dy32), _ptr(dy16), z32.data_ptr(), _ptr(res), gamma.data_ptr(), beta.data_ptr(),
              stats.data_ptr(), _ptr(dres), dz16.data_ptr(), int(mask_z), dgamma.data_ptr(), dbeta.data_ptr(),
              dbias.data_ptr(), scratch.data_ptr(), rows, C, int(relu_ln), _lib.stream_ptr())
    return dres, dz16, dgamma, dbeta, dbias


class LinearLNFn(torch.autograd.Function):
    """(y fp32, y bf16) = LN(res + x W^T + b) (+ReLU) -- attention projection + layernorm1 of the post-LN Transformer
    layer (module_utils.py:133-135) as one autograd node: the LayerNorm backward kernel hands the projection's backward
    its bf16 operand and bias gradient directly (as_layer_norm_bwd_fused)."""

    @staticmethod
    def forward(ctx, x, res, weight, bias, gamma, beta, relu_ln):
        ctx.set_materialize_grads(False)
        Kd, N = x.shape[-1], weight.shape[0]
        x16 = to_bf16(x.reshape(-1, Kd))
        M = x16.shape[0]
        w16 = weight_bf16(weight)
        z32 = torch.empty((M, N), dtype=torch.float32, device=x.device)
        gemm(x16, w16, z32, M, N, x16.shape[1], lda=x16.shape[1], ldb=w16.shape[1], bias=bias)
        res2 = res.reshape(M, N).contiguous().float()
        y32, y16, stats = _ln_fwd(z32, res2, gamma, beta, relu_ln)
        ctx.relu_ln, ctx.xshape, ctx.x_bf16 = relu_ln, x.shape, x.dtype == torch.bfloat16
        ctx.save_for_backward(x16, w16, z32, res2, gamma, beta, stats)
        shp = (*x.shape[:-1], N)
        return y32.view(shp), y16.view(shp)

    @staticmethod
    def backward(ctx, dy32, dy16):
        x16, w16, z32, res2, gamma, beta, stats = ctx.saved_tensors
        if dy32 is None and dy16 is None:
            return (None,) * 7
        M, N = z32.shape
        K8 = w16.shape[1]
        dres, dz16, dgamma, dbeta, dbias = _ln_bwd_fused(dy32, dy16, z32, res2, gamma, beta, stats, ctx.relu_ln, False,
                                                         ctx.needs_input_grad[1])
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K8), dtype=torch.bfloat16 if ctx.x_bf16 else torch.float32, device=z32.device)
            gemm(dz16, w16, dx, M, K8, N, lda=N, b_mn=True, ldb=K8)
            dx = dx.reshape(ctx.xshape)
        if ctx.needs_input_grad[2]:
            dw = _wgrad(dz16, x16, N, K8)
        return dx, (None if dres is None else dres.view(*ctx.xshape[:-1], N)), dw, dbias, dgamma, dbeta, None


def _relu_bits_gemm(x16, w16, h16, bias, M, Hd, Kd):
    """h16 = relu(x16 W^T + b) and the sign bits of h16 ([M, Hd / 32] words) in one launch: the backward's masked
    data-gradient GEMM then reads 1 bit instead of 16 per hidden element."""
    assert Hd % 32 == 0
    bits = torch.empty((M, Hd // 32), dtype=torch.int32, device=x16.device)
    _lib.call('as_gemm_bf16_relu_bits', x16.data_ptr(), w16.data_ptr(), h16.data_ptr(), _ptr(bias), bits.data_ptr(), M, Hd, Kd,
              Kd, Kd, Hd, _lib.stream_ptr())
    return bits


class MLPLNFn(torch.autograd.Function):
    """(y fp32, y bf16) = LN(res + ReLU(ReLU(x W1^T + b1) W2^T + b2)) (+ReLU) -- the FFN (ReLU on BOTH layers,
    module_utils.py:121-123) + layernorm2 of the post-LN Transformer layer (:136-139) as one autograd node.  Backward:
    the LayerNorm kernel emits the masked bf16 gradient and bias gradient of the second layer; the data-gradient GEMM of
    the second layer applies the first layer's ReLU mask in its epilogue (as_gemm_bf16_bitmasked, sign bits written by the
    forward GEMM), so the [rows, hidden]
    gradient is written once and never re-read by an element-wise pass."""

    @staticmethod
    def forward(ctx, x, res, w1, b1, w2, b2, gamma, beta, relu_ln):
        ctx.set_materialize_grads(False)
        Kd, Hd, N = x.shape[-1], w1.shape[0], w2.shape[0]
        x16 = to_bf16(x.reshape(-1, Kd))
        M = x16.shape[0]
        w1_16, w2_16 = weight_bf16(w1), weight_bf16(w2)
        h16 = torch.empty((M, Hd), dtype=torch.bfloat16, device=x.device)
        hbits = _relu_bits_gemm(x16, w1_16, h16, b1, M, Hd, Kd)
        z32 = torch.empty((M, N), dtype=torch.float32, device=x.device)
        gemm(h16, w2_16, z32, M, N, Hd, lda=Hd, ldb=Hd, bias=b2, relu=True)
        res2 = res.reshape(M, N).contiguous().float()
        y32, y16, stats = _ln_fwd(z32, res2, gamma, beta, relu_ln)
        ctx.relu_ln, ctx.xshape, ctx.x_bf16 = relu_ln, x.shape, x.dtype == torch.bfloat16
        ctx.save_for_backward(x16, w1_16, w2_16, h16, z32, res2, gamma, beta, stats, hbits)
        shp = (*x.shape[:-1], N)
        return y32.view(shp), y16.view(shp)

    @staticmethod
    def backward(ctx, dy32, dy16):
        x16, w1_16, w2_16, h16, z32, res2, gamma, beta, stats, hbits = ctx.saved_tensors
        if dy32 is None and dy16 is None:
            return (None,) * 9
        M, N = z32.shape
        Hd, Kd = w1_16.shape
        dev = z32.device
        dres, dz2, dgamma, dbeta, db2 = _ln_bwd_fused(dy32, dy16, z32, res2, gamma, beta, stats, ctx.relu_ln, True,
                                                      ctx.needs_input_grad[1])
        # d hidden = (dz2 W2) * (h > 0): mask applied in the GEMM epilogue
        dz1 = torch.empty((M, Hd), dtype=torch.bfloat16, device=dev)
        _lib.call('as_gemm_bf16_bitmasked', dz2.data_ptr(), w2_16.data_ptr(), dz1.data_ptr(), hbits.data_ptr(), M, Hd, N, 0, 1,
                  N, Hd, Hd, 1, 1.0, _lib.stream_ptr())
        db1 = column_sum(dz1) if ctx.needs_input_grad[3] else None
        dw2 = _wgrad(dz2, h16, N, Hd) if ctx.needs_input_grad[4] else None
        dw1 = _wgrad(dz1, x16, Hd, Kd) if ctx.needs_input_grad[2] else None
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, Kd), dtype=torch.bfloat16 if ctx.x_bf16 else torch.float32, device=dev)
            gemm(dz1, w1_16, dx, M, Kd, Hd, lda=Hd, b_mn=True, ldb=Kd)
            dx = dx.reshape(ctx.xshape)
        return (dx, (None if dres is None else dres.view(*ctx.xshape[:-1], N)), dw1, db1, dw2, db2, dgamma, dbeta, None)


def _ln16_fwd(z16, res16, gamma, beta, relu):
    rows = z16.shape[0]
    y16 = torch.empty_like(z16)
    stats = torch.empty((rows, 2), dtype=torch.float32, device=z16.device)
    _lib.call('as_ln256_bf16_fwd', z16.data_ptr(), _ptr(res16), gamma.data_ptr(), beta.data_ptr(), y16.data_ptr(),
              stats.data_ptr(), rows, int(relu), _lib.stream_ptr())
    return y16, stats


def _ln16_bwd(dya, dyb, z16, res16, gamma, beta, stats, relu_ln, mask_z, want_dres):
    rows, C = z16.shape
    dev = z16.device
    if dya is None:
        dya, dyb = dyb, None
    dya = dya.contiguous().reshape(rows, C)
    dyb = None if dyb is None else dyb.contiguous().reshape(rows, C)
    if dya.dtype != torch.bfloat16:
        dya = dya.to(torch.bfloat16)
    if dyb is not None and dyb.dtype != torch.bfloat16:
        dyb = dyb.to(torch.bfloat16)
    dres = torch.empty_like(z16) if want_dres else None
    dz16 = torch.empty_like(z16)
    dgamma, dbeta, dbias = (torch.empty(C, dtype=torch.float32, device=dev) for _ in range(3))
    scratch = torch.empty(_lib.lib().as_ln256_bf16_bwd_scratch_floats(), dtype=torch.float32, device=dev)
    _lib.call('as_ln256_bf16_bwd', dya.data_ptr(), _ptr(dyb), z16.data_ptr(), _ptr(res16), gamma.data_ptr(), beta.data_ptr(),
              stats.data_ptr(), _ptr(dres), dz16.data_ptr(), int(mask_z), dgamma.data_ptr(), dbeta.data_ptr(),
              dbias.data_ptr(), scratch.data_ptr(), rows, int(relu_ln), _lib.stream_ptr())
    return dres, dz16, dgamma, dbeta, dbias


class LinearLN16Fn(torch.autograd.Function):
    """bf16-stream form of LinearLNFn: (y, y') = LN(res + x W^T + b) with x, res, the GEMM output and y in bf16.  The two
    returned tensors are the same data: the caller hands one to the residual path and one to the next GEMM, so that their
    gradients come back separately and are summed inside the LayerNorm backward kernel."""

    @staticmethod
    def forward(ctx, x, res, weight, bias, gamma, beta, relu_ln):
        ctx.set_materialize_grads(False)
        Kd, N = x.shape[-1], weight.shape[0]
        x16 = to_bf16(x.reshape(-1, Kd))
        M = x16.shape[0]
        w16 = weight_bf16(weight)
        z16 = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
        gemm(x16, w16, z16, M, N, x16.shape[1], lda=x16.shape[1], ldb=w16.shape[1], bias=bias)
        res16 = to_bf16(res.reshape(M, N))
        y16, stats = _ln16_fwd(z16, res16, gamma, beta, relu_ln)
        ctx.relu_ln, ctx.xshape = relu_ln, x.shape
        ctx.save_for_backward(x16, w16, z16, res16, gamma, beta, stats)
        shp = (*x.shape[:-1], N)
        return y16.view(shp), y16.view(shp)

    @staticmethod
    def backward(ctx, dya, dyb):
        x16, w16, z16, res16, gamma, beta, stats = ctx.saved_tensors
        if dya is None and dyb is None:
            return (None,) * 7
        M, N = z16.shape
        K8 = w16.shape[1]
        dres, dz16, dgamma, dbeta, dbias = _ln16_bwd(dya, dyb, z16, res16, gamma, beta, stats, ctx.relu_ln, False,
                                                     ctx.needs_input_grad[1])
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K8), dtype=torch.bfloat16, device=z16.device)
            gemm(dz16, w16, dx, M, K8, N, lda=N, b_mn=True, ldb=K8)
            dx = dx.reshape(ctx.xshape)
        if ctx.needs_input_grad[2]:
            dw = _wgrad(dz16, x16, N, K8)
        return dx, (None if dres is None else dres.view(*ctx.xshape[:-1], N)), dw, dbias, dgamma, dbeta, None


class MLPLN16Fn(torch.autograd.Function):
    """bf16-stream form of MLPLNFn (FFN with ReLU on both layers + layernorm2, module_utils.py:121-123,136-139)."""

    @staticmethod
    def forward(ctx, x, res, w1, b1, w2, b2, gamma, beta, relu_ln):
        ctx.set_materialize_grads(False)
        Kd, Hd, N = x.shape[-1], w1.shape[0], w2.shape[0]
        x16 = to_bf16(x.reshape(-1, Kd))
        M = x16.shape[0]
        w1_16, w2_16 = weight_bf16(w1), weight_bf16(w2)
        h16 = torch.empty((M, Hd), dtype=torch.bfloat16, device=x.device)
        hbits = _relu_bits_gemm(x16, w1_16, h16, b1, M, Hd, Kd)
        z16 = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
        gemm(h16, w2_16, z16, M, N, Hd, lda=Hd, ldb=Hd, bias=b2, relu=True)
        res16 = to_bf16(res.reshape(M, N))
        y16, stats = _ln16_fwd(z16, res16, gamma, beta, relu_ln)
        ctx.relu_ln, ctx.xshape = relu_ln, x.shape
        ctx.save_for_backward(x16, w1_16, w2_16, h16, z16, res16, gamma, beta, stats, hbits)
        shp = (*x.shape[:-1], N)
        return y16.view(shp), y16.view(shp)

    @staticmethod
    def backward(ctx, dya, dyb):
        x16, w1_16, w2_16, h16, z16, res16, gamma, beta, stats, hbits = ctx.saved_tensors
        if dya is None and dyb is None:
            return (None,) * 9
        M, N = z16.shape
        Hd, Kd = w1_16.shape
        dev = z16.device
        dres, dz2, dgamma, dbeta, db2 = _ln16_bwd(dya, dyb, z16, res16, gamma, beta, stats, ctx.relu_ln, True,
                                                  ctx.needs_input_grad[1])
        dz1 = torch.empty((M, Hd), dtype=torch.bfloat16, device=dev)
        _lib.call('as_gemm_bf16_bitmasked', dz2.data_ptr(), w2_16.data_ptr(), dz1.data_ptr(), hbits.data_ptr(), M, Hd, N, 0, 1,
                  N, Hd, Hd, 1, 1.0, _lib.stream_ptr())
        db1 = column_sum(dz1) if ctx.needs_input_grad[3] else None
        dw2 = _wgrad(dz2, h16, N, Hd) if ctx.needs_input_grad[4] else None
        dw1 = _wgrad(dz1, x16, Hd, Kd) if ctx.needs_input_grad[2] else None
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, Kd), dtype=torch.bfloat16, device=dev)
            gemm(dz1, w1_16, dx, M, Kd, Hd, lda=Hd, b_mn=True, ldb=Kd)
            dx = dx.reshape(ctx.xshape)
        return (dx, (None if dres is None else dres.view(*ctx.xshape[:-1], N)), dw1, db1, dw2, db2, dgamma, dbeta, None)


class MaskedMean16Fn(torch.autograd.Function):
    """sum over valid entities / entity_num of a bf16 stream [R, N, 256] -> fp32 [R, 256]; bf16 gradient."""

    @staticmethod
    def forward(ctx, x, entity_num):
        R, N, Cc = x.shape
        x = x.contiguous()
        entity_num = entity_num.contiguous()
        out = torch.empty((R, Cc), dtype=torch.float32, device=x.device)
        _lib.call('as_masked_mean256_bf16_fwd', x.data_ptr(), entity_num.data_ptr(), out.data_ptr(), R, N, _lib.stream_ptr())
        ctx.dims = (R, N, Cc)
        ctx.save_for_backward(entity_num)
        return out

    @staticmethod
    def backward(ctx, dout):
        (entity_num,) = ctx.saved_tensors
        R, N, Cc = ctx.dims
        dout = dout.contiguous().float()
        dx = torch.empty((R, N, Cc), dtype=torch.bfloat16, device=dout.device)
        _lib.call('as_masked_mean256_bf16_bwd', dout.data_ptr(), entity_num.data_ptr(), dx.data_ptr(), R, N, _lib.stream_ptr())
        return dx, None


def fused_ln_ok(dim, *weights):
    return dim in (256, 384) and all(w.dim() == 2 and w.shape[1] % 8 == 0 and w.shape[0] % 32 == 0 for w in weights)


# dtype of the materialised attention scores S = QK^T/sqrt(d) and of dP = dO V^T.  fp32 keeps the reference's fp32
# scores; bf16 (what the reference's own autocast path stores) halves the HBM traffic of four passes per layer.
SCORE_DTYPE = torch.bfloat16 if os.environ.get('APPLESTAR_ATTN_BF16_SCORES', '1') == '1' else torch.float32


ATTN_FUSED = os.environ.get('APPLESTAR_ATTN_FUSED', '1') != '0'
ATTN_SMALL = os.environ.get('APPLESTAR_ATTN_SMALL', '1') != '0'


class AttentionFn(torch.autograd.Function):
    """softmax(Q K^T / sqrt(d), keys >= key_len masked) V over a packed bf16 qkv buffer [R, N, 3*H*D].
    Q K^T, P V and the four backward products are batched tcgen05 GEMMs reading Q/K/V (and dO) in place through
    strided / MN-major tensor maps; the softmax (forward and backward) is a warp-shuffle kernel."""

    @staticmethod
    def forward(ctx, qkv, key_len, heads, head_dim):
        R, N, _ = qkv.shape
        H, D = heads, head_dim
        ld = 3 * H * D
        qkv = qkv.contiguous()
        q, k, v = qkv[:, :, :H * D], qkv[:, :, H * D:2 * H * D], qkv[:, :, 2 * H * D:]
        scale = 1.0 / math.sqrt(D)
        kl = None if key_len is None else key_len.contiguous().long()
        if ATTN_FUSED and ATTN_SMALL and _lib.lib().as_attn_small_ok(N, D):
            # short sequence / narrow heads: one warp per (row, head), nothing saved but the inputs (attn_small.cu)
            O = torch.empty((R, N, H * D), dtype=torch.bfloat16, device=qkv.device)
            _lib.call('as_attn_small_fwd', qkv.data_ptr(), _ptr(kl), O.data_ptr(), R, N, H, D, scale, _lib.stream_ptr())
            ctx.dims = (R, N, N, H, D)
            ctx.fused = 'small'
            ctx.kl = kl
            ctx.save_for_backward(qkv)
            return O
        if ATTN_FUSED and _lib.lib().as_attn_fwd_ok(N, D):
            # scores stay in tensor memory (attn.cu); P is written only when a backward pass will read it
            need_p = ctx.needs_input_grad[0]
            P = torch.empty((R, H, N, N), dtype=torch.bfloat16, device=qkv.device) if need_p else None
            O = torch.empty((R, N, H * D), dtype=torch.bfloat16, device=qkv.device)
            _lib.call('as_attn_fwd', qkv.data_ptr(), _ptr(kl), O.data_ptr(), _ptr(P), R, N, H, D, scale, _lib.stream_ptr())
            ctx.dims = (R, N, N, H, D)
            ctx.fused = True
            if need_p:
                ctx.save_for_backward(qkv, P, O)
            return O
        Np = (N + 7) // 8 * 8                       # pitch of the score rows (bf16 P must be 16-byte pitched)
        S = torch.empty((R, H, N, Np), dtype=SCORE_DTYPE, device=qkv.device)
        gemm(q, k, S, N, N, D, lda=ld, ldb=ld, ldc=Np, batch=(H, R), a_strides=(D, N * ld), b_strides=(D, N * ld),
             c_strides=(N * Np, H * N * Np), alpha=scale)
        P = torch.empty((R, H, N, Np), dtype=torch.bfloat16, device=qkv.device)
        if Np != N:
            P.zero_()
        _softmax_rows(S, kl, P, R * H, N, N, Np, H)
        O = torch.empty((R, N, H * D), dtype=torch.bfloat16, device=qkv.device)
        gemm(P, v, O, N, D, N, lda=Np, b_mn=True, ldb=ld, ldc=H * D, batch=(H, R),
             a_strides=(N * Np, H * N * Np), b_strides=(D, N * ld), c_strides=(D, N * H * D))
        ctx.dims = (R, N, Np, H, D)
        ctx.fused = False
        ctx.save_for_backward(qkv, P)
        return O

    @staticmethod
    def backward(ctx, dO):
        R, N, Np, H, D = ctx.dims
        if ctx.fused == 'small':
            qkv, = ctx.saved_tensors
            dO = to_bf16(dO.contiguous().reshape(R * N, H * D))
            dqkv = torch.empty_like(qkv)
            _lib.call('as_attn_small_bwd', qkv.data_ptr(), _ptr(ctx.kl), dO.data_ptr(), dqkv.data_ptr(), R, N, H, D,
                      1.0 / math.sqrt(D), _lib.stream_ptr())
            return dqkv, None, None, None
        qkv, P = ctx.saved_tensors[:2]
        ld = 3 * H * D
        scale = 1.0 / math.sqrt(D)
        dO = to_bf16(dO.contiguous().reshape(R * N, H * D)).view(R, N, H * D)
        q, k, v = qkv[:, :, :H * D], qkv[:, :, H * D:2 * H * D], qkv[:, :, 2 * H * D:]
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = dqkv[:, :, :H * D], dqkv[:, :, H * D:2 * H * D], dqkv[:, :, 2 * H * D:]
        sP, sO = (N * Np, H * N * Np), (D, N * H * D)
        sQ = (D, N * ld)
        if ctx.fused:
            # dP stays in tensor memory: one launch gives dS and dQ; dV and dK are the two query-summed products
            O = ctx.saved_tensors[2]
            dS = torch.empty((R, H, N, N), dtype=torch.bfloat16, device=qkv.device)
            _lib.call('as_attn_bwd_dq', qkv.data_ptr(), O.data_ptr(), dO.data_ptr(), P.data_ptr(), dS.data_ptr(),
                      dqkv.data_ptr(), R, N, H, D, scale, _lib.stream_ptr())
            gemm(P, dO, dv, N, D, N, a_mn=True, lda=Np, b_mn=True, ldb=H * D, ldc=ld, batch=(H, R), a_strides=sP,
                 b_strides=sO, c_strides=sQ)
            gemm(dS, q, dk, N, D, N, a_mn=True, lda=Np, b_mn=True, ldb=ld, ldc=ld, batch=(H, R), a_strides=sP,
                 b_strides=sQ, c_strides=sQ, alpha=scale)
            return dqkv, None, None, None
        # dP = dO V^T
        dP = torch.empty((R, H, N, Np), dtype=SCORE_DTYPE, device=qkv.device)
        gemm(dO, v, dP, N, N, D, lda=H * D, ldb=ld, ldc=Np, batch=(H, R), a_strides=sO, b_strides=sQ, c_strides=sP)
        # dV = P^T dO   (both operands MN-major)
        gemm(P, dO, dv, N, D, N, a_mn=True, lda=Np, b_mn=True, ldb=H * D, ldc=ld, batch=(H, R), a_strides=sP,
             b_strides=sO, c_strides=sQ)
        dS = torch.empty((R, H, N, Np), dtype=torch.bfloat16, device=qkv.device)
        if Np != N:
            dS.zero_()
        _softmax_bwd_rows(P, dP, dS, R * H * N, N, Np)
        # dQ = scale * dS K ; dK = scale * dS^T Q
        gemm(dS, k, dq, N, D, N, lda=Np, b_mn=True, ldb=ld, ldc=ld, batch=(H, R), a_strides=sP, b_strides=sQ,
             c_strides=sQ, alpha=scale)
        gemm(dS, q, dk, N, D, N, a_mn=True, lda=Np, b_mn=True, ldb=ld, ldc=ld, batch=(H, R), a_strides=sP,
             b_strides=sQ, c_strides=sQ, alpha=scale)
        return dqkv, None, None, None


def _softmax_rows(S, key_len, P, groups, Nq, Nk, pitch, groups_per_len):
    if pitch == Nk:
        _lib.call('as_softmax_fwd', S.data_ptr(), _ptr(key_len), P.data_ptr(), groups, Nq, Nk, groups_per_len,
                  int(S.dtype == torch.bfloat16), _lib.stream_ptr())
    else:   # padded pitch: treat the pad columns as masked keys
        kl = key_len if key_len is not None else torch.full((groups // groups_per_len,), Nk, dtype=torch.int64,
                                                            device=S.device)
        _lib.call('as_softmax_fwd', S.data_ptr(), kl.data_ptr(), P.data_ptr(), groups, Nq, pitch, groups_per_len,
                  int(S.dtype == torch.bfloat16), _lib.stream_ptr())


def _softmax_bwd_rows(P, dP, dS, rows, Nk, pitch):
    _lib.call('as_softmax_bwd', P.data_ptr(), dP.data_ptr(), dS.data_ptr(), rows, pitch, int(dP.dtype == torch.bfloat16),
              _lib.stream_ptr())


def attention(qkv, key_len, heads, head_dim):
    if qkv.dtype != torch.bfloat16:
        qkv = to_bf16(qkv.reshape(-1, qkv.shape[-1])).view(qkv.shape)
    return AttentionFn.apply(qkv, key_len, heads, head_dim)


_KIND = {'one_hot': 0, 'binary': 1, 'unsqueeze': 2}


def entity_features(entity_info, fields, pitch):
    """bf16 [R*N, pitch] feature rows (997 real columns) straight from the narrow per-field arrays."""
    import ctypes
    n = len(fields)
    first = entity_info[fields[0][0]]
    rows = first.numel()
    keep, col = [], 0
    data = (ctypes.c_void_p * n)()
    kind, num, cols, dts = (ctypes.c_int * n)(), (ctypes.c_int * n)(), (ctypes.c_int * n)(), (ctypes.c_int * n)()
    for i, (name, _, arc, k) in enumerate(fields):
        t = entity_info[name].contiguous()
        if t.dtype not in _DT_CODE:
            t = t.float()
        keep.append(t)
        data[i], kind[i], num[i], cols[i], dts[i] = t.data_ptr(), _KIND[arc], k, col, _DT_CODE[t.dtype]
        col += k
    out = torch.empty((rows, pitch), dtype=torch.bfloat16, device=first.device)
    neg = torch.zeros(1, dtype=torch.int32, device=first.device)
    _lib.call('as_entity_features', data, kind, num, cols, dts, n, out.data_ptr(), rows, pitch, neg.data_ptr(),
              _lib.stream_ptr())
    return out, neg


def entity_embed(entity_info, weight, bias, layout, bf16_only=False):
    """ReLU(features @ W^T + b) -> (fp32, bf16) [R, N, 256].  The one-hot concat exists only as a bf16 GEMM operand."""
    from . import config as C
    first = entity_info['x']
    R, N = first.shape
    feats, neg = entity_features(entity_info, C.ENTITY_FIELDS, 1000)
    entity_embed.last_neg_flag = neg          # checked lazily by the caller (no host sync on the hot path)
    if bf16_only:
        return linear(feats, weight, bias, relu=True, out_bf16=True).view(R, N, -1)
    y32, y16 = linear_dual(feats, weight, bias, relu=True)
    return y32.view(R, N, -1), y16.view(R, N, -1)


class LNLSTMLayerFn(torch.autograd.Function):
    """One LayerNorm-LSTM layer over the whole unroll in a single persistent kernel (one CTA per trajectory)."""

    @staticmethod
    def forward(ctx, gi, h0, c0, whh, gh_w, gh_b, gc_w, gc_b):
        ctx.set_materialize_grads(False)
        T, B, G = gi.shape
        H = G // 4
        gi, h0, c0 = gi.contiguous(), h0.contiguous().float(), c0.contiguous().float()
        whh16 = weight_bf16(whh)
        dev = gi.device
        y = torch.empty((T, B, H), dtype=torch.float32, device=dev)
        c_all = torch.empty((T, B, H), dtype=torch.float32, device=dev)
        u_all = torch.empty((T, B, G), dtype=torch.float32, device=dev)
        z_all = torch.empty((T, B, H), dtype=torch.float32, device=dev)
        stats = torch.empty((T, B, 4), dtype=torch.float32, device=dev)
        _lib.call('as_lnlstm_fwd', gi.data_ptr(), h0.data_ptr(), c0.data_ptr(), whh16.data_ptr(), gh_w.data_ptr(),
                  gh_b.data_ptr(), gc_w.data_ptr(), gc_b.data_ptr(), y.data_ptr(), c_all.data_ptr(), u_all.data_ptr(),
                  z_all.data_ptr(), stats.data_ptr(), T, B, H, _lib.stream_ptr())
        ctx.dims = (T, B, H)
        ctx.save_for_backward(gi, h0, c0, y, c_all, u_all, z_all, stats, whh, gh_w, gh_b, gc_w, gc_b)
        return y, y[-1].clone(), c_all[-1].clone()

    @staticmethod
    def backward(ctx, dy, dhT, dcT):
        gi, h0, c0, y, c_all, u_all, z_all, stats, whh, gh_w, gh_b, gc_w, gc_b = ctx.saved_tensors
        T, B, H = ctx.dims
        G = 4 * H
        dev = gi.device
        tag = (_SHADOW_GEN[0], whh._version, whh.data_ptr())
        hit = getattr(whh, '_as_wt16', None)
        if hit is None or hit[0] != tag:
            hit = (tag, whh.detach().t().contiguous().to(torch.bfloat16))
            try:
                whh._as_wt16 = hit
            except AttributeError:
                pass
        whh_t16 = hit[1]
        dy = None if dy is None else dy.contiguous().float()
        dhT = None if dhT is None else dhT.contiguous().float()
        dcT = None if dcT is None else dcT.contiguous().float()
        dgi = torch.empty_like(gi)
        du = torch.empty_like(gi)
        dh0 = torch.empty_like(h0)
        dc0 = torch.empty_like(c0)
        dparam = torch.empty((B, 10 * H), dtype=torch.float32, device=dev)
        _lib.call('as_lnlstm_bwd', gi.data_ptr(), h0.data_ptr(), c0.data_ptr(), y.data_ptr(), c_all.data_ptr(),
                  u_all.data_ptr(), z_all.data_ptr(), stats.data_ptr(), whh_t16.data_ptr(), gh_w.data_ptr(),
                  gh_b.data_ptr(), gc_w.data_ptr(), gc_b.data_ptr(), _ptr(dy), _ptr(dhT), _ptr(dcT), dgi.data_ptr(),
                  du.data_ptr(), dh0.data_ptr(), dc0.data_ptr(), dparam.data_ptr(), T, B, H, _lib.stream_ptr())
        dp = dparam.sum(dim=0)
        hprev = torch.cat([h0.unsqueeze(0), y[:-1]], dim=0).view(T * B, H)
        dwhh = _wgrad(to_bf16(du.view(T * B, G)), to_bf16(hprev), G, H)
        return (dgi, dh0, dc0, dwhh, dp[:G], dp[G:2 * G], dp[2 * G:2 * G + H], dp[2 * G + H:])


def lnlstm_layer(x, h, c, cell):
    """x [T,B,I] -> (y [T,B,H], h_T, c_T).  Input projection for all timesteps = one GEMM + one LayerNorm launch."""
    li, lh, lc = cell.layernorm_i, cell.layernorm_h, cell.layernorm_c
    gi = layer_norm(linear(x, cell.weight_ih, None), li.weight, li.bias)
    return LNLSTMLayerFn.apply(gi, h, c, cell.weight_hh, lh.weight, lh.bias, lc.weight, lc.bias)
