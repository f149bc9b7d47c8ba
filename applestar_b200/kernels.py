"""torch.autograd bindings of the C-ABI kernels (applestar_b200/_lib.py).

Tensors stay torch tensors only so that torch owns device memory, streams and the autograd tape; every
function here hands raw `data_ptr()`s to `libapplestar_b200.so` on `torch.cuda.current_stream()`.
"""
import math

import torch
import torch.nn.functional as F

from . import _lib

NUM_SMS = 148


def _ptr(t):
    return 0 if t is None else t.data_ptr()


# ---------------------------------------------------------------------------------------------
# bf16 staging
def to_bf16(x):
    """fp32 [rows, K] -> bf16 [rows, K8] (K padded to a multiple of 8 with zeros so TMA pitches are 16-byte)."""
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
    view of the flat bf16 shadow the fused Adam kernel keeps fresh; otherwise the copy is cached on the Parameter
    object itself (never keyed by address: freed addresses get reused by other models)."""
    sh = getattr(w, '_as_shadow', None)
    if sh is not None and w.shape[-1] % 8 == 0 and w.dim() == 2:
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
         alpha=1.0, batch=(1, 1), a_strides=(0, 0), b_strides=(0, 0), c_strides=(0, 0), splits=1, ws=None):
    """Raw tcgen05 GEMM launch: C[b][m][n] = act(alpha * sum_k A[b][m][k] B[b][n][k] + bias[n])."""
    lda = lda if lda is not None else (M if a_mn else K)
    ldb = ldb if ldb is not None else (N if b_mn else K)
    ldc = ldc if ldc is not None else N
    _lib.call('as_gemm_bf16', A.data_ptr(), B.data_ptr(), _ptr(C), _ptr(bias), M, N, K, int(a_mn), int(b_mn),
              lda, ldb, ldc, batch[0], batch[1], a_strides[0], a_strides[1], b_strides[0], b_strides[1],
              c_strides[0], c_strides[1], int(C is not None and C.dtype == torch.bfloat16), int(relu), float(alpha),
              splits, _ptr(ws), _lib.stream_ptr())


def _wgrad(dy16, x16, n_out, k_in):
    """dW[n_out, k_in] = dY^T X with both operands read MN-major (no transposes); split-K over the rows."""
    M = dy16.shape[0]
    tiles = ((n_out + 127) // 128) * ((k_in + 255) // 256 if k_in > 128 else 1)
    want = max(1, min((M + 63) // 64, (2 * NUM_SMS) // max(tiles, 1)))
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
    """y = act(x W^T + b): forward, dgrad and wgrad all on the tcgen05 GEMM."""

    @staticmethod
    def forward(ctx, x, weight, bias, relu):
        K = x.shape[-1]
        N = weight.shape[0]
        x2 = x.reshape(-1, K)
        M = x2.shape[0]
        x16 = to_bf16(x2)
        w16 = weight_bf16(weight)
        y = torch.empty((M, N), dtype=torch.float32, device=x.device)
        gemm(x16, w16, y, M, N, x16.shape[1], lda=x16.shape[1], ldb=w16.shape[1], bias=bias, relu=relu)
        ctx.relu = relu
        ctx.k = K
        ctx.xshape = x.shape
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x16, w16, y if relu else None)
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x16, w16, y = ctx.saved_tensors
        N = w16.shape[0]
        K8 = w16.shape[1]
        dy2 = dy.reshape(-1, N)
        if ctx.relu:
            dy2 = dy2 * (y > 0)
        M = dy2.shape[0]
        dy16 = to_bf16(dy2)                      # [M, N8]
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K8), dtype=torch.float32, device=dy.device)
            # dX[m,k] = sum_n dY[m,n] W[n,k]: A = dY (K-major over n), B = W read MN-major (k contiguous)
            gemm(dy16, w16, dx, M, K8, N, lda=dy16.shape[1], b_mn=True, ldb=K8)
            if K8 != ctx.k:
                dx = dx[:, :ctx.k]
            dx = dx.reshape(ctx.xshape)
        if ctx.needs_input_grad[1]:
            dw = _wgrad(dy16, x16, N, K8)
            if K8 != ctx.k:
                dw = dw[:, :ctx.k]
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy2.sum(dim=0)
        return dx, dw, db, None


def linear(x, weight, bias, relu=False):
    return LinearFn.apply(x, weight, bias, relu)


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
# [library] pieces still on ATen (replacement queue: DESIGN.md section 6)
def layer_norm(x, weight, bias, residual=None, relu=False):
    if residual is not None:
        x = x + residual
    y = F.layer_norm(x, (x.shape[-1],), weight, bias, 1e-5)
    return F.relu(y) if relu else y


def attention(qkv, key_len, heads, head_dim):
    R, N, _ = qkv.shape
    q, k, v = qkv.chunk(3, dim=2)
    q = q.view(R, N, heads, head_dim).permute(0, 2, 1, 3)
    k = k.view(R, N, heads, head_dim).permute(0, 2, 3, 1)
    v = v.view(R, N, heads, head_dim).permute(0, 2, 1, 3)
    s = torch.matmul(q, k) / math.sqrt(head_dim)
    if key_len is not None:
        mask = torch.arange(N, device=qkv.device).view(1, 1, 1, N) < key_len.view(R, 1, 1, 1)
        s = s.masked_fill(~mask, -1e9)
    o = torch.matmul(torch.softmax(s, dim=-1), v)
    return o.permute(0, 2, 1, 3).reshape(R, N, heads * head_dim)


def entity_embed(entity_info, weight, bias, layout):
    Wt = weight.t()                                               # [997, 256]
    idx = torch.stack([entity_info[n].long() for n, _, _ in layout['one_hot']], dim=-1)
    if (idx < 0).any():
        raise RuntimeError('negative one-hot index in entity_info')          # entity_encoder.py:68-71
    idx = torch.minimum(idx, layout['oh_max']) + layout['oh_off']
    R, N, Fh = idx.shape
    y = F.embedding_bag(idx.view(R * N, Fh), Wt, mode='sum')
    bits = []
    for n, _, nb in layout['binary']:
        shifts = torch.arange(nb - 1, -1, -1, device=idx.device)
        bits.append(((entity_info[n].long().unsqueeze(-1) >> shifts) & 1).float())
    y = y + torch.cat(bits, dim=-1).view(R * N, -1) @ Wt[layout['bin_cols']]
    sc = torch.stack([entity_info[n].float() for n, _ in layout['scalar']], dim=-1)
    y = y + sc.view(R * N, -1) @ Wt[layout['sc_cols']]
    return F.relu(y + bias).view(R, N, -1)


def lnlstm_layer(x, h, c, cell):
    T = x.shape[0]
    wi, wh = cell.weight_ih, cell.weight_hh
    li, lh, lc = cell.layernorm_i, cell.layernorm_h, cell.layernorm_c
    Hs = h.shape[-1]
    gi_all = F.layer_norm(linear(x, wi, None), (4 * Hs,), li.weight, li.bias, 1e-5)   # all timesteps at once
    ys = []
    for t in range(T):
        gh = F.layer_norm(h @ wh.t(), (4 * Hs,), lh.weight, lh.bias, 1e-5)
        i, f, g, o = (gi_all[t] + gh).chunk(4, 1)
        c = F.layer_norm(torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g), (Hs,), lc.weight, lc.bias, 1e-5)
        h = torch.sigmoid(o) * torch.tanh(c)
        ys.append(h)
    return torch.stack(ys), h, c
