"""Operator layer of the hot path.

Every function here runs on the GPU only.  Ops marked [kernel] launch this repo's sm_100a kernels
through the C-ABI (applestar_b200/_lib.py -> libapplestar_b200.so); the few small element-wise / gather
ops that still go through ATen (scalar-encoder embeddings, gates, glue) are listed in DESIGN.md §7.  Nothing
falls back to the CPU: `require_cuda` raises, and `_lib` raises if the shared library is missing.
"""
import math
import os

import torch
import torch.nn.functional as F

from . import config as C
from . import _lib
from . import kernels as K


def require_cuda(t):
    if not t.is_cuda:
        raise RuntimeError('applestar_b200 runs on CUDA tensors only (no CPU / MPS fallback); got device %s' % t.device)
    _lib.lib()      # raises loudly when the C-ABI library is missing


# ---------------------------------------------------------------------------------------------
# dense layers
def linear(x, weight, bias, relu=False, out_bf16=False):
    """[kernel] y = act(x @ W^T + b) on the tcgen05 GEMM (fp32 or bf16 output)."""
    require_cuda(x)
    return K.linear(x, weight, bias, relu, out_bf16)


def layer_norm(x, weight, bias, residual=None, relu=False):
    """[kernel] LN(x (+ residual)) (+ReLU), eps 1e-5, fp32 statistics."""
    return K.layer_norm(x, weight, bias, residual, relu)


def layer_norm_dual(x, weight, bias, residual=None, relu=False):
    """[kernel] same, returning (fp32, bf16) copies."""
    return K.layer_norm_dual(x, weight, bias, residual, relu)


def transformer_layer_fused_ok(layer):
    pj, f1, f2 = layer.attention.project[0], layer.mlp[0][0], layer.mlp[1][0]
    return (os.environ.get('APPLESTAR_FUSED_LAYER', '1') != '0'
            and K.fused_ln_ok(pj.weight.shape[0], pj.weight, f1.weight, f2.weight))


def linear_ln(x, res, weight, bias, gamma, beta, relu_ln=False):
    """[kernel] (fp32, bf16) = LN(res + x W^T + b) as one autograd node (kernels.LinearLNFn)."""
    require_cuda(x)
    return K.LinearLNFn.apply(x, res, weight, bias, gamma, beta, relu_ln)


def mlp_ln(x, res, w1, b1, w2, b2, gamma, beta, relu_ln=False):
    """[kernel] (fp32, bf16) = LN(res + ReLU(ReLU(x W1^T + b1) W2^T + b2)) as one autograd node (kernels.MLPLNFn)."""
    require_cuda(x)
    return K.MLPLNFn.apply(x, res, w1, b1, w2, b2, gamma, beta, relu_ln)


# Residual stream of the entity Transformer in bf16 (LayerNorm traffic 0.4 - 0.6x, measured -1.07 ms of 60.2 ms per learner
# step).  OFF by default: at N = 512 the target-unit head's gradient aggregate rose from 0.087 to 0.169 relative L2 against
# the reference (budget 0.128 = 2 x the reference's own bf16-autocast error), so the fp32 stream stays the product path.
ENTITY_BF16 = os.environ.get('APPLESTAR_ENTITY_BF16', '0') == '1'


def linear_ln16(x, res, weight, bias, gamma, beta, relu_ln=False):
    """[kernel] bf16-stream LN(res + x W^T + b) -> two handles of the same bf16 tensor (kernels.LinearLN16Fn)."""
    require_cuda(x)
    return K.LinearLN16Fn.apply(x, res, weight, bias, gamma, beta, relu_ln)


def mlp_ln16(x, res, w1, b1, w2, b2, gamma, beta, relu_ln=False):
    """[kernel] bf16-stream LN(res + FFN(x)) -> two handles of the same bf16 tensor (kernels.MLPLN16Fn)."""
    require_cuda(x)
    return K.MLPLN16Fn.apply(x, res, w1, b1, w2, b2, gamma, beta, relu_ln)


def masked_mean16(x, entity_num):
    """[kernel] masked mean of a bf16 [R, N, 256] stream (entity_encoder.py:83-85)."""
    require_cuda(x)
    return K.MaskedMean16Fn.apply(x, entity_num.long())


def add_relu(a, b):
    return F.relu(a + b)


def sigmoid_gate(gate_pre, x):
    return torch.sigmoid(gate_pre) * x


def atan_squash(v):
    return (2.0 / math.pi) * torch.atan((math.pi / 2.0) * v)


# ---------------------------------------------------------------------------------------------
# entity path
_ENTITY_LAYOUT = None


def _entity_layout(device):
    """Column offsets of the 997-wide feature vector (entity_encoder.py:60-78)."""
    global _ENTITY_LAYOUT
    if _ENTITY_LAYOUT is None or _ENTITY_LAYOUT['dev'] != device:
        off = 0
        one_hot, binary, scalar = [], [], []
        for name, _, arc, num in C.ENTITY_FIELDS:
            if arc == 'one_hot':
                one_hot.append((name, off, num))
            elif arc == 'binary':
                binary.append((name, off, num))
            else:
                scalar.append((name, off))
            off += num
        _ENTITY_LAYOUT = {
            'dev': device, 'one_hot': one_hot, 'binary': binary, 'scalar': scalar,
            'oh_off': torch.tensor([o for _, o, _ in one_hot], device=device),
            'oh_max': torch.tensor([n - 1 for _, _, n in one_hot], device=device),
            'bin_cols': torch.tensor([o + i for _, o, n in binary for i in range(n)], device=device),
            'sc_cols': torch.tensor([o for _, o in scalar], device=device),
        }
    return _ENTITY_LAYOUT


def entity_embed(entity_info, weight, bias, bf16_only=False):
    """[kernel] ReLU(onehot_concat(entity_info) @ W^T + b) -> (fp32, bf16), or the bf16 copy alone.  The 997-wide concat
    exists only as a bf16 GEMM operand written by one kernel from the narrow per-field arrays (never as 2 MB/obs of fp32)."""
    x0 = entity_info['x']
    require_cuda(x0)
    return K.entity_embed(entity_info, weight, bias, _entity_layout(x0.device), bf16_only)


def raise_if_negative_index():
    """Deferred form of the reference's RuntimeError on a negative one-hot index (entity_encoder.py:68-71): the flag
    of the previous call is inspected at the next entry point, so the hot path never waits on the host."""
    flag = getattr(K.entity_embed, 'last_neg_flag', None)
    if flag is not None:
        K.entity_embed.last_neg_flag = None
        if int(flag.item()) != 0:
            raise RuntimeError('negative one-hot index in entity_info')


class _MaskedMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, entity_num):
        R, N, Cc = x.shape
        x = x.contiguous()
        entity_num = entity_num.contiguous()
        out = torch.empty((R, Cc), dtype=torch.float32, device=x.device)
        _lib.call('as_masked_mean_fwd', x.data_ptr(), entity_num.data_ptr(), out.data_ptr(), R, N, Cc, _lib.stream_ptr())
        ctx.dims = (R, N, Cc)
        ctx.save_for_backward(entity_num)
        return out

    @staticmethod
    def backward(ctx, dout):
        (entity_num,) = ctx.saved_tensors
        R, N, Cc = ctx.dims
        dout = dout.contiguous()
        dx = torch.empty((R, N, Cc), dtype=torch.float32, device=dout.device)
        _lib.call('as_masked_mean_bwd', dout.data_ptr(), entity_num.data_ptr(), dx.data_ptr(), R, N, Cc, _lib.stream_ptr())
        return dx, None


def masked_mean(x, entity_num):
    """[kernel] sum over valid entities / entity_num (entity_encoder.py:83-85)."""
    require_cuda(x)
    if x.dtype != torch.float32 or x.shape[-1] % 4 or x.shape[-1] > 1024:
        raise RuntimeError('masked_mean: fp32 rows with C % 4 == 0 expected')
    return _MaskedMean.apply(x, entity_num.long())


def attention(qkv, key_len, heads, head_dim):
    """softmax(Q K^T / sqrt(d), keys >= key_len masked) V over packed qkv [R,N,3*heads*head_dim]."""
    return K.attention(qkv, key_len, heads, head_dim)


def scatter_connection(proj, x, y, entity_num, H, W, scatter_type='add'):
    """[kernel] ScatterConnection (module_utils.py:11-34): sort-then-segment, atomic-free, ascending
    entity order per cell.  proj [R,N,D] fp32, x/y [R,N] integer -> [R,D,H,W] fp32."""
    require_cuda(proj)
    return K.ScatterConnection.apply(proj, x, y, entity_num, H, W, scatter_type)


# ---------------------------------------------------------------------------------------------
# scalar path
def _index_tensor(t):
    t = t.contiguous()
    if t.dtype not in K._DT_CODE:
        t = t.long()
    return t


def one_hot_bf16(idx, V):
    """[kernel] bf16 one-hot rows [R, V8] (index clamped to V - 1): a GEMM operand."""
    idx = _index_tensor(idx).reshape(-1)
    v8 = (V + 7) // 8 * 8
    out = torch.empty((idx.numel(), v8), dtype=torch.bfloat16, device=idx.device)
    _lib.call('as_one_hot_bf16', idx.data_ptr(), K._DT_CODE[idx.dtype], out.data_ptr(), idx.numel(), V, v8, _lib.stream_ptr())
    return out


class _EmbedRelu(torch.autograd.Function):
    """[kernel] relu(table[min(idx, V-1)]) (scalar_encoder.py:112-116): exact fp32 gather forward; the table gradient is the
    tensor-core product (masked dy)^T . one_hot (split-K, deterministic) instead of an index_put."""

    @staticmethod
    def forward(ctx, idx, table):
        idx = _index_tensor(idx)
        table = table.contiguous()
        R = idx.numel()
        V, D = table.shape
        out = torch.empty((R, D), dtype=torch.float32, device=table.device)
        _lib.call('as_embed_relu_fwd', idx.data_ptr(), K._DT_CODE[idx.dtype], table.data_ptr(), out.data_ptr(), R, V, D,
                  _lib.stream_ptr())
        ctx.save_for_backward(idx, out)
        ctx.vd = (V, D)
        return out

    @staticmethod
    def backward(ctx, dy):
        idx, out = ctx.saved_tensors
        V, D = ctx.vd
        dy = dy.contiguous().float()
        R = idx.numel()
        dz16 = torch.empty((R, D), dtype=torch.bfloat16, device=dy.device)
        _lib.call('as_linear_bwd_prep2', dy.data_ptr(), 0, out.data_ptr(), 0, dz16.data_ptr(), 0, 0, R, D, _lib.stream_ptr())
        dt = K._wgrad(dz16, one_hot_bf16(idx, V), D, (V + 7) // 8 * 8)           # [D, V8]
        return None, dt.t()[:V].contiguous()


def embedding_relu(idx, table):
    require_cuda(table)
    return _EmbedRelu.apply(idx, table)


def embedding(idx, table):
    """[kernel] plain nn.Embedding(idx) (value_encoder.py:53-55) as a tensor-core product one_hot . table, so that the
    forward and the table gradient are ordinary GEMMs -> [idx.numel(), D]."""
    require_cuda(table)
    return K.linear(one_hot_bf16(idx, table.shape[0]), table.t(), None, False)


def time_encode(t, position_array):
    """[kernel] sinusoidal time code (scalar_encoder.py:91-97)."""
    require_cuda(position_array)
    t = _index_tensor(t) if t.dtype != torch.float32 else t.contiguous()
    out = torch.empty((t.numel(), position_array.shape[0]), dtype=torch.float32, device=position_array.device)
    _lib.call('as_time_encode', t.data_ptr(), K._DT_CODE[t.dtype], position_array.detach().contiguous().data_ptr(),
              out.data_ptr(), t.numel(), position_array.shape[0], _lib.stream_ptr())
    return out


def build_order_embed(order, location, weight, bias):
    """[kernel] ReLU(W . [one_hot(order,174) | eye(20) | bits10(x) | bits10(y)] + b) (scalar_encoder.py:41-50 + the
    transformer's embedding layer): the token features are written once as a bf16 operand (as_bo_features) and the
    product, its ReLU and both gradients run on the tcgen05 GEMM like every other linear layer."""
    require_cuda(weight)
    R, L = order.shape
    order, location = _index_tensor(order), _index_tensor(location)
    Kin = weight.shape[1]
    A = C.NUM_BEGINNING_ORDER_ACTIONS
    NB = (Kin - A - L) // 2
    k8 = (Kin + 7) // 8 * 8
    feats = torch.empty((R * L, k8), dtype=torch.bfloat16, device=weight.device)
    _lib.call('as_bo_features', order.data_ptr(), K._DT_CODE[order.dtype], location.data_ptr(), K._DT_CODE[location.dtype],
              feats.data_ptr(), R * L, L, A, NB, C.SPATIAL_X, k8, _lib.stream_ptr())
    return K.linear(feats, weight, bias, relu=True).view(R, L, -1)


# ---------------------------------------------------------------------------------------------
# recurrent core
def lnlstm_layer(x, h, c, cell):
    """One LayerNorm-LSTM layer over time (lstm.py:138-167).  x [T,B,I] -> ([T,B,H], h_T, c_T)."""
    return K.lnlstm_layer(x, h, c, cell)


# ---------------------------------------------------------------------------------------------
# heads
def gumbel_argmax(logits, noise):
    """[kernel] argmax_k(fp32 logit_k + g_k), lowest index on ties."""
    return K.gumbel_argmax(logits, noise)


def fresh_gumbel(rows, n_entity, device):
    def gum(*shape):
        u = torch.rand(*shape, device=device).clamp_(1e-10, 1.0 - 1e-7)
        return -torch.log(-torch.log(u))
    return {'action_type': gum(rows, C.NUM_ACTIONS), 'delay': gum(rows, 128), 'queued': gum(rows, 2),
            'selected_units': gum(C.MAX_SELECTED_UNITS_NUM, rows, n_entity + 1),
            'target_unit': gum(rows, n_entity), 'target_location': gum(rows, C.SPATIAL_Y * C.SPATIAL_X)}


def one_hot_linear(idx, weight, bias, relu=False):
    """[kernel] fc over a one-hot input (head/action_type_head.py:60-61, action_arg_head.py:48-50, 81-83): the one-hot rows
    are written as a bf16 operand and the layer is an ordinary tcgen05 linear (forward, data-free backward: dW, db)."""
    require_cuda(weight)
    shape = idx.shape
    y = K.linear(one_hot_bf16(idx, weight.shape[1]), weight, bias, relu)
    return y.view(*shape, -1)


def categorical_log_prob(logits, action):
    return torch.log_softmax(logits, dim=-1).gather(-1, action.unsqueeze(-1)).squeeze(-1)


class _PrefixFork(torch.autograd.Function):
    """x -> (alias of x, x[:n]).  The learner runs the encoder on (T + 1) * B rows but the policy heads only on the first
    T * B (model.py:128-137 slices `[:-batch_size]`); autograd would zero-fill a full-size buffer for the slice's gradient,
    copy the heads' gradient into it and add it to the gradient of the tensor's other consumers in a third pass.  With
    both consumers behind this node the sum is one pass: out[:n] = g_full[:n] + g_head (as_add_act_bf16), out[n:] = g_full[n:]."""

    @staticmethod
    def forward(ctx, x, n):
        ctx.set_materialize_grads(False)
        ctx.n = n
        ctx.rows = x.shape[0]
        return x.view_as(x), x.narrow(0, 0, n)

    @staticmethod
    def backward(ctx, g_full, g_head):
        n = ctx.n
        if g_head is None:
            return g_full, None
        g_head = g_head.contiguous()
        if g_full is None:
            out = g_head.new_zeros((ctx.rows,) + tuple(g_head.shape[1:])) if hasattr(ctx, 'rows') else None
            if out is None:
                raise RuntimeError('prefix_fork: the full-size alias received no gradient')
            out[:n].copy_(g_head)
            return out, None
        g_full = g_full.contiguous()
        out = torch.empty_like(g_full)
        numel = g_head.numel()
        if g_full.dtype == torch.bfloat16 and g_head.dtype == torch.bfloat16 and numel % 8 == 0 and g_full.is_cuda:
            _lib.call('as_add_act_bf16', g_full.data_ptr(), g_head.data_ptr(), out.data_ptr(), numel, 0, _lib.stream_ptr())
        else:
            torch.add(g_full[:n], g_head.to(g_full.dtype), out=out[:n])
        if out.shape[0] > n:
            out[n:].copy_(g_full[n:])
        return out, None


def prefix_fork(x, n):
    """(x for the consumers that take every row, x[:n] for the policy heads); see _PrefixFork."""
    if n is None or n >= x.shape[0] or not x.requires_grad:
        return x, (x if n is None else x[:n])
    return _PrefixFork.apply(x, int(n))


class _PointerKeys(torch.autograd.Function):
    """[kernel] as_pointer_keys_fwd / _bwd"""

    @staticmethod
    def forward(ctx, key, end_embedding, entity_num):
        R, N, D = key.shape
        key = key.contiguous().float()
        end = end_embedding.reshape(-1).contiguous().float()
        en = entity_num.contiguous().long()
        out = torch.empty((R, N + 1, D), dtype=torch.float32, device=key.device)
        _lib.call('as_pointer_keys_fwd', key.data_ptr(), end.data_ptr(), en.data_ptr(), out.data_ptr(), R, N, D, _lib.stream_ptr())
        ctx.save_for_backward(en)
        ctx.dims = (R, N, D, end_embedding.shape)
        return out

    @staticmethod
    def backward(ctx, dout):
        (en,) = ctx.saved_tensors
        R, N, D, eshape = ctx.dims
        dout = dout.contiguous().float()
        dkey = torch.empty((R, N, D), dtype=torch.float32, device=dout.device) if ctx.needs_input_grad[0] else None
        dend = torch.empty(D, dtype=torch.float32, device=dout.device) if ctx.needs_input_grad[1] else None
        _lib.call('as_pointer_keys_bwd', dout.data_ptr(), en.data_ptr(), K._ptr(dkey), K._ptr(dend), R, N, D, _lib.stream_ptr())
        return dkey, (None if dend is None else dend.view(eshape)), None


def pointer_keys(key, end_embedding, entity_num):
    """[kernel] Append a row and place the end-flag embedding at index entity_num (action_arg_head.py:118-129)."""
    require_cuda(key)
    return _PointerKeys.apply(key, end_embedding, entity_num)


class _PointerLogits(torch.autograd.Function):
    """[kernel] out[r,s,n] = n < key_len[r] ? scale * <query[r,s], key[r,n]> : fill (as_pointer_logits_fwd / _bwd)."""

    @staticmethod
    def forward(ctx, query, key, key_len, scale, fill):
        R, S, D = query.shape
        N = key.shape[1]
        query, key = query.contiguous().float(), key.contiguous().float()
        kl = None if key_len is None else key_len.contiguous().long()
        out = torch.empty((R, S, N), dtype=torch.float32, device=query.device)
        _lib.call('as_pointer_logits_fwd', query.data_ptr(), key.data_ptr(), K._ptr(kl), out.data_ptr(), R, S, N, D,
                  float(scale), float(fill), _lib.stream_ptr())
        ctx.dims = (R, S, N, D, float(scale))
        ctx.save_for_backward(query, key, kl)
        return out

    @staticmethod
    def backward(ctx, dout):
        query, key, kl = ctx.saved_tensors
        R, S, N, D, scale = ctx.dims
        dout = dout.contiguous().float()
        dq = torch.empty_like(query) if ctx.needs_input_grad[0] else None
        dk = torch.empty_like(key) if ctx.needs_input_grad[1] else None
        if dq is not None or dk is not None:
            _lib.call('as_pointer_logits_bwd', dout.data_ptr(), query.data_ptr(), key.data_ptr(), K._ptr(kl), K._ptr(dq),
                      K._ptr(dk), R, S, N, D, scale, _lib.stream_ptr())
        return dq, dk, None, None, None


def pointer_logits(query, key, entity_num, temperature=1.0):
    """[kernel] logits[r,n] = <query[r], key[r,n]> / temperature, keys n >= entity_num masked to -1e9 / temperature
    (head/action_arg_head.py:343-363)."""
    require_cuda(query)
    inv_t = 1.0 / temperature
    return _PointerLogits.apply(query.unsqueeze(1), key, entity_num, inv_t, -1e9 * inv_t).squeeze(1)


def _su_step(head, ae, h, c):
    q = head.query_fc2(head.query_fc1(ae))
    out, h, c = K.lnlstm_layer(q.unsqueeze(0), h, c, head.lstm.layers[0].cell)
    return out[0], h, c


SU_DEFER_TRIM = None     # a list while an actor step is being captured into a CUDA graph: the sampling loop appends its
                         # device-side step count there and returns untrimmed [R, 64, ...] tensors (learner.ActorStep)
SU_STEPS_HINT = None     # host-side max(selected_units_num) supplied by the caller (e.g. the collate step, which has it
                         # on the CPU for free); lets the learner step run without any device->host sync


class _SUPrefix(torch.autograd.Function):
    """[kernel] embedding of the units picked so far, for all S steps at once (as_su_prefix_fwd/bwd)."""

    @staticmethod
    def forward(ctx, key, sel, entity_num, su_num, S):
        R, N1, D = key.shape
        key = key.contiguous()
        sel = sel.contiguous()
        emb = torch.empty((R, S, D), dtype=torch.float32, device=key.device)
        cnt = torch.empty((R, S), dtype=torch.float32, device=key.device)
        isnew = torch.empty((R, S), dtype=torch.uint8, device=key.device)
        _lib.call('as_su_prefix_fwd', key.data_ptr(), sel.data_ptr(), entity_num.data_ptr(), su_num.data_ptr(),
                  emb.data_ptr(), cnt.data_ptr(), isnew.data_ptr(), R, N1, D, sel.shape[1], S, _lib.stream_ptr())
        ctx.dims = (R, N1, D, S)
        ctx.save_for_backward(sel, su_num, cnt, isnew)
        return emb

    @staticmethod
    def backward(ctx, demb):
        sel, su_num, cnt, isnew = ctx.saved_tensors
        R, N1, D, S = ctx.dims
        demb = demb.contiguous()
        dkey = torch.zeros((R, N1, D), dtype=torch.float32, device=demb.device)
        _lib.call('as_su_prefix_bwd', demb.data_ptr(), sel.data_ptr(), su_num.data_ptr(), cnt.data_ptr(),
                  isnew.data_ptr(), dkey.data_ptr(), R, N1, D, sel.shape[1], S, _lib.stream_ptr())
        return dkey, None, None, None, None


class _SUMask(torch.autograd.Function):
    """[kernel] logits mask of every step + 1/temperature (as_su_mask_logits); masked entries are -1e9/temperature."""

    @staticmethod
    def forward(ctx, x, sel, entity_num, S, scale):
        R, _, N1 = x.shape
        x = x.contiguous()
        sel = sel.contiguous()
        out = torch.empty_like(x)
        _lib.call('as_su_mask_logits', x.data_ptr(), sel.data_ptr(), entity_num.data_ptr(), out.data_ptr(), R, N1,
                  sel.shape[1], S, float(scale), float(-1e9 * scale), _lib.stream_ptr())
        ctx.args = (R, N1, S, float(scale))
        ctx.save_for_backward(sel, entity_num)
        return out

    @staticmethod
    def backward(ctx, dy):
        sel, entity_num = ctx.saved_tensors
        R, N1, S, scale = ctx.args
        dy = dy.contiguous()
        dx = torch.empty_like(dy)
        _lib.call('as_su_mask_logits', dy.data_ptr(), sel.data_ptr(), entity_num.data_ptr(), dx.data_ptr(), R, N1,
                  sel.shape[1], S, scale, 0.0, _lib.stream_ptr())
        return dx, None, None, None, None


def selected_units_train(head, key, ae0, entity_num, selected_units, selected_units_num, temperature):
    """Teacher-forced pointer loop (action_arg_head.py:168-216) WITHOUT the loop: the picks are given, so the embedding
    after every step (as_su_prefix_fwd) and the logits mask of every step (as_su_mask_logits) depend on
    `selected_units` only; the embedding MLP, the query MLP, the LayerNorm-LSTM and the key dot products then run once
    over all S steps.  Same results as selected_units_train_loop (tests/test_model_gpu.py)."""
    R, N1, D = key.shape
    S = int(SU_STEPS_HINT) if SU_STEPS_HINT is not None else max(int(selected_units_num.max()), 1)
    S = max(1, min(S, selected_units.shape[1]))
    entity_num = entity_num.contiguous()
    su_num = selected_units_num.contiguous()
    emb = _SUPrefix.apply(key, selected_units, entity_num, su_num, S)                    # [R,S,32] after each step
    delta = head.embed_fc2(head.embed_fc1(emb.reshape(R * S, D))).view(R, S, -1)
    ae_seq = ae0.unsqueeze(1) + delta                                                    # embedding after step i
    x = torch.cat([ae0.unsqueeze(1), ae_seq[:, :S - 1]], dim=1)                          # embedding fed to step i
    x = x.transpose(0, 1).contiguous()                                                   # time-major [S,R,1024]
    q = head.query_fc2(head.query_fc1(x.view(S * R, -1))).view(S, R, -1)
    zeros = key.new_zeros(R, q.shape[-1])
    out, _, _ = K.lnlstm_layer(q, zeros, zeros, head.lstm.layers[0].cell)               # [S,R,32]
    logits = _PointerLogits.apply(out.transpose(0, 1), key, None, 1.0, 0.0)              # [kernel] [R,S,N1]
    logits = _SUMask.apply(logits, selected_units, entity_num, S, 1.0 / temperature)
    return logits, selected_units, ae_seq[:, S - 1], selected_units_num, None


def selected_units_train_loop(head, key, ae0, entity_num, selected_units, selected_units_num, temperature):
    """Step-by-step version of selected_units_train (kept as the readable restatement and test reference).  The loop length is the maximum needed by the
    batch: taken from SU_STEPS_HINT when the caller knows it on the host, else read once from the device (one
    sync per call instead of the reference's 2-4 per step).  Any length >= max(selected_units_num) gives identical
    valid logits and embedding (ended rows keep a frozen selection)."""
    R, N1, D = key.shape
    dev = key.device
    ar = torch.arange(R, device=dev)
    S = int(SU_STEPS_HINT) if SU_STEPS_HINT is not None else max(int(selected_units_num.max()), 1)
    S = max(1, min(S, selected_units.shape[1]))
    h = key.new_zeros(R, 32)
    c = key.new_zeros(R, 32)
    # masks are built from one-hot tensors (no scalar index_put: keeps the loop free of host->device copies, so it
    # can be captured in a CUDA graph)
    end_oh = F.one_hot(entity_num, N1).bool()
    lm = (torch.arange(N1, device=dev).unsqueeze(0) < (entity_num + 1).unsqueeze(1)) & ~end_oh
    one_hot = key.new_zeros(R, N1)
    end_flag = torch.zeros(R, dtype=torch.bool, device=dev)
    has_units = (selected_units_num != 0).unsqueeze(1)
    ae = ae0
    logits = []
    for i in range(S):
        if i > 0:
            if i == 1:
                lm = lm | end_oh                                     # end flag selectable from the second pick on
            lm = lm & ~F.one_hot(selected_units[:, i - 1], N1).bool()   # mask the unit picked at the previous step
        q, h, c = _su_step(head, ae, h, c)
        logits.append(torch.einsum('rk,rnk->rn', q, key).masked_fill(~lm, -1e9))
        end_flag = end_flag | (selected_units[:, i] == entity_num)
        pick = F.one_hot(selected_units[:, i], N1).to(one_hot.dtype) * (~end_flag).unsqueeze(1)
        one_hot = torch.maximum(one_hot, pick)
        emb = torch.einsum('rn,rnk->rk', one_hot, key)
        emb = torch.where(has_units, emb / one_hot.sum(dim=1, keepdim=True), emb)
        ae = ae0 + head.embed_fc2(head.embed_fc1(emb))
    logits = torch.stack(logits, dim=1)
    if temperature != 1.0:
        logits = logits / temperature
    return logits, selected_units, ae, selected_units_num, None


def _su_sample_weights(head):
    """fp32 operands of the fused sampling kernel, derived from the head's weights and cached until they change:
    mt = (Wq1 Wf2)^T, c1 = Wq1 bf2 (see su_sample.cu) and the transposed small matrices."""
    cell = head.lstm.layers[0].cell
    src = (head.query_fc1[0].weight, head.query_fc1[0].bias, head.query_fc2[0].weight, head.query_fc2[0].bias,
           head.embed_fc1[0].weight, head.embed_fc1[0].bias, head.embed_fc2[0].weight, head.embed_fc2[0].bias,
           cell.weight_ih, cell.weight_hh)
    tag = (K._SHADOW_GEN[0],) + tuple((t._version, t.data_ptr()) for t in src)
    hit = getattr(head, '_su_sample_cache', None)
    if hit is not None and hit[0] == tag:
        return hit[1]
    with torch.no_grad():
        wq1, bq1, wq2, bq2, wf1, bf1, wf2, bf2, wih, whh = (t.detach().float() for t in src)
        w = {'mt': (wq1 @ wf2).t().contiguous(), 'c1': (wq1 @ bf2).contiguous(), 'wq1t': wq1.t().contiguous(), 'bq1': bq1,
             'wf1t': wf1.t().contiguous(), 'bf1': bf1.contiguous(), 'wq2t': wq2.t().contiguous(), 'bq2': bq2.contiguous(),
             'wih_t': wih.t().contiguous(), 'whh_t': whh.t().contiguous()}
    head._su_sample_cache = (tag, w)
    return w


def selected_units_sample(head, key, ae0, entity_num, su_mask, noise, temperature, extra_units):
    """Sampling pointer loop (action_arg_head.py:262-313): one launch of the fused kernel (su_sample.cu) for all 64
    steps, then the reference's early-break trim.  `APPLESTAR_SU_FUSED=0` selects the step-by-step loop below."""
    if os.environ.get('APPLESTAR_SU_FUSED', '1') == '0' or torch.is_grad_enabled() and (key.requires_grad or ae0.requires_grad):
        return selected_units_sample_loop(head, key, ae0, entity_num, su_mask, noise, temperature, extra_units)
    require_cuda(key)
    R, N1, D = key.shape
    S = C.MAX_SELECTED_UNITS_NUM
    dev = key.device
    w = _su_sample_weights(head)
    cell = head.lstm.layers[0].cell
    key = key.float().contiguous()
    c0 = torch.addmm(w['bq1'], ae0.float(), w['wq1t'])                    # query_fc1(ae0) pre-activation, fp32
    noise = noise.float().contiguous()
    entity_num = entity_num.long().contiguous()
    mask8 = su_mask.to(torch.uint8).contiguous()
    logits = torch.empty(R, S, N1, device=dev)
    units = torch.zeros(R, S, dtype=torch.long, device=dev)     # steps after the batch-wide end are not computed: keep their
    num = torch.empty(R, dtype=torch.long, device=dev)          # indices valid for the (trimmed-away) gathers downstream
    extra = torch.empty(R, N1, device=dev)
    emb = torch.empty(R, D, device=dev)
    steps_row = torch.empty(R, dtype=torch.int32, device=dev)
    ended_rows = torch.zeros(2, dtype=torch.int32, device=dev)          # [ended rows, latest end step]: early loop exit
    li, lh, lc = cell.layernorm_i, cell.layernorm_h, cell.layernorm_c
    a = _lib.AsSuSampleArgs()
    for name, t in (('key', key), ('c0', c0), ('c1', w['c1']), ('mt', w['mt']), ('wf1t', w['wf1t']), ('bf1', w['bf1']),
                    ('wq2t', w['wq2t']), ('bq2', w['bq2']), ('wih_t', w['wih_t']), ('whh_t', w['whh_t']),
                    ('ln_i_w', li.weight), ('ln_i_b', li.bias), ('ln_h_w', lh.weight), ('ln_h_b', lh.bias),
                    ('ln_c_w', lc.weight), ('ln_c_b', lc.bias), ('entity_num', entity_num), ('su_mask', mask8),
                    ('noise', noise), ('logits', logits), ('units', units), ('num', num), ('extra', extra), ('emb', emb),
                    ('steps_row', steps_row), ('ended_rows', ended_rows)):
        assert t.is_contiguous() and t.device == dev
        setattr(a, name, t.data_ptr())
    a.R, a.N1, a.S, a.extra_units, a.temperature = R, N1, S, int(bool(extra_units)), float(temperature)
    import ctypes
    _lib.call('as_su_sample', ctypes.byref(a), _lib.stream_ptr())
    ae = ae0 + head.embed_fc2(head.embed_fc1(emb))
    steps_dev = steps_row.max().clamp(min=1).long()
    if SU_DEFER_TRIM is not None:       # graph capture: keep all S steps, hand the count out as a device scalar
        SU_DEFER_TRIM.append(steps_dev)
    else:
        steps = int(steps_dev.item())                                    # one host read
        logits, units = logits[:, :steps], units[:, :steps]
    return logits, units, ae, num, extra


def selected_units_sample_loop(head, key, ae0, entity_num, su_mask, noise, temperature, extra_units):
    """Step-by-step form of the sampling pointer loop (action_arg_head.py:262-313) with Gumbel-max picks; runs all 64 steps on the
    device without the per-step `end_flag.all()` host sync, then trims to the reference's step count."""
    R, N1, D = key.shape
    dev = key.device
    ar = torch.arange(R, device=dev)
    S = C.MAX_SELECTED_UNITS_NUM
    h = key.new_zeros(R, 32)
    c = key.new_zeros(R, 32)
    lm = torch.arange(N1, device=dev).unsqueeze(0) < (entity_num + 1).unsqueeze(1)
    end_col = entity_num.long().unsqueeze(1)
    lm.scatter_(1, end_col, False)              # scalar scatter: no host tensor, safe under graph capture
    num = torch.where(su_mask, torch.full((R,), S, dtype=torch.long, device=dev), torch.zeros(R, dtype=torch.long, device=dev))
    end_flag = ~su_mask
    one_hot = key.new_zeros(R, N1)
    ae = ae0
    logits, units, all_ended = [], [], []
    last_logits = torch.zeros(R, N1, device=dev)
    done = torch.zeros((), dtype=torch.bool, device=dev)       # "reference loop has broken" flag
    result = None
    for i in range(S):
        if i > 0:
            if i == 1:
                lm.scatter_(1, end_col, True)
            lm.scatter_(1, result.unsqueeze(1), False)
        q, h, c = _su_step(head, ae, h, c)
        step_logits = torch.einsum('rk,rnk->rn', q, key).masked_fill(~lm, -1e9)
        if temperature != 1.0:
            step_logits = step_logits / temperature
        result = K.gumbel_argmax(step_logits, noise[i])
        hit = result == entity_num
        num = torch.where(hit & ~end_flag & ~done, torch.full_like(num, i + 1), num)
        end_flag_new = end_flag | hit
        live = (~end_flag_new & ~done).to(one_hot.dtype).unsqueeze(1)
        one_hot = torch.maximum(one_hot, F.one_hot(result, N1).to(one_hot.dtype) * live)
        cnt = one_hot.sum(dim=1, keepdim=True)
        emb = torch.einsum('rn,rnk->rk', one_hot, key)
        emb = torch.where(cnt != 0, emb / cnt.clamp(min=1), emb)
        ae_new = ae0 + head.embed_fc2(head.embed_fc1(emb))
        ae = torch.where(done, ae, ae_new)
        last_logits = torch.where(done, last_logits, step_logits)
        end_flag = torch.where(done, end_flag, end_flag_new)
        logits.append(step_logits)
        units.append(result)
        done = done | end_flag.all()
        all_ended.append(done)
    steps_dev = ((~torch.stack(all_ended)).sum() + 1).clamp(max=S)
    if SU_DEFER_TRIM is not None:       # graph capture: keep all S steps, hand the count out as a device scalar
        SU_DEFER_TRIM.append(steps_dev)
        steps = S
    else:
        steps = int(steps_dev.item())                            # one host read for the whole loop
    logits = torch.stack(logits[:steps], dim=1)
    units = torch.stack(units[:steps], dim=1)
    extra = key.new_zeros(R, N1)
    if extra_units:
        end_logit = last_logits[ar, entity_num]
        extra = ((last_logits > end_logit.unsqueeze(1)) & ~end_flag.unsqueeze(1)).float()
    return logits, units, ae, num, extra
