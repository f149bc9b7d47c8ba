"""Host-side mirror of the reference `Model` (distar/agent/default/model/model.py:22-189).

Same constructor, same five entry methods with the same signatures, same `state_dict()` key names
and shapes so reference checkpoints load and the reference agent / learner can
hold this object unchanged. All arithmetic goes through `applestar_b200.ops`, which launches the
sm_100a kernels behind the C-ABI on `torch.cuda.current_stream()`; there is no CPU path.

Module tree (names are the state_dict contract; do not rename):
  encoder.{scalar_encoder, spatial_encoder, entity_encoder, scatter_project}
  policy.{action_type_head, delay_head, queued_head, selected_units_head, target_unit_head, location_head}
  value_networks.<field>, core_lstm
"""
import math
import os

import torch
import torch.nn as nn

from . import config as C
from . import ops, spatial, tables


# ----------------------------------------------------------------------------------------------
# building blocks (ctools/torch_utils/network/nn_module.py:119-174,231-314; res_block.py)
def _xavier_normal(w):
    nn.init.xavier_normal_(w)


class FC(nn.Sequential):
    """fc_block: `.0` = Linear; ReLU fused into the GEMM epilogue when `relu`."""

    def __init__(self, cin, cout, relu=False, out_bf16=False):
        super().__init__(nn.Linear(cin, cout))
        _xavier_normal(self[0].weight)
        self.relu, self.out_bf16 = relu, out_bf16

    def forward(self, x):
        return ops.linear(x, self[0].weight, self[0].bias, self.relu, self.out_bf16)


class FCLN(nn.Sequential):
    """fc_block with norm_type='LN': `.0` Linear, `.1` LayerNorm, optional ReLU."""

    def __init__(self, cin, cout, relu=False):
        super().__init__(nn.Linear(cin, cout), nn.LayerNorm(cout))
        _xavier_normal(self[0].weight)
        self.relu = relu

    def forward(self, x, residual=None):
        y = ops.linear(x, self[0].weight, self[0].bias, False)
        return ops.layer_norm(y, self[1].weight, self[1].bias, relu=self.relu)


class Conv(nn.Sequential):
    """conv2d_block (no norm): `.0` = Conv2d, run as an implicit GEMM on the tcgen05 kernel over a zero-padded NHWC bf16
    map (spatial.py).  The one convolution built with `nhwc=False` (the location head's final 32 -> 1 layer) only holds
    the parameters: its arithmetic is spatial.conv_c32_to_1."""

    def __init__(self, cin, cout, k, pad, relu=False, nhwc=False):
        super().__init__(nn.Conv2d(cin, cout, k, 1, pad))
        _xavier_normal(self[0].weight)
        self.relu, self.pad, self.nhwc = relu, pad, nhwc

    def forward(self, x, grad_premasked=False, mask_input_grad=False):
        """grad_premasked / mask_input_grad: see spatial.ConvNHWC (who applies the ReLU backward mask of a conv -> ReLU ->
        single consumer chain: the consumer's data-gradient epilogue instead of a separate pass)."""
        if not self.nhwc:
            raise RuntimeError('this convolution is evaluated by its owner (no library convolution path exists)')
        return spatial.conv(x, self[0].weight, self[0].bias, self.relu, grad_premasked, mask_input_grad)

    def fork(self, x, grad_premasked=False):
        """-> (self(x), x'): x' aliases x and must replace it in x's OTHER consumers, so that their gradient is added to
        this convolution's data gradient in its GEMM epilogue (spatial.ConvNHWC)."""
        return spatial.conv_fork(x, self[0].weight, self[0].bias, self.relu, grad_premasked)


def _eye_table(n):
    return nn.Embedding.from_pretrained(torch.eye(n), freeze=True, padding_idx=None)


def _binary_table(bits):
    v = torch.arange(2 ** bits)
    shifts = torch.arange(bits - 1, -1, -1)
    return nn.Embedding.from_pretrained(((v.unsqueeze(-1) >> shifts) & 1).float(), freeze=True, padding_idx=None)


class GLU(nn.Module):
    """module_utils.py:508-524: gate on the input, context through layer1."""

    def __init__(self, input_dim, output_dim, context_dim):
        super().__init__()
        self.layer1 = FC(context_dim, input_dim)
        self.layer2 = FC(input_dim, output_dim)

    def forward(self, x, context):
        return self.layer2(ops.sigmoid_gate(self.layer1(context), x))


# ----------------------------------------------------------------------------------------------
class Attention(nn.Module):
    def __init__(self, dim, head_dim, head_num):
        super().__init__()
        self.head_num, self.head_dim = head_num, head_dim
        self.attention_pre = FC(dim, head_dim * head_num * 3, out_bf16=True)
        self.project = FC(head_dim * head_num, dim)

    def forward(self, x, key_len):
        qkv = self.attention_pre(x)
        return self.project(ops.attention(qkv, key_len, self.head_num, self.head_dim))


class TransformerLayer(nn.Module):
    def __init__(self, dim, head_dim, hidden_dim, head_num, post_ln):
        super().__init__()
        self.attention = Attention(dim, head_dim, head_num)
        self.layernorm1 = nn.LayerNorm(dim)
        self.mlp = nn.Sequential(FC(dim, hidden_dim, relu=True, out_bf16=True), FC(hidden_dim, dim, relu=True))
        self.layernorm2 = nn.LayerNorm(dim)
        self.post_ln = post_ln

    def forward(self, x32, x16, key_len, final_relu=False):
        """(fp32 residual stream, bf16 GEMM operand) in and out."""
        l1, l2 = self.layernorm1, self.layernorm2
        if self.post_ln and x32 is None:            # bf16 stream: x16 = (residual handle, GEMM-operand handle) of one tensor
            xa, xb = x16
            att = self.attention
            a = ops.attention(att.attention_pre(xb), key_len, att.head_num, att.head_dim)
            pj, f1, f2 = att.project[0], self.mlp[0][0], self.mlp[1][0]
            xa, xb = ops.linear_ln16(a, xa, pj.weight, pj.bias, l1.weight, l1.bias)
            xa, xb = ops.mlp_ln16(xb, xa, f1.weight, f1.bias, f2.weight, f2.bias, l2.weight, l2.bias, final_relu)
            return None, (xa, xb)
        if self.post_ln and ops.transformer_layer_fused_ok(self):     # module_utils.py:133-139, two fused autograd nodes
            att = self.attention
            a = ops.attention(att.attention_pre(x16), key_len, att.head_num, att.head_dim)
            pj, f1, f2 = att.project[0], self.mlp[0][0], self.mlp[1][0]
            x32, x16 = ops.linear_ln(a, x32, pj.weight, pj.bias, l1.weight, l1.bias)
            x32, x16 = ops.mlp_ln(x16, x32, f1.weight, f1.bias, f2.weight, f2.bias, l2.weight, l2.bias, final_relu)
        elif self.post_ln:
            x32, x16 = ops.layer_norm_dual(self.attention(x16, key_len), l1.weight, l1.bias, residual=x32)
            x32, x16 = ops.layer_norm_dual(self.mlp(x16), l2.weight, l2.bias, residual=x32, relu=final_relu)
        else:                # :140-148
            x32 = x32 + self.attention(ops.layer_norm_dual(x32, l1.weight, l1.bias)[1], key_len)
            x32 = x32 + self.mlp(ops.layer_norm_dual(x32, l2.weight, l2.bias)[1])
            x16 = None
        return x32, x16


class Transformer(nn.Module):
    """module_utils.py:154-199; the input embedding layer is applied by the caller-specific front end."""

    def __init__(self, input_dim, head_dim, hidden_dim, output_dim, head_num, layer_num, post_ln):
        super().__init__()
        self.embedding = FC(input_dim, output_dim, relu=True)
        self.layers = nn.ModuleList([TransformerLayer(output_dim, head_dim, hidden_dim, head_num, post_ln)
                                     for _ in range(layer_num)])

    def run_layers(self, x32, x16, key_len=None, final_relu=False):
        for i, layer in enumerate(self.layers):
            x32, x16 = layer(x32, x16, key_len, final_relu and i == len(self.layers) - 1)
        return x32, x16


class EntityEncoder(nn.Module):
    """obs_encoder/entity_encoder.py:20-96. The 997-wide one-hot concat is never materialised: the
    embedding GEMM over one-hot columns is evaluated as a gather-sum of weight columns (ops.entity_embed)."""

    def __init__(self):
        super().__init__()
        self.encode_modules = nn.ModuleDict()
        for name, _, arc, num in C.ENTITY_FIELDS:
            if arc == 'one_hot':
                self.encode_modules[name] = _eye_table(num)
            elif arc == 'binary':
                self.encode_modules[name] = _binary_table(num)
        self.transformer = Transformer(C.ENTITY_INPUT_DIM, 128, 1024, 256, 2, 3, post_ln=True)
        self.entity_fc = FC(256, 256, relu=True, out_bf16=True)
        self.embed_fc = FC(256, 256, relu=True)

    def forward(self, entity_info, entity_num):
        emb = self.transformer.embedding[0]
        layer0 = self.transformer.layers[0]
        if ops.ENTITY_BF16 and ops.transformer_layer_fused_ok(layer0) and emb.weight.shape[0] == 256:
            # bf16 residual stream (what the reference's autocast path holds): LayerNorm traffic 0.4 - 0.6x
            x16 = ops.entity_embed(entity_info, emb.weight, emb.bias, bf16_only=True)
            _, (xa, xb) = self.transformer.run_layers(None, (x16, x16), entity_num, final_relu=True)
            return self.entity_fc(xb), self.embed_fc(ops.masked_mean16(xa, entity_num))
        x32, x16 = ops.entity_embed(entity_info, emb.weight, emb.bias)    # [R,N,256], ReLU applied
        # the reference's in-place ReLU after the transformer feeds both entity_fc and the mean (A1): fused
        # into the last LayerNorm
        x32, x16 = self.transformer.run_layers(x32, x16, entity_num, final_relu=True)
        entity_embeddings = self.entity_fc(x16)                           # bf16: consumed only by GEMMs
        pooled = ops.masked_mean(x32, entity_num)
        return entity_embeddings, self.embed_fc(pooled)


class ResBlock(nn.Module):
    def __init__(self, ch):
        super().__init__()
        self.conv1 = Conv(ch, ch, 3, 1, relu=True, nhwc=True)
        self.conv2 = Conv(ch, ch, 3, 1, nhwc=True)

    def forward(self, x):
        h, x = self.conv1.fork(x, grad_premasked=True)           # h feeds conv2 only: conv2's dgrad applies h's ReLU mask
        return spatial.add_relu(self.conv2(h, mask_input_grad=True), x)


class SpatialEncoder(nn.Module):
    """obs_encoder/spatial_encoder.py:8-90."""

    def __init__(self):
        super().__init__()
        self.project = Conv(56, 32, 1, 0, relu=True, nhwc=True)
        self.encode_modules = nn.ModuleDict({n: _eye_table(k) for n, arc, k in C.SPATIAL_FIELDS if arc == 'one_hot'})
        self.downsample = nn.ModuleList([Conv(32, 64, 3, 1, True, nhwc=True), Conv(64, 128, 3, 1, True, nhwc=True),
                                         Conv(128, 128, 3, 1, True, nhwc=True)])
        self.res = nn.ModuleList([ResBlock(128) for _ in range(4)])
        self.fc = FC(128 * (C.SPATIAL_Y // 8) * (C.SPATIAL_X // 8), 256, relu=True)

    def forward(self, spatial_info, x, head_rows=None):
        """`x` = the 2x2 max-pooled output of the 1x1 project conv + ReLU as a padded NHWC bf16 map (spatial.spatial_front,
        which also holds the ScatterConnection and the 56-channel input).
        -> (embedded_spatial [R,256], map_skip).  map_skip[0..2] (pre-pool maps, only read by the U-Net variant of
        the location head, which the default config disables) are not kept; map_skip[3..6] are the 19x20x128 maps
        as zero-padded NHWC bf16.  Every convolution is an implicit GEMM on the tcgen05 kernel."""
        skip = [None, None, None]
        for i, conv in enumerate(self.downsample):                # 3 x (2x2 max-pool, 3x3 conv + ReLU); the first pool is
            if i > 0:                                             # part of the fused front kernel
                x = spatial.max_pool2(x, relu_input=True)         # the pool's backward applies the conv's ReLU mask
            x = conv(x, grad_premasked=(i < 2))                   # outputs 0 and 1 feed only the next pool
        for block in self.res:
            x, x_heads = ops.prefix_fork(x, head_rows)       # the location head reads the first head_rows rows only
            skip.append(x_heads)
            x = block(x)
        H8, W8 = C.SPATIAL_Y // 8, C.SPATIAL_X // 8
        return spatial.spatial_fc(x, self.fc[0].weight, self.fc[0].bias, 128, H8, W8), skip


class BeginningBuildOrderEncoder(nn.Module):
    """obs_encoder/scalar_encoder.py:19-53 (pre-LN transformer, 20 tokens, no mask)."""

    def __init__(self):
        super().__init__()
        self.transformer = Transformer(174 + 20 + 20, 8, 128, 64, 2, 3, post_ln=False)
        self.embedd_fc = FC(64, 64, relu=True)
        self.action_one_hot = _eye_table(174)
        self.order_one_hot = _eye_table(20)
        self.location_binary = _binary_table(10)

    def forward(self, order, location):
        emb = self.transformer.embedding[0]
        x = ops.build_order_embed(order, location, emb.weight, emb.bias)   # [R,20,64], ReLU applied
        x, _ = self.transformer.run_layers(x, None, None)
        return self.embedd_fc(x.mean(dim=1))


class ScalarEncoder(nn.Module):
    """obs_encoder/scalar_encoder.py:56-132."""

    def __init__(self):
        super().__init__()
        self.encode_modules = nn.ModuleDict()
        for name, arc, cin, cout, _, _ in C.SCALAR_FIELDS:
            if arc == 'one_hot':
                e = nn.Embedding(cin, cout)
                nn.init.xavier_uniform_(e.weight)
                self.encode_modules[name] = e
            elif arc == 'fc':
                self.encode_modules[name] = FC(cin, cout, relu=True)
        from .synth import time_position_array
        self.position_array = nn.Parameter(time_position_array(32), requires_grad=False)
        self.encode_modules['beginning_order'] = BeginningBuildOrderEncoder()

    def forward(self, s):
        emb, ctx, base = [], [], []
        for name, arc, cin, cout, is_ctx, is_base in C.SCALAR_FIELDS:
            if arc == 'identity':
                continue
            if arc == 'one_hot':
                e = ops.embedding_relu(s[name], self.encode_modules[name].weight)
            elif arc == 'transformer':
                e = self.encode_modules[name](s[name], s['bo_location'])
            else:
                e = self.encode_modules[name](s[name])            # narrow storage dtypes go straight to the bf16 operand
            emb.append(e)
            if is_ctx:
                ctx.append(e)
            if is_base:
                base.append(e)
        emb.append(ops.time_encode(s['time'], self.position_array))
        return torch.cat(emb, 1), torch.cat(ctx, 1), torch.cat(base, 1)


_SIDE_STREAMS = {}


class side_stream:
    """Runs the enclosed (small-kernel) branch on a second CUDA stream, forked from and later joined to the current
    one: the value towers and the scalar encoder are chains of ~100 launches that each fill a few SMs, and they do
    not depend on the big kernels issued next to them.  Autograd replays the backward of the branch on the same
    stream, so the overlap holds in both directions; CUDA-graph capture records the fork / join as graph edges."""

    def __init__(self, device, enabled=True):
        self.enabled = enabled and device.type == 'cuda' and os.environ.get('APPLESTAR_SIDE_STREAM', '1') == '1'
        if self.enabled:
            if device.index not in _SIDE_STREAMS:
                _SIDE_STREAMS[device.index] = torch.cuda.Stream(device=device)
            self.side = _SIDE_STREAMS[device.index]
            self.main = torch.cuda.current_stream(device)

    def __enter__(self):
        if self.enabled:
            self.side.wait_stream(self.main)
            self.ctx = torch.cuda.stream(self.side)
            self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.enabled:
            self.ctx.__exit__(*exc)
        return False

    def join(self, tensors):
        """Call on the main stream before the branch's results are consumed."""
        if self.enabled:
            self.main.wait_stream(self.side)
            for t in tensors:
                t.record_stream(self.main)


class Encoder(nn.Module):
    """model/encoder.py:15-45."""

    def __init__(self):
        super().__init__()
        self.encoder = nn.ModuleDict()
        self.scalar_encoder = ScalarEncoder()
        self.spatial_encoder = SpatialEncoder()
        self.entity_encoder = EntityEncoder()
        self.scatter_project = FC(256, 32, relu=True)

    def forward(self, spatial_info, entity_info, scalar_info, entity_num, head_rows=None):
        """head_rows: the learner's policy heads read only the first `head_rows` rows of entity_embeddings / map_skip
        (model.py:128-137); they are then returned already cut (ops.prefix_fork)."""
        br = side_stream(entity_num.device)
        with br:
            embedded_scalar, scalar_context, baseline_feature = self.scalar_encoder(scalar_info)
        entity_embeddings, embedded_entity = self.entity_encoder(entity_info, entity_num)
        ent_full, entity_embeddings = ops.prefix_fork(entity_embeddings, head_rows)
        proj = self.scatter_project(ent_full)
        # ScatterConnection ('add') + one-hot / effect planes + 1x1 project conv + ReLU in padded NHWC bf16
        pw = self.spatial_encoder.project[0]
        projected = spatial.spatial_front(proj, entity_info['x'], entity_info['y'], entity_num, spatial_info,
                                          pw.weight, pw.bias)
        embedded_spatial, map_skip = self.spatial_encoder(spatial_info, projected, head_rows)
        br.join([embedded_scalar, scalar_context, baseline_feature])
        lstm_input = torch.cat([embedded_scalar, embedded_entity, embedded_spatial], dim=-1)
        return lstm_input, scalar_context, baseline_feature, entity_embeddings, map_skip


# ----------------------------------------------------------------------------------------------
class LayerNormLSTMCell(nn.Module):
    """lstm.py:120-153 parameter holder (randn weights as the reference)."""

    def __init__(self, input_size, hidden_size):
        super().__init__()
        self.weight_ih = nn.Parameter(torch.randn(4 * hidden_size, input_size))
        self.weight_hh = nn.Parameter(torch.randn(4 * hidden_size, hidden_size))
        self.layernorm_i = nn.LayerNorm(4 * hidden_size)
        self.layernorm_h = nn.LayerNorm(4 * hidden_size)
        self.layernorm_c = nn.LayerNorm(hidden_size)


class LSTMLayer(nn.Module):
    def __init__(self, input_size, hidden_size):
        super().__init__()
        self.cell = LayerNormLSTMCell(input_size, hidden_size)


class StackedLNLSTM(nn.Module):
    """lstm.py:215-234: input [T,B,I], states = list over layers of (h, c)."""

    def __init__(self, input_size, hidden_size, num_layers):
        super().__init__()
        self.layers = nn.ModuleList([LSTMLayer(input_size if i == 0 else hidden_size, hidden_size)
                                     for i in range(num_layers)])

    def forward(self, x, states):
        out_states = []
        for layer, (h, c) in zip(self.layers, states):
            x, h, c = ops.lnlstm_layer(x, h, c, layer.cell)
            out_states.append((h, c))
        return x, out_states


# ----------------------------------------------------------------------------------------------
class ResFCBlock(nn.Module):
    """res_block.py:68-108 with LN."""

    def __init__(self, dim):
        super().__init__()
        self.fc1 = FCLN(dim, dim, relu=True)
        self.fc2 = FCLN(dim, dim)

    def forward(self, x):
        return ops.add_relu(self.fc2(self.fc1(x)), x)


class ActionTypeHead(nn.Module):
    """head/action_type_head.py:17-67."""

    def __init__(self, common_type):
        super().__init__()
        self.project = FC(384, 256, relu=True)
        self.res = nn.Sequential(ResFCBlock(256), ResFCBlock(256))
        self.action_fc = GLU(256, C.NUM_ACTIONS, 448)
        self.action_map_fc1 = FC(C.NUM_ACTIONS, 256, relu=True)
        self.action_map_fc2 = FC(256, 256)
        self.glu1 = GLU(256, 1024, 448)
        self.glu2 = GLU(384, 1024, 448)
        self.use_mask = common_type == 'play'
        self.race = 'zerg'

    def forward(self, lstm_output, scalar_context, temperature, action_type=None, noise=None):
        x = self.action_fc(self.res(self.project(lstm_output)), scalar_context)
        if temperature != 1.0:
            x = x / temperature
        if self.use_mask:
            m = torch.tensor(tables.flags(tables.ACTION_RACE_MASK_BITS[self.race]), device=x.device)
            x = x.masked_fill(~m.unsqueeze(0), -1e9)
        if action_type is None:
            action_type = ops.gumbel_argmax(x, noise)
        w1 = self.action_map_fc1[0]
        e1 = ops.one_hot_linear(action_type, w1.weight, w1.bias, relu=True)     # fc over a one-hot = column gather
        e1 = self.glu1(self.action_map_fc2(e1), scalar_context)
        return x, action_type, e1 + self.glu2(lstm_output, scalar_context)


class ArgHead(nn.Module):
    """DelayHead / QueuedHead (head/action_arg_head.py:27-86)."""

    def __init__(self, dim, use_temperature):
        super().__init__()
        self.fc1 = FC(1024, 256, relu=True)
        self.fc2 = FC(256, 256, relu=True)
        self.fc3 = FC(256, dim)
        self.embed_fc1 = FC(dim, 256, relu=True)
        self.embed_fc2 = FC(256, 1024)
        self.use_temperature = use_temperature

    def forward(self, embedding, temperature, action=None, noise=None):
        x = self.fc3(self.fc2(self.fc1(embedding)))
        if self.use_temperature and temperature != 1.0:
            x = x / temperature
        if action is None:
            action = ops.gumbel_argmax(x, noise)
        w = self.embed_fc1[0]
        e = self.embed_fc2(ops.one_hot_linear(action, w.weight, w.bias, relu=True))
        return x, action, embedding + e


class SelectedUnitsHead(nn.Module):
    """head/action_arg_head.py:89-328 ('selected_units_num' reduce type)."""

    def __init__(self, extra_units):
        super().__init__()
        self.key_fc = FC(256, 32)
        self.query_fc1 = FC(1024, 256, relu=True)
        self.query_fc2 = FC(256, 32)
        self.embed_fc1 = FC(32, 256, relu=True)
        self.embed_fc2 = FC(256, 1024)
        self.lstm = StackedLNLSTM(32, 32, 1)
        self.end_embedding = nn.Parameter(torch.empty(1, 32).uniform_(-1 / math.sqrt(32), 1 / math.sqrt(32)))
        self.extra_units = extra_units

    def forward(self, embedding, entity_embeddings, entity_num, temperature, selected_units_num=None,
                selected_units=None, su_mask=None, noise=None):
        key = ops.pointer_keys(self.key_fc(entity_embeddings), self.end_embedding, entity_num)   # [R,N+1,32]
        if selected_units is not None:
            return ops.selected_units_train(self, key, embedding, entity_num, selected_units, selected_units_num,
                                            temperature)
        return ops.selected_units_sample(self, key, embedding, entity_num, su_mask, noise, temperature,
                                         self.extra_units)


class TargetUnitHead(nn.Module):
    """head/action_arg_head.py:331-363."""

    def __init__(self):
        super().__init__()
        self.key_fc = FC(256, 32)
        self.query_fc1 = FC(1024, 32, relu=True)
        self.query_fc2 = FC(32, 32)

    def forward(self, embedding, entity_embeddings, entity_num, temperature, target_unit=None, noise=None):
        key = self.key_fc(entity_embeddings)
        query = self.query_fc2(self.query_fc1(embedding))
        logits = ops.pointer_logits(query, key, entity_num, temperature)
        if target_unit is None:
            target_unit = ops.gumbel_argmax(logits, noise)
        return logits, target_unit


class GatedResBlock(nn.Module):
    """module_utils.py:204-231."""

    def __init__(self, ch):
        super().__init__()
        self.conv1 = Conv(ch, ch, 3, 1, relu=True, nhwc=True)
        self.conv2 = Conv(ch, ch, 3, 1, nhwc=True)
        self.GateWeightG = nn.Sequential(Conv(ch, ch, 1, 0, True, nhwc=True), Conv(ch, ch, 1, 0, True, nhwc=True),
                                         Conv(ch, ch, 1, 0, True, nhwc=True), Conv(ch, ch, 1, 0, nhwc=True))
        self.UpdateSP = nn.Parameter(torch.full((1,), 0.1))

    def forward(self, x):
        h, x = self.conv1.fork(x, grad_premasked=True)           # single-consumer ReLU chains: the consumer's data-gradient
        g, x = self.GateWeightG[0].fork(x, grad_premasked=True)  # epilogue applies the producer's ReLU mask
        gates = list(self.GateWeightG)
        for j, layer in enumerate(gates[1:], 1):
            g = layer(g, grad_premasked=(j < len(gates) - 1), mask_input_grad=True)
        return spatial.gated_residual(self.conv2(h, mask_input_grad=True), g, self.UpdateSP, x)


class LocationHead(nn.Module):
    """head/action_arg_head.py:366-450 (bilinear upsample, gated res blocks, no FiLM / U-Net)."""

    def __init__(self):
        super().__init__()
        self.conv1 = Conv(128 + 4, 128, 1, 0, relu=True, nhwc=True)
        self.project_embed = FC(1024, (C.SPATIAL_Y // 8) * (C.SPATIAL_X // 8) * 4, relu=True)
        self.res = nn.ModuleList([GatedResBlock(128) for _ in range(4)])
        self.upsample = nn.ModuleList([Conv(128, 64, 3, 1, True, nhwc=True), Conv(64, 32, 3, 1, True, nhwc=True),
                                       Conv(32, 1, 3, 1)])

    def forward(self, embedding, map_skip, temperature, location=None, noise=None):
        # relu(cat[reshape(project_embed), map_skip[-1]]) as a 192-channel padded NHWC map (132 real channels)
        x = spatial.loc_cat(self.project_embed(embedding), map_skip[-1], 192)
        x = self.conv1(x)
        for i, block in enumerate(self.res):
            x = block(spatial.add(x, map_skip[len(map_skip) - i - 1]))
        x = self.upsample[0](spatial.upsample2(x))
        x = self.upsample[1](spatial.upsample2(x), grad_premasked=True)     # only consumer: the fused tail, which masks
        last = self.upsample[2][0]      # bilinear x2 + 3x3 conv 32 -> 1 + flatten + 1/temperature: one kernel, no 152x160x32 map
        logits = spatial.upconv_c32_to_1(x, last.weight, last.bias, 1.0 / temperature, relu_input=True)
        if location is None:
            location = ops.gumbel_argmax(logits, noise)
        return logits, location


class Policy(nn.Module):
    """model/policy.py:10-73."""

    def __init__(self, common_type, extra_units):
        super().__init__()
        self.action_type_head = ActionTypeHead(common_type)
        self.delay_head = ArgHead(128, use_temperature=False)
        self.queued_head = ArgHead(2, use_temperature=True)
        self.selected_units_head = SelectedUnitsHead(extra_units)
        self.target_unit_head = TargetUnitHead()
        self.location_head = LocationHead()
        self.register_buffer('_su_table', torch.tensor(tables.flags(tables.ACTION_HAS_SELECTED_UNITS)),
                             persistent=False)

    def forward(self, lstm_output, entity_embeddings, map_skip, scalar_context, entity_num, temperature, noise):
        action, logit = {}, {}
        logit['action_type'], action['action_type'], emb = self.action_type_head(
            lstm_output, scalar_context, temperature, None, noise['action_type'])
        logit['delay'], action['delay'], emb = self.delay_head(emb, temperature, None, noise['delay'])
        logit['queued'], action['queued'], emb = self.queued_head(emb, temperature, None, noise['queued'])
        su_mask = self._su_table[action['action_type']]
        logit['selected_units'], action['selected_units'], emb, sun, extra = self.selected_units_head(
            emb, entity_embeddings, entity_num, temperature, None, None, su_mask, noise['selected_units'])
        logit['target_unit'], action['target_unit'] = self.target_unit_head(
            emb, entity_embeddings, entity_num, temperature, None, noise['target_unit'])
        logit['target_location'], action['target_location'] = self.location_head(
            emb, map_skip, temperature, None, noise['target_location'])
        return action, sun, logit, extra

    def train_forward(self, lstm_output, entity_embeddings, map_skip, scalar_context, entity_num, action_info,
                      selected_units_num, temperature):
        action, logit = {}, {}
        logit['action_type'], action['action_type'], emb = self.action_type_head(
            lstm_output, scalar_context, temperature, action_info['action_type'])
        logit['delay'], action['delay'], emb = self.delay_head(emb, temperature, action_info['delay'])
        logit['queued'], action['queued'], emb = self.queued_head(emb, temperature, action_info['queued'])
        logit['selected_units'], action['selected_units'], emb, sun, _ = self.selected_units_head(
            emb, entity_embeddings, entity_num, temperature, selected_units_num, action_info['selected_units'])
        logit['target_unit'], action['target_unit'] = self.target_unit_head(
            emb, entity_embeddings, entity_num, temperature, action_info['target_unit'])
        logit['target_location'], action['target_location'] = self.location_head(
            emb, map_skip, temperature, action_info['target_location'])
        return action, sun, logit


class ResFCBlock2(nn.Module):
    """res_block.py:110-141."""

    def __init__(self, dim):
        super().__init__()
        self.fc1 = FC(dim, dim, relu=True)
        self.fc2 = FC(dim, dim)
        self.norm = nn.LayerNorm(dim)

    def forward(self, x):
        return ops.layer_norm(self.fc2(self.fc1(x)), self.norm.weight, self.norm.bias, residual=x)


class ValueEncoder(nn.Module):
    """obs_encoder/value_encoder.py:12-74 (`learner.use_value_feature`, self-play critics): opponent statistics through
    small fc layers, a second ScatterConnection (own + enemy units, 8 channels, 'add') feeding a 10-channel spatial tower
    (1x1 project, 3 x (max-pool, 3x3 conv), 4 ResBlocks at 32 channels, fc) and the build-order Transformer ->
    [rows, 544].  Same kernels as the observation encoder: tcgen05 GEMM / implicit-GEMM convolutions, as_scatter_connection."""

    FC_FIELDS = (('enemy_unit_counts_bow', 260, 64), ('enemy_unit_type_bool', 260, 64), ('enemy_agent_statistics', 10, 64),
                 ('enemy_upgrades', 90, 32), ('cumulative_stat', 167, 128))
    ONE_HOT_FIELDS = (('unit_alliance', 2, 16), ('unit_type', 260, 48))

    def __init__(self):
        super().__init__()
        self.encode_modules = nn.ModuleDict()
        order = ('enemy_unit_counts_bow', 'enemy_unit_type_bool', 'enemy_agent_statistics', 'enemy_upgrades', 'unit_alliance',
                 'unit_type', 'cumulative_stat')                       # yaml order (registration order = state_dict order)
        fc = {n: (i, o) for n, i, o in self.FC_FIELDS}
        oh = {n: (v, d) for n, v, d in self.ONE_HOT_FIELDS}
        for name in order:
            if name in fc:
                self.encode_modules[name] = FC(fc[name][0], fc[name][1], relu=True)
            else:
                self.encode_modules[name] = nn.Embedding(oh[name][0], oh[name][1])
        self.encode_modules['beginning_order'] = BeginningBuildOrderEncoder()
        self.scatter_project = FC(64, 8, relu=True)
        self.project = Conv(10, 16, 1, 0, relu=True, nhwc=True)
        self.downsample = nn.Sequential(nn.Identity(), Conv(16, 16, 3, 1, True, nhwc=True), nn.Identity(),
                                        Conv(16, 32, 3, 1, True, nhwc=True), nn.Identity(), Conv(32, 32, 3, 1, True, nhwc=True))
        self.res = nn.ModuleList([ResBlock(32) for _ in range(4)])
        self.spatial_fc = FC(32 * (C.SPATIAL_Y // 8) * (C.SPATIAL_X // 8), 128, relu=True)

    def forward(self, x):
        fc_emb = [self.encode_modules[n](x[n]) for n, _, _ in self.FC_FIELDS]
        unit = [ops.embedding(x[n], self.encode_modules[n].weight) for n, _, _ in self.ONE_HOT_FIELDS]   # [R*U, d]
        R, U = x['unit_type'].shape
        proj = self.scatter_project(torch.cat(unit, dim=-1)).view(R, U, -1)
        H, W = C.SPATIAL_Y, C.SPATIAL_X
        smap = ops.scatter_connection(proj, x['unit_x'], x['unit_y'], x['total_unit_count'], H, W, 'add')
        planes = torch.cat([smap, x['own_units_spatial'].reshape(R, 1, H, W).float(),
                            x['enemy_units_spatial'].reshape(R, 1, H, W).float()], dim=1)
        m = self.project(spatial.to_pnhwc(planes, 16))
        for i in (1, 3, 5):
            m = self.downsample[i](spatial.max_pool2(m))
        for block in self.res:
            m = block(m)
        sp = spatial.spatial_fc(m, self.spatial_fc[0].weight, self.spatial_fc[0].bias, 32, H // 8, W // 8)
        bo = self.encode_modules['beginning_order'](x['beginning_order'], x['bo_location'])
        return torch.cat(fc_emb + [sp, bo], dim=-1)


class ValueBaseline(nn.Module):
    """model/value.py:9-39."""

    def __init__(self, param, use_value_feature=False):
        super().__init__()
        param = dict(param)
        if use_value_feature:                    # value.py:20-21: + value-encoder output (544) + baseline feature (512)
            param['input_dim'] = param['input_dim'] + 1056
        self.project = FC(param['input_dim'], param['res_dim'], relu=True)
        self.res = nn.Sequential(*[ResFCBlock2(param['res_dim']) for _ in range(param['res_num'])])
        self.value_fc = FC(param['res_dim'], 1)
        nn.init.xavier_uniform_(self.value_fc[0].weight, 0.1)
        nn.init.constant_(self.value_fc[0].bias, 0.0)
        self.atan = bool(param['atan'])

    def forward(self, x):
        v = self.value_fc(self.res(self.project(x))).squeeze(1)
        return ops.atan_squash(v) if self.atan else v


# ----------------------------------------------------------------------------------------------
class Model(nn.Module):
    """Drop-in for `distar.agent.default.model.Model` (model.py:22-189)."""

    def __init__(self, cfg=None, use_value_network=False, temperature=None):
        super().__init__()
        self.whole_cfg = C.AttrDict(C.deep_merge(C.DEFAULT_MODEL_CONFIG, cfg or {}))
        if temperature is not None:
            self.whole_cfg.model.temperature = temperature
        self.cfg = self.whole_cfg.model
        self._use_value_feature = bool(self.whole_cfg.get('learner', {}).get('use_value_feature', False))
        self.encoder = Encoder()
        extra_units = bool(self.whole_cfg.get('agent', {}).get('extra_units', False))
        self.policy = Policy(self.whole_cfg.common.type, extra_units)
        if use_value_network:
            if self._use_value_feature:                      # model.py:33-34
                self.value_encoder = ValueEncoder()
            self.value_networks = nn.ModuleDict()
            for k, v in self.cfg.value.items():
                if k in self.cfg.enable_baselines:
                    self.value_networks[v['name']] = ValueBaseline(v['param'], self._use_value_feature)
        self.only_update_baseline = self.cfg.get('only_update_baseline', False)
        lc = self.cfg.encoder.core_lstm
        self.core_lstm = StackedLNLSTM(lc.input_size, lc.hidden_size, lc.num_layers)

    # -- helpers -------------------------------------------------------------------------------
    def _noise(self, kwargs, rows, n_entity, device):
        noise = kwargs.get('gumbel_noise')
        if noise is None:      # fresh Gumbel(0,1) draws == sampling from softmax (multinomial in the reference)
            noise = ops.fresh_gumbel(rows, n_entity, device)
        return noise

    def _encode_core(self, spatial_info, entity_info, scalar_info, entity_num, hidden_state, defer_checks=False):
        if not defer_checks:
            ops.raise_if_negative_index()
        lstm_input, ctx, base, ent, skip = self.encoder(spatial_info, entity_info, scalar_info, entity_num)
        out, state = self.core_lstm(lstm_input.unsqueeze(0), hidden_state)
        return out.squeeze(0), state, ctx, base, ent, skip

    # -- reference entry points ---------------------------------------------------------------
    def forward(self, spatial_info, entity_info, scalar_info, entity_num, hidden_state, **kwargs):
        out, state, ctx, _, ent, skip = self._encode_core(spatial_info, entity_info, scalar_info, entity_num,
                                                          hidden_state)
        noise = self._noise(kwargs, out.shape[0], ent.shape[1], out.device)
        action, sun, _, _ = self.policy(out, ent, skip, ctx, entity_num, self.cfg.temperature, noise)
        return action, sun, state

    def compute_logp_action(self, spatial_info, entity_info, scalar_info, entity_num, hidden_state, **kwargs):
        """model.py:82-116.  `defer_checks` / `defer_trim` (used by learner.ActorStep) keep the call free of host reads
        so that it can be captured into a CUDA graph: the selected-units tensors then keep all 64 steps and the
        reference's step count comes back as the device scalar `selected_units_steps`."""
        out, state, ctx, _, ent, skip = self._encode_core(spatial_info, entity_info, scalar_info, entity_num,
                                                          hidden_state, kwargs.get('defer_checks', False))
        noise = self._noise(kwargs, out.shape[0], ent.shape[1], out.device)
        ops.SU_DEFER_TRIM = [] if kwargs.get('defer_trim', False) else None
        try:
            action, sun, logit, extra = self.policy(out, ent, skip, ctx, entity_num, self.cfg.temperature, noise)
            steps = ops.SU_DEFER_TRIM[0] if ops.SU_DEFER_TRIM else None
        finally:
            ops.SU_DEFER_TRIM = None
        logp = {k: ops.categorical_log_prob(logit[k], action[k]) for k in action}
        res = {'action_info': action, 'action_logp': logp, 'selected_units_num': sun, 'entity_num': entity_num,
               'hidden_state': state, 'logit': logit, 'extra_units': extra}
        if steps is not None:
            res['selected_units_steps'] = steps
        return res

    def compute_teacher_logit(self, spatial_info, entity_info, scalar_info, entity_num, hidden_state,
                              selected_units_num, action_info, **kwargs):
        out, state, ctx, _, ent, skip = self._encode_core(spatial_info, entity_info, scalar_info, entity_num,
                                                          hidden_state)
        _, sun, logit = self.policy.train_forward(out, ent, skip, ctx, entity_num, action_info, selected_units_num,
                                                  self.cfg.temperature)
        return {'logit': logit, 'hidden_state': state, 'entity_num': entity_num, 'selected_units_num': sun}

    def rl_learner_forward(self, spatial_info, entity_info, scalar_info, entity_num, hidden_state, action_info,
                           selected_units_num, behaviour_logp, teacher_logit, mask, reward, step, batch_size,
                           unroll_len, **kwargs):
        B, T = batch_size, unroll_len
        if not kwargs.get('defer_checks', False):
            ops.raise_if_negative_index()
        flat_action = {k: v.flatten(0, 1) for k, v in action_info.items()}
        lstm_input, ctx, base, ent, skip = self.encoder(spatial_info, entity_info, scalar_info, entity_num,
                                                        head_rows=entity_num.shape[0] - B)
        hs = hidden_state[0][0].shape[-1]
        state0 = [tuple(s.view(-1, B, hs)[0] for s in layer) for layer in hidden_state]     # t=0 state only (A12)
        out, _ = self.core_lstm(lstm_input.view(-1, B, lstm_input.shape[-1]), state0)
        out = out.reshape(-1, out.shape[-1])
        critic_input = out.detach() if self.only_update_baseline else out
        br = side_stream(out.device)
        with br:                                                       # value towers next to the policy heads
            if self._use_value_feature:                                # model.py:138-144
                bf = base.detach() if self.only_update_baseline else base
                critic_input = torch.cat([critic_input, self.value_encoder(kwargs['value_feature']), bf], dim=1)
            values = {k: v(critic_input).view(T + 1, B) for k, v in self.value_networks.items()}
        ops.SU_STEPS_HINT = kwargs.get('selected_units_max')       # optional host-side int: avoids a device sync
        try:
            _, _, logits = self.policy.train_forward(out[:-B], ent, skip, ctx[:-B], entity_num[:-B], flat_action,
                                                     selected_units_num.flatten(0, 1), self.cfg.temperature)
        finally:
            ops.SU_STEPS_HINT = None
        br.join(list(values.values()))
        logits = {k: v.view(T, B, *v.shape[1:]) for k, v in logits.items()}
        su = logits['selected_units']
        if su.shape[2] < C.MAX_SELECTED_UNITS_NUM:                                           # model.py:156-158
            logits['selected_units'] = torch.nn.functional.pad(
                su, (0, 0, 0, C.MAX_SELECTED_UNITS_NUM - su.shape[2]), 'constant', -1e9)
        return {'unroll_len': T, 'batch_size': B, 'selected_units_num': selected_units_num,
                'target_logit': logits, 'value': values, 'action_log_prob': behaviour_logp,
                'teacher_logit': teacher_logit, 'mask': mask, 'action': action_info, 'reward': reward, 'step': step}

    def sl_train(self, spatial_info, entity_info, scalar_info, entity_num, selected_units_num, traj_lens,
                 hidden_state, action_info, **kwargs):
        """model.py:170-189 (same kernels, batch-major trajectories)."""
        batch_size = len(traj_lens)
        lstm_input, ctx, base, ent, skip = self.encoder(spatial_info, entity_info, scalar_info, entity_num)
        x = lstm_input.view(-1, lstm_input.shape[0] // batch_size, lstm_input.shape[-1]).permute(1, 0, 2)
        out, state = self.core_lstm(x.contiguous(), hidden_state)
        out = out.permute(1, 0, 2).contiguous().view(-1, out.shape[-1])
        action, sun, logits = self.policy.train_forward(out, ent, skip, ctx, entity_num, action_info,
                                                        selected_units_num, self.cfg.temperature)
        return logits, action, state
