"""TEST INFRASTRUCTURE — CPU fp32 restatement of the reference's policy/value forward, RL loss and supervised loss.

Parity pinned: every function here is checked against outputs of the executed reference (fixtures under tests/golden/,
written by oracle/make_golden.py from the imported reference; the reference holds no golden vectors of its own).

This file is the *checker* for the CUDA path: only tests/, __graft_entry__.smoke() and the
`cpu_baseline` / `--impl reference` legs of bench.py may import it. It is never the thing
shipped or measured as product. It is a functional re-derivation (weights come in as a flat
`state_dict` with the reference's key names); every function cites the
reference lines it follows. Pinned against the executed reference by oracle/make_golden.py
-> tests/golden/*.pt and tests/test_oracle_golden.py (and, in the build container, directly
against the imported reference by tests/test_oracle_vs_reference.py, marker `reference`).

Floating point only: plain torch ops with autograd, so gradients of the oracle come for free.
"""
import math

import torch
import torch.nn.functional as F

# ---------------------------------------------------------------------------------------------
# constants of the path (lib/features.py:31-38; actor_critic_default_config.yaml)
H, W = 152, 160
MAX_SU = 64

ENTITY_FIELDS = [  # yaml order, (name, arc, num) -- entity_encoder.py:60-78
    ('unit_type', 'one_hot', 260), ('alliance', 'one_hot', 5), ('cargo_space_taken', 'one_hot', 9),
    ('build_progress', 'unsqueeze', 1), ('health_ratio', 'unsqueeze', 1), ('shield_ratio', 'unsqueeze', 1),
    ('energy_ratio', 'unsqueeze', 1), ('display_type', 'one_hot', 5), ('x', 'binary', 11), ('y', 'binary', 11),
    ('cloak', 'one_hot', 5), ('is_blip', 'one_hot', 2), ('is_powered', 'one_hot', 2),
    ('mineral_contents', 'unsqueeze', 1), ('vespene_contents', 'unsqueeze', 1), ('cargo_space_max', 'one_hot', 9),
    ('assigned_harvesters', 'one_hot', 24), ('weapon_cooldown', 'one_hot', 32), ('order_length', 'one_hot', 9),
    ('order_id_0', 'one_hot', 327), ('order_id_1', 'one_hot', 49), ('is_hallucination', 'one_hot', 2),
    ('buff_id_0', 'one_hot', 50), ('buff_id_1', 'one_hot', 50), ('addon_unit_type', 'one_hot', 9),
    ('is_active', 'one_hot', 2), ('order_progress_0', 'unsqueeze', 1), ('order_progress_1', 'unsqueeze', 1),
    ('order_id_2', 'one_hot', 49), ('order_id_3', 'one_hot', 49), ('is_in_cargo', 'one_hot', 2),
    ('attack_upgrade_level', 'one_hot', 4), ('armor_upgrade_level', 'one_hot', 4),
    ('shield_upgrade_level', 'one_hot', 4), ('last_selected_units', 'one_hot', 2),
    ('last_targeted_unit', 'one_hot', 2)]
SPATIAL_ONE_HOT = [('visibility_map', 4), ('creep', 2), ('player_relative', 5), ('alerts', 2), ('pathable', 2),
                   ('buildable', 2)]
EFFECTS = ['effect_PsiStorm', 'effect_NukeDot', 'effect_LiberatorDefenderZone', 'effect_BlindingCloud',
           'effect_CorrosiveBile', 'effect_LurkerSpines']
SCALAR_FIELDS = [  # (name, arc, num_embeddings, context, baseline) -- yaml order
    ('agent_statistics', 'fc', 0, False, True), ('home_race', 'one_hot', 5, True, False),
    ('away_race', 'one_hot', 5, True, False), ('upgrades', 'fc', 0, False, True), ('time', 'time', 0, False, False),
    ('unit_counts_bow', 'fc', 0, False, True), ('last_delay', 'one_hot', 128, False, False),
    ('last_queued', 'one_hot', 2, False, False), ('last_action_type', 'one_hot', 327, False, False),
    ('cumulative_stat', 'fc', 0, True, True), ('beginning_order', 'bo', 0, True, True),
    ('unit_type_bool', 'fc', 0, True, False), ('enemy_unit_type_bool', 'fc', 0, True, False),
    ('unit_order_type', 'fc', 0, True, False)]
HEADS = ['action_type', 'delay', 'queued', 'selected_units', 'target_unit', 'target_location']


def lin(sd, p, x, relu=False):
    """fc_block: Linear (+ReLU) -- ctools/torch_utils/network/nn_module.py:231-270 (`<p>.0.*`)."""
    y = F.linear(x, sd[p + '.0.weight'], sd[p + '.0.bias'])
    return F.relu(y) if relu else y


def conv(sd, p, x, relu=False, pad=0):
    """conv2d_block without norm -- nn_module.py:119-174."""
    y = F.conv2d(x, sd[p + '.0.weight'], sd[p + '.0.bias'], padding=pad)
    return F.relu(y) if relu else y


def ln(sd, p, x):
    return F.layer_norm(x, (x.shape[-1],), sd[p + '.weight'], sd[p + '.bias'], 1e-5)


def seq_mask(lengths, max_len):
    """rnn.py:20-36: mask[b, j] = j < lengths[b]."""
    return torch.arange(max_len, device=lengths.device).unsqueeze(0) < lengths.reshape(-1, 1)


def bits_msb_first(v, nbits):
    """entity_encoder.py:12-17 binary table row v: nbits bits, most significant first."""
    shifts = torch.arange(nbits - 1, -1, -1, device=v.device)
    return ((v.long().unsqueeze(-1) >> shifts) & 1).float()


# ---------------------------------------------------------------------------------------------
def transformer(sd, p, x, mask, n_layer, heads, head_dim, post_ln):
    """module_utils.py:154-199 (+ layer :114-151, attention :71-111). `mask` = key-padding [R,N] or None."""
    x = lin(sd, p + '.embedding', x, relu=True)
    R, N, _ = x.shape
    for i in range(n_layer):
        q = '%s.layers.%d' % (p, i)

        def attn(z):
            qkv = lin(sd, q + '.attention.attention_pre', z)
            a, b, c = qkv.chunk(3, dim=2)
            a = a.view(R, N, heads, head_dim).permute(0, 2, 1, 3)
            b = b.view(R, N, heads, head_dim).permute(0, 2, 3, 1)
            c = c.view(R, N, heads, head_dim).permute(0, 2, 1, 3)
            s = torch.matmul(a, b) / math.sqrt(head_dim)
            if mask is not None:
                s = s.masked_fill(~mask.view(R, 1, 1, N), -1e9)
            o = torch.matmul(F.softmax(s, dim=-1), c).permute(0, 2, 1, 3).reshape(R, N, heads * head_dim)
            return lin(sd, q + '.attention.project', o)

        def mlp(z):  # ReLU on both layers (module_utils.py:121-123)
            return lin(sd, q + '.mlp.1', lin(sd, q + '.mlp.0', z, relu=True), relu=True)
        if post_ln:
            x = ln(sd, q + '.layernorm1', x + attn(x))
            x = ln(sd, q + '.layernorm2', x + mlp(x))
        else:
            x = x + attn(ln(sd, q + '.layernorm1', x))
            x = x + mlp(ln(sd, q + '.layernorm2', x))
    return x


def entity_features(entity_info):
    """entity_encoder.py:60-78: 997 columns, one-hot indices clamped to num-1, negatives raise."""
    cols = []
    for name, arc, num in ENTITY_FIELDS:
        v = entity_info[name]
        if arc == 'one_hot':
            if (v < 0).any():
                raise RuntimeError('negative one-hot index in %s' % name)
            cols.append(F.one_hot(v.long().clamp(max=num - 1), num).float())
        elif arc == 'binary':
            cols.append(bits_msb_first(v, num))
        else:
            cols.append(v.float().unsqueeze(-1))
    return torch.cat(cols, dim=-1)


def entity_encoder(sd, entity_info, entity_num):
    """entity_encoder.py:59-96."""
    p = 'encoder.entity_encoder'
    x = entity_features(entity_info)
    mask = seq_mask(entity_num, x.shape[1])
    x = transformer(sd, p + '.transformer', x, mask, 3, 2, 128, post_ln=True)
    x = F.relu(x)                                    # in-place ReLU also feeds the mean (App. A1)
    entity_embeddings = lin(sd, p + '.entity_fc', x, relu=True)
    pooled = (x * mask.unsqueeze(2)).sum(dim=1) / entity_num.unsqueeze(-1)
    embedded_entity = lin(sd, p + '.embed_fc', pooled, relu=True)
    return entity_embeddings, embedded_entity, mask


def scatter_connection(project_embeddings, x, y, scatter_type='add'):
    """module_utils.py:11-34. project_embeddings [R,N,D]; x,y [R,N] -> [R,D,H,W].
    'add': fp32 sum in ascending entity order per cell (SURVEY §8c bit-exact subset)."""
    R, N, D = project_embeddings.shape
    cell = y.long().clamp(0, H - 1) * W + x.long().clamp(0, W - 1)
    flat = torch.zeros(R, H * W, D, dtype=project_embeddings.dtype, device=project_embeddings.device)
    idx = cell.unsqueeze(-1).expand(R, N, D)
    if scatter_type == 'add':
        flat.scatter_add_(1, idx, project_embeddings)
    elif scatter_type == 'cover':
        flat.scatter_(1, idx, project_embeddings)
    else:
        raise NotImplementedError(scatter_type)
    return flat.view(R, H, W, D).permute(0, 3, 1, 2)


def spatial_input(spatial_info, scatter_map):
    """spatial_encoder.py:52-71: 56-channel input incl. the effect-plane (0,0) quirk (App. A7/A8)."""
    R = scatter_map.shape[0]
    planes = [spatial_info['height_map'].float().unsqueeze(1) / 256]
    for name, num in SPATIAL_ONE_HOT:
        planes.append(F.one_hot(spatial_info[name].long(), num).permute(0, 3, 1, 2).float())
    for name in EFFECTS:
        e = torch.zeros(R, H * W, device=scatter_map.device)
        e.scatter_(1, spatial_info[name].long(), 1.0)
        planes.append(e.view(R, 1, H, W))
    planes.append(scatter_map)
    return torch.cat(planes, dim=1)


def spatial_encoder(sd, spatial_info, scatter_map):
    """spatial_encoder.py:51-90 (+ ResBlock res_block.py:61-65)."""
    p = 'encoder.spatial_encoder'
    x = conv(sd, p + '.project', spatial_input(spatial_info, scatter_map), relu=True)
    skip = []
    for i in range(3):
        skip.append(x)
        x = conv(sd, '%s.downsample.%d' % (p, i), F.max_pool2d(x, 2, 2), relu=True, pad=1)
    for i in range(4):
        skip.append(x)
        r = conv(sd, '%s.res.%d.conv1' % (p, i), x, relu=True, pad=1)
        r = conv(sd, '%s.res.%d.conv2' % (p, i), r, pad=1)
        x = F.relu(r + x)
    return lin(sd, p + '.fc', x.reshape(x.shape[0], -1), relu=True), skip


def build_order_encoder(sd, q, order, bo_location):
    """BeginningBuildOrderEncoder (scalar_encoder.py:19-53): pre-LN transformer over 20 tokens, mean, fc."""
    Rn = order.shape[0]
    loc = bo_location.long()
    seq = torch.eye(20, device=order.device).unsqueeze(0).expand(Rn, 20, 20)
    tok = torch.cat([F.one_hot(order.long(), 174).float(), seq, bits_msb_first(loc % W, 10), bits_msb_first(loc // W, 10)], dim=2)
    t = transformer(sd, q + '.transformer', tok, None, 3, 2, 8, post_ln=False)
    return lin(sd, q + '.embedd_fc', t.mean(dim=1), relu=True)


def scalar_encoder(sd, scalar_info):
    """scalar_encoder.py:99-132 (+ time :91-97, build-order transformer :19-53, pre-LN, App. A3)."""
    p = 'encoder.scalar_encoder'
    emb, ctx, base = [], [], []
    for name, arc, num, is_ctx, is_base in SCALAR_FIELDS:
        if arc == 'time':
            continue
        v = scalar_info[name]
        if arc == 'one_hot':
            e = F.relu(F.embedding(v.long().clamp(max=num - 1), sd['%s.encode_modules.%s.weight' % (p, name)]))
        elif arc == 'bo':
            e = build_order_encoder(sd, '%s.encode_modules.beginning_order' % p, v, scalar_info['bo_location'])
        else:
            e = lin(sd, '%s.encode_modules.%s' % (p, name), v.float(), relu=True)
        emb.append(e)
        if is_ctx:
            ctx.append(e)
        if is_base:
            base.append(e)
    pos = sd[p + '.position_array']
    t = scalar_info['time'].float().unsqueeze(1)
    te = torch.zeros(t.shape[0], pos.shape[0], device=t.device)
    te[:, 0::2] = torch.sin(t * pos[0::2])
    te[:, 1::2] = torch.cos(t * pos[1::2])
    emb.append(te)
    return torch.cat(emb, 1), torch.cat(ctx, 1), torch.cat(base, 1)


def encoder(sd, spatial_info, entity_info, scalar_info, entity_num):
    """encoder.py:28-45."""
    embedded_scalar, scalar_context, baseline_feature = scalar_encoder(sd, scalar_info)
    entity_embeddings, embedded_entity, mask = entity_encoder(sd, entity_info, entity_num)
    proj = lin(sd, 'encoder.scatter_project', entity_embeddings, relu=True) * mask.unsqueeze(2)
    smap = scatter_connection(proj, entity_info['x'], entity_info['y'], 'add')
    embedded_spatial, map_skip = spatial_encoder(sd, spatial_info, smap)
    lstm_input = torch.cat([embedded_scalar, embedded_entity, embedded_spatial], dim=-1)
    return lstm_input, scalar_context, baseline_feature, entity_embeddings, map_skip


# ---------------------------------------------------------------------------------------------
def lnlstm_cell(sd, p, x, h, c):
    """lstm.py:138-153: gates i,f,g,o; carried cell state is the *normalised* one (App. A11)."""
    gi = ln(sd, p + '.layernorm_i', x @ sd[p + '.weight_ih'].t())
    gh = ln(sd, p + '.layernorm_h', h @ sd[p + '.weight_hh'].t())
    i, f, g, o = (gi + gh).chunk(4, 1)
    c2 = ln(sd, p + '.layernorm_c', torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g))
    h2 = torch.sigmoid(o) * torch.tanh(c2)
    return h2, c2


def lnlstm(sd, p, x, state, n_layer):
    """lstm.py:215-234,156-167: x [T,B,I]; layer-major then time."""
    out_state = []
    for l in range(n_layer):
        h, c = state[l]
        ys = []
        for t in range(x.shape[0]):
            h, c = lnlstm_cell(sd, '%s.layers.%d.cell' % (p, l), x[t], h, c)
            ys.append(h)
        x = torch.stack(ys)
        out_state.append((h, c))
    return x, out_state


# ---------------------------------------------------------------------------------------------
def gumbel_pick(logits, noise):
    """Sampling contract: argmax_k(fp32 logit_k + g_k), lowest index on ties."""
    return torch.argmax(logits.float() + noise, dim=-1)


def glu(sd, p, x, context):
    """module_utils.py:519-524."""
    return lin(sd, p + '.layer2', torch.sigmoid(lin(sd, p + '.layer1', context)) * x)


def action_type_head(sd, lstm_output, scalar_context, action_type=None, noise=None, temperature=1.0,
                     race_mask=None):
    """action_type_head.py:48-67 (ResFCBlock res_block.py:68-108: fc-LN-ReLU, fc-LN, +res, ReLU)."""
    p = 'policy.action_type_head'
    x = lin(sd, p + '.project', lstm_output, relu=True)
    for i in range(2):
        q = '%s.res.%d' % (p, i)
        r = F.relu(ln(sd, q + '.fc1.1', F.linear(x, sd[q + '.fc1.0.weight'], sd[q + '.fc1.0.bias'])))
        r = ln(sd, q + '.fc2.1', F.linear(r, sd[q + '.fc2.0.weight'], sd[q + '.fc2.0.bias']))
        x = F.relu(r + x)
    logits = glu(sd, p + '.action_fc', x, scalar_context) / temperature
    if race_mask is not None:
        logits = logits.masked_fill(~race_mask.unsqueeze(0), -1e9)
    if action_type is None:
        action_type = gumbel_pick(logits, noise)
    one_hot = F.one_hot(action_type.long(), 327).float()
    e1 = lin(sd, p + '.action_map_fc2', lin(sd, p + '.action_map_fc1', one_hot, relu=True))
    emb = glu(sd, p + '.glu1', e1, scalar_context) + glu(sd, p + '.glu2', lstm_output, scalar_context)
    return logits, action_type, emb


def small_head(sd, p, emb, dim, action=None, noise=None, temperature=None):
    """DelayHead / QueuedHead -- action_arg_head.py:41-53,73-86 (delay has no temperature)."""
    x = lin(sd, p + '.fc3', lin(sd, p + '.fc2', lin(sd, p + '.fc1', emb, relu=True), relu=True))
    if temperature is not None:
        x = x / temperature
    if action is None:
        action = gumbel_pick(x, noise)
    one_hot = F.one_hot(action.long(), dim).float()
    return x, action, emb + lin(sd, p + '.embed_fc2', lin(sd, p + '.embed_fc1', one_hot, relu=True))


def su_keys(sd, entity_embeddings, entity_num):
    """action_arg_head.py:118-143: key rows, end-embedding row at index entity_num, logits mask."""
    p = 'policy.selected_units_head'
    R, N, _ = entity_embeddings.shape
    key = torch.cat([lin(sd, p + '.key_fc', entity_embeddings), torch.zeros(R, 1, 32, device=entity_embeddings.device)], 1)
    is_end = torch.arange(N + 1, device=key.device).unsqueeze(0) == entity_num.unsqueeze(1)
    key = torch.where(is_end.unsqueeze(2), sd[p + '.end_embedding'].view(1, 1, 32), key)
    return key, seq_mask(entity_num + 1, N + 1)


def su_head_train(sd, ae0, entity_embeddings, entity_num, selected_units, selected_units_num):
    """Teacher-forced pointer loop -- action_arg_head.py:168-216 ('selected_units_num' reduce type)."""
    p = 'policy.selected_units_head'
    key, mask0 = su_keys(sd, entity_embeddings, entity_num)
    R = ae0.shape[0]
    ar = torch.arange(R, device=ae0.device)
    S = max(int(selected_units_num.max()), 1)
    h = torch.zeros(R, 32, device=ae0.device)
    c = torch.zeros(R, 32, device=ae0.device)
    lm = mask0.clone()
    lm[ar, entity_num] = False
    one_hot = torch.zeros(R, key.shape[1], device=ae0.device)
    end_flag = torch.zeros(R, dtype=torch.bool, device=ae0.device)
    ae = ae0
    logits = []
    for i in range(S):
        if i > 0:
            lm = lm.clone()
            if i == 1:
                lm[ar, entity_num] = True
            lm[ar, selected_units[:, i - 1]] = False
        q = lin(sd, p + '.query_fc2', lin(sd, p + '.query_fc1', ae, relu=True))
        h, c = lnlstm_cell(sd, p + '.lstm.layers.0.cell', q, h, c)
        logits.append(torch.einsum('rk,rnk->rn', h, key).masked_fill(~lm, -1e9))
        end_flag = end_flag | (selected_units[:, i] == entity_num)
        pick = F.one_hot(selected_units[:, i], key.shape[1]).to(one_hot.dtype) * (~end_flag).unsqueeze(1)
        one_hot = torch.maximum(one_hot, pick)
        su_emb = torch.einsum('rn,rnk->rk', one_hot, key)
        cnt = one_hot.sum(dim=1, keepdim=True)
        su_emb = torch.where((selected_units_num != 0).unsqueeze(1), su_emb / cnt, su_emb)
        ae = ae0 + lin(sd, p + '.embed_fc2', lin(sd, p + '.embed_fc1', su_emb, relu=True))
    return torch.stack(logits, dim=1), ae


def su_head_sample(sd, ae0, entity_embeddings, entity_num, su_mask, noise, temperature=1.0, extra_units=True):
    """Sampling pointer loop -- action_arg_head.py:262-313 (noise [64,R,N+1])."""
    p = 'policy.selected_units_head'
    key, lm = su_keys(sd, entity_embeddings, entity_num)
    R = ae0.shape[0]
    ar = torch.arange(R, device=ae0.device)
    h = torch.zeros(R, 32, device=ae0.device)
    c = torch.zeros(R, 32, device=ae0.device)
    lm = lm.clone()
    lm[ar, entity_num] = False
    num = torch.full((R,), MAX_SU, dtype=torch.long, device=ae0.device)
    end_flag = ~su_mask.clone()
    num[~su_mask] = 0
    one_hot = torch.zeros(R, key.shape[1], device=ae0.device)
    ae = ae0
    logits, units = [], []
    step_logits = None
    result = None
    for i in range(MAX_SU):
        if i > 0:
            if i == 1:
                lm[ar, entity_num] = True
            lm[ar, result] = False
        q = lin(sd, p + '.query_fc2', lin(sd, p + '.query_fc1', ae, relu=True))
        h, c = lnlstm_cell(sd, p + '.lstm.layers.0.cell', q, h, c)
        step_logits = torch.einsum('rk,rnk->rn', h, key).masked_fill(~lm, -1e9) / temperature
        result = gumbel_pick(step_logits, noise[i])
        hit = result == entity_num
        num[hit & ~end_flag] = i + 1
        end_flag = end_flag | hit
        units.append(result)
        logits.append(step_logits)
        one_hot[ar[~end_flag], result[~end_flag]] = 1
        su_emb = torch.einsum('rn,rnk->rk', one_hot, key)
        cnt = one_hot.sum(dim=1, keepdim=True)
        su_emb = torch.where(cnt != 0, su_emb / cnt.clamp(min=1), su_emb)
        ae = ae0 + lin(sd, p + '.embed_fc2', lin(sd, p + '.embed_fc1', su_emb, relu=True))
        if bool(end_flag.all()):
            break
    extra = torch.zeros(R, key.shape[1], device=ae0.device)
    if extra_units:
        end_logit = step_logits[ar, entity_num]
        extra = ((step_logits > end_logit.unsqueeze(1)) & ~end_flag.unsqueeze(1)).float()
    return torch.stack(logits, 1), torch.stack(units, 1), ae, num, extra


def target_unit_head(sd, emb, entity_embeddings, entity_num, target_unit=None, noise=None, temperature=1.0):
    """action_arg_head.py:343-363."""
    p = 'policy.target_unit_head'
    key = lin(sd, p + '.key_fc', entity_embeddings)
    q = lin(sd, p + '.query_fc2', lin(sd, p + '.query_fc1', emb, relu=True))
    logits = torch.einsum('rk,rnk->rn', q, key)
    logits = logits.masked_fill(~seq_mask(entity_num, key.shape[1]), -1e9) / temperature
    if target_unit is None:
        target_unit = gumbel_pick(logits, noise)
    return logits, target_unit


def location_head(sd, emb, map_skip, location=None, noise=None, temperature=1.0):
    """action_arg_head.py:417-450 (+ GatedResBlock module_utils.py:224-231), bilinear variant."""
    p = 'policy.location_head'
    R = emb.shape[0]
    x = lin(sd, p + '.project_embed', emb, relu=True).reshape(R, 4, H // 8, W // 8)
    x = conv(sd, p + '.conv1', F.relu(torch.cat([x, map_skip[-1]], dim=1)), relu=True)
    for i in range(4):
        x = x + map_skip[len(map_skip) - i - 1]
        q = '%s.res.%d' % (p, i)
        r = conv(sd, q + '.conv2', conv(sd, q + '.conv1', x, relu=True, pad=1), pad=1)
        g = x
        for j in range(4):
            g = conv(sd, '%s.GateWeightG.%d' % (q, j), g, relu=(j < 3))
        x = F.relu(torch.tanh(r * torch.sigmoid(g)) * sd[q + '.UpdateSP'] + x)
    for i in range(3):
        x = F.interpolate(x, scale_factor=2., mode='bilinear')
        x = conv(sd, '%s.upsample.%d' % (p, i), x, relu=(i < 2), pad=1)
    logits = x.reshape(R, -1) / temperature
    if location is None:
        location = gumbel_pick(logits, noise)
    return logits, location


def value_baseline(sd, field, x, atan):
    """value.py:31-39 (+ ResFCBlock2 res_block.py:110-141)."""
    p = 'value_networks.' + field
    x = lin(sd, p + '.project', x, relu=True)
    for i in range(16):
        q = '%s.res.%d' % (p, i)
        x = ln(sd, q + '.norm', lin(sd, q + '.fc2', lin(sd, q + '.fc1', x, relu=True)) + x)
    v = lin(sd, p + '.value_fc', x).squeeze(1)
    return (2.0 / math.pi) * torch.atan((math.pi / 2.0) * v) if atan else v


# ---------------------------------------------------------------------------------------------
def value_encoder(sd, vf):
    """obs_encoder/value_encoder.py:48-74 (`learner.use_value_feature`)."""
    p = 'value_encoder'
    fc = [lin(sd, '%s.encode_modules.%s' % (p, k), vf[k].float(), relu=True)
          for k in ('enemy_unit_counts_bow', 'enemy_unit_type_bool', 'enemy_agent_statistics', 'enemy_upgrades', 'cumulative_stat')]
    unit = torch.cat([F.embedding(vf[k].long(), sd['%s.encode_modules.%s.weight' % (p, k)]) for k in ('unit_alliance', 'unit_type')], -1)
    proj = lin(sd, p + '.scatter_project', unit, relu=True)
    proj = proj * seq_mask(vf['total_unit_count'], proj.shape[1]).unsqueeze(2)
    smap = scatter_connection(proj, vf['unit_x'], vf['unit_y'], 'add')
    Rn = smap.shape[0]
    x = torch.cat([smap, vf['own_units_spatial'].reshape(Rn, 1, H, W).float(), vf['enemy_units_spatial'].reshape(Rn, 1, H, W).float()], 1)
    x = conv(sd, p + '.project', x, relu=True)
    for i in (1, 3, 5):
        x = conv(sd, '%s.downsample.%d' % (p, i), F.max_pool2d(x, 2, 2), relu=True, pad=1)
    for i in range(4):
        q = '%s.res.%d' % (p, i)
        x = F.relu(conv(sd, q + '.conv2', conv(sd, q + '.conv1', x, relu=True, pad=1), pad=1) + x)
    sp = lin(sd, p + '.spatial_fc', x.reshape(Rn, -1), relu=True)
    bo = build_order_encoder(sd, p + '.encode_modules.beginning_order', vf['beginning_order'], vf['bo_location'])
    return torch.cat(fc + [sp, bo], -1)


def policy_train(sd, lstm_out, entity_embeddings, map_skip, scalar_context, entity_num, action, sun,
                 temperature=1.0):
    """policy.py:50-73."""
    logit = {}
    logit['action_type'], _, emb = action_type_head(sd, lstm_out, scalar_context, action['action_type'],
                                                    temperature=temperature)
    logit['delay'], _, emb = small_head(sd, 'policy.delay_head', emb, 128, action['delay'])
    logit['queued'], _, emb = small_head(sd, 'policy.queued_head', emb, 2, action['queued'], temperature=temperature)
    su_logits, emb = su_head_train(sd, emb, entity_embeddings, entity_num, action['selected_units'], sun)
    logit['selected_units'] = su_logits / temperature
    logit['target_unit'], _ = target_unit_head(sd, emb, entity_embeddings, entity_num, action['target_unit'],
                                               temperature=temperature)
    logit['target_location'], _ = location_head(sd, emb, map_skip, action['target_location'],
                                                temperature=temperature)
    return logit


def compute_logp_action(sd, spatial_info, entity_info, scalar_info, entity_num, hidden_state, noise,
                        su_table, temperature=1.0, extra_units=True, **kw):
    """model.py:56-74 with Gumbel-max sampling under supplied noise (App. A24)."""
    lstm_input, ctx, _, ent, skip = encoder(sd, spatial_info, entity_info, scalar_info, entity_num)
    out, state = lnlstm(sd, 'core_lstm', lstm_input.unsqueeze(0), hidden_state, 3)
    out = out.squeeze(0)
    logit, action = {}, {}
    logit['action_type'], action['action_type'], emb = action_type_head(sd, out, ctx, None, noise['action_type'],
                                                                        temperature)
    logit['delay'], action['delay'], emb = small_head(sd, 'policy.delay_head', emb, 128, None, noise['delay'])
    logit['queued'], action['queued'], emb = small_head(sd, 'policy.queued_head', emb, 2, None, noise['queued'],
                                                        temperature)
    su_mask = su_table[action['action_type']]
    logit['selected_units'], action['selected_units'], emb, sun, extra = su_head_sample(
        sd, emb, ent, entity_num, su_mask, noise['selected_units'], temperature, extra_units)
    logit['target_unit'], action['target_unit'] = target_unit_head(sd, emb, ent, entity_num, None,
                                                                   noise['target_unit'], temperature)
    logit['target_location'], action['target_location'] = location_head(sd, emb, skip, None,
                                                                        noise['target_location'], temperature)
    logp = {k: torch.distributions.Categorical(logits=logit[k]).log_prob(action[k]) for k in action}
    return {'action_info': action, 'action_logp': logp, 'selected_units_num': sun, 'entity_num': entity_num,
            'hidden_state': state, 'logit': logit, 'extra_units': extra}


def compute_teacher_logit(sd, spatial_info, entity_info, scalar_info, entity_num, hidden_state,
                          selected_units_num, action_info, temperature=1.0, **kw):
    """model.py:76-93."""
    lstm_input, ctx, _, ent, skip = encoder(sd, spatial_info, entity_info, scalar_info, entity_num)
    out, state = lnlstm(sd, 'core_lstm', lstm_input.unsqueeze(0), hidden_state, 3)
    logit = policy_train(sd, out.squeeze(0), ent, skip, ctx, entity_num, action_info, selected_units_num, temperature)
    return {'logit': logit, 'hidden_state': state, 'entity_num': entity_num, 'selected_units_num': selected_units_num}


def rl_learner_forward(sd, fields, spatial_info, entity_info, scalar_info, entity_num, hidden_state, action_info,
                       selected_units_num, behaviour_logp, teacher_logit, mask, reward, step, batch_size,
                       unroll_len, temperature=1.0, value_feature=None, **kw):
    """model.py:95-168. `fields` = enabled baselines (value_networks keys); `value_feature` (learner.use_value_feature):
    the critics see cat[lstm_output, ValueEncoder(value_feature), baseline_feature] (:141-144)."""
    B, T = batch_size, unroll_len
    flat_action = {k: v.flatten(0, 1) for k, v in action_info.items()}
    lstm_input, ctx, base_feat, ent, skip = encoder(sd, spatial_info, entity_info, scalar_info, entity_num)
    state0 = [(h.view(-1, B, h.shape[-1])[0], c.view(-1, B, c.shape[-1])[0]) for h, c in hidden_state]
    out, _ = lnlstm(sd, 'core_lstm', lstm_input.view(-1, B, lstm_input.shape[-1]), state0, 3)
    out = out.reshape(-1, out.shape[-1])
    logit = policy_train(sd, out[:-B], ent[:-B], [m[:-B] for m in skip], ctx[:-B], entity_num[:-B], flat_action,
                         selected_units_num.flatten(0, 1), temperature)
    logit = {k: v.view(T, B, *v.shape[1:]) for k, v in logit.items()}
    su = logit['selected_units']
    logit['selected_units'] = F.pad(su, (0, 0, 0, MAX_SU - su.shape[2]), 'constant', -1e9)
    critic_in = out if value_feature is None else torch.cat([out, value_encoder(sd, value_feature), base_feat], dim=1)
    value = {f: value_baseline(sd, f, critic_in, f == 'winloss').view(T + 1, B) for f in fields}
    return {'unroll_len': T, 'batch_size': B, 'selected_units_num': selected_units_num, 'target_logit': logit,
            'value': value, 'action_log_prob': behaviour_logp, 'teacher_logit': teacher_logit, 'mask': mask,
            'action': action_info, 'reward': reward, 'step': step}


# ---------------------------------------------------------------------------------------------
# RL loss (rl_loss.py:33-185, as_rl_utils.py)
def lambda_returns(rewards, gamma, values, lam):
    """as_rl_utils.py:157-218: reverse scan; `lam` may be a [T,B] tensor or a float."""
    Tn = rewards.shape[0]
    lam = lam if isinstance(lam, torch.Tensor) else torch.full_like(rewards, lam)
    boot = values[1:]
    res = [None] * Tn
    res[Tn - 1] = rewards[Tn - 1] + gamma * boot[Tn - 1]
    for t in range(Tn - 2, -1, -1):
        d = gamma * lam[t]
        res[t] = rewards[t] + d * res[t + 1] + (gamma - d) * boot[t]
    return torch.stack(res)


def vtrace_adv(rho, rewards, values, gamma=1.0, lam=1.0):
    """as_rl_utils.py:284-312 with clipped_cs = clipped_pg_rhos = rho."""
    Tn = rewards.shape[0]
    delta = rho * (rewards + gamma * values[1:] - values[:-1])
    vs = [None] * (Tn + 1)
    vs[Tn] = values[Tn]
    for t in range(Tn - 1, -1, -1):
        vs[t] = values[t] + delta[t] + gamma * lam * rho[t] * (vs[t + 1] - values[t + 1])
    vs = torch.stack(vs)
    return rho * (rewards + gamma * vs[1:] - values[:-1])


def upgo_ret(rewards, values):
    """as_rl_utils.py:265-281."""
    lam = ((rewards + values[1:]) >= values[:-1]).float()
    lam = torch.cat([lam[1:], torch.ones_like(lam[-1:])], 0)
    return lambda_returns(rewards, 1.0, values, lam)


def compute_loss(cfg, inputs, only_update_value=False, use_dapo=False):
    """rl_loss.py:33-185. `cfg` = merged learner loss config (dict). Returns the reference's info dict
    (python floats) with 'total_loss' as an autograd tensor. Mutates value[-1] in place like the reference.
    `use_dapo`: adds the DAPO term KL(successive || pi) of as_rl_utils.py:105-127 (inputs['successive_logit'])."""
    logits, values = inputs['target_logit'], inputs['value']
    mu_logp, teacher = inputs['action_log_prob'], inputs['teacher_logit']
    mask, action, reward, step = inputs['mask'], inputs['action'], inputs['reward'], inputs['step']
    flag = (reward['winloss'][-1] == 0)
    values = {f: torch.cat([v[:-1], (v[-1] * flag).unsqueeze(0)], 0) for f, v in values.items()}
    info = {}
    su_mask = mask['selected_units_mask']
    logp_all, probs_all, logp_a, rho = {}, {}, {}, {}
    for h in HEADS:
        lp = F.log_softmax(logits[h], dim=-1)
        la = lp.gather(-1, action[h].unsqueeze(-1)).squeeze(-1)
        with torch.no_grad():
            lr = la - mu_logp[h]
            if h == 'selected_units':
                lr = (lr * su_mask).sum(-1)
            rho[h] = torch.exp(lr).clamp(max=1)
        if h == 'selected_units':
            la = la.masked_fill(~su_mask, 0).sum(-1)
        logp_all[h], probs_all[h], logp_a[h] = lp, lp.exp(), la

    def head_mask(h):
        return 1.0 if h in ('action_type', 'delay') else mask['actions_mask'][h]

    lw = cfg['loss_weights']
    total_pg = 0.
    for f, v in values.items():
        fsum = 0.
        for h in HEADS:
            with torch.no_grad():
                adv = vtrace_adv(rho[h], reward[f], v, 1.0, 1.0)
            l = -adv * logp_a[h] * head_mask(h)
            if f in ('build_order', 'built_unit', 'effect'):
                l = l * mask[f + '_mask']
            l = l.mean()
            fsum = fsum + l * cfg['pg_head_weights'][h]
            info['%s/%s' % (f, h)] = l.item()
        info[f + '/total'] = float(fsum.detach())
        total_pg = total_pg + lw['pg'][f] * fsum
    total_upgo = 0.
    v = values['winloss']
    for h in HEADS:
        with torch.no_grad():
            adv = rho[h] * (upgo_ret(reward['winloss'], v) - v[:-1])
        l = (-adv * logp_a[h] * head_mask(h)).mean()
        total_upgo = total_upgo + l * cfg['upgo_head_weights'][h]
        info['upgo/' + h] = l.item()
    info['upgo/total'] = float(total_upgo.detach())
    total_upgo = total_upgo * lw['upgo']['winloss']
    total_critic = 0.
    for f, v in values.items():
        with torch.no_grad():
            ret = lambda_returns(reward[f], cfg['gammas']['baseline'][f], v, 0.8)
        l = 0.5 * (ret - v[:-1]) ** 2
        if f in ('build_order', 'built_unit', 'effect'):
            l = l * mask[f + '_mask']
        l = l.mean()
        total_critic = total_critic + lw['baseline'][f] * l
        info[f + '/td'] = l.item()
        info[f + '/reward'] = reward[f].float().mean().item()
        info[f + '/value'] = v.mean().item()
    info['battle/reward'] = reward['battle'].float().mean().item()
    total_ent = 0.
    for h in HEADS:
        e = -(probs_all[h] * logp_all[h]).sum(-1)
        if h == 'selected_units':
            e = e / (1e-9 + torch.log(mask['selected_units_logits_mask'].float().sum(-1) + 1).unsqueeze(-1))
            e = (e * su_mask).sum(-1) / (su_mask.sum(-1) + 1e-9)
        elif h == 'target_unit':
            e = e / (1e-9 + torch.log(mask['target_units_logits_mask'].float().sum(-1) + 1))
        else:
            e = e / torch.log(torch.tensor([float(logits[h].shape[-1])], device=e.device))
        e = (e * head_mask(h)).mean()
        info['entropy/' + h] = e.item()
        total_ent = total_ent - e * cfg['entropy_head_weights'][h]
    info['entropy/total'] = float(total_ent.detach())
    total_ent = total_ent * lw['entropy']
    total_kl = 0.
    at_kl = None
    for h in ['action_type', 'queued', 'delay', 'selected_units', 'target_unit', 'target_location']:
        tl = F.log_softmax(teacher[h], dim=-1)
        kl = (tl.exp() * (tl - logp_all[h])).sum(-1)
        if h == 'selected_units':
            kl = (kl * su_mask).sum(-1)
        kl = kl * head_mask(h)
        if h == 'action_type':
            at_kl = (kl * (step < cfg['kl']['action_type_kl_steps']) * mask['cum_action_mask']).mean()
            info['kl/extra_at'] = at_kl.item()
        l = kl.mean()
        total_kl = total_kl + l * cfg['kl_head_weights'][h]
        info['kl/' + h] = l.item()
    info['kl/total'] = float(total_kl.detach())
    total_kl = total_kl * lw['kl']
    at_kl = at_kl * lw['action_type_kl']
    total_dapo = 0.
    if use_dapo:
        flag = (step < cfg['dapo']['dapo_steps'])
        for h in HEADS:
            sl = F.log_softmax(inputs['successive_logit'][h], dim=-1)
            kl = (sl.exp() * (sl - logp_all[h])).sum(-1)
            if h == 'selected_units':
                kl = (kl * su_mask).sum(-1)
            if h not in ('action_type', 'delay'):
                kl = kl * mask['actions_mask'][h]
            l = (kl * flag).mean()
            total_dapo = total_dapo + l * cfg['dapo_head_weights'][h]
            info['battle/' + h] = l.item()
        info['battle/total'] = float(total_dapo.detach())
        total_dapo = total_dapo * lw['dapo']
    if only_update_value:
        info['total_loss'] = total_critic
    else:
        info['total_loss'] = total_pg + total_upgo + total_critic + total_ent + total_kl + at_kl + total_dapo
    return info


# ---------------------------------------------------------------------------------------------------------------------
# Supervised loss (reference: agent/default/sl_training/sl_loss.py; weights: default_supervised_loss.yaml)
SL_LOSS_WEIGHT = {'action_type': 30.0, 'delay': 9.0, 'queued': 1.0, 'selected_units': 4.0, 'target_unit': 4.0,
                  'target_location': 8.0}


def _sl_ce(logits, labels, smoothing):
    """sl_loss.py:16-36 (label smoothing) / torch CrossEntropyLoss(reduction='none') (:58-59)"""
    logp = torch.log_softmax(logits, dim=-1)
    nll = -logp.gather(-1, labels.unsqueeze(-1)).squeeze(-1)
    if smoothing > 0:
        return (1.0 - smoothing) * nll + smoothing * (-logp.mean(dim=-1))
    return nll


def _sl_masked_mean(loss_rows, mask):
    """sl_loss.py:134-139: sum over valid rows / number of valid rows (0 when none)"""
    n = mask.sum()
    return loss_rows.sum() / n if n > 0 else loss_rows.sum() * 0


def supervised_loss(logits, actions, actions_mask, selected_units_num, entity_num, infer_selected_units, label_smooth=False,
                    su_mask=True, loss_weight=None):
    """SupervisedLoss.compute_loss (sl_loss.py:102-121) without cross_rank_loss."""
    w = dict(SL_LOSS_WEIGHT, **(loss_weight or {}))
    sm = 0.1 if label_smooth else 0.0
    out = {}
    # single-label heads (:123-185, :246-285)
    for head in ('action_type', 'delay', 'queued', 'target_unit', 'target_location'):
        x, y, m = logits[head], actions[head].long(), actions_mask[head].float()
        out[head + '_loss'] = _sl_masked_mean(_sl_ce(x, y, sm) * m, m)
        pred = x.argmax(dim=-1)
        if head == 'action_type':
            out['action_type_acc'] = (pred == y).float().sum() / len(y)
        elif head == 'delay':
            out['delay_distance_L1'] = ((pred - y).abs() * m).sum() / (m.sum() + 1e-6)
        elif head == 'queued':
            out['queued_acc'] = ((pred - y).abs() * m).sum() / (m.sum() + 1e-6)
        elif head == 'target_unit':
            out['target_unit_acc'] = ((pred == y) * m).sum() / (m.sum() + 1e-6)
        else:
            dx, dy = pred % 160 - y % 160, pred // 160 - y // 160
            out['target_location_distance_L2'] = ((dx * dx + dy * dy).float().sqrt() * m).sum() / (m.sum() + 1e-6)
    # selected units (:187-244)
    x, m, lengths = logits['selected_units'], actions_mask['selected_units'].float(), selected_units_num.long()
    b, s, n = x.shape
    y = actions['selected_units'][:, :s].long()
    if su_mask:                                   # :190-204: the sample's other picks (end flag excluded) leave the candidates
        steps = torch.arange(s)
        picked = steps.unsqueeze(0) < (lengths - 1).clamp(min=0).unsqueeze(1)               # [b, s]
        hide = torch.zeros(b, n + 1, dtype=torch.bool)
        hide.scatter_(1, torch.where(picked, y, torch.full_like(y, n)), True)
        hide = hide[:, :n].unsqueeze(1).repeat(1, s, 1)                                     # same set for every step ...
        own = torch.zeros(b, s, n + 1, dtype=torch.bool)
        own.scatter_(2, torch.where(picked, y, torch.full_like(y, n)).unsqueeze(2), True)
        hide = hide & ~own[:, :, :n]                                                        # ... except the step's own label
        x = x.masked_fill(hide, -1e9)
    in_seq = torch.arange(s).unsqueeze(0) < lengths.unsqueeze(1)
    rows = _sl_ce(x.reshape(b * s, n), y.reshape(-1), 0.0).view(b, s)
    rows = rows.masked_fill(~in_seq, 0) * m.unsqueeze(1)
    out['selected_units_loss'] = rows.sum() / b
    out['selected_units_loss_norm'] = rows.sum() / (lengths.sum() + 1e-6)
    out['selected_units_end_flag_loss'] = rows[torch.arange(b), lengths - 1].mean()
    if infer_selected_units is not None:          # :223-241: IoU of the predicted and labelled unit sets
        p = infer_selected_units[:, :s].long()
        iou_rows = torch.zeros(b)
        for i in range(b):
            hits = (p[i] == entity_num[i]).nonzero()
            first = int(hits[0]) if len(hits) else 0
            stop = first + 1 + (s if first == 0 else 0)
            pset = set((p[i, :min(stop, s)]).tolist())
            lset = set(y[i, :int(lengths[i])].tolist())
            inter, union = len(pset & lset), len(pset | lset)
            iou_rows[i] = inter / (union + 1e-6)
        out['selected_units_iou'] = (iou_rows * m).sum() / (m.sum() + 1e-6)
    else:
        out['selected_units_iou'] = 0.
    out['total_loss'] = sum(out[h + '_loss'] * w[h] for h in ('action_type', 'delay', 'queued', 'selected_units', 'target_unit',
                                                              'target_location'))
    return out
