"""Drop-in for `distar.agent.default.rl_training.rl_loss.ReinforcementLoss` (rl_loss.py:8-199).

`compute_loss(inputs)` takes the dict `Model.rl_learner_forward` returns and gives back the reference's
log dict (same keys, python floats) with `'total_loss'` as an autograd tensor.  All arithmetic runs in the
fused CUDA kernels of csrc/loss.cu: one stats pass over the logits, one scan kernel for V-trace / UPGO /
TD(lambda), one gradient pass.  The ~65 `.item()` host syncs of the reference become a single 384-byte D2H copy.
"""
import ctypes

import torch

from . import _lib
from . import config as C

HEADS = C.HEADS
_MASKED_HEADS = {'queued': 2, 'selected_units': 3, 'target_unit': 4, 'target_location': 5}
_MASK_FIELDS = ('build_order', 'built_unit', 'effect')


class _FusedLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, owner, inputs, fields, *tensors):
        nh = len(HEADS)
        logits = [t.contiguous() for t in tensors[:nh]]
        values = [t.contiguous() for t in tensors[nh:]]
        dev = logits[0].device
        T, B = inputs['reward']['winloss'].shape
        TB = T * B
        S = logits[3].shape[2]
        a = _lib.AsLossArgs()
        a.T, a.B, a.F, a.S = T, B, len(fields), S
        a.N1, a.NE = logits[3].shape[-1], logits[4].shape[-1]
        keep = []

        def dptr(t, dtype=None):
            if dtype is torch.uint8 and t.dtype == torch.bool:
                t = t.contiguous().view(torch.uint8)          # same bytes: no conversion pass
            else:
                t = t.contiguous() if dtype is None else t.to(dtype).contiguous()
            keep.append(t)
            return t.data_ptr()

        mask = inputs['mask']
        dlogits, stats, coefs = [], [], []
        for h, name in enumerate(HEADS):
            lg = logits[h]
            rows = TB * S if h == 3 else TB
            a.head_k[h] = lg.shape[-1]
            a.logit[h] = lg.data_ptr()
            a.teacher[h] = dptr(inputs['teacher_logit'][name], torch.float32)
            a.successive[h] = dptr(inputs['successive_logit'][name], torch.float32) if owner.use_dapo else None
            a.action[h] = dptr(inputs['action'][name], torch.int64)
            a.mu_logp[h] = dptr(inputs['action_log_prob'][name], torch.float32)
            a.actions_mask[h] = dptr(mask['actions_mask'][name], torch.int64) if name in _MASKED_HEADS else None
            dl = torch.empty_like(lg)
            st = torch.empty((rows, 8), dtype=torch.float32, device=dev)
            cf = torch.empty((TB, 4), dtype=torch.float32, device=dev)      # written in full by the scan kernel
            dlogits.append(dl); stats.append(st); coefs.append(cf)
            a.dlogit[h], a.row_stats[h], a.coef[h] = dl.data_ptr(), st.data_ptr(), cf.data_ptr()
        su_step = torch.empty((TB, 8), dtype=torch.float32, device=dev)
        keep.append(su_step)
        a.su_step = su_step.data_ptr()
        a.su_mask = dptr(mask['selected_units_mask'], torch.uint8)
        a.su_logits_mask = dptr(mask['selected_units_logits_mask'], torch.uint8)
        a.tu_logits_mask = dptr(mask['target_units_logits_mask'], torch.uint8)
        a.cum_action_mask = dptr(mask['cum_action_mask'], torch.float32)
        a.step = dptr(inputs['step'], torch.float32)
        a.at_kl_steps = float(owner.action_type_kl_steps)
        a.dapo_steps = float(owner.dapo_steps)
        dvalues = []
        lw = owner.loss_weights
        for f, name in enumerate(fields):
            a.value[f] = values[f].data_ptr()
            dv = torch.empty_like(values[f])
            dvalues.append(dv)
            a.dvalue[f] = dv.data_ptr()
            a.reward[f] = dptr(inputs['reward'][name], torch.float32)
            a.field_mask[f] = dptr(mask[name + '_mask'], torch.float32) if name in _MASK_FIELDS else None
            a.gamma_td[f] = float(owner.gammas['baseline'][name])
            a.w_pg[f] = float(lw['pg'][name])
            a.w_td[f] = float(lw['baseline'][name])
        a.winloss_reward = dptr(inputs['reward']['winloss'], torch.float32)
        a.battle_reward = dptr(inputs['reward']['battle'], torch.float32)
        a.upgo_field = fields.index('winloss')
        for h, name in enumerate(HEADS):
            a.hw_pg[h] = float(owner.pg_head_weights[name])
            a.hw_upgo[h] = float(owner.upgo_head_weights[name])
            a.hw_ent[h] = float(owner.entropy_head_weights[name])
            a.hw_kl[h] = float(owner.kl_head_weights[name])
            a.hw_dapo[h] = float(owner.dapo_head_weights[name])
        only_value = bool(owner.only_update_value)
        a.w_upgo = float(lw['upgo']['winloss'])
        a.w_ent = float(lw['entropy'])
        a.w_kl = float(lw['kl'])
        a.w_atkl = float(lw['action_type_kl'])
        a.w_dapo = float(lw['dapo']) if owner.use_dapo else 0.0
        F_ = len(fields)
        sw = torch.zeros(_lib.AS_LOSS_MAX_STATS, dtype=torch.float32)
        base = F_ * 10
        for f in range(F_):
            sw[f * 10 + 6] = a.w_td[f]
        if only_value:                      # rl_loss.py:170-171: critic loss only -> policy coefficients vanish
            for f in range(F_):
                a.w_pg[f] = 0.0
            a.w_upgo = a.w_ent = a.w_kl = a.w_atkl = a.w_dapo = 0.0
        else:
            for f in range(F_):
                for h in range(6):
                    sw[f * 10 + h] = a.w_pg[f] * a.hw_pg[h]
            for h in range(6):
                sw[base + h] = a.w_upgo * a.hw_upgo[h]
                sw[base + 6 + h] = -a.w_ent * a.hw_ent[h]
                sw[base + 12 + h] = a.w_kl * a.hw_kl[h]
            sw[base + 18] = a.w_atkl
            for h in range(6):
                sw[base + 20 + h] = a.w_dapo * a.hw_dapo[h]
        key = (str(dev), tuple(sw.tolist()))
        cache = owner._sw_cache
        if key not in cache:                       # uploaded once: no H2D copy inside a captured step
            cache.clear()
            cache[key] = sw.to(dev)
        sw = cache[key]
        keep.append(sw)
        a.stat_weight = sw.data_ptr()
        a.grad_scale = None
        partial = torch.empty((B, _lib.AS_LOSS_MAX_STATS), dtype=torch.float32, device=dev)
        out = torch.zeros(_lib.AS_LOSS_MAX_STATS, dtype=torch.float32, device=dev)
        a.partial, a.out = partial.data_ptr(), out.data_ptr()
        st_ptr = _lib.stream_ptr()
        _lib.call('as_rl_loss_forward', ctypes.byref(a), st_ptr)
        _lib.LAUNCHES += 4
        ctx.args = a
        ctx.keep = keep + logits + values + stats + coefs + [partial, out]
        ctx.dlogits, ctx.dvalues = dlogits, dvalues
        ctx.mark_non_differentiable(out)
        total = out[_lib.AS_LOSS_MAX_STATS - 1].clone()
        return total, out

    @staticmethod
    def backward(ctx, gtotal, _gout):
        # d(objective)/d(total_loss) (1 for loss.backward()) stays on the device: the gradient kernels read it
        # through a pointer, so there is neither a host sync nor a separate multiply pass over the 240 MB of dlogits
        gs = gtotal.detach().reshape(1).float().contiguous()
        ctx.args.grad_scale = gs.data_ptr()
        _lib.call('as_rl_loss_backward', ctypes.byref(ctx.args), 1.0, _lib.stream_ptr())
        _lib.LAUNCHES += 2
        ctx.keep.append(gs)
        grads = list(ctx.dlogits) + [g * gs for g in ctx.dvalues]          # the [T+1, B] value gradients are tiny
        return (None, None, None) + tuple(grads)


class LossInfo(dict):
    """The reference's log dict, filled lazily: `['total_loss']` is available at once (autograd tensor); the ~63
    python floats are fetched with ONE device->host copy the first time any other key is read (e.g. when the
    learner does `log_buffer.update(info)` after launching backward), so nothing blocks between forward and backward."""

    def __init__(self, owner, fields, total, stats):
        super().__init__()
        dict.__setitem__(self, 'total_loss', total)
        self._pending = (owner, fields, total, stats)
        self._stats_keep = stats

    def stats_tensor(self):
        return self._stats_keep

    def _fill(self):
        if self._pending is not None:
            owner, fields, total, stats = self._pending
            self._pending = None
            dict.update(self, owner._unpack(fields, total, stats.tolist()))   # the one D2H copy of the step

    def __getitem__(self, k):
        if k != 'total_loss':
            self._fill()
        return dict.__getitem__(self, k)

    def get(self, k, default=None):
        if k != 'total_loss':
            self._fill()
        return dict.get(self, k, default)

    def __contains__(self, k):
        self._fill()
        return dict.__contains__(self, k)

    def __iter__(self):
        self._fill()
        return dict.__iter__(self)

    def __len__(self):
        self._fill()
        return dict.__len__(self)

    def keys(self):
        self._fill()
        return dict.keys(self)

    def items(self):
        self._fill()
        return dict.items(self)

    def values(self):
        self._fill()
        return dict.values(self)

    def __eq__(self, other):
        self._fill()
        return dict.__eq__(self, other)

    def __repr__(self):
        self._fill()
        return dict.__repr__(self)


class ReinforcementLoss:
    def __init__(self, learner_cfg, player_id):
        cfg = C.AttrDict(C.deep_merge(C.DEFAULT_LOSS_CONFIG, dict(learner_cfg or {})))
        self.cfg = cfg
        self._bind(player_id)
        self.only_update_value = False
        self._sw_cache = {}

    def _bind(self, player_id=None):
        cfg = self.cfg
        self.gammas = cfg.gammas
        self.loss_weights = cfg.loss_weights
        self.action_type_kl_steps = cfg.kl.action_type_kl_steps
        self.dapo_steps = cfg.dapo.dapo_steps
        self.use_dapo = cfg.use_dapo
        if player_id is not None and 'MP' not in player_id:      # rl_loss.py:22-24
            self.use_dapo = False
            self.loss_weights.dapo = 0.0
        self.dapo_head_weights = cfg.dapo_head_weights
        self.pg_head_weights = cfg.pg_head_weights
        self.upgo_head_weights = cfg.upgo_head_weights
        self.entropy_head_weights = cfg.entropy_head_weights
        self.kl_head_weights = cfg.kl_head_weights
        self.use_total_rhos = cfg.get('use_total_rhos', False)

    def reset(self, learner_cfg):
        """rl_loss.py:187-199: re-reads the weights; `use_dapo`, `dapo_steps` and the DAPO head weights keep their
        constructor values there, so they do here."""
        keep = (self.use_dapo, self.dapo_steps, self.dapo_head_weights, float(self.loss_weights.get('dapo', 0.0)))
        self.cfg = C.AttrDict(C.deep_merge(self.cfg, dict(learner_cfg or {})))
        self._bind()
        self.use_dapo, self.dapo_steps, self.dapo_head_weights = keep[:3]
        self.only_update_value = False

    def compute_loss(self, inputs):
        if self.use_dapo and 'successive_logit' not in inputs:
            raise KeyError("use_dapo needs inputs['successive_logit'] (rl_loss.py:40)")
        fields = list(inputs['value'].keys())
        if 'winloss' not in fields:
            raise KeyError("the UPGO term needs the 'winloss' baseline (rl_loss.py:112-114)")
        logits = [inputs['target_logit'][h] for h in HEADS]
        if not logits[0].is_cuda:
            raise RuntimeError('applestar_b200.ReinforcementLoss runs on CUDA tensors only (no CPU fallback)')
        values = [inputs['value'][f] for f in fields]
        total, out = _FusedLoss.apply(self, inputs, fields, *logits, *values)
        return LossInfo(self, fields, total, out)

    def _unpack(self, fields, total, s):
        info = {}
        base = len(fields) * 10
        pg_hw, upgo_hw, ent_hw, kl_hw = (self.pg_head_weights, self.upgo_head_weights, self.entropy_head_weights,
                                         self.kl_head_weights)
        for f, name in enumerate(fields):
            tot = 0.0
            for h, head in enumerate(HEADS):
                info['%s/%s' % (name, head)] = s[f * 10 + h]
                tot += s[f * 10 + h] * float(pg_hw[head])
            info[name + '/total'] = tot
            info[name + '/td'] = s[f * 10 + 6]
            info[name + '/reward'] = s[f * 10 + 7]
            info[name + '/value'] = s[f * 10 + 8]
        info['battle/reward'] = s[base + 19]
        for key, off, hw, sign in (('upgo', 0, upgo_hw, 1.0), ('entropy', 6, ent_hw, -1.0), ('kl', 12, kl_hw, 1.0)):
            tot = 0.0
            for h, head in enumerate(HEADS):
                info['%s/%s' % (key, head)] = s[base + off + h]
                tot += sign * s[base + off + h] * float(hw[head])
            info[key + '/total'] = tot
        info['kl/extra_at'] = s[base + 18]
        if self.use_dapo:                        # as_rl_utils.py:105-127 logs the DAPO terms under 'battle/<head>'
            tot = 0.0
            for h, head in enumerate(HEADS):
                info['battle/%s' % head] = s[base + 20 + h]
                tot += s[base + 20 + h] * float(self.dapo_head_weights[head])
            info['battle/total'] = tot
        info['total_loss'] = total
        return info
