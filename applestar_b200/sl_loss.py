"""SupervisedLoss of the SL learner (reference: agent/default/sl_training/sl_loss.py:39-285, called from
sl_learner.py:50-53 on the logits `Model.sl_train` returns): same constructor, `compute_loss` signature, loss weights
(default_supervised_loss.yaml) and log keys.  The per-row work over the logits (log-softmax, label term, arg-max, the
SelectedUnits label mask, and the gradient) runs in two kernels (csrc/ce_rows.cu); what is left in torch are reductions
over per-row vectors and the index bookkeeping of the logged accuracy statistics."""
import copy

import torch

from . import _lib
from .ops import require_cuda

DEFAULT_LOSS_WEIGHT = {'action_type': 30.0, 'delay': 9.0, 'queued': 1.0, 'selected_units': 4.0,
                       'select_unit_num_logits': 8.0, 'target_unit': 4.0, 'target_location': 8.0}
HEADS = ('action_type', 'delay', 'queued', 'selected_units', 'target_unit', 'target_location')


def _get(cfg, key, default=None):
    if cfg is None:
        return default
    if isinstance(cfg, dict):
        return cfg.get(key, default)
    return getattr(cfg, key, default)


class _CERows(torch.autograd.Function):
    """loss_row, argmax = CE(logits [R, K], labels [R]); optional SelectedUnits label mask (excl [R/S, S], excl_len [R/S])."""

    @staticmethod
    def forward(ctx, logits, labels, excl, excl_len, S, smoothing):
        require_cuda(logits)
        assert logits.dim() == 2 and logits.dtype == torch.float32
        logits = logits if logits.stride(1) == 1 else logits.contiguous()
        R, K = logits.shape
        labels = labels.reshape(-1).long().contiguous()
        if excl is not None:
            excl, excl_len = excl.long().contiguous(), excl_len.long().contiguous()
        loss = torch.empty(R, dtype=torch.float32, device=logits.device)
        lse = torch.empty_like(loss)
        arg = torch.empty(R, dtype=torch.int32, device=logits.device)
        _lib.call('as_ce_rows_fwd', logits.data_ptr(), logits.stride(0), labels.data_ptr(),
                  excl.data_ptr() if excl is not None else None, excl_len.data_ptr() if excl is not None else None, S, K,
                  float(smoothing), loss.data_ptr(), lse.data_ptr(), arg.data_ptr(), R, _lib.stream_ptr())
        ctx.save_for_backward(logits, labels, lse, *((excl, excl_len) if excl is not None else ()))
        ctx.S, ctx.smoothing = S, float(smoothing)
        ctx.mark_non_differentiable(arg)
        return loss, arg

    @staticmethod
    def backward(ctx, g, _):
        logits, labels, lse = ctx.saved_tensors[:3]
        excl, excl_len = ctx.saved_tensors[3:] if len(ctx.saved_tensors) > 3 else (None, None)
        R, K = logits.shape
        d = torch.empty((R, K), dtype=torch.float32, device=logits.device)
        g = g.contiguous().float()
        _lib.call('as_ce_rows_bwd', logits.data_ptr(), logits.stride(0), labels.data_ptr(),
                  excl.data_ptr() if excl is not None else None, excl_len.data_ptr() if excl is not None else None, ctx.S, K,
                  ctx.smoothing, lse.data_ptr(), g.data_ptr(), d.data_ptr(), R, _lib.stream_ptr())
        return d, None, None, None, None, None


def ce_rows(logits, labels, smoothing=0.0, excl=None, excl_len=None, S=1):
    return _CERows.apply(logits, labels, excl, excl_len, S, smoothing)


def _sequence_mask(lengths, max_len):
    """ctools/torch_utils/rnn.py:20-36"""
    return torch.arange(max_len, device=lengths.device).unsqueeze(0) < lengths.unsqueeze(1)


class SupervisedLoss:
    def __init__(self, cfg=None):
        learner = copy.deepcopy(_get(cfg, 'learner', None))
        self.whole_cfg = cfg
        self.cfg = learner
        self.loss_weight = dict(DEFAULT_LOSS_WEIGHT)
        lw = _get(learner, 'loss_weight', None)
        if lw:
            self.loss_weight.update({k: float(lw[k]) for k in lw})
        self.smoothing = 0.1 if _get(learner, 'label_smooth', False) else 0.0      # sl_loss.py:56-59, :19
        self.su_mask = bool(_get(learner, 'su_mask', True))
        self.cross_rank_loss = bool(_get(learner, 'cross_rank_loss', False))
        self.total_batch_size = None
        dist_on = torch.distributed.is_available() and torch.distributed.is_initialized()
        self.rank = torch.distributed.get_rank() if dist_on else 0
        self.world_size = torch.distributed.get_world_size() if dist_on else 1

    def register_stats(self, record, tb_logger):
        """sl_loss.py:70-100"""
        names = ['total_loss', 'action_type_loss', 'delay_loss', 'queued_loss', 'selected_units_loss',
                 'selected_units_loss_norm', 'selected_units_end_flag_loss', 'target_unit_loss', 'target_location_loss',
                 'action_type_acc', 'delay_distance_L1', 'queued_acc', 'selected_units_iou', 'target_unit_acc',
                 'target_location_distance_L2']
        for n in names:
            record.register_var(n)
            tb_logger.register_var(n)

    # sl_loss.py:102-121
    def compute_loss(self, policy_logits, actions, actions_mask, selected_units_num, entity_num, infer_action_info):
        if self.cross_rank_loss:
            self.total_batch_size = torch.tensor(entity_num.shape[0], device=entity_num.device)
            if self.world_size > 1:
                torch.distributed.all_reduce(self.total_batch_size)
        out = {}
        for name in HEADS:
            if name == 'selected_units':
                out.update(self._selected_units_loss(policy_logits[name], actions[name], actions_mask[name], selected_units_num,
                                                     entity_num, infer_action_info['selected_units']))
            else:
                out.update(self._head_loss(name, policy_logits[name], actions[name], actions_mask[name]))
        total = 0.
        for name in HEADS:
            total = total + out[name + '_loss'] * self.loss_weight[name]
        out['total_loss'] = total
        return out

    def _reduce(self, loss_tmp, labels, mask):
        """sl_loss.py:130-139 (identical in every single-label head)"""
        if self.cross_rank_loss:
            return (labels.shape[0] / self.total_batch_size * self.world_size) * loss_tmp.mean()
        valid = mask.sum()
        return torch.where(valid > 0, loss_tmp.sum() / valid.clamp(min=1e-30), loss_tmp.sum() * 0)

    def _head_loss(self, name, logits, labels, mask):
        mask = mask.float()
        labels = labels.long()
        row, arg = ce_rows(logits, labels, self.smoothing)
        loss = self._reduce(row * mask, labels, mask)
        arg = arg.long()
        with torch.no_grad():
            if name == 'action_type':                                         # sl_loss.py:124-126
                return {'action_type_loss': loss, 'action_type_acc': (arg == labels).float().sum() / len(labels)}
            if name == 'delay':                                               # :162-164
                return {'delay_loss': loss, 'delay_distance_L1': ((arg - labels).abs() * mask).sum() / (mask.sum() + 1e-6)}
            if name == 'queued':                                              # :182-184 (an L1 distance, named acc)
                return {'queued_loss': loss, 'queued_acc': ((arg - labels).abs() * mask).sum() / (mask.sum() + 1e-6)}
            if name == 'target_unit':                                         # :257-259
                return {'target_unit_loss': loss, 'target_unit_acc': ((arg == labels) * mask).sum() / (mask.sum() + 1e-6)}
            W = 160                                                           # :262, :276-285
            d2 = ((arg % W - labels % W) ** 2 + (arg // W - labels // W) ** 2).float().sqrt()
            return {'target_location_loss': loss, 'target_location_distance_L2': (d2 * mask).sum() / (mask.sum() + 1e-6)}

    # sl_loss.py:187-244
    def _selected_units_loss(self, logits, labels, mask, lengths, entity_num, selected_units):
        b, s, n = logits.shape
        mask = mask.float()
        lengths = lengths.long()
        labels_s = labels[:, :s].long().contiguous()
        excl = excl_len = None
        if self.su_mask:
            excl, excl_len = labels_s, (lengths - 1).clamp(min=0).clamp(max=s)
        row, _ = ce_rows(logits.reshape(b * s, n), labels_s.reshape(-1), 0.0, excl, excl_len, s)
        select_mask = _sequence_mask(lengths, s)
        loss_tmp = row.view(b, s) * select_mask * mask.unsqueeze(1)
        loss = loss_tmp.sum() / b
        if self.cross_rank_loss:
            loss = (labels.shape[0] / self.total_batch_size * self.world_size) * loss
        loss_norm = loss_tmp.sum() / (lengths.sum() + 1e-6)
        end_flag_loss = loss_tmp[torch.arange(b, device=logits.device), lengths - 1].mean()
        with torch.no_grad():
            if selected_units is not None:
                preds = selected_units[:, :s].long()
                # first occurrence of the end flag (the reference's CPU sort is stable; a CUDA sort needs to be told)
                end_idx = torch.sort((preds == entity_num.unsqueeze(1)).long(), dim=-1, descending=True, stable=True)[1][:, 0]
                invalid = end_idx == 0
                end_idx = end_idx + 1
                end_idx[invalid] += s
                preds = (preds + 1) * _sequence_mask(end_idx, s)
                lab = (labels_s + 1) * select_mask
                pz = torch.zeros(b, n + 1, device=logits.device, dtype=torch.bool).scatter_(1, preds, True)
                lz = torch.zeros(b, n + 1, device=logits.device, dtype=torch.bool).scatter_(1, lab, True)
                inter = (pz & lz)[:, 1:].sum(1)
                union = (pz | lz)[:, 1:].sum(1)
                iou = (inter / (union + 1e-6) * mask).sum() / (mask.sum() + 1e-6)
            else:
                iou = 0.
        return {'selected_units_loss': loss, 'selected_units_loss_norm': loss_norm,
                'selected_units_end_flag_loss': end_flag_loss, 'selected_units_iou': iou}
