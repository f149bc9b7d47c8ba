"""Trajectory staging between the data loader and the learner step (reference: agent/default/rl_training/
rl_dataloader.py:45-76,206-245: `collate_fn` pads every trajectory tensor to the batch maximum and the learner copies the
whole padded batch to the device).

The one tensor that is almost entirely padding is the teacher's selected-units logits: [T, B, 64, N+1] fp32, 134 MB of
the 456 MB of a 64 x 16 x 512 batch, of which only the rows the teacher emitted (max(selected_units_num, 1) per step) are
data; the rest is the -1e9 fill of `padding_entity_info`.  `pack` (host side, part of collation) keeps only those rows and
one int32 per dense row; `load` copies the compact batch from pinned memory and rebuilds the dense tensor on the device
(as_expand_rows_f32), so the learner step sees exactly the tensors the reference learner gets.  Everything else keeps
the reference's narrow storage dtypes and is copied as is.
"""
import torch

from . import _lib

_SU = ('teacher_logit', 'selected_units')
FILL = -1e9


def pack(batch, pin=True):
    """Collation side: -> a batch dict whose teacher selected-units logits are `{'rows': [n, N+1] fp32, 'row_of': [T*B*64]
    int32, 'shape': (T, B, 64, N+1)}`.  Lossless: a dense row is dropped only if every entry equals the fill value."""
    out = dict(batch)
    out['teacher_logit'] = dict(batch['teacher_logit'])
    dense = batch['teacher_logit']['selected_units']
    shape = tuple(dense.shape)
    flat = dense.reshape(-1, shape[-1])
    keep = (flat != FILL).any(dim=1)
    idx = keep.nonzero().squeeze(1)
    row_of = torch.full((flat.shape[0],), -1, dtype=torch.int32)
    row_of[idx] = torch.arange(idx.numel(), dtype=torch.int32)
    rows = flat[idx].contiguous()
    if pin and torch.cuda.is_available():
        rows, row_of = rows.pin_memory(), row_of.pin_memory()
    out['teacher_logit']['selected_units'] = {'rows': rows, 'row_of': row_of, 'shape': shape}
    return out


def packed_bytes(packed):
    """Host -> device bytes of a packed batch."""
    total = [0]

    def walk(x):
        if isinstance(x, torch.Tensor):
            total[0] += x.numel() * x.element_size()
        elif isinstance(x, dict):
            for v in x.values():
                walk(v)
        elif isinstance(x, (list, tuple)):
            for v in x:
                walk(v)
    walk(packed)
    return total[0]


class Stager:
    """Device side: staging buffers for one packed batch shape.  `load(packed, dst)` issues the H2D copies (non-blocking,
    on the current stream) and expands the compact teacher logits into `dst['teacher_logit']['selected_units']`."""

    def __init__(self, packed, device):
        c = packed['teacher_logit']['selected_units']
        self.capacity = max(1, int(c['row_of'].numel()))            # worst case: every dense row is data
        self.K = c['shape'][-1]
        self.rows = torch.empty((self.capacity, self.K), dtype=torch.float32, device=device)
        self.row_of = torch.empty(c['row_of'].numel(), dtype=torch.int32, device=device)

    def load(self, packed, dst):
        from .learner import _tree_copy_
        c = packed['teacher_logit']['selected_units']
        n = c['rows'].shape[0]
        rest = dict(packed)
        rest['teacher_logit'] = {k: v for k, v in packed['teacher_logit'].items() if k != 'selected_units'}
        dst_rest = dict(dst)
        dst_rest['teacher_logit'] = {k: v for k, v in dst['teacher_logit'].items() if k != 'selected_units'}
        _tree_copy_(dst_rest, rest)
        if n:
            self.rows[:n].copy_(c['rows'], non_blocking=True)
        self.row_of.copy_(c['row_of'], non_blocking=True)
        out = dst['teacher_logit']['selected_units']
        assert out.is_contiguous() and out.dtype == torch.float32 and tuple(out.shape) == tuple(c['shape'])
        _lib.call('as_expand_rows_f32', self.rows.data_ptr(), self.row_of.data_ptr(), out.data_ptr(), self.row_of.numel(),
                  self.K, FILL, _lib.stream_ptr())
        return dst
