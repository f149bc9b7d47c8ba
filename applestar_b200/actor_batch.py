"""GPU side of the actor's batched inference (reference: distar/agent/default/agent.py:37-89 `copy_input_data` /
`copy_output_data`, :781-805 `Agent.gpu_batch_inference`, distar/actor/actor.py:268-299 `_gpu_inference_loop`).

Protocol kept from the reference: every environment process owns one row of a set of shared-memory input tensors, writes
its observation there (`copy_input_data`), raises its flag in `signals` and polls the flag; the GPU loop serves all
raised rows with one batched model call, writes their results into the shared output tensors and clears the flags.

What is different underneath: the shared tensors are page-locked once (so the batch travels with asynchronous copies
into device buffers that live as long as the object), the model call is `learner.ActorStep` (one CUDA-graph replay for
the whole batch, all 64 sampling steps inside) and the results of the raised rows are gathered on the device and come
back with one transfer per tensor.  The environment side needs SC2 and is not part of this repository; `tests/` drive
the protocol with synthetic observations.
"""
import torch

from .learner import ActorStep, _tree_copy_


def copy_input_data(shared, step_data, idx):
    """Environment side: write one observation into row `idx` of the shared input tensors (agent.py:37-67: entity rows
    beyond entity_num, selected units beyond selected_units_num and map area beyond the game's h x w keep their padding)."""
    entity_num = int(step_data['entity_num'])
    sun = int(step_data.get('selected_units_num', 0))
    for key, val in step_data.items():
        if key == 'value_feature':
            continue
        if key == 'hidden_state':
            for layer, (h, c) in enumerate(val):
                shared['hidden_state'][layer][0][idx].copy_(h.reshape(-1))
                shared['hidden_state'][layer][1][idx].copy_(c.reshape(-1))
        elif isinstance(val, torch.Tensor):
            shared[key][idx].copy_(val)
        elif isinstance(val, dict):
            for sub, t in val.items():
                dst = shared[key][sub]
                if key == 'action_info' and sub == 'selected_units':
                    if sun > 0:
                        dst[idx, :sun].copy_(t)
                elif key == 'entity_info':
                    dst[idx, :entity_num].copy_(t)
                elif key == 'spatial_info' and 'effect' not in sub:
                    h, w = t.shape
                    dst[idx].zero_()
                    dst[idx, :h, :w].copy_(t)
                else:
                    dst[idx].copy_(t)


def copy_output_data(shared, out, rows):
    """GPU side: rows `rows` (bool mask or index tensor) of the model output -> the shared output tensors (agent.py:70-89).
    The raised rows are selected on the device, so only they cross the bus."""
    rows = rows.nonzero().squeeze(1) if rows.dtype == torch.bool else rows
    if rows.numel() == 0:
        return

    def put(dst, src):
        picked = src.index_select(0, rows.to(src.device)).cpu()
        view = dst
        for d in range(1, picked.dim()):
            view = view.narrow(d, 0, picked.shape[d])      # the model may return fewer steps / entities than the buffer holds
        view.index_copy_(0, rows.cpu(), picked.to(dst.dtype))

    for key, val in out.items():
        if key not in shared:
            continue
        if key == 'hidden_state':
            for layer, (h, c) in enumerate(val):
                put(shared['hidden_state'][layer][0], h)
                put(shared['hidden_state'][layer][1], c)
        elif isinstance(val, dict):
            for sub, t in val.items():
                if sub in shared[key]:
                    put(shared[key][sub], t)
        elif isinstance(val, torch.Tensor):
            put(shared[key], val)


def _tree_apply(x, fn):
    if isinstance(x, torch.Tensor):
        return fn(x)
    if isinstance(x, dict):
        return {k: _tree_apply(v, fn) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_tree_apply(v, fn) for v in x)
    return x


def _page_lock(t):
    """cudaHostRegister on a (shared-memory) CPU tensor so that copies from it are asynchronous; best effort."""
    if t.is_pinned() or t.numel() == 0:
        return True
    rc = torch.cuda.cudart().cudaHostRegister(t.data_ptr(), t.numel() * t.element_size(), 0)
    return int(rc) == 0


class SharedBatchInference:
    """`Agent.gpu_batch_inference` (agent.py:781-805) for one model.  `teacher=True` serves `compute_teacher_logit`
    (the shared input then also carries `action_info` and `selected_units_num`)."""

    def __init__(self, model, shared_input, shared_output, signals, teacher=False):
        self.model = model.eval()
        self.teacher = teacher
        self.shared_input, self.shared_output, self.signals = shared_input, shared_output, signals
        self.device = next(model.parameters()).device
        self.locked = all(_tree_leaves(_tree_apply(shared_input, _page_lock)))
        self.dev_input = _tree_apply(shared_input, lambda t: torch.empty_like(t, device=self.device))
        self.step = None if teacher else ActorStep(model)
        self.served = 0

    def poll(self, gumbel_noise=None):
        """One turn of the GPU loop: serve every raised row, clear its flag; returns how many rows were served."""
        raised = self.signals.clone().bool()
        n = int(raised.sum())
        if n == 0:
            return 0
        _tree_copy_(self.dev_input, self.shared_input)            # whole batch, as the reference (rows are fixed slots)
        d = self.dev_input
        with torch.no_grad():
            if self.teacher:
                out = self.model.compute_teacher_logit(d['spatial_info'], d['entity_info'], d['scalar_info'], d['entity_num'],
                                                       d['hidden_state'], d['selected_units_num'], d['action_info'])
            else:
                out = self.step(d['spatial_info'], d['entity_info'], d['scalar_info'], d['entity_num'], d['hidden_state'],
                                gumbel_noise=gumbel_noise)
        copy_output_data(self.shared_output, out, raised)
        self.signals[raised] = 0
        self.served += n
        return n


def _tree_leaves(x):
    if isinstance(x, dict):
        for v in x.values():
            yield from _tree_leaves(v)
    elif isinstance(x, (list, tuple)):
        for v in x:
            yield from _tree_leaves(v)
    else:
        yield x
