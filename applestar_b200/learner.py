"""Data-parallel learner step around `Model` + `ReinforcementLoss`.

Mirrors the compute part of the reference learner: `RLLearner._train` (agent/default/rl_learner.py:82-145:
forward -> loss -> zero_grad -> backward -> sync_gradients -> grad clip -> Adam step) and `DistModule`
(ctools/utils/dist_helper.py:372-439), redesigned for one process per B200:

* all trainable parameters live in ONE flat fp32 buffer (+ flat grad, Adam m/v and a bf16 shadow the GEMMs
  read), so broadcast is one call instead of 432, the optimizer is one fused kernel instead of ~400, and
* gradients are all-reduced per ~25 MB bucket of the flat grad buffer from post-accumulate-grad hooks, on a side
  stream, while the rest of backward is still running (the reference issues 394 blocking per-tensor all_reduces
  after backward, dist_helper.py:421-431).  Trajectories are independent, so there is no other collective.
"""
import os

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _lib
from . import kernels as K


class FlatParams:
    """Re-homes every trainable parameter of `module` into flat buffers (views keep the nn.Parameter objects)."""

    ALIGN = 8          # elements: keeps every tensor 32-byte (fp32) / 16-byte (bf16 shadow) aligned for TMA
    CHUNK = 1 << 16    # elements per CTA of the gradient packing kernel

    def __init__(self, module, order=('encoder', 'core_lstm', 'policy', 'value_networks'), with_shadow=True):
        named = [(k, p) for k, p in module.named_parameters() if p.requires_grad]
        rank = {pfx: i for i, pfx in enumerate(order)}
        named.sort(key=lambda kp: rank.get(kp[0].split('.')[0], len(order)))       # stable: forward order
        self.names = [k for k, _ in named]
        self.params = [p for _, p in named]
        dev = self.params[0].device
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        self.numel = off
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(off, dtype=torch.float32, device=dev)
        self.shadow = torch.zeros(off, dtype=torch.bfloat16, device=dev) if (with_shadow and dev.type == 'cuda') else None
        for p, o in zip(self.params, self.offsets):
            n = p.numel()
            self.flat[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.flat[o:o + n].view(p.shape)
            p.grad = self.grad[o:o + n].view(p.shape)
            if self.shadow is not None:
                p._as_shadow = self.shadow[o:o + n].view(p.shape[0], -1) if p.dim() >= 2 else None
        # CUDA: autograd hands every parameter its gradient as a fresh tensor (p.grad = None before backward, so
        # AccumulateGrad keeps the incoming tensor instead of launching `grad += g` per parameter) and ONE kernel per
        # all-reduce bucket gathers them into the flat buffer (`pack`); CPU (gloo tests): gradients accumulate into views
        self.pack_mode = dev.type == 'cuda' and os.environ.get('APPLESTAR_GRAD_PACK', '1') != '0'
        self._pack_keep = []
        # per-parameter optimizer bookkeeping (torch.optim.Adam skips a parameter whose .grad is None): which tensor an
        # 8-element block of the flat buffer belongs to, and which tensors received a gradient in the current step
        self.block_param = None
        self.active = None
        if self.pack_mode and len(self.params) < 65536:
            bp = torch.empty(off // self.ALIGN, dtype=torch.int16)
            for i, (p, o) in enumerate(zip(self.params, self.offsets)):
                end = self.offsets[i + 1] if i + 1 < len(self.offsets) else off
                bp[o // self.ALIGN:end // self.ALIGN] = i if i < 32768 else i - 65536
            self.block_param = bp.to(dev)
            self._active_host = torch.ones(len(self.params), dtype=torch.int32).pin_memory()
            self.active = torch.ones(len(self.params), dtype=torch.int32, device=dev)
        self.refresh_shadow()

    def pack(self, idx=None):
        """Copy the .grad tensors of the parameters `idx` (default: all) into their slots of the flat gradient buffer;
        a parameter that received no gradient gets zeros."""
        rows, keep = [], []
        for i in (range(len(self.params)) if idx is None else idx):
            p, o, n = self.params[i], self.offsets[i], self.params[i].numel()
            g = p.grad
            if self.active is not None:
                self._active_host[i] = 0 if g is None else 1
            if g is not None and (g.dtype != torch.float32 or not g.is_contiguous()):
                g = g.float().contiguous()
            if g is not None and g.data_ptr() == self.grad.data_ptr() + 4 * o:
                continue                                            # already a view of its slot
            keep.append(g)
            src = 0 if g is None else g.data_ptr()
            for c in range(0, n, self.CHUNK):
                rows.append((src + 4 * c if src else 0, o + c, min(self.CHUNK, n - c)))
        if not rows:
            return
        if not self.grad.is_cuda:                                   # host tensors (gloo tests): plain copies
            pos = 0
            for i in (range(len(self.params)) if idx is None else idx):
                p, o, n = self.params[i], self.offsets[i], self.params[i].numel()
                if p.grad is not None and p.grad.data_ptr() == self.grad.data_ptr() + 4 * o:
                    continue
                g = keep[pos]
                pos += 1
                if g is None:
                    self.grad[o:o + n].zero_()
                else:
                    self.grad[o:o + n].copy_(g.reshape(-1))
            return
        host = torch.tensor(rows, dtype=torch.int64).pin_memory()
        table = host.to(self.grad.device, non_blocking=True)
        capturing = torch.cuda.is_current_stream_capturing()
        if self.active is not None:                                 # all buckets are packed before the optimizer runs
            flags = self._active_host.clone().pin_memory() if capturing else self._active_host
            self.active.copy_(flags, non_blocking=True)
            if capturing:
                self._pack_keep.append((flags,))
        if capturing:                                               # the captured copy re-reads `host` at every replay
            self._pack_keep.append((host, table, keep))
        _lib.call('as_pack_f32', table.data_ptr(), len(rows), self.grad.data_ptr(), _lib.stream_ptr())

    def attach_views(self):
        """Point every p.grad at its slot of the flat buffer (what callers of the reference API read after
        `sync_gradients`)."""
        for p, o in zip(self.params, self.offsets):
            if p.grad is None or p.grad.data_ptr() != self.grad.data_ptr() + 4 * o:
                p.grad = self.grad[o:o + p.numel()].view(p.shape)

    def refresh_shadow(self):
        """Re-cast the whole bf16 shadow from the fp32 master buffer and tag every parameter as in sync
        (kernels.weight_bf16 re-casts a single parameter whose `_version` moved since: torch optimizers,
        `load_state_dict`)."""
        if self.shadow is not None:
            _lib.call('as_cast_f32_bf16', self.flat.data_ptr(), self.shadow.data_ptr(), self.numel, _lib.stream_ptr())
        for p in self.params:
            p._as_shadow_ver = p._version
        K.weights_changed()

    def zero_grad(self):
        if self.pack_mode:
            for p in self.params:
                p.grad = None
            return
        self.grad.zero_()
        self.attach_views()                                    # autograd may have replaced .grad

    def buckets(self, bucket_elems):
        """Contiguous [start, end) ranges of the flat grad buffer, filled from the END (grads of the heads / value
        nets arrive first in backward) -> list of (start, end, param_indices)."""
        out, end, idx = [], self.numel, []
        for i in range(len(self.params) - 1, -1, -1):
            idx.append(i)
            if end - self.offsets[i] >= bucket_elems or i == 0:
                out.append((self.offsets[i], end, idx))
                end, idx = self.offsets[i], []
        return out


class DistModule(nn.Module):
    """Reference-compatible wrapper (`.module`, forwarded attributes, `sync_gradients`, `broadcast_params`)."""

    def __init__(self, module, sync=True, bucket_mb=25.0, flat=None):
        super().__init__()
        self.module = module
        for attr in ('forward', 'state_dict', 'load_state_dict', 'named_parameters', 'sl_train', 'rl_learner_forward',
                     'compute_logp_action', 'compute_teacher_logit'):
            if hasattr(module, attr):
                setattr(self, attr, getattr(module, attr))
        self.flat = flat if flat is not None else FlatParams(module)
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.sync = sync
        self._works = []
        self._buckets = self.flat.buckets(int(bucket_mb * 1024 * 1024 / 4))
        self._pending = [0] * len(self._buckets)
        self._bucket_of = {}
        for b, (_, _, idx) in enumerate(self._buckets):
            for i in idx:
                self._bucket_of[i] = b
        self._comm_stream = torch.cuda.Stream() if self.flat.flat.is_cuda else None
        self._grad_events = {}          # parameter index -> event recorded on the stream that produced its gradient
        self.local_only = False         # True: skip the collectives (graph warm-up of one rank, see LearnerStep._capture)
        self._avg = dist.ReduceOp.AVG if (self.world > 1 and dist.get_backend() == 'nccl') else None
        self.broadcast_params()
        if self.world > 1:
            for i, p in enumerate(self.flat.params):
                p.register_post_accumulate_grad_hook(self._make_hook(i))
            if self._comm_stream is not None:
                # one eager all-reduce per bucket shape (the gradient buffer is still zero): NCCL sets up the channels
                # of every algorithm / protocol it will use, so that a later CUDA-graph capture only records kernels
                self._comm_stream.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(self._comm_stream):
                    for start, end, _ in self._buckets:
                        dist.all_reduce(self.flat.grad[start:end], op=self._avg or dist.ReduceOp.SUM)
                torch.cuda.current_stream().wait_stream(self._comm_stream)
        self.prepare_backward()

    # -- collectives ------------------------------------------------------------------------------
    def broadcast_params(self):
        """One flat broadcast from rank 0 (reference: one broadcast per state_dict tensor, dist_helper.py:433-439)."""
        if self.world > 1:
            dist.broadcast(self.flat.flat, 0)
            self.flat.refresh_shadow()

    def prepare_backward(self):
        """Arm the per-bucket counters for the next backward pass.  `sync_gradients` re-arms them itself, so callers
        that follow the reference's `zero_grad / backward / sync_gradients` sequence never need to call this."""
        self._pending = [len(idx) for _, _, idx in self._buckets]
        self._works = []
        self._grad_events = {}

    def _make_hook(self, i):
        def hook(_param):
            if self.local_only:
                return
            if self._comm_stream is not None:
                # gradients of one bucket are produced on different streams (value towers / scalar encoder run on the
                # side stream): remember where this one was written so that the bucket launch can wait for it
                ev = torch.cuda.Event()
                ev.record()
                self._grad_events[i] = ev
            b = self._bucket_of[i]
            self._pending[b] -= 1
            if self._pending[b] == 0:
                self._launch(b)
        return hook

    def _launch(self, b):
        start, end, idx = self._buckets[b]
        if self._comm_stream is not None:
            cur = torch.cuda.current_stream()
            for i in idx:
                ev = self._grad_events.get(i)
                if ev is not None:
                    cur.wait_event(ev)
        self.flat.pack(idx)                 # gathers whatever is not already a view of its slot (p.grad = None -> zeros)
        chunk = self.flat.grad[start:end]
        if self._comm_stream is not None:
            ev = torch.cuda.Event()
            ev.record()
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                work = dist.all_reduce(chunk, op=self._avg or dist.ReduceOp.SUM, async_op=True)
        else:
            work = dist.all_reduce(chunk, op=dist.ReduceOp.SUM, async_op=True)
        self._works.append((work, chunk))

    def reduce_stats(self, stats):
        """Average the step's packed statistics vector over the ranks, in place, with ONE collective (the reference's
        LogReduceHook moves every logged scalar to the device, all-reduces it and moves it back, ~70 round trips per step:
        ctools/worker/learner/learner_hook.py:313-321).  On CUDA the reduction runs on the communication stream next to
        backward; `sync_gradients` waits for it."""
        if self.world <= 1 or self.local_only:
            return stats
        if self._comm_stream is not None:
            ev = torch.cuda.Event()
            ev.record()
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                work = dist.all_reduce(stats, op=self._avg or dist.ReduceOp.SUM, async_op=True)
            stats.record_stream(self._comm_stream)
        else:
            work = dist.all_reduce(stats, op=dist.ReduceOp.SUM, async_op=True)
        self._works.append((work, stats))
        return stats

    def sync_gradients(self):
        """Average gradients across ranks.  With hooks active this only waits for the in-flight bucket reductions."""
        if self.world <= 1 or self.local_only:
            self.flat.pack()
            self.flat.attach_views()
            return
        for b, n in enumerate(self._pending):
            if n > 0:                       # parameters that received no gradient this step: reduce what is there
                self._pending[b] = 0
                self._launch(b)
        for work, chunk in self._works:
            work.wait()
            if self._avg is None:
                chunk.div_(self.world)
        self.flat.attach_views()            # callers of the reference API read p.grad after sync_gradients
        self.prepare_backward()             # re-arm for the next zero_grad / backward / sync_gradients round


class FusedAdam:
    """clip_grad_norm_(max_norm) + Adam(lr, betas=(0, 0.99), eps=1e-5) as two kernels over the flat buffers
    (reference: rl_learner.py:73-80,125,132 + grad_clip.py:140-143).

    On CUDA (gradient pack mode) the update follows torch.optim.Adam per parameter tensor: a tensor whose `.grad` was
    None this step keeps its weights and moments, and every tensor has its own step count for the bias corrections
    (as_clip_adam_per_param).  Without pack mode (CPU tests) every element is updated with the shared step counter: a
    parameter that received no gradient is then treated as having a zero gradient, i.e. its second moment decays and
    the shared step counter advances for it.  With betas = (0, 0.99) a zero gradient moves no weight (m = 0); it
    matters for a parameter that is frozen for a while and trained later (the policy after an `only_update_baseline`
    phase, rl_learner.py:149-161): torch Adam would start its bias correction at that parameter's own first step,
    here bc2 is already ~1, so its first updates are larger (up to 1 / sqrt(1 - beta2) = 10x for the very first one,
    back to normal after ~1 / (1 - beta2) = 100 steps).  Callers that need the reference behaviour there keep the
    reference's own torch optimizer on the DistModule-wrapped model.  `state_dict` /
    `load_state_dict` carry m, v and the step counter."""

    def __init__(self, flat, lr=1e-5, betas=(0.0, 0.99), eps=1e-5, weight_decay=0.0, max_norm=1.0):
        self.flat, self.lr, self.betas, self.eps, self.wd, self.max_norm = flat, lr, betas, eps, weight_decay, max_norm
        self.m = torch.zeros_like(flat.flat)
        self.v = torch.zeros_like(flat.flat)
        self.t = 0
        self.t_dev = torch.zeros(1, dtype=torch.int32, device=flat.flat.device)
        self.scratch = torch.empty(_lib.lib().as_clip_adam_scratch_bytes() // 4, dtype=torch.float32,
                                   device=flat.flat.device)
        self.grad_norm = torch.zeros(1, dtype=torch.float32, device=flat.flat.device)
        # torch.optim.Adam keeps one step count per parameter and skips parameters without a gradient
        self.p_step = torch.zeros(len(flat.params), dtype=torch.int32, device=flat.flat.device) \
            if flat.active is not None else None

    def step(self):
        """One optimizer step.  The step count lives on the device (incremented by the kernel) so that a captured
        CUDA graph of the learner step replays with correct Adam bias corrections."""
        self.t += 1
        f = self.flat
        if self.p_step is not None:
            _lib.call('as_clip_adam_per_param', f.flat.data_ptr(), f.grad.data_ptr(), self.m.data_ptr(), self.v.data_ptr(),
                      0 if f.shadow is None else f.shadow.data_ptr(), f.numel, self.max_norm, self.lr, self.betas[0],
                      self.betas[1], self.eps, self.wd, f.block_param.data_ptr(), f.active.data_ptr(), self.p_step.data_ptr(),
                      len(f.params), self.t_dev.data_ptr(), self.scratch.data_ptr(), self.grad_norm.data_ptr(),
                      _lib.stream_ptr())
            _lib.LAUNCHES += 1
            K.weights_changed()
            return
        _lib.call('as_clip_adam', f.flat.data_ptr(), f.grad.data_ptr(), self.m.data_ptr(), self.v.data_ptr(),
                  0 if f.shadow is None else f.shadow.data_ptr(), f.numel, self.max_norm, self.lr, self.betas[0],
                  self.betas[1], self.eps, self.wd, 0, self.t_dev.data_ptr(), self.scratch.data_ptr(),
                  self.grad_norm.data_ptr(), _lib.stream_ptr())
        _lib.LAUNCHES += 1
        K.weights_changed()

    def state_dict(self):
        """Adam moments and the device-side step counter (the reference checkpoints `optimizer.state_dict()`,
        ctools/torch_utils/checkpoint_helper.py:85-141), plus the per-parameter step counts."""
        return {'m': self.m.clone(), 'v': self.v.clone(), 't': int(self.t_dev.item()),
                'param_steps': None if self.p_step is None else self.p_step.clone(), 'names': list(self.flat.names),
                'offsets': list(self.flat.offsets), 'lr': self.lr, 'betas': tuple(self.betas), 'eps': self.eps,
                'weight_decay': self.wd, 'max_norm': self.max_norm}

    def load_state_dict(self, state):
        if list(state['names']) != list(self.flat.names) or list(state['offsets']) != list(self.flat.offsets):
            raise ValueError('optimizer state belongs to a different parameter layout')
        self.m.copy_(state['m'])
        self.v.copy_(state['v'])
        self.t = int(state['t'])
        if self.p_step is not None:
            if state.get('param_steps') is not None:
                self.p_step.copy_(state['param_steps'])
            else:
                self.p_step.fill_(self.t)
        self.t_dev.fill_(self.t)
        self.lr, self.betas, self.eps = state['lr'], tuple(state['betas']), state['eps']
        self.wd, self.max_norm = state['weight_decay'], state['max_norm']


def _tree_copy_(dst, src):
    """In-place copy of every tensor of `src` into the same-structured `dst` (H2D when src is pinned host memory)."""
    if isinstance(dst, torch.Tensor):
        dst.copy_(src, non_blocking=True)
    elif isinstance(dst, dict):
        for k in dst:
            _tree_copy_(dst[k], src[k])
    elif isinstance(dst, (list, tuple)):
        for d, s_ in zip(dst, src):
            _tree_copy_(d, s_)


def tree_arena(x, dev):
    """A copy of the tensor tree `x` whose tensors are views of ONE flat device buffer (256-byte aligned pieces), so
    that a whole batch moves between two such trees of the same shapes with a single copy of the flat buffers instead of
    one launch per tensor.  Returns (flat uint8 tensor, tree)."""
    leaves = []

    def walk(t):
        if isinstance(t, torch.Tensor):
            leaves.append(t)
        elif isinstance(t, dict):
            for v in t.values():
                walk(v)
        elif isinstance(t, (list, tuple)):
            for v in t:
                walk(v)
    walk(x)
    offsets, total = [], 0
    for t in leaves:
        offsets.append(total)
        total += (t.numel() * t.element_size() + 255) // 256 * 256
    flat = torch.empty(max(total, 256), dtype=torch.uint8, device=dev)
    it = iter(offsets)

    def build(t):
        if isinstance(t, torch.Tensor):
            off = next(it)
            n = t.numel() * t.element_size()
            view = flat[off:off + n].view(t.dtype).view(t.shape)
            view.copy_(t, non_blocking=True)
            return view
        if isinstance(t, dict):
            return {k: build(v) for k, v in t.items()}
        if isinstance(t, (list, tuple)):
            return type(t)(build(v) for v in t)
        return t
    return flat, build(x)


def _tree_to(x, dev):
    if isinstance(x, torch.Tensor):
        return x.to(dev, non_blocking=True).clone() if x.device == dev else x.to(dev, non_blocking=True)
    if isinstance(x, dict):
        return {k: _tree_to(v, dev) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_tree_to(v, dev) for v in x)
    return x


class LearnerStep:
    """One data-parallel RL learner iteration: forward -> fused loss -> backward -> bucketed all-reduce -> fused
    clip + Adam.  With `use_graph` the whole iteration is captured once per selected-units loop length into a CUDA
    graph and replayed (fixed shapes, no host synchronisation inside), which removes the Python / launch overhead
    of the ~1.4 k kernel launches of a step; inputs are copied into the graph's static buffers each call."""

    def __init__(self, model, loss, lr=1e-5, max_norm=1.0, bucket_mb=25.0, use_graph=True):
        self.model = model
        self.loss = loss
        bucket_mb = float(os.environ.get('APPLESTAR_BUCKET_MB', bucket_mb))       # measurement knob (bucket size of the gradient all-reduce)
        self.dist_model = DistModule(model, bucket_mb=bucket_mb)
        self.flat = self.dist_model.flat
        self.optimizer = FusedAdam(self.flat, lr=lr, max_norm=max_norm)
        self.use_graph = use_graph and self.flat.flat.is_cuda
        self._graphs = {}
        self._pool = None
        self.launches_per_step = None

    # -- eager ----------------------------------------------------------------------------------------
    def _run(self, data, defer_checks=False):
        self.flat.zero_grad()
        self.dist_model.prepare_backward()
        out = self.model.rl_learner_forward(**data, defer_checks=defer_checks)
        info = self.loss.compute_loss(out)
        info['total_loss'].backward()
        self.dist_model.reduce_stats(info.stats_tensor())      # logged scalars: one packed all-reduce (rank mean)
        self.dist_model.sync_gradients()
        self.optimizer.step()
        return info

    # -- graph ----------------------------------------------------------------------------------------
    SU_STEPS = (1, 2, 4, 6, 8, 12, 16, 24, 32, 48, 64)    # captured loop lengths of the selected-units head

    def _key(self, data):
        """Graph key of a batch, or None (eager).  The selected-units loop length is rounded up to one of SU_STEPS
        (any length >= the batch maximum gives the same valid logits) and, when several ranks train together, the
        (any length >= the batch maximum gives the same valid logits).  Ranks may pick different graphs for the same
        step: every graph holds the same sequence of bucket all-reduces, and a rank that has to capture first
        issues no collective while doing so (`_capture`), so the ranks' collectives always pair up.
        Loss settings that are baked into captured launch arguments are part of the key."""
        su = data.get('selected_units_max')
        if su is None:
            return None
        su = int(su)
        su = next((s for s in self.SU_STEPS if s >= su), self.SU_STEPS[-1])
        loss = self.loss
        loss_tag = (bool(getattr(loss, 'only_update_value', False)), bool(getattr(self.model, 'only_update_baseline', False)),
                    repr(sorted((k, repr(v)) for k, v in getattr(loss, 'cfg', {}).items())))
        return (su, int(data['batch_size']), int(data['unroll_len']), tuple(data['entity_info']['x'].shape), loss_tag)

    def _capture(self, data, key):
        dev = self.flat.flat.device
        static_flat, static = tree_arena(data, dev)          # one flat buffer: a staged batch arrives with one copy
        static['selected_units_max'] = key[0]
        self._static_flat = getattr(self, '_static_flat', {})
        self._static_flat[key] = static_flat
        # warm-up (allocator, lazy initialisation) must leave no trace: the weights, the Adam moments and the step
        # counter are put back afterwards, so that the batch that triggers a capture gets exactly one optimizer
        # update like every other batch; and it runs without collectives (a peer may be replaying an existing graph
        # at the same time; NCCL's channels were set up by DistModule.__init__)
        opt = self.optimizer
        self.dist_model.local_only = True
        state = [self.flat.flat, opt.m, opt.v, opt.t_dev] + ([opt.p_step] if opt.p_step is not None else [])
        keep = [t.clone() for t in state]
        keep_t = opt.t
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                        # warm-up on a side stream
            for _ in range(2):
                self._run(static, defer_checks=True)
        torch.cuda.current_stream().wait_stream(side)
        self.dist_model.local_only = False
        for dst, src in zip(state, keep):
            dst.copy_(src)
        opt.t = keep_t
        self.flat.refresh_shadow()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        n0 = _lib.LAUNCHES
        with torch.cuda.graph(graph, pool=self._pool):
            info = self._run(static, defer_checks=True)
        opt.t = keep_t                                       # capturing does not execute: nothing was stepped
        self.launches_per_step = _lib.LAUNCHES - n0
        if self._pool is None:
            self._pool = graph.pool()
        entry = (graph, static, info)
        self._graphs[key] = entry
        return entry

    def __call__(self, data):
        key = self._key(data) if self.use_graph else None
        if key is None:
            return self._run(data)
        entry = self._graphs.get(key)
        if entry is None:
            entry = self._capture(data, key)
        graph, static, info = entry
        if data is not static:
            _tree_copy_(static, data)
        graph.replay()
        self.optimizer.t += 1
        K.weights_changed()                                  # the replay moved the weights behind Python's back
        from .rl_loss import LossInfo
        return LossInfo(self.loss, list(self.model.value_networks.keys()), info['total_loss'], info.stats_tensor())

    def static_inputs(self, data):
        """The graph's own input buffers for `data`'s shape (captures on first use): callers that fill them in place
        (e.g. with H2D copies from pinned memory) skip one device-to-device copy per step."""
        key = self._key(data) if self.use_graph else None
        if key is None:
            return None
        entry = self._graphs.get(key) or self._capture(data, key)
        return entry[1]

    def static_arena(self, data):
        """The flat buffer behind `static_inputs(data)` (learner.tree_arena layout): a staging tree built with
        tree_arena from a batch of the same shapes can be moved in with `static_arena(data).copy_(staging_flat)`."""
        key = self._key(data) if self.use_graph else None
        if key is None:
            return None
        if key not in self._graphs:
            self._capture(data, key)
        return self._static_flat[key]


class ActorStep:
    """Batched actor inference (`Model.compute_logp_action`, model.py:82-116 of the reference) replayed from a CUDA
    graph.  The sampling pointer loop always runs its 64 steps inside the graph; the reference's early `break`
    (action_arg_head.py:309-310) is reproduced after the replay by one host read of the loop's step count and a trim
    of the selected-units tensors, so the results equal the eager call's.  One graph per (batch, entity) shape.

    The returned tensors are views of the graph's static output buffers: they are overwritten by the next call, so a
    caller that keeps them (e.g. to build a trajectory) clones what it needs.  `hidden_state` may be fed back directly:
    inputs are copied into separate static buffers before the replay."""

    def __init__(self, model, use_graph=True):
        self.model = model.eval()
        self.use_graph = use_graph
        self._graphs = {}
        self._pool = None
        self.launches_per_call = None

    def _run(self, static, defer):
        with torch.no_grad():
            return self.model.compute_logp_action(static['spatial_info'], static['entity_info'], static['scalar_info'],
                                                  static['entity_num'], static['hidden_state'],
                                                  gumbel_noise=static['gumbel_noise'], defer_checks=defer,
                                                  defer_trim=defer)

    def _capture(self, data, key):
        dev = next(self.model.parameters()).device
        static = _tree_to(data, dev)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self._run(static, True)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        n0 = _lib.LAUNCHES
        with torch.cuda.graph(graph, pool=self._pool):
            out = self._run(static, True)
        self.launches_per_call = _lib.LAUNCHES - n0
        if self._pool is None:
            self._pool = graph.pool()
        out['_neg_flag'] = K.entity_embed.last_neg_flag            # zeroed and written inside the graph
        K.entity_embed.last_neg_flag = None
        self._graphs[key] = (graph, static, out)
        return self._graphs[key]

    def __call__(self, spatial_info, entity_info, scalar_info, entity_num, hidden_state, gumbel_noise=None):
        from . import ops
        if gumbel_noise is None:
            gumbel_noise = ops.fresh_gumbel(entity_num.shape[0], entity_info['x'].shape[1], entity_num.device)
        data = {'spatial_info': spatial_info, 'entity_info': entity_info, 'scalar_info': scalar_info,
                'entity_num': entity_num, 'hidden_state': hidden_state, 'gumbel_noise': gumbel_noise}
        if not self.use_graph:
            return self._run(data, False)
        key = (int(entity_num.shape[0]), tuple(gumbel_noise['selected_units'].shape))
        entry = self._graphs.get(key) or self._capture(data, key)
        graph, static, out = entry
        _tree_copy_(static, data)
        graph.replay()
        host = torch.stack([out['selected_units_steps'].to(torch.int64), out['_neg_flag'].view(-1)[0].to(torch.int64)])
        steps, neg = host.tolist()                                # the call's one host read
        if neg != 0:
            raise RuntimeError('negative one-hot index in entity_info')
        res = {k: v for k, v in out.items() if k not in ('selected_units_steps', '_neg_flag')}
        for name in ('logit', 'action_info', 'action_logp'):
            res[name] = dict(out[name])
            res[name]['selected_units'] = out[name]['selected_units'][:, :steps]
        return res
