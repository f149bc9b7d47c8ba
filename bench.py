#!/usr/bin/env python
"""Benchmark of the hot path (BASELINE.json metric: RL learner traj-steps/s).

  python bench.py --gpus N --steps K --warmup W            # this repo, one rank per GPU (torchrun for N > 1)
  python bench.py --impl reference ...                      # the reference algorithm's CPU path (oracle port) on host cores

A "step" = one learner iteration over a synthetic batch of `--traj` trajectories x `--unroll` steps per GPU:
forward (encoder on (T+1)*B observations, LN-LSTM core, teacher-forced heads, baselines) -> fused RL loss ->
backward -> bucketed gradient all-reduce (N > 1) -> fused clip + Adam.  Weak scaling: 16 trajectories per GPU,
so N = 8 is the 128-trajectory data-parallel configuration of BASELINE.json.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--traj', type=int, default=16, help='trajectories per GPU')
    ap.add_argument('--unroll', type=int, default=64)
    ap.add_argument('--entities', type=int, default=512)
    ap.add_argument('--cpu-sample', default='8x2', help='unroll x trajectories of the bounded CPU sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-inference', action='store_true', help='skip the batched actor-inference figure (N = 1 only)')
    ap.add_argument('--no-extras', action='store_true',
                    help='skip the secondary measurements of the JSON line (library baseline, encoder microbench, HBM rooflines)')
    ap.add_argument('--workload', default='learner', choices=['learner', 'encoder_microbench'],
                    help='encoder_microbench: BASELINE.json configs[1] alone (entity Transformer + ScatterConnection + '
                         'spatial encoder, batch 1024) as its own JSON line')
    return ap.parse_args()


def load_spec():
    with open(os.path.join(ROOT, 'tests', 'golden', 'state_dict_spec.json')) as f:
        return [(k, tuple(s), t) for k, s, t in json.load(f)['winloss']]


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get('hbm_gbs', 6650.0), d.get('bf16_tflops', 1590.0), d.get('bf16_tflops_sustained', 1400.0), 'measured'
    return 6650.0, 1590.0, 1400.0, 'fallback'


# --------------------------------------------------------------------------------------------- CPU arm
def cpu_learner_throughput(T, B, N, steps, warmup, threads):
    """Oracle port (oracle/restate.py = CPU restatement of the reference algorithm) forward + loss + backward."""
    from applestar_b200 import synth, config as C
    from oracle import restate as R
    torch.set_num_threads(threads)
    sd = synth.full_state_dict(load_spec(), 0)
    sd = {k: v.requires_grad_(True) for k, v in sd.items()}
    data = synth.gen_trajectories(T, B, N, 1234)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        out = R.rl_learner_forward(sd, ['winloss'], **synth.tree_map(lambda t: t.clone(), data))
        info = R.compute_loss(C.DEFAULT_LOSS_CONFIG, out)
        for v in sd.values():
            v.grad = None
        info['total_loss'].backward()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    dt = sum(times) / len(times)
    return T * B / dt, dt


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    if 'WORLD_SIZE' in os.environ or 'OMP_NUM_THREADS' in os.environ:
        # torchrun pins OMP_NUM_THREADS=1 in its workers: time the CPU arm in a child with the launcher's variables
        # removed so that it can use every host core
        import subprocess
        drop = ('RANK', 'LOCAL_RANK', 'WORLD_SIZE', 'LOCAL_WORLD_SIZE', 'GROUP_RANK', 'ROLE_RANK', 'ROLE_WORLD_SIZE',
                'MASTER_ADDR', 'MASTER_PORT', 'OMP_NUM_THREADS', 'GROUP_WORLD_SIZE', 'ROLE_NAME')
        env = {k: v for k, v in os.environ.items() if k not in drop and not k.startswith('TORCHELASTIC')}
        out = subprocess.run([sys.executable, os.path.abspath(__file__)] + sys.argv[1:], env=env, capture_output=True, text=True)
        sys.stderr.write(out.stderr[-2000:])
        sys.stdout.write(out.stdout)
        sys.stdout.flush()
        return
    T, B = (int(x) for x in args.cpu_sample.split('x'))
    threads = os.cpu_count() or 1
    steps, warmup = max(1, min(args.steps, 3)), min(args.warmup, 1)
    v, dt = cpu_learner_throughput(T, B, args.entities, steps, warmup, threads)
    line = {
        'impl': 'reference', 'metric': 'rl_learner_traj_steps_per_sec', 'value': v, 'unit': 'traj-steps/s',
        'n_gpus': args.gpus, 'steps': steps, 'warmup': warmup, 'ms_per_step': dt * 1e3, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': 'rl_learner_step_T%d_B%d_N%d' % (args.unroll, args.traj, args.entities),
                   'sample': 'T=%d B=%d N=%d per step' % (T, B, args.entities)},
        'cpu_baseline': {'value': v, 'unit': 'traj-steps/s', 'cores': threads, 'kind': 'port',
                         'sample': 'oracle/restate.py forward+loss+backward, T=%d B=%d N=%d, %d steps' % (
                             T, B, args.entities, steps)},
        'e2e': {'value': v, 'unit': 'traj-steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------- GPU arm
class ClockSampler:
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.proc, self.lines = None, []
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for ts, ln in self.lines:
            if ts < t0 or ts > t1:
                continue
            parts = [p.strip() for p in ln.split(',')]
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except (ValueError, IndexError):
                continue
            for nm, val in zip(names, parts[2:]):
                if val.lower().startswith('active'):
                    reasons.add(nm)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons),
                'samples': len(sm)}


def gemm_flops_hook(records):
    """Wrap _lib.call so every as_gemm_bf16 / as_conv_nhwc_bf16 / as_attn_* launch is bracketed by CUDA events (used for ONE
    extra instrumented step after the timed region; the timed steps run unhooked)."""
    from applestar_b200 import _lib
    orig = _lib.call

    def hooked(name, *a):
        if name not in ('as_gemm_bf16', 'as_gemm_bf16_masked', 'as_gemm_bf16_relu_bits', 'as_gemm_bf16_bitmasked', 'as_conv_nhwc_bf16', 'as_conv_nhwc_bf16_add',
                        'as_conv_nhwc_bf16_masked', 'as_conv_nhwc_bf16_relu_bits', 'as_conv_nhwc_bf16_bitmasked', 'as_conv_wgrad_nhwc_bf16', 'as_attn_fwd', 'as_attn_bwd_dq'):
            return orig(name, *a)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(120000)         # keep the stream busy while the host enqueues, so that e0 -> e1 is the kernel's
        e0.record()                       # own duration and not the host's launch latency (tiny GEMMs of the heads)
        orig(name, *a)
        e1.record()
        if name == 'as_gemm_bf16':
            M, N, K, nb = a[5], a[6], a[7], a[13] * a[14]
            fl = 2.0 * M * N * K * nb
        elif name in ('as_gemm_bf16_masked', 'as_gemm_bf16_bitmasked'):   # (A, B, C, mask, M, N, K, ...)
            fl = 2.0 * a[4] * a[5] * a[6]
        elif name == 'as_gemm_bf16_relu_bits':            # (A, B, C, bias, bits, M, N, K, ...)
            fl = 2.0 * a[5] * a[6] * a[7]
        elif name in ('as_conv_nhwc_bf16_add', 'as_conv_nhwc_bf16_masked', 'as_conv_nhwc_bf16_relu_bits', 'as_conv_nhwc_bf16_bitmasked'):   # (x, w, y, bias|addend, addend|mask, R, H, W, Cin, Cout, taps, ...)
            R, H, W, Cin, Cout, taps = a[5], a[6], a[7], a[8], a[9], a[10]
            fl = 2.0 * R * (H + 2) * (W + 2) * Cin * Cout * taps
        elif name == 'as_conv_wgrad_nhwc_bf16':           # (dy, x, dw, ws, R, H, W, Cin, Cout, taps, ...)
            R, H, W, Cin, Cout, taps = a[4], a[5], a[6], a[7], a[8], a[9]
            fl = 2.0 * R * (H + 2) * (W + 2) * Cin * Cout * taps
        elif name == 'as_attn_fwd':                       # Q K^T and P V: 2 x (2 N N D) per (row, head)
            R, N, H, D = a[4], a[5], a[6], a[7]
            fl = 4.0 * R * H * N * N * D
        elif name == 'as_attn_bwd_dq':                    # dO V^T and dS K
            R, N, H, D = a[6], a[7], a[8], a[9]
            fl = 4.0 * R * H * N * N * D
        else:
            R, H, W, Cin, Cout, taps = a[4], a[5], a[6], a[7], a[8], a[9]
            fl = 2.0 * R * (H + 2) * (W + 2) * Cin * Cout * taps
        records.append((e0, e1, fl, (name,) + tuple(a[4:16])))
    _lib.call = hooked
    return orig


def actor_inference(model, dev, batch=256, iters=8, cpu_baseline=True):
    """BASELINE.json's second metric (reported per GPU): `compute_logp_action` on `batch` observations x 512 entities
    through learner.ActorStep (CUDA graph, all 64 sampling steps inside).  `value`: observations resident on the
    device; `e2e`: every call copies its observations (and the Gumbel noise it samples with) from pinned host memory
    and reads the sampled actions + log-probs back to the host, inside the timed region."""
    from applestar_b200 import synth
    from applestar_b200.learner import ActorStep
    _, _, tf_sus, which = peaks()
    actor = ActorStep(model)
    host_obs = synth.tree_map(lambda t: t.pin_memory(), synth.gen_observations(batch, 512, 7))
    host_noise = synth.tree_map(lambda t: t.pin_memory(), synth.gen_gumbel(batch, 11))
    obs = synth.tree_map(lambda t: t.to(dev), host_obs)
    noise = synth.tree_map(lambda t: t.to(dev), host_noise)
    hidden = [(torch.zeros(batch, 384, device=dev), torch.zeros(batch, 384, device=dev)) for _ in range(3)]
    nbytes = [0, 0]

    def call(o=obs, n=noise):
        return actor(o['spatial_info'], o['entity_info'], o['scalar_info'], o['entity_num'], hidden, gumbel_noise=n)

    # end to end: the observations (and noise) of call i+1 travel from pinned host memory on a copy stream while call i
    # computes (two device staging sets, as an actor loop with two groups of environments would run it); the sampled
    # actions and log-probs of every call are read back to the host inside the timed region
    from applestar_b200.learner import _tree_copy_
    host_in = [{k: host_obs[k] for k in ('spatial_info', 'entity_info', 'scalar_info', 'entity_num')}, host_noise]
    stage = [synth.tree_map(lambda t: torch.empty_like(t, device=dev), host_in) for _ in range(2)]
    copy_stream = torch.cuda.Stream(device=dev)
    ev_ready = [torch.cuda.Event(), torch.cuda.Event()]
    ev_done = [torch.cuda.Event(), torch.cuda.Event()]
    turn = [0]

    def upload(slot):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_done[slot])
            _tree_copy_(stage[slot], host_in)
            ev_ready[slot].record(copy_stream)

    def call_e2e(count=False):
        slot = turn[0] & 1
        turn[0] += 1
        cur = torch.cuda.current_stream()
        cur.wait_event(ev_ready[slot])
        upload(slot ^ 1)
        res = call(stage[slot][0], stage[slot][1])
        ev_done[slot].record(cur)
        back = [res['selected_units_num']] + list(res['action_info'].values()) + list(res['action_logp'].values())
        back = [t.cpu() for t in back]
        if count:
            tot = [0]
            synth.tree_map(lambda t: tot.__setitem__(0, tot[0] + t.numel() * t.element_size()), host_in)
            nbytes[0], nbytes[1] = tot[0], sum(t.numel() * t.element_size() for t in back)

    for _ in range(3):
        call()
    torch.cuda.synchronize()
    e0, e1 = _events()
    e0.record()
    for _ in range(iters):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    for ev in ev_done:
        ev.record()
    upload(0)
    call_e2e(count=True)
    call_e2e()
    torch.cuda.synchronize()
    e0, e1 = _events()
    e0.record()
    for _ in range(iters):
        call_e2e()
    e1.record()
    torch.cuda.synchronize()
    ms_e2e = e0.elapsed_time(e1) / iters
    obs_s = batch / ms * 1e3
    line = {'metric': 'inference_obs_per_sec_per_gpu', 'value': obs_s, 'unit': 'obs/s', 'ms_per_call': ms,
            'config': {'workload': 'compute_logp_action_B%d_N512' % batch, 'cuda_graph': True,
                       'e2e_pipeline': 'H2D of call i+1 on a copy stream under the compute of call i; D2H of every result',
                       'launches_per_call': actor.launches_per_call},
            'e2e': {'value': batch / ms_e2e * 1e3, 'unit': 'obs/s', 'ms_per_call': ms_e2e,
                    'h2d_bytes_per_step': nbytes[0], 'd2h_bytes_per_step': nbytes[1]},
            'roofline': {'bound': 'tensor', 'achieved': obs_s * 6.81e9 / 1e12, 'peak': tf_sus, 'unit': 'TFLOP/s',
                         'frac': obs_s * 6.81e9 / 1e12 / tf_sus, 'flop_per_obs': 6.81e9,
                         'peak_source': which + ' bf16_tflops_sustained',
                         'note': 'whole call incl. the 64 sequential pointer steps and the small-M heads (latency-bound)'}}
    if cpu_baseline:
        line['cpu_baseline'] = cpu_inference_throughput(16, os.cpu_count() or 1)
    return line


def cpu_inference_throughput(batch, threads):
    """Oracle port of `compute_logp_action` (oracle/restate.py) on the host cores: bounded sample of `batch` observations."""
    from applestar_b200 import synth, tables
    from oracle import restate as R
    torch.set_num_threads(threads)
    sd = synth.full_state_dict(load_spec(), 0)
    data = synth.gen_observations(batch, 512, 7)
    noise = synth.gen_gumbel(batch, 11)
    su_table = torch.tensor(tables.flags(tables.ACTION_HAS_SELECTED_UNITS))
    times = []
    for i in range(3):
        t0 = time.perf_counter()
        with torch.no_grad():
            R.compute_logp_action(sd, noise=noise, su_table=su_table, **synth.tree_map(lambda t: t.clone(), data))
        times.append(time.perf_counter() - t0)
    dt = min(times[1:])
    return {'value': batch / dt, 'unit': 'obs/s', 'cores': threads, 'kind': 'port',
            'sample': 'oracle/restate.py compute_logp_action, B=%d N=512, best of 2 calls' % batch}


def algorithmic_flops(T, B, su_steps_mean, fields=1):
    """dense algorithmic FLOPs of one learner step (forward x 3 for forward + backward)."""
    fwd = (T + 1) * B * 5.138e9 + T * B * (1.574e9 + 0.024e9 + su_steps_mean * 0.00116e9) + (T + 1) * B * fields * 0.0044e9
    return 3.0 * fwd


def _events():
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def gpu_library_baseline(T, B, N, dev, hbm_guard_gb=150.0):
    """SURVEY §8(d) 'GPU reference baseline': the SAME algorithm through stock ATen / cuBLAS / cuDNN kernels on the same
    B200 -- the oracle restatement of the reference (plain torch ops, its autograd) moved to the GPU, fp32 with TF32
    off and under autocast(bfloat16): forward + loss + backward.  This is the 'existing GPU path' every kernel of this
    repo has to beat; it is a reported baseline, never part of the product path.  The largest trajectory count that
    fits is used (and named)."""
    from applestar_b200 import synth, config as C
    from oracle import restate as R
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sd = {k: v.to(dev).requires_grad_(True) for k, v in synth.full_state_dict(load_spec(), 0).items()}
    res = {}
    for mode in ('fp32', 'autocast_bf16'):
        b = B
        while b >= 1:
            try:
                data = synth.tree_map(lambda t: t.to(dev), synth.gen_trajectories(T, b, N, 1234))
                times = []
                for i in range(3):
                    torch.cuda.synchronize()
                    e0, e1 = _events()
                    e0.record()
                    with torch.autocast('cuda', dtype=torch.bfloat16, enabled=(mode != 'fp32')):
                        out = R.rl_learner_forward(sd, ['winloss'], **synth.tree_map(lambda t: t.clone(), data))
                        out['target_logit'] = {k: v.float() for k, v in out['target_logit'].items()}
                        out['value'] = {k: v.float() for k, v in out['value'].items()}
                    info = R.compute_loss(C.DEFAULT_LOSS_CONFIG, out)
                    for v in sd.values():
                        v.grad = None
                    info['total_loss'].backward()
                    e1.record()
                    torch.cuda.synchronize()
                    times.append(e0.elapsed_time(e1))
                    del out, info
                ms = min(times[1:])
                res[mode] = {'value': T * b / ms * 1e3, 'unit': 'traj-steps/s', 'ms_per_step': ms, 'trajectories': b,
                             'unroll_len': T, 'peak_mem_gb': torch.cuda.max_memory_allocated(dev) / 1e9}
                break
            except torch.OutOfMemoryError:
                b //= 2
            finally:
                data = None
                for v in sd.values():
                    v.grad = None
                torch.cuda.empty_cache()
        if mode not in res:
            res[mode] = {'unavailable': 'out of memory down to 1 trajectory'}
    res['what'] = ('oracle/restate.py (plain torch restatement of the reference) on this GPU: ATen/cuBLAS/cuDNN kernels, '
                   'forward + loss + backward, TF32 off; best of 2 steps after 1 warm-up')
    return res


def encoder_microbench(model, dev, batch=1024, chunk=256, iters=3):
    """BASELINE.json configs[1]: entity Transformer + ScatterConnection + spatial encoder, batch 1024 observations x 512
    entities, 152x160 map, forward, observations resident on the device (processed in chunks of `chunk`).  The
    stand-alone ScatterConnection kernel (as_scatter_connection_fwd: [R,512,32] fp32 -> [R,32,152,160] fp32) is timed
    on its own against its algorithmic bytes, 3.179 MB per observation; inside the model the scatter is
    fused into the kernel that builds the spatial encoder's input."""
    from applestar_b200 import synth, ops, spatial, config as C
    hbm, _, tf_sus, which = peaks()
    enc = model.encoder
    obs = synth.tree_map(lambda t: t.to(dev), synth.gen_observations(chunk, 512, 77, entity_num_range=(512, 512), hidden=False))
    n_chunks = batch // chunk
    stages = {'entity_encoder': 0.0, 'scatter_project': 0.0, 'scatter_connection': 0.0, 'spatial_encoder': 0.0}

    def run(record):
        for _ in range(n_chunks):
            marks = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
            with torch.no_grad():
                marks[0].record()
                ent, _pooled = enc.entity_encoder(obs['entity_info'], obs['entity_num'])
                marks[1].record()
                proj = enc.scatter_project(ent)
                marks[2].record()
                smap = ops.scatter_connection(proj.float(), obs['entity_info']['x'], obs['entity_info']['y'],
                                              obs['entity_num'], C.SPATIAL_Y, C.SPATIAL_X)
                marks[3].record()
                pw = enc.spatial_encoder.project[0]
                x = spatial.spatial_front(proj, obs['entity_info']['x'], obs['entity_info']['y'], obs['entity_num'],
                                          obs['spatial_info'], pw.weight, pw.bias)
                enc.spatial_encoder(obs['spatial_info'], x)
                marks[4].record()
            del smap
            if record:
                torch.cuda.synchronize()
                for name, a, b in zip(stages, marks[:-1], marks[1:]):
                    stages[name] += a.elapsed_time(b)
    run(False)
    for _ in range(iters):
        run(True)
    stages = {k: v / iters for k, v in stages.items()}
    total = sum(stages.values())
    model_ms = total - stages['scatter_connection']          # the model path (scatter fused into the feature kernel)
    sc_bytes = batch * (512 * 32 * 4 + 512 * 2 + 32 * C.SPATIAL_Y * C.SPATIAL_X * 4)
    sc_gbs = sc_bytes / (stages['scatter_connection'] * 1e-3) / 1e9
    flops = batch * 5.127e9
    return {'metric': 'encoder_microbench_obs_per_sec', 'value': batch / model_ms * 1e3, 'unit': 'obs/s',
            'config': {'workload': 'entity_transformer+scatter_connection+spatial_encoder_B%d_N512_152x160' % batch,
                       'chunk': chunk, 'direction': 'forward', 'data': 'synthetic, resident'},
            'ms_per_stage': stages, 'ms_model_path': model_ms,
            'roofline': {'bound': 'tensor', 'achieved': flops / (model_ms * 1e-3) / 1e12, 'peak': tf_sus, 'unit': 'TFLOP/s',
                         'frac': flops / (model_ms * 1e-3) / 1e12 / tf_sus, 'flop_per_obs': 5.127e9,
                         'peak_source': which + ' bf16_tflops_sustained'},
            'scatter_connection': {'bound': 'hbm', 'kernel': 'as::scatter_fwd_kernel (stand-alone op)', 'achieved': sc_gbs,
                                   'peak': hbm, 'unit': 'GB/s', 'frac': sc_gbs / hbm, 'bytes_per_obs': sc_bytes / batch,
                                   'ms': stages['scatter_connection'], 'peak_source': which + ' hbm_gbs'}}


def hbm_rooflines(model, loss_fn, resident, dev):
    """CUDA-event time of the two HBM-bound kernel families the north star names, against their ALGORITHMIC bytes
    (SURVEY §8d): the ScatterConnection kernel at the learner's row count, and the fused loss (forward statistics pass
    + scan + gradient pass) on the step's own logits: 3 x 4 B per VALID logit (target + teacher read, gradient
    written; the -1e9 padding is not counted)."""
    from applestar_b200 import ops, config as C
    hbm, _, _, which = peaks()
    out = {}
    R = resident['entity_num'].shape[0]
    proj = torch.randn(R, 512, 32, device=dev)
    x, y, en = resident['entity_info']['x'], resident['entity_info']['y'], resident['entity_num']
    with torch.no_grad():
        for _ in range(2):
            ops.scatter_connection(proj, x, y, en, C.SPATIAL_Y, C.SPATIAL_X)
        e0, e1 = _events()
        e0.record()
        for _ in range(5):
            ops.scatter_connection(proj, x, y, en, C.SPATIAL_Y, C.SPATIAL_X)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    nbytes = R * (512 * 32 * 4 + 512 * 2 + 32 * C.SPATIAL_Y * C.SPATIAL_X * 4)
    out['scatter_connection'] = {'kernel': 'as::scatter_fwd_kernel', 'rows': R, 'ms': ms, 'bytes': nbytes,
                                 'achieved': nbytes / (ms * 1e-3) / 1e9, 'peak': hbm, 'unit': 'GB/s',
                                 'frac': nbytes / (ms * 1e-3) / 1e9 / hbm}
    del proj
    with torch.no_grad():
        fwd = model.rl_learner_forward(**resident, defer_checks=True)
    T, B = fwd['unroll_len'], fwd['batch_size']
    en_tb = resident['entity_num'][:T * B].view(T, B)
    sun = resident['selected_units_num']
    valid = T * B * (327 + 128 + 2 + C.SPATIAL_Y * C.SPATIAL_X) + int(en_tb.sum()) + int((torch.clamp(sun, min=1) * (en_tb + 1)).sum())
    leaf = dict(fwd)
    leaf['target_logit'] = {k: v.detach().clone().requires_grad_(True) for k, v in fwd['target_logit'].items()}
    leaf['value'] = {k: v.detach().clone().requires_grad_(True) for k, v in fwd['value'].items()}

    def once():
        info = loss_fn.compute_loss(leaf)
        info['total_loss'].backward()
        for d in (leaf['target_logit'], leaf['value']):
            for v in d.values():
                v.grad = None

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        once()
        once()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()                   # as in the learner step: the 6 launches replay from a CUDA graph
    with torch.cuda.graph(graph):
        once()
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = _events()
    e0.record()
    for _ in range(5):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    nbytes = 12.0 * valid
    out['rl_loss'] = {'kernel': 'as::loss_stats_small/big + loss_scan + loss_finalize + loss_grad_small/big (forward + backward)',
                      'valid_logits': valid, 'ms': ms, 'bytes': nbytes, 'achieved': nbytes / (ms * 1e-3) / 1e9,
                      'peak': hbm, 'unit': 'GB/s', 'frac': nbytes / (ms * 1e-3) / 1e9 / hbm,
                      'note': 'compute_loss + backward (6 launches) replayed from a CUDA graph, CUDA events around 5 replays'}
    out['peak_source'] = which + ' hbm_gbs'
    return out


def run_ours(args):
    import torch.distributed as dist
    from applestar_b200 import Model, ReinforcementLoss, synth, _lib
    from applestar_b200.learner import LearnerStep

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py (impl=ours) needs a B200; there is no CPU fallback')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    T, B, N = args.unroll, args.traj, args.entities

    model = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
    model.load_state_dict(synth.full_state_dict(load_spec(), 0), strict=False)
    model = model.to(dev)
    if args.workload == 'encoder_microbench':
        if rank == 0:
            mb = encoder_microbench(model, dev)
            mb.update({'n_gpus': 1, 'steps': 3, 'warmup': 1, 'ms_per_step': mb['ms_model_path'], 'higher_is_better': True,
                       'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic', 'gpu_launches': None})
            print(json.dumps(mb))
        return
    step_fn = LearnerStep(model, ReinforcementLoss({}, 'MP0'))
    host = synth.gen_trajectories(T, B, N, 1234 + rank)
    host = synth.tree_map(lambda t: t.pin_memory(), host)
    h2d_bytes = [0]

    def count(t):
        h2d_bytes[0] += t.numel() * t.element_size()
        return t
    synth.tree_map(count, host)
    dense_bytes = h2d_bytes[0]
    resident = synth.tree_map(lambda t: t.to(dev, non_blocking=True), host)
    # compact trajectory layout (applestar_b200/staging.py): the teacher's selected-units logits travel as their emitted
    # rows only and are expanded to the reference's dense -1e9-padded tensor on the device
    from applestar_b200 import staging as traj_staging
    packed = traj_staging.pack(host)
    h2d_bytes[0] = traj_staging.packed_bytes(packed)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        w1 = time.perf_counter()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item(), w0, w1

    from applestar_b200.learner import _tree_copy_
    step_fn(resident)                                     # first call captures the CUDA graph of the step
    static = step_fn.static_inputs(resident)

    def step_resident():
        step_fn(static if static is not None else resident)

    # End-to-end arm: every step's inputs travel host -> device from pinned memory and every step's statistics are read
    # back, all inside the timed region.  As in any learner with a prefetching data loader the copy of batch i+1
    # (copy stream, into a staging set of device buffers) overlaps the compute of batch i; a device-to-device copy
    # moves the staged batch into the step's (CUDA-graph) input buffers at the start of the step.
    copy_stream = torch.cuda.Stream(device=dev)
    # staging set with the layout of the graph's input arena: the staged batch moves in with ONE device-to-device copy
    from applestar_b200.learner import tree_arena
    static_flat = step_fn.static_arena(resident) if static is not None else None
    staging_flat, staging = tree_arena(resident, dev) if static is not None else (None, None)
    stager = traj_staging.Stager(packed, dev)
    ev_copied, ev_free = torch.cuda.Event(), torch.cuda.Event()

    def prefetch():
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_free)               # the previous staged batch has been consumed
            stager.load(packed, staging)                  # H2D, pinned -> device, non-blocking; + row expansion
            ev_copied.record(copy_stream)

    # the step's statistics (loss terms) travel device -> host every step: copied into pinned memory right after the
    # replay, consumed by the host while the next step runs (one step of lag, as an asynchronous logger would do it),
    # so the host never idles the GPU between two graph launches
    stats_host = [torch.empty(_lib.AS_LOSS_MAX_STATS, dtype=torch.float32).pin_memory() for _ in range(2)]
    stats_ev = [torch.cuda.Event(), torch.cuda.Event()]
    turn = [0]
    last_read = [0.0]

    def read_stats(slot):
        stats_ev[slot].synchronize()
        last_read[0] = float(stats_host[slot][0])

    def step_e2e():
        slot = turn[0] & 1
        if static is not None:
            cur = torch.cuda.current_stream()
            cur.wait_event(ev_copied)                     # this step's inputs have landed
            static_flat.copy_(staging_flat, non_blocking=True)   # D2D into the graph's input buffers (one flat copy)
            ev_free.record(cur)
            prefetch()                                    # next step's H2D runs under this step's compute
            info = step_fn(static)
        else:
            fresh = synth.tree_map(lambda t: torch.empty_like(t, device=dev), host)
            info = step_fn(stager.load(packed, fresh))
        stats_host[slot].copy_(info.stats_tensor(), non_blocking=True)
        stats_ev[slot].record()
        if turn[0] > 0:
            read_stats(slot ^ 1)                          # the previous step's statistics, now on the host
        turn[0] += 1

    if static is not None:
        ev_free.record(torch.cuda.current_stream())
        prefetch()                                        # prime the pipeline (one batch in flight at any time)

    for _ in range(max(args.warmup, 3)):
        step_resident()
    sampler = ClockSampler(local) if rank == 0 else None
    launches0 = _lib.LAUNCHES
    ms, w0, w1 = timed(step_resident, args.steps)
    launches = step_fn.launches_per_step or (_lib.LAUNCHES - launches0) // max(args.steps, 1)
    clocks = sampler.stop(w0, w1) if sampler else None
    step_e2e()
    ms_e2e, _, _ = timed(step_e2e, args.steps)
    read_stats((turn[0] - 1) & 1)                          # the last step's statistics (already on the host)

    # dominant kernel family = the tcgen05 GEMM: one instrumented step (outside the timed region)
    records = []
    orig = gemm_flops_hook(records)
    # eager (un-captured) so that every launch can be bracketed, and on ONE stream: a launch of the side-stream
    # branches (value towers, scalar encoder) would otherwise be timed while it waits for SMs held by the main stream.
    # The optimizer state is put back afterwards (this extra step is not part of the run).
    os.environ['APPLESTAR_SIDE_STREAM'] = '0'
    opt_state = [step_fn.flat.flat, step_fn.optimizer.m, step_fn.optimizer.v, step_fn.optimizer.t_dev] + (
        [step_fn.optimizer.p_step] if getattr(step_fn.optimizer, 'p_step', None) is not None else [])
    keep = [t.clone() for t in opt_state]
    step_fn._run(resident)
    torch.cuda.synchronize()
    for dst, src in zip(opt_state, keep):
        dst.copy_(src)
    step_fn.flat.refresh_shadow()
    del keep
    os.environ['APPLESTAR_SIDE_STREAM'] = '1'
    _lib.call = orig
    g_ms = sum(r[0].elapsed_time(r[1]) for r in records)
    g_fl = sum(r[2] for r in records)
    su_mean = float(torch.clamp(host['selected_units_num'], min=1).float().mean())
    alg_fl = algorithmic_flops(T, B, su_mean)
    if os.environ.get('APPLESTAR_GEMM_DUMP') and rank == 0:
        agg = {}
        for r in records:
            k = r[3]
            t = agg.setdefault(k, [0, 0.0, 0.0])
            t[0] += 1; t[1] += r[0].elapsed_time(r[1]); t[2] += r[2]
        for k, t in sorted(agg.items(), key=lambda kv: -kv[1][1])[:120]:
            print('GEMM %-22s args=%s n=%d ms=%.3f TF/s=%.0f' % (k[0], k[1:], t[0], t[1], t[2] / t[1] / 1e9), file=sys.stderr)
    hbm, tf_burst, tf_sus, which = peaks()
    achieved = alg_fl / (g_ms * 1e-3) / 1e12 if g_ms > 0 else 0.0            # ALGORITHMIC flops
    executed = g_fl / (g_ms * 1e-3) / 1e12 if g_ms > 0 else 0.0              # incl. zero borders / padded channels
    traffic = traffic_note = None             # DRAM bytes of the same kernel family from the committed ncu capture
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'profiles', 'tensor_family_traffic.json')) as f:
            tj = json.load(f)
        traffic, traffic_note = tj['dram_bytes_per_launch_mean'], '%s: %.1f GB per step over %d launches' % (
            tj['source'], tj['dram_bytes_per_step'] / 1e9, tj['launches'])
    except (OSError, KeyError, ValueError):
        pass

    if rank == 0:
        total_steps = T * B * world
        value = total_steps / (ms / args.steps * 1e-3)
        e2e = total_steps / (ms_e2e / args.steps * 1e-3)
        line = {
            'metric': 'rl_learner_traj_steps_per_sec', 'value': value, 'unit': 'traj-steps/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': ms / args.steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
            'config': {'workload': 'rl_learner_step_T%d_B%d_N%d' % (T, B, N), 'trajectories_per_gpu': B,
                       'unroll_len': T, 'entities': N, 'baselines': ['winloss'], 'parallelism': 'dp%d' % world,
                       'l2_policy': 'inputs_larger_than_l2 (%.0f MB per step)' % (h2d_bytes[0] / 1e6),
                       'cuda_graph': bool(step_fn.use_graph and step_fn.launches_per_step),
                       'e2e_pipeline': 'H2D of batch i+1 on a copy stream under the compute of batch i (staging buffers + D2D); the statistics of step i are copied to pinned memory after its replay and read by the host during step i+1',
                       'e2e_layout': 'compact teacher selected-units rows, expanded on the device (%.0f MB instead of %.0f MB)' % (
                           h2d_bytes[0] / 1e6, dense_bytes / 1e6)},
            'e2e': {'value': e2e, 'unit': 'traj-steps/s', 'h2d_bytes_per_step': h2d_bytes[0],
                    'd2h_bytes_per_step': 4 * _lib.AS_LOSS_MAX_STATS},
            'gpu_launches': int(launches),
            'clocks': clocks,
            'roofline': {'bound': 'tensor', 'kernel': 'tcgen05 kernels as::gemm_kernel + as::attn_kernel + as::wgrad_slab_kernel (all %d launches of one step)' % len(records),
                         'achieved': achieved, 'peak': tf_sus, 'unit': 'TFLOP/s', 'frac': achieved / tf_sus,
                         'flops': 'algorithmic: %.2f TFLOP per step (SURVEY 8d: dense forward x 3, H x W pixels, real K)' % (alg_fl / 1e12),
                         'achieved_executed': executed, 'frac_executed': executed / tf_sus,
                         'frac_of_whole_step': alg_fl / (ms / args.steps * 1e-3) / 1e12 / tf_sus,
                         'peak_source': which + ' bf16_tflops_sustained', 'gemm_ms_per_step': g_ms,
                         'gemm_share_of_step': g_ms / (ms / args.steps), 'traffic': traffic, 'traffic_source': traffic_note},
        }
        if not args.no_cpu_baseline and world == 1:      # the CPU baseline is an N = 1 figure
            Tc, Bc = (int(x) for x in args.cpu_sample.split('x'))
            threads = os.cpu_count() or 1
            v, dt = cpu_learner_throughput(Tc, Bc, N, 2, 1, threads)
            line['cpu_baseline'] = {'value': v, 'unit': 'traj-steps/s', 'cores': threads, 'kind': 'port',
                                    'sample': 'oracle/restate.py forward+loss+backward, T=%d B=%d N=%d, 2 steps' % (Tc, Bc, N)}
        if not args.no_inference and world == 1:
            line['inference'] = actor_inference(model, dev, cpu_baseline=not args.no_cpu_baseline)
        if not args.no_extras and world == 1:
            line['roofline_hbm'] = hbm_rooflines(model, step_fn.loss, resident, dev)
            line['encoder_microbench'] = encoder_microbench(model, dev)
            step_fn._graphs.clear()                       # free the graph pool before the library path needs its memory
            torch.cuda.empty_cache()
            line['gpu_library_baseline'] = gpu_library_baseline(T, B, N, dev)
        print(json.dumps(line))
    if world > 1:
        # tear-down: the graphs hold captured NCCL kernels, so they go first (all ranks, after a common barrier),
        # then the process group
        dist.barrier()
        torch.cuda.synchronize()
        step_fn._graphs.clear()
        step_fn._pool = None
        import gc
        gc.collect()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        # a tear-down that does not finish within a minute (a peer that already left) must not hang the job: the
        # numbers are out, so the watchdog ends the process
        guard = threading.Timer(60.0, lambda: os._exit(0))
        guard.daemon = True
        guard.start()
        dist.destroy_process_group()
        guard.cancel()


def main():
    args = parse()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
