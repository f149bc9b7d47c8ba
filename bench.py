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
        if name not in ('as_gemm_bf16', 'as_conv_nhwc_bf16', 'as_conv_nhwc_bf16_add', 'as_conv_wgrad_nhwc_bf16', 'as_attn_fwd',
                        'as_attn_bwd_dq'):
            return orig(name, *a)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(120000)         # keep the stream busy while the host enqueues, so that e0 -> e1 is the kernel's
        e0.record()                       # own duration and not the host's launch latency (tiny GEMMs of the heads)
        orig(name, *a)
        e1.record()
        if name == 'as_gemm_bf16':
            M, N, K, nb = a[5], a[6], a[7], a[13] * a[14]
            fl = 2.0 * M * N * K * nb
        elif name == 'as_conv_nhwc_bf16_add':             # (x, w, y, bias, addend, R, H, W, Cin, Cout, taps, ...)
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


def actor_inference(model, dev, batch=256, iters=8):
    """BASELINE.json's second metric (reported per GPU): `compute_logp_action` on `batch` observations x 512 entities
    through learner.ActorStep (CUDA graph, all 64 sampling steps inside), observations resident on the device."""
    from applestar_b200 import synth
    from applestar_b200.learner import ActorStep
    actor = ActorStep(model)
    obs = synth.tree_map(lambda t: t.to(dev), synth.gen_observations(batch, 512, 7))
    hidden = [(torch.zeros(batch, 384, device=dev), torch.zeros(batch, 384, device=dev)) for _ in range(3)]
    noise = synth.tree_map(lambda t: t.to(dev), synth.gen_gumbel(batch, 11))

    def call():
        return actor(obs['spatial_info'], obs['entity_info'], obs['scalar_info'], obs['entity_num'], hidden, gumbel_noise=noise)

    for _ in range(3):
        call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    return {'metric': 'inference_obs_per_sec_per_gpu', 'value': batch / ms * 1e3, 'unit': 'obs/s', 'ms_per_call': ms,
            'config': {'workload': 'compute_logp_action_B%d_N512' % batch, 'cuda_graph': True,
                       'launches_per_call': actor.launches_per_call}}


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
    step_fn = LearnerStep(model, ReinforcementLoss({}, 'MP0'))
    host = synth.gen_trajectories(T, B, N, 1234 + rank)
    host = synth.tree_map(lambda t: t.pin_memory(), host)
    h2d_bytes = [0]

    def count(t):
        h2d_bytes[0] += t.numel() * t.element_size()
        return t
    synth.tree_map(count, host)
    resident = synth.tree_map(lambda t: t.to(dev, non_blocking=True), host)

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
    staging = synth.tree_map(lambda t: torch.empty_like(t, device=dev), host) if static is not None else None
    ev_copied, ev_free = torch.cuda.Event(), torch.cuda.Event()

    def prefetch():
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_free)               # the previous staged batch has been consumed
            _tree_copy_(staging, host)                    # H2D, pinned -> device, non-blocking
            ev_copied.record(copy_stream)

    def step_e2e():
        if static is not None:
            cur = torch.cuda.current_stream()
            cur.wait_event(ev_copied)                     # this step's inputs have landed
            _tree_copy_(static, staging)                  # D2D into the graph's input buffers
            ev_free.record(cur)
            prefetch()                                    # next step's H2D runs under this step's compute
            info = step_fn(static)
        else:
            info = step_fn(synth.tree_map(lambda t: t.to(dev, non_blocking=True), host))
        return info['winloss/td']                         # forces the device->host read of the step's result

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

    # dominant kernel family = the tcgen05 GEMM: one instrumented step (outside the timed region)
    records = []
    orig = gemm_flops_hook(records)
    step_fn._run(resident)                                # eager (un-captured) so that every launch can be bracketed
    torch.cuda.synchronize()
    _lib.call = orig
    g_ms = sum(r[0].elapsed_time(r[1]) for r in records)
    g_fl = sum(r[2] for r in records)
    if os.environ.get('APPLESTAR_GEMM_DUMP') and rank == 0:
        agg = {}
        for r in records:
            k = r[3]
            t = agg.setdefault(k, [0, 0.0, 0.0])
            t[0] += 1; t[1] += r[0].elapsed_time(r[1]); t[2] += r[2]
        for k, t in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
            print('GEMM %-22s args=%s n=%d ms=%.3f TF/s=%.0f' % (k[0], k[1:], t[0], t[1], t[2] / t[1] / 1e9), file=sys.stderr)
    hbm, tf_burst, tf_sus, which = peaks()
    achieved = g_fl / (g_ms * 1e-3) / 1e12 if g_ms > 0 else 0.0
    traffic = traffic_note = None             # DRAM bytes of the same kernel family from the committed ncu capture
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'profiles', 'r01_tensor_family_traffic.json')) as f:
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
                       'e2e_pipeline': 'H2D of batch i+1 on a copy stream under the compute of batch i (staging buffers + D2D)'},
            'e2e': {'value': e2e, 'unit': 'traj-steps/s', 'h2d_bytes_per_step': h2d_bytes[0],
                    'd2h_bytes_per_step': 4 * _lib.AS_LOSS_MAX_STATS},
            'gpu_launches': int(launches),
            'clocks': clocks,
            'roofline': {'bound': 'tensor', 'kernel': 'tcgen05 kernels as::gemm_kernel + as::attn_kernel + as::wgrad_slab_kernel (all %d launches of one step)' % len(records),
                         'achieved': achieved, 'peak': tf_sus, 'unit': 'TFLOP/s', 'frac': achieved / tf_sus,
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
            line['inference'] = actor_inference(model, dev)
        print(json.dumps(line))
    if world > 1:
        # tear-down: captured NCCL kernels inside CUDA graphs can stall destroy_process_group(); the numbers are out,
        # so synchronise, drop the graphs and leave without the collective tear-down
        dist.barrier()
        torch.cuda.synchronize()
        step_fn._graphs.clear()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    args = parse()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
