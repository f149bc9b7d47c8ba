"""Per-tensor parity table of the CUDA learner path against tests/golden/learner_n512.pt (the executed reference at
N = 512): logits (max-abs, RMS), loss scalars and every stored gradient tensor (relative L2, cosine) next to the
reference's own bf16-autocast error.  Writes the table to stdout; `profiles/r02_parity_n512.txt` is its output.
    python tools/parity_report.py [T8]      (T8: all 394 gradients against the oracle on the CPU, T = 8, B = 2)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
from applestar_b200 import Model, ReinforcementLoss, config as C, synth  # noqa: E402
from util import golden, grad_group_check, grad_sample, load_spec  # noqa: E402


def main():
    t8 = len(sys.argv) > 1 and sys.argv[1] == 'T8'
    g = golden('learner_n512_t8_budget.pt' if t8 else 'learner_n512.pt')
    sd = synth.full_state_dict(load_spec('winloss'), 0)
    m = Model({'model': {'enable_baselines': ['winloss']}, 'agent': {'extra_units': True}}, use_value_network=True)
    m.load_state_dict(sd, strict=False)
    m = m.cuda()
    data = synth.gen_trajectories(8, 2, 512, 8512, special=True) if t8 else \
        synth.gen_trajectories(g['T'], g['B'], g['N'], g['seed'], special=True)
    out = m.rl_learner_forward(**synth.tree_map(lambda t: t.cuda(), data))
    info = ReinforcementLoss({}, 'MP0').compute_loss(out)
    info['total_loss'].backward()
    params = dict(m.named_parameters())
    if t8:
        from oracle import restate as R
        sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        ref = R.rl_learner_forward(sdr, ['winloss'], **synth.tree_map(lambda t: t.clone(), data))
        R.compute_loss(C.DEFAULT_LOSS_CONFIG, ref)['total_loss'].backward()
        refs = {k: (v.grad if v.grad is not None else torch.zeros_like(v)).reshape(-1) for k, v in sdr.items()
                if k in params and params[k].requires_grad}
        mine = {k: params[k].grad.detach().float().cpu().reshape(-1) for k in refs}
    else:
        print('%-18s %10s %10s %10s %10s' % ('logits', 'max-abs', 'ref bf16', 'rms', 'ref bf16'))
        for k in C.HEADS:
            a, r64 = out['target_logit'][k].detach().double().cpu(), g['logit64'][k]
            d = (a - r64)[r64 > -1e8]
            print('%-18s %10.4f %10.4f %10.5f %10.5f' % (k, d.abs().max(), g['bf16_autocast_err'][k][0],
                                                        d.pow(2).mean().sqrt(), g['bf16_autocast_err'][k][1]))
        for f, v in out['value'].items():
            d = v.detach().double().cpu() - g['value64'][f]
            print('%-18s %10.4f %10.4f %10.5f %10.5f' % ('value/' + f, d.abs().max(), g['bf16_autocast_err']['value/' + f][0],
                                                        d.pow(2).mean().sqrt(), g['bf16_autocast_err']['value/' + f][1]))
        print('\n%-28s %12s %12s %12s' % ('loss scalar', 'ours', 'reference', 'ref bf16'))
        for k, v in g['loss_info'].items():
            print('%-28s %12.6f %12.6f %12.6f' % (k, float(info[k].detach() if hasattr(info[k], 'detach') else info[k]), v, g['bf16_autocast_loss_info'][k]))
        refs = g['grads']
        mine = {k: grad_sample(params[k].grad) for k in refs}
    rows = []
    for k, r in refs.items():
        a, b = mine[k].double(), r.double()
        rel = ((a - b).norm() / b.norm().clamp(min=1e-30)).item()
        cos = (torch.dot(a, b) / (a.norm() * b.norm()).clamp(min=1e-30)).item()
        rb, cb = (float(x) for x in g['bf16_autocast_grad_err'][k])
        rows.append((rel / max(rb, 1e-9), k, rel, rb, cos, cb, b.norm().item()))
    rows.sort(reverse=True)
    print('\n%d gradient tensors; relative L2 / cosine of ours vs fp32 reference, next to the reference bf16-autocast run'
          % len(rows))
    print('%-75s %8s %8s %8s %8s %10s' % ('tensor', 'rel', 'ref rel', 'cos', 'ref cos', '|g|'))
    for ratio, k, rel, rb, cos, cb, n in rows:
        print('%-75s %8.4f %8.4f %8.4f %8.4f %10.3e' % (k, rel, rb, cos, cb, n))
    print('\nmodule-level aggregates (relative L2 of the concatenated gradient; budget = 2 x reference bf16)')
    for grp, rel, rel_b in grad_group_check(mine, refs, g['bf16_autocast_grad_err'], 0.0)[1]:
        print('%-40s ours %8.4f   reference bf16 %8.4f' % (grp, rel, rel_b))
    import statistics
    print('\nmedian rel ours %.4f, reference bf16 %.4f' % (statistics.median(r[2] for r in rows),
                                                         statistics.median(r[3] for r in rows)))


if __name__ == '__main__':
    main()
