"""Location head in isolation: gradients of the CUDA path and of the oracle under CPU bf16 autocast, both against the
fp32 oracle, on the same random inputs (diagnostic for the parity budget of the location-head tensors)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
from applestar_b200 import Model, spatial, synth  # noqa: E402
from oracle import restate as R  # noqa: E402
from util import load_spec  # noqa: E402


def loss_of(logits, act, coef):
    lp = torch.log_softmax(logits.float(), -1)
    return -(coef * lp.gather(-1, act.unsqueeze(-1)).squeeze(-1)).mean() + 1e-3 * (lp.exp() * lp).sum(-1).mean()


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    sd = synth.full_state_dict(load_spec('winloss'), 0)
    g = torch.Generator().manual_seed(1)
    emb = torch.randn(M, 1024, generator=g) * 0.5
    skips = [torch.relu(torch.randn(M, 128, 19, 20, generator=g)) for _ in range(4)]
    act = torch.randint(0, 24320, (M,), generator=g)
    coef = torch.randn(M, generator=g) * 0.1
    keys = [k for k in sd if k.startswith('policy.location_head.')]

    def run_oracle(autocast):
        s = {k: (v.clone().requires_grad_(True) if k in keys else v) for k, v in sd.items()}
        e = emb.clone().requires_grad_(True)
        sk = [None, None, None] + [t.clone().requires_grad_(True) for t in skips]
        with torch.autocast('cpu', dtype=torch.bfloat16, enabled=autocast):
            logits, _ = R.location_head(s, e, sk, location=act)
        loss_of(logits, act, coef).backward()
        out = {k: s[k].grad for k in keys}
        out['emb'] = e.grad
        for i in range(4):
            out['skip%d' % i] = sk[3 + i].grad
        return out, logits.detach().float()

    ref, ref_logits = run_oracle(False)
    bf, bf_logits = run_oracle(True)
    m = Model({'model': {'enable_baselines': ['winloss']}}, use_value_network=True)
    m.load_state_dict(sd, strict=False)
    head = m.policy.location_head.cuda()
    e = emb.cuda().requires_grad_(True)
    sk32 = [t.cuda().requires_grad_(True) for t in skips]
    sk = [None, None, None] + [spatial.to_pnhwc(t) for t in sk32]
    logits, _ = head(e, sk, 1.0, act.cuda())
    loss_of(logits, act.cuda(), coef.cuda()).backward()
    mine = {'policy.location_head.' + k: p.grad.cpu() for k, p in head.named_parameters()}
    mine['emb'] = e.grad.cpu()
    for i in range(4):
        mine['skip%d' % i] = sk32[i].grad.cpu()
    print('logits: ours max %.4f, autocast max %.4f' % ((logits.float().cpu() - ref_logits).abs().max(),
                                                       (bf_logits - ref_logits).abs().max()))
    print('%-55s %9s %9s' % ('gradient', 'ours rel', 'autocast'))
    for k in list(ref):
        r = ref[k].double()
        a = ((mine[k].double() - r).norm() / r.norm()).item()
        b = ((bf[k].double() - r).norm() / r.norm()).item()
        print('%-55s %9.4f %9.4f' % (k, a, b))


if __name__ == '__main__':
    main()
