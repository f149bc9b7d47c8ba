"""Diagnostics: error of the learner forward vs the fp64 reference, as a fraction of the tolerance budget
(2 x the reference's own bf16-autocast error) -- run on the GPU box."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch
from applestar_b200 import config as C, synth, kernels
from test_model_gpu import _model, golden
g = golden('learner_small.pt')
model, sd = _model(g['baselines'], 'four')
data = synth.gen_trajectories(g['T'], g['B'], g['N'], g['seed'])
out = model.rl_learner_forward(**synth.tree_map(lambda t: t.cuda(), data))
print('score dtype', kernels.SCORE_DTYPE)
for k in C.HEADS:
    a = out['target_logit'][k].detach().double().cpu(); ref = g['logit64'][k]; m = ref <= -1e8
    err = (a - ref)[~m].abs().max().item(); bud = 2.0 * g['bf16_autocast_err'][k][0].item()
    print('%-18s err %.4g  budget %.4g  ratio %.2f' % (k, err, bud, err / bud))
for f, v in out['value'].items():
    err = (v.detach().double().cpu() - g['value64'][f]).abs().max().item(); bud = 2.0 * g['bf16_autocast_err']['value/' + f][0].item()
    print('%-18s err %.4g  budget %.4g  ratio %.2f' % ('value/' + f, err, bud, err / bud))
