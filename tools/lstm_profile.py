"""Diagnostics: cycle counts of the cluster LN-LSTM forward phases (run on the GPU box)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from applestar_b200 import _lib, kernels as K
from applestar_b200.model import LayerNormLSTMCell
lib = _lib.lib()
T, B, H = 64, int(os.environ.get('B', 16)), 384
cell = LayerNormLSTMCell(H, H).cuda()
x = torch.randn(T, B, H, device='cuda')
h0 = torch.zeros(B, H, device='cuda'); c0 = torch.zeros(B, H, device='cuda')
for _ in range(2):
    K.lnlstm_layer(x, h0, c0, cell)
nblk = ((B + 1) // 2 if B > 15 else B) * 8
buf = torch.zeros(nblk * 8, dtype=torch.int64, device='cuda')
lib.as_debug_lstm_profile.argtypes = [ctypes.c_void_p]
lib.as_debug_lstm_profile(ctypes.c_void_p(buf.data_ptr()))
K.lnlstm_layer(x, h0, c0, cell)
torch.cuda.synchronize()
lib.as_debug_lstm_profile(ctypes.c_void_p(0))
v = buf.view(nblk, 8).double().cpu() / T
names = ['matvec+shfl', 'stage+sync', 'remote stores', 'cluster barrier', 'LN_h sums', 'cell+LN_c+store+sync']
print('cycles per step (mean over CTAs, min, max):')
for i, n in enumerate(names):
    print('  %-24s %8.0f %8.0f %8.0f' % (n, v[:, i].mean(), v[:, i].min(), v[:, i].max()))
print('  total %.0f cycles/step' % v[:, :6].sum(1).mean())
