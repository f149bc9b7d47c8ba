"""Fused attention forward (as_attn_fwd) check + timing against the composed GEMM/softmax path.
usage: python tools/attn_check.py [R] [N]"""
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from applestar_b200 import _lib, kernels as K  # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 64
N = int(sys.argv[2]) if len(sys.argv) > 2 else 512
H, D = 2, 128
g = torch.Generator().manual_seed(1)
qkv = (torch.randn(R, N, 3 * H * D, generator=g) * 0.7).cuda().to(torch.bfloat16)
key_len = torch.randint(N // 2, N + 1, (R,), generator=g).cuda()
scale = 1.0 / math.sqrt(D)
O = torch.empty(R, N, H * D, dtype=torch.bfloat16, device='cuda')
P = torch.empty(R, H, N, N, dtype=torch.bfloat16, device='cuda')
torch.cuda.synchronize()
_lib.call('as_attn_fwd', qkv.data_ptr(), key_len.data_ptr(), O.data_ptr(), P.data_ptr(), R, N, H, D, scale, _lib.stream_ptr())
torch.cuda.synchronize()
print('launched ok')
Rc = min(R, 8)
q, k, v = qkv[:Rc].float().chunk(3, dim=2)
q = q.view(Rc, N, H, D).permute(0, 2, 1, 3)
k = k.view(Rc, N, H, D).permute(0, 2, 3, 1)
v = v.view(Rc, N, H, D).permute(0, 2, 1, 3)
s = torch.matmul(q, k) * scale
m = torch.arange(N, device='cuda').view(1, 1, 1, N) < key_len[:Rc].view(Rc, 1, 1, 1)
pr = torch.softmax(s.masked_fill(~m, -1e9), -1)
orf = torch.matmul(pr, v).permute(0, 2, 1, 3).reshape(Rc, N, H * D)
print('P max err', (P[:Rc].float() - pr).abs().max().item(), 'O max err', (O[:Rc].float() - orf).abs().max().item())


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


t_p = timeit(lambda: _lib.call('as_attn_fwd', qkv.data_ptr(), key_len.data_ptr(), O.data_ptr(), P.data_ptr(), R, N, H, D, scale,
                               _lib.stream_ptr()))
t_np = timeit(lambda: _lib.call('as_attn_fwd', qkv.data_ptr(), key_len.data_ptr(), O.data_ptr(), 0, R, N, H, D, scale,
                                _lib.stream_ptr()))
K.ATTN_FUSED = False
with torch.no_grad():
    t_old = timeit(lambda: K.attention(qkv, key_len, H, D))
flops = 4.0 * R * H * N * N * D
print('fused+P %.3f ms (%.0f TF/s)  fused %.3f ms (%.0f TF/s)  composed %.3f ms' % (t_p, flops / t_p / 1e9, t_np, flops / t_np / 1e9, t_old))

prof = torch.zeros(8, dtype=torch.int64, device='cuda')
_lib.lib().as_debug_attn_profile.argtypes = [__import__('ctypes').c_void_p]
_lib.lib().as_debug_attn_profile(prof.data_ptr())
_lib.call('as_attn_fwd', qkv.data_ptr(), key_len.data_ptr(), O.data_ptr(), P.data_ptr(), R, N, H, D, scale, _lib.stream_ptr())
torch.cuda.synchronize()
_lib.lib().as_debug_attn_profile(0)
pr = prof.tolist()
n = max(pr[0], 1)
print('fwd CTA0 items %d: pass1(incl. wait S0) %.0f  exchange+pass2 %.0f  wait O %.0f  store %.0f  gap to next item %.0f clk'
      % (pr[0], pr[1] / n, pr[2] / n, pr[3] / n, pr[4] / n, pr[5] / max(n - 1, 1)))
