"""Per-shape timing of the tcgen05 GEMM on the learner step's dominant shapes (CUDA events, L2 flushed between
iterations).  usage: python tools/bench_gemm.py > gpurun_out/gemm_shapes.txt"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from applestar_b200 import _lib, kernels as K  # noqa: E402

dev = torch.device('cuda:0')
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)


def timeit(fn, iters=5):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def bf(*shape):
    return torch.randn(*shape, device=dev).to(torch.bfloat16)


rows = []


def run(name, M, N, Kd, a_mn=False, b_mn=False, out_bf16=True, splits=1, batch=(1, 1)):
    nb = batch[0] * batch[1]
    A = bf(nb, Kd, M) if a_mn else bf(nb, M, Kd)
    B = bf(nb, Kd, N) if b_mn else bf(nb, N, Kd)
    C = torch.empty((nb, M, N), device=dev, dtype=torch.bfloat16 if out_bf16 else torch.float32)
    ws = None
    if splits > 1:
        splits = _lib.lib().as_gemm_splits(Kd, splits)
        ws = torch.empty((splits, nb, M, N), device=dev)
    sa = (M * Kd, M * Kd * batch[0])
    sb = (N * Kd, N * Kd * batch[0])
    sc = (M * N, M * N * batch[0])

    def fn():
        K.gemm(A, B, None if splits > 1 else C, M, N, Kd, a_mn=a_mn, b_mn=b_mn, batch=batch, a_strides=sa, b_strides=sb,
               c_strides=sc, splits=splits, ws=ws)
    ms = timeit(fn)
    fl = 2.0 * M * N * Kd * nb
    by = nb * (M * Kd * 2 + N * Kd * 2 + (M * N * (2 if out_bf16 else 4) if splits == 1 else splits * M * N * 4))
    rows.append({'name': name, 'M': M, 'N': N, 'K': Kd, 'batch': nb, 'a_mn': a_mn, 'b_mn': b_mn, 'splits': splits,
                 'ms': ms, 'tflops': fl / ms / 1e9, 'gbs': by / ms / 1e6})
    print('%-34s M=%-7d N=%-5d K=%-7d nb=%-5d %s%s sp=%-3d %8.3f ms %8.1f TF/s %8.1f GB/s' % (
        name, M, N, Kd, nb, 'T' if a_mn else 'N', 'T' if b_mn else 'N', splits, ms, fl / ms / 1e9, by / ms / 1e6), flush=True)


R = 1040
M = R * 512
run('square 8192 (peak check)', 8192, 8192, 8192)
run('entity embed  K=1000', M, 256, 1000, out_bf16=False)
run('qkv           256->768', M, 768, 256)
run('attn proj     256->256 f32', M, 256, 256, out_bf16=False)
run('ffn1          256->1024', M, 1024, 256)
run('ffn2          1024->256 f32', M, 256, 1024, out_bf16=False)
run('scatter_proj  256->32 f32', M, 32, 256, out_bf16=False)
run('dgrad qkv     768->256 (B mn)', M, 256, 768, b_mn=True)
run('dgrad ffn1    1024->256 (B mn)', M, 256, 1024, b_mn=True)
run('wgrad qkv     [768x256] K=M', 768, 256, M, a_mn=True, b_mn=True, out_bf16=False, splits=24)
run('wgrad ffn1    [1024x256] K=M', 1024, 256, M, a_mn=True, b_mn=True, out_bf16=False, splits=18)
run('attn QK^T     512x512x128', 512, 512, 128, out_bf16=False, batch=(2, R))
run('attn PV       512x128x512 (B mn)', 512, 128, 512, b_mn=True, batch=(2, R))
run('attn dV       P^T dO (A mn,B mn)', 512, 128, 512, a_mn=True, b_mn=True, batch=(2, R))
run('spatial fc    59136->256', R, 256, 59136, out_bf16=False)
run('lstm gates    1536->1536', R, 1536, 1536, out_bf16=False)
run('value fc      256->256 M=1040', R, 256, 256, out_bf16=False)
# implicit-GEMM conv 3x3 128->128 at 19x20 (padded 21x22)
x = bf(R, 21, 22, 128)
w = bf(128, 9 * 128)
y = torch.empty_like(x)


def conv():
    _lib.call('as_conv_nhwc_bf16', x.data_ptr(), w.data_ptr(), y.data_ptr(), 0, R, 19, 20, 128, 128, 9, 0, 1, 1,
              _lib.stream_ptr())


ms = timeit(conv)
fl = 2.0 * R * 462 * 128 * 128 * 9
print('%-34s %8.3f ms %8.1f TF/s (padded pixels)' % ('conv3x3 128->128 19x20', ms, fl / ms / 1e9))
rows.append({'name': 'conv3x3 128->128 19x20', 'ms': ms, 'tflops': fl / ms / 1e9})
print(json.dumps(rows))
