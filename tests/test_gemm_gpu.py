"""GPU parity of the tcgen05 GEMM (all operand-major combinations, batching, split-K, epilogues, conv mode)
against a plain PyTorch fp32 reference computed from the same bf16-rounded operands.
Tolerance: fp32 accumulation of bf16 products -> |err| <= 2e-3 * sqrt(K) * scale (stated per test)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(shape, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(shape, generator=g).cuda().to(torch.bfloat16)


def _tol(K):
    return 4e-3 * math.sqrt(K)


@pytest.mark.parametrize('a_mn', [False, True])
@pytest.mark.parametrize('b_mn', [False, True])
@pytest.mark.parametrize('shape', [(256, 256, 256), (300, 72, 136), (128, 768, 64), (1000, 327 + 1, 448), (64, 32, 32)])
def test_gemm_majors(a_mn, b_mn, shape):
    from applestar_b200 import kernels as K
    M, N, Kd = shape
    if (b_mn and N % 8) or (a_mn and M % 8):
        pytest.skip('pitch of the MN-major operand must be a multiple of 8 elements')
    A = _mk((Kd, M) if a_mn else (M, Kd), 1)
    B = _mk((Kd, N) if b_mn else (N, Kd), 2)
    C = torch.full((M, N), float('nan'), device='cuda')
    K.gemm(A, B, C, M, N, Kd, a_mn=a_mn, b_mn=b_mn)
    Af = A.float().t() if a_mn else A.float()
    Bf = B.float() if b_mn else B.float().t()
    ref = Af @ Bf
    torch.cuda.synchronize()
    assert torch.isfinite(C).all()
    assert (C - ref).abs().max().item() < _tol(Kd)


def test_gemm_epilogue_bias_relu_bf16_alpha():
    from applestar_b200 import kernels as K
    M, N, Kd = 384, 200, 512
    A, B = _mk((M, Kd), 3), _mk((N, Kd), 4)
    bias = torch.randn(N, device='cuda')
    C = torch.empty((M, N), device='cuda', dtype=torch.bfloat16)
    K.gemm(A, B, C, M, N, Kd, bias=bias, relu=True, alpha=0.5)
    ref = torch.relu(0.5 * (A.float() @ B.float().t()) + bias)
    assert (C.float() - ref).abs().max().item() < _tol(Kd) + 0.01 * ref.abs().max().item()


def test_gemm_batched_strided_attention_layout():
    """Q K^T per (obs, head) straight out of a packed qkv buffer, and P V with V read MN-major."""
    from applestar_b200 import kernels as K
    R, Nn, H, D = 3, 256, 2, 128
    qkv = _mk((R, Nn, 3 * H * D), 5)
    S = torch.empty((R, H, Nn, Nn), device='cuda')
    q = qkv[:, :, :H * D]
    k = qkv[:, :, H * D:2 * H * D]
    v = qkv[:, :, 2 * H * D:]
    ld = 3 * H * D
    K.gemm(q, k, S, Nn, Nn, D, lda=ld, ldb=ld, ldc=Nn, batch=(H, R), a_strides=(D, Nn * ld), b_strides=(D, Nn * ld),
           c_strides=(Nn * Nn, H * Nn * Nn), alpha=1.0 / math.sqrt(D))
    ref = torch.einsum('rqhd,rkhd->rhqk', q.float().view(R, Nn, H, D), k.float().view(R, Nn, H, D)) / math.sqrt(D)
    assert (S - ref).abs().max().item() < _tol(D)
    P = torch.softmax(S, dim=-1).to(torch.bfloat16)
    O = torch.empty((R, Nn, H * D), device='cuda')
    K.gemm(P, v, O, Nn, D, Nn, lda=Nn, b_mn=True, ldb=ld, ldc=H * D, batch=(H, R),
           a_strides=(Nn * Nn, H * Nn * Nn), b_strides=(D, Nn * ld), c_strides=(D, Nn * H * D))
    ref_o = torch.einsum('rhqk,rkhd->rqhd', P.float(), v.float().view(R, Nn, H, D)).reshape(R, Nn, H * D)
    assert (O - ref_o).abs().max().item() < 2e-2


def test_gemm_splitk_wgrad():
    from applestar_b200 import kernels as K
    M, N, Kd = 5000, 96, 256            # dW[N,Kd] = dY^T X over M rows
    dy, x = _mk((M, N), 6), _mk((M, Kd), 7)
    dw = K._wgrad(dy, x, N, Kd)
    ref = dy.float().t() @ x.float()
    assert (dw - ref).abs().max().item() < _tol(M)


def test_linear_autograd_matches_torch():
    from applestar_b200 import kernels as K
    g = torch.Generator().manual_seed(8)
    x = torch.randn((77, 5, 90), generator=g).cuda().requires_grad_(True)        # K = 90 -> padded to 96
    w = (torch.randn((130, 90), generator=g) * 0.1).cuda().requires_grad_(True)
    b = torch.randn(130, generator=g).cuda().requires_grad_(True)
    y = K.linear(x, w, b, True)
    gy = torch.randn(y.shape, generator=torch.Generator().manual_seed(9)).cuda()
    y.backward(gy)
    xr = x.detach().to(torch.bfloat16).float().requires_grad_(True)
    wr = w.detach().to(torch.bfloat16).float().requires_grad_(True)
    br = b.detach().clone().requires_grad_(True)
    yr = torch.relu(torch.nn.functional.linear(xr, wr, br))
    yr.backward(gy)
    assert (y - yr).abs().max().item() < 2e-2
    assert (x.grad - xr.grad).abs().max().item() < 5e-2
    assert (w.grad - wr.grad).abs().max().item() < 0.25      # bf16-rounded dY, 385 rows
    assert (b.grad - br.grad).abs().max().item() < 1e-3


@pytest.mark.parametrize('taps', [9, 1])
def test_conv_nhwc_matches_conv2d(taps):
    from applestar_b200 import _lib
    R, H, W, Cin, Cout = 3, 19, 20, 128, 128
    g = torch.Generator().manual_seed(10)
    x = torch.randn((R, Cin, H, W), generator=g).cuda()
    k = 3 if taps == 9 else 1
    w = (torch.randn((Cout, Cin, k, k), generator=g) * 0.05).cuda()
    b = torch.randn(Cout, generator=g).cuda()
    xp = torch.zeros((R, H + 2, W + 2, Cin), device='cuda', dtype=torch.bfloat16)
    xp[:, 1:-1, 1:-1] = x.permute(0, 2, 3, 1).to(torch.bfloat16)
    w2 = w.permute(0, 2, 3, 1).reshape(Cout, taps * Cin).contiguous().to(torch.bfloat16)     # [co][tap][ci]
    y = torch.full((R, H + 2, W + 2, Cout), float('nan'), device='cuda')
    _lib.call('as_conv_nhwc_bf16', xp.data_ptr(), w2.data_ptr(), y.data_ptr(), b.data_ptr(), R, H, W, Cin, Cout,
              taps, 0, 0, 1, _lib.stream_ptr())
    ref = torch.relu(torch.nn.functional.conv2d(x.to(torch.bfloat16).float(), w.to(torch.bfloat16).float(), b,
                                                padding=k // 2))
    assert torch.isfinite(y).all()
    assert y[:, 0].abs().max() == 0 and y[:, -1].abs().max() == 0 and y[:, :, 0].abs().max() == 0
    assert (y[:, 1:-1, 1:-1].permute(0, 3, 1, 2) - ref).abs().max().item() < _tol(taps * Cin) * 0.05 * 4
