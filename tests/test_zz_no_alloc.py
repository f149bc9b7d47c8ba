"""Collected last on purpose: checks the allocation-free column-sum entry point against the regular one.

`as_colsum_bf16_static` exists for hosts that are not allowed to allocate device memory while the learner or an
actor is running -- a process that handed its whole device budget to an arena at start-up, or a step loop that
cannot take the device-wide synchronisation a cudaMalloc / cudaFree would cause.  It runs the same kernels as
`as_colsum_bf16` but keeps the per-chunk partial sums in a scratch area built into the library (so C <= 256).
The test calls both through the C-ABI on two different inputs in a row: the second call reuses the shared area that
still holds the first call's partial sums, so a stale-scratch bug would show up there.  Same kernels, same summation
order => the results must be identical, which is the tolerance the other bit-exact tests here use."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(lib_call, lib, x, rows, C):
    from applestar_b200 import _lib
    out_ref = torch.full((C,), float('nan'), device='cuda')
    out_new = torch.full((C,), float('nan'), device='cuda')
    scratch = torch.empty(lib.as_colsum_scratch_floats(C), dtype=torch.float32, device='cuda')
    st = _lib.stream_ptr()
    lib_call('as_colsum_bf16', x.data_ptr(), out_ref.data_ptr(), scratch.data_ptr(), rows, C, st)
    lib_call('as_colsum_bf16_static', x.data_ptr(), out_new.data_ptr(), rows, C, st)
    torch.cuda.synchronize()
    return out_ref, out_new


def test_static_scratch_colsum_matches_regular_entry_point_twice():
    from applestar_b200 import _lib
    lib = _lib.lib()
    g = torch.Generator().manual_seed(41)
    for rows, C in ((70000, 128), (333, 256)):                  # second call: other shape, shared area still dirty
        x = torch.randn(rows, C, generator=g).cuda().to(torch.bfloat16)
        ref, new = _both(_lib.call, lib, x, rows, C)
        assert torch.isfinite(new).all()
        assert torch.equal(ref, new)
        exact = x.float().sum(dim=0)
        assert (new - exact).abs().max().item() <= 1e-3 * max(1.0, exact.abs().max().item())


def test_static_scratch_colsum_rejects_rows_wider_than_its_area():
    from applestar_b200 import _lib
    lib = _lib.lib()
    x = torch.zeros(8, 512, device='cuda', dtype=torch.bfloat16)
    out = torch.zeros(512, device='cuda')
    rc = lib.as_colsum_bf16_static(x.data_ptr(), out.data_ptr(), 8, 512, _lib.stream_ptr())
    assert rc == -2                                              # AS_ESIZE: C is beyond the 256-column built-in scratch
