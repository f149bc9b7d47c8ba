"""CPU suite: the C-ABI library builds for sm_100a, loads, and exports every symbol include/*.h declares."""
import os
import re

from applestar_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    header = open(os.path.join(ROOT, 'include', 'applestar_b200.h')).read()
    declared = set(re.findall(r'\b(as_[a-z0-9_]+)\s*\(', header))
    assert declared, 'no declarations found'
    lib = _lib.lib()
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(_lib.EXPORTS)
    import subprocess
    nm = subprocess.run(['nm', '-D', '--defined-only', _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {ln.split()[-1] for ln in nm.splitlines() if ln.split() and ln.split()[-1].startswith('as_')}
    assert exported == declared, sorted(exported ^ declared)        # nothing undeclared is exported either
    assert lib.as_abi_version() >= 1
    assert lib.as_gemm_splits(64 * 10, 4) == 4
    assert lib.as_gemm_splits(64 * 10, 7) == 5     # ceil(10/7)=2 k-blocks per split -> 5 non-empty splits


def test_product_refuses_cpu_tensors():
    import pytest
    import torch
    from applestar_b200 import ops
    with pytest.raises(RuntimeError):
        ops.linear(torch.zeros(2, 8), torch.zeros(4, 8), None)


def test_missing_extension_fails_loudly(monkeypatch):
    """No CPU / library fallback: without the built .so every product call raises (it never routes around the kernels)."""
    import pytest
    from applestar_b200 import _lib
    monkeypatch.setattr(_lib, '_lib', None)
    monkeypatch.setattr(_lib, 'LIB_PATH', '/nonexistent/libapplestar_b200.so')
    with pytest.raises(RuntimeError, match='no fallback'):
        _lib.lib()
    with pytest.raises(RuntimeError):
        _lib.call('as_abi_version')
