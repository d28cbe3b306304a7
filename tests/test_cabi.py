"""CPU-side checks of the drop-in boundary: libnr_sm100.so loads and exports every symbol
include/nr_sm100.h declares; argument validation errors are reported, never swallowed."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, 'include', 'nr_sm100.h')
SO = os.path.join(ROOT, 'newsrecommendation_b200', 'libnr_sm100.so')


def declared_symbols():
    src = open(HDR).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(nr_[a-z0-9_]+)\s*\(', src)))


@pytest.fixture(scope='module')
def lib():
    if not os.path.exists(SO):
        import __graft_entry__ as g
        g.build()
    return ctypes.CDLL(SO)


def test_header_symbols_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f'{s} declared in include/nr_sm100.h but not exported'


def test_python_binding_covers_header(lib):
    from newsrecommendation_b200 import _lib
    assert sorted(_lib.EXPORTS) == declared_symbols()
    assert _lib.lib.nr_abi_version() == 1


def test_errors_are_loud(lib):
    from newsrecommendation_b200 import _lib
    import torch
    # NULL pointers -> error code + message, no crash
    rc = _lib.lib.nr_gather_rows(None, 1, None, 1, 4, None, None)
    assert rc != 0 and 'NULL' in _lib.last_error()
    with pytest.raises(RuntimeError, match='CUDA tensor required'):
        _lib.ptr(torch.zeros(4))          # there is no CPU path
