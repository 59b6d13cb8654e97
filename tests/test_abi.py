"""The C-ABI library loads and exports every symbol include/epde_b200.h declares (no GPU)."""
import os
import re

import pytest

import epde_b200
from epde_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, 'include', 'epde_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(epde_[a-z0-9_]+)\s*\(', src)))


def test_library_builds_and_loads():
    lib = epde_b200.load_library()
    assert os.path.exists(epde_b200.library_path())
    assert lib.epde_version().decode().endswith('sm_100a')


def test_every_declared_symbol_is_exported_and_bound():
    lib = epde_b200.load_library()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f'{s} declared in include/epde_b200.h but not exported'
        assert s in _lib.SIGNATURES, f'{s} has no ctypes signature in epde_b200/_lib.py'
    for s in _lib.SIGNATURES:
        assert s in syms, f'{s} bound in _lib.py but not declared in the header'


def test_no_cpu_fallback(has_gpu):
    if has_gpu:
        pytest.skip('GPU present')
    with pytest.raises(epde_b200.EpdeB200Error, match='no CPU fallback'):
        epde_b200.Engine((16, 16))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'epde_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.h', '.cuh')):
                text = open(os.path.join(dirpath, f)).read()
                assert 'import oracle' not in text and 'from oracle' not in text, f
                assert '/root/reference' not in text, f


def test_encode_csr():
    enc = epde_b200.encode([[[(0, 1)], [(1, 2), (2, 1, 1, 3.5)]], [[(3, 1)]]])
    assert enc.n_eq == 2 and enc.n_eval == 3
    assert enc.term_off.tolist() == [0, 2, 3]
    assert enc.fac_off.tolist() == [0, 1, 3, 4]
    assert enc.fac_token.tolist() == [0, 1, 2, 3]
    assert enc.fac_kind.tolist() == [0, 0, 1, 0]
    assert enc.fac_param.tolist() == [0.0, 0.0, 3.5, 0.0]
    assert enc.eval_eq.tolist() == [0, 0, 1] and enc.eval_target.tolist() == [0, 1, 0]
