"""CPU checks of the C-ABI boundary: the library loads, exports every symbol the header declares and the Python
binding expects.  No compute call is made (no GPU here)."""
import ctypes
import os
import re

import pytest

from pycsou_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, 'include', 'pcsb200.h')).read()
    txt = re.sub(r'/\*.*?\*/', '', txt, flags=re.S)
    names = set()
    for m in re.finditer(r'^PCS_DECL_REAL\(\s*(pcs_\w+)\s*,', txt, flags=re.M):
        names.update({m.group(1) + '_f32', m.group(1) + '_f64'})
    for m in re.finditer(r'^\s*(?:int|int64_t|const char\*)\s+(pcs_\w+)\s*\(', txt, flags=re.M):
        names.add(m.group(1))
    return names


def test_library_is_built_and_loads():
    assert os.path.exists(_lib.LIB_PATH), 'run __graft_entry__.build() first'
    lib = _lib.lib()
    assert lib.pcs_version() >= 100


def test_every_declared_symbol_is_exported():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    hs = header_symbols()
    assert len(hs) > 60
    for name in sorted(hs):
        assert hasattr(lib, name), f'{name} declared in include/pcsb200.h but not exported'


def test_binding_and_header_agree():
    hs, bs = header_symbols(), set(_lib.declared_symbols())
    assert bs <= hs, f'bound but undeclared: {sorted(bs - hs)}'
    assert hs <= bs, f'declared but unbound: {sorted(hs - bs)}'


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, '_lib', None)
    monkeypatch.setattr(_lib, 'LIB_PATH', '/nonexistent/libpcsb200.so')
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        _lib.lib()
