"""CPU-only: the C-ABI library loads and exports every symbol include/deltaconv_b200.h declares,
and the ctypes table matches the header's parameter counts.  No compute call is made."""
import ctypes
import os
import re

import pytest

from deltaconv_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'deltaconv_b200.h')


def _declarations():
    text = open(HEADER).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    decls = {}
    for m in re.finditer(r'\b(dc_\w+)\s*\(([^;]*?)\)\s*;', text, flags=re.S):
        args = m.group(2).strip()
        n = 0 if args in ('', 'void') else len(args.split(','))
        decls[m.group(1)] = n
    return decls


def test_header_declares_something():
    assert len(_declarations()) >= 15


@pytest.mark.skipif(not os.path.exists(_lib.LIB_PATH), reason='library not built')
def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declarations():
        assert hasattr(lib, name), f'{name} declared in the header but not exported'


def test_ctypes_table_matches_header():
    decls = _declarations()
    table = dict(_lib._SIGNATURES)
    assert set(table) | {'dc_last_error'} == set(decls), set(decls) ^ (set(table) | {'dc_last_error'})
    for name, argtypes in table.items():
        assert len(argtypes) == decls[name], name


@pytest.mark.skipif(not os.path.exists(_lib.LIB_PATH), reason='library not built')
def test_loader_and_version():
    lib = _lib.load()
    assert lib.dc_abi_version() == 2
    assert lib.dc_last_error() is not None


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, '_lib', None)
    monkeypatch.setattr(_lib, 'LIB_PATH', '/nonexistent/libdeltaconv_b200.so')
    with pytest.raises(_lib.DeltaConvLibraryError):
        _lib.load()
