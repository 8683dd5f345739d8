"""CPU: the C-ABI shared library loads and exports every symbol include/nds_b200.h declares;
without a GPU the engine refuses to start instead of falling back to the CPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'nds_b200.h')).read()
    return re.findall(r'NDS_API[^;(]*?\b(nds_\w+)\s*\(', text)


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    from spark_rapids_benchmarks_b200 import engine
    lib = engine.load_library()
    names = declared_symbols()
    assert len(names) >= 11 and set(names) == set(engine.SYMBOLS)
    for n in names:
        assert getattr(lib, n) is not None
    assert b'sm_100a' in lib.nds_version()


def test_engine_fails_loudly_without_gpu():
    from spark_rapids_benchmarks_b200 import engine
    lib = engine.load_library()
    h = ctypes.c_void_p()
    rc = lib.nds_engine_create(b'{"device": 9999}', ctypes.byref(h))
    assert rc != 0 and lib.nds_last_error()          # no such device / no device: never a silent CPU path
    with pytest.raises(engine.EngineError):
        engine.load_library(os.path.join(ROOT, 'no_such_dir', 'libndsb200.so'))
