"""GPU: nds_join_probe_i64_noalloc, the hash-join lookup for callers that may not allocate GPU memory.

Why it exists: when this library runs inside a host that already owns the device's memory - the
spark-rapids plugin gives nearly all HBM to its RMM pool when an executor starts - a cudaMalloc made
by a library fails or slips past the pool's spill accounting, so such a caller needs a join whose
hash table is not allocated per call.  This entry point keeps the table in a fixed array inside the
library (at most NDS_NOALLOC_MAX_BUILD_KEYS = 65536 build keys).

The test drives it through the C ABI with caller-owned device buffers and checks it, bit for bit, first
against a plain dictionary lookup on the host (an oracle that shares no code with the kernels) and then
against the regular entry point (nds_engine_execute running a join plan), twice in a row: the table is
shared between calls and still holds the first call's keys when the second one starts.
This file sorts last on purpose (pytest collects files alphabetically).
"""
import ctypes

import numpy as np
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu

MAX_BUILD = 65536


def _reference_via_engine(eng, build, probe):
    """Same join through nds_engine_execute: inner join of (fk, rid) with (k, brow), back to dense form."""
    eng.register_arrow('na_build', pa.table({'k': pa.array(build), 'brow': pa.array(np.arange(len(build), dtype=np.int64))}))
    eng.register_arrow('na_probe', pa.table({'fk': pa.array(probe), 'rid': pa.array(np.arange(len(probe), dtype=np.int64))}))
    plan = {'op': 'join', 'probe': {'op': 'scan', 'table': 'na_probe', 'columns': ['fk', 'rid']},
            'build': {'op': 'scan', 'table': 'na_build', 'columns': ['k', 'brow']}, 'probe_key': 'fk', 'build_key': 'k'}
    got = eng.execute(plan)
    out = np.full(len(probe), -1, dtype=np.int32)
    rid = got.column(got.schema.get_field_index('rid')).to_numpy()
    brow = got.column(got.schema.get_field_index('brow')).to_numpy()
    out[rid] = brow.astype(np.int32)
    eng.drop_table('na_build')
    eng.drop_table('na_probe')
    return out


def _reference_on_host(build, probe):
    """The definition itself: position of each probe key among the (unique) build keys, -1 if absent."""
    where = {int(k): i for i, k in enumerate(build.tolist())}
    return np.array([where.get(int(k), -1) for k in probe.tolist()], dtype=np.int32)


def _call_noalloc(lib, torch, build, probe):
    d_build = torch.from_numpy(build).cuda()
    d_probe = torch.from_numpy(probe).cuda()
    d_out = torch.full((len(probe),), -7, dtype=torch.int32, device='cuda')
    torch.cuda.synchronize()
    rc = lib.nds_join_probe_i64_noalloc(ctypes.c_void_p(d_build.data_ptr()), len(build), ctypes.c_void_p(d_probe.data_ptr()),
                                        len(probe), ctypes.c_void_p(d_out.data_ptr()), None)
    assert rc == 0, lib.nds_last_error().decode()
    return d_out.cpu().numpy()


def test_noalloc_join_matches_engine_twice_in_a_row():
    import torch
    from spark_rapids_benchmarks_b200 import engine
    lib = engine.load_library()
    eng = engine.Engine(device=0)
    torch.cuda.set_device(0)
    rng = np.random.default_rng(11)
    # call 1: a full-size build side; call 2: a small, different key set (stale keys of call 1 must be gone)
    cases = [(rng.permutation(np.arange(10, 10 + 4 * MAX_BUILD, 4, dtype=np.int64)), rng.integers(0, 5 * MAX_BUILD, 300001).astype(np.int64)),
             (np.array([-5, 7, 2 ** 40, 14, 18], dtype=np.int64), rng.integers(-10, 60, 4099).astype(np.int64))]
    for build, probe in cases:
        got = _call_noalloc(lib, torch, build, probe)
        oracle = _reference_on_host(build, probe)
        assert got.dtype == oracle.dtype and np.array_equal(got, oracle)   # integer indices: bit-exact
        exp = _reference_via_engine(eng, build, probe)
        assert got.dtype == exp.dtype and np.array_equal(got, exp)      # and identical to the allocating entry point
        assert (got >= 0).sum() > 0 and (got < 0).sum() > 0
    # empty sides
    e = np.empty(0, dtype=np.int64)
    assert np.array_equal(_call_noalloc(lib, torch, e, np.array([1, 2, 3], dtype=np.int64)), np.array([-1, -1, -1], dtype=np.int32))
    assert _call_noalloc(lib, torch, np.array([1], dtype=np.int64), e).size == 0
    eng.close()


def test_noalloc_join_refuses_what_it_cannot_hold():
    import torch
    from spark_rapids_benchmarks_b200 import engine
    lib = engine.load_library()
    torch.cuda.set_device(0)
    too_many = torch.arange(MAX_BUILD + 1, dtype=torch.int64, device='cuda')
    probe = torch.zeros(8, dtype=torch.int64, device='cuda')
    out = torch.zeros(8, dtype=torch.int32, device='cuda')
    rc = lib.nds_join_probe_i64_noalloc(ctypes.c_void_p(too_many.data_ptr()), MAX_BUILD + 1, ctypes.c_void_p(probe.data_ptr()), 8,
                                        ctypes.c_void_p(out.data_ptr()), None)
    assert rc != 0 and b'65536' in lib.nds_last_error()
    dup = torch.tensor([3, 4, 3], dtype=torch.int64, device='cuda')
    rc = lib.nds_join_probe_i64_noalloc(ctypes.c_void_p(dup.data_ptr()), 3, ctypes.c_void_p(probe.data_ptr()), 8,
                                        ctypes.c_void_p(out.data_ptr()), None)
    assert rc != 0 and b'duplicate' in lib.nds_last_error()
    # and it still works afterwards
    ok = torch.tensor([0, 9], dtype=torch.int64, device='cuda')
    rc = lib.nds_join_probe_i64_noalloc(ctypes.c_void_p(ok.data_ptr()), 2, ctypes.c_void_p(probe.data_ptr()), 8,
                                        ctypes.c_void_p(out.data_ptr()), None)
    assert rc == 0 and out.cpu().tolist() == [0] * 8
