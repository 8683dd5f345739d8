"""GPU parity for the page decoder on hand-built Parquet files: every bit width, RLE-heavy and
bit-packed-heavy index streams, all-null / no-null / single-row pages, maximum page size, PLAIN
fallbacks, variable-length strings, several row groups and files - bit exact vs pyarrow."""
import os

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

pytestmark = pytest.mark.gpu

KW = dict(compression='NONE', data_page_version='1.0', store_decimal_as_integer=True)


def _roundtrip(eng, tmp_path, name, table, files=1, **kw):
    d = tmp_path / name
    os.makedirs(d)
    n = table.num_rows
    per = (n + files - 1) // files if n else 0
    for f in range(files):
        part = table.slice(f * per, per) if files > 1 else table
        args = dict(KW)
        args.update(kw)
        pq.write_table(part, d / ('part-%05d.parquet' % f), **args)
    eng.register_table(name, str(d))
    got = eng.execute({'op': 'scan', 'table': name, 'columns': table.schema.names})
    got_general = eng.execute({'op': 'scan', 'table': name, 'columns': table.schema.names}, options={'split_decode': False})
    assert got_general.equals(got), name + ': the general kernel and the split def/rows kernels disagree'
    exp = pa.concat_tables([pq.ParquetFile(str(d / ('part-%05d.parquet' % f))).read() for f in range(files)]).combine_chunks()
    got_t = pa.Table.from_batches([got])
    assert got_t.num_rows == exp.num_rows
    for c in table.schema.names:
        assert got_t[c].combine_chunks().equals(exp[c].combine_chunks()), (name, c)
    eng.drop_table(name)
    return got_t


def test_every_bit_width(gpu_engine, tmp_path):
    rng = np.random.default_rng(5)
    n = 30011
    cols = {}
    for bw in list(range(1, 21)):
        card = min(1 << bw, 200000)
        lo = (1 << (bw - 1)) if bw > 1 else 1
        k = int(rng.integers(lo, card)) + 1 if card > lo else 2
        vals = rng.integers(0, k, n)
        cols['w%02d' % bw] = pa.array(vals.astype(np.int32 if bw % 2 else np.int64), mask=rng.random(n) < 0.07)
    _roundtrip(gpu_engine, tmp_path, 'bitwidths', pa.table(cols), use_dictionary=True, dictionary_pagesize_limit=8 << 20)


def test_run_structures(gpu_engine, tmp_path):
    rng = np.random.default_rng(6)
    n = 100003
    sorted_low = np.sort(rng.integers(0, 40, n))                 # long RLE runs
    short_runs = np.repeat(rng.integers(0, 500, n // 9 + 1), 9)[:n]   # thousands of short RLE runs per page
    alternating = np.arange(n) % 2                               # pure bit-packed, bw 1
    mixed = np.where(rng.random(n) < 0.5, 7, rng.integers(0, 1000, n))
    t = pa.table({
        'sorted_low': pa.array(sorted_low, mask=rng.random(n) < 0.02),
        'short_runs': pa.array(short_runs.astype(np.int32)),
        'short_runs_nullable': pa.array(short_runs, mask=rng.random(n) < 0.3),
        'alternating': pa.array(alternating.astype(np.int32), mask=rng.random(n) < 0.5),
        'mixed': pa.array(mixed),
        'constant': pa.array(np.full(n, 42, dtype=np.int64)),
        'all_null': pa.array([None] * n, type=pa.int32()),
        'no_null_but_nullable': pa.array(rng.integers(-2 ** 40, 2 ** 40, n)),
        'null_blocks': pa.array(rng.integers(0, 9, n), mask=(np.arange(n) // 700) % 3 == 0),
    })
    _roundtrip(gpu_engine, tmp_path, 'runs', t, use_dictionary=True)


def test_page_sizes_and_row_groups(gpu_engine, tmp_path):
    rng = np.random.default_rng(7)
    for n in (1, 2, 31, 32, 33, 255, 256, 257, 4095, 20000, 20001, 32768, 70001):
        t = pa.table({'a': pa.array(rng.integers(0, 50, n).astype(np.int32), mask=rng.random(n) < 0.1),
                      'b': pa.array(rng.integers(0, 10 ** 9, n)),
                      's': pa.array(['v%d' % (i % 17) for i in range(n)])})
        _roundtrip(gpu_engine, tmp_path, 'n%d' % n, t, use_dictionary=True, max_rows_per_page=32768, write_batch_size=32768,
                   data_page_size=8 << 20)
    n = 250007
    t = pa.table({'a': pa.array(rng.integers(0, 5000, n).astype(np.int32), mask=rng.random(n) < 0.04),
                  'd': pa.array(rng.random(n)), 's': pa.array(['k%05d' % (i % 3000) for i in range(n)])})
    _roundtrip(gpu_engine, tmp_path, 'rowgroups', t, files=3, use_dictionary=True, row_group_size=40000)


def test_plain_and_strings(gpu_engine, tmp_path):
    rng = np.random.default_rng(8)
    n = 60013
    words = ['', 'a', 'bb', 'longer string with spaces', 'ünïcødé ✓', 'x' * 300]
    var = [None if rng.random() < 0.05 else words[int(rng.integers(0, len(words)))] + str(i) for i in range(n)]
    fixed = ['ID%014d' % i for i in range(n)]
    t = pa.table({'plain_i32': pa.array(rng.integers(-2 ** 31, 2 ** 31, n).astype(np.int32), mask=rng.random(n) < 0.1),
                  'plain_i64': pa.array(rng.integers(-2 ** 62, 2 ** 62, n)),
                  'plain_f64': pa.array(rng.standard_normal(n), mask=rng.random(n) < 0.1),
                  'var_str': pa.array(var, type=pa.string()),
                  'fixed_str': pa.array(fixed)})
    _roundtrip(gpu_engine, tmp_path, 'plain', t, use_dictionary=False)
    # dictionary overflow: starts dictionary-encoded, falls back to PLAIN pages inside the same column chunk
    _roundtrip(gpu_engine, tmp_path, 'fallback', t.select(['fixed_str', 'var_str', 'plain_i64']), use_dictionary=True,
               dictionary_pagesize_limit=64 << 10)


def test_pages_beyond_the_limit_are_refused(gpu_engine, tmp_path):
    from spark_rapids_benchmarks_b200.engine import EngineError
    n = 40000
    d = tmp_path / 'big'
    os.makedirs(d)
    pq.write_table(pa.table({'a': pa.array(np.arange(n) % 7)}), d / 'p.parquet', max_rows_per_page=40000, write_batch_size=40000,
                   data_page_size=64 << 20, **KW)
    with pytest.raises(EngineError) as ei:
        gpu_engine.register_table('big', str(d))
    assert '32768' in str(ei.value)
    pq.write_table(pa.table({'a': pa.array(np.arange(100))}), d / 'p.parquet', compression='SNAPPY')
    with pytest.raises(EngineError) as ei:
        gpu_engine.register_table('big', str(d))
    assert 'compress' in str(ei.value)
