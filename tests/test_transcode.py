"""nds_transcode (pyarrow): raw `|`-delimited text -> the reference's Parquet layout.  CPU tests on text dumped from the
tiny synthetic warehouse; the GPU test at the bottom runs a query on the transcoded copy."""
import glob
import os
import subprocess
import sys

import pyarrow as pa
import pyarrow.parquet as pq
import pytest

from spark_rapids_benchmarks_b200 import nds_schema, nds_transcode

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TABLES = ['store_sales', 'date_dim', 'item', 'promotion', 'customer_demographics', 'store', 'customer', 'customer_address']


def read_table(warehouse, table):
    """Whole table of the generated warehouse, partition column restored, in file order."""
    schema = nds_schema.get_schemas(True)[table]
    pcol = nds_schema.TABLE_PARTITIONING.get(table)
    parts = []
    for f in sorted(glob.glob(os.path.join(warehouse, table, '**', '*.parquet'), recursive=True)):
        t = pq.ParquetFile(f).read()
        if pcol:
            v = os.path.basename(os.path.dirname(f)).split('=', 1)[1]
            val = None if v == nds_transcode.HIVE_NULL else int(v)
            t = t.append_column(pcol, pa.array([val] * t.num_rows, type=pa.int32()))
        parts.append(t.select(schema.names).cast(schema))
    return pa.concat_tables(parts)


def dump_dat(t, path, nfiles=1, trailing=True):
    """dsdgen-style text: `|` between fields and after the last one, empty field for NULL, ISO-8859-1."""
    os.makedirs(path, exist_ok=True)
    cols = [c.to_pylist() for c in t.columns]
    rows = list(zip(*cols))
    per = (len(rows) + nfiles - 1) // nfiles if rows else 1
    for k in range(nfiles):
        with open(os.path.join(path, '%s_%d_%d.dat' % (os.path.basename(path), k + 1, nfiles)), 'w', encoding='latin1') as f:
            for r in rows[k * per:(k + 1) * per]:
                f.write('|'.join('' if v is None else str(v) for v in r) + ('|' if trailing else '') + '\n')


@pytest.fixture(scope='module')
def raw(warehouse, tmp_path_factory):
    d = str(tmp_path_factory.mktemp('raw'))
    originals = {}
    for name in TABLES:
        t = read_table(warehouse, name)
        originals[name] = t
        dump_dat(t, os.path.join(d, name), nfiles=3 if name == 'store_sales' else 1, trailing=name != 'store')
    return d, originals


def test_transcode_layout_values_and_report(raw, tmp_path):
    d, originals = raw
    out, report = str(tmp_path / 'wh'), str(tmp_path / 'load.txt')
    r = subprocess.run([sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_transcode', d, out, report, '--tables', ','.join(TABLES)],
                       cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    for name in TABLES:
        got = read_table(out, name)
        exp = originals[name]
        assert got.schema.equals(exp.schema), name
        if name in nds_schema.TABLE_PARTITIONING:
            # rows are regrouped by partition; inside a partition the input order is kept
            pcol = nds_schema.TABLE_PARTITIONING[name]
            key = [(v is None, v) for v in exp.column(pcol).to_pylist()]
            order = sorted(range(exp.num_rows), key=lambda i: key[i])
            exp = exp.take(pa.array(order))
            got_dirs = sorted(os.listdir(os.path.join(out, name)))
            assert nds_transcode.HIVE_NULL in ''.join(got_dirs) and all(g.startswith(pcol + '=') for g in got_dirs)
            f0 = glob.glob(os.path.join(out, name, '*', '*.parquet'))[0]
            assert pcol not in pq.ParquetFile(f0).schema_arrow.names          # hive style: the key lives in the path
        else:
            assert len(glob.glob(os.path.join(out, name, '*.parquet'))) == 1   # coalesce(1)
        assert got.equals(exp), name
    # physical layout the engine needs: no compression, v1 pages, decimals as integers
    md = pq.ParquetFile(glob.glob(os.path.join(out, 'store_sales', '*', '*.parquet'))[0]).metadata
    col = md.row_group(0).column(md.schema.names.index('ss_list_price'))
    assert col.compression == 'UNCOMPRESSED' and col.physical_type == 'INT32'
    text = open(report).read()
    assert text.startswith('Load Test Time: ') and 'RNGSEED used: ' in text and "Time to convert 'store_sales' was " in text
    assert float(text.split('\n')[0].split(':')[1].split(' ')[1]) >= 0        # how nds_bench.get_load_time reads it
    # save modes
    r2 = subprocess.run([sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_transcode', d, out, report, '--tables', 'store'],
                        cwd=ROOT, capture_output=True, text=True)
    assert r2.returncode != 0 and 'already exists' in r2.stderr
    r3 = subprocess.run([sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_transcode', d, out, report, '--tables', 'store',
                         '--output_mode', 'overwrite'], cwd=ROOT, capture_output=True, text=True)
    assert r3.returncode == 0, r3.stderr[-1000:]


def test_transcode_rejects_what_it_cannot_do(raw, tmp_path):
    d, _ = raw
    for extra in (['--output_format', 'iceberg'], ['--input_format', 'orc'], ['--hive']):
        r = subprocess.run([sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_transcode', d, str(tmp_path / 'x'), str(tmp_path / 'r.txt'),
                            '--tables', 'store'] + extra, cwd=ROOT, capture_output=True, text=True)
        assert r.returncode != 0
    r = subprocess.run([sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_transcode', d, str(tmp_path / 'y'), str(tmp_path / 'r.txt'),
                        '--tables', 'no_such_table'], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode != 0 and 'invalid table name' in r.stderr


@pytest.mark.gpu
def test_engine_reads_the_transcoded_warehouse(raw, warehouse, tmp_path):
    """Raw text -> nds_transcode -> engine: same query answer as on the generated warehouse (and the oracle's)."""
    from oracle import nds_queries as oq
    from spark_rapids_benchmarks_b200 import engine, planner, streams
    d, _ = raw
    out = str(tmp_path / 'wh')
    r = subprocess.run([sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_transcode', d, out, str(tmp_path / 'load.txt'),
                        '--tables', ','.join(TABLES)], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    eng = engine.Engine(device=0)
    for t in TABLES:
        eng.register_table(t, os.path.join(out, t))
    try:
        for q in ('query3', 'query7', 'query19'):
            sql = streams.render(q, overrides={'MANUFACT': 606, 'MANAGER': 63, 'MONTH': 1, 'YEAR': 1998})
            got = pa.Table.from_batches([eng.execute({'root': planner.plan_sql(sql)['root']})])
            exp = oq.run_query(q, sql, warehouse)
            assert got.num_rows == exp.num_rows
            for name in exp.schema.names:
                g, e = got.column(name).to_pylist(), exp.column(name).to_pylist()
                if pa.types.is_floating(exp.schema.field(name).type):
                    assert all((a is None and b is None) or abs(a - b) <= 1e-5 * max(1.0, abs(b)) for a, b in zip(g, e)), (q, name)
                else:
                    assert g == e, (q, name)
    finally:
        eng.close()
