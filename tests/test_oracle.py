"""CPU: the oracles against independent implementations.
 * oracle/parquet_pages.py (Thrift + RLE/bit-packed hybrid restatement) vs pyarrow.parquet
 * oracle/nds_queries.py: Acero path vs the plain-Python path
The reference ships no golden vectors for the query results: parity with Spark is
unpinned; these tests pin the oracle against pyarrow and against itself."""
import os

import pyarrow.parquet as pq
import pytest

from oracle import parquet_pages as pp
from oracle import nds_queries as oq
from spark_rapids_benchmarks_b200 import streams


def _files(warehouse, table):
    for root, _, files in os.walk(os.path.join(warehouse, table)):
        for f in sorted(files):
            if f.endswith('.parquet'):
                yield os.path.join(root, f)


@pytest.mark.parametrize('table', ['item', 'promotion', 'store', 'customer_address', 'store_sales'])
def test_page_oracle_matches_pyarrow(warehouse, table):
    n = 0
    for path in _files(warehouse, table):
        buf = open(path, 'rb').read()
        md = pp.file_metadata(buf)
        t = pq.ParquetFile(path).read()
        schema = md[2]
        for rg_i, rg in enumerate(md[4]):
            assert len(md[4]) == 1
            for ci, cc in enumerate(rg[1]):
                cm = cc[3]
                max_def = 1 if schema[ci + 1].get(3, 0) == 1 else 0
                got = pp.decode_column_chunk(buf, cm, max_def)
                col = t.column(ci)
                if cm[1] == 6:
                    exp = [None if v is None else v.encode() for v in col.to_pylist()]
                elif str(col.type).startswith('decimal'):
                    exp = [None if v is None else int(v.scaleb(col.type.scale)) for v in col.to_pylist()]
                elif str(col.type) == 'date32[day]':
                    exp = col.cast('int32').to_pylist()
                else:
                    exp = col.to_pylist()
                assert got == exp, (path, cm[3])
        n += 1
        if n >= 3:
            break
    assert n > 0


@pytest.mark.parametrize('query', streams.SUBSET)
def test_query_oracle_two_paths_agree(warehouse, query):
    for ov in ({'MANUFACT': 606, 'MANAGER': 63, 'MONTH': 1, 'YEAR': 1998}, {'MONTH': 2, 'YEAR': 1998, 'MANAGER': 1, 'MANUFACT': 1}):
        sql = streams.render(query, overrides=ov)
        a = oq.run_query(query, sql, warehouse)
        b = oq.run_query_slow(query, sql, warehouse)
        assert a.to_pylist() == b


def test_stream_generator_roundtrip(tmp_path):
    from spark_rapids_benchmarks_b200 import nds_power
    p = str(tmp_path / 'query_0.sql')
    names = streams.generate_stream(p)
    d = nds_power.gen_sql_from_stream(p)
    assert list(d.keys()) == names == streams.SUBSET
    for k, v in d.items():
        assert v.startswith('-- start query') and (k + '.tpl') in v
