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


def test_query98_oracle_against_plain_python(warehouse):
    """The Acero restatement of query98 against dict / Decimal loops over the same files (an independent second path, like
    run_query_slow for the six bench statements): same groups, same revenue, same share of the class's revenue, same order."""
    import datetime
    import decimal
    sql = streams.render('query98')
    a = oq.run_query('query98', sql, warehouse).to_pylist()
    d0 = datetime.date(1998, 1, 5)
    days = {r['d_date_sk'] for r in oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_date']).to_pylist()
            if r['d_date'] is not None and d0 <= r['d_date'] <= d0 + datetime.timedelta(days=30)}
    items = {r['i_item_sk']: r for r in oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_item_id', 'i_item_desc', 'i_category', 'i_class',
                                                                       'i_current_price']).to_pylist()
             if r['i_category'] in ('Sports', 'Books', 'Home')}
    groups = {}
    for r in oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_item_sk', 'ss_ext_sales_price']).to_pylist():
        if r['ss_sold_date_sk'] in days and r['ss_item_sk'] in items:
            it = items[r['ss_item_sk']]
            k = (it['i_item_id'], it['i_item_desc'], it['i_category'], it['i_class'], it['i_current_price'])
            s = groups.setdefault(k, [None])
            if r['ss_ext_sales_price'] is not None:
                s[0] = r['ss_ext_sales_price'] if s[0] is None else s[0] + r['ss_ext_sales_price']
    total = {}
    for k, (s,) in groups.items():
        if s is not None:
            total[k[3]] = total.get(k[3], decimal.Decimal(0)) + s
    b = {k: (s, (float(s * 100) / float(total[k[3]])) if s is not None and total.get(k[3]) else None) for k, (s,) in groups.items()}
    assert len(a) == len(b) > 20
    for r in a:
        k = (r['i_item_id'], r['i_item_desc'], r['i_category'], r['i_class'], r['i_current_price'])
        assert b[k][0] == r['itemrevenue'] and (b[k][1] == r['revenueratio'] or abs(b[k][1] - r['revenueratio']) < 1e-9 * abs(b[k][1]))
    keys = [(r['i_category'], r['i_class'], r['i_item_id']) for r in a]
    none_first = lambda t: tuple((v is not None, v or '') for v in t)
    assert keys == sorted(keys, key=none_first)
