"""GPU parity for the fused fact scan (page decode -> join-chain probe -> late materialisation, fused_scan.cu) on hand-built
partition files: PLAIN pages, dictionary pages with a closing RLE run, a one-entry dictionary (bit width 0), an all-NULL
key column, a REQUIRED key column, several pages per column chunk, an empty partition, a NULL partition - and files the
fused route must decline (RLE-heavy key columns, key columns cut into pages at different rows), where the ordinary
decode + probe route has to give the same rows.  Expected rows: pyarrow joins / group-bys over the same files."""
import os

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc
import pyarrow.parquet as pq
import pytest

pytestmark = pytest.mark.gpu

KW = dict(compression='NONE', data_page_version='1.0', store_decimal_as_integer=True)
N_ITEM, N_CD = 500, 3000
DAY0 = 2450816


def _fact(rng, n, cd_mode='random', item_required=False):
    item = rng.integers(1, N_ITEM + 1, n).astype(np.int32)
    if cd_mode == 'random':
        cd = pa.array(rng.integers(1, N_CD + 1, n).astype(np.int32), mask=rng.random(n) < 0.05)
    elif cd_mode == 'null':
        cd = pa.array(np.zeros(n, np.int32), mask=np.ones(n, bool))
    elif cd_mode == 'constant':
        cd = pa.array(np.full(n, 77, np.int32), mask=rng.random(n) < 0.1)
    elif cd_mode == 'runs':                   # long repeats: RLE runs in the index stream -> not directly addressable
        cd = pa.array(np.repeat(rng.integers(1, N_CD + 1, n // 50 + 1), 50)[:n].astype(np.int32), mask=rng.random(n) < 0.02)
    fields = [pa.field('ss_item_sk', pa.int32(), nullable=not item_required), pa.field('ss_cdemo_sk', pa.int32()),
              pa.field('ss_quantity', pa.int64()), pa.field('ss_sales_price', pa.decimal128(7, 2)), pa.field('ss_ticket_number', pa.int64())]
    qty = pa.array(rng.integers(1, 101, n), mask=rng.random(n) < 0.04)
    import decimal
    price = pa.array([None if rng.random() < 0.04 else decimal.Decimal(int(v)).scaleb(-2) for v in rng.integers(0, 200000, n)], type=pa.decimal128(7, 2))
    return pa.table([pa.array(item), cd, qty, price, pa.array(np.arange(n, dtype=np.int64))], schema=pa.schema(fields))


def _write_warehouse(root, parts, seed=3):
    """parts: [(partition value | None, rows, cd_mode, writer kwargs, item_required)]"""
    rng = np.random.default_rng(seed)
    tables = []
    for i, (pval, n, mode, kw, req) in enumerate(parts):
        t = _fact(rng, n, mode, req)
        d = os.path.join(root, 'store_sales', 'ss_sold_date_sk=%s' % ('__HIVE_DEFAULT_PARTITION__' if pval is None else pval))
        os.makedirs(d, exist_ok=True)
        args = dict(KW)
        args.update(kw)
        pq.write_table(t, os.path.join(d, 'part-%05d.parquet' % i), **args)
        loose = pa.schema([pa.field(f.name, f.type) for f in t.schema])        # REQUIRED or not, the expected table is all nullable
        tables.append(t.cast(loose).append_column('ss_sold_date_sk', pa.array([pval] * n, type=pa.int32())))
    item = pa.table({'i_item_sk': pa.array(np.arange(1, N_ITEM + 1, dtype=np.int32)),
                     'i_manager_id': pa.array(rng.integers(1, 11, N_ITEM).astype(np.int64)),
                     'i_category': pa.array(['cat%d' % (k % 7) for k in range(N_ITEM)])})
    cd = pa.table({'cd_demo_sk': pa.array(np.arange(1, N_CD + 1, dtype=np.int32)),
                   'cd_gender': pa.array([['M', 'F'][k % 2] for k in range(N_CD)])})
    dd = pa.table({'d_date_sk': pa.array(np.arange(DAY0, DAY0 + 40, dtype=np.int32)), 'd_moy': pa.array((np.arange(40) // 20 + 1).astype(np.int64))})
    for name, t in (('item', item), ('customer_demographics', cd), ('date_dim', dd)):
        os.makedirs(os.path.join(root, name))
        pq.write_table(t, os.path.join(root, name, 'part-00000.parquet'), **KW)
    return pa.concat_tables(tables), item, cd, dd


SQL = ("select i_category, ss_item_sk, count(*) n, sum(ss_quantity) q, sum(ss_sales_price) p, min(ss_ticket_number) t "
       "from store_sales, item, customer_demographics, date_dim where ss_item_sk = i_item_sk and ss_cdemo_sk = cd_demo_sk "
       "and ss_sold_date_sk = d_date_sk and cd_gender = 'F' and i_manager_id <= 6 and d_moy = 1 "
       "group by i_category, ss_item_sk order by i_category, ss_item_sk limit 100000")


def _expected(fact, item, cd, dd):
    it = item.filter(pc.less_equal(item['i_manager_id'], 6))
    c2 = cd.filter(pc.equal(cd['cd_gender'], 'F')).select(['cd_demo_sk'])
    d2 = dd.filter(pc.equal(dd['d_moy'], 1)).select(['d_date_sk'])
    j = fact.join(it, keys='ss_item_sk', right_keys='i_item_sk', join_type='inner')
    j = j.join(c2, keys='ss_cdemo_sk', right_keys='cd_demo_sk', join_type='left semi').join(d2, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='left semi')
    g = j.group_by(['i_category', 'ss_item_sk']).aggregate([([], 'count_all'), ('ss_quantity', 'sum'), ('ss_sales_price', 'sum'), ('ss_ticket_number', 'min')])
    return sorted((r['i_category'], r['ss_item_sk'], r['count_all'], r['ss_quantity_sum'], r['ss_sales_price_sum'], r['ss_ticket_number_min'])
                  for r in g.to_pylist())


def _run(root, expect_fused):
    from spark_rapids_benchmarks_b200 import engine
    from tests.test_gpu_queries import _Sess
    eng = engine.Engine(device=0, keep_host=True)
    try:
        for t in ('store_sales', 'item', 'customer_demographics', 'date_dim'):
            eng.register_table(t, os.path.join(root, t))
        out = {}
        for source in ('hbm', 'host'):
            s = _Sess(eng, source)
            got = s.sql(SQL).toArrow()
            ops = [o['op'] for o in eng.last_stats()['ops']]
            assert ('scan_probe' in ops) == expect_fused, ops
            out[source] = [tuple(r.values()) for r in got.to_pylist()]
        # and the ordinary route on the same files
        from spark_rapids_benchmarks_b200 import planner
        plain = pa.Table.from_batches([eng.execute({'root': planner.plan_sql(SQL)['root']}, options={'fused_scan': False})])
        out['plain'] = [tuple(r.values()) for r in plain.to_pylist()]
        # the persistent probe kernel with the first join's keys as a filter in shared memory (an option: measured slower than the plain kernel so far)
        pers = pa.Table.from_batches([eng.execute({'root': planner.plan_sql(SQL)['root']}, options={'probe_bloom': True, 'probe_bloom_chunks': 1})])
        out['persistent'] = [tuple(r.values()) for r in pers.to_pylist()]
        # the route large scans take by themselves: first join tested per dictionary entry (dict_match_kernel + scan_probe_dm_kernel)
        for source in (None, 'host'):
            r = pa.Table.from_batches([eng.execute({'root': planner.plan_sql(SQL)['root']}, source=source, options={'dict_match_chunks': 1})])
            out['dict_match/%s' % source] = [tuple(x.values()) for x in r.to_pylist()]
        return out
    finally:
        eng.close()


def test_fused_scan_over_mixed_page_shapes(tmp_path):
    root = str(tmp_path)
    parts = [(DAY0 + 0, 20011, 'random', dict(use_dictionary=True), False),                       # dictionary pages, 20000-row page + 11-row page
             (DAY0 + 1, 9001, 'random', dict(use_dictionary=False), False),                       # PLAIN pages
             (DAY0 + 2, 7003, 'null', dict(use_dictionary=True), False),                          # key column all NULL
             (DAY0 + 3, 8009, 'constant', dict(use_dictionary=True), False),                      # one-entry dictionary: bit width 0
             (DAY0 + 4, 12345, 'random', dict(use_dictionary=True), True),                        # REQUIRED item key: no def levels
             (DAY0 + 5, 1, 'random', dict(use_dictionary=True), False),                           # a single row
             (DAY0 + 6, 50021, 'random', dict(use_dictionary=True, data_page_size=1 << 20), False),   # three pages per chunk
             (DAY0 + 25, 3000, 'random', dict(use_dictionary=True), False),                       # d_moy = 2: pruned away
             (None, 4000, 'random', dict(use_dictionary=True), False)]                            # NULL partition: never joins
    fact, item, cd, dd = _write_warehouse(root, parts)
    exp = _expected(fact, item, cd, dd)
    got = _run(root, expect_fused=True)
    assert len(exp) > 200
    for k, rows in got.items():
        assert rows == exp, k


def test_fused_scan_declines_and_the_ordinary_route_agrees(tmp_path):
    root = str(tmp_path)
    parts = [(DAY0 + 0, 20011, 'random', dict(use_dictionary=True), False),
             (DAY0 + 1, 30000, 'runs', dict(use_dictionary=True), False)]                         # RLE runs in the key's index stream
    fact, item, cd, dd = _write_warehouse(root, parts, seed=4)
    got = _run(root, expect_fused=False)
    exp = _expected(fact, item, cd, dd)
    for k, rows in got.items():
        assert rows == exp, k


def test_fused_scan_with_differently_paged_key_columns(tmp_path):
    """A tiny data page size cuts the key columns into pages by bytes: their bit widths differ, so do the page boundaries,
    and the fused scan - which walks the key columns page by page together - must step aside."""
    root = str(tmp_path)
    parts = [(DAY0 + 0, 60000, 'random', dict(use_dictionary=True, data_page_size=4096, write_batch_size=64), False)]
    fact, item, cd, dd = _write_warehouse(root, parts, seed=5)
    from spark_rapids_benchmarks_b200 import engine
    eng = engine.Engine(device=0)
    eng.register_table('store_sales', os.path.join(root, 'store_sales'))
    info = {c['name']: c['pages'] for c in eng.table_info('store_sales')['columns']}
    eng.close()
    got = _run(root, expect_fused=info['ss_item_sk'] == info['ss_cdemo_sk'] and info['ss_item_sk'] <= 3)
    exp = _expected(fact, item, cd, dd)
    for k, rows in got.items():
        assert rows == exp, k
