"""GPU parity at a BASELINE.json scale: the six-query stream on an SF1 warehouse (2,880,404 store_sales rows) whose
sales fall on 182 days, so that a partition file holds ~15.8 k rows - the page shape of the SF10 / SF100 bench
warehouses (one 15.8 k-row page per column chunk, index streams of back-to-back 504-value bit-packed runs).
The tiny SF 0.01 warehouse of the other tests never reaches these routes:

  * the cursor-free rows path of the split page decode and the fused scan -> probe kernels over full-size pages,
  * dictionaries larger than the shared-memory stage (customer keys, measures),
  * the top-k prefilter (query7: ~9000 groups > the rank sort's 8192-row limit),
  * the group-by with more groups than its shared-memory cache.

Every statement is compared with the CPU oracle on both sources (pages resident in HBM / copied from host
buffers inside the query): bit exact for integers, strings, decimals and row order, doubles at rel 1e-5.
"""
import math
import os

import pytest

pytestmark = pytest.mark.gpu

SALES_DAYS = 182
# the same rows on 18 days: ~152 k rows and eight 20000-row pages per column chunk, dictionaries of ~0.5 MB whose index
# bit width grows from page to page - the shape of the SF100 bench warehouse's partition files
FEW_DAYS = 18
PARAMS = [{'MANUFACT': 606, 'MANAGER': 63, 'MONTH': 3, 'YEAR': 1998, 'GEN': 'M', 'MS': 'S', 'ES': 'College'},
          {'MANUFACT': 17, 'MANAGER': 8, 'MONTH': 5, 'YEAR': 1998, 'GEN': 'F', 'MS': 'W', 'ES': 'Primary'}]


@pytest.fixture(scope='module', params=[SALES_DAYS, FEW_DAYS])
def sf1(request, tmp_path_factory):
    from spark_rapids_benchmarks_b200 import datagen, engine
    days = request.param
    d = (os.environ.get('NDS_TEST_WAREHOUSE_SF1') or str(tmp_path_factory.mktemp('nds_sf1'))) + '_d%d' % days
    tables = list(datagen.SUBSET_TABLES)
    rows = datagen.ensure(d, 1, tables=tables, sales_days=days)
    assert rows['store_sales'] == 2880404
    eng = engine.Engine(device=0, resident=True, keep_host=True)
    for t in tables:
        eng.register_table(t, os.path.join(d, t))
    yield d, eng, days
    eng.close()


def rows_equal(got, exp):
    if len(got) != len(exp):
        return False
    for g, e in zip(got, exp):
        if g.keys() != e.keys():
            return False
        for k in g:
            if isinstance(e[k], float) and g[k] is not None:
                if not math.isclose(g[k], e[k], rel_tol=1e-5):
                    return False
            elif g[k] != e[k]:
                return False
    return True


@pytest.mark.parametrize('pi', range(len(PARAMS)))
def test_six_query_stream_at_sf1_matches_oracle(sf1, pi):
    import pyarrow as pa
    from oracle import nds_queries as oq
    from spark_rapids_benchmarks_b200 import planner, streams
    from tests.test_gpu_queries import _Sess
    wh, eng, days = sf1
    info = eng.table_info('store_sales')
    assert info['rows'] == 2880404 and info['files'] == days + 1
    pages = [c['pages'] for c in info['columns'] if c['name'] == 'ss_item_sk'][0]
    if days == SALES_DAYS:
        assert info['files'] <= pages <= info['files'] + 8    # one ~15.8 k-row page per date file, a few 20 k-row pages in the NULL partition
    else:
        assert pages >= 7 * days                              # eight pages per date file
    for query in streams.SUBSET:
        # the 18-day warehouse only has sales in January 1998
        sql = streams.render(query, overrides=PARAMS[pi] if days == SALES_DAYS else dict(PARAMS[pi], MONTH=1))
        exp = oq.run_query(query, sql, wh)
        assert exp.num_rows > 0, query
        for source in ('hbm', 'host'):
            s = _Sess(eng, source)
            got = s.sql(sql).toArrow()
            assert got.schema.names == exp.schema.names
            assert [str(t) for t in got.schema.types] == [str(t) for t in exp.schema.types], (got.schema, exp.schema)
            assert rows_equal(got.to_pylist(), exp.to_pylist()), (query, source, got.to_pylist()[:3], exp.to_pylist()[:3])
            ops = [o['op'] for o in eng.last_stats()['ops']]
            assert 'scan_probe' in ops and not any(o.startswith('decode:store_sales') for o in ops), (query, ops)   # the fused fact scan ran
        # the first join tested per dictionary entry and file (dict_match_kernel; on by itself from 4096 chunks = 8 M rows)
        got = pa.Table.from_batches([eng.execute(planner.plan_sql(sql), options={'dict_match_chunks': 1})])
        assert rows_equal(got.to_pylist(), exp.to_pylist()), query
        assert 'dict_match' in [o['op'] for o in eng.last_stats()['ops']], query


def test_query7_at_sf1_takes_the_large_routes(sf1):
    """The routes the small warehouse cannot reach are really taken here: thousands of groups (more than the group-by's
    shared-memory cache and than the rank sort's direct limit)."""
    from spark_rapids_benchmarks_b200 import planner, streams
    wh, eng, days = sf1
    sql = streams.render('query7', overrides=PARAMS[0])
    plan = planner.plan_sql(sql)
    node = plan['root']
    while node['op'] != 'aggregate':
        node = node['child']
    groups = eng.execute({'root': node})
    assert groups.num_rows > 8192, groups.num_rows
