"""GPU parity, end to end: SQL text -> planner -> C ABI -> CUDA kernels, against the CPU oracle on the
same seeded warehouse.  Bit exact for integers, strings, decimals and row order (ties broken by the
remaining columns on both sides); AVG(bigint) doubles compared with nds_validate's epsilon 1e-5."""
import contextlib
import glob
import io
import math
import os

import pytest

from oracle import nds_queries as oq
from spark_rapids_benchmarks_b200 import nds_power, nds_validate, streams
from spark_rapids_benchmarks_b200.session import Session

pytestmark = pytest.mark.gpu

PARAMS = [{'MANUFACT': 606, 'MANAGER': 63, 'MONTH': 1, 'YEAR': 1998},
          {'MANUFACT': 1, 'MANAGER': 1, 'MONTH': 2, 'YEAR': 1998},
          {'MANUFACT': 9999, 'MANAGER': 9999, 'MONTH': 1, 'YEAR': 1998},      # empty dimension side
          {'MONTH': 7, 'YEAR': 2001, 'MANUFACT': 606, 'MANAGER': 63}]         # no partition survives pruning


class _Sess(Session):
    """Session over the shared test engine (tables already registered)."""

    def __init__(self, eng, source='hbm'):
        self.conf, self.app_name, self.version, self.application_id = {}, 't', 'test', 'b200-test'
        self.rank, self.world, self.comm, self.source, self.engine, self._stats = 0, 1, None, source, eng, None

    def stop(self):
        pass


def rows_equal(got, exp):
    if len(got) != len(exp):
        return False
    for g, e in zip(got, exp):
        if g.keys() != e.keys():
            return False
        for k in g:
            if isinstance(e[k], float) and e[k] is not None and g[k] is not None:
                if not math.isclose(g[k], e[k], rel_tol=1e-5):
                    return False
            elif g[k] != e[k]:
                return False
    return True


def test_query98_matches_the_oracle(gpu_engine, warehouse):
    """A seventh NDS statement beyond the bench stream: query98 as the Spark-dialect template has it (date constant + INTERVAL,
    five group keys incl. the 200-character item description, sum(sum(x)) OVER (PARTITION BY i_class), ORDER BY without LIMIT)."""
    sql = streams.render('query98')
    exp = oq.run_query('query98', sql, warehouse)
    assert exp.num_rows > 20
    for source in ('hbm', 'host'):
        got = _Sess(gpu_engine, source).sql(sql).toArrow()
        assert got.schema.names == exp.schema.names
        assert [str(t) for t in got.schema.types] == [str(t) for t in exp.schema.types], (got.schema, exp.schema)
        assert rows_equal(got.to_pylist(), exp.to_pylist()), (source, got.to_pylist()[:2], exp.to_pylist()[:2])
        ops = [o['op'] for o in gpu_engine.last_stats()['ops']]
        assert 'scan_probe' in ops and 'window' in ops, ops


@pytest.mark.parametrize('query', streams.SUBSET)
@pytest.mark.parametrize('pi', range(len(PARAMS)))
def test_query_matches_oracle(gpu_engine, warehouse, query, pi):
    sql = streams.render(query, overrides=PARAMS[pi])
    exp = oq.run_query(query, sql, warehouse)
    for source in ('hbm', 'host'):
        got = _Sess(gpu_engine, source).sql(sql).toArrow()
        assert got.schema.names == exp.schema.names
        assert [str(t) for t in got.schema.types] == [str(t) for t in exp.schema.types], (got.schema, exp.schema)
        assert rows_equal(got.to_pylist(), exp.to_pylist()), (query, source, got.to_pylist()[:3], exp.to_pylist()[:3])


def test_power_run_then_validate_against_oracle(gpu_engine, warehouse, tmp_path):
    """nds_power-compatible run on the GPU, oracle outputs written beside it, nds_validate-compatible
    comparison of the two folders - the reference's own correctness procedure."""
    import pyarrow.parquet as pq
    stream = str(tmp_path / 'query_0.sql')
    streams.generate_stream(stream, overrides=PARAMS[0])
    qd = nds_power.gen_sql_from_stream(stream)
    out_gpu, out_cpu, js = str(tmp_path / 'gpu'), str(tmp_path / 'cpu'), str(tmp_path / 'json')
    sess = _Sess(gpu_engine)
    sess.register_table = lambda *a: None
    with contextlib.redirect_stdout(io.StringIO()):
        rows = nds_power.run_query_stream(warehouse, None, qd, str(tmp_path / 'time.csv'), None, None, output_path=out_gpu,
                                          json_summary_folder=js, session=sess)
    assert [r[1] for r in rows if r[1].startswith('query')] == streams.SUBSET
    for q, sql in qd.items():
        t = oq.run_query(q, sql, warehouse)
        os.makedirs(os.path.join(out_cpu, q))
        df = nds_power.ensure_valid_column_names(__import__('spark_rapids_benchmarks_b200.session', fromlist=['DataFrame']).DataFrame(t))
        pq.write_table(df.toArrow(), os.path.join(out_cpu, q, 'part-0.parquet'))
    with contextlib.redirect_stdout(io.StringIO()):
        assert nds_validate.main([out_gpu, out_cpu, stream, '--json_summary_folder', js]) == []
        assert nds_validate.main([out_gpu, out_cpu, stream, '--ignore_ordering']) == []
    import json
    for q in qd:
        f = glob.glob(os.path.join(js, '*%s-*.json' % q))
        assert len(f) == 1 and json.load(open(f[0]))['queryValidationStatus'] == ['Pass']


def test_unsupported_statement_is_reported_failed(gpu_engine, tmp_path):
    sess = _Sess(gpu_engine)
    sess.register_table = lambda *a: None
    stream = tmp_path / 'q.sql'
    stream.write_text('-- start query 1 in stream 0 using template query98.tpl\n'
                      'select i_item_id, sum(i_item_sk) over (partition by i_class order by i_item_id) from item\n;\n'
                      '-- end query 1 in stream 0 using template query98.tpl\n')
    qd = nds_power.gen_sql_from_stream(str(stream))
    with contextlib.redirect_stdout(io.StringIO()):
        with pytest.raises(SystemExit):
            nds_power.run_query_stream(str(tmp_path), None, qd, str(tmp_path / 't.csv'), None, None, session=sess)


@pytest.mark.parametrize('query', streams.SUBSET)
def test_two_rank_emulation_on_one_gpu(warehouse, query):
    """Distributed plan without collectives: two engines (rank 0/2 and 1/2) on the same GPU produce
    partial states for their file shards; merging them gives the single-GPU answer and the oracle's."""
    import os
    from spark_rapids_benchmarks_b200 import dist as ndist, engine, nds_schema, planner
    import pyarrow as pa
    sql = streams.render(query, overrides=PARAMS[0])
    plan = planner.plan_sql(sql, distributed=True)
    partials, engines = [], []
    for r in range(2):
        eng = engine.Engine(device=0, rank=r, world=2)
        for t in nds_schema.get_schemas(True):
            eng.register_table(t, os.path.join(warehouse, t))
        engines.append(eng)
        partials.append(pa.Table.from_batches([eng.execute({'root': plan['partial']})]))
    info = [e.table_info('store_sales') for e in engines]
    assert info[0]['files'] + info[1]['files'] == 41 and abs(info[0]['files'] - info[1]['files']) <= 1
    assert info[0]['rows'] + info[1]['rows'] == 28804
    got = pa.Table.from_batches([ndist.merge_and_finish(engines[0], plan, partials)])
    exp = oq.run_query(query, sql, warehouse)
    assert rows_equal(got.to_pylist(), exp.to_pylist())
    for e in engines:
        e.close()


def test_power_and_validate_cli_as_subprocesses(warehouse, tmp_path):
    """The drop-in surface itself: `nds_power.py <input_prefix> <stream> <time_log> ...` and
    `nds_validate.py <input1> <input2> <stream> ...` run as scripts, like the reference's spark-submit templates do."""
    import json
    import subprocess
    import sys
    import pyarrow.parquet as pq
    from spark_rapids_benchmarks_b200.session import DataFrame
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    stream = str(tmp_path / 'query_0.sql')
    streams.generate_stream(stream, overrides=PARAMS[0])
    props = tmp_path / 'b200.properties'
    props.write_text('b200.source=host\nspark.sql.adaptive.enabled=true\n')
    out_gpu, out_cpu, js, log = str(tmp_path / 'gpu'), str(tmp_path / 'cpu'), str(tmp_path / 'json'), str(tmp_path / 'time.csv')
    r = subprocess.run([sys.executable, os.path.join(root, 'spark_rapids_benchmarks_b200', 'nds_power.py'), warehouse, stream, log,
                        '--output_prefix', out_gpu, '--json_summary_folder', js, '--property_file', str(props),
                        '--sub_queries', 'query3,query7,query19'], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = open(log).read().strip().split('\n')
    assert lines[0] == 'application_id,query,time/milliseconds'
    assert [l.split(',')[1] for l in lines[-7:]] == ['query3', 'query7', 'query19', 'Power Start Time', 'Power End Time',
                                                    'Power Test Time', 'Total Time']
    summaries = sorted(os.listdir(js))
    assert len(summaries) == 3 and all(s.startswith('b200-query') for s in summaries)     # '<property file stem>-<query>-<start>.json'
    summ = json.load(open(os.path.join(js, summaries[0])))
    assert summ['queryStatus'] == ['Completed'] and summ['env']['sparkConf']['b200.source'] == 'host'
    qd = nds_power.gen_sql_from_stream(stream)
    for q in ('query3', 'query7', 'query19'):
        t = oq.run_query(q, qd[q], warehouse)
        os.makedirs(os.path.join(out_cpu, q))
        pq.write_table(nds_power.ensure_valid_column_names(DataFrame(t)).toArrow(), os.path.join(out_cpu, q, 'part-0.parquet'))
    v = subprocess.run([sys.executable, os.path.join(root, 'spark_rapids_benchmarks_b200', 'nds_validate.py'), out_gpu, out_cpu, stream,
                        '--sub_queries', 'query3,query7,query19', '--json_summary_folder', js], capture_output=True, text=True,
                       cwd=str(tmp_path))
    assert v.returncode == 0, v.stdout[-2000:] + v.stderr[-2000:]
    assert 'Unmatch Queries' not in v.stdout and v.stdout.count('Results match') == 3
    assert json.load(open(os.path.join(js, summaries[0])))['queryValidationStatus'] == ['Pass']


@pytest.mark.parametrize('query', streams.SUBSET)
def test_fused_exchange_plan_on_one_engine(gpu_engine, warehouse, query):
    """The single-plan form of the distributed query (partial aggregate -> exchange -> merge aggregate -> sort):
    with one rank the exchange passes through, with exchange_loopback=2 every partial state arrives twice, which
    doubles SUM/COUNT and leaves AVG and the row order alone - checked against the oracle either way."""
    from spark_rapids_benchmarks_b200 import planner
    import pyarrow as pa
    sql = streams.render(query, overrides=PARAMS[0])
    plan = planner.plan_sql(sql, distributed=True)
    exp = oq.run_query(query, sql, warehouse)
    got = pa.Table.from_batches([gpu_engine.execute({'root': plan['fused']})])
    assert rows_equal(got.to_pylist(), exp.to_pylist())
    twice = pa.Table.from_batches([gpu_engine.execute({'root': plan['fused']}, options={'exchange_loopback': 2})])
    assert twice.num_rows == exp.num_rows
    if query == 'query7':          # averages only: unchanged
        assert rows_equal(twice.to_pylist(), exp.to_pylist())


def test_exchange_over_nccl_two_gpus(warehouse, tmp_path):
    """The real thing: two processes, two GPUs, engines joined by nds_engine_comm_init, every query of the subset
    through Session.sql() with the exchange in device memory - compared with the oracle on rank 0.  Skipped on a
    one-GPU box (run it with `gpurun --gpus 2`)."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, NDS_TEST_WAREHOUSE=warehouse)
    # own process group, killed as a whole on a timeout: a worker left behind in a collective would keep both GPUs busy
    import signal
    p = subprocess.Popen([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
                          '--master-port', '29533', os.path.join(root, 'tests', 'nccl_exchange_worker.py')],
                         cwd=root, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, start_new_session=True)
    try:
        out, errs = p.communicate(timeout=300)
    except subprocess.TimeoutExpired:
        os.killpg(p.pid, signal.SIGKILL)
        out, errs = p.communicate()
        raise AssertionError('the two-GPU exchange run did not finish in 300 s\n' + out[-2000:] + errs[-3000:])

    class r:
        returncode, stdout, stderr = p.returncode, out, errs
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert 'EXCHANGE-OK 6' in r.stdout


def test_throughput_two_streams_share_one_gpu(warehouse, tmp_path):
    """Throughput mode (reference `nds-throughput`): two power-run processes at once on one GPU, each with its own
    engine and stream; both must complete, both results must equal the oracle's, and the elapsed time comes out of the
    two time logs the way nds_bench computes it."""
    import subprocess
    import sys
    import pyarrow.parquet as pq
    from spark_rapids_benchmarks_b200 import nds_throughput as thr
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    params = {1: PARAMS[0], 2: PARAMS[1]}
    for k, ov in params.items():
        streams.generate_stream(str(tmp_path / ('query_%d.sql' % k)), stream=k, overrides=ov)
    r = subprocess.run([sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_throughput', '1,2',
                        sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_power', warehouse, str(tmp_path / 'query_{}.sql'),
                        str(tmp_path / 'time_{}.csv'), '--output_prefix', str(tmp_path / 'out_{}')],
                       capture_output=True, text=True, cwd=root)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    for k in (1, 2):
        qd = nds_power.gen_sql_from_stream(str(tmp_path / ('query_%d.sql' % k)))
        assert len(qd) == len(streams.SUBSET)
        for q, sql in qd.items():
            got = pq.ParquetDataset(str(tmp_path / ('out_%d' % k) / q)).read()
            exp = oq.run_query(q, sql, warehouse)
            assert rows_equal(got.to_pylist(), [dict(zip(got.schema.names, row.values())) for row in exp.to_pylist()]), (k, q)
    elapsed = thr.get_throughput_time(str(tmp_path / 'time'), 5, 1)        # 5 streams in total -> first test = streams 1, 2
    # the time log carries whole epoch seconds (like the reference's), so two sub-second streams may give 0.0
    assert 0 <= elapsed < 600 and abs(elapsed * 10 - round(elapsed * 10)) < 1e-9


def test_repeated_queries_do_not_leak(gpu_engine, warehouse):
    """The power run's loop in miniature: 150 executions of two plans (one of them failing every few rounds) must
    leave device memory, the event pool and the side streams where they were - and keep giving the same answer."""
    import torch
    from spark_rapids_benchmarks_b200 import planner
    from spark_rapids_benchmarks_b200.engine import EngineError
    import pyarrow as pa
    plans = [planner.plan_sql(streams.render(q, overrides=PARAMS[0])) for q in ('query7', 'query42')]
    bad = {'op': 'join', 'type': 'inner', 'probe_key': 'ss_item_sk', 'build_key': 'i_manufact_id',       # duplicate build keys -> error
           'probe': {'op': 'scan', 'table': 'store_sales', 'columns': ['ss_item_sk']},
           'build': {'op': 'scan', 'table': 'item', 'columns': ['i_manufact_id']}}
    first = [pa.Table.from_batches([gpu_engine.execute({'root': p['root']})]).to_pylist() for p in plans]
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for it in range(75):
        for p, exp in zip(plans, first):
            got = pa.Table.from_batches([gpu_engine.execute({'root': p['root']}, source='host' if it % 2 else None)]).to_pylist()
            assert got == exp
        if it % 10 == 0:
            with pytest.raises(EngineError):
                gpu_engine.execute(bad)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    assert free0 - free1 < (64 << 20), (free0, free1)       # the async pool may keep a few blocks; nothing grows per query


def test_small_slab_overflows_then_grows(warehouse):
    """An engine whose per-query slab is far too small: the first query falls back to stream-ordered allocations, the
    slab is grown before the next one - same answers throughout."""
    from spark_rapids_benchmarks_b200 import engine, nds_schema, planner
    import pyarrow as pa
    eng = engine.Engine(device=0, slab_mb=1)
    try:
        for t in ('store_sales', 'date_dim', 'item', 'customer_demographics', 'promotion'):
            eng.register_table(t, os.path.join(warehouse, t))
        sql = streams.render('query7', overrides=PARAMS[0])
        exp = oq.run_query('query7', sql, warehouse)
        plan = planner.plan_sql(sql)
        for _ in range(3):
            got = pa.Table.from_batches([eng.execute({'root': plan['root']})])
            assert rows_equal(got.to_pylist(), exp.to_pylist())
    finally:
        eng.close()


def test_predicate_forms_and_min_max_count_through_sql(gpu_engine, warehouse):
    """SQL shapes outside the six templates, through the whole path (parser -> planner -> engine): IN lists, BETWEEN, OR,
    explicit JOIN ... ON, MIN / MAX / COUNT(*) / COUNT(col) next to SUM - against pyarrow on the same files."""
    import pyarrow as pa
    import pyarrow.compute as pc
    sql = '''select i_category_id, d_moy, sum(ss_ext_sales_price) s, min(ss_sales_price) lo, max(ss_sales_price) hi,
                    count(*) n, count(ss_coupon_amt) nc
             from store_sales join date_dim on ss_sold_date_sk = d_date_sk join item on ss_item_sk = i_item_sk
             where d_moy in (1, 2, 12) and d_year between 1998 and 1999 and (i_category_id = 3 or i_manager_id between 10 and 40)
             group by i_category_id, d_moy
             order by i_category_id, d_moy limit 100'''
    got = _Sess(gpu_engine).sql(sql).toArrow()
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_moy'])
    dd = dd.filter(pc.and_(pc.is_in(dd['d_moy'], value_set=pa.array([1, 2, 12])),
                           pc.and_(pc.greater_equal(dd['d_year'], 1998), pc.less_equal(dd['d_year'], 1999))))
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category_id', 'i_manager_id'])
    keep = pc.or_kleene(pc.equal(it['i_category_id'], 3),
                        pc.and_kleene(pc.greater_equal(it['i_manager_id'], 10), pc.less_equal(it['i_manager_id'], 40)))
    it = it.filter(pc.fill_null(keep, False))
    fact = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk',
                        ['ss_sold_date_sk', 'ss_item_sk', 'ss_ext_sales_price', 'ss_sales_price', 'ss_coupon_amt'], dd['d_date_sk'].to_pylist())
    j = fact.join(dd, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='inner').join(it, keys='ss_item_sk', right_keys='i_item_sk',
                                                                                               join_type='inner')
    g = j.group_by(['i_category_id', 'd_moy']).aggregate([('ss_ext_sales_price', 'sum'), ('ss_sales_price', 'min'), ('ss_sales_price', 'max'),
                                                         ([], 'count_all'), ('ss_coupon_amt', 'count')])
    rows = sorted(g.to_pylist(), key=lambda r: ((r['i_category_id'] is not None, r['i_category_id']), (r['d_moy'] is not None, r['d_moy'])))[:100]
    assert got.num_rows == len(rows) and got.num_rows > 0
    exp_cols = ['i_category_id', 'd_moy', 'ss_ext_sales_price_sum', 'ss_sales_price_min', 'ss_sales_price_max', 'count_all', 'ss_coupon_amt_count']
    for name, ename in zip(got.schema.names, exp_cols):
        assert got.column(name).to_pylist() == [r[ename] for r in rows], name


def test_aggregates_over_expressions_through_sql(gpu_engine, warehouse):
    """Arithmetic inside and around aggregates (decimal - decimal, bigint * decimal, CASE WHEN, int + literal, MAX - MIN,
    CAST to double, decimal * literal) through the SQL path; expectation from plain Python Decimal / int arithmetic over
    the rows pyarrow joins.  Exact for integers and decimals, 1e-9 relative for the double."""
    import decimal
    import pyarrow as pa
    import pyarrow.compute as pc
    sql = '''select i_category_id,
                    sum(ss_ext_sales_price - ss_coupon_amt) a, sum(ss_quantity * ss_sales_price) b,
                    sum(case when i_manager_id < 50 then ss_ext_sales_price else 0 end) c, sum(ss_quantity + 1) d,
                    max(ss_quantity) - min(ss_quantity) e, sum(cast(ss_quantity as double)) f, sum(ss_ext_sales_price * 2) g
             from store_sales, date_dim, item
             where ss_sold_date_sk = d_date_sk and ss_item_sk = i_item_sk and d_year = 1998 and d_moy = 1
             group by i_category_id order by i_category_id limit 100'''
    got = _Sess(gpu_engine).sql(sql).toArrow().to_pylist()
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_moy'])
    dd = dd.filter(pc.and_(pc.equal(dd['d_year'], 1998), pc.equal(dd['d_moy'], 1)))
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category_id', 'i_manager_id'])
    fact = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk',
                        ['ss_sold_date_sk', 'ss_item_sk', 'ss_ext_sales_price', 'ss_sales_price', 'ss_coupon_amt', 'ss_quantity'],
                        dd['d_date_sk'].to_pylist())
    j = fact.join(dd, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='inner').join(it, keys='ss_item_sk', right_keys='i_item_sk',
                                                                                               join_type='inner')
    groups = {}
    for r in j.to_pylist():
        groups.setdefault(r['i_category_id'], []).append(r)

    def ssum(vals):
        vals = [v for v in vals if v is not None]
        return sum(vals) if vals else None

    def both(x, y, f):
        return None if x is None or y is None else f(x, y)
    exp = []
    for cat in sorted(groups, key=lambda c: (c is not None, c)):
        rs = groups[cat]
        q = [r['ss_quantity'] for r in rs if r['ss_quantity'] is not None]
        exp.append({
            'i_category_id': cat,
            'a': ssum(both(r['ss_ext_sales_price'], r['ss_coupon_amt'], lambda x, y: x - y) for r in rs),
            'b': ssum(both(r['ss_quantity'], r['ss_sales_price'], lambda x, y: x * y) for r in rs),
            'c': ssum((r['ss_ext_sales_price'] if (r['i_manager_id'] is not None and r['i_manager_id'] < 50) else decimal.Decimal(0)) for r in rs),
            'd': ssum(both(r['ss_quantity'], 1, lambda x, y: x + y) for r in rs),
            'e': (max(q) - min(q)) if q else None,
            'f': float(sum(q)) if q else None,
            'g': ssum(both(r['ss_ext_sales_price'], 2, lambda x, y: x * y) for r in rs),
        })
    exp = exp[:100]
    assert len(got) == len(exp) and len(got) > 3
    for g, e in zip(got, exp):
        for k in e:
            if k == 'f' and e[k] is not None:
                assert g[k] is not None and abs(g[k] - e[k]) <= 1e-9 * max(1.0, abs(e[k])), (k, g, e)
            else:
                assert g[k] == e[k], (k, g[k], e[k], g['i_category_id'])


def test_non_aggregating_and_single_table_queries_through_sql(gpu_engine, warehouse):
    """Shapes without GROUP BY or without a join: a filtered star join ordered by a measure (top-k over raw rows), a
    group-by over a computed string key on one dimension table, and a plain ORDER BY ... DESC NULLS FIRST - vs pyarrow."""
    import pyarrow as pa
    import pyarrow.compute as pc
    sess = _Sess(gpu_engine)
    # 1. top-k over joined rows
    got = sess.sql('''select ss_ticket_number, ss_item_sk, ss_quantity, i_brand from store_sales, item, date_dim
                      where ss_item_sk = i_item_sk and ss_sold_date_sk = d_date_sk and d_year = 1998 and d_moy = 2 and ss_quantity > 90
                      order by ss_quantity desc, ss_ticket_number, ss_item_sk limit 50''').toArrow().to_pylist()
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_moy'])
    dd = dd.filter(pc.and_(pc.equal(dd['d_year'], 1998), pc.equal(dd['d_moy'], 2)))
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_brand'])
    fact = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_item_sk', 'ss_ticket_number', 'ss_quantity'],
                        dd['d_date_sk'].to_pylist())
    fact = fact.filter(pc.fill_null(pc.greater(fact['ss_quantity'], 90), False))
    j = fact.join(it, keys='ss_item_sk', right_keys='i_item_sk', join_type='inner')
    rows = sorted(j.select(['ss_ticket_number', 'ss_item_sk', 'ss_quantity', 'i_brand']).to_pylist(),
                  key=lambda r: (-r['ss_quantity'], r['ss_ticket_number'], r['ss_item_sk']))[:50]
    assert got == rows and len(got) > 5
    # 2. group by a computed string key, one table
    got = sess.sql('''select substr(i_brand, 1, 6) b, count(*) n from item where i_manager_id < 20
                      group by substr(i_brand, 1, 6) order by n desc, b limit 10''').toArrow().to_pylist()
    it = oq.read_dim(warehouse, 'item', ['i_brand', 'i_manager_id'])
    counts = {}
    for r in it.to_pylist():
        if r['i_manager_id'] is not None and r['i_manager_id'] < 20:
            k = None if r['i_brand'] is None else r['i_brand'][:6]
            counts[k] = counts.get(k, 0) + 1
    exp = sorted(({'b': k, 'n': v} for k, v in counts.items()), key=lambda r: (-r['n'], (r['b'] is not None, r['b'])))[:10]   # NULLS FIRST asc
    assert got == exp and len(got) > 1
    # 3. ORDER BY ... DESC NULLS FIRST on strings
    got = sess.sql('''select i_brand, i_class from item where i_category_id = 2
                      order by i_brand desc nulls first, i_class limit 20''').toArrow().to_pylist()
    it = oq.read_dim(warehouse, 'item', ['i_brand', 'i_class', 'i_category_id'])
    rs = [{'i_brand': r['i_brand'], 'i_class': r['i_class']} for r in it.to_pylist() if r['i_category_id'] == 2]
    import functools

    def cmp(x, y):
        a, b = x['i_brand'], y['i_brand']
        if a != b:
            if a is None:
                return -1
            if b is None:
                return 1
            return -1 if a.encode() > b.encode() else 1            # descending
        a, b = x['i_class'], y['i_class']
        if a == b:
            return 0
        if a is None:
            return -1                                               # ascending: NULLS FIRST
        if b is None:
            return 1
        return -1 if a.encode() < b.encode() else 1
    exp = sorted(rs, key=functools.cmp_to_key(cmp))[:20]
    assert got == exp and len(got) > 1


def test_string_predicates_through_sql(gpu_engine, warehouse):
    """String IN lists, <>, >= / < (byte order) and IS [NOT] NULL on dictionary-encoded string columns, vs plain Python."""
    sess = _Sess(gpu_engine)
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category', 'i_class', 'i_brand']).to_pylist()
    cats = sorted({r['i_category'] for r in it if r['i_category'] is not None})
    classes = sorted({r['i_class'] for r in it if r['i_class'] is not None})
    brands = sorted({r['i_brand'] for r in it if r['i_brand'] is not None}, key=lambda s: s.encode())
    pick, bad_class, mid = cats[:2] + cats[-1:], classes[len(classes) // 2], brands[len(brands) // 2]
    sql = """select i_item_sk, i_category, i_class, i_brand from item
             where i_category in (%s) and i_class <> '%s' and (i_brand >= '%s' or i_brand is null) and i_class is not null
             order by i_item_sk limit 100000""" % (', '.join("'%s'" % c.replace("'", "''") for c in pick), bad_class.replace("'", "''"),
                                                    mid.replace("'", "''"))
    got = sess.sql(sql).toArrow().to_pylist()
    exp = [r for r in it if r['i_category'] in pick and r['i_class'] is not None and r['i_class'] != bad_class and
           (r['i_brand'] is None or r['i_brand'].encode() >= mid.encode())]
    exp.sort(key=lambda r: r['i_item_sk'])
    assert got == exp and len(got) > 0


@pytest.mark.parametrize('fact,pcol,sql,dims,measure', [
    ('catalog_sales', 'cs_sold_date_sk',
     '''select i_category_id k, sum(cs_ext_sales_price) s, count(*) n from catalog_sales, date_dim, item
        where cs_sold_date_sk = d_date_sk and cs_item_sk = i_item_sk and d_year = 1998 and d_moy = 1
        group by i_category_id order by i_category_id limit 100''',
     {'year': 1998, 'moy': 1, 'item_key': 'cs_item_sk', 'group': 'i_category_id'}, 'cs_ext_sales_price'),
    ('web_returns', 'wr_returned_date_sk',
     '''select d_moy k, sum(wr_return_amt) s, count(*) n from web_returns, date_dim
        where wr_returned_date_sk = d_date_sk and d_year = 1998 group by d_moy order by d_moy limit 100''',
     {'year': 1998, 'moy': None, 'item_key': None, 'group': 'd_moy'}, 'wr_return_amt'),
    ('inventory', 'inv_date_sk',
     '''select inv_warehouse_sk k, sum(inv_quantity_on_hand) s, count(*) n from inventory, date_dim
        where inv_date_sk = d_date_sk and d_year = 1998 and d_moy = 1 group by inv_warehouse_sk order by inv_warehouse_sk limit 100''',
     {'year': 1998, 'moy': 1, 'item_key': None, 'group': 'inv_warehouse_sk'}, 'inv_quantity_on_hand'),
])
def test_other_fact_tables_prune_and_aggregate(gpu_engine, warehouse, fact, pcol, sql, dims, measure):
    """The engine is not store_sales-specific: the same star join + partition pruning + group-by on catalog_sales,
    web_returns and inventory (each partitioned by its own date key), vs pyarrow."""
    import pyarrow.compute as pc
    got = _Sess(gpu_engine).sql(sql).toArrow().to_pylist()
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_moy'])
    cond = pc.equal(dd['d_year'], dims['year'])
    if dims['moy'] is not None:
        cond = pc.and_(cond, pc.equal(dd['d_moy'], dims['moy']))
    dd = dd.filter(cond)
    cols = [pcol, measure] + ([dims['item_key']] if dims['item_key'] else []) + ([dims['group']] if dims['group'].startswith('inv_') else [])
    f = oq.read_fact(warehouse, fact, pcol, cols, dd['d_date_sk'].to_pylist())
    j = f.join(dd, keys=pcol, right_keys='d_date_sk', join_type='inner')
    if dims['item_key']:
        it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category_id'])
        j = j.join(it, keys=dims['item_key'], right_keys='i_item_sk', join_type='inner')
    groups = {}
    for r in j.to_pylist():
        g = groups.setdefault(r[dims['group']], [None, 0])
        g[1] += 1
        if r[measure] is not None:
            g[0] = r[measure] if g[0] is None else g[0] + r[measure]
    exp = [{'k': k, 's': v[0], 'n': v[1]} for k, v in sorted(groups.items(), key=lambda kv: (kv[0] is not None, kv[0]))][:100]
    assert got == exp and len(got) > 0


def test_null_group_keys_through_sql(gpu_engine, warehouse):
    """GROUP BY over nullable fact columns: NULL keys form their own groups and sort first (ascending); vs plain Python."""
    import pyarrow.compute as pc
    got = _Sess(gpu_engine).sql('''select ss_store_sk, ss_promo_sk, count(*) n, sum(ss_net_paid) s, count(ss_net_paid) c
                                   from store_sales, date_dim where ss_sold_date_sk = d_date_sk and d_year = 1998 and d_moy = 1
                                   group by ss_store_sk, ss_promo_sk order by ss_store_sk, ss_promo_sk limit 1000''').toArrow().to_pylist()
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_moy'])
    dd = dd.filter(pc.and_(pc.equal(dd['d_year'], 1998), pc.equal(dd['d_moy'], 1)))
    f = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_store_sk', 'ss_promo_sk', 'ss_net_paid'],
                     dd['d_date_sk'].to_pylist())
    groups = {}
    for r in f.to_pylist():
        g = groups.setdefault((r['ss_store_sk'], r['ss_promo_sk']), [0, None, 0])
        g[0] += 1
        if r['ss_net_paid'] is not None:
            g[1] = r['ss_net_paid'] if g[1] is None else g[1] + r['ss_net_paid']
            g[2] += 1
    key = lambda kv: tuple((v is not None, v if v is not None else 0) for v in kv[0])
    exp = [{'ss_store_sk': k[0], 'ss_promo_sk': k[1], 'n': v[0], 's': v[1], 'c': v[2]} for k, v in sorted(groups.items(), key=key)][:1000]
    assert got == exp
    assert any(r['ss_store_sk'] is None for r in got) and any(r['ss_promo_sk'] is None for r in got)


def test_host_source_equals_resident_source_on_other_shapes(gpu_engine):
    """Pages copied from pinned host memory inside the query (`b200.source=host`) give the same rows as pages resident in
    HBM, for fact tables and query shapes the six templates do not touch."""
    sqls = ['''select i_category_id k, sum(cs_ext_sales_price) s, count(*) n from catalog_sales, date_dim, item
               where cs_sold_date_sk = d_date_sk and cs_item_sk = i_item_sk and d_year = 1998 group by i_category_id order by i_category_id limit 100''',
            '''select inv_warehouse_sk k, sum(inv_quantity_on_hand) s from inventory, date_dim where inv_date_sk = d_date_sk and d_moy = 1
               group by inv_warehouse_sk order by inv_warehouse_sk limit 100''',
            '''select sr_item_sk, sr_ticket_number, sr_return_amt from store_returns where sr_return_quantity > 95
               order by sr_return_amt desc, sr_ticket_number, sr_item_sk limit 40''']
    for sql in sqls:
        a = _Sess(gpu_engine, 'hbm').sql(sql).toArrow()
        b = _Sess(gpu_engine, 'host').sql(sql).toArrow()
        assert a.num_rows > 0 and a.equals(b), sql


@pytest.mark.parametrize('query', streams.SUBSET)
@pytest.mark.parametrize('opts', [{'fused_scan': False}, {'dense_tables': False}, {'fused_scan': False, 'dense_tables': False},
                                  {'speculate': False}, {'topk_stream': False}, {'probe_bloom': True, 'probe_bloom_chunks': 1},
                                  {'dict_match_chunks': 1}, {'dict_match_chunks': 1, 'dense_tables': False}, {'dict_match': False},
                                  {'build_filter': False}])
def test_routes_agree_with_the_oracle(gpu_engine, warehouse, query, opts):
    """The fact scan has two routes - fused page decode -> probe with late materialisation (fused_scan.cu), and decode
    every column then probe (decode_fast.cu + ops.cu) - and a join table two layouts (direct-address over the key
    range the column statistics give, open addressing).  Every combination must give the oracle's rows."""
    from spark_rapids_benchmarks_b200 import planner
    import pyarrow as pa
    sql = streams.render(query, overrides=PARAMS[0])
    plan = planner.plan_sql(sql)
    exp = oq.run_query(query, sql, warehouse)
    for source in (None, 'host'):
        got = pa.Table.from_batches([gpu_engine.execute({'root': plan['root']}, source=source, options=opts)])
        assert rows_equal(got.to_pylist(), exp.to_pylist()), (query, opts, source)
        ops = [o['op'] for o in gpu_engine.last_stats()['ops']]
        assert ('scan_probe' in ops) == opts.get('fused_scan', True), ops
        assert ('dict_match' in ops) == ('dict_match_chunks' in opts), ops


def test_fused_scan_keeps_key_columns_the_query_reads(gpu_engine, warehouse):
    """A probe key that is also grouped on / selected is a payload column too: the fused scan must hand it on."""
    import pyarrow as pa
    import pyarrow.compute as pc
    sql = ('select ss_item_sk, ss_sold_date_sk, count(*) cnt, sum(ss_quantity) q from store_sales, item, date_dim '
           'where ss_item_sk = i_item_sk and ss_sold_date_sk = d_date_sk and i_manager_id = 63 and d_moy = 1 '
           'group by ss_item_sk, ss_sold_date_sk order by ss_item_sk, ss_sold_date_sk limit 100')
    got = _Sess(gpu_engine).sql(sql).toArrow()
    assert any(o['op'] == 'scan_probe' for o in gpu_engine.last_stats()['ops'])
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_manager_id'])
    it = it.filter(pc.equal(it['i_manager_id'], 63).fill_null(False))
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_moy'])
    dd = dd.filter(pc.equal(dd['d_moy'], 1).fill_null(False))
    ss = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_item_sk', 'ss_quantity'], dd['d_date_sk'].to_pylist())
    j = ss.join(it, keys='ss_item_sk', right_keys='i_item_sk', join_type='inner').join(dd, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='inner')
    g = j.group_by(['ss_item_sk', 'ss_sold_date_sk']).aggregate([([], 'count_all'), ('ss_quantity', 'sum')])
    exp = sorted(((r['ss_item_sk'], r['ss_sold_date_sk'], r['count_all'], r['ss_quantity_sum']) for r in g.to_pylist()))[:100]
    assert [(r['ss_item_sk'], r['ss_sold_date_sk'], r['cnt'], r['q']) for r in got.to_pylist()] == exp


def test_keyless_aggregate_over_a_star_join(gpu_engine, warehouse):
    """SELECT aggregates without GROUP BY through SQL: one row; and an empty selection still gives one row."""
    import decimal
    import pyarrow.compute as pc
    sql = ('select count(*) n, sum(ss_quantity) q, avg(ss_sales_price) p, min(ss_item_sk) lo, max(ss_item_sk) hi '
           'from store_sales, date_dim where ss_sold_date_sk = d_date_sk and d_moy = %d')
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_moy'])
    for moy in (1, 7):                      # the tiny warehouse has sales in January and February only
        keys = dd.filter(pc.equal(dd['d_moy'], moy).fill_null(False))['d_date_sk'].to_pylist()
        ss = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_quantity', 'ss_sales_price', 'ss_item_sk'], keys)
        got = _Sess(gpu_engine).sql(sql % moy).toArrow().to_pylist()
        if ss.num_rows == 0:
            assert got == [{'n': 0, 'q': None, 'p': None, 'lo': None, 'hi': None}]
            continue
        total, cnt = pc.sum(ss['ss_sales_price']).as_py(), pc.count(ss['ss_sales_price']).as_py()
        avg = (decimal.Decimal(total) / cnt).quantize(decimal.Decimal('0.000001'), rounding=decimal.ROUND_HALF_UP)
        assert got == [{'n': ss.num_rows, 'q': pc.sum(ss['ss_quantity']).as_py(), 'p': avg, 'lo': pc.min(ss['ss_item_sk']).as_py(),
                        'hi': pc.max(ss['ss_item_sk']).as_py()}]
