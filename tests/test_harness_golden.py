"""CPU: the host-side mirror of the reference harness against golden vectors recorded from the
reference's own Python (tests/golden/make_reference_goldens.py, run where /root/reference exists)."""
import contextlib
import decimal
import glob
import io
import json
import os

import pyarrow as pa
import pytest

from spark_rapids_benchmarks_b200 import nds_power, nds_validate
from spark_rapids_benchmarks_b200.PysparkBenchReport import PysparkBenchReport
from spark_rapids_benchmarks_b200.nds_gen_query_stream import split_special_query
from spark_rapids_benchmarks_b200.session import DataFrame

G = json.load(open(os.path.join(os.path.dirname(__file__), 'golden', 'reference_harness.json')))


def dec(v):
    if isinstance(v, dict):
        return decimal.Decimal(v['dec']) if 'dec' in v else float(v['f'])
    return v


def test_gen_sql_from_stream(tmp_path):
    p = tmp_path / 'query_0.sql'
    p.write_text(G['stream_text'])
    got = nds_power.gen_sql_from_stream(str(p))
    assert [list(kv) for kv in got.items()] == G['gen_sql_from_stream']


def test_split_special_query():
    assert list(split_special_query(G['split_special_query']['input'])) == G['split_special_query']['output']


def test_get_query_subset(tmp_path):
    p = tmp_path / 'query_0.sql'
    p.write_text(G['stream_text'])
    d = nds_power.gen_sql_from_stream(str(p))
    assert list(nds_power.get_query_subset(d, G['get_query_subset']['subset']).keys()) == G['get_query_subset']['keys']
    with pytest.raises(Exception) as ei:
        nds_power.get_query_subset(d, ['query1'])
    assert str(ei.value) == G['get_query_subset']['missing_error']


def test_ensure_valid_column_names():
    for case in G['ensure_valid_column_names']:
        df = DataFrame(pa.table([pa.array([1])] * len(case['in']), names=case['in']))
        assert nds_power.ensure_valid_column_names(df).columns == case['out']


def test_load_properties(tmp_path):
    p = tmp_path / 'x.properties'
    p.write_text(G['load_properties']['text'])
    assert nds_power.load_properties(str(p)) == G['load_properties']['out']


def test_compare_matrix():
    assert len(G['compare']) > 900
    for a, b, eps, want in G['compare']:
        a, b = dec(a), dec(b)
        try:
            got = bool(nds_validate.compare(a, b, eps))
        except Exception as e:
            got = 'EXC:' + type(e).__name__
        assert got == want, (a, b, eps, got, want)


def test_row_equal_and_q78():
    for c in G['rowEqual']['cases']:
        assert bool(nds_validate.rowEqual(list(c['row1']), list(c['row2']), c['epsilon'], c['is_q78'], c['col'])) == c['out'], c
    with pytest.raises(Exception) as ei:
        nds_validate.rowEqual([1, 2, 3], [1, 2, 3], 0.00001, True, 3)
    assert str(ei.value) == G['rowEqual']['bad_col_error']
    assert nds_validate.check_nth_col_problematic_q78(G['check_nth_col_problematic_q78']['text']) == G['check_nth_col_problematic_q78']['out']


class _Sess:
    version = '3.1.2-fake'

    def conf_items(self):
        return {'spark.app.name': 'x', 'k': 'v'}.items()


def test_bench_report(tmp_path, monkeypatch):
    g = G['bench_report']
    monkeypatch.setenv('TOKEN', 'secret-token')
    monkeypatch.setenv('MY_TOKEN', 'kept')
    rep = PysparkBenchReport(_Sess(), 'query3')
    s = rep.report_on(lambda a, b: None, 1, 2)
    assert sorted(s.keys()) == g['summary_keys']
    assert sorted(s['env'].keys()) == g['env_keys']
    assert s['queryStatus'] == g['status_ok'] and rep.is_success() == g['is_success_ok']
    assert ('TOKEN' not in s['env']['envVars']) == g['token_redacted']
    assert ('MY_TOKEN' in s['env']['envVars']) == g['my_token_kept']
    assert s['env']['sparkConf'] == g['spark_conf'] and s['env']['sparkVersion'] == g['spark_version']
    assert len(s['queryTimes']) == g['times_len'] and isinstance(s['queryTimes'][0], int) and isinstance(s['startTime'], int)
    prefix = str(tmp_path / 'aqe-on')
    rep.write_summary(prefix=prefix)
    files = glob.glob(prefix + '-query3-*.json')
    assert len(files) == 1
    written = json.load(open(files[0]))
    assert sorted(written.keys()) == g['written_keys']
    assert os.path.basename(written['filename']) == 'aqe-on-query3-%d.json' % written['startTime']
    assert open(files[0]).read().startswith('{\n  "env"')          # indent=2, insertion order
    rep2 = PysparkBenchReport(_Sess(), 'query7')

    def boom():
        raise ValueError('boom')
    with contextlib.redirect_stdout(io.StringIO()) as out:
        s2 = rep2.report_on(boom)
    assert s2['queryStatus'] == g['status_fail'] and s2['exceptions'] == g['exceptions_fail']
    assert rep2.is_success() == g['is_success_fail']
    assert 'ERROR BEGIN' in out.getvalue() and 'ERROR END' in out.getvalue()


def test_update_summary(tmp_path):
    for q, st in (('query3', ['Completed']), ('query7', ['Failed']), ('query19', ['CompletedWithTaskFailures']), ('query42', ['Completed'])):
        json.dump({'query': q, 'queryStatus': st}, open(tmp_path / ('p-%s-123.json' % q), 'w'))
    nds_validate.update_summary(str(tmp_path), ['query7', 'query19', 'query42'], queries={q: '' for q in G['update_summary']})
    got = {q: json.load(open(tmp_path / ('p-%s-123.json' % q)))['queryValidationStatus'] for q in G['update_summary']}
    assert got == G['update_summary']
    json.dump({}, open(tmp_path / 'q-query3-9.json', 'w'))
    with pytest.raises(Exception):
        nds_validate.update_summary(str(tmp_path), [], queries={'query3': ''})


class StubSession:
    application_id = 'b200-test'
    version = 'stub'

    def __init__(self, fail=()):
        self.fail = set(fail)
        self.registered = []

    def conf_items(self):
        return {}.items()

    def register_table(self, name, path, fmt):
        self.registered.append(name)

    def sql(self, q):
        for f in self.fail:
            if f + '.tpl' in q:
                raise RuntimeError('unsupported plan node for ' + f)
        return DataFrame(pa.table({'sum(x)': [1, 2], 'a': ['p', 'q'], 'a ': [1.5, None]}))

    def stop(self):
        pass


def test_time_log_layout_and_exit_code(tmp_path):
    p = tmp_path / 'query_0.sql'
    p.write_text(G['stream_text'])
    qd = nds_power.gen_sql_from_stream(str(p))
    log, out, js = tmp_path / 'time.csv', tmp_path / 'out', tmp_path / 'json'
    sess = StubSession()
    with contextlib.redirect_stdout(io.StringIO()):
        nds_power.run_query_stream(str(tmp_path), None, qd, str(log), str(tmp_path / 'extra.csv'), None, output_path=str(out),
                                   json_summary_folder=str(js), session=sess)
    lines = log.read_text().strip().split('\n')
    assert lines[0] == G['time_log']['header']
    assert [l.split(',')[1] for l in lines[1:]] == G['time_log']['labels']
    assert len(sess.registered) == 24
    assert (tmp_path / 'extra.csv').read_text() == log.read_text()
    start, end = int(lines[-4].split(',')[2]), int(lines[-3].split(',')[2])
    assert 1.6e9 < start <= end < 4e9                                  # epoch seconds
    for q in qd:
        files = os.listdir(out / q)
        assert any(f.endswith('.parquet') for f in files)
        assert len(glob.glob(str(js / ('-%s-*.json' % q)))) == 1        # no property file -> '<folder>/' prefix + '-'
    import pyarrow.parquet as pq
    t = pq.read_table(glob.glob(str(out / 'query3' / '*.parquet'))[0])
    assert t.schema.names == ['sum_x_', 'a', 'a_']
    # a failing query -> status Failed, exit code 1 unless --allow_failure
    with contextlib.redirect_stdout(io.StringIO()) as so:
        with pytest.raises(SystemExit) as ei:
            nds_power.run_query_stream(str(tmp_path), None, qd, str(log), None, None, session=StubSession(fail=['query98']))
    assert ei.value.code == 1 and 'query98 status: [\'Failed\']' in so.getvalue()
    with contextlib.redirect_stdout(io.StringIO()):
        nds_power.run_query_stream(str(tmp_path), None, qd, str(log), None, ['query3'], allow_failure=True,
                                   session=StubSession(fail=['query3']))
    # non-empty json folder is refused (reference check.py:136-145)
    with contextlib.redirect_stdout(io.StringIO()):
        with pytest.raises(Exception):
            nds_power.run_query_stream(str(tmp_path), None, qd, str(log), None, None, json_summary_folder=str(js), session=StubSession())
    # non-parquet input fails loudly
    with pytest.raises(Exception):
        nds_power.run_query_stream(str(tmp_path), None, qd, str(log), None, None, input_format='orc', session=StubSession())


def test_power_cli_arguments_match_reference():
    p = nds_power.build_parser()
    a = p.parse_args(['in', 'stream.sql', 'time.csv', '--input_format', 'parquet', '--output_prefix', 'o', '--output_format', 'parquet',
                      '--property_file', 'x', '--floats', '--json_summary_folder', 'j', '--delta_unmanaged', '--keep_sc', '--hive',
                      '--extra_time_log', 'e', '--sub_queries', 'query1, query2', '--allow_failure'])
    assert a.sub_queries == ['query1', 'query2'] and a.input_prefix == 'in' and a.time_log == 'time.csv'
    v = nds_validate.build_parser().parse_args(['a', 'b', 's', '--input1_format', 'parquet', '--input2_format', 'parquet', '--max_errors', '3',
                                                '--epsilon', '0.1', '--ignore_ordering', '--use_iterator', '--floats',
                                                '--json_summary_folder', 'j', '--sub_queries', 'query3'])
    assert v.max_errors == 3 and v.epsilon == 0.1 and v.ignore_ordering


def test_validate_two_outputs(tmp_path):
    import pyarrow.parquet as pq
    p = tmp_path / 'query_0.sql'
    p.write_text(G['stream_text'])
    t1 = pa.table({'k': [3, 1, 2, None], 'v': [1.0, 2.0, float('nan'), None], 's': ['a', 'b', None, 'c'],
                   'd': [decimal.Decimal('1.10'), None, decimal.Decimal('2.00'), decimal.Decimal('3.00')]})
    t2 = t1.take([1, 3, 0, 2])
    t3 = t1.set_column(1, 'v', pa.array([1.0, 2.1, float('nan'), None]))
    for name, t in (('r1', t1), ('r2', t2), ('r3', t3)):
        for q in ('query3', 'query14_part1', 'query14_part2', 'query98'):
            os.makedirs(tmp_path / name / q)
            pq.write_table(t, tmp_path / name / q / 'part-0.parquet')
    with contextlib.redirect_stdout(io.StringIO()):
        assert nds_validate.main([str(tmp_path / 'r1'), str(tmp_path / 'r2'), str(p), '--ignore_ordering']) == []
        assert nds_validate.main([str(tmp_path / 'r1'), str(tmp_path / 'r2'), str(p)]) != []
        assert nds_validate.main([str(tmp_path / 'r1'), str(tmp_path / 'r3'), str(p), '--sub_queries', 'query3']) == ['query3']
        assert nds_validate.main([str(tmp_path / 'r1'), str(tmp_path / 'r3'), str(p), '--sub_queries', 'query3', '--epsilon', '0.1']) == []


def test_throughput_arithmetic_matches_reference(tmp_path):
    """nds_throughput's stream ranges, rounding and elapsed time against values recorded from the reference's
    nds_bench.py (get_stream_range, round_up_to_nearest_10_percent, get_throughput_time)."""
    from spark_rapids_benchmarks_b200 import nds_throughput as thr
    for c in G['get_stream_range']:
        assert thr.get_stream_range(c['n'], c['which']) == c['out'], c
    for c in G['round_up_to_nearest_10_percent']:
        assert thr.round_up_to_nearest_10_percent(c['in']) == c['out'], c
    base = str(tmp_path / 'thr')
    for k, (a, b) in G['get_throughput_time']['times'].items():
        with open(base + '_%s.csv' % k, 'w') as f:
            f.write('application_id,query,time/milliseconds\n')
            f.write('app,query1,12\napp,Power Start Time,%r\napp,Power End Time,%r\napp,Power Test Time,10\n' % (a, b))
    for c in G['get_throughput_time']['cases']:
        assert thr.get_throughput_time(base, c['n'], c['which']) == c['out'], c


def test_throughput_runner_substitutes_and_reports_failures(tmp_path):
    """`nds_throughput 1,2,3 <cmd with {}>`: one process per stream, `{}` replaced everywhere, all run concurrently,
    exit code 123 when any stream fails (what xargs returns for the reference script)."""
    import subprocess
    import sys
    import time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / 'w.py'
    script.write_text('import sys, time\nopen(sys.argv[1], "w").write(sys.argv[2]); time.sleep(1.0); sys.exit(int(sys.argv[3]))\n')
    t0 = time.time()
    r = subprocess.run([sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_throughput', '1,2,3', sys.executable, str(script),
                        str(tmp_path / 'out_{}.txt'), 's{}', '0'], cwd=root, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert time.time() - t0 < 2.9            # three one-second streams ran side by side
    assert r.stdout.splitlines()[0] == '3'
    assert [(tmp_path / ('out_%d.txt' % k)).read_text() for k in (1, 2, 3)] == ['s1', 's2', 's3']
    r = subprocess.run([sys.executable, '-m', 'spark_rapids_benchmarks_b200.nds_throughput', '7', sys.executable, str(script),
                        str(tmp_path / 'f_{}.txt'), 'x', '5'], cwd=root, capture_output=True, text=True)
    assert r.returncode == 123


def test_load_report_is_what_the_reference_parsers_read(tmp_path):
    """nds_transcode's report: first line `Load Test Time: <s> seconds`, an 11-digit `RNGSEED used:` line - the two
    fields the reference's nds_bench.get_load_time / get_load_end_timestamp extract (recorded acceptance in G)."""
    from spark_rapids_benchmarks_b200 import nds_schema, nds_transcode
    assert G['load_report']['reference_parsers_accept_it'] is True
    ncol = len(nds_schema.get_schemas(True)['store'].names)
    raw = tmp_path / 'raw' / 'store'
    raw.mkdir(parents=True)
    (raw / 'store_1_1.dat').write_text('|'.join(['1', 'AAAAAAAABAAAAAAA'] + [''] * (ncol - 2)) + '|\n')
    rep = str(tmp_path / 'report.txt')
    args = nds_transcode.build_parser().parse_args([str(tmp_path / 'raw'), str(tmp_path / 'wh'), rep, '--tables', 'store'])
    with contextlib.redirect_stdout(io.StringIO()):
        nds_transcode.transcode(args)
    lines = open(rep).read().split('\n')
    assert lines[0].split(':')[0] == G['load_report']['first_line_prefix'] and lines[0].endswith(' seconds')
    seed = [l for l in lines if 'RNGSEED used:' in l][0].split(':')[1].strip()
    assert seed.isdigit() and len(seed) == G['load_report']['rngseed_digits']


def test_composite_metric_and_report_parsers(tmp_path):
    """nds_bench.get_perf_metric (reference nds_bench.py:334-357) and the report parsers beside it, against values the
    reference's own functions returned."""
    from spark_rapids_benchmarks_b200 import nds_bench
    for c in G['get_perf_metric']:
        assert nds_bench.get_perf_metric(*c['args']) == c['out'], c
    for c in G['get_throughput_stream_nums']:
        assert nds_bench.get_throughput_stream_nums(c['n'], c['which']) == c['out']
    rep = tmp_path / 'load.txt'
    rep.write_text('Load Test Time: 1234 seconds\nLoad Test Finished at: x\nRNGSEED used: 07291122510\n')
    assert nds_bench.get_load_time(str(rep)) == '1234' and nds_bench.get_load_end_timestamp(str(rep)) == '07291122510'
    log = tmp_path / 'time.csv'
    log.write_text('application_id,query,time/milliseconds\napp,query3,12\napp,Power Start Time,100\napp,Power End Time,111\napp,Power Test Time,11838\n')
    assert nds_bench.get_power_time(str(log)) == '11838'
    with pytest.raises(Exception):
        nds_bench.get_power_time(str(rep))
    out = tmp_path / 'metrics.csv'
    nds_bench.write_metrics_report(str(out), {'scale_factor': 100, 'perf_metric': 42})
    assert out.read_text() == 'scale_factor,100\nperf_metric,42\n'


@pytest.mark.parametrize('which', ['source_decimal', 'source_double', 'maintenance_decimal', 'maintenance_double'])
def test_schema_types_and_nullability_match_the_reference(which):
    """Every column's name, type and nullability for the 24 source and 12 maintenance tables, as the reference's
    get_schemas(use_decimal) / get_maintenance_schemas(use_decimal) build them (recorded by executing
    /root/reference/nds/nds_schema.py under a stub pyspark - independent of the text extraction behind _schema_data.py)."""
    from spark_rapids_benchmarks_b200 import nds_schema
    use_decimal = which.endswith('decimal')
    mine = nds_schema.get_schemas(use_decimal) if which.startswith('source') else nds_schema.get_maintenance_schemas(use_decimal)
    ref = dict((t, cols) for t, cols in G['schemas'][which])
    assert list(mine) == [t for t, _ in G['schemas'][which]]      # same tables, same order
    for table, cols in ref.items():
        sch = mine[table]
        assert sch.names == [c[0] for c in cols], table
        for (name, typ, nullable), field in zip(cols, sch):
            assert field.nullable == nullable, (table, name)
            assert field.type == nds_schema.arrow_type(typ, True), (table, name, typ, field.type)
    if which == 'source_decimal':
        spec = {t: [(n, ty, nu) for n, ty, nu in nds_schema.column_spec(t)] for t in ref}
        assert spec == {t: [tuple(c) for c in cols] for t, cols in ref.items()}
