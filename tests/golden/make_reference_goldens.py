"""Generates tests/golden/reference_harness.json by IMPORTING THE REFERENCE'S OWN PYTHON
(/root/reference/nds/*.py) in the build container and recording what its pure-Python functions
return on fixed inputs.  pyspark is not installed, so the modules are imported under a stub
`pyspark` package; only functions that never touch Spark are exercised:

    nds_power.gen_sql_from_stream / ensure_valid_column_names / load_properties / get_query_subset
    nds_gen_query_stream.split_special_query
    nds_validate.compare / rowEqual / check_nth_col_problematic_q78 / update_summary
    PysparkBenchReport.report_on / write_summary / is_success   (with a fake session object)
    nds_bench.get_power_time / get_start_end_time               (parsing a time log written by THIS repo)

Run:  python tests/golden/make_reference_goldens.py      (needs /root/reference; not run on the GPU box)
"""
import decimal
import glob
import io
import json
import math
import os
import sys
import tempfile
import types
import contextlib

REF = '/root/reference/nds'
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def install_stub_pyspark():
    def mod(name):
        m = types.ModuleType(name)
        sys.modules[name] = m
        return m
    ps = mod('pyspark')
    sql = mod('pyspark.sql')
    typ = mod('pyspark.sql.types')
    fun = mod('pyspark.sql.functions')
    jg = mod('pyspark.java_gateway')
    ps.sql = sql
    sql.types = typ
    sql.functions = fun

    class _T:
        def __init__(self, *a, **k):
            self.a = a
            self.k = k

        @classmethod
        def typeName(cls):
            return cls.__name__.replace('Type', '').lower()
    for n in ['StructType', 'StructField', 'IntegerType', 'LongType', 'DoubleType', 'FloatType', 'StringType', 'DateType',
              'DecimalType', 'VarcharType', 'CharType', 'TimestampType', 'ShortType', 'ByteType', 'BooleanType']:
        setattr(typ, n, type(n, (_T,), {}))
    typ.__all__ = [n for n in dir(typ) if n.endswith('Type') or n == 'StructField']
    sql.SparkSession = type('SparkSession', (), {})
    sql.DataFrame = type('DataFrame', (), {})
    ps.SparkContext = type('SparkContext', (), {})
    fun.col = lambda n: n
    jg.ensure_callback_server_started = lambda gw=None: None


class FakeDF:
    def __init__(self, columns):
        self.columns = list(columns)

    def toDF(self, *names):
        return FakeDF(names)


STREAM = """-- start query 1 in stream 0 using template query3.tpl
select a, b from t where x = 1
 LIMIT 100;

-- end query 1 in stream 0 using template query3.tpl
-- start query 2 in stream 0 using template query14.tpl
with cs as (select 1) select * from cs limit 100;
 select k, sum(v) from u group by k
 limit 100;

-- end query 2 in stream 0 using template query14.tpl
-- start query 3 in stream 0 using template query98.tpl
select 'x;y' from w order by 1;

-- end query 3 in stream 0 using template query98.tpl
"""

PROPS = """#
# comment line = with equals
spark.sql.adaptive.enabled=true
spark.executor.memory = 16g
b200.source=host
novalue
k=a=b
"""

Q78 = """-- start query 78 in stream 0 using template query78.tpl
with ws as (select 1 from web_sales), ss as (select 2 from store_sales)
select ss_sold_year, ss_item_sk, ss_customer_sk,
round(ss_qty/(coalesce(ws_qty,0)+coalesce(cs_qty,0)),2) ratio,
ss_qty store_qty
from ss limit 100;
"""


def main():
    install_stub_pyspark()
    sys.path.insert(0, REF)
    import nds_power as ref_power
    import nds_validate as ref_validate
    import nds_gen_query_stream as ref_stream
    import PysparkBenchReport as ref_report
    import nds_bench as ref_bench
    out = {'_generated_by': 'tests/golden/make_reference_goldens.py', '_reference': 'bilalbari/spark-rapids-benchmarks @ /root/reference'}

    with tempfile.TemporaryDirectory() as tmp:
        # ---- gen_sql_from_stream / split_special_query
        sp = os.path.join(tmp, 'query_0.sql')
        open(sp, 'w').write(STREAM)
        d = ref_power.gen_sql_from_stream(sp)
        out['stream_text'] = STREAM
        out['gen_sql_from_stream'] = list(d.items())
        block = STREAM.split('-- start')[2]
        out['split_special_query'] = {'input': block, 'output': list(ref_stream.split_special_query(block))}
        # ---- get_query_subset
        out['get_query_subset'] = {'subset': ['query98', 'query14_part2'],
                                   'keys': list(ref_power.get_query_subset(d, ['query98', 'query14_part2']).keys())}
        try:
            ref_power.get_query_subset(d, ['query1'])
            out['get_query_subset']['missing_error'] = None
        except Exception as e:
            out['get_query_subset']['missing_error'] = str(e)
        # ---- ensure_valid_column_names
        cases = [['d_year', 'brand_id', 'sum(ss_ext_sales_price)'], ['a', 'a', 'b', 'a'], ['1abc', 'x-y', '_ok', 'c d e'],
                 ['count(1)', 'count(1)', 'avg(x)'], ['ünï', 'col#1', 'col#1'], ['a', 'a0', 'a']]
        out['ensure_valid_column_names'] = [{'in': c, 'out': ref_power.ensure_valid_column_names(FakeDF(c)).columns} for c in cases]
        # ---- load_properties
        pp = os.path.join(tmp, 'x.properties')
        open(pp, 'w').write(PROPS)
        out['load_properties'] = {'text': PROPS, 'out': ref_power.load_properties(pp)}
        # ---- compare
        D = decimal.Decimal
        vals = [1, 2, 1.0, 1.000001, 1.1, float('nan'), float('inf'), 0.0, -0.0, 1e-12, 'a', 'b', '', None, D('1.00'), D('1.000009'),
                D('1.01'), D('0'), D('100000.00'), D('100000.99'), True, False]

        def enc(v):
            if isinstance(v, D):
                return {'dec': str(v)}
            if isinstance(v, float):
                return {'f': repr(v)}
            return v
        cmp_rows = []
        for a in vals:
            for b in vals:
                for eps in (0.00001, 0.01):
                    try:
                        r = bool(ref_validate.compare(a, b, eps))
                    except Exception as e:   # the reference raises on Decimal vs float isclose? record it
                        r = 'EXC:' + type(e).__name__
                    cmp_rows.append([enc(a), enc(b), eps, r])
        out['compare'] = cmp_rows
        # ---- rowEqual incl. q78
        rows = [([1, 'a', 2.0], [1, 'a', 2.00001], 0.00001, False, 2),
                ([1, 'a', 2.0], [1, 'b', 2.0], 0.00001, False, 2),
                ([1, 0.50, 3, 4.0], [1, 0.505, 3, 4.0], 0.00001, True, 2),
                ([1, 0.50, 3, 4.0], [1, 0.52, 3, 4.0], 0.00001, True, 2),
                ([1, None, 3, 4.0], [1, None, 3, 4.0], 0.00001, True, 2),
                ([1, None, 3, 4.0], [1, 0.1, 3, 4.0], 0.00001, True, 2),
                ([1, 2, 3, 0.25], [1, 2, 3, 0.2599], 0.00001, True, 4),
                ([1, 0.0, 3, 4.0], [1, 0.0, 3, 4.0], 0.00001, True, 2)]
        re_out = []
        for r1, r2, eps, q78, col in rows:
            re_out.append({'row1': r1, 'row2': r2, 'epsilon': eps, 'is_q78': q78, 'col': col,
                           'out': bool(ref_validate.rowEqual(list(r1), list(r2), eps, q78, col))})
        try:
            ref_validate.rowEqual([1, 2, 3], [1, 2, 3], 0.00001, True, 3)
            bad = None
        except Exception as e:
            bad = str(e)
        out['rowEqual'] = {'cases': re_out, 'bad_col_error': bad}
        out['check_nth_col_problematic_q78'] = {'text': Q78, 'out': ref_validate.check_nth_col_problematic_q78(Q78)}
        # ---- PysparkBenchReport with a fake session
        class FakeConf:
            def getAll(self):
                return [('spark.app.name', 'x'), ('k', 'v')]

        class FakeSC:
            _conf = FakeConf()

        class FakeSession:
            sparkContext = FakeSC()
            version = '3.1.2-fake'
        os.environ['TOKEN'] = 'secret-token'
        os.environ['MY_TOKEN'] = 'kept'

        # python_listener.PythonListener().register() needs a JVM: make it raise TypeError like a missing jar does
        import python_listener
        python_listener.PythonListener.register = lambda self: (_ for _ in ()).throw(TypeError("'JavaPackage' object is not callable"))
        rep = ref_report.PysparkBenchReport(FakeSession(), 'query3')
        with contextlib.redirect_stdout(io.StringIO()):
            s = rep.report_on(lambda a, b: None, 1, 2)
        ok_keys = sorted(s.keys())
        prefix = os.path.join(tmp, 'aqe-on')
        rep.write_summary(prefix=prefix)
        written = json.load(open(glob.glob(prefix + '-query3-*.json')[0]))
        rep2 = ref_report.PysparkBenchReport(FakeSession(), 'query7')
        with contextlib.redirect_stdout(io.StringIO()):
            s2 = rep2.report_on(lambda: (_ for _ in ()).throw(ValueError('boom')))
        out['bench_report'] = {
            'summary_keys': ok_keys, 'env_keys': sorted(s['env'].keys()), 'status_ok': s['queryStatus'], 'is_success_ok': rep.is_success(),
            'token_redacted': 'TOKEN' not in s['env']['envVars'], 'my_token_kept': 'MY_TOKEN' in s['env']['envVars'],
            'written_keys': sorted(written.keys()), 'filename_pattern_ok': os.path.basename(written['filename']) ==
            'aqe-on-query3-%d.json' % written['startTime'],
            'status_fail': s2['queryStatus'], 'exceptions_fail': s2['exceptions'], 'is_success_fail': rep2.is_success(),
            'times_len': len(s['queryTimes']), 'spark_conf': s['env']['sparkConf'], 'spark_version': s['env']['sparkVersion']}
        # ---- update_summary
        jf = os.path.join(tmp, 'json')
        os.makedirs(jf)
        for q, st in (('query3', ['Completed']), ('query7', ['Failed']), ('query19', ['CompletedWithTaskFailures']), ('query42', ['Completed'])):
            json.dump({'query': q, 'queryStatus': st}, open(os.path.join(jf, 'p-%s-123.json' % q), 'w'))
        ref_validate.query_dict = {'query3': '', 'query7': '', 'query19': '', 'query42': ''}
        with contextlib.redirect_stdout(io.StringIO()):
            ref_validate.update_summary(jf, ['query7', 'query19', 'query42'])
        out['update_summary'] = {q: json.load(open(os.path.join(jf, 'p-%s-123.json' % q)))['queryValidationStatus']
                                 for q in ('query3', 'query7', 'query19', 'query42')}
        # ---- the reference's own time-log parsers on a log written by THIS repo's nds_power
        sys.path.insert(0, ROOT)
        import pyarrow as pa
        from spark_rapids_benchmarks_b200 import nds_power as my_power
        from spark_rapids_benchmarks_b200.session import DataFrame

        class StubSession:
            application_id = 'b200-test'
            version = 'stub'

            def conf_items(self):
                return {}.items()

            def register_table(self, *a):
                pass

            def sql(self, q):
                return DataFrame(pa.table({'x': [1]}))

            def stop(self):
                pass
        log = os.path.join(tmp, 'time.csv')
        mine = my_power.gen_sql_from_stream(sp)
        with contextlib.redirect_stdout(io.StringIO()):
            my_power.run_query_stream(tmp, None, mine, log, None, None, session=StubSession())
        text = open(log).read()
        pt = ref_bench.get_power_time(log)
        st, en = ref_bench.get_start_end_time(log)
        assert pt.isdigit() and st.isdigit() and en.isdigit() and int(en) >= int(st)
        lines = text.strip().split('\n')
        out['time_log'] = {'header': lines[0], 'labels': [l.split(',')[1] for l in lines[1:]],
                           'reference_parsers_accept_it': True,
                           'n_rows': len(lines) - 1}
        # ---- throughput test arithmetic (nds_bench.py:127-157, 207-208)
        out['get_stream_range'] = [{'n': n, 'which': w, 'out': ref_bench.get_stream_range(n, w)}
                                   for n in (3, 5, 9, 10) for w in (1, 2)]
        out['round_up_to_nearest_10_percent'] = [{'in': v, 'out': ref_bench.round_up_to_nearest_10_percent(v)}
                                                 for v in (0.0, 0.01, 1.0, 1.04, 1.05, 12.345, 99.91, 100.0)]
        base = os.path.join(tmp, 'thr')
        times = {1: (1000.25, 1010.5), 2: (1000.75, 1013.26), 3: (999.5, 1001.0), 4: (1002.0, 1012.0)}
        for k, (a, b) in times.items():
            with open(base + '_%d.csv' % k, 'w') as f:
                f.write('application_id,query,time/milliseconds\n')
                f.write('app,query1,12\napp,Power Start Time,%r\napp,Power End Time,%r\napp,Power Test Time,10\n' % (a, b))
        out['get_throughput_time'] = {'times': {str(k): list(v) for k, v in times.items()},
                                      'cases': [{'n': n, 'which': w, 'out': ref_bench.get_throughput_time(base, n, w)}
                                                for n, w in ((5, 1), (5, 2), (3, 1))]}
        # ---- the reference's load-report parsers on a report written by THIS repo's nds_transcode
        from spark_rapids_benchmarks_b200 import nds_transcode as my_transcode
        rawdir = os.path.join(tmp, 'raw', 'store')
        os.makedirs(rawdir)
        ncol = len(__import__('spark_rapids_benchmarks_b200.nds_schema', fromlist=['x']).get_schemas(True)['store'].names)
        with open(os.path.join(rawdir, 'store_1_1.dat'), 'w') as f:
            f.write('|'.join(['1', 'AAAAAAAABAAAAAAA'] + [''] * (ncol - 2)) + '|\n')
        rep = os.path.join(tmp, 'load_report.txt')
        targs = my_transcode.build_parser().parse_args([os.path.join(tmp, 'raw'), os.path.join(tmp, 'wh'), rep, '--tables', 'store'])
        with contextlib.redirect_stdout(io.StringIO()):
            my_transcode.transcode(targs)
        lt, seed = ref_bench.get_load_time(rep), ref_bench.get_load_end_timestamp(rep)
        assert float(lt) >= 0 and seed.isdigit() and len(seed) == 11
        out['load_report'] = {'reference_parsers_accept_it': True, 'rngseed_digits': len(seed),
                              'first_line_prefix': open(rep).readline().split(':')[0]}
        # ---- composite metric (nds_bench.py:334-357)
        pm_cases = [('100', 4, 1200.0, 3600.0, 5000.0, 5200.0, 300.0, 320.0), ('1000', 2, 86400.0, 7200.5, 9000.0, 9100.25, 1000.0, 900.0),
                    ('10', 1, 50.0, 12.5, 30.0, 31.0, 4.0, 5.0), ('3000', 8, 100000.0, 50000.0, 120000.0, 118000.0, 20000.0, 21000.0)]
        out['get_perf_metric'] = [{'args': list(c), 'out': ref_bench.get_perf_metric(*c)} for c in pm_cases]
        out['get_throughput_stream_nums'] = [{'n': n, 'which': w, 'out': ref_bench.get_throughput_stream_nums(n, w)}
                                             for n in (3, 9) for w in (1, 2)]
        # ---- schema catalogue: types and nullability as the reference's get_schemas builds them (nds_schema.py:49-716)
        import nds_schema as ref_schema

        def type_str(t):
            n = type(t).__name__
            if n == 'IntegerType':
                return 'int32'
            if n == 'LongType':
                return 'int64'
            if n == 'DoubleType':
                return 'double'
            if n == 'DateType':
                return 'date'
            if n == 'StringType':
                return 'string'
            if n == 'TimestampType':
                return 'timestamp'
            if n == 'DecimalType':
                return 'decimal(%d,%d)' % (t.a[0], t.a[1])
            if n in ('CharType', 'VarcharType'):
                return '%s(%d)' % (n[:-4].lower(), t.a[0])
            raise SystemExit('unmapped reference type ' + n)

        def dump(schemas):
            # a list of [table, columns] pairs: the JSON file is written with sorted keys and the table order matters
            return [[tbl, [[f.a[0], type_str(f.a[1]), bool(f.k.get('nullable', f.a[2] if len(f.a) > 2 else True))] for f in st.a[0]]]
                    for tbl, st in schemas.items()]
        out['schemas'] = {'source_decimal': dump(ref_schema.get_schemas(True)), 'source_double': dump(ref_schema.get_schemas(False)),
                          'maintenance_decimal': dump(ref_schema.get_maintenance_schemas(True)),
                          'maintenance_double': dump(ref_schema.get_maintenance_schemas(False))}
    with open(os.path.join(HERE, 'reference_harness.json'), 'w') as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print('wrote', os.path.join(HERE, 'reference_harness.json'), len(out['compare']), 'compare cases')


if __name__ == '__main__':
    main()
