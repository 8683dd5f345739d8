#!/usr/bin/env python3
"""Power run - drop-in for the reference's nds_power.py (/root/reference/nds/nds_power.py): same
positional arguments and flags, same CSV time log rows, same per-query JSON summaries, same
`<output_prefix>/<query_name>/` result layout, same exit-code rule.  The statements run on the
B200 engine instead of Spark; everything the engine cannot do raises inside the timed call and
is reported as `Failed` exactly like a Spark exception would be (reference PysparkBenchReport :94-102).
"""
import argparse
import csv
import os
import sys
import time
from collections import OrderedDict

if __package__ in (None, ''):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    __package__ = 'spark_rapids_benchmarks_b200'

from .PysparkBenchReport import PysparkBenchReport
from .check import check_json_summary_folder, check_query_subset_exists, check_version
from .nds_gen_query_stream import split_special_query
from .nds_schema import get_schemas
from .session import Session

check_version()


def gen_sql_from_stream(query_stream_file_path):
    """Split a query stream into an ordered {query_name: statement text} dict
    (reference nds_power.py:50-77).

    The stream is cut at every '-- start'; the name is the text between 'template ' and '.tpl';
    a block whose second ';'-separated piece contains 'select' is a two-statement template and is
    split into <name>_part1 / <name>_part2; every value gets its '-- start' prefix back."""
    with open(query_stream_file_path, 'r') as f:
        stream = f.read()
    queries = OrderedDict()
    for block in stream.split('-- start')[1:]:
        name = block[block.find('template') + 9: block.find('.tpl')]
        if 'select' in block.split(';')[1]:
            one, two = split_special_query(block)
            queries[name + '_part1'] = one
            queries[name + '_part2'] = two
        else:
            queries[name] = block
    for name in queries:
        queries[name] = '-- start' + queries[name]
    return queries


def setup_tables(spark_session, input_prefix, input_format, use_decimal, execution_time_list):
    """Register the 24 source tables and record one `CreateTempView <table>` row each
    (reference nds_power.py:79-106).  Registration parses the Parquet footers and stages the
    encoded column chunks in HBM; it is outside Power Test Time like the reference's (:246-254)."""
    spark_app_id = spark_session.application_id
    for table_name in get_schemas(False).keys():
        start = int(time.time() * 1000)
        table_path = input_prefix + '/' + table_name
        spark_session.register_table(table_name, table_path, input_format)
        end = int(time.time() * 1000)
        print("====== Creating TempView for table {} ======".format(table_name))
        print("Time taken: {} millis for table {}".format(end - start, table_name))
        execution_time_list.append((spark_app_id, "CreateTempView {}".format(table_name), end - start))
    return execution_time_list


def run_one_query(spark_session, query, query_name, output_path, output_format):
    """reference nds_power.py:125-135"""
    df = spark_session.sql(query)
    if not output_path:
        df.collect()
    else:
        ensure_valid_column_names(df).write.format(output_format).mode('overwrite').save(
            output_path + '/' + query_name)


def ensure_valid_column_names(df):
    """Make result column names writable (reference nds_power.py:137-174): a name must start with a
    letter or '_' and continue with letters, digits or '_' - every other character becomes '_';
    a name occurring more than once gets its occurrence index appended."""
    def ok_first(ch):
        return ch.isalpha() or ch == '_'

    def ok_rest(ch):
        return ch.isalpha() or ch.isdigit() or ch == '_'

    def fix(name):
        if len(name) > 0 and ok_first(name[0]) and all(ok_rest(c) for c in name[1:]):
            return name
        return (name[0] if ok_first(name[0]) else '_') + ''.join(c if ok_rest(c) else '_' for c in name[1:])

    fixed = [fix(c) for c in df.columns]
    final = [n + str(fixed[:i].count(n)) if fixed.count(n) > 1 else n for i, n in enumerate(fixed)]
    return df.toDF(*final)


def get_query_subset(query_dict, subset):
    """reference nds_power.py:176-181"""
    check_query_subset_exists(query_dict, subset)
    return dict((k, query_dict[k]) for k in subset)


def run_query_stream(input_prefix,
                     property_file,
                     query_dict,
                     time_log_output_path,
                     extra_time_log_output_path,
                     sub_queries,
                     input_format="parquet",
                     use_decimal=True,
                     output_path=None,
                     output_format="parquet",
                     json_summary_folder=None,
                     delta_unmanaged=False,
                     keep_sc=False,
                     hive_external=False,
                     allow_failure=False,
                     session=None):
    """Run the stream and write the time log (reference nds_power.py:184-322).

    CSV rows, in order: `CreateTempView <t>` per table, one row per query (ms),
    `Power Start Time` / `Power End Time` (epoch seconds), `Power Test Time` / `Total Time` (ms)."""
    queries_reports = []
    execution_time_list = []
    total_time_start = time.time()
    app_name = "NDS - " + list(query_dict.keys())[0] if len(query_dict) == 1 else "NDS - Power Run"
    if input_format != 'parquet':
        raise Exception("input format '%s' is not supported by the B200 engine: transcode to uncompressed "
                        "parquet (nds_transcode --output_format parquet --compression none)" % input_format)
    if hive_external or delta_unmanaged:
        raise Exception("--hive / --delta_unmanaged need a Spark metastore; not supported by the B200 engine")
    conf = load_properties(property_file) if property_file else {}
    own_session = session is None
    spark_session = session or Session(app_name, conf, rank=int(os.environ.get('RANK', 0)),
                                       world=int(os.environ.get('WORLD_SIZE', 1)))
    spark_app_id = spark_session.application_id
    execution_time_list = setup_tables(spark_session, input_prefix, input_format, use_decimal, execution_time_list)

    check_json_summary_folder(json_summary_folder)
    if sub_queries:
        query_dict = get_query_subset(query_dict, sub_queries)
    power_start = int(time.time())
    for query_name, q_content in query_dict.items():
        print("====== Run {} ======".format(query_name))
        q_report = PysparkBenchReport(spark_session, query_name)
        summary = q_report.report_on(run_one_query, spark_session, q_content, query_name, output_path, output_format)
        print(f"Time taken: {summary['queryTimes']} millis for {query_name}")
        execution_time_list.append((spark_app_id, query_name, summary['queryTimes'][0]))
        queries_reports.append(q_report)
        if json_summary_folder:
            # property_file e.g.: "property/aqe-on.properties" -> prefix "<folder>/aqe-on"
            if property_file:
                summary_prefix = os.path.join(json_summary_folder, os.path.basename(property_file).split('.')[0])
            else:
                summary_prefix = os.path.join(json_summary_folder, '')
            q_report.write_summary(prefix=summary_prefix)
    power_end = int(time.time())
    power_elapse = int((power_end - power_start) * 1000)
    if not keep_sc and own_session:
        spark_session.stop()
    total_elapse = int((time.time() - total_time_start) * 1000)
    print("====== Power Test Time: {} milliseconds ======".format(power_elapse))
    print("====== Total Time: {} milliseconds ======".format(total_elapse))
    execution_time_list.append((spark_app_id, "Power Start Time", power_start))
    execution_time_list.append((spark_app_id, "Power End Time", power_end))
    execution_time_list.append((spark_app_id, "Power Test Time", power_elapse))
    execution_time_list.append((spark_app_id, "Total Time", total_elapse))

    header = ["application_id", "query", "time/milliseconds"]
    print(header)
    for row in execution_time_list:
        print(row)
    for path in (time_log_output_path, extra_time_log_output_path):
        if not path:
            continue
        with open(path, 'w', encoding='UTF8') as f:
            writer = csv.writer(f)
            writer.writerow(header)
            writer.writerows(execution_time_list)

    exit_code = 0
    for q in queries_reports:
        if not q.is_success():
            if exit_code == 0:
                print("====== Queries with failure ======")
            print("{} status: {}".format(q.summary['query'], q.summary['queryStatus']))
            exit_code = 1
    if exit_code:
        print("Above queries failed or completed with failed tasks. Please check the logs for the detailed reason.")
    if not allow_failure and exit_code:
        sys.exit(exit_code)
    return execution_time_list


def load_properties(filename):
    """k=v lines -> dict, split at the first '=' (reference nds_power.py:324-330; like the reference,
    lines without '=' become keys with an empty value)."""
    props = {}
    with open(filename) as fh:
        for line in fh:
            key, _, value = line.partition("=")
            props[key.strip()] = value.strip()
    return props


def build_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument('input_prefix', help='text to prepend to every input file path')
    parser.add_argument('query_stream_file', help='query stream file that contains NDS queries in specific order')
    parser.add_argument('time_log', help='path to execution time log, only support local path.', default="")
    parser.add_argument('--input_format', help='type for input data source; the B200 engine reads parquet only',
                        choices=['parquet', 'orc', 'avro', 'csv', 'json', 'iceberg', 'delta'], default='parquet')
    parser.add_argument('--output_prefix', help='text to prepend to every output file')
    parser.add_argument('--output_format', help='type of query output', default='parquet')
    parser.add_argument('--property_file', help='property file for engine configuration (k=v lines).')
    parser.add_argument('--floats', action='store_true',
                        help='kept for CLI compatibility; only affects text inputs, which the engine rejects')
    parser.add_argument('--json_summary_folder',
                        help='Empty folder/path (will create if not exist) to save JSON summary file for each query.')
    parser.add_argument('--delta_unmanaged', action='store_true', help='not supported by the B200 engine')
    parser.add_argument('--keep_sc', action='store_true', help='keep the engine session alive after the run')
    parser.add_argument('--hive', action='store_true', help='not supported by the B200 engine')
    parser.add_argument('--extra_time_log', help='extra (local) path to save the time log')
    parser.add_argument('--sub_queries', type=lambda s: [x.strip() for x in s.split(',')],
                        help='comma separated list of queries to run, e.g. "query1,query2"; use _part1/_part2 suffixes '
                             'for query14, query23, query24, query39')
    parser.add_argument('--allow_failure', action='store_true',
                        help='Do not exit with non zero when any query failed')
    return parser


if __name__ == "__main__":
    args = build_parser().parse_args()
    query_dict = gen_sql_from_stream(args.query_stream_file)
    run_query_stream(args.input_prefix,
                     args.property_file,
                     query_dict,
                     args.time_log,
                     args.extra_time_log,
                     args.sub_queries,
                     args.input_format,
                     not args.floats,
                     args.output_prefix,
                     args.output_format,
                     args.json_summary_folder,
                     args.delta_unmanaged,
                     args.keep_sc,
                     args.hive,
                     args.allow_failure)
