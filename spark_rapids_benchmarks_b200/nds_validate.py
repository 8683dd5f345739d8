#!/usr/bin/env python3
"""Result validation - drop-in for the reference's nds_validate.py
(/root/reference/nds/nds_validate.py): same arguments, same comparison rules, same
`queryValidationStatus` stamping of the JSON summaries.  The two runs' outputs are read with
pyarrow instead of a SparkSession; row values become the same Python types pyspark's collect()
yields (int, float, str, Decimal, date, None), so `compare` follows reference :194-215 case by case.
"""
import argparse
import functools
import glob
import json
import math
import os
import re
import sys
import time
from decimal import Decimal

import pyarrow as pa
import pyarrow.csv as pacsv
import pyarrow.parquet as pq

if __package__ in (None, ''):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    __package__ = 'spark_rapids_benchmarks_b200'

from .nds_power import gen_sql_from_stream, get_query_subset

query_dict = {}   # module-global like the reference's (update_summary reads it, reference :279)


def load_output(path, fmt):
    """One query's output folder -> pyarrow.Table."""
    if fmt == 'parquet':
        files = sorted(f for f in glob.glob(os.path.join(path, '*.parquet')))
        if not files:
            raise FileNotFoundError('no parquet output under %s' % path)
        return pa.concat_tables([pq.ParquetFile(f).read() for f in files])
    if fmt == 'csv':
        files = sorted(f for f in glob.glob(os.path.join(path, '*.csv')))
        return pa.concat_tables([pacsv.read_csv(f) for f in files])
    raise Exception("validation input format '%s' is not supported (parquet, csv)" % fmt)


def compare_results(spark_session, input1, input2, input1_format, input2_format, ignore_ordering, is_q78,
                    q78_problematic_col, use_iterator=False, max_errors=10, epsilon=0.00001) -> bool:
    """Row-by-row, value-by-value comparison of two outputs (reference :48-114).
    spark_session is accepted for signature compatibility and unused."""
    df1 = load_output(input1, input1_format)
    df2 = load_output(input2, input2_format)
    count1, count2 = df1.num_rows, df2.num_rows
    if count1 != count2:
        print(f"DataFrame row counts do not match: {count1} != {count2}")
        return False
    result1 = collect_results(df1, ignore_ordering, use_iterator)
    result2 = collect_results(df2, ignore_ordering, use_iterator)
    errors = 0
    i = 0
    while i < count1 and errors < max_errors:
        lhs = next(result1)
        rhs = next(result2)
        if not rowEqual(list(lhs), list(rhs), epsilon, is_q78, q78_problematic_col):
            print(f"Row {i}: \n{list(lhs)}\n{list(rhs)}\n")
            errors += 1
        i += 1
    print(f"Processed {i} rows")
    if errors == max_errors:
        print(f"Aborting comparison after reaching maximum of {max_errors} errors")
        return False
    if errors == 0:
        print("Results match")
        return True
    print(f"There were {errors} errors")
    return False


def _null_first_cmp(a, b):
    if a is None and b is None:
        return 0
    if a is None:
        return -1
    if b is None:
        return 1
    if isinstance(a, float) and isinstance(b, float):
        # Spark orders NaN above every other double
        an, bn = math.isnan(a), math.isnan(b)
        if an or bn:
            return 0 if (an and bn) else (1 if an else -1)
    if isinstance(a, str):
        a, b = a.encode('utf-8'), b.encode('utf-8')
    return -1 if a < b else (1 if a > b else 0)


def collect_results(df, ignore_ordering, use_iterator):
    """Rows as tuples; with ignore_ordering they are first sorted ascending (NULLS FIRST) by the
    non-floating columns, then by the floating ones (reference :116-144)."""
    names = df.schema.names
    is_float = [pa.types.is_floating(f.type) for f in df.schema]
    print("Collecting rows from DataFrame")
    t1 = time.time()
    cols = [c.to_pylist() for c in df.columns]
    rows = list(zip(*cols)) if cols else []
    if ignore_ordering:
        order = [i for i in range(len(names)) if not is_float[i]] + [i for i in range(len(names)) if is_float[i]]

        def cmp(x, y):
            for i in order:
                c = _null_first_cmp(x[i], y[i])
                if c:
                    return c
            return 0
        rows.sort(key=functools.cmp_to_key(cmp))
    t2 = time.time()
    print(f"Collected {len(rows)} rows in {t2-t1} seconds")
    return iter(rows)


def check_nth_col_problematic_q78(q78_content: str) -> int:
    """1-based position of the select-list item containing 'ratio' in query78's last
    select ... from (reference :146-164)."""
    tail = q78_content.split("select")[-1].split("from")[0]
    parts = re.split(', |,\n', tail)
    nth = -1
    for index, item in enumerate(parts):
        if 'ratio' in item:
            nth = index
    if nth == -1:
        raise Exception(f"Cannot find the problematic column in the query78 content. Please check the content.\n{q78_content}")
    return nth + 1


def rowEqual(row1, row2, epsilon, is_q78, q78_problematic_col):
    """All values of a row must `compare` equal; query78's rounded ratio column (2nd or 4th) is taken
    out and allowed an absolute difference of 0.01001 (reference :166-192)."""
    if is_q78:
        if q78_problematic_col != 2 and q78_problematic_col != 4:
            raise Exception(f"q78 problematic column should be 2nd or 4th, but get {q78_problematic_col}")
        v1 = row1.pop(q78_problematic_col - 1)
        v2 = row2.pop(q78_problematic_col - 1)
        if all([v1, v2]):
            special_ok = abs(v1 - v2) <= 0.01001
        elif v1 == None and v2 == None:   # noqa: E711 (mirrors the reference's == None)
            special_ok = True
        else:
            special_ok = False
        return special_ok and all([compare(lhs, rhs, epsilon) for lhs, rhs in zip(row1, row2)])
    return all([compare(lhs, rhs, epsilon) for lhs, rhs in zip(row1, row2)])


def compare(expected, actual, epsilon=0.00001):
    """reference :194-215 - floats: both NaN or isclose(rel_tol=epsilon); str: ==; None only equals None;
    Decimal: isclose(rel_tol=epsilon); everything else: =="""
    if isinstance(expected, float) and isinstance(actual, float):
        if math.isnan(expected) and math.isnan(actual):
            return True
        return math.isclose(expected, actual, rel_tol=epsilon)
    if isinstance(expected, str) and isinstance(actual, str):
        return expected == actual
    if expected is None and actual is None:
        return True
    if expected is None or actual is None:
        return False
    if isinstance(expected, Decimal) and isinstance(actual, Decimal):
        return math.isclose(expected, actual, rel_tol=epsilon)
    return expected == actual


def iterate_queries(spark_session, input1, input2, input1_format, input2_format, ignore_ordering, query_dict,
                    use_iterator=False, max_errors=10, epsilon=0.00001, is_float=False):
    """Compare every query folder; query65 is always skipped and query67 when --floats
    (reference :217-260)."""
    unmatch_queries = []
    for query_name in query_dict.keys():
        if query_name == 'query65':
            continue
        if query_name == 'query67' and is_float:
            continue
        sub_input1 = input1 + '/' + query_name
        sub_input2 = input2 + '/' + query_name
        print(f"=== Comparing Query: {query_name} ===")
        problematic_col = 2
        if query_name == 'query78':
            problematic_col = check_nth_col_problematic_q78(query_dict[query_name])
        same = compare_results(spark_session, sub_input1, sub_input2, input1_format, input2_format, ignore_ordering,
                               query_name == 'query78', q78_problematic_col=problematic_col, use_iterator=use_iterator,
                               max_errors=max_errors, epsilon=epsilon)
        if same == False:   # noqa: E712
            unmatch_queries.append(query_name)
    if len(unmatch_queries) != 0:
        print(f"=== Unmatch Queries: {unmatch_queries} ===")
    return unmatch_queries


def update_summary(prefix, unmatch_queries, queries=None):
    """Stamp queryValidationStatus into each query's JSON summary: 'Pass' when it matched; otherwise
    'Fail' when the query had completed and 'NotAttempted' when it had failed (reference :262-296).
    Exactly one file `<prefix>/*<query>-*.json` must exist per query."""
    queries = query_dict if queries is None else queries
    if not os.path.exists(prefix):
        raise Exception("The json summary folder doesn't exist.")
    print(f"Updating queryValidationStatus in folder {prefix}.")
    for query_name in queries.keys():
        found = glob.glob(prefix + f'/*{query_name}-*.json')
        if len(found) > 1:
            raise Exception(f"More than one summary file found for query {query_name} in folder {prefix}.")
        if len(found) == 0:
            raise Exception(f"No summary file found for query {query_name} in folder {prefix}.")
        filename = found[0]
        with open(filename, 'r') as f:
            summary = json.load(f)
        if query_name in unmatch_queries:
            done = 'Completed' in summary['queryStatus'] or 'CompletedWithTaskFailures' in summary['queryStatus']
            summary['queryValidationStatus'] = ['Fail'] if done else ['NotAttempted']
        else:
            summary['queryValidationStatus'] = ['Pass']
        with open(filename, 'w') as f:
            json.dump(summary, f, indent=2)


def build_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument('input1', help='path of the first input data.')
    parser.add_argument('input2', help='path of the second input data.')
    parser.add_argument('query_stream_file', help='query stream file that contains NDS queries in specific order.')
    parser.add_argument('--input1_format', default='parquet', help='data source type for the first input data.')
    parser.add_argument('--input2_format', default='parquet', help='data source type for the second input data.')
    parser.add_argument('--max_errors', help='Maximum number of differences to report.', type=int, default=10)
    parser.add_argument('--epsilon', type=float, default=0.00001,
                        help='Allow for differences in precision when comparing floating point values.')
    parser.add_argument('--ignore_ordering', action='store_true',
                        help='Sort the data collected from the DataFrames before comparing them.')
    parser.add_argument('--use_iterator', action='store_true', help='kept for CLI compatibility (results are small)')
    parser.add_argument('--floats', action='store_true',
                        help='the inputs hold float data instead of decimal data (query67 is then skipped)')
    parser.add_argument('--json_summary_folder', help='path of a folder that contains json summary file for each query.')
    parser.add_argument('--sub_queries', type=lambda s: [x.strip() for x in s.split(',')],
                        help='comma separated list of queries to compare')
    return parser


def main(argv=None):
    global query_dict
    args = build_parser().parse_args(argv)
    query_dict = gen_sql_from_stream(args.query_stream_file)
    if args.sub_queries:
        query_dict = get_query_subset(query_dict, args.sub_queries)
    unmatch_queries = iterate_queries(None, args.input1, args.input2, args.input1_format, args.input2_format,
                                      args.ignore_ordering, query_dict, use_iterator=args.use_iterator,
                                      max_errors=args.max_errors, epsilon=args.epsilon, is_float=args.floats)
    if args.json_summary_folder:
        update_summary(args.json_summary_folder, unmatch_queries)
    return unmatch_queries


if __name__ == "__main__":
    main()
