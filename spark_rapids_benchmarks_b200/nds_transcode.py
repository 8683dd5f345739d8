"""Load test: raw dsdgen text -> the Parquet warehouse the power run reads.

Same command line, table layout and report file as the reference's Spark job
(/root/reference/nds/nds_transcode.py: `load` :56-68, `store` :69-152, `transcode` :154-226, CLI :229-315),
done with pyarrow on the host - there is no Spark on these boxes and the transcode is not on the
timed query path.

  * input: `<input_prefix>/<table>` holding `|`-delimited, header-less, ISO-8859-1 text (a directory of
    `.dat` files as nds_gen_data.py leaves them, or one file); a trailing `|` per line is tolerated;
    an empty field is NULL (Spark's csv reader default);
  * output: `<output_prefix>/<table>/`; the seven fact tables are hive-partitioned by their date key
    (`<col>=<value>/`, NULL -> `__HIVE_DEFAULT_PARTITION__`, the column itself is not stored) with rows
    kept in input order inside a partition; every other table becomes a single file (`coalesce(1)`);
  * Parquet written the way the engine reads it: data page v1, dictionary encoding, decimals with
    precision <= 18 as INT32/INT64 (what Spark's writer does too), and NO compression unless
    `--compression` says otherwise (the engine rejects compressed pages - DESIGN.md section 8);
  * report file: the lines nds_bench.py parses (`Load Test Time: <s> seconds`, `RNGSEED used: ...`) and
    one `Time to convert '<table>' was ...s` line per table.

Only csv -> parquet is implemented; any other format combination, `--hive`, iceberg and delta raise.
CHAR(n) values are stored as read: Spark >= 3.1 pads them with spaces to n on write [EXT]; pass
`--pad_char` to do the same.
"""
import argparse
import os
import shutil
import sys
import timeit
from datetime import datetime

import pyarrow as pa
import pyarrow.compute as pc
import pyarrow.csv as pacsv
import pyarrow.parquet as pq

if __package__ in (None, ''):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from spark_rapids_benchmarks_b200 import nds_schema
else:
    from . import nds_schema

HIVE_NULL = '__HIVE_DEFAULT_PARTITION__'
WRITE_KW = dict(use_dictionary=True, data_page_version='1.0', store_decimal_as_integer=True, write_statistics=True)


def input_files(path):
    if os.path.isdir(path):
        fs = sorted(os.path.join(path, f) for f in os.listdir(path) if not f.startswith(('.', '_')))
        if not fs:
            raise ValueError('no input files under %s' % path)
        return fs
    if os.path.exists(path):
        return [path]
    raise ValueError('input path %s does not exist' % path)


def load(path, schema, delimiter='|', char_lengths=None, pad_char=False):
    """One raw file -> pyarrow Table with the table's schema (reference `load`, csv branch)."""
    names = list(schema.names)
    with open(path, 'rb') as f:
        first = f.readline().rstrip(b'\r\n')
    nfields = first.count(delimiter.encode()) + 1 if first else len(names)
    if nfields not in (len(names), len(names) + 1):
        raise ValueError('%s: %d fields per line, the schema has %d columns' % (path, nfields, len(names)))
    cols = names + (['__trailing__'] if nfields == len(names) + 1 else [])
    types = {f.name: f.type for f in schema}
    t = pacsv.read_csv(
        path,
        read_options=pacsv.ReadOptions(column_names=cols, encoding='latin1', block_size=64 << 20),
        parse_options=pacsv.ParseOptions(delimiter=delimiter, quote_char=False, escape_char=False),
        convert_options=pacsv.ConvertOptions(column_types=types, null_values=[''], strings_can_be_null=True,
                                             include_columns=names))
    if pad_char and char_lengths:
        for name, n in char_lengths.items():
            i = t.schema.get_field_index(name)
            t = t.set_column(i, name, pc.utf8_rpad(t.column(name), n, ' '))
    return t.cast(schema)


def _prepare(path, output_mode):
    """Spark save modes; returns False when the table is to be skipped."""
    if os.path.exists(path):
        if output_mode in ('error', 'errorifexists'):
            raise FileExistsError('path %s already exists' % path)
        if output_mode == 'ignore':
            return False
        if output_mode == 'overwrite':
            shutil.rmtree(path)
    os.makedirs(path, exist_ok=True)
    return True


def store(tables, name, output_mode, compression, prefix, part_no=0):
    """Write the pieces of one table (reference `store`, plain-format branch). `tables` yields pyarrow Tables."""
    out = os.path.join(prefix, name)
    if not _prepare(out, output_mode):
        return 0
    kw = dict(WRITE_KW, compression=(compression or 'none').upper() if (compression or 'none').lower() != 'none' else 'NONE')
    pcol = nds_schema.TABLE_PARTITIONING.get(name)
    rows = 0
    if pcol is None:
        t = pa.concat_tables(list(tables))
        pq.write_table(t, os.path.join(out, 'part-%05d-b200transcode.parquet' % part_no), row_group_size=1 << 20, **kw)
        return t.num_rows
    for k, t in enumerate(tables):
        rows += t.num_rows
        key = t.column(pcol).combine_chunks()
        order = pc.sort_indices(pa.table({'k': key}), sort_keys=[('k', 'ascending')], null_placement='at_end')   # stable
        t = t.take(order).drop_columns([pcol])
        skey = key.take(order)
        runs = pc.run_end_encode(skey) if len(skey) else None
        start = 0
        if runs is not None:
            for v, end in zip(runs.values.to_pylist(), runs.run_ends.to_pylist()):
                d = os.path.join(out, '%s=%s' % (pcol, HIVE_NULL if v is None else v))
                os.makedirs(d, exist_ok=True)
                pq.write_table(t.slice(start, end - start), os.path.join(d, 'part-%05d-b200transcode.parquet' % (part_no + k)),
                               row_group_size=1 << 20, **kw)
                start = end
    return rows


def transcode(args):
    if args.input_format != 'csv' or args.output_format != 'parquet':
        raise ValueError('this transcoder converts csv to parquet only (asked: %s -> %s)' % (args.input_format, args.output_format))
    if args.hive or args.delta_unmanaged:
        raise ValueError('--hive / --delta_unmanaged need a metastore; not supported')
    schemas = nds_schema.get_maintenance_schemas(not args.floats) if args.update else nds_schema.get_schemas(not args.floats)
    spec = nds_schema.MAINTENANCE_TABLES if args.update else nds_schema.SOURCE_TABLES
    if args.tables:
        for t in args.tables:
            if t not in schemas:
                raise Exception('invalid table name: %s. Valid tables are: %s' % (t, list(schemas.keys())))
        schemas = {t: schemas[t] for t in args.tables}
    results = {}
    start_time = datetime.now()
    print('Load Test Start Time: %s' % start_time)
    for name, schema in schemas.items():
        chars = {c: int(typ[5:-1]) for c, typ, _ in spec[name] if typ.startswith('char(')}

        def one(name=name, schema=schema, chars=chars):
            files = input_files(os.path.join(args.input_prefix, name))
            store((load(f, schema, char_lengths=chars, pad_char=args.pad_char) for f in files), name, args.output_mode,
                  args.compression, args.output_prefix)
        results[name] = timeit.timeit(one, number=1)
    end_time = datetime.now()
    delta = (end_time - start_time).total_seconds()
    print('Load Test Finished at: %s' % end_time)
    print('Load Test Time: %s seconds' % delta)
    end_time_formatted = end_time.strftime('%m%d%H%M%S%f')[:-5]      # TPC-DS spec 4.3.1, as the reference formats it
    print('RNGSEED used :%s' % end_time_formatted)
    text = 'Load Test Time: %s seconds\n' % delta
    text += 'Load Test Finished at: %s\n' % end_time
    text += 'RNGSEED used: %s\n' % end_time_formatted
    for table, duration in results.items():
        text += "Time to convert '%s' was %.04fs\n" % (table, duration)
    text += '\n\n\nSpark configuration follows:\n\n'
    with open(args.report_file, 'w') as report:
        report.write(text)
        print(text)
        for conf in (('b200.transcoder', 'pyarrow ' + pa.__version__), ('b200.compression', args.compression or 'none')):
            report.write(str(conf) + '\n')
            print(conf)


def build_parser():
    p = argparse.ArgumentParser()
    p.add_argument('input_prefix', help='folder holding one sub-folder (or file) of raw text per table')
    p.add_argument('output_prefix', help='folder the converted tables are written under')
    p.add_argument('report_file', help='location to store a performance report(local)')
    p.add_argument('--output_mode', choices=['overwrite', 'append', 'ignore', 'error', 'errorifexists'], default='errorifexists')
    p.add_argument('--input_format', choices=['csv', 'parquet', 'orc', 'avro', 'json'], default='csv')
    p.add_argument('--output_format', choices=['parquet', 'orc', 'avro', 'json', 'iceberg', 'delta'], default='parquet')
    p.add_argument('--tables', type=lambda s: s.split(','), help="comma separated table names, e.g. 'catalog_page,catalog_sales'")
    p.add_argument('--log_level', default='INFO', help='accepted for compatibility; there is no Spark driver log')
    p.add_argument('--floats', action='store_true', help='replace DecimalType with DoubleType when saving parquet files')
    p.add_argument('--update', action='store_true', help='transcode the update (maintenance) data instead of the source data')
    p.add_argument('--iceberg_write_format', choices=['parquet', 'orc', 'avro'], default='parquet')
    p.add_argument('--compression', help='parquet codec; default none, which is what the B200 engine reads')
    p.add_argument('--delta_unmanaged', action='store_true')
    p.add_argument('--hive', action='store_true')
    p.add_argument('--database', default='default')
    p.add_argument('--pad_char', action='store_true', help='pad CHAR(n) values with spaces to n, as Spark >= 3.1 does on write')
    return p


if __name__ == '__main__':
    transcode(build_parser().parse_args())
