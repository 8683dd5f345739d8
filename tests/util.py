import os

import pyarrow as pa
import pyarrow.parquet as pq

from spark_rapids_benchmarks_b200 import nds_schema

HIVE_NULL = '__HIVE_DEFAULT_PARTITION__'


def read_table_like_engine(warehouse, table, columns=None):
    """pyarrow restatement of the engine's scan order: files in sorted directory order, the
    hive partition column appended as int32 (null for __HIVE_DEFAULT_PARTITION__)."""
    tdir = os.path.join(warehouse, table)
    pcol = nds_schema.TABLE_PARTITIONING.get(table)
    parts = []
    for e in sorted(os.listdir(tdir)):
        p = os.path.join(tdir, e)
        if e.startswith('_') or e.startswith('.'):
            continue
        if os.path.isdir(p):
            val = e.split('=', 1)[1]
            for f in sorted(os.listdir(p)):
                if not f.endswith('.parquet'):
                    continue
                t = pq.ParquetFile(os.path.join(p, f)).read()
                v = None if val == HIVE_NULL else int(val)
                t = t.append_column(pcol, pa.array([v] * t.num_rows, type=pa.int32()))
                parts.append(t)
        elif e.endswith('.parquet'):
            parts.append(pq.ParquetFile(p).read())
    t = pa.concat_tables(parts)
    if columns:
        t = t.select(columns)
    return t


def scan_plan(table, columns):
    return {'op': 'scan', 'table': table, 'columns': list(columns)}
