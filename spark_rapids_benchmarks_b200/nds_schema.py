"""Schema catalogue: mirror of the reference's `nds_schema.get_schemas(use_decimal)`
(/root/reference/nds/nds_schema.py:43-568) and `get_maintenance_schemas` (:570-716),
expressed as pyarrow schemas instead of pyspark StructTypes.

Type mapping (the engine's device layout in brackets):
  IntegerType -> int32 [i32]      LongType -> int64 [i64]      DateType -> date32 [i32 days]
  DecimalType(p,s) -> decimal128(p,s), stored in Parquet as INT32 (p<=9) / INT64 (p<=18)
                      [unscaled i32/i64]; with use_decimal=False -> float64 (reference :43-47)
  CharType(n)/VarcharType(n) -> string [string views]
"""
import re
import pyarrow as pa

from ._schema_data import SOURCE_TABLES, MAINTENANCE_TABLES

# /root/reference/nds/nds_transcode.py:45-53
TABLE_PARTITIONING = {
    'catalog_sales': 'cs_sold_date_sk',
    'catalog_returns': 'cr_returned_date_sk',
    'inventory': 'inv_date_sk',
    'store_sales': 'ss_sold_date_sk',
    'store_returns': 'sr_returned_date_sk',
    'web_sales': 'ws_sold_date_sk',
    'web_returns': 'wr_returned_date_sk',
}

_DEC = re.compile(r'decimal\((\d+),(\d+)\)')
_CHR = re.compile(r'(char|varchar)\((\d+)\)')


def arrow_type(t, use_decimal=True):
    if t == 'int32':
        return pa.int32()
    if t == 'int64':
        return pa.int64()
    if t == 'date':
        return pa.date32()
    if t == 'double':
        return pa.float64()
    if t == 'timestamp':
        return pa.timestamp('us')
    if t == 'string' or _CHR.match(t):
        return pa.string()
    m = _DEC.match(t)
    if m:
        return pa.decimal128(int(m.group(1)), int(m.group(2))) if use_decimal else pa.float64()
    raise ValueError('unknown column type %r' % t)


def decimalType(use_decimal, precision, scale):
    """reference nds_schema.py:43-47"""
    return pa.decimal128(precision, scale) if use_decimal else pa.float64()


def _build(spec, use_decimal):
    out = {}
    for table, cols in spec.items():
        out[table] = pa.schema([pa.field(n, arrow_type(t, use_decimal), nullable=nullable)
                                for n, t, nullable in cols])
    return out


def get_schemas(use_decimal):
    """{table_name: pyarrow.Schema} for the 24 source tables, in the reference's order
    (reference nds_schema.py:49-568)."""
    return _build(SOURCE_TABLES, use_decimal)


def get_maintenance_schemas(use_decimal):
    """reference nds_schema.py:570-716"""
    return _build(MAINTENANCE_TABLES, use_decimal)


def column_spec(table):
    """[(name, type-string, nullable)] - the raw facts, used by the generator and planner."""
    return SOURCE_TABLES[table]


_CACHE = {}


def column_types():
    """{column_name: type-string} over all source tables (TPC-DS column names are unique)."""
    if 'types' in _CACHE:
        return _CACHE['types']
    out = _CACHE['types'] = {}
    for cols in SOURCE_TABLES.values():
        for n, t, _ in cols:
            out[n] = t
    return out


def column_table():
    if 'table' in _CACHE:
        return _CACHE['table']
    out = _CACHE['table'] = {}
    for tbl, cols in SOURCE_TABLES.items():
        for n, _, _ in cols:
            out[n] = tbl
    return out


if __name__ == "__main__":
    # reference :718-722 prints the schema dicts
    print(get_schemas(False))
    print(get_schemas(True))
    print(get_maintenance_schemas(False))
    print(get_maintenance_schemas(True))
