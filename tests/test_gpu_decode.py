"""GPU parity: Parquet page decode (RLE/bit-packed def levels, dictionary + PLAIN values,
partition-column injection) against pyarrow.parquet - bit exact, every column of every table."""
import pyarrow as pa
import pytest

from spark_rapids_benchmarks_b200 import nds_schema
from tests.util import read_table_like_engine, scan_plan

pytestmark = pytest.mark.gpu

TABLES = list(nds_schema.get_schemas(True))


def _normalize(col):
    col = col.combine_chunks() if isinstance(col, pa.ChunkedArray) else col
    return col


@pytest.mark.parametrize('table', TABLES)
@pytest.mark.parametrize('source', ['hbm', 'host'])
def test_decode_matches_pyarrow(gpu_engine, warehouse, table, source):
    expected = read_table_like_engine(warehouse, table)
    cols = expected.schema.names
    got = gpu_engine.execute(scan_plan(table, cols), source=source)
    assert got.num_rows == expected.num_rows
    for name in cols:
        e = _normalize(expected.column(name))
        g = got.column(got.schema.get_field_index(name))
        assert g.type == e.type, (name, g.type, e.type)
        assert g.null_count == e.null_count, name
        assert g.equals(e), 'column %s of %s differs' % (name, table)
    st = gpu_engine.last_stats()
    assert st['launches'] >= 1


@pytest.mark.parametrize('table', ['store_sales', 'item', 'date_dim', 'customer_demographics', 'customer'])
def test_general_kernel_alone_matches_pyarrow(gpu_engine, warehouse, table):
    """split_decode=false sends every page through decode_pages_kernel (the route RLE-heavy pages always take)."""
    expected = read_table_like_engine(warehouse, table)
    cols = expected.schema.names
    got = gpu_engine.execute(scan_plan(table, cols), options={'split_decode': False})
    for name in cols:
        assert got.column(got.schema.get_field_index(name)).equals(_normalize(expected.column(name))), name
    routes = [o['op'] for o in gpu_engine.last_stats()['ops'] if o['op'].startswith('decode_routes')]
    assert routes and all('def=0 rows=0' in r for r in routes)


def test_decode_column_subset_and_order(gpu_engine, warehouse):
    cols = ['ss_ext_sales_price', 'ss_sold_date_sk', 'ss_item_sk']
    expected = read_table_like_engine(warehouse, 'store_sales', cols)
    got = gpu_engine.execute(scan_plan('store_sales', cols))
    assert got.schema.names == cols
    assert pa.Table.from_batches([got]).equals(expected.combine_chunks().cast(got.schema))


def test_unsupported_format_fails_loudly(gpu_engine, warehouse):
    from spark_rapids_benchmarks_b200.engine import EngineError
    with pytest.raises(EngineError):
        gpu_engine.register_table('x', warehouse + '/item', 'orc')
    with pytest.raises(EngineError):
        gpu_engine.execute({'op': 'window', 'child': scan_plan('item', ['i_item_sk'])})
    with pytest.raises(EngineError):
        gpu_engine.execute(scan_plan('item', ['no_such_column']))
