"""Seeded synthetic NDS warehouse generator.

Stand-in for `nds_gen_data.py` (TPC dsdgen, absent here) + `nds_transcode.py`: emits the
24 source tables with the column types of the reference's schema catalogue
(/root/reference/nds/nds_schema.py:49-568) in the on-disk layout the reference's
`store()` produces (/root/reference/nds/nds_transcode.py:131-150):

  * the 7 fact tables of TABLE_PARTITIONING are hive-partitioned by their date_sk column
    (`<table>/<col>=<value>/part-*.parquet`, NULL -> `__HIVE_DEFAULT_PARTITION__`), the
    partition column is NOT stored in the files, rows sorted by it within a partition;
  * every other table is one file (`coalesce(1)`).

Files are written uncompressed (`--compression none` in the reference's CLI) with
dictionary + RLE/bit-packed hybrid pages, data page v1, decimals as INT32/INT64.

This is NOT dsdgen data: value distributions follow SURVEY.md section 8(d).
"""
import os
import zlib
import math
import datetime
import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
from concurrent.futures import ProcessPoolExecutor

from . import nds_schema

SEED = 20261106
NULL_FRAC = 0.04
FIRST_DATE_SK = 2415022            # 1900-01-02 in TPC-DS
DATE_DIM_ROWS = 73049
SALES_FIRST_SK = 2450816           # 1998-01-02
SALES_DAYS = 1823
_CFG = {'sales_days': SALES_DAYS}    # tests shrink the partition count; bench keeps 1823
HIVE_NULL = '__HIVE_DEFAULT_PARTITION__'
EPOCH = datetime.date(1970, 1, 1)
DAY0 = (datetime.date(1900, 1, 2) - EPOCH).days

# rows at SF 1 / 10 / 100 / 1000 ([EXT] TPC-DS spec table 3-2)
_ROWS = {
    'store_sales': {1: 2880404, 10: 28800991, 100: 287997024, 1000: 2879987999},
    'item': {1: 18000, 10: 102000, 100: 204000, 1000: 300000},
    'customer': {1: 100000, 10: 500000, 100: 2000000, 1000: 12000000},
    'customer_address': {1: 50000, 10: 250000, 100: 1000000, 1000: 6000000},
    'customer_demographics': {1: 1920800, 10: 1920800, 100: 1920800, 1000: 1920800},
    'store': {1: 12, 10: 102, 100: 402, 1000: 1002},
    'promotion': {1: 300, 10: 500, 100: 1000, 1000: 1500},
    'date_dim': {1: DATE_DIM_ROWS, 10: DATE_DIM_ROWS, 100: DATE_DIM_ROWS, 1000: DATE_DIM_ROWS},
}
# tables the star-join subset (q3,q7,q19,q42,q52,q55) touches; the rest are generated at a
# token size (schema-complete, so the 24 CreateTempView registrations all succeed).
SUBSET_TABLES = tuple(_ROWS)
_TOKEN_ROWS = {'time_dim': 86400, 'household_demographics': 7200, 'income_band': 20,
               'reason': 35, 'ship_mode': 20, 'warehouse': 5, 'call_center': 6, 'web_site': 30,
               'web_page': 60, 'catalog_page': 11718}

EDUCATION = ['Primary', 'Secondary', 'College', '2 yr Degree', '4 yr Degree', 'Advanced Degree',
             'Unknown']
MARITAL = ['M', 'S', 'D', 'W', 'U']
CATEGORIES = ['Books', 'Children', 'Electronics', 'Home', 'Jewelry', 'Men', 'Music', 'Shoes',
              'Sports', 'Women']
_SYL = ['amalg', 'edu pack', 'export', 'import', 'scholar', 'brand', 'corp', 'maxi', 'univ',
        'nameless']


def table_rows(table, sf):
    if table in _ROWS:
        d = _ROWS[table]
        if sf in d:
            return d[sf]
        if sf < 1:
            if table == 'date_dim':
                return DATE_DIM_ROWS
            return max(40 if table not in ('store',) else 6, int(d[1] * sf))
        # log-linear interpolation between the tabulated scale factors
        ks = sorted(d)
        lo = max(k for k in ks if k <= sf) if sf >= ks[0] else ks[0]
        hi = min(k for k in ks if k >= sf) if sf <= ks[-1] else ks[-1]
        if lo == hi:
            return int(d[lo] * sf / lo)
        t = (math.log(sf) - math.log(lo)) / (math.log(hi) - math.log(lo))
        return int(round(math.exp(math.log(d[lo]) * (1 - t) + math.log(d[hi]) * t)))
    if table in _TOKEN_ROWS:
        return _TOKEN_ROWS[table] if sf >= 1 else max(5, int(_TOKEN_ROWS[table] * max(sf, 0.01)))
    # non-subset fact tables: token size
    return max(200, int(2000 * min(sf, 10)))


def _rng(seed, table, part=0):
    return np.random.default_rng([seed, zlib.crc32(table.encode()), part])


def _mask(rng, n, nullable, frac=NULL_FRAC):
    if not nullable or n == 0:
        return None
    return rng.random(n) < frac


def _fk_target(col):
    for suf, tbl in (('_date_sk', 'date_dim'), ('_time_sk', 'time_dim'), ('_item_sk', 'item'),
                     ('_customer_sk', 'customer'), ('_cdemo_sk', 'customer_demographics'),
                     ('_hdemo_sk', 'household_demographics'), ('_addr_sk', 'customer_address'),
                     ('_store_sk', 'store'), ('_promo_sk', 'promotion'),
                     ('_warehouse_sk', 'warehouse'), ('_ship_mode_sk', 'ship_mode'),
                     ('_reason_sk', 'reason'), ('_call_center_sk', 'call_center'),
                     ('_web_page_sk', 'web_page'), ('_web_site_sk', 'web_site'),
                     ('_catalog_page_sk', 'catalog_page'), ('_income_band_sk', 'income_band')):
        if col.endswith(suf):
            return tbl
    return None


def _strings(values, mask):
    return pa.array(values, type=pa.string(), mask=mask)


def _vocab_col(rng, n, vocab, mask):
    idx = rng.integers(0, len(vocab), n)
    arr = pa.DictionaryArray.from_arrays(pa.array(idx.astype(np.int32), mask=mask),
                                         pa.array(vocab, type=pa.string()))
    return arr.cast(pa.string())


def _item_id(k):
    # 16 chars: 'AAAAAAAA' + 8 base-16 letters, least significant first (dsdgen-like look)
    s = []
    for _ in range(8):
        s.append(chr(ord('A') + (k & 15)))
        k >>= 4
    return 'AAAAAAAA' + ''.join(s)


def _brand_vocab(n=1000):
    return ['%s%s #%d' % (_SYL[i % 10], _SYL[(i // 10) % 10], i // 100 + 1) for i in range(n)]


def _gen_column(rng, table, name, typ, nullable, n, sf, row0=0):
    """Generic typed column; special-cased columns are handled before this is called."""
    m = _mask(rng, n, nullable)
    if typ == 'int32':
        tgt = _fk_target(name)
        if tgt == 'date_dim':
            v = rng.integers(SALES_FIRST_SK, SALES_FIRST_SK + _CFG['sales_days'], n)
        elif tgt is not None:
            v = rng.integers(1, table_rows(tgt, sf) + 1, n)
        else:
            v = rng.integers(1, 1001, n)
        return pa.array(v.astype(np.int32), mask=m)
    if typ == 'int64':
        return pa.array(rng.integers(1, 101, n).astype(np.int64), mask=m)
    if typ == 'double':
        return pa.array(rng.random(n) * 1000.0, mask=m)
    if typ == 'date':
        v = rng.integers(DAY0 + 35000, DAY0 + 38000, n).astype(np.int32)
        return pa.array(v, type=pa.int32(), mask=m).cast(pa.date32())
    md = nds_schema._DEC.match(typ)
    if md:
        p, s = int(md.group(1)), int(md.group(2))
        hi = min(2_000_000, 10 ** p)
        v = rng.integers(0, hi, n).astype(np.int64)
        it = pa.int32() if p <= 9 else pa.int64()
        return _decimal_from_unscaled(pa.array(v, type=pa.int64(), mask=m).cast(it), p, s)
    mc = nds_schema._CHR.match(typ)
    if mc:
        w = int(mc.group(2))
        if w == 1:
            return _vocab_col(rng, n, ['Y', 'N'], m)
        if w == 2:
            return _vocab_col(rng, n, ['AL', 'CA', 'GA', 'IL', 'NY', 'OH', 'TN', 'TX', 'VA', 'WA'], m)
        vocab = [('%s_%03d' % (name, j))[:w] for j in range(50)]
        return _vocab_col(rng, n, vocab, m)
    raise ValueError(typ)


def _decimal_from_unscaled(int_arr, p, s):
    """int32/int64 unscaled -> decimal128(p, s) (exact)."""
    return _rescale(int_arr, p, s)


def _rescale(int_arr, p, s):
    # arrow has no "reinterpret unscaled" cast: build the decimal buffer directly
    v = int_arr.cast(pa.int64())
    vals = np.asarray(v.fill_null(0)).astype(np.int64)
    buf = np.empty((len(vals), 2), dtype=np.int64)
    buf[:, 0] = vals
    buf[:, 1] = vals >> 63
    validity = v.buffers()[0] if v.null_count else None
    if validity is not None and v.offset != 0:
        validity = pa.array(np.asarray(v.is_valid())).buffers()[1]
    return pa.Array.from_buffers(pa.decimal128(p, s), len(vals),
                                 [validity, pa.py_buffer(buf.tobytes())], null_count=v.null_count)


def _zip_pool(seed):
    r = np.random.default_rng([seed, 99991])
    return ['%05d' % z for z in r.choice(90000, size=400, replace=False) + 10000]


def _gen_table(table, n, sf, seed, part=0, row0=0, skip_cols=()):
    """One pyarrow Table of n rows (dimension table, or one partition of a fact table)."""
    rng = _rng(seed, table, part)
    cols = {}
    spec = nds_schema.column_spec(table)
    idx = np.arange(row0, row0 + n, dtype=np.int64)
    brand_vocab = _brand_vocab()
    for ci, (name, typ, nullable) in enumerate(spec):
        if name in skip_cols:
            continue
        crng = np.random.default_rng([seed, zlib.crc32(table.encode()), part, ci])
        # ---- primary keys -------------------------------------------------------------
        if ci == 0 and not nullable and name.endswith('_sk') and table not in nds_schema.TABLE_PARTITIONING:
            base = FIRST_DATE_SK if table == 'date_dim' else 1
            cols[name] = pa.array((idx + base).astype(np.int32))
            continue
        # ---- date_dim -----------------------------------------------------------------
        if table == 'date_dim' and name in ('d_date', 'd_year', 'd_moy', 'd_dom', 'd_qoy', 'd_dow',
                                            'd_month_seq', 'd_week_seq', 'd_quarter_seq',
                                            'd_date_id'):
            days = (idx + DAY0).astype('datetime64[D]')
            y = days.astype('datetime64[Y]').astype(np.int64) + 1970
            mth = days.astype('datetime64[M]').astype(np.int64) % 12 + 1
            dom = (days - days.astype('datetime64[M]')).astype(np.int64) + 1
            if name == 'd_date':
                cols[name] = pa.array((idx + DAY0).astype(np.int32), type=pa.int32()).cast(pa.date32())
            elif name == 'd_year':
                cols[name] = pa.array(y)
            elif name == 'd_moy':
                cols[name] = pa.array(mth)
            elif name == 'd_dom':
                cols[name] = pa.array(dom)
            elif name == 'd_qoy':
                cols[name] = pa.array((mth - 1) // 3 + 1)
            elif name == 'd_dow':
                cols[name] = pa.array((idx + 2) % 7)
            elif name == 'd_month_seq':
                cols[name] = pa.array((y - 1900) * 12 + mth - 1)
            elif name == 'd_week_seq':
                cols[name] = pa.array(idx // 7 + 1)
            elif name == 'd_quarter_seq':
                cols[name] = pa.array((y - 1900) * 4 + (mth - 1) // 3 + 1)
            else:
                cols[name] = _strings([_item_id(int(k)) for k in idx], None)
            continue
        # ---- item ---------------------------------------------------------------------
        if table == 'item':
            if name == 'i_item_id':
                cols[name] = _strings([_item_id(int(k) // 2) for k in idx], None)
                continue
            if name in ('i_brand_id', 'i_brand'):
                r2 = np.random.default_rng([seed, zlib.crc32(b'item'), part, 7777])
                bidx = r2.integers(0, 1000, n)
                if name == 'i_brand_id':
                    cols[name] = pa.array((1001001 + bidx * 1001 % 9000000).astype(np.int64),
                                          mask=_mask(crng, n, nullable))
                else:
                    flip = r2.random(n) < 0.1
                    nidx = np.where(flip, r2.integers(0, 1000, n), bidx)
                    arr = pa.DictionaryArray.from_arrays(
                        pa.array(nidx.astype(np.int32), mask=_mask(crng, n, nullable)),
                        pa.array(brand_vocab, type=pa.string()))
                    cols[name] = arr.cast(pa.string())
                continue
            if name in ('i_manufact_id', 'i_manufact'):
                r2 = np.random.default_rng([seed, zlib.crc32(b'item'), part, 8888])
                midx = r2.integers(0, 1000, n)
                if name == 'i_manufact_id':
                    cols[name] = pa.array((midx + 1).astype(np.int64), mask=_mask(crng, n, nullable))
                else:
                    vocab = ['%s%s%d' % (_SYL[(i // 7) % 10], _SYL[i % 10], i % 13) for i in range(1000)]
                    flip = r2.random(n) < 0.05
                    nidx = np.where(flip, r2.integers(0, 1000, n), midx)
                    arr = pa.DictionaryArray.from_arrays(
                        pa.array(nidx.astype(np.int32), mask=_mask(crng, n, nullable)),
                        pa.array(vocab, type=pa.string()))
                    cols[name] = arr.cast(pa.string())
                continue
            if name == 'i_manager_id':
                cols[name] = pa.array(crng.integers(1, 101, n).astype(np.int64),
                                      mask=_mask(crng, n, nullable))
                continue
            if name in ('i_category_id', 'i_category'):
                r2 = np.random.default_rng([seed, zlib.crc32(b'item'), part, 9999])
                cidx = r2.integers(0, 10, n)
                if name == 'i_category_id':
                    cols[name] = pa.array((cidx + 1).astype(np.int64), mask=_mask(crng, n, nullable))
                else:
                    arr = pa.DictionaryArray.from_arrays(
                        pa.array(cidx.astype(np.int32), mask=_mask(crng, n, nullable)),
                        pa.array(CATEGORIES, type=pa.string()))
                    cols[name] = arr.cast(pa.string())
                continue
            if name == 'i_item_desc':
                # free text: long, mostly unique -> exceeds the dictionary page, PLAIN fallback
                words = ['able', 'bar', 'cally', 'dote', 'eing', 'ese', 'ought', 'pri', 'anti', 'n st']
                w = crng.integers(0, 10, (n, 6))
                vals = [' '.join(words[j] for j in row) + ' %d' % (row0 + i)
                        for i, row in enumerate(w)]
                cols[name] = _strings(vals, _mask(crng, n, nullable))
                continue
        # ---- customer_demographics: mixed-radix cross product like dsdgen -------------------
        if table == 'customer_demographics' and name in ('cd_gender', 'cd_marital_status',
                                                         'cd_education_status'):
            if name == 'cd_gender':
                vocab, k = ['M', 'F'], idx % 2
            elif name == 'cd_marital_status':
                vocab, k = MARITAL, (idx // 2) % 5
            else:
                vocab, k = EDUCATION, (idx // 10) % 7
            arr = pa.DictionaryArray.from_arrays(pa.array(k.astype(np.int32),
                                                          mask=_mask(crng, n, nullable, 0.01)),
                                                 pa.array(vocab, type=pa.string()))
            cols[name] = arr.cast(pa.string())
            continue
        # ---- zips: shared pool so substr(ca_zip,1,5) = substr(s_zip,1,5) happens ---------
        if name in ('ca_zip', 's_zip'):
            pool = _zip_pool(seed)
            k = crng.integers(0, len(pool), n)
            if name == 'ca_zip':
                plus4 = crng.random(n) < 0.1
                vals = [pool[a] + ('-%04d' % (a * 7 % 10000) if p4 else '') for a, p4 in zip(k, plus4)]
            else:
                vals = [pool[a] for a in k]
            cols[name] = _strings(vals, _mask(crng, n, nullable))
            continue
        if table == 'store_sales' and name == 'ss_ticket_number':
            cols[name] = pa.array(idx // 8 + 1 + part * 100000000)
            continue
        cols[name] = _gen_column(crng, table, name, typ, nullable, n, sf, row0)
    sch = nds_schema.get_schemas(True)[table]
    return pa.table(cols).cast(pa.schema([sch.field(c) for c in cols]))


WRITE_KW = dict(compression='NONE', use_dictionary=True, data_page_version='1.0',
                store_decimal_as_integer=True, write_statistics=True)


def _write_partition(args):
    table, out_dir, n, sf, seed, part, pcol, pval, row_group_size, sales_days = args
    _CFG['sales_days'] = sales_days
    t = _gen_table(table, n, sf, seed, part=part, row0=0, skip_cols=(pcol,))
    d = os.path.join(out_dir, '%s=%s' % (pcol, pval))
    os.makedirs(d, exist_ok=True)
    pq.write_table(t, os.path.join(d, 'part-%05d-b200gen.parquet' % part),
                   row_group_size=row_group_size, **WRITE_KW)
    return n


def partition_sizes(table, n, seed):
    """rows per date partition (uniform over SALES_DAYS) + the NULL partition (~4 %)."""
    rng = _rng(seed, table, 1 << 30)
    n_null = int(rng.binomial(n, NULL_FRAC)) if n >= 100 else 0
    days = _CFG['sales_days']
    counts = rng.multinomial(n - n_null, np.full(days, 1.0 / days))
    return counts, n_null


def generate(out_prefix, sf, seed=SEED, tables=None, workers=None, row_group_size=1 << 20,
             verbose=False, sales_days=SALES_DAYS):
    """Write the warehouse under out_prefix. Returns {table: rows}."""
    _CFG['sales_days'] = sales_days
    os.makedirs(out_prefix, exist_ok=True)
    schemas = nds_schema.get_schemas(True)
    names = list(schemas) if tables is None else list(tables)
    workers = workers or min(16, os.cpu_count() or 1)
    report = {}
    for table in names:
        n = table_rows(table, sf)
        out_dir = os.path.join(out_prefix, table)
        os.makedirs(out_dir, exist_ok=True)
        if table in nds_schema.TABLE_PARTITIONING:
            pcol = nds_schema.TABLE_PARTITIONING[table]
            counts, n_null = partition_sizes(table, n, seed)
            jobs = [(table, out_dir, int(c), sf, seed, i, pcol, str(SALES_FIRST_SK + i), row_group_size, sales_days)
                    for i, c in enumerate(counts) if c > 0]
            if n_null:
                jobs.append((table, out_dir, n_null, sf, seed, sales_days, pcol, HIVE_NULL,
                             row_group_size, sales_days))
            if workers > 1 and n > 200000:
                with ProcessPoolExecutor(workers) as ex:
                    list(ex.map(_write_partition, jobs, chunksize=16))
            else:
                for j in jobs:
                    _write_partition(j)
        else:
            # one file, generated in row blocks so each block has its own substream
            block = 1 << 20
            parts = [_gen_table(table, min(block, n - r0), sf, seed, part=bi, row0=r0)
                     for bi, r0 in enumerate(range(0, n, block))]
            t = pa.concat_tables(parts)
            pq.write_table(t, os.path.join(out_dir, 'part-00000-b200gen.parquet'),
                           row_group_size=row_group_size, **WRITE_KW)
        report[table] = n
        if verbose:
            print('generated %s: %d rows' % (table, n))
    with open(os.path.join(out_prefix, '_b200gen.json'), 'w') as f:
        import json
        json.dump({'sf': sf, 'seed': seed, 'rows': report, 'generator': 'b200gen-1',
                   'sales_days': sales_days}, f)
    return report


def ensure(out_prefix, sf, seed=SEED, tables=None, sales_days=SALES_DAYS, **kw):
    """Generate unless an identical warehouse (same sf/seed/tables) already exists."""
    import json
    marker = os.path.join(out_prefix, '_b200gen.json')
    want = list(nds_schema.get_schemas(True)) if tables is None else list(tables)
    if os.path.exists(marker):
        with open(marker) as f:
            m = json.load(f)
        if m.get('sf') == sf and m.get('seed') == seed and all(t in m.get('rows', {}) for t in want) \
                and m.get('generator') == 'b200gen-1' and m.get('sales_days') == sales_days:
            return m['rows']
    return generate(out_prefix, sf, seed, tables, sales_days=sales_days, **kw)


if __name__ == '__main__':
    import argparse
    ap = argparse.ArgumentParser(description='seeded synthetic NDS warehouse (not dsdgen data)')
    ap.add_argument('out_prefix')
    ap.add_argument('--sf', type=float, default=1.0)
    ap.add_argument('--seed', type=int, default=SEED)
    ap.add_argument('--tables', type=lambda s: s.split(','))
    ap.add_argument('--sales_days', type=int, default=SALES_DAYS)
    a = ap.parse_args()
    sf = int(a.sf) if a.sf == int(a.sf) else a.sf
    print(generate(a.out_prefix, sf, a.seed, a.tables, verbose=True, sales_days=a.sales_days))
