"""ORACLE / TEST INFRASTRUCTURE -- not part of the product path.

CPU restatement of the six star-join NDS query shapes (q3, q7, q19, q42,
q52, q55) over the warehouse layout of /root/reference/nds/nds_transcode.py:131-150, in two
independent forms:

  * `run_query`      - pyarrow / Acero (multi-threaded): the checker for GPU parity tests and the
                       timed "pyarrow-Acero CPU, not Spark" baseline of bench.py;
  * `run_query_slow` - plain Python dict/Decimal loops, for cross-checking the Acero path on tiny data.

PARITY UNPINNED: the reference executes these statements inside Apache Spark
(/root/reference/nds/nds_power.py:125-135), which is not available on the build or GPU box, and it
ships no golden results.  The semantics restated here are SQL's plus the Spark rules SURVEY.md
marks [EXT]: SUM(decimal(p,s)) -> decimal(p+10,s); AVG(decimal(p,s)) -> decimal(p+4,s+4) rounded
HALF_UP; AVG(bigint) -> double; inner equi-joins drop NULL keys; GROUP BY groups NULLs together;
ORDER BY is NULLS FIRST ascending / NULLS LAST descending with UTF-8 byte order for strings; ties
are broken by the remaining output columns ascending (this repo's rule, applied on both sides).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import decimal
import functools
import os
import re

import pyarrow as pa
import pyarrow.compute as pc
import pyarrow.dataset as ds
import pyarrow.parquet as pq

HIVE_NULL = '__HIVE_DEFAULT_PARTITION__'


# ------------------------------------------------------------------------------------------ parameters
def _int(sql, pat):
    m = re.search(pat, sql)
    if not m:
        raise ValueError('oracle: parameter %r not found' % pat)
    return int(m.group(1))


def _str(sql, pat):
    m = re.search(pat, sql)
    if not m:
        raise ValueError('oracle: parameter %r not found' % pat)
    return m.group(1)


def template_of(sql):
    m = re.search(r'using template (\w+)\.tpl', sql)
    return m.group(1) if m else None


# ------------------------------------------------------------------------------------------ readers
def read_dim(warehouse, table, columns):
    d = os.path.join(warehouse, table)
    files = sorted(f for f in os.listdir(d) if f.endswith('.parquet'))
    return pa.concat_tables([pq.ParquetFile(os.path.join(d, f)).read(columns=columns) for f in files])


def read_fact(warehouse, table, part_col, columns, part_values=None, shard=None):
    """Read a hive-partitioned fact table, keeping only partitions whose value is in part_values
    (what Spark's dynamic partition pruning does for these joins).  shard=(rank, world) keeps every
    world-th file like the engine's multi-GPU sharding.  The selected files are read by a pool of
    threads (pyarrow releases the GIL while it decodes), so the scan uses the host's cores like the
    joins and the group-by do."""
    d = os.path.join(warehouse, table)
    file_idx = 0
    want = None if part_values is None else set(int(v) for v in part_values)
    cols = [c for c in columns if c != part_col]
    todo = []
    for e in sorted(os.listdir(d)):
        p = os.path.join(d, e)
        if not os.path.isdir(p):
            continue
        val = e.split('=', 1)[1]
        v = None if val == HIVE_NULL else int(val)
        for f in sorted(os.listdir(p)):
            if not f.endswith('.parquet'):
                continue
            mine = shard is None or file_idx % shard[1] == shard[0]
            file_idx += 1
            if not mine:
                continue
            if want is not None and (v is None or v not in want):
                continue
            todo.append((os.path.join(p, f), v))

    def read_one(job):
        path, v = job
        t = pq.ParquetFile(path).read(columns=cols, use_threads=False)
        if part_col in columns:
            t = t.append_column(part_col, pa.array([v] * t.num_rows, type=pa.int32()))
        return t

    if len(todo) > 4:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max(1, min(pa.cpu_count(), 64))) as ex:
            parts = list(ex.map(read_one, todo))
    else:
        parts = [read_one(j) for j in todo]
    if not parts:
        sch = pq.ParquetFile(_any_file(d)).schema_arrow
        fields = [sch.field(c) for c in cols] + ([pa.field(part_col, pa.int32())] if part_col in columns else [])
        return pa.schema(fields).empty_table()
    return pa.concat_tables(parts)


def _any_file(d):
    for root, _, files in os.walk(d):
        for f in files:
            if f.endswith('.parquet'):
                return os.path.join(root, f)
    raise FileNotFoundError(d)


# ------------------------------------------------------------------------------------------ ordering
def _cmp_val(a, b, desc):
    if a is None and b is None:
        return 0
    if a is None:
        return 1 if desc else -1          # NULLS FIRST asc, NULLS LAST desc
    if b is None:
        return -1 if desc else 1
    if isinstance(a, str):
        a, b = a.encode('utf-8'), b.encode('utf-8')
    if a == b:
        return 0
    r = -1 if a < b else 1
    return -r if desc else r


def order_rows(rows, names, order, limit):
    """rows: list of dicts. order: [(name, desc)]. Remaining columns are tie-breakers (asc)."""
    full = list(order) + [(n, False) for n in names if n not in [o[0] for o in order]]

    def cmp(x, y):
        for n, desc in full:
            c = _cmp_val(x[n], y[n], desc)
            if c:
                return c
        return 0
    out = sorted(rows, key=functools.cmp_to_key(cmp))
    return out if limit is None else out[:limit]


def _avg_dec(total, count):
    if count == 0 or total is None:
        return None
    return (decimal.Decimal(total) / decimal.Decimal(count)).quantize(decimal.Decimal('0.000001'),
                                                                     rounding=decimal.ROUND_HALF_UP)


def _to_table(rows, schema):
    return pa.Table.from_pylist(rows, schema=schema)


DEC172 = pa.decimal128(17, 2)
DEC116 = pa.decimal128(11, 6)


# ------------------------------------------------------------------------------------------ Acero path
def _star(warehouse, fact_cols, joins, date_filter, shard=None):
    """joins: [(fact_key, dim_table_filtered(pa.Table), dim_key)] applied in order. date_filter: filtered date_dim."""
    keys = date_filter['d_date_sk'].to_pylist()
    fact = read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', fact_cols, keys, shard)
    fact = fact.join(date_filter, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='inner')
    for fk, dim, dk in joins:
        fact = fact.join(dim, keys=fk, right_keys=dk, join_type='inner')
    return fact


def _group_sum(t, keys, value='ss_ext_sales_price'):
    g = t.group_by(keys, use_threads=True).aggregate([(value, 'sum'), (value, 'count')])
    rows = []
    for r in g.to_pylist():
        d = {k: r[k] for k in keys}
        d['_sum'] = r[value + '_sum'] if r[value + '_count'] else None
        rows.append(d)
    return rows


def _brand_query(sql, warehouse, with_year, shard, item_filter, date_filter, order, names, rename):
    dd = read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_moy'])
    dd = dd.filter(date_filter(dd))
    it = read_dim(warehouse, 'item', ['i_item_sk', 'i_brand_id', 'i_brand', 'i_manufact_id', 'i_manager_id'])
    it = it.filter(item_filter(it))
    t = _star(warehouse, ['ss_sold_date_sk', 'ss_item_sk', 'ss_ext_sales_price'], [('ss_item_sk', it, 'i_item_sk')], dd, shard)
    keys = (['d_year'] if with_year else []) + ['i_brand', 'i_brand_id']
    rows = []
    for r in _group_sum(t, keys):
        rows.append({rename.get(k, k): v for k, v in r.items()})
    return rows


def q3(sql, warehouse, shard=None, partial=False):
    manufact = _int(sql, r'i_manufact_id\s*=\s*(\d+)')
    moy = _int(sql, r'd_moy\s*=\s*(\d+)')
    rows = _brand_query(sql, warehouse, True, shard,
                        lambda it: pc.equal(it['i_manufact_id'], manufact).fill_null(False),
                        lambda dd: pc.equal(dd['d_moy'], moy).fill_null(False), None, None,
                        {'i_brand_id': 'brand_id', 'i_brand': 'brand', '_sum': 'sum_agg'})
    names = ['d_year', 'brand_id', 'brand', 'sum_agg']
    schema = pa.schema([('d_year', pa.int64()), ('brand_id', pa.int64()), ('brand', pa.string()), ('sum_agg', DEC172)])
    if partial:
        return rows, names, schema
    return _to_table(order_rows(rows, names, [('d_year', False), ('sum_agg', True), ('brand_id', False)], 100), schema)


def q52(sql, warehouse, shard=None, partial=False):
    moy = _int(sql, r'd_moy\s*=\s*(\d+)')
    year = _int(sql, r'd_year\s*=\s*(\d+)')
    mgr = _int(sql, r'i_manager_id\s*=\s*(\d+)')
    rows = _brand_query(sql, warehouse, True, shard,
                        lambda it: pc.equal(it['i_manager_id'], mgr).fill_null(False),
                        lambda dd: pc.and_(pc.equal(dd['d_moy'], moy), pc.equal(dd['d_year'], year)).fill_null(False), None, None,
                        {'i_brand_id': 'brand_id', 'i_brand': 'brand', '_sum': 'ext_price'})
    names = ['d_year', 'brand_id', 'brand', 'ext_price']
    schema = pa.schema([('d_year', pa.int64()), ('brand_id', pa.int64()), ('brand', pa.string()), ('ext_price', DEC172)])
    if partial:
        return rows, names, schema
    return _to_table(order_rows(rows, names, [('d_year', False), ('ext_price', True), ('brand_id', False)], 100), schema)


def q55(sql, warehouse, shard=None, partial=False):
    moy = _int(sql, r'd_moy\s*=\s*(\d+)')
    year = _int(sql, r'd_year\s*=\s*(\d+)')
    mgr = _int(sql, r'i_manager_id\s*=\s*(\d+)')
    rows = _brand_query(sql, warehouse, False, shard,
                        lambda it: pc.equal(it['i_manager_id'], mgr).fill_null(False),
                        lambda dd: pc.and_(pc.equal(dd['d_moy'], moy), pc.equal(dd['d_year'], year)).fill_null(False), None, None,
                        {'i_brand_id': 'brand_id', 'i_brand': 'brand', '_sum': 'ext_price'})
    names = ['brand_id', 'brand', 'ext_price']
    schema = pa.schema([('brand_id', pa.int64()), ('brand', pa.string()), ('ext_price', DEC172)])
    if partial:
        return rows, names, schema
    return _to_table(order_rows(rows, names, [('ext_price', True), ('brand_id', False)], 100), schema)


def q42(sql, warehouse, shard=None, partial=False):
    moy = _int(sql, r'd_moy\s*=\s*(\d+)')
    year = _int(sql, r'd_year\s*=\s*(\d+)')
    mgr = _int(sql, r'i_manager_id\s*=\s*(\d+)')
    dd = read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_moy'])
    dd = dd.filter(pc.and_(pc.equal(dd['d_moy'], moy), pc.equal(dd['d_year'], year)).fill_null(False))
    it = read_dim(warehouse, 'item', ['i_item_sk', 'i_category_id', 'i_category', 'i_manager_id'])
    it = it.filter(pc.equal(it['i_manager_id'], mgr).fill_null(False))
    t = _star(warehouse, ['ss_sold_date_sk', 'ss_item_sk', 'ss_ext_sales_price'], [('ss_item_sk', it, 'i_item_sk')], dd, shard)
    sname = 'sum(ss_ext_sales_price)'
    rows = [{'d_year': r['d_year'], 'i_category_id': r['i_category_id'], 'i_category': r['i_category'], sname: r['_sum']}
            for r in _group_sum(t, ['d_year', 'i_category_id', 'i_category'])]
    names = ['d_year', 'i_category_id', 'i_category', sname]
    schema = pa.schema([('d_year', pa.int64()), ('i_category_id', pa.int64()), ('i_category', pa.string()), (sname, DEC172)])
    if partial:
        return rows, names, schema
    return _to_table(order_rows(rows, names, [(sname, True), ('d_year', False), ('i_category_id', False), ('i_category', False)], 100), schema)


def q19(sql, warehouse, shard=None, partial=False):
    moy = _int(sql, r'd_moy\s*=\s*(\d+)')
    year = _int(sql, r'd_year\s*=\s*(\d+)')
    mgr = _int(sql, r'i_manager_id\s*=\s*(\d+)')
    dd = read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_moy'])
    dd = dd.filter(pc.and_(pc.equal(dd['d_moy'], moy), pc.equal(dd['d_year'], year)).fill_null(False))
    it = read_dim(warehouse, 'item', ['i_item_sk', 'i_brand_id', 'i_brand', 'i_manufact_id', 'i_manufact', 'i_manager_id'])
    it = it.filter(pc.equal(it['i_manager_id'], mgr).fill_null(False))
    cu = read_dim(warehouse, 'customer', ['c_customer_sk', 'c_current_addr_sk'])
    ca = read_dim(warehouse, 'customer_address', ['ca_address_sk', 'ca_zip'])
    st = read_dim(warehouse, 'store', ['s_store_sk', 's_zip'])
    cu = cu.join(ca, keys='c_current_addr_sk', right_keys='ca_address_sk', join_type='inner')
    t = _star(warehouse, ['ss_sold_date_sk', 'ss_item_sk', 'ss_customer_sk', 'ss_store_sk', 'ss_ext_sales_price'],
              [('ss_item_sk', it, 'i_item_sk'), ('ss_customer_sk', cu, 'c_customer_sk'), ('ss_store_sk', st, 's_store_sk')], dd, shard)
    ne = pc.not_equal(pc.utf8_slice_codeunits(t['ca_zip'], 0, 5), pc.utf8_slice_codeunits(t['s_zip'], 0, 5))
    t = t.filter(ne.combine_chunks().fill_null(False) if isinstance(ne, pa.ChunkedArray) else ne.fill_null(False))
    rows = [{'brand_id': r['i_brand_id'], 'brand': r['i_brand'], 'i_manufact_id': r['i_manufact_id'], 'i_manufact': r['i_manufact'],
             'ext_price': r['_sum']} for r in _group_sum(t, ['i_brand', 'i_brand_id', 'i_manufact_id', 'i_manufact'])]
    names = ['brand_id', 'brand', 'i_manufact_id', 'i_manufact', 'ext_price']
    schema = pa.schema([('brand_id', pa.int64()), ('brand', pa.string()), ('i_manufact_id', pa.int64()),
                        ('i_manufact', pa.string()), ('ext_price', DEC172)])
    if partial:
        return rows, names, schema
    return _to_table(order_rows(rows, names, [('ext_price', True), ('brand', False), ('brand_id', False),
                                             ('i_manufact_id', False), ('i_manufact', False)], 100), schema)


def q7(sql, warehouse, shard=None, partial=False):
    year = _int(sql, r'd_year\s*=\s*(\d+)')
    gen = _str(sql, r"cd_gender\s*=\s*'([^']*)'")
    ms = _str(sql, r"cd_marital_status\s*=\s*'([^']*)'")
    es = _str(sql, r"cd_education_status\s*=\s*'([^']*)'")
    dd = read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year'])
    dd = dd.filter(pc.equal(dd['d_year'], year).fill_null(False))
    it = read_dim(warehouse, 'item', ['i_item_sk', 'i_item_id'])
    cd = read_dim(warehouse, 'customer_demographics', ['cd_demo_sk', 'cd_gender', 'cd_marital_status', 'cd_education_status'])
    cd = cd.filter(pc.and_(pc.and_(pc.equal(cd['cd_gender'], gen), pc.equal(cd['cd_marital_status'], ms)),
                           pc.equal(cd['cd_education_status'], es)).fill_null(False))
    pr = read_dim(warehouse, 'promotion', ['p_promo_sk', 'p_channel_email', 'p_channel_event'])
    pr = pr.filter(pc.or_kleene(pc.equal(pr['p_channel_email'], 'N'), pc.equal(pr['p_channel_event'], 'N')).fill_null(False))
    t = _star(warehouse, ['ss_sold_date_sk', 'ss_item_sk', 'ss_cdemo_sk', 'ss_promo_sk', 'ss_quantity', 'ss_list_price',
                          'ss_coupon_amt', 'ss_sales_price'],
              [('ss_cdemo_sk', cd, 'cd_demo_sk'), ('ss_promo_sk', pr, 'p_promo_sk'), ('ss_item_sk', it, 'i_item_sk')], dd, shard)
    vals = ['ss_quantity', 'ss_list_price', 'ss_coupon_amt', 'ss_sales_price']
    g = t.group_by(['i_item_id'], use_threads=True).aggregate([(v, f) for v in vals for f in ('sum', 'count')])
    rows = []
    for r in g.to_pylist():
        qs, qc = r['ss_quantity_sum'], r['ss_quantity_count']
        row = {'i_item_id': r['i_item_id']}
        if partial:
            for v in vals:
                row[v + '_sum'] = r[v + '_sum'] if r[v + '_count'] else None
                row[v + '_count'] = r[v + '_count']
        else:
            row['agg1'] = (float(qs) / float(qc)) if qc else None
            row['agg2'] = _avg_dec(r['ss_list_price_sum'], r['ss_list_price_count'])
            row['agg3'] = _avg_dec(r['ss_coupon_amt_sum'], r['ss_coupon_amt_count'])
            row['agg4'] = _avg_dec(r['ss_sales_price_sum'], r['ss_sales_price_count'])
        rows.append(row)
    names = ['i_item_id', 'agg1', 'agg2', 'agg3', 'agg4']
    schema = pa.schema([('i_item_id', pa.string()), ('agg1', pa.float64()), ('agg2', DEC116), ('agg3', DEC116), ('agg4', DEC116)])
    if partial:
        return rows, names, schema
    return _to_table(order_rows(rows, names, [('i_item_id', False)], 100), schema)


def q98(sql, warehouse, shard=None, partial=False):
    """TPC-DS query98 (Spark dialect of the NDS templates): item revenue over 31 days and its share of the class's revenue -
    sum(sum(x)) over (partition by i_class) restated as a second pass over the groups.  The ratio is a double (the engine's
    `/`; Spark would give a decimal), compared with a relative tolerance by the tests."""
    import datetime
    cats = re.search(r"i_category\s+in\s*\(\s*'([^']*)'\s*,\s*'([^']*)'\s*,\s*'([^']*)'\s*\)", sql).groups()
    d0 = datetime.date.fromisoformat(_str(sql, r"cast\('([^']*)'\s+as\s+date\)"))
    days = _int(sql, r'interval\s+(\d+)\s+days')
    d1 = d0 + datetime.timedelta(days=days)
    dd = read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_date'])
    dd = dd.filter(pc.and_(pc.greater_equal(dd['d_date'], pa.scalar(d0, type=dd.schema.field('d_date').type)),
                           pc.less_equal(dd['d_date'], pa.scalar(d1, type=dd.schema.field('d_date').type))).fill_null(False))
    dd = dd.select(['d_date_sk'])
    it = read_dim(warehouse, 'item', ['i_item_sk', 'i_item_id', 'i_item_desc', 'i_category', 'i_class', 'i_current_price'])
    it = it.filter(pc.is_in(it['i_category'], value_set=pa.array(list(cats))).fill_null(False))
    t = _star(warehouse, ['ss_sold_date_sk', 'ss_item_sk', 'ss_ext_sales_price'], [('ss_item_sk', it, 'i_item_sk')], dd, shard)
    keys = ['i_item_id', 'i_item_desc', 'i_category', 'i_class', 'i_current_price']
    groups = _group_sum(t, keys)
    total = {}
    for r in groups:
        if r['_sum'] is not None:
            total[r['i_class']] = total.get(r['i_class'], 0) + r['_sum']
    rows = []
    for r in groups:
        s, tot = r['_sum'], total.get(r['i_class'])
        ratio = (float(s * 100) / float(tot)) if (s is not None and tot) else None
        row = {k: r[k] for k in keys}
        row['itemrevenue'] = s
        row['revenueratio'] = ratio
        rows.append(row)
    names = keys + ['itemrevenue', 'revenueratio']
    schema = pa.schema([('i_item_id', pa.string()), ('i_item_desc', pa.string()), ('i_category', pa.string()), ('i_class', pa.string()),
                        ('i_current_price', pa.decimal128(7, 2)), ('itemrevenue', DEC172), ('revenueratio', pa.float64())])
    order = [('i_category', False), ('i_class', False), ('i_item_id', False), ('i_item_desc', False), ('revenueratio', False)]
    return _to_table(order_rows(rows, names, order, None), schema)


QUERIES = {'query3': q3, 'query7': q7, 'query19': q19, 'query42': q42, 'query52': q52, 'query55': q55, 'query98': q98}


def run_query(name, sql, warehouse, shard=None):
    """Acero restatement. name: template name (query3 ...)."""
    if name not in QUERIES:
        raise KeyError('oracle has no restatement of %s' % name)
    return QUERIES[name](sql, warehouse, shard)


# ------------------------------------------------------------------------------------------ slow path
def run_query_slow(name, sql, warehouse):
    """Plain-Python restatement (dicts, Decimal, explicit loops) - independent of Acero's join / group-by."""
    def rows_of(table, cols, fact=False):
        if fact:
            return read_fact(warehouse, table, 'ss_sold_date_sk', cols).to_pylist()
        return read_dim(warehouse, table, cols).to_pylist()

    def index(rows, key):
        return {r[key]: r for r in rows if r[key] is not None}

    p = {}
    for k, pat in (('moy', r'd_moy\s*=\s*(\d+)'), ('year', r'd_year\s*=\s*(\d+)'), ('mgr', r'i_manager_id\s*=\s*(\d+)'),
                   ('manufact', r'i_manufact_id\s*=\s*(\d+)')):
        m = re.search(pat, sql)
        p[k] = int(m.group(1)) if m else None
    dd = index([r for r in rows_of('date_dim', ['d_date_sk', 'd_year', 'd_moy'])
                if (p['moy'] is None or r['d_moy'] == p['moy']) and (p['year'] is None or r['d_year'] == p['year'])], 'd_date_sk')
    item_cols = ['i_item_sk', 'i_item_id', 'i_brand_id', 'i_brand', 'i_manufact_id', 'i_manufact', 'i_manager_id', 'i_category_id', 'i_category']
    it = index([r for r in rows_of('item', item_cols)
                if (p['mgr'] is None or r['i_manager_id'] == p['mgr']) and (p['manufact'] is None or r['i_manufact_id'] == p['manufact'])],
               'i_item_sk')
    fact_cols = ['ss_sold_date_sk', 'ss_item_sk', 'ss_customer_sk', 'ss_store_sk', 'ss_cdemo_sk', 'ss_promo_sk', 'ss_quantity',
                 'ss_list_price', 'ss_coupon_amt', 'ss_sales_price', 'ss_ext_sales_price']
    fact = rows_of('store_sales', fact_cols, fact=True)
    joined = []
    if name == 'query7':
        gen = _str(sql, r"cd_gender\s*=\s*'([^']*)'")
        ms = _str(sql, r"cd_marital_status\s*=\s*'([^']*)'")
        es = _str(sql, r"cd_education_status\s*=\s*'([^']*)'")
        cd = index([r for r in rows_of('customer_demographics', ['cd_demo_sk', 'cd_gender', 'cd_marital_status', 'cd_education_status'])
                    if r['cd_gender'] == gen and r['cd_marital_status'] == ms and r['cd_education_status'] == es], 'cd_demo_sk')
        pr = index([r for r in rows_of('promotion', ['p_promo_sk', 'p_channel_email', 'p_channel_event'])
                    if r['p_channel_email'] == 'N' or r['p_channel_event'] == 'N'], 'p_promo_sk')
    if name == 'query19':
        ca = index(rows_of('customer_address', ['ca_address_sk', 'ca_zip']), 'ca_address_sk')
        cu = index(rows_of('customer', ['c_customer_sk', 'c_current_addr_sk']), 'c_customer_sk')
        st = index(rows_of('store', ['s_store_sk', 's_zip']), 's_store_sk')
    for f in fact:
        d = dd.get(f['ss_sold_date_sk'])
        i = it.get(f['ss_item_sk'])
        if d is None or i is None:
            continue
        r = dict(f)
        r.update(d)
        r.update(i)
        if name == 'query7':
            if f['ss_cdemo_sk'] not in cd or f['ss_promo_sk'] not in pr:
                continue
        if name == 'query19':
            c = cu.get(f['ss_customer_sk'])
            s = st.get(f['ss_store_sk'])
            if c is None or s is None:
                continue
            a = ca.get(c['c_current_addr_sk'])
            if a is None or a['ca_zip'] is None or s['s_zip'] is None or a['ca_zip'][:5] == s['s_zip'][:5]:
                continue
        joined.append(r)
    spec = {
        'query3': (['d_year', 'i_brand', 'i_brand_id'], {'d_year': 'd_year', 'i_brand_id': 'brand_id', 'i_brand': 'brand'}, 'sum_agg',
                   ['d_year', 'brand_id', 'brand', 'sum_agg'], [('d_year', False), ('sum_agg', True), ('brand_id', False)]),
        'query52': (['d_year', 'i_brand', 'i_brand_id'], {'d_year': 'd_year', 'i_brand_id': 'brand_id', 'i_brand': 'brand'}, 'ext_price',
                    ['d_year', 'brand_id', 'brand', 'ext_price'], [('d_year', False), ('ext_price', True), ('brand_id', False)]),
        'query55': (['i_brand', 'i_brand_id'], {'i_brand_id': 'brand_id', 'i_brand': 'brand'}, 'ext_price',
                    ['brand_id', 'brand', 'ext_price'], [('ext_price', True), ('brand_id', False)]),
        'query42': (['d_year', 'i_category_id', 'i_category'], {'d_year': 'd_year', 'i_category_id': 'i_category_id', 'i_category': 'i_category'},
                    'sum(ss_ext_sales_price)', ['d_year', 'i_category_id', 'i_category', 'sum(ss_ext_sales_price)'],
                    [('sum(ss_ext_sales_price)', True), ('d_year', False), ('i_category_id', False), ('i_category', False)]),
        'query19': (['i_brand', 'i_brand_id', 'i_manufact_id', 'i_manufact'],
                    {'i_brand_id': 'brand_id', 'i_brand': 'brand', 'i_manufact_id': 'i_manufact_id', 'i_manufact': 'i_manufact'}, 'ext_price',
                    ['brand_id', 'brand', 'i_manufact_id', 'i_manufact', 'ext_price'],
                    [('ext_price', True), ('brand', False), ('brand_id', False), ('i_manufact_id', False), ('i_manufact', False)]),
    }
    if name == 'query7':
        groups = {}
        for r in joined:
            groups.setdefault(r['i_item_id'], []).append(r)
        rows = []
        for k, rs in groups.items():
            def avg_dec(c):
                v = [x[c] for x in rs if x[c] is not None]
                return _avg_dec(sum(v), len(v)) if v else None
            qv = [x['ss_quantity'] for x in rs if x['ss_quantity'] is not None]
            rows.append({'i_item_id': k, 'agg1': (float(sum(qv)) / len(qv)) if qv else None, 'agg2': avg_dec('ss_list_price'),
                         'agg3': avg_dec('ss_coupon_amt'), 'agg4': avg_dec('ss_sales_price')})
        return order_rows(rows, ['i_item_id', 'agg1', 'agg2', 'agg3', 'agg4'], [('i_item_id', False)], 100)
    keys, rename, sname, names, order = spec[name]
    groups = {}
    for r in joined:
        groups.setdefault(tuple(r[k] for k in keys), []).append(r['ss_ext_sales_price'])
    rows = []
    for k, vs in groups.items():
        row = {rename[kn]: kv for kn, kv in zip(keys, k)}
        nn = [v for v in vs if v is not None]
        row[sname] = sum(nn) if nn else None
        rows.append(row)
    return order_rows(rows, names, order, 100)
