"""GPU parity for the SQL shapes beyond the inner star join: LEFT OUTER JOIN, [NOT] EXISTS and
[NOT] IN subqueries (semi / anti joins), joins on keys that repeat on the build side, SELECT DISTINCT and HAVING - SQL text
through the planner and the C ABI, expected rows restated with pyarrow joins / group-bys on the same files."""
import decimal

import pyarrow as pa
import pyarrow.compute as pc
import pytest

from oracle import nds_queries as oq
from tests.test_gpu_queries import _Sess

pytestmark = pytest.mark.gpu


def _rows(t, names):
    return [tuple(r[n] for n in names) for r in t.to_pylist()]


def _key(row):
    return tuple((v is not None, v.encode() if isinstance(v, str) else (v if v is not None else 0)) for v in row)


def test_left_join_repeated_keys_having(gpu_engine, warehouse):
    sql = ("select i_category, count(*) c, count(p_promo_sk) cp, sum(p_cost) pc from item left join promotion "
           "on p_item_sk = i_item_sk and p_channel_tv = 'N' where i_manager_id < 70 group by i_category "
           "having count(*) > 2 order by i_category limit 1000")
    got = _Sess(gpu_engine).sql(sql).toArrow()
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category', 'i_manager_id'])
    it = it.filter(pc.less(it['i_manager_id'], 70).fill_null(False))
    pr = oq.read_dim(warehouse, 'promotion', ['p_promo_sk', 'p_item_sk', 'p_cost', 'p_channel_tv'])
    pr = pr.filter(pc.equal(pr['p_channel_tv'], 'N').fill_null(False)).filter(pc.is_valid(pr.filter(pc.equal(pr['p_channel_tv'], 'N').fill_null(False))['p_item_sk']))
    j = it.join(pr, keys='i_item_sk', right_keys='p_item_sk', join_type='left outer')
    g = j.group_by('i_category').aggregate([([], 'count_all'), ('p_promo_sk', 'count'), ('p_cost', 'sum')])
    exp = sorted([(r['i_category'], r['count_all'], r['p_promo_sk_count'], r['p_cost_sum']) for r in g.to_pylist() if r['count_all'] > 2], key=_key)[:1000]
    assert _rows(got, ['i_category', 'c', 'cp', 'pc']) == exp
    assert len(exp) > 3 and any(r[2] > 0 for r in exp) and j.num_rows > it.num_rows


def test_left_join_where_on_the_outer_side(gpu_engine, warehouse):
    """A WHERE condition on the left-joined table runs after the join: `p_promo_sk is null` keeps exactly the unmatched rows."""
    sql = ("select i_item_sk from item left join promotion on p_item_sk = i_item_sk where p_promo_sk is null and i_manager_id = 7 "
           "order by i_item_sk")
    got = _Sess(gpu_engine).sql(sql).toArrow()
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_manager_id'])
    it = it.filter(pc.equal(it['i_manager_id'], 7).fill_null(False))
    pr = oq.read_dim(warehouse, 'promotion', ['p_item_sk'])
    have = set(v for v in pr['p_item_sk'].to_pylist() if v is not None)
    assert got['i_item_sk'].to_pylist() == sorted(v for v in it['i_item_sk'].to_pylist() if v not in have)


def test_exists_and_not_in_subqueries(gpu_engine, warehouse):
    sql = ("select c_customer_sk, c_current_addr_sk from customer where exists (select 1 from store_sales where ss_customer_sk = c_customer_sk "
           "and ss_quantity > 95) and c_current_addr_sk not in (select ca_address_sk from customer_address where ca_zip > '5') "
           "order by c_customer_sk limit 100000")
    got = _Sess(gpu_engine).sql(sql).toArrow()
    cu = oq.read_dim(warehouse, 'customer', ['c_customer_sk', 'c_current_addr_sk'])
    ss = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_customer_sk', 'ss_quantity'])
    buyers = set(ss.filter(pc.greater(ss['ss_quantity'], 95).fill_null(False))['ss_customer_sk'].to_pylist()) - {None}
    ca = oq.read_dim(warehouse, 'customer_address', ['ca_address_sk', 'ca_zip'])
    far = set(ca.filter(pc.greater(ca['ca_zip'], '5').fill_null(False))['ca_address_sk'].to_pylist())
    exp = sorted((c, a) for c, a in zip(cu['c_customer_sk'].to_pylist(), cu['c_current_addr_sk'].to_pylist())
                 if c in buyers and a is not None and a not in far)
    assert _rows(got, ['c_customer_sk', 'c_current_addr_sk']) == exp and len(exp) > 0


def test_not_exists_over_a_fact_scan(gpu_engine, warehouse):
    """Anti and semi joins above the fact table's star join (two subqueries over the same dimension), then an aggregate."""
    base = ("select d_year, count(*) n, sum(ss_quantity) q from store_sales, date_dim where ss_sold_date_sk = d_date_sk and d_moy = 1 "
            "and not exists (select 1 from item where i_item_sk = ss_item_sk and i_manager_id < 50)%s group by d_year order by d_year")
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_moy'])
    dd = dd.filter(pc.equal(dd['d_moy'], 1).fill_null(False))
    ss = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_item_sk', 'ss_quantity'], dd['d_date_sk'].to_pylist())
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_manager_id', 'i_category_id'])
    low = it.filter(pc.less(it['i_manager_id'], 50).fill_null(False)).select(['i_item_sk'])
    cat = it.filter(pc.less_equal(it['i_category_id'], 5).fill_null(False)).select(['i_item_sk'])
    j = ss.join(dd, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='inner').join(low, keys='ss_item_sk', right_keys='i_item_sk', join_type='left anti')
    for extra, jj in (('', j), (' and ss_item_sk in (select i_item_sk from item i2 where i2.i_category_id <= 5)',
                            j.join(cat, keys='ss_item_sk', right_keys='i_item_sk', join_type='left semi'))):
        got = _Sess(gpu_engine).sql(base % extra).toArrow()
        g = jj.group_by('d_year').aggregate([([], 'count_all'), ('ss_quantity', 'sum')]).sort_by('d_year')
        assert _rows(got, ['d_year', 'n', 'q']) == [(r['d_year'], r['count_all'], r['ss_quantity_sum']) for r in g.to_pylist()]
        assert got.num_rows > 0


def test_select_distinct(gpu_engine, warehouse):
    sql = "select distinct i_category, i_manager_id from item where i_manager_id < 6 order by i_category, i_manager_id"
    got = _Sess(gpu_engine).sql(sql).toArrow()
    it = oq.read_dim(warehouse, 'item', ['i_category', 'i_manager_id'])
    it = it.filter(pc.less(it['i_manager_id'], 6).fill_null(False))
    exp = sorted(set(zip(it['i_category'].to_pylist(), it['i_manager_id'].to_pylist())), key=_key)
    assert _rows(got, ['i_category', 'i_manager_id']) == exp


def test_inner_join_on_a_repeated_build_key(gpu_engine, warehouse):
    """promotion.p_item_sk is not a key of promotion: an item joins every promotion that names it."""
    sql = "select i_item_sk, p_promo_sk from item, promotion where p_item_sk = i_item_sk and i_manager_id < 60 order by i_item_sk, p_promo_sk"
    got = _Sess(gpu_engine).sql(sql).toArrow()
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_manager_id'])
    it = it.filter(pc.less(it['i_manager_id'], 60).fill_null(False))
    pr = oq.read_dim(warehouse, 'promotion', ['p_promo_sk', 'p_item_sk'])
    j = it.join(pr.filter(pc.is_valid(pr['p_item_sk'])), keys='i_item_sk', right_keys='p_item_sk', join_type='inner')
    exp = sorted(zip(j['i_item_sk'].to_pylist(), j['p_promo_sk'].to_pylist()))
    assert _rows(got, ['i_item_sk', 'p_promo_sk']) == exp and len(exp) > 0


def test_count_distinct(gpu_engine, warehouse):
    """count(DISTINCT x) next to other aggregates: two-level aggregation (group by (keys, x), then count / merge per key);
    NULL x is not counted; also without GROUP BY."""
    sql = ("select i_category, count(distinct i_manager_id) nm, count(*) n, sum(i_current_price) p, min(i_item_sk) lo "
           "from item group by i_category order by i_category")
    got = _Sess(gpu_engine).sql(sql).toArrow()
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category', 'i_manager_id', 'i_current_price'])
    g = it.group_by('i_category').aggregate([('i_manager_id', 'count_distinct'), ([], 'count_all'), ('i_current_price', 'sum'), ('i_item_sk', 'min')])
    exp = sorted(((r['i_category'], r['i_manager_id_count_distinct'], r['count_all'], r['i_current_price_sum'], r['i_item_sk_min'])
                  for r in g.to_pylist()), key=_key)
    assert _rows(got, ['i_category', 'nm', 'n', 'p', 'lo']) == exp
    got = _Sess(gpu_engine).sql("select count(distinct ss_customer_sk) c, count(ss_customer_sk) n from store_sales").toArrow().to_pylist()
    ss = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_customer_sk'])
    assert got == [{'c': pc.count_distinct(ss['ss_customer_sk']).as_py(), 'n': pc.count(ss['ss_customer_sk']).as_py()}]


def _rollup_expected(table, keys, aggs, finish):
    """Rows of GROUP BY ROLLUP(keys) restated with pyarrow: one group_by per level, NULL for the keys a level rolls up."""
    rows = []
    for depth in range(len(keys), -1, -1):
        ks = keys[:depth]
        if ks:
            g = table.group_by(ks).aggregate(aggs)
        else:
            if table.num_rows == 0:
                continue
            g = table.append_column('_all', pa.array([0] * table.num_rows)).group_by(['_all']).aggregate(aggs)
        for r in g.to_pylist():
            rows.append(tuple(r.get(k) if k in ks else None for k in keys) + finish(r))
    return sorted(rows, key=_key)


def test_group_by_rollup(gpu_engine, warehouse):
    """GROUP BY ROLLUP over a dimension (string keys, NULL groups beside rolled-up rows, avg = merged sum / merged count) and over
    the fused fact scan (row counts of the levels only exist on the device); and no row at all when no row qualifies."""
    sql = ("select i_category, i_class, count(*) n, count(i_current_price) c, sum(i_current_price) p, avg(i_current_price) a, "
           "min(i_item_sk) lo, max(i_item_sk) hi from item where i_manager_id < 60 group by rollup(i_category, i_class) "
           "order by i_category, i_class, n, lo limit 100000")
    got = _Sess(gpu_engine).sql(sql).toArrow()
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category', 'i_class', 'i_current_price', 'i_manager_id'])
    it = it.filter(pc.less(it['i_manager_id'], 60).fill_null(False))

    def fin(r):
        s, c = r['i_current_price_sum'], r['i_current_price_count']
        avg = None if not c else (decimal.Decimal(s) / c).quantize(decimal.Decimal('0.000001'), rounding=decimal.ROUND_HALF_UP)
        return (r['count_all'], c, s, avg, r['i_item_sk_min'], r['i_item_sk_max'])
    exp = _rollup_expected(it, ['i_category', 'i_class'],
                           [([], 'count_all'), ('i_current_price', 'count'), ('i_current_price', 'sum'), ('i_item_sk', 'min'), ('i_item_sk', 'max')], fin)
    rows = _rows(got, ['i_category', 'i_class', 'n', 'c', 'p', 'a', 'lo', 'hi'])
    assert sorted(rows, key=_key) == exp
    assert len(exp) > 20 and sum(1 for r in exp if r[0] is None and r[1] is None) >= 1
    assert 'rollup_union' in [o['op'] for o in gpu_engine.last_stats()['ops']]
    # the grand total is the last level: count(*) over everything that passed the filter
    assert max(r[2] for r in rows) == it.num_rows

    sql = ("select d_moy, i_category, sum(ss_quantity) q, count(*) n from store_sales, date_dim, item where ss_sold_date_sk = d_date_sk "
           "and ss_item_sk = i_item_sk and i_manager_id < 30 and d_year = 1998 group by rollup(d_moy, i_category) "
           "order by d_moy, i_category, n, q limit 100000")
    got = _Sess(gpu_engine).sql(sql).toArrow()
    ss = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_item_sk', 'ss_quantity'])
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_moy', 'd_year'])
    dd = dd.filter(pc.equal(dd['d_year'], 1998).fill_null(False)).select(['d_date_sk', 'd_moy'])
    i2 = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category', 'i_manager_id'])
    i2 = i2.filter(pc.less(i2['i_manager_id'], 30).fill_null(False)).select(['i_item_sk', 'i_category'])
    j = ss.join(dd, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='inner').join(i2, keys='ss_item_sk', right_keys='i_item_sk', join_type='inner')
    exp = _rollup_expected(j, ['d_moy', 'i_category'], [('ss_quantity', 'sum'), ([], 'count_all')], lambda r: (r['ss_quantity_sum'], r['count_all']))
    assert sorted(_rows(got, ['d_moy', 'i_category', 'q', 'n']), key=_key) == exp and len(exp) > 5
    assert 'scan_probe' in [o['op'] for o in gpu_engine.last_stats()['ops']]

    got = _Sess(gpu_engine).sql("select i_category, count(*) n from item where i_manager_id < 0 group by rollup(i_category) order by i_category").toArrow()
    assert got.num_rows == 0


def test_window_aggregates_over_partitions(gpu_engine, warehouse):
    """fn(x) OVER (PARTITION BY ...): over the joined rows of a dimension (two partition lists, NULL keys are a partition of their
    own, decimal avg) and over the groups of a star join (NDS query98's revenue ratio: sum(sum(x)) over (partition by i_class))."""
    import math
    sql = ("select i_item_sk, i_class, avg(i_current_price) over (partition by i_class) a, count(*) over (partition by i_class) n, "
           "count(i_current_price) over (partition by i_class) c, max(i_item_sk) over (partition by i_category, i_class) hi, "
           "min(i_current_price) over (partition by i_category, i_class) lo from item where i_manager_id < 70 order by i_item_sk limit 100000")
    got = _Sess(gpu_engine).sql(sql).toArrow().to_pylist()
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category', 'i_class', 'i_current_price', 'i_manager_id'])
    it = it.filter(pc.less(it['i_manager_id'], 70).fill_null(False)).to_pylist()
    by_class, by_cc = {}, {}
    for r in it:
        by_class.setdefault(r['i_class'], []).append(r)
        by_cc.setdefault((r['i_category'], r['i_class']), []).append(r)
    exp = []
    for r in sorted(it, key=lambda r: r['i_item_sk']):
        g = by_class[r['i_class']]
        prices = [x['i_current_price'] for x in g if x['i_current_price'] is not None]
        avg = (sum(prices) / len(prices)).quantize(decimal.Decimal('0.000001'), rounding=decimal.ROUND_HALF_UP) if prices else None
        g2 = by_cc[(r['i_category'], r['i_class'])]
        p2 = [x['i_current_price'] for x in g2 if x['i_current_price'] is not None]
        exp.append({'i_item_sk': r['i_item_sk'], 'i_class': r['i_class'], 'a': avg, 'n': len(g), 'c': len(prices),
                    'hi': max(x['i_item_sk'] for x in g2), 'lo': min(p2) if p2 else None})
    assert got == exp and len(exp) > 50 and any(r['n'] > 1 for r in exp)
    assert [o['op'] for o in gpu_engine.last_stats()['ops']].count('window') == 2

    sql = ("select i_item_id, i_class, sum(ss_ext_sales_price) itemrevenue, "
           "sum(ss_ext_sales_price) * 100 / sum(sum(ss_ext_sales_price)) over (partition by i_class) revenueratio "
           "from store_sales, item, date_dim where ss_item_sk = i_item_sk and i_category in ('Sports', 'Books', 'Home') "
           "and ss_sold_date_sk = d_date_sk and d_year = 1998 group by i_item_id, i_class "
           "order by i_class, i_item_id, itemrevenue limit 100000")
    got = _Sess(gpu_engine).sql(sql).toArrow().to_pylist()
    ss = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_item_sk', 'ss_ext_sales_price'])
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year'])
    dd = dd.filter(pc.equal(dd['d_year'], 1998).fill_null(False)).select(['d_date_sk'])
    i2 = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_item_id', 'i_class', 'i_category'])
    i2 = i2.filter(pc.is_in(i2['i_category'], value_set=pa.array(['Sports', 'Books', 'Home'])).fill_null(False)).select(['i_item_sk', 'i_item_id', 'i_class'])
    j = ss.join(dd, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='left semi').join(i2, keys='ss_item_sk', right_keys='i_item_sk', join_type='inner')
    g = j.group_by(['i_item_id', 'i_class']).aggregate([('ss_ext_sales_price', 'sum')]).to_pylist()
    tot = {}
    for r in g:
        if r['ss_ext_sales_price_sum'] is not None:
            tot[r['i_class']] = tot.get(r['i_class'], 0) + r['ss_ext_sales_price_sum']
    exp = sorted(((r['i_class'], r['i_item_id'], r['ss_ext_sales_price_sum']) for r in g), key=_key)
    assert [(r['i_class'], r['i_item_id'], r['itemrevenue']) for r in got] == exp and len(exp) > 20
    for r in got:
        if r['itemrevenue'] is None or not tot.get(r['i_class']):
            assert r['revenueratio'] is None
        else:
            assert math.isclose(r['revenueratio'], float(r['itemrevenue'] * 100) / float(tot[r['i_class']]), rel_tol=1e-9), r
    ops = [o['op'] for o in gpu_engine.last_stats()['ops']]
    assert 'window' in ops and 'scan_probe' in ops


def test_derived_table_with_window_filter(gpu_engine, warehouse):
    """NDS query53's shape: a star join grouped by (manufacturer, day of week; the small warehouse sells in one quarter only), the manufacturer's average over them as a window
    aggregate, and an outer query over that derived table keeping the quarters that deviate by more than 10 %."""
    sql = ("select * from (select i_manufact_id, sum(ss_sales_price) sum_sales, "
           "avg(sum(ss_sales_price)) over (partition by i_manufact_id) avg_quarterly_sales "
           "from item, store_sales, date_dim where ss_item_sk = i_item_sk and ss_sold_date_sk = d_date_sk "
           "and d_year = 1998 group by i_manufact_id, d_dow) tmp1 "
           "where case when avg_quarterly_sales > 0 then abs(sum_sales - avg_quarterly_sales) / avg_quarterly_sales else null end > 0.1 "
           "order by avg_quarterly_sales, sum_sales, i_manufact_id limit 100")
    got = _Sess(gpu_engine).sql(sql).toArrow()
    ss = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_item_sk', 'ss_sales_price'])
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_dow'])
    dd = dd.filter(pc.equal(dd['d_year'], 1998).fill_null(False)).select(['d_date_sk', 'd_dow'])
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_manufact_id'])
    j = ss.join(dd, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='inner').join(it, keys='ss_item_sk', right_keys='i_item_sk', join_type='inner')
    g = j.group_by(['i_manufact_id', 'd_dow']).aggregate([('ss_sales_price', 'sum')]).to_pylist()
    per = {}
    for r in g:
        per.setdefault(r['i_manufact_id'], []).append(r['ss_sales_price_sum'])
    rows = []
    for r in g:
        sums = [x for x in per[r['i_manufact_id']] if x is not None]
        avg = (sum(sums) / len(sums)).quantize(decimal.Decimal('0.000001'), rounding=decimal.ROUND_HALF_UP) if sums else None
        s = r['ss_sales_price_sum']
        if avg is not None and avg > 0 and s is not None and abs(float(s - avg)) / float(avg) > 0.1:
            rows.append((avg, s, r['i_manufact_id']))
    exp = sorted(rows, key=_key)[:100]
    assert _rows(got, ['avg_quarterly_sales', 'sum_sales', 'i_manufact_id']) == exp and len(exp) > 10
    assert got.schema.names == ['i_manufact_id', 'sum_sales', 'avg_quarterly_sales']


def test_ranking_windows_and_nested_derived_tables(gpu_engine, warehouse):
    """rank() / row_number() / dense_rank() OVER (PARTITION BY ... ORDER BY ...) over groups with many ties, filtered by an outer query
    on the rank; the same over a star join; a CTE referenced once and nested derived tables."""
    sql = ("select * from (select i_category, i_class, sum(i_current_price) s, count(*) n, "
           "rank() over (partition by i_category order by count(*) desc) rk, "
           "row_number() over (partition by i_category order by count(*) desc, i_class) rn, "
           "dense_rank() over (partition by i_category order by count(*) desc) dr "
           "from item group by i_category, i_class) t where rk <= 5 order by i_category, rn")
    got = _Sess(gpu_engine).sql(sql).toArrow()
    it = oq.read_dim(warehouse, 'item', ['i_category', 'i_class', 'i_current_price'])
    g = it.group_by(['i_category', 'i_class']).aggregate([('i_current_price', 'sum'), ([], 'count_all')]).to_pylist()
    cats = {}
    for r in g:
        cats.setdefault(r['i_category'], []).append(r)
    exp = []
    for cat, rows in cats.items():
        ns = sorted({r['count_all'] for r in rows}, reverse=True)
        order = sorted(rows, key=lambda r: (-r['count_all'], r['i_class'] is not None, r['i_class'] or ''))
        for rn, r in enumerate(order, 1):
            rk = 1 + sum(1 for x in rows if x['count_all'] > r['count_all'])
            if rk <= 5:
                exp.append((cat, r['i_class'], r['i_current_price_sum'], r['count_all'], rk, rn, 1 + ns.index(r['count_all'])))
    exp.sort(key=lambda r: ((r[0] is not None, r[0] or ''), r[5]))
    assert _rows(got, ['i_category', 'i_class', 's', 'n', 'rk', 'rn', 'dr']) == exp
    assert len(exp) > 20 and any(r[4] != r[5] for r in exp) and any(r[4] != r[6] for r in exp)       # ties: rank, row_number and dense_rank differ
    assert [o['op'] for o in gpu_engine.last_stats()['ops']].count('rank') == 2

    sql = ("select * from (select i_category, d_dow, sum(ss_quantity) q, rank() over (partition by i_category order by sum(ss_quantity) desc) rk "
           "from store_sales, item, date_dim where ss_item_sk = i_item_sk and ss_sold_date_sk = d_date_sk and d_year = 1998 "
           "group by i_category, d_dow) t where rk <= 2 order by i_category, rk, d_dow")
    got = _Sess(gpu_engine).sql(sql).toArrow()
    ss = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', ['ss_sold_date_sk', 'ss_item_sk', 'ss_quantity'])
    dd = oq.read_dim(warehouse, 'date_dim', ['d_date_sk', 'd_year', 'd_dow'])
    dd = dd.filter(pc.equal(dd['d_year'], 1998).fill_null(False)).select(['d_date_sk', 'd_dow'])
    i2 = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category'])
    j = ss.join(dd, keys='ss_sold_date_sk', right_keys='d_date_sk', join_type='inner').join(i2, keys='ss_item_sk', right_keys='i_item_sk', join_type='inner')
    g = j.group_by(['i_category', 'd_dow']).aggregate([('ss_quantity', 'sum')]).to_pylist()
    exp = []
    for r in g:
        q = r['ss_quantity_sum']
        # ORDER BY q DESC: NULL sums last; rank = 1 + the groups of the category that sort before
        before = sum(1 for x in g if x['i_category'] == r['i_category'] and x['ss_quantity_sum'] is not None and (q is None or x['ss_quantity_sum'] > q))
        if before + 1 <= 2:
            exp.append((r['i_category'], r['d_dow'], q, before + 1))
    exp.sort(key=lambda r: ((r[0] is not None, r[0] or ''), r[3], r[1]))
    assert _rows(got, ['i_category', 'd_dow', 'q', 'rk']) == exp and len(exp) > 5

    a = _Sess(gpu_engine).sql("with t as (select i_item_sk k, i_class c from item where i_manager_id < 30) "
                              "select k from (select k, c from t where c > 'i_class_03') v where k < 5000 order by k").toArrow()
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_class', 'i_manager_id']).to_pylist()
    want = sorted(r['i_item_sk'] for r in it if r['i_manager_id'] is not None and r['i_manager_id'] < 30 and r['i_class'] is not None
                  and r['i_class'] > 'i_class_03' and r['i_item_sk'] < 5000)
    assert a['k'].to_pylist() == want and len(want) > 3


def test_like_prefix_and_coalesce(gpu_engine, warehouse):
    sql = ("select i_item_sk, coalesce(i_current_price, 0) p, coalesce(i_manager_id, i_manufact_id, -1) m from item "
           "where i_category like 'S%' and i_category not like 'Sports' order by i_item_sk")
    got = _Sess(gpu_engine).sql(sql).toArrow().to_pylist()
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_category', 'i_current_price', 'i_manager_id', 'i_manufact_id']).to_pylist()
    exp = []
    for r in sorted(it, key=lambda r: r['i_item_sk']):
        if r['i_category'] is not None and r['i_category'].startswith('S') and r['i_category'] != 'Sports':
            m = r['i_manager_id'] if r['i_manager_id'] is not None else (r['i_manufact_id'] if r['i_manufact_id'] is not None else -1)
            exp.append({'i_item_sk': r['i_item_sk'], 'p': r['i_current_price'] if r['i_current_price'] is not None else decimal.Decimal('0.00'), 'm': m})
    assert got == exp and len(exp) > 5 and any(r['i_current_price'] is None for r in it)


def test_order_by_outside_the_select_list(gpu_engine, warehouse):
    """ORDER BY a column / an aggregate / an expression over a select alias that is not an output column: computed beside the
    outputs, sorted on, dropped (project -> sort -> project)."""
    it = oq.read_dim(warehouse, 'item', ['i_item_sk', 'i_class', 'i_category', 'i_current_price', 'i_manager_id']).to_pylist()
    got = _Sess(gpu_engine).sql("select i_item_sk k, i_current_price p from item where i_manager_id < 40 "
                                "order by i_class desc, k limit 50").toArrow()
    rows = [r for r in it if r['i_manager_id'] is not None and r['i_manager_id'] < 40]
    rows.sort(key=lambda r: r['i_item_sk'])
    rows.sort(key=lambda r: (r['i_class'] is not None, (r['i_class'] or '').encode()), reverse=True)      # DESC: NULLS LAST, stable
    assert got.schema.names == ['k', 'p']
    assert _rows(got, ['k', 'p']) == [(r['i_item_sk'], r['i_current_price']) for r in rows[:50]]

    got = _Sess(gpu_engine).sql("select i_category c, sum(i_current_price) s from item group by i_category "
                                "order by count(*) desc, case when s > 0 then c end limit 100").toArrow()
    groups = {}
    for r in it:
        g = groups.setdefault(r['i_category'], [0, None])
        g[0] += 1
        if r['i_current_price'] is not None:
            g[1] = r['i_current_price'] if g[1] is None else g[1] + r['i_current_price']
    exp = [(c, s, n) for c, (n, s) in groups.items()]
    exp.sort(key=lambda t: (t[0] is not None, (t[0] or '').encode(), t[1] is not None, t[1] or 0))     # tie-break: the outputs, ascending, NULLS FIRST
    exp.sort(key=lambda t: ((t[0] if (t[1] is not None and t[1] > 0) else None) is not None,
                            ((t[0] if (t[1] is not None and t[1] > 0) else None) or '').encode()))
    exp.sort(key=lambda t: -t[2])
    assert got.schema.names == ['c', 's']
    assert _rows(got, ['c', 's']) == [(c, s) for c, s, _ in exp]
