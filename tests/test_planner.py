"""CPU: SQL front-end and planner (host logic standing in for Catalyst on this path)."""
import json

import pytest

from spark_rapids_benchmarks_b200 import planner, sql, streams


def _walk(n, f):
    f(n)
    for k in ('child', 'probe', 'build'):
        if k in n:
            _walk(n[k], f)


@pytest.mark.parametrize('query', streams.SUBSET)
def test_plans_are_star_joins_with_pruning(query):
    p = planner.plan_sql(streams.render(query))
    ops, prune, scans = [], [], []
    _walk(p['root'], lambda n: (ops.append(n['op']), prune.append(n.get('prune_partitions', False)),
                                scans.append(n.get('table')) if n['op'] == 'scan' else None))
    assert ops[0] == 'sort' and 'aggregate' in ops and 'join' in ops
    assert any(prune), 'the date join must carry dynamic partition pruning'
    assert 'store_sales' in scans
    json.dumps(p)     # serialisable


def test_output_names_and_types():
    p = planner.plan_sql(streams.render('query42'))
    assert [o['name'] for o in p['output']] == ['d_year', 'i_category_id', 'i_category', 'sum(ss_ext_sales_price)']
    assert p['output'][-1]['type'] == 'decimal(17,2)'
    p7 = planner.plan_sql(streams.render('query7'))
    assert [o['type'] for o in p7['output']] == ['string', 'double', 'decimal(11,6)', 'decimal(11,6)', 'decimal(11,6)']


def test_snowflake_join_goes_through_customer():
    p = planner.plan_sql(streams.render('query19'))
    found = []
    _walk(p['root'], lambda n: found.append((n.get('probe_key'), n.get('build_key'))) if n['op'] == 'join' else None)
    assert ('c_current_addr_sk', 'ca_address_sk') in found and ('ss_customer_sk', 'c_customer_sk') in found


def test_distributed_plan_splits_at_the_aggregate():
    p = planner.plan_sql(streams.render('query7'), distributed=True)
    assert p['partial']['op'] == 'aggregate'
    fns = [a['fn'] for a in p['partial']['aggs']]
    assert 'avg' not in fns and 'sum' in fns and 'count' in fns
    scans = []
    _walk(p['root'], lambda n: scans.append(n['table']) if n['op'] == 'scan' else None)
    assert scans == ['__partial__']


@pytest.mark.parametrize('text', [
    'select * from item',
    'select * from (select i_item_sk from item) t, store',                     # derived table joined with another table
    'select x from (select i_item_sk from item) t',                            # unknown column of the derived table
    'select count(*) from (select i_item_sk from item) t',                     # aggregation over a derived table
    'select i_item_sk from item where i_item_sk in (select ss_item_sk from store_sales, date_dim where ss_sold_date_sk = d_date_sk)',   # multi-table subquery
    'select i_item_sk from item where i_item_sk not in (select ss_customer_sk from store_sales)',   # NOT IN over a nullable column
    'select i_item_sk from item where exists (select 1 from item where i_manager_id = 3)',          # uncorrelated / same table
    'select i_item_sk from item where i_item_sk = (select max(ss_item_sk) from store_sales)',      # scalar subquery
    'select i_item_sk, sum(i_item_sk) over (partition by i_brand order by i_item_sk) from item',     # running total
    'select i_item_sk, rank() over (partition by i_brand) from item',                               # ranking without ORDER BY
    'select i_item_sk, sum(i_item_sk) over (partition by i_brand rows between 1 preceding and current row) from item',
    'select i_brand, count(distinct i_item_sk) from item group by rollup(i_brand)',
    'with x as (select i_item_sk from item) select a.i_item_sk from x a, x b',    # a CTE referenced twice
    'select i_item_sk from item, store',                                     # cross join
    'select ss_item_sk from store_sales, store_returns where ss_item_sk = sr_item_sk',   # fact-fact join
    'select distinct i_brand, count(*) from item group by i_brand',
    'select i_brand from item having count(*) > 1 and i_brand > \'a\'',       # HAVING over a non-grouped column
    'select i_brand from item left join promotion on p_item_sk = i_item_sk left join store on s_store_sk = p_promo_sk',   # chained outer joins
    'select i_brand from item where i_brand like \'%a\'',                      # only a literal prefix + % is supported
    'select i_brand from item where i_brand like \'a_c%\'',
    'select i_brand, sum(distinct i_item_sk) from item group by i_brand',
    'select i_brand, count(distinct i_item_sk), count(distinct i_class) from item group by i_brand',
    'select nope from item',
])
def test_unsupported_sql_raises(text):
    with pytest.raises(planner.PlanError):
        planner.plan_sql(text)


def test_parser_details():
    q = sql.parse("select a.x y, sum(b) from t a where x between 1 and 2 and s not in ('p','q') or z is not null "
                  "order by y desc nulls first, 2 limit 5")
    assert q['limit'] == 5 and q['from'] == [('t', 'a')] and q['select'][0][1] == 'y'
    assert q['order'][0][1] is True and q['order'][0][2] is True
    assert q['where'][0] == 'bin' and q['where'][1] == 'or'


def test_collect_matches_to_pylist():
    """DataFrame.collect() converts columns through numpy / raw buffers; it must return exactly what pyarrow's
    to_pylist() returns (ints, doubles, decimals, dates, strings, nulls, sliced and chunked arrays)."""
    import datetime
    import decimal
    import numpy as np
    import pyarrow as pa
    from spark_rapids_benchmarks_b200.session import DataFrame, _column_to_pylist
    rng = np.random.default_rng(3)
    n = 257
    mask = rng.random(n) < 0.2
    cols = {
        'i32': pa.array(rng.integers(-2**31, 2**31, n).astype(np.int32), mask=mask),
        'i64': pa.array(rng.integers(-2**62, 2**62, n), mask=mask),
        'i64nn': pa.array(rng.integers(-2**62, 2**62, n)),
        'f64': pa.array(rng.normal(size=n), mask=mask),
        'dec': pa.array([None if m else decimal.Decimal(int(v)).scaleb(-2) for v, m in zip(rng.integers(-10**15, 10**15, n), mask)], type=pa.decimal128(17, 2)),
        'dec0': pa.array([decimal.Decimal(int(v)) for v in rng.integers(-10**6, 10**6, n)], type=pa.decimal128(7, 0)),
        'str': pa.array([None if m else ('s%d-é中' % v if v % 3 else '') for v, m in zip(rng.integers(0, 1000, n), mask)], type=pa.string()),
        'date': pa.array([None if m else datetime.date(1990, 1, 1) + datetime.timedelta(days=int(v)) for v, m in zip(rng.integers(0, 20000, n), mask)], type=pa.date32()),
        'bool': pa.array(rng.random(n) < 0.5, mask=mask),      # falls back to to_pylist
    }
    t = pa.table(cols)
    for variant in (t, t.slice(13, 100), pa.concat_tables([t.slice(0, 50), t.slice(200, 57)]), t.slice(0, 0)):
        for name in variant.schema.names:
            got = _column_to_pylist(variant.column(name))
            exp = variant.column(name).to_pylist()
            assert got == exp, name
            assert [type(x) for x in got] == [type(x) for x in exp], name
        rows = DataFrame(variant).collect()
        assert len(rows) == variant.num_rows
        if variant.num_rows:
            assert tuple(rows[0]) == tuple(variant.column(c)[0].as_py() for c in variant.schema.names)


def test_distributed_plan_has_partial_merge_and_fused_forms():
    """plan_sql(distributed=True): 'partial' (scan -> joins -> partial aggregate), 'root' (merge over __partial__) and
    'fused' (the same merge tree with the __partial__ scan replaced by exchange{partial}); AVG travels as SUM + COUNT."""
    from spark_rapids_benchmarks_b200 import planner, streams

    def ops(node, out):
        out.append(node['op'])
        for k in ('child', 'probe', 'build'):
            if k in node:
                ops(node[k], out)
        return out
    for q in streams.SUBSET:
        p = planner.plan_sql(streams.render(q), distributed=True)
        assert set(p) >= {'root', 'partial', 'fused', 'output'}
        assert p['partial']['op'] == 'aggregate'
        root_ops, fused_ops = ops(p['root'], []), ops(p['fused'], [])
        assert 'exchange' not in root_ops and root_ops.count('scan') == 1
        i = fused_ops.index('exchange')
        assert fused_ops[:i] == root_ops[:root_ops.index('scan')]              # same merge tree above the exchange
        assert fused_ops[i + 1:] == ops(p['partial'], [])                      # the partial plan below it
        merge = p['root']
        while merge['op'] != 'aggregate':
            merge = merge['child']
        assert {a['fn'] for a in merge['aggs']} <= {'sum', 'min', 'max'}       # counts are summed when merging
        assert merge['child'] == {'op': 'scan', 'table': '__partial__',
                                  'columns': [k['name'] for k in merge['keys']] + [a['name'] for a in merge['aggs']]}
    single = planner.plan_sql(streams.render('query3'))
    assert 'partial' not in single and 'fused' not in single


def test_explicit_inner_join_syntax_plans_like_comma_joins():
    """`a JOIN b ON a.x = b.y` is the comma join with the ON condition as one more WHERE conjunct: same plan. Outer joins raise."""
    import pytest
    from spark_rapids_benchmarks_b200 import planner
    comma = '''select d_year, i_brand_id, sum(ss_ext_sales_price) s from date_dim, store_sales, item
               where d_date_sk = ss_sold_date_sk and ss_item_sk = i_item_sk and i_manufact_id = 7 and d_moy = 11
               group by d_year, i_brand_id order by d_year, s desc, i_brand_id limit 100'''
    explicit = '''select d_year, i_brand_id, sum(ss_ext_sales_price) s
                  from date_dim join store_sales on d_date_sk = ss_sold_date_sk inner join item on ss_item_sk = i_item_sk
                  where i_manufact_id = 7 and d_moy = 11
                  group by d_year, i_brand_id order by d_year, s desc, i_brand_id limit 100'''
    assert planner.plan_sql(explicit) == planner.plan_sql(comma)
    mixed = comma.replace('from date_dim, store_sales, item\n               where d_date_sk = ss_sold_date_sk and',
                          'from date_dim join store_sales on d_date_sk = ss_sold_date_sk, item\n               where')
    assert planner.plan_sql(mixed) == planner.plan_sql(comma)
    for kind in ('right', 'full outer'):
        with pytest.raises(planner.PlanError, match='OUTER JOIN is not supported'):
            planner.plan_sql(explicit.replace('inner join item', kind + ' join item'))
    # LEFT [OUTER] JOIN plans a left join above the inner chain; the WHERE condition on the joined table runs after it
    for kind in ('left', 'left outer'):
        p = planner.plan_sql(explicit.replace('inner join item', kind + ' join item'))
        node = p['root']
        while node['op'] != 'aggregate':
            node = node['child']
        flt = node['child']
        assert flt['op'] == 'filter' and flt['child']['op'] == 'join' and flt['child']['type'] == 'left'
        assert flt['child']['build'] == {'op': 'scan', 'table': 'item', 'columns': ['i_item_sk', 'i_brand_id', 'i_manufact_id']}


def test_subqueries_distinct_having_plan_shapes():
    from spark_rapids_benchmarks_b200 import planner

    def find(node, op):
        out = []
        if node.get('op') == op:
            out.append(node)
        for k in ('child', 'probe', 'build'):
            if k in node:
                out += find(node[k], op)
        return out
    p = planner.plan_sql('select c_customer_sk from customer where exists (select 1 from store_sales ss where ss.ss_customer_sk = c_customer_sk '
                         'and ss_quantity > 90) and c_current_addr_sk not in (select ca_address_sk from customer_address) order by c_customer_sk')
    joins = find(p['root'], 'join')
    assert [(j['type'], j['probe_key'], j['build_key']) for j in joins] == [('anti', 'c_current_addr_sk', 'ca_address_sk'),
                                                                             ('semi', 'c_customer_sk', 'ss_customer_sk')]
    assert joins[0]['probe']['op'] == 'filter' and joins[0]['probe']['cond']['fn'] == 'isnotnull'      # NOT IN drops NULL outer values
    p = planner.plan_sql('select distinct i_category, i_class from item')
    agg = find(p['root'], 'aggregate')[0]
    assert [k['name'] for k in agg['keys']] == ['i_category', 'i_class'] and agg['aggs'] == []
    p = planner.plan_sql('select i_brand, sum(i_current_price) s from item group by i_brand having sum(i_current_price) > 10 and count(*) > 1',
                         distributed=True, sharded={'item'})
    flt = find(p['root'], 'filter')[0]
    assert flt['child']['op'] == 'aggregate' and flt['child']['child']['table'] == '__partial__'       # HAVING runs over the merged groups
    # a join on a column that is not the build table's primary key chains repeated keys
    p = planner.plan_sql('select i_brand, count(*) c from item, promotion where p_item_sk = i_item_sk group by i_brand')
    assert find(p['root'], 'join')[0].get('unique_build') is False


def test_table_aliases_and_qualified_columns_plan_identically():
    from spark_rapids_benchmarks_b200 import planner
    plain = '''select d_year, i_brand_id, sum(ss_ext_sales_price) s from date_dim, store_sales, item
               where d_date_sk = ss_sold_date_sk and ss_item_sk = i_item_sk and i_manufact_id = 7 and d_moy = 11
               group by d_year, i_brand_id order by d_year, s desc, i_brand_id limit 100'''
    aliased = '''select dt.d_year, it.i_brand_id, sum(ss.ss_ext_sales_price) s from date_dim dt, store_sales as ss, item it
                 where dt.d_date_sk = ss.ss_sold_date_sk and ss.ss_item_sk = it.i_item_sk and it.i_manufact_id = 7 and dt.d_moy = 11
                 group by dt.d_year, it.i_brand_id order by dt.d_year, s desc, it.i_brand_id limit 100'''
    assert planner.plan_sql(aliased) == planner.plan_sql(plain)


def test_dataframe_writer_formats_and_modes(tmp_path):
    """The slice of DataFrameWriter the harness uses: parquet and csv outputs, overwrite vs the default errorifexists,
    anything else raises (reference nds_power.py:133-135 writes `df.write.format(f).mode('overwrite').save(path)`)."""
    import decimal
    import os
    import pyarrow as pa
    import pyarrow.parquet as pq
    import pytest
    from spark_rapids_benchmarks_b200.engine import EngineError
    from spark_rapids_benchmarks_b200.session import DataFrame
    t = pa.table({'k': pa.array([1, 2, None]), 'd': pa.array([decimal.Decimal('1.50'), None, decimal.Decimal('-2.25')], type=pa.decimal128(7, 2)),
                  's': pa.array(['a', None, 'c,d'])})
    df = DataFrame(t)
    p = str(tmp_path / 'out_parquet')
    df.write.format('parquet').mode('overwrite').save(p)
    assert os.path.exists(os.path.join(p, '_SUCCESS'))
    assert pq.ParquetDataset(p).read().equals(t)
    df.toDF('x', 'y', 'z').write.format('parquet').mode('overwrite').save(p)           # overwrite replaces the folder
    assert pq.ParquetDataset(p).read().schema.names == ['x', 'y', 'z']
    with pytest.raises(EngineError):
        df.write.format('parquet').save(p)                                              # default mode: errorifexists
    c = str(tmp_path / 'out_csv')
    df.write.format('csv').mode('overwrite').save(c)
    text = open([os.path.join(c, f) for f in os.listdir(c) if f.endswith('.csv')][0]).read()
    assert text.splitlines()[0].replace('"', '') == 'k,d,s' and '"c,d"' in text
    with pytest.raises(EngineError):
        df.write.format('orc').mode('overwrite').save(str(tmp_path / 'out_orc'))


def test_distributed_plan_of_a_replicated_root_is_not_split():
    """A query whose root table every rank holds in full (a dimension table, or a fact table that was stored
    unpartitioned) must not be planned as partial -> exchange -> merge: each rank's 'partial' would be the whole
    answer and the merge would multiply every SUM / COUNT by the world size."""
    from spark_rapids_benchmarks_b200 import planner, streams
    p = planner.plan_sql('select i_category, count(*) cnt, sum(i_manager_id) s from item group by i_category order by i_category', distributed=True)
    assert p.get('replicated') is True and 'partial' not in p and 'fused' not in p
    # a fact table the engines do not hold in shards (sharded = what Session learned from table_info)
    sql = streams.render('query3')
    assert 'partial' in planner.plan_sql(sql, distributed=True)
    assert 'partial' in planner.plan_sql(sql, distributed=True, sharded={'store_sales'})
    q = planner.plan_sql(sql, distributed=True, sharded=set())
    assert q.get('replicated') is True and 'partial' not in q
    assert q['root'] == planner.plan_sql(sql)['root']


def test_merge_aggregates_carry_the_declared_type():
    """sum(decimal(7,2)) is decimal(17,2) on one GPU; the merge of partial sums must declare the same type instead of
    widening it to decimal(27,2) (the Arrow / Parquet result schema must not depend on the GPU count)."""
    from spark_rapids_benchmarks_b200 import planner, streams
    p = planner.plan_sql(streams.render('query7'), distributed=True)
    merge = p['root']
    while merge['op'] != 'aggregate':
        merge = merge['child']
    types = {a['name']: a['type'] for a in merge['aggs']}
    assert sorted(set(types.values())) == ['decimal(17,2)', 'int64']
    assert all('type' not in a for a in p['partial']['aggs'])


def test_rollup_plan_shape():
    """GROUP BY ROLLUP(a, b): one `rollup` node whose first level is the finest aggregate over the child and whose further levels
    aggregate the level before them in merge form (SUM of counts ...); the grand total groups by a constant."""
    p = planner.plan_sql('select i_category, i_class, count(*) n, avg(i_current_price) a, min(i_item_sk) lo from item '
                         'group by rollup(i_category, i_class) order by i_category, i_class limit 10')
    node = p['root']
    while node['op'] != 'rollup':
        node = node['child']
    lv = node['levels']
    assert [[k['name'] for k in l['keys']] for l in lv] == [['i_category', 'i_class'], ['i_category'], ['_rollup_all']]
    assert {a['fn'] for a in lv[0]['aggs']} == {'count_star', 'sum', 'count', 'min'}
    assert {a['fn'] for a in lv[1]['aggs']} == {'sum', 'min'} and all(a['expr'] == {'col': a['name']} for a in lv[2]['aggs'])
    assert node['child']['op'] == 'scan'
    # every output column breaks ties: a rolled-up row has the same keys as the NULL group of the key it rolled up
    sort = p['root']
    assert sort['op'] == 'sort' and len(sort['keys']) == 5
    with pytest.raises(planner.PlanError):
        planner.plan_sql('select i_category, count(*) from store_sales, item where ss_item_sk = i_item_sk group by rollup(i_category)', distributed=True)


def test_window_plan_shape():
    """fn(x) OVER (PARTITION BY k): one `window` node per partition list above the aggregate (sum(sum(x)) over ...: the inner sum is
    an ordinary aggregate) or above the joins; ORDER BY / frames, rank() and an empty OVER () raise."""
    p = planner.plan_sql('select i_class, i_brand, sum(i_current_price) s, sum(sum(i_current_price)) over (partition by i_class) t, '
                         'count(*) over (partition by i_class) n, max(sum(i_item_sk)) over (partition by i_brand) m '
                         'from item group by i_class, i_brand order by i_class, i_brand limit 10')
    node = p['root']
    seen = []
    while 'child' in node:
        seen.append(node['op'])
        if node['op'] == 'window':
            assert node['keys'][0]['expr'] in ({'col': 'i_class'}, {'col': 'i_brand'})
        node = node['child']
    assert seen == ['sort', 'project', 'window', 'window', 'aggregate']
    p = planner.plan_sql('select i_item_sk, avg(i_current_price) over (partition by i_class) a from item where i_manager_id = 1')
    w = p['root']['child']
    assert w['op'] == 'window' and {a['fn'] for a in w['aggs']} == {'sum', 'count'} and w['child']['op'] == 'filter'
    for bad in ('select i_item_sk, sum(i_item_sk) over () from item',
                'select i_item_sk, sum(i_item_sk) over (partition by i_class order by i_item_sk) from item',
                'select i_item_sk, ntile(4) over (partition by i_class order by i_item_sk) from item',
                'select i_item_sk from item where sum(i_item_sk) over (partition by i_class) > 3'):
        with pytest.raises(planner.PlanError):
            planner.plan_sql(bad)


def test_derived_table_plan_shape():
    """select ... from (subquery) alias where ... order by ... limit: the subquery's plan with a filter / project / sort of the outer
    query on top (NDS query53 / 63 / 89); CASE ... ELSE NULL takes the type of its branch; abs(x) = if(x < 0, -x, x)."""
    p = planner.plan_sql('select * from (select i_manufact_id, sum(i_current_price) s, avg(sum(i_current_price)) over (partition by i_manufact_id) a '
                         'from item group by i_manufact_id, i_class) t where case when a > 0 then abs(s - a) / a else null end > 0.1 '
                         'order by a, s, i_manufact_id limit 100')
    ops = []
    node = p['root']
    while 'child' in node:
        ops.append(node['op'])
        node = node['child']
    assert ops == ['sort', 'project', 'filter', 'project', 'window', 'aggregate'] and node['op'] == 'scan'
    assert [o['name'] for o in p['output']] == ['i_manufact_id', 's', 'a']
    cond = p['root']['child']['child']['cond']
    case = cond['args'][0]
    assert cond['fn'] == 'gt' and case['fn'] == 'if' and case['args'][2] == {'lit': None, 'type': 'double'}
    assert case['args'][1]['fn'] == 'div' and case['args'][1]['args'][0]['fn'] == 'if'      # abs()
    p = planner.plan_sql('select t.i_item_sk k, i_class from (select i_item_sk, i_class from item where i_manager_id = 1) as t order by k')
    assert [o['name'] for o in p['output']] == ['k', 'i_class']
    # a CTE that is referenced once is a derived table; derived tables nest
    a = planner.plan_sql('with t as (select i_item_sk, i_class from item where i_manager_id = 1) select t.i_item_sk k, i_class from t order by k')
    assert a == p
    n = planner.plan_sql('select k from (select i_item_sk k, c from (select i_item_sk, i_class c from item) u where c > \'m\') v where k < 10 order by k')
    ops = []
    node = n['root']
    while 'child' in node:
        ops.append(node['op'])
        node = node['child']
    assert ops == ['sort', 'project', 'filter', 'project', 'filter', 'project'] and node['op'] == 'scan'


def test_ranking_window_plan_shape():
    """rank() / row_number() / dense_rank() OVER (PARTITION BY p ORDER BY o): a `rank` node whose keys are the partition keys followed
    by the order keys; functions over the same window share it; in a derived table the outer query can filter on the rank."""
    p = planner.plan_sql('select * from (select i_category, i_class, sum(i_current_price) s, '
                         'rank() over (partition by i_category order by sum(i_current_price) desc) rk, '
                         'row_number() over (partition by i_category order by sum(i_current_price) desc) rn, '
                         'dense_rank() over (order by i_class) dr from item group by i_category, i_class) t where rk <= 3 '
                         'order by i_category, rk, i_class')
    node = p['root']
    ranks = []
    while 'child' in node:
        if node['op'] == 'rank':
            ranks.append(node)
        node = node['child']
    assert len(ranks) == 2
    by_np = {r['npart']: r for r in ranks}
    assert [f['fn'] for f in by_np[1]['funcs']] == ['rank', 'row_number']
    assert by_np[1]['keys'] == [{'expr': {'col': 'i_category'}, 'desc': False}, {'expr': {'col': '_agg0'}, 'desc': True}]
    assert by_np[0]['keys'] == [{'expr': {'col': 'i_class'}, 'desc': False}] and by_np[0]['funcs'][0]['fn'] == 'dense_rank'
    assert [o['name'] for o in p['output']] == ['i_category', 'i_class', 's', 'rk', 'rn', 'dr']


def test_query98_template_plans():
    """query98 of the NDS templates: `cast('..' as date) + interval 30 days` folds to a date constant; the window sits between
    the aggregate and the projection; no LIMIT."""
    from spark_rapids_benchmarks_b200 import streams
    p = planner.plan_sql(streams.render('query98'))
    assert p['root']['op'] == 'sort' and p['root']['limit'] == -1
    ops = []
    node = p['root']
    while 'child' in node:
        ops.append(node['op'])
        node = node['child']
    assert ops == ['sort', 'project', 'window', 'aggregate'] and node['op'] == 'join'
    dates = node['build']['cond']['args']
    assert dates[0]['args'][1] == {'lit': 10231, 'type': 'date'} and dates[1]['args'][1] == {'lit': 10261, 'type': 'date'}
    assert [o['type'] for o in p['output']][-2:] == ['decimal(17,2)', 'double']
    with pytest.raises(planner.PlanError):
        planner.plan_sql('select d_date_sk from date_dim where d_date + interval 3 days > cast(\'1998-01-01\' as date)')


def test_like_prefix_and_coalesce():
    p = planner.plan_sql("select i_item_sk, coalesce(i_current_price, 0) p from item where i_brand like 'amalgam%' and i_class not like 'x' order by i_item_sk")
    f = p['root']['child']['child']
    assert f['op'] == 'filter'
    a, b = f['cond']['args']
    assert a == {'fn': 'eq', 'args': [{'fn': 'substr', 'args': [{'col': 'i_brand'}], 'start': 1, 'len': 7}, {'lit': 'amalgam', 'type': 'string'}]}
    assert b == {'fn': 'not', 'args': [{'fn': 'eq', 'args': [{'col': 'i_class'}, {'lit': 'x', 'type': 'string'}]}]}
    e = p['root']['child']['exprs'][1]['expr']
    assert e['fn'] == 'if' and e['args'][0] == {'fn': 'isnotnull', 'args': [{'col': 'i_current_price'}]} and p['output'][1]['type'] == 'decimal(7,2)'
