"""Catalyst physical-plan JSON -> plan IR (spark_rapids_benchmarks_b200/catalyst.py), pinned against a HAND-WRITTEN
`executedPlan.toJSON` sample of query3 (tests/golden/make_catalyst_fixture.py; unverified against Spark: no JVM on either box).
CPU: the mapping; GPU: the translated plan runs and returns the oracle's rows."""
import json
import os

import pytest

from spark_rapids_benchmarks_b200 import catalyst, planner, streams

FIXTURE = os.path.join(os.path.dirname(__file__), 'golden', 'catalyst_q3_executed_plan.json')
Q3 = {'MANUFACT': 606, 'MONTH': 1}


def _walk(node, out):
    out.append(node)
    for k in ('child', 'probe', 'build'):
        if k in node:
            _walk(node[k], out)
    return out


def test_q3_plan_maps_to_the_star_join_ir():
    ir = catalyst.translate(open(FIXTURE).read())['root']
    nodes = _walk(ir, [])
    assert [n['op'] for n in nodes] == ['sort', 'project', 'aggregate', 'join', 'join', 'scan', 'filter', 'scan', 'filter', 'scan']
    assert ir['limit'] == 100
    assert [(k['expr']['col'], k['desc'], k['nulls_first']) for k in ir['keys']] == [('d_year', False, True), ('sum_agg', True, False), ('brand_id', False, True)]
    assert [(e['name'], e['expr']) for e in nodes[1]['exprs']] == [('d_year', {'col': 'd_year'}), ('brand_id', {'col': 'i_brand_id'}),
                                                                   ('brand', {'col': 'i_brand'}), ('sum_agg', {'col': '_agg0'})]
    agg = nodes[2]
    assert [k['name'] for k in agg['keys']] == ['d_year', 'i_brand', 'i_brand_id']
    assert agg['aggs'] == [{'name': '_agg0', 'fn': 'sum', 'expr': {'col': 'ss_ext_sales_price'}}]      # partial + final pair -> one aggregate
    item_join, date_join = nodes[3], nodes[4]
    assert (item_join['probe_key'], item_join['build_key'], item_join['type']) == ('ss_item_sk', 'i_item_sk', 'inner')
    assert (date_join['probe_key'], date_join['build_key']) == ('ss_sold_date_sk', 'd_date_sk')
    assert date_join.get('prune_partitions') is True and 'prune_partitions' not in item_join       # DynamicPruningExpression on the fact scan
    assert date_join['probe'] == {'op': 'scan', 'table': 'store_sales', 'columns': ['ss_item_sk', 'ss_ext_sales_price', 'ss_sold_date_sk']}
    # same scans, same predicates (up to Catalyst's inferred isnotnull conjuncts) as this repo's own planner
    mine = _walk(planner.plan_sql(streams.render('query3', overrides=Q3))['root'], [])
    assert sorted(n['table'] for n in mine if n['op'] == 'scan') == sorted(n['table'] for n in nodes if n['op'] == 'scan')
    text = json.dumps(ir)
    assert '{"fn": "eq", "args": [{"col": "i_manufact_id"}, {"lit": 606, "type": "int64"}]}' in text
    assert '{"fn": "eq", "args": [{"col": "d_moy"}, {"lit": 1, "type": "int64"}]}' in text


def test_unsupported_plan_nodes_raise():
    arr = json.load(open(FIXTURE))
    arr[0] = dict(arr[0], **{'class': 'org.apache.spark.sql.execution.window.WindowExec'})
    with pytest.raises(catalyst.CatalystError, match='WindowExec'):
        catalyst.translate(arr)
    with pytest.raises(catalyst.CatalystError):
        catalyst.translate('[]')


@pytest.mark.gpu
def test_translated_plan_runs_and_matches_the_oracle(gpu_engine, warehouse):
    import pyarrow as pa
    from oracle import nds_queries as oq
    sql = streams.render('query3', overrides=Q3)
    exp = oq.run_query('query3', sql, warehouse)
    got = pa.Table.from_batches([gpu_engine.execute(catalyst.translate(open(FIXTURE).read()))])
    assert got.schema.names == exp.schema.names
    assert got.to_pylist() == exp.to_pylist() and got.num_rows > 0
    assert any(o['op'] == 'scan_probe' for o in gpu_engine.last_stats()['ops'])       # the fused fact scan took it
