"""Writes tests/golden/catalyst_q3_executed_plan.json: a HAND-WRITTEN sample of what
`spark.sql(query3).queryExecution.executedPlan.toJSON` looks like in Spark 3.x (TreeNode.toJSON: pre-order node arrays,
"class" / "num-children" / constructor fields, case-class values with "product-class").  No JVM exists on the build or GPU
box and the reference ships no plan dumps, so this layout is restated from memory and UNVERIFIED against Spark; it pins the
Catalyst -> IR mapping of spark_rapids_benchmarks_b200/catalyst.py until a real dump can replace it.

The statement: streams.render('query3', overrides={'MANUFACT': 606, 'MONTH': 1}).   Run: python tests/golden/make_catalyst_fixture.py
"""
import json
import os

X = 'org.apache.spark.sql.catalyst.expressions.'
P = 'org.apache.spark.sql.execution.'
_ids = {}


def eid(name):
    _ids.setdefault(name, len(_ids) + 1)
    return {'product-class': X + 'ExprId', 'id': _ids[name], 'jvmId': '5c6a9a0e-0000-4000-8000-000000000000'}


def attr(name, dtype, nullable=True, key=None):
    return [{'class': X + 'AttributeReference', 'num-children': 0, 'name': name, 'dataType': dtype, 'nullable': nullable, 'metadata': {},
             'exprId': eid(key or name), 'qualifier': []}]


def lit(value, dtype):
    return [{'class': X + 'Literal', 'num-children': 0, 'value': str(value), 'dataType': dtype}]


def un(cls, child, **kw):
    return [dict({'class': X + cls, 'num-children': 1, 'child': 0}, **kw)] + child


def bi(cls, l, r):
    return [{'class': X + cls, 'num-children': 2, 'left': 0, 'right': 1}] + l + r


def conj(*xs):
    out = xs[0]
    for x in xs[1:]:
        out = bi('And', out, x)
    return out


def alias(child, name):
    return [{'class': X + 'Alias', 'num-children': 1, 'child': 0, 'name': name, 'exprId': eid(name), 'qualifier': [], 'explicitMetadata': {},
             'nonInheritableMetadataKeys': []}] + child


def node(cls, nchildren, **kw):
    return dict({'class': P + cls, 'num-children': nchildren}, **kw)


def scan(table, cols, partition_filters=(), data_filters=()):
    return [node('FileSourceScanExec', 0, relation=None, output=[attr(*c) for c in cols], requiredSchema={'type': 'struct', 'fields': []},
                 partitionFilters=list(partition_filters), optionalBucketSet=None, optionalNumCoalescedBuckets=None,
                 dataFilters=list(data_filters), tableIdentifier={'product-class': 'org.apache.spark.sql.catalyst.TableIdentifier', 'table': table},
                 disableBucketedScan=False)]


def wrap(cls, child, **kw):
    return [node(cls, 1, child=0, **kw)] + child


def bhj(left_keys, right_keys, build_side, left, right):
    return [node('joins.BroadcastHashJoinExec', 2, leftKeys=left_keys, rightKeys=right_keys,
                 joinType={'object': 'org.apache.spark.sql.catalyst.plans.Inner$'},
                 buildSide={'object': 'org.apache.spark.sql.catalyst.optimizer.' + build_side + '$'}, condition=None, left=0, right=1,
                 isNullAwareAntiJoin=False)] + left + right


def agg_expr(fn_cls, arg, mode, result_key):
    return [{'class': X + 'aggregate.AggregateExpression', 'num-children': 1, 'aggregateFunction': 0,
             'mode': {'object': X + 'aggregate.' + mode + '$'}, 'isDistinct': False, 'filter': None, 'resultId': eid(result_key)},
            {'class': X + 'aggregate.' + fn_cls, 'num-children': 1, 'child': 0}] + arg


def sort_order(child, direction, nulls):
    return [{'class': X + 'SortOrder', 'num-children': 1, 'child': 0, 'direction': {'object': X + direction + '$'},
             'nullOrdering': {'object': X + nulls + '$'}, 'sameOrderExpressions': []}] + child


def main():
    dyear, dsk, dmoy = ('d_year', 'long'), ('d_date_sk', 'integer', False), ('d_moy', 'long')
    ssitem, ssprice, ssdate = ('ss_item_sk', 'integer', False), ('ss_ext_sales_price', 'decimal(7,2)'), ('ss_sold_date_sk', 'integer')
    isk, ibid, ibrand, imanu = ('i_item_sk', 'integer', False), ('i_brand_id', 'long'), ('i_brand', 'string'), ('i_manufact_id', 'long')
    date_side = wrap('exchange.BroadcastExchangeExec',
                     wrap('ProjectExec',
                          wrap('FilterExec', scan('date_dim', [dsk, dyear, dmoy]),
                               condition=conj(un('IsNotNull', attr(*dmoy)), bi('EqualTo', attr(*dmoy), lit(1, 'long')), un('IsNotNull', attr(*dsk)))),
                          projectList=[attr(*dsk), attr(*dyear)]),
                     mode={'object': 'org.apache.spark.sql.execution.joins.HashedRelationBroadcastMode'}, runId='00000000-0000-0000-0000-000000000001')
    dyn = [{'class': X + 'DynamicPruningExpression', 'num-children': 1, 'child': 0},
           {'class': X + 'InSubqueryExec', 'num-children': 1, 'child': 0, 'plan': None, 'exprId': eid('dynamicpruning'), 'resultBroadcast': None}] + attr(*ssdate)
    fact = wrap('FilterExec',
                wrap('ColumnarToRowExec', scan('store_sales', [ssitem, ssprice, ssdate], partition_filters=[un('IsNotNull', attr(*ssdate)), dyn],
                                               data_filters=[un('IsNotNull', attr(*ssitem))])),
                condition=un('IsNotNull', attr(*ssitem)))
    j1 = wrap('ProjectExec', bhj([attr(*dsk)], [attr(*ssdate)], 'BuildLeft', date_side, fact), projectList=[attr(*dyear), attr(*ssitem), attr(*ssprice)])
    item_side = wrap('exchange.BroadcastExchangeExec',
                     wrap('ProjectExec',
                          wrap('FilterExec', scan('item', [isk, ibid, ibrand, imanu]),
                               condition=conj(un('IsNotNull', attr(*imanu)), bi('EqualTo', attr(*imanu), lit(606, 'long')), un('IsNotNull', attr(*isk)))),
                          projectList=[attr(*isk), attr(*ibid), attr(*ibrand)]),
                     mode={'object': 'org.apache.spark.sql.execution.joins.HashedRelationBroadcastMode'}, runId='00000000-0000-0000-0000-000000000002')
    j2 = wrap('ProjectExec', bhj([attr(*ssitem)], [attr(*isk)], 'BuildRight', j1, item_side),
              projectList=[attr(*dyear), attr(*ssprice), attr(*ibid), attr(*ibrand)])
    keys = [attr(*dyear), attr(*ibrand), attr(*ibid)]
    summed = un('UnscaledValue', attr(*ssprice))
    partial = wrap('aggregate.HashAggregateExec', j2, requiredChildDistributionExpressions=None, groupingExpressions=keys,
                   aggregateExpressions=[agg_expr('Sum', summed, 'Partial', 'sum_result')], aggregateAttributes=[attr('sum', 'long', key='sum_buf')],
                   initialInputBufferOffset=0, resultExpressions=keys + [attr('sum', 'long', key='sum_buf')])
    shuffled = wrap('exchange.ShuffleExchangeExec', partial,
                    outputPartitioning={'product-class': 'org.apache.spark.sql.catalyst.plans.physical.HashPartitioning', 'numPartitions': 200},
                    shuffleOrigin={'object': 'org.apache.spark.sql.execution.exchange.ENSURE_REQUIREMENTS$'})
    result = [attr(*dyear), alias(attr(*ibid), 'brand_id'), alias(attr(*ibrand), 'brand'),
              alias(un('MakeDecimal', attr('sum(UnscaledValue(ss_ext_sales_price))', 'long', key='sum_result'), precision=17, scale=2, nullOnOverflow=True),
                    'sum_agg')]
    final = wrap('aggregate.HashAggregateExec', shuffled, requiredChildDistributionExpressions=keys, groupingExpressions=keys,
                 aggregateExpressions=[agg_expr('Sum', summed, 'Final', 'sum_result')],
                 aggregateAttributes=[attr('sum(UnscaledValue(ss_ext_sales_price))', 'long', key='sum_result')], initialInputBufferOffset=3,
                 resultExpressions=result)
    out_attrs = [attr(*dyear), attr('brand_id', 'long'), attr('brand', 'string'), attr('sum_agg', 'decimal(17,2)')]
    top = wrap('TakeOrderedAndProjectExec', final, limit=100,
               sortOrder=[sort_order(attr(*dyear), 'Ascending', 'NullsFirst'), sort_order(attr('sum_agg', 'decimal(17,2)'), 'Descending', 'NullsLast'),
                          sort_order(attr('brand_id', 'long'), 'Ascending', 'NullsFirst')],
               projectList=out_attrs)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'catalyst_q3_executed_plan.json')
    with open(path, 'w') as f:
        json.dump(top, f, indent=1)
    print('wrote', path, len(top), 'nodes')


if __name__ == '__main__':
    main()
