"""Catalyst physical plan (JSON) -> the engine's plan IR.

`north_star`: "Spark runs on the host CPU only to parse and plan each NDS query and export its Catalyst physical plan";
the reference reaches Catalyst through `spark_session.sql(query)` (/root/reference/nds/nds_power.py:130).  With a JVM the
host side would call `df._jdf.queryExecution().executedPlan().toJSON()` and hand the text to `translate()` instead of
going through this repo's own SQL front-end (sql.py / planner.py).

UNVERIFIED AGAINST SPARK: neither the build nor the GPU box has a JVM, and the reference ships no plan dumps.  The JSON
layout below is `TreeNode.toJSON` as Spark 3.x writes it, restated from memory ([EXT] in SURVEY.md's terms): a plan is a
JSON array of node objects in PRE-ORDER, each with "class", "num-children" and its constructor fields; a field holding
expressions is a list of such pre-order arrays (one per expression); case-class values (ExprId, TableIdentifier, ...) are
objects with "product-class".  tests/golden/catalyst_q3_executed_plan.json is a HAND-WRITTEN sample in that layout for
query3; the mapping is pinned against it and must be re-checked against a real dump before it is trusted.

Supported operators: WholeStageCodegen / InputAdapter / ColumnarToRow / AdaptiveSparkPlan / ShuffleQueryStage /
BroadcastQueryStage / ReusedExchange (transparent), FileSourceScan, Filter, Project, BroadcastHashJoin (inner, left outer,
left semi, left anti), BroadcastExchange, ShuffleExchange, HashAggregate (partial + final pairs), Sort, TakeOrderedAndProject,
GlobalLimit / LocalLimit.  Anything else raises CatalystError: an unsupported plan node fails loudly.
"""
import json


class CatalystError(ValueError):
    pass


def _short(cls):
    return cls.rsplit('.', 1)[-1]


class _Tree:
    """A pre-order array turned back into a tree: node dict + children."""

    def __init__(self, arr):
        self.arr = arr
        self.pos = 0

    def build(self):
        node = self.arr[self.pos]
        self.pos += 1
        kids = [self.build() for _ in range(int(node.get('num-children', 0)))]
        return (node, kids)


def _tree(arr):
    if not isinstance(arr, list) or not arr:
        raise CatalystError('expected a non-empty pre-order node array')
    t = _Tree(arr)
    root = t.build()
    if t.pos != len(arr):
        raise CatalystError('trailing nodes in a pre-order array')
    return root


_DTYPES = {'integer': 'int32', 'long': 'int64', 'double': 'double', 'string': 'string', 'date': 'date', 'boolean': 'bool'}


def _dtype(t):
    if isinstance(t, str):
        if t in _DTYPES:
            return _DTYPES[t]
        if t.startswith('decimal('):
            return t
        if t.startswith('char(') or t.startswith('varchar('):
            return 'string'
    raise CatalystError('unsupported Catalyst data type %r' % (t,))


_CMP = {'EqualTo': 'eq', 'LessThan': 'lt', 'LessThanOrEqual': 'le', 'GreaterThan': 'gt', 'GreaterThanOrEqual': 'ge'}
_ARITH = {'Add': 'add', 'Subtract': 'sub', 'Multiply': 'mul', 'Divide': 'div'}
_PASS = ('PromotePrecision', 'CheckOverflow', 'KnownFloatingPointNormalized', 'NormalizeNaNAndZero', 'KnownNotNull')


class Translator:
    def __init__(self):
        self.names = {}          # exprId -> column name in the IR
        self.agg_outputs = {}    # exprId of an aggregate result attribute -> IR aggregate name
        self.nagg = 0

    # ---- expressions ---------------------------------------------------------------------------------
    def attr_name(self, node):
        eid = node.get('exprId', {}).get('id')
        if eid in self.names:
            return self.names[eid]
        self.names[eid] = node['name']
        return node['name']

    def expr(self, arr):
        return self._expr(_tree(arr))

    def _expr(self, t):
        node, kids = t
        c = _short(node['class'])
        if c == 'AttributeReference':
            return {'col': self.attr_name(node)}
        if c == 'Alias':
            return self._expr(kids[0])
        if c in _PASS:
            return self._expr(kids[0])
        if c == 'Literal':
            typ = _dtype(node['dataType'])
            v = node.get('value')
            if v is None:
                return {'lit': None, 'type': typ}
            if typ == 'string':
                return {'lit': str(v), 'type': 'string'}
            if typ == 'double':
                return {'lit': float(v), 'type': 'double'}
            if typ.startswith('decimal('):
                scale = int(typ[:-1].split(',')[1])
                import decimal
                return {'lit': int(decimal.Decimal(str(v)).scaleb(scale)), 'type': typ}
            return {'lit': int(v), 'type': 'int64' if typ in ('int32', 'int64') else typ}
        if c in _CMP:
            return {'fn': _CMP[c], 'args': [self._expr(kids[0]), self._expr(kids[1])]}
        if c in _ARITH:
            return {'fn': _ARITH[c], 'args': [self._expr(kids[0]), self._expr(kids[1])]}
        if c in ('And', 'Or'):
            return {'fn': c.lower(), 'args': [self._expr(kids[0]), self._expr(kids[1])]}
        if c == 'Not':
            return {'fn': 'not', 'args': [self._expr(kids[0])]}
        if c == 'IsNotNull':
            return {'fn': 'isnotnull', 'args': [self._expr(kids[0])]}
        if c == 'IsNull':
            return {'fn': 'isnull', 'args': [self._expr(kids[0])]}
        if c == 'In':
            x = self._expr(kids[0])
            items = [{'fn': 'eq', 'args': [x, self._expr(k)]} for k in kids[1:]]
            return items[0] if len(items) == 1 else {'fn': 'or', 'args': items}
        if c == 'Substring':
            pos, ln = self._expr(kids[1]), self._expr(kids[2])
            return {'fn': 'substr', 'args': [self._expr(kids[0])], 'start': pos['lit'], 'len': ln['lit']}
        if c == 'Cast':
            typ = _dtype(node['dataType'])
            if typ == 'double':
                return {'fn': 'cast_double', 'args': [self._expr(kids[0])]}
            return self._expr(kids[0])                 # widening integer / decimal casts keep the value
        if c == 'UnscaledValue' or c == 'MakeDecimal':
            return self._expr(kids[0])                 # Spark's long arithmetic over small decimals: values are unscaled here anyway
        if c == 'DynamicPruningExpression':
            return None
        raise CatalystError('unsupported Catalyst expression %s' % c)

    # ---- plan nodes ------------------------------------------------------------------------------------
    def plan(self, t):
        node, kids = t
        c = _short(node['class'])
        if c in ('WholeStageCodegenExec', 'InputAdapter', 'ColumnarToRowExec', 'AdaptiveSparkPlanExec', 'ShuffleQueryStageExec',
                 'BroadcastQueryStageExec', 'ReusedExchangeExec', 'BroadcastExchangeExec', 'ShuffleExchangeExec', 'LocalLimitExec'):
            return self.plan(kids[0])
        if c == 'FileSourceScanExec':
            table = node.get('tableIdentifier', {}).get('table')
            if not table:
                raise CatalystError('FileSourceScanExec without a table identifier')
            cols = [self.attr_name(a[0]) for a in node['output']]
            scan = {'op': 'scan', 'table': table, 'columns': cols}
            pruned = any('DynamicPruning' in json.dumps(f) for f in node.get('partitionFilters', []))
            conds = [self.expr(f) for f in node.get('dataFilters', [])] if node.get('pushedDownFiltersAreExact') else []
            conds = [x for x in conds if x is not None]
            if conds:
                scan = {'op': 'filter', 'cond': conds[0] if len(conds) == 1 else {'fn': 'and', 'args': conds}, 'child': scan}
            scan['_pruned'] = pruned
            return scan
        if c == 'FilterExec':
            child = self.plan(kids[0])
            cond = self.expr(node['condition'])
            pruned = child.pop('_pruned', False) if isinstance(child, dict) else False
            out = {'op': 'filter', 'cond': cond, 'child': child}
            out['_pruned'] = pruned
            return out
        if c == 'ProjectExec':
            child = self.plan(kids[0])
            pruned = child.pop('_pruned', False)
            exprs = []
            passthrough = True
            for e in node['projectList']:
                root = e[0]
                ir = self.expr(e)
                name = self.attr_name(root) if _short(root['class']) in ('AttributeReference', 'Alias') else None
                if _short(root['class']) == 'Alias':
                    self.names[root['exprId']['id']] = root['name']
                    name = root['name']
                    passthrough = False
                exprs.append({'name': name, 'expr': ir})
            if passthrough:
                child['_pruned'] = pruned
                return child                 # a pure column-pruning projection: the IR's scans / joins carry columns by name
            out = {'op': 'project', 'exprs': exprs, 'child': child}
            out['_pruned'] = pruned
            return out
        if c == 'BroadcastHashJoinExec':
            lk, rk = node['leftKeys'], node['rightKeys']
            if len(lk) != 1 or len(rk) != 1:
                raise CatalystError('multi-column join keys are not supported')
            jt = node.get('joinType', {})
            jt = jt.get('object', jt) if isinstance(jt, dict) else jt
            kind = {'Inner': 'inner', 'LeftOuter': 'left', 'LeftSemi': 'semi', 'LeftAnti': 'anti'}.get(_short(str(jt)).rstrip('$'))
            if kind is None:
                raise CatalystError('unsupported join type %r' % (jt,))
            bs = node.get('buildSide', {})
            bs = bs.get('object', bs) if isinstance(bs, dict) else bs
            build_right = 'BuildRight' in str(bs)
            left, right = self.plan(kids[0]), self.plan(kids[1])
            lkey, rkey = self.expr(lk[0]), self.expr(rk[0])
            if 'col' not in lkey or 'col' not in rkey:
                raise CatalystError('join keys must be plain columns')
            probe, build, pkey, bkey = (left, right, lkey, rkey) if build_right else (right, left, rkey, lkey)
            if kind != 'inner' and not build_right:
                raise CatalystError('an outer / semi / anti join that builds its left side is not supported')
            pruned = probe.pop('_pruned', False)
            build.pop('_pruned', None)
            out = {'op': 'join', 'type': kind, 'probe': probe, 'build': build, 'probe_key': pkey['col'], 'build_key': bkey['col']}
            if pruned and self._is_partition_key(probe, pkey['col']):
                out['prune_partitions'] = True
            else:
                out['_pruned'] = pruned
            return out
        if c == 'HashAggregateExec':
            child = self.plan(kids[0])
            child.pop('_pruned', None)
            modes = set()
            aggs = []
            for ae in node.get('aggregateExpressions', []):
                (an, akids) = _tree(ae)
                if _short(an['class']) != 'AggregateExpression':
                    raise CatalystError('expected AggregateExpression')
                mode = an.get('mode', {})
                mode = mode.get('object', mode) if isinstance(mode, dict) else mode
                modes.add(_short(str(mode)).rstrip('$'))
                fn_node, fn_kids = akids[0]
                fname = _short(fn_node['class'])
                rid = an.get('resultId', {}).get('id')
                aggs.append((fname, fn_kids, rid, bool(an.get('isDistinct'))))
            if modes <= {'Partial'}:
                # the partial half of a partial / final pair: remember the grouping, the final half emits the IR node
                self._partial = (node, child)
                child['_partial'] = True
                return child
            if not modes <= {'Final', 'Complete'}:
                raise CatalystError('unsupported aggregate modes %s' % sorted(modes))
            if child.pop('_partial', False):
                pnode, child = self._partial
            keys = []
            for g in node.get('groupingExpressions', []):
                ir = self.expr(g)
                keys.append({'name': ir['col'] if 'col' in ir else '_gk%d' % len(keys), 'expr': ir})
            specs = []
            for fname, fn_kids, rid, distinct in aggs:
                if distinct:
                    raise CatalystError('DISTINCT aggregates are not supported')
                name = '_agg%d' % self.nagg
                self.nagg += 1
                if fname == 'Count':
                    arg = self._expr(fn_kids[0])
                    spec = {'name': name, 'fn': 'count_star'} if 'lit' in arg else {'name': name, 'fn': 'count', 'expr': arg}
                elif fname in ('Sum', 'Min', 'Max'):
                    spec = {'name': name, 'fn': fname.lower(), 'expr': self._expr(fn_kids[0])}
                elif fname == 'Average':
                    raise CatalystError('Average must be planned as Sum / Count (this translator does not split it yet)')
                else:
                    raise CatalystError('unsupported aggregate function %s' % fname)
                specs.append(spec)
                self.agg_outputs[rid] = name
                self.names[rid] = name
            out = {'op': 'aggregate', 'keys': keys, 'aggs': specs, 'child': child}
            # resultExpressions name the output attributes: aliases of aggregate results keep the IR aggregate name
            for e in node.get('resultExpressions', []):
                root = e[0]
                if _short(root['class']) == 'Alias' and len(e) > 1:
                    inner = self.expr(e)
                    if 'col' in inner:
                        self.names[root['exprId']['id']] = inner['col']
                        self._aliases = getattr(self, '_aliases', {})
                        self._aliases[root['exprId']['id']] = (root['name'], inner['col'])
            return out
        if c in ('TakeOrderedAndProjectExec', 'SortExec'):
            child = self.plan(kids[0])
            child.pop('_pruned', None)
            if c == 'TakeOrderedAndProjectExec':
                exprs = []
                for e in node['projectList']:
                    root = e[0]
                    ir = self.expr(e)
                    out_name = root['name']
                    alias = getattr(self, '_aliases', {}).get(root.get('exprId', {}).get('id'))
                    if alias:
                        out_name = alias[0]
                    exprs.append({'name': out_name, 'expr': ir})
                renamed = {x['expr']['col']: x['name'] for x in exprs if 'col' in x['expr']}
                child = {'op': 'project', 'exprs': exprs, 'child': child}
            keys = []
            for so in node['sortOrder']:
                (sn, sk) = _tree(so)
                ir = self._expr(sk[0])
                if c == 'TakeOrderedAndProjectExec' and 'col' in ir:
                    ir = {'col': renamed.get(ir['col'], ir['col'])}
                d = sn.get('direction', {})
                d = d.get('object', d) if isinstance(d, dict) else d
                no = sn.get('nullOrdering', {})
                no = no.get('object', no) if isinstance(no, dict) else no
                keys.append({'expr': ir, 'desc': 'Descending' in str(d), 'nulls_first': 'NullsFirst' in str(no)})
            return {'op': 'sort', 'keys': keys, 'limit': int(node.get('limit', -1)), 'child': child}
        if c == 'GlobalLimitExec':
            child = self.plan(kids[0])
            if child.get('op') == 'sort':
                child['limit'] = int(node['limit'])
                return child
            return {'op': 'sort', 'keys': [], 'limit': int(node['limit']), 'child': child}
        raise CatalystError("unsupported Catalyst plan node %s (no CPU fallback exists)" % c)

    @staticmethod
    def _is_partition_key(probe, col):
        from . import nds_schema
        node = probe
        while node.get('op') in ('filter', 'project'):
            node = node['child']
        if node.get('op') == 'join':
            return Translator._is_partition_key(node['probe'], col)
        return node.get('op') == 'scan' and nds_schema.TABLE_PARTITIONING.get(node['table']) == col


def _strip(node):
    if isinstance(node, dict):
        node.pop('_pruned', None)
        node.pop('_partial', None)
        for k in ('child', 'probe', 'build'):
            if k in node:
                _strip(node[k])
    return node


def _drop_inferred_notnull(node, keys=()):
    """Catalyst infers isnotnull(k) for every inner / semi join key and places it over the scan; the joins drop NULL keys
    themselves, and without the filter the fact scan under a join chain can take the engine's fused scan -> probe route."""
    if not isinstance(node, dict):
        return node
    op = node.get('op')
    if op == 'join':
        inner = node.get('type', 'inner') in ('inner', 'semi')
        node['probe'] = _drop_inferred_notnull(node['probe'], tuple(keys) + ((node['probe_key'],) if inner else ()))
        node['build'] = _drop_inferred_notnull(node['build'], (node['build_key'],) if inner else ())
        return node
    if op == 'filter' and node['child'].get('op') == 'scan' and keys:
        def conj(e):
            return conj(e['args'][0]) + conj(e['args'][1]) if e.get('fn') == 'and' and len(e['args']) == 2 else [e]
        rest = [e for e in conj(node['cond']) if not (e.get('fn') == 'isnotnull' and e['args'][0].get('col') in keys)]
        if not rest:
            return node['child']
        node['cond'] = rest[0] if len(rest) == 1 else {'fn': 'and', 'args': rest}
        return node
    if 'child' in node:
        node['child'] = _drop_inferred_notnull(node['child'], keys if op in ('filter', 'project') else ())
    return node


def translate(plan_json):
    """`executedPlan.toJSON` text (or the parsed array) -> {'root': IR}."""
    arr = json.loads(plan_json) if isinstance(plan_json, (str, bytes)) else plan_json
    tr = Translator()
    root = tr.plan(_tree(arr))
    return {'root': _drop_inferred_notnull(_strip(root))}
