"""CPU restatement of the plan IR's operator semantics (TEST INFRASTRUCTURE - only tests may import this).

The engine (libndsb200.so) executes the JSON plans `planner.py` writes; this module executes the same plans row by row in
plain Python over pyarrow-read files, so that the CPU test suite (`-m "not gpu"`) can check SQL -> planner -> plan against
the hand-written query restatements of `oracle/nds_queries.py` without a GPU, and so that the expectations of the GPU
SQL tests are checked against a second, independent executor (tests/test_plan_interp.py runs those test functions with
`FakeEngine` in place of the engine).

What it restates (reference boundary: the statements the reference hands to Spark at /root/reference/nds/nds_power.py:130;
operator semantics as DESIGN.md section 4 states them, [EXT] items included):
  scan / filter / project / join (inner, left, semi, anti; unique or repeated build keys) / aggregate (sum, count,
  count_star, min, max; keyless; DISTINCT) / rollup / window / rank / sort (+ limit) / exchange (pass-through, one rank),
  and the expression functions of exec.cu's compile() with its result types (decimal scale alignment, SUM / AVG widening,
  `/` -> double, x / 0 -> NULL, three-valued logic).
Not restated: dynamic partition pruning, row maps, device-side row counts - they do not change results.
"""
import decimal
import functools
import os
import re

import pyarrow as pa

from . import nds_queries as oq

_DEC = re.compile(r'decimal\((\d+),\s*(\d+)\)')

# ---- types: ('int', is32) | ('double',) | ('string',) | ('bool',) | ('date', is32) | ('decimal', p, s)
INT64, INT32, DOUBLE, STRING, BOOL = ('int', False), ('int', True), ('double',), ('string',), ('bool',)


def parse_type(s):
    if s in ('int', 'int64', 'long', 'bigint'):
        return INT64
    if s == 'int32':
        return INT32
    if s == 'double':
        return DOUBLE
    if s == 'string':
        return STRING
    if s == 'bool':
        return BOOL
    if s == 'date':
        return ('date', True)
    m = _DEC.match(s)
    if m:
        return ('decimal', int(m.group(1)), int(m.group(2)))
    raise ValueError('unknown type %r' % s)


def _arrow_to_type(t):
    if pa.types.is_int32(t):
        return INT32
    if pa.types.is_integer(t):
        return INT64
    if pa.types.is_floating(t):
        return DOUBLE
    if pa.types.is_string(t) or pa.types.is_large_string(t):
        return STRING
    if pa.types.is_date32(t):
        return ('date', True)
    if pa.types.is_decimal(t):
        return ('decimal', t.precision, t.scale)
    if pa.types.is_boolean(t):
        return BOOL
    raise ValueError('unsupported column type %s' % t)


def _type_to_arrow(t):
    k = t[0]
    if k == 'int':
        return pa.int32() if t[1] else pa.int64()
    if k == 'bool':
        return pa.int64()
    if k == 'double':
        return pa.float64()
    if k == 'string':
        return pa.string()
    if k == 'date':
        return pa.date32()
    return pa.decimal128(t[1], t[2])


def _no32(t):
    if t[0] == 'int':
        return INT64
    if t[0] == 'date':
        return ('date', False)
    return t


class Batch:
    """Columns as Python lists: ints (decimals unscaled, dates as days), floats, str, None."""
    def __init__(self):
        self.names, self.types, self.cols = [], [], []

    @property
    def n(self):
        return len(self.cols[0]) if self.cols else self._n

    _n = 0

    def add(self, name, t, values):
        self.names.append(name)
        self.types.append(t)
        self.cols.append(values)

    def find(self, name):
        return self.names.index(name) if name in self.names else -1

    def take(self, idx):
        out = Batch()
        out._n = len(idx)
        for nme, t, c in zip(self.names, self.types, self.cols):
            out.add(nme, t, [None if i is None else c[i] for i in idx])
        return out


def _stable_hash(key):
    import zlib
    return zlib.crc32(repr(key).encode())


def _scan_table(warehouse, table, columns, shard=None):
    from spark_rapids_benchmarks_b200 import nds_schema
    part = nds_schema.TABLE_PARTITIONING.get(table)
    if part and os.path.isdir(os.path.join(warehouse, table)) and any('=' in e for e in os.listdir(os.path.join(warehouse, table))):
        t = oq.read_fact(warehouse, table, part, list(columns), None, shard)
    else:
        t = oq.read_dim(warehouse, table, list(columns))
    b = Batch()
    b._n = t.num_rows
    for name in columns:
        col = t[name]
        ty = _arrow_to_type(col.type)
        vals = col.to_pylist()
        if ty[0] == 'decimal':
            vals = [None if v is None else int(v.scaleb(ty[2])) for v in vals]
        elif ty[0] == 'date':
            vals = col.cast(pa.int32()).to_pylist()
        b.add(name, ty, vals)
    return b


# ---------------------------------------------------------------- expressions: compile to (type, fn(row_index) -> value)
def _to_double(t, f):
    if t[0] == 'double':
        return f
    if t[0] == 'decimal' and t[2] > 0:
        sc = 10.0 ** t[2]
        return lambda i: (lambda v: None if v is None else float(v) / sc)(f(i))
    return lambda i: (lambda v: None if v is None else float(v))(f(i))


def _scale_up(f, k):
    if k <= 0:
        return f
    m = 10 ** k
    return lambda i: (lambda v: None if v is None else v * m)(f(i))


def _half_up_div(a, b):
    """a / b rounded half away from zero (Spark HALF_UP), b > 0."""
    q, r = divmod(abs(a), b)
    if 2 * r >= b:
        q += 1
    return q if a >= 0 else -q


def compile_expr(x, b):
    if 'col' in x:
        ci = b.find(x['col'])
        if ci < 0:
            raise ValueError("unknown column '%s'" % x['col'])
        col = b.cols[ci]
        return b.types[ci], (lambda i: col[i])
    if 'lit' in x:
        v = x['lit']
        if v is None:
            return (parse_type(x['type']) if 'type' in x else INT64), (lambda i: None)
        t = parse_type(x.get('type', 'string' if isinstance(v, str) else ('int' if isinstance(v, int) else 'double')))
        if t[0] == 'double':
            v = float(v)
        return t, (lambda i: v)
    fn, args = x['fn'], x['args']
    out = None
    if fn in ('eq', 'ne', 'lt', 'le', 'gt', 'ge'):
        (ta, fa), (tb, fb) = compile_expr(args[0], b), compile_expr(args[1], b)
        if (ta[0] == 'string') != (tb[0] == 'string'):
            raise ValueError('comparison between string and non-string')
        if ta[0] == 'string':
            ga, gb = (lambda i: (lambda v: None if v is None else v.encode())(fa(i))), (lambda i: (lambda v: None if v is None else v.encode())(fb(i)))
        elif ta[0] == 'double' or tb[0] == 'double':
            ga, gb = _to_double(ta, fa), _to_double(tb, fb)
        else:
            sa, sb = (ta[2] if ta[0] == 'decimal' else 0), (tb[2] if tb[0] == 'decimal' else 0)
            ga, gb = _scale_up(fa, sb - sa), _scale_up(fb, sa - sb)
        op = {'eq': lambda p, q: p == q, 'ne': lambda p, q: p != q, 'lt': lambda p, q: p < q, 'le': lambda p, q: p <= q,
              'gt': lambda p, q: p > q, 'ge': lambda p, q: p >= q}[fn]

        def cmp(i):
            p, q = ga(i), gb(i)
            return None if p is None or q is None else op(p, q)
        out = BOOL, cmp
    elif fn in ('and', 'or'):
        fs = [compile_expr(a, b)[1] for a in args]
        if fn == 'and':
            def f(i):
                res = True
                for g in fs:
                    v = g(i)
                    if v is None:
                        res = None
                    elif not v:
                        return False
                return res
        else:
            def f(i):
                res = False
                for g in fs:
                    v = g(i)
                    if v is not None and v:
                        return True
                    if v is None:
                        res = None
                return res
        out = BOOL, f
    elif fn == 'not':
        _, g = compile_expr(args[0], b)
        out = BOOL, (lambda i: (lambda v: None if v is None else not v)(g(i)))
    elif fn in ('isnull', 'isnotnull'):
        _, g = compile_expr(args[0], b)
        out = BOOL, ((lambda i: g(i) is None) if fn == 'isnull' else (lambda i: g(i) is not None))
    elif fn in ('add', 'sub', 'mul'):
        (ta, fa), (tb, fb) = compile_expr(args[0], b), compile_expr(args[1], b)
        if ta[0] == 'string' or tb[0] == 'string':
            raise ValueError('arithmetic on strings')
        pyop = {'add': lambda p, q: p + q, 'sub': lambda p, q: p - q, 'mul': lambda p, q: p * q}[fn]
        if ta[0] == 'double' or tb[0] == 'double':
            ga, gb = _to_double(ta, fa), _to_double(tb, fb)
            t = DOUBLE
        else:
            sa, sb = (ta[2] if ta[0] == 'decimal' else 0), (tb[2] if tb[0] == 'decimal' else 0)
            dec = ta[0] == 'decimal' or tb[0] == 'decimal'
            pa_ = ta[1] if ta[0] == 'decimal' else (10 if ta[1] else 20)
            pb_ = tb[1] if tb[0] == 'decimal' else (10 if tb[1] else 20)
            ga, gb = (fa, fb) if fn == 'mul' else (_scale_up(fa, sb - sa), _scale_up(fb, sa - sb))
            if dec:
                t = ('decimal', min(38, pa_ + pb_ + 1), sa + sb) if fn == 'mul' else \
                    ('decimal', min(38, max(pa_ - sa, pb_ - sb) + max(sa, sb) + 1), max(sa, sb))
            else:
                t = INT64

        def arith(i):
            p, q = ga(i), gb(i)
            return None if p is None or q is None else pyop(p, q)
        out = t, arith
    elif fn == 'neg':
        t, g = compile_expr(args[0], b)
        out = _no32(t), (lambda i: (lambda v: None if v is None else -v)(g(i)))
    elif fn == 'substr':
        t, g = compile_expr(args[0], b)
        start, ln = int(x.get('start', 1)), int(x.get('len', 1 << 30))

        def sub(i):
            v = g(i)
            if v is None:
                return None
            s0 = max(start, 1) - 1
            return v[s0:s0 + max(ln, 0)]
        out = STRING, sub
    elif fn == 'if':
        _, gc = compile_expr(args[0], b)
        (tt, gt), (te, ge) = compile_expr(args[1], b), compile_expr(args[2], b)
        if (tt[0] == 'string') != (te[0] == 'string') or (tt[0] == 'double') != (te[0] == 'double'):
            raise ValueError('if() branches of different type families')
        # the engine does not align decimal scales between the branches; the planner only builds equal-scale branches
        out = _no32(tt), (lambda i: gt(i) if gc(i) else ge(i))
    elif fn == 'decavg':
        (ts, gs), (_, gcnt) = compile_expr(args[0], b), compile_expr(args[1], b)
        up = int(x.get('up', 4))
        m = 10 ** up

        def avg(i):
            s, c = gs(i), gcnt(i)
            return None if s is None or not c else _half_up_div(s * m, c)
        out = ('decimal', min(38, (ts[1] if ts[0] == 'decimal' else 20) - 10 + up), (ts[2] if ts[0] == 'decimal' else 0) + up), avg
    elif fn in ('divi2f', 'div'):
        (ta, fa), (tb, fb) = compile_expr(args[0], b), compile_expr(args[1], b)
        ga, gb = (_to_double(ta, fa), _to_double(tb, fb)) if fn == 'div' else \
                 ((lambda i: (lambda v: None if v is None else float(v))(fa(i))), (lambda i: (lambda v: None if v is None else float(v))(fb(i))))

        def div(i):
            p, q = ga(i), gb(i)
            return None if p is None or q is None or q == 0 else p / q
        out = DOUBLE, div
    elif fn == 'cast_double':
        t, g = compile_expr(args[0], b)
        out = DOUBLE, _to_double(t, g)
    else:
        raise ValueError("unsupported expression function '%s'" % fn)
    if 'type' in x:
        forced = parse_type(x['type'])
        out = _no32(forced), out[1]
    return out


def _out_type(t):
    return INT64 if t[0] == 'bool' else t


def _out_val(t, v):
    return (None if v is None else int(bool(v))) if t[0] == 'bool' else v


# ---------------------------------------------------------------- operators
class Interp:
    """rank / world: one rank of `world` - a hive-partitioned table is read as this rank's shard (every world-th file, like the
    engine), an `exchange` runs its child for every rank and gathers the rows (mode hash: keeps the rows whose keys hash to
    this rank), so that the distributed plan forms can be checked on the CPU."""
    def __init__(self, warehouse, rank=0, world=1):
        self.wh = warehouse
        self.rank, self.world = rank, world
        self.ops = []

    def run(self, node):
        op = node['op']
        return getattr(self, 'op_' + op)(node)

    def op_scan(self, node):
        return _scan_table(self.wh, node['table'], node['columns'], (self.rank, self.world) if self.world > 1 else None)

    def op_exchange(self, node):
        if self.world <= 1:
            return self.run(node['child'])
        parts = [Interp(self.wh, r, self.world).run(node['child']) for r in range(self.world)]
        out = Batch()
        out._n = sum(p.n for p in parts)
        for c, name in enumerate(parts[0].names):
            out.add(name, parts[0].types[c], [v for p in parts for v in p.cols[c]])
        if node.get('mode', 'gather') == 'hash':
            kc = [out.cols[out.find(k)] for k in node['keys']]
            mine = [i for i in range(out.n) if _stable_hash(tuple(col[i] for col in kc)) % self.world == self.rank]
            out = out.take(mine)
        return out

    def op_filter(self, node):
        b = self.run(node['child'])
        t, f = compile_expr(node['cond'], b)
        if t != BOOL:
            raise ValueError('filter condition is not boolean')
        return b.take([i for i in range(b.n) if f(i)])

    def op_project(self, node):
        b = self.run(node['child'])
        out = Batch()
        out._n = b.n
        for e in node['exprs']:
            if 'col' in e['expr'] and len(e['expr']) == 1:
                ci = b.find(e['expr']['col'])
                if ci < 0:
                    raise ValueError("unknown column '%s'" % e['expr']['col'])
                out.add(e['name'], _out_type(b.types[ci]), b.cols[ci])
                continue
            t, f = compile_expr(e['expr'], b)
            out.add(e['name'], _out_type(t), [_out_val(t, f(i)) for i in range(b.n)])
        return out

    def op_join(self, node):
        kind = node.get('type', 'inner')
        build = self.run(node['build'])
        probe = self.run(node['probe'])
        if probe.find(node['probe_key']) >= 0 and self._is_fact_scan(node['probe']):
            self.ops.append('scan_probe')
        bk, pk = build.cols[build.find(node['build_key'])], probe.cols[probe.find(node['probe_key'])]
        table = {}
        for i, k in enumerate(bk):
            if k is not None:
                table.setdefault(k, []).append(i)
        if node.get('unique_build', True) and kind in ('inner', 'left'):
            if any(len(v) > 1 for v in table.values()):
                raise ValueError('duplicate build keys in a join declared unique')
        pi, bi = [], []
        for i, k in enumerate(pk):
            m = table.get(k) if k is not None else None
            if kind == 'semi':
                if m:
                    pi.append(i)
            elif kind == 'anti':
                if not m:
                    pi.append(i)
            elif m:
                for j in m:
                    pi.append(i)
                    bi.append(j)
            elif kind == 'left':
                pi.append(i)
                bi.append(None)
        out = probe.take(pi)
        if kind in ('inner', 'left'):
            bt = build.take(bi)
            for nme, t, c in zip(bt.names, bt.types, bt.cols):
                if out.find(nme) < 0:
                    out.add(nme, t, c)
        return out

    def _is_fact_scan(self, node):
        from spark_rapids_benchmarks_b200 import nds_schema
        while node['op'] == 'join':
            node = node['probe']
        return node['op'] == 'scan' and node['table'] in nds_schema.TABLE_PARTITIONING

    # ---- aggregation
    def _agg_specs(self, aggs, b):
        specs = []
        for a in aggs:
            fn = a['fn']
            if fn == 'count_star':
                specs.append((a['name'], fn, None, INT64))
                continue
            t, f = compile_expr(a['expr'], b)
            if fn == 'count':
                ot = INT64
            elif t[0] == 'string':
                raise ValueError(fn + ' over strings is not supported')
            elif fn == 'sum':
                ot = DOUBLE if t[0] == 'double' else (('decimal', min(38, t[1] + 10), t[2]) if t[0] == 'decimal' else INT64)
            elif fn in ('min', 'max'):
                ot = _no32(t)
            else:
                raise ValueError("unsupported aggregate function '%s'" % fn)
            if 'type' in a:
                ot = _no32(parse_type(a['type']))
            specs.append((a['name'], fn, f, ot))
        return specs

    @staticmethod
    def _agg_init(specs):
        return [[0, None] for _ in specs]       # [count, state]

    @staticmethod
    def _agg_add(specs, st, i):
        for (name, fn, f, _), s in zip(specs, st):
            if fn == 'count_star':
                s[0] += 1
                continue
            v = f(i)
            if v is None:
                continue
            s[0] += 1
            if fn == 'sum':
                s[1] = v if s[1] is None else s[1] + v
            elif fn == 'min':
                s[1] = v if s[1] is None else min(s[1], v)
            elif fn == 'max':
                s[1] = v if s[1] is None else max(s[1], v)

    @staticmethod
    def _agg_value(spec, s):
        return s[0] if spec[1] in ('count', 'count_star') else (s[1] if s[0] > 0 else None)

    def aggregate_batch(self, node, b):
        keys = [(k['name'],) + compile_expr(k['expr'], b) for k in node['keys']]
        specs = self._agg_specs(node['aggs'], b)
        groups, order = {}, []
        for i in range(b.n):
            kv = tuple(f(i) for _, _, f in keys)
            st = groups.get(kv)
            if st is None:
                st = groups[kv] = self._agg_init(specs)
                order.append(kv)
            self._agg_add(specs, st, i)
        if not keys and not order:
            order.append(())
            groups[()] = self._agg_init(specs)
        out = Batch()
        out._n = len(order)
        for j, (name, t, _) in enumerate(keys):
            out.add(name, _out_type(t), [_out_val(t, kv[j]) for kv in order])
        for j, sp in enumerate(specs):
            out.add(sp[0], sp[3], [self._agg_value(sp, groups[kv][j]) for kv in order])
        return out

    def op_aggregate(self, node):
        return self.aggregate_batch(node, self.run(node['child']))

    def op_rollup(self, node):
        lv = [self.aggregate_batch(node['levels'][0], self.run(node['child']))]
        for spec in node['levels'][1:]:
            lv.append(self.aggregate_batch(spec, lv[-1]))
        self.ops.append('rollup_union')
        out = Batch()
        out._n = sum(x.n for x in lv)
        for c, name in enumerate(lv[0].names):
            vals = []
            for x in lv:
                ci = x.find(name)
                vals += x.cols[ci] if ci >= 0 else [None] * x.n
            out.add(name, lv[0].types[c], vals)
        return out

    def op_window(self, node):
        b = self.run(node['child'])
        self.ops.append('window')
        keys = [compile_expr(k['expr'], b)[1] for k in node['keys']]
        specs = self._agg_specs(node['aggs'], b)
        groups, kvs = {}, []
        for i in range(b.n):
            kv = tuple(f(i) for f in keys)
            kvs.append(kv)
            st = groups.get(kv)
            if st is None:
                st = groups[kv] = self._agg_init(specs)
            self._agg_add(specs, st, i)
        out = b.take(list(range(b.n)))
        for j, sp in enumerate(specs):
            out.add(sp[0], sp[3], [self._agg_value(sp, groups[kv][j]) for kv in kvs])
        return out

    # ---- ordering
    @staticmethod
    def _sort_idx(b, keys):
        comp = []
        for k in keys:
            t, f = compile_expr(k['expr'], b)
            desc = bool(k.get('desc', False))
            comp.append(([f(i) for i in range(b.n)], t, desc, bool(k.get('nulls_first', not desc))))

        def cmp(x, y):
            for vals, t, desc, nf in comp:
                p, q = vals[x], vals[y]
                if p is None or q is None:
                    if p is None and q is None:
                        continue
                    return (-1 if nf else 1) if p is None else (1 if nf else -1)
                if t[0] == 'string':
                    p, q = p.encode(), q.encode()
                if p != q:
                    r = -1 if p < q else 1
                    return -r if desc else r
            return -1 if x < y else (1 if x > y else 0)       # ties by row id: the order is total
        return sorted(range(b.n), key=functools.cmp_to_key(cmp)), comp

    def op_sort(self, node):
        b = self.run(node['child'])
        idx, _ = self._sort_idx(b, node['keys'])
        lim = node.get('limit', -1)
        if lim is not None and lim >= 0:
            idx = idx[:lim]
        out = b.take(idx)
        # like the engine: a sort materialises its columns (int32 stays int32, types unchanged)
        return out

    def op_rank(self, node):
        b = self.run(node['child'])
        self.ops.append('rank')
        idx, comp = self._sort_idx(b, node['keys'])
        npart = int(node.get('npart', 0))
        out = b.take(idx)
        rank, rown, dense = [], [], []
        ps = ts = 0
        d = 0
        for j, i in enumerate(idx):
            newp = j == 0 or any(comp[k][0][i] != comp[k][0][idx[j - 1]] for k in range(npart))
            newo = newp or any(comp[k][0][i] != comp[k][0][idx[j - 1]] for k in range(npart, len(comp)))
            if newp:
                ps, d = j, 0
            if newo:
                ts = j
                d += 1
            rank.append(ts - ps + 1)
            rown.append(j - ps + 1)
            dense.append(d)
        for f in node['funcs']:
            out.add(f['name'], INT64, {'rank': rank, 'row_number': rown, 'dense_rank': dense}[f['fn']])
        return out


def to_arrow(b):
    arrays, fields = [], []
    for name, t, col in zip(b.names, b.types, b.cols):
        at = _type_to_arrow(t)
        if t[0] == 'decimal':
            col = [None if v is None else decimal.Decimal(v).scaleb(-t[2]) for v in col]
            arrays.append(pa.array(col, type=at))
        elif t[0] == 'date':
            arrays.append(pa.array(col, type=pa.int32()).cast(pa.date32()))
        else:
            arrays.append(pa.array(col, type=at))
        fields.append(pa.field(name, at))
    if not arrays:
        return pa.RecordBatch.from_arrays([], schema=pa.schema([]))
    return pa.RecordBatch.from_arrays(arrays, schema=pa.schema(fields))


def execute(plan, warehouse, rank=0, world=1):
    """plan: {'root': node} or a node. Returns (pyarrow.RecordBatch, [operator names the tests look for])."""
    it = Interp(warehouse, rank, world)
    b = it.run(plan['root'] if 'root' in plan else plan)
    return to_arrow(b), it.ops


class FakeEngine:
    """Stands in for spark_rapids_benchmarks_b200.engine.Engine in the SQL test functions (CPU suite): same calls, the plan
    executed by this module.  `options` that select engine routes are accepted and ignored - every route must give these rows."""
    def __init__(self, warehouse):
        self.wh = warehouse
        self._ops = []
        self.has_comm = False

    def encode_plan(self, plan, source=None, profile=False, decode_timing=False, options=None):
        import json
        if isinstance(plan, (bytes, bytearray)):
            return bytes(plan)
        return json.dumps(plan if 'root' in plan else {'root': plan}).encode()

    def execute(self, plan, source=None, profile=False, decode_timing=False, options=None):
        import json
        if isinstance(plan, (bytes, bytearray, str)):
            plan = json.loads(plan)
        rb, ops = execute(plan, self.wh)
        self._ops = ops
        return rb

    def last_stats(self):
        return {'ops': [{'op': o, 'ms': 0.0, 'bytes': 0, 'rows': 0} for o in self._ops], 'launches': 0, 'device_ms': 0.0,
                'speculation_retries': 0}
