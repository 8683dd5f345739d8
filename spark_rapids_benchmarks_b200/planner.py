"""SQL AST -> physical plan IR for the B200 engine (star / snowflake join + aggregate + top-k).

Stands in for Catalyst, which the reference reaches through `spark_session.sql(query)`
(/root/reference/nds/nds_power.py:130).  The produced IR is the JSON the C ABI's
`nds_engine_execute` takes (DESIGN.md "Plan IR").  Shapes outside the supported subset raise
PlanError: an unsupported query fails loudly, it is never run elsewhere.

Planning rules (all result-preserving for inner equi-joins):
  * single-table conjuncts are pushed below the joins; col = col conjuncts between two tables
    become hash joins; other multi-table conjuncts run after the joins;
  * the hive-partitioned fact table is the probe side; every dimension (with its own
    snowflake children joined in first) is a unique-key build side;
  * joins are ordered most-selective first by a static estimate; the join on the fact table's
    partition column is marked `prune_partitions` (Spark's dynamic partition pruning) and goes last;
  * `t LEFT JOIN u ON t.a = u.b [AND <u-only conditions>]` becomes a left outer join above t's inner joins (WHERE
    conditions on u run after it); `[NOT] EXISTS (select .. from u where u.b = t.a [and ..])` and
    `t.a [NOT] IN (select b from u [where ..])` become semi / anti joins (NOT IN needs a NOT NULL subquery column and
    drops NULL t.a, as SQL's three-valued logic does); a build key that is not its table's primary key makes the join
    `unique_build: false` (repeated keys are chained and expand the probe rows);
  * SELECT DISTINCT is a GROUP BY over the select list without aggregates; HAVING is a filter above the aggregate;
  * AVG is planned as SUM and COUNT plus a final divide (decimal: scale + 4, HALF_UP - Spark's
    Average for DecimalType; integer: double);
  * ORDER BY keys are followed by every remaining output column as tie-breakers so that the
    LIMIT cut is deterministic.
"""
import datetime
import decimal
import json
import re

from . import nds_schema
from .sql import parse, SqlError


class PlanError(ValueError):
    pass


AGG_FUNCS = ('sum', 'avg', 'count', 'min', 'max')
_DEC = re.compile(r'decimal\((\d+),(\d+)\)')


def _col(n):
    return {'col': n}


def _fn(name, *args, **kw):
    d = {'fn': name, 'args': list(args)}
    d.update(kw)
    return d


def _engine_type(t):
    """schema type-string -> engine type-string"""
    if t in ('int32', 'int64', 'date', 'double', 'string', 'bool'):
        return t
    if _DEC.match(t):
        return t
    if t.startswith('char(') or t.startswith('varchar('):
        return 'string'
    raise PlanError('unsupported column type %r' % t)


def _is_num(t):
    return t in ('int32', 'int64') or bool(_DEC.match(t))


class Planner:
    def __init__(self, ast, catalog=None):
        self.q = ast
        self.col_types = catalog or nds_schema.column_types()
        self.col_table = nds_schema.column_table()
        self.alias = {}
        if any(e == ('star',) for e, _ in ast['select']):
            raise PlanError('SELECT * is only supported over a derived table')
        for name, alias in ast['from']:
            if isinstance(name, tuple):
                raise PlanError('a derived table joined with other tables is not supported yet')
            if name not in nds_schema.SOURCE_TABLES:
                raise PlanError('unknown table %r' % name)
            if alias in self.alias:
                raise PlanError('duplicate table alias %r (self joins are not supported yet)' % alias)
            self.alias[alias] = name
        self.tables = [t for t, _ in ast['from']]
        if len(set(self.tables)) != len(self.tables):
            raise PlanError('a table appears twice in FROM (self joins are not supported yet)')

    # ---- name resolution ----------------------------------------------------------------------------
    def resolve(self, node):
        _, qual, name = node
        if qual is not None:
            if qual not in self.alias:
                raise PlanError('unknown table alias %r' % qual)
            tbl = self.alias[qual]
            if self.col_table.get(name) != tbl:
                raise PlanError('column %r is not in table %r' % (name, tbl))
            return name
        tbl = self.col_table.get(name)
        if tbl is None or tbl not in self.tables:
            raise PlanError('unknown column %r' % name)
        return name

    def tables_of(self, ast, out=None):
        out = set() if out is None else out
        if not isinstance(ast, tuple):
            return out
        if ast[0] == 'col':
            out.add(self.col_table[self.resolve(ast)])
            return out
        for x in ast[1:]:
            if isinstance(x, tuple):
                self.tables_of(x, out)
            elif isinstance(x, list):
                for y in x:
                    if isinstance(y, tuple) and y and isinstance(y[0], str):
                        self.tables_of(y, out)
                    elif isinstance(y, tuple):
                        for z in y:
                            self.tables_of(z, out)
        return out

    def columns_of(self, ast, out=None):
        out = set() if out is None else out
        if not isinstance(ast, tuple):
            return out
        if ast[0] == 'col':
            out.add(self.resolve(ast))
            return out
        for x in ast[1:]:
            if isinstance(x, tuple):
                self.columns_of(x, out)
            elif isinstance(x, list):
                for y in x:
                    if isinstance(y, tuple) and y and isinstance(y[0], str):
                        self.columns_of(y, out)
                    elif isinstance(y, tuple):
                        for z in y:
                            self.columns_of(z, out)
        return out

    # ---- expression translation ------------------------------------------------------------------------
    def expr(self, ast, agg_map=None, key_map=None):
        """AST -> (IR, type-string). agg_map/key_map: replace aggregate calls / group expressions by
        references to aggregate output columns."""
        if key_map is not None:
            k = self.ast_key(ast)
            if k in key_map:
                name, t = key_map[k]
                return _col(name), t
        kind = ast[0]
        if kind == 'col':
            name = self.resolve(ast)
            if key_map is not None:
                raise PlanError('column %r must appear in GROUP BY or inside an aggregate' % name)
            return _col(name), _engine_type(self.col_types[name])
        if kind == 'num':
            txt = ast[1]
            if '.' in txt:
                d = decimal.Decimal(txt)
                scale = -d.as_tuple().exponent
                unscaled = int(d.scaleb(scale))
                t = 'decimal(%d,%d)' % (max(len(str(abs(unscaled))), scale, 1), scale)
                return {'lit': unscaled, 'type': t}, t
            return {'lit': int(txt), 'type': 'int64'}, 'int64'
        if kind == 'str':
            return {'lit': ast[1], 'type': 'string'}, 'string'
        if kind == 'null':
            return {'lit': None, 'type': 'int64'}, 'int64'
        if kind == 'neg':
            x, t = self.expr(ast[1], agg_map, key_map)
            if 'lit' in x and isinstance(x['lit'], int):
                return {'lit': -x['lit'], 'type': x['type']}, t
            return _fn('neg', x), t
        if kind == 'not':
            x, _ = self.expr(ast[1], agg_map, key_map)
            return _fn('not', x), 'bool'
        if kind == 'isnull':
            x, _ = self.expr(ast[1], agg_map, key_map)
            return _fn('isnotnull' if ast[2] else 'isnull', x), 'bool'
        if kind == 'between':
            x, _ = self.expr(ast[1], agg_map, key_map)
            lo, _ = self.expr(ast[2], agg_map, key_map)
            hi, _ = self.expr(ast[3], agg_map, key_map)
            r = _fn('and', _fn('ge', x, lo), _fn('le', x, hi))
            return (_fn('not', r) if ast[4] else r), 'bool'
        if kind == 'in':
            x, _ = self.expr(ast[1], agg_map, key_map)
            items = [self.expr(i, agg_map, key_map)[0] for i in ast[2]]
            r = _fn('or', *[_fn('eq', x, i) for i in items]) if len(items) > 1 else _fn('eq', x, items[0])
            return (_fn('not', r) if ast[3] else r), 'bool'
        if kind == 'case':
            other = self.expr(ast[2], agg_map, key_map) if ast[2] is not None else ({'lit': None, 'type': 'int64'}, 'int64')
            res, t = other
            for cond, val in reversed(ast[1]):
                c, _ = self.expr(cond, agg_map, key_map)
                v, tv = self.expr(val, agg_map, key_map)
                if 'lit' in res and res['lit'] is None:          # ELSE NULL / no ELSE: a NULL of the branch's type
                    res = {'lit': None, 'type': tv}
                elif 'lit' in v and v['lit'] is None:
                    v, tv = {'lit': None, 'type': t}, t
                res, t = _fn('if', c, v, res), tv
            return res, t
        if kind == 'cast':
            tname = ast[2].lower()
            if tname == 'date' and ast[1][0] == 'str':
                d = datetime.date.fromisoformat(ast[1][1])
                return {'lit': (d - datetime.date(1970, 1, 1)).days, 'type': 'date'}, 'date'
            x, t = self.expr(ast[1], agg_map, key_map)
            if tname in ('double', 'float'):
                return _fn('cast_double', x), 'double'
            raise PlanError('CAST to %r is not supported yet' % tname)
        if kind == 'bin':
            op = ast[1]
            if op in ('and', 'or'):
                l, _ = self.expr(ast[2], agg_map, key_map)
                r, _ = self.expr(ast[3], agg_map, key_map)
                return _fn(op, l, r), 'bool'
            if op in ('add', 'sub') and isinstance(ast[3], tuple) and ast[3][0] == 'interval':
                # date literal +/- INTERVAL n DAYS: folded (the NDS templates only add intervals to constants)
                l, tl = self.expr(ast[2], agg_map, key_map)
                if tl != 'date' or 'lit' not in l or l['lit'] is None:
                    raise PlanError('INTERVAL arithmetic is only supported on date constants')
                return {'lit': l['lit'] + (ast[3][1] if op == 'add' else -ast[3][1]), 'type': 'date'}, 'date'
            l, tl = self.expr(ast[2], agg_map, key_map)
            r, tr = self.expr(ast[3], agg_map, key_map)
            if op in ('eq', 'ne', 'lt', 'le', 'gt', 'ge'):
                if (tl == 'string') != (tr == 'string'):
                    raise PlanError('comparison between string and non-string')
                return _fn(op, l, r), 'bool'
            if op in ('add', 'sub', 'mul'):
                if tl == 'string' or tr == 'string':
                    raise PlanError('arithmetic on strings')
                if tl == 'double' or tr == 'double':
                    return _fn(op, l, r), 'double'
                ml, mr = _DEC.match(tl), _DEC.match(tr)
                if ml or mr:
                    pl, sl = (int(ml.group(1)), int(ml.group(2))) if ml else (20, 0)
                    pr, sr = (int(mr.group(1)), int(mr.group(2))) if mr else (20, 0)
                    if op == 'mul':
                        return _fn(op, l, r), 'decimal(%d,%d)' % (min(38, pl + pr + 1), sl + sr)
                    s = max(sl, sr)
                    return _fn(op, l, r), 'decimal(%d,%d)' % (min(38, max(pl - sl, pr - sr) + s + 1), s)
                return _fn(op, l, r), 'int64'
            if op == 'div':
                return _fn('div', l, r), 'double'
            raise PlanError('operator %r' % op)
        if kind == 'like':
            # x LIKE 'abc%' (a prefix) = substr(x, 1, 3) = 'abc'; x LIKE 'abc' (no wildcard) = equality.  Other patterns raise.
            x, t = self.expr(ast[1], agg_map, key_map)
            pat = ast[2]
            if t != 'string':
                raise PlanError('LIKE over a non-string')
            body = pat[:-1] if pat.endswith('%') else pat
            if '%' in body or '_' in body or '\\' in body or not body:
                raise PlanError('LIKE pattern %r is not supported yet (a literal prefix followed by %% is)' % pat)
            lhs = _fn('substr', x, start=1, len=len(body)) if pat.endswith('%') else x
            r = _fn('eq', lhs, {'lit': body, 'type': 'string'})
            return (_fn('not', r) if ast[3] else r), 'bool'
        if kind == 'interval':
            raise PlanError('INTERVAL is only supported as date constant +/- INTERVAL n DAYS')
        if kind == 'wcall':
            return self.win_ref(ast, agg_map, key_map)
        if kind == 'call':
            fname = ast[1]
            if fname in AGG_FUNCS:
                if agg_map is None:
                    raise PlanError('aggregate %s() is not allowed here' % fname)
                return self.agg_ref(ast, agg_map)
            if fname in ('substr', 'substring'):
                x, t = self.expr(ast[2][0], agg_map, key_map)
                if t != 'string':
                    raise PlanError('substr over a non-string')
                start = self.const_int(ast[2][1])
                ln = self.const_int(ast[2][2]) if len(ast[2]) > 2 else (1 << 30)
                return _fn('substr', x, start=start, len=ln), 'string'
            if fname == 'coalesce' and len(ast[2]) >= 2:
                # coalesce(a, b, ...) = if(a is not null, a, coalesce(b, ...)); the arguments must be of one type family
                parts = [self.expr(a, agg_map, key_map) for a in ast[2]]
                if any(tv == 'string' for _, tv in parts):
                    raise PlanError('coalesce() over strings is not supported yet (a projected string must come from one column)')
                res, t = parts[-1]
                for v, tv in reversed(parts[:-1]):
                    if (tv == 'string') != (t == 'string') or (tv == 'double') != (t == 'double'):
                        raise PlanError('coalesce() over different types')
                    if 'lit' in res and res['lit'] is None:
                        res = {'lit': None, 'type': tv}
                    res, t = _fn('if', _fn('isnotnull', v), v, res), tv
                return res, t
            if fname == 'abs' and len(ast[2]) == 1:
                x, t = self.expr(ast[2][0], agg_map, key_map)
                if not _is_num(t) and t != 'double':
                    raise PlanError('abs() over %s' % t)
                return _fn('if', _fn('lt', x, {'lit': 0, 'type': 'int64'}), _fn('neg', x), x), t
            raise PlanError('function %s() is not supported yet' % fname)
        raise PlanError('unsupported expression %r' % (kind,))

    def const_int(self, ast):
        if ast[0] == 'num' and '.' not in ast[1]:
            return int(ast[1])
        raise PlanError('expected an integer constant')

    def ast_key(self, ast):
        """structural key with resolved column names (dt.d_year == d_year)"""
        if isinstance(ast, tuple):
            if ast and ast[0] == 'col':
                return ('col', self.resolve(ast))
            return tuple(self.ast_key(x) for x in ast)
        if isinstance(ast, list):
            return tuple(self.ast_key(x) for x in ast)
        return ast

    # ---- aggregates ----------------------------------------------------------------------------------------
    def agg_ref(self, ast, agg_map):
        fname, args, distinct = ast[1], ast[2], ast[3]
        if distinct:
            # count(DISTINCT x): planned as two aggregates - group by (keys, x) first, then count x per key (plan())
            if fname != 'count' or len(args) != 1 or args[0] == ('star',):
                raise PlanError('%s(DISTINCT ...) is not supported yet (count(DISTINCT x) is)' % fname)
            x, _ = self.expr(args[0])
            return _col(agg_map.add('count_distinct', x, 'int64')), 'int64'
        if fname == 'count' and (not args or args[0] == ('star',)):
            name = agg_map.add('count_star', None, 'int64')
            return _col(name), 'int64'
        if len(args) != 1:
            raise PlanError('%s() takes one argument' % fname)
        x, t = self.expr(args[0])
        if fname == 'count':
            return _col(agg_map.add('count', x, 'int64')), 'int64'
        if not _is_num(t) and t != 'double':
            raise PlanError('%s() over %s is not supported' % (fname, t))
        m = _DEC.match(t)
        if fname in ('min', 'max'):
            rt = t if t != 'int32' else 'int64'
            return _col(agg_map.add(fname, x, rt)), rt
        sum_t = 'decimal(%d,%s)' % (min(38, int(m.group(1)) + 10), m.group(2)) if m else ('double' if t == 'double' else 'int64')
        sname = agg_map.add('sum', x, sum_t)
        if fname == 'sum':
            return _col(sname), sum_t
        cname = agg_map.add('count', x, 'int64')
        if m:   # Spark Average(DecimalType(p,s)) -> DecimalType(p+4, s+4)
            rt = 'decimal(%d,%d)' % (min(38, int(m.group(1)) + 4), int(m.group(2)) + 4)
            return _fn('decavg', _col(sname), _col(cname), up=4, type=rt), rt
        if t == 'double':
            return _fn('div', _col(sname), _col(cname)), 'double'
        return _fn('divi2f', _col(sname), _col(cname)), 'double'

    # ---- window aggregates: fn(x) OVER (PARTITION BY keys) ---------------------------------------------------
    def win_ref(self, ast, agg_map, key_map):
        """The argument and the partition keys are expressions over the rows the window sees: the groups of an aggregating
        query (sum(sum(x)) over (partition by k): the inner sum is an ordinary aggregate), else the joined rows."""
        fname, args, part, order = ast[1], ast[2], ast[3], ast[4]
        wm = getattr(self, '_win', None)
        if wm is None:
            raise PlanError('window function %s() is not allowed here (select list only)' % fname)
        if fname in ('rank', 'row_number', 'dense_rank'):
            # ranking functions: the rows sorted by (partition keys, order keys), positions from neighbour comparisons
            if args or not order:
                raise PlanError('%s() takes no argument and needs OVER (... ORDER BY ...)' % fname)
            self._win = None
            try:
                pirs = [self.expr(p, agg_map, key_map)[0] for p in part]
                okeys = []
                for e, desc, nf in order:
                    k = {'expr': self.expr(e, agg_map, key_map)[0], 'desc': bool(desc)}
                    if nf is not None:
                        k['nulls_first'] = bool(nf)
                    okeys.append(k)
            finally:
                self._win = wm
            return _col(wm.add_rank(pirs, okeys, fname)), 'int64'
        if fname not in ('sum', 'avg', 'min', 'max', 'count'):
            raise PlanError('window function %s() is not supported yet' % fname)
        if order:
            raise PlanError('aggregate window functions with ORDER BY (running totals) are not supported yet')
        if not part:
            raise PlanError('a window without PARTITION BY is not supported yet')
        self._win = None                     # no window inside a window's argument
        try:
            pirs = [self.expr(p, agg_map, key_map)[0] for p in part]
            if fname == 'count' and (not args or args[0] == ('star',)):
                return _col(wm.add(pirs, 'count_star', None, 'int64')), 'int64'
            if len(args) != 1:
                raise PlanError('%s() takes one argument' % fname)
            x, t = self.expr(args[0], agg_map, key_map)
        finally:
            self._win = wm
        if fname == 'count':
            return _col(wm.add(pirs, 'count', x, 'int64')), 'int64'
        if not _is_num(t) and t != 'double':
            raise PlanError('%s() over %s is not supported' % (fname, t))
        m = _DEC.match(t)
        if fname in ('min', 'max'):
            rt = t if t != 'int32' else 'int64'
            return _col(wm.add(pirs, fname, x, rt)), rt
        sum_t = 'decimal(%d,%s)' % (min(38, int(m.group(1)) + 10), m.group(2)) if m else ('double' if t == 'double' else 'int64')
        sname = wm.add(pirs, 'sum', x, sum_t)
        if fname == 'sum':
            return _col(sname), sum_t
        cname = wm.add(pirs, 'count', x, 'int64')
        if m:
            rt = 'decimal(%d,%d)' % (min(38, int(m.group(1)) + 4), int(m.group(2)) + 4)
            return _fn('decavg', _col(sname), _col(cname), up=4, type=rt), rt
        if t == 'double':
            return _fn('div', _col(sname), _col(cname)), 'double'
        return _fn('divi2f', _col(sname), _col(cname)), 'double'

    # ---- SQL text of an expression, for unaliased output names (Spark style) ----------------------------
    def text(self, ast):
        k = ast[0]
        if k == 'col':
            return self.resolve(ast)
        if k == 'num':
            return ast[1]
        if k == 'str':
            return ast[1]
        if k == 'call':
            return '%s(%s)' % (ast[1], ', '.join('*' if a == ('star',) else self.text(a) for a in ast[2]))
        if k == 'bin':
            sym = {'add': '+', 'sub': '-', 'mul': '*', 'div': '/', 'eq': '=', 'ne': '<>', 'lt': '<', 'le': '<=', 'gt': '>',
                   'ge': '>=', 'and': 'AND', 'or': 'OR'}[ast[1]]
            return '(%s %s %s)' % (self.text(ast[2]), sym, self.text(ast[3]))
        if k == 'neg':
            return '(- %s)' % self.text(ast[1])
        return k

    # ---- the plan ---------------------------------------------------------------------------------------------
    def plan(self, distributed=False, sharded=None):
        """distributed: plan partial aggregate -> exchange -> merge for engines that each hold a shard of the root table.
        sharded: names of the tables the engines really hold in shards (None: the hive-partitioned fact tables of
        TABLE_PARTITIONING).  A query whose root table is replicated on every rank is planned as a single-GPU plan
        and marked 'replicated': every rank computes the whole result and nothing is exchanged - summing the ranks'
        identical partial states would multiply every SUM and COUNT by the world size."""
        q = self.q
        # ORDER BY expressions may use select aliases (Spark resolves them against the select list first): an alias INSIDE an
        # expression is replaced by the select expression it names; a bare alias stays a reference to the output column
        alias_ast = {a: e for e, a in q['select'] if a and self.col_table.get(a) not in self.tables}

        def subst(a):
            if not isinstance(a, tuple):
                return a
            if a[0] == 'col' and a[1] is None and a[2] in alias_ast:
                return alias_ast[a[2]]
            return tuple([subst(y) if isinstance(y, tuple) else y for y in x] if isinstance(x, list) else subst(x) for x in a)
        if alias_ast:
            q = dict(q, order=[((e if (e[0] == 'col' and e[1] is None) else subst(e)), d, nf) for e, d, nf in q['order']])
            self.q = q
        conj = []

        def split(e):
            if e is None:
                return
            if e[0] == 'bin' and e[1] == 'and':
                split(e[2])
                split(e[3])
            else:
                conj.append(e)
        split(q['where'])
        edges, table_filters, post = [], {}, []
        # ---- subquery conjuncts -> semi / anti joins: (outer column, subquery table, its key column, its filters, type)
        sub_joins = []
        plain = []
        for c in conj:
            if c[0] in ('exists', 'in_sub'):
                sub_joins.append(self.plan_subquery(c))
            else:
                plain.append(c)
        conj = plain
        # ---- LEFT JOINs: (outer column, joined table, its key column, its ON-clause filters)
        left_joins = []
        left_tables = set()
        for alias, on in q.get('left_joins', []):
            tl = self.alias[alias]
            parts = []

            def split_on(e):
                if e[0] == 'bin' and e[1] == 'and':
                    split_on(e[2])
                    split_on(e[3])
                else:
                    parts.append(e)
            split_on(on)
            key, flt = None, []
            for c in parts:
                tabs = self.tables_of(c)
                if c[0] == 'bin' and c[1] == 'eq' and c[2][0] == 'col' and c[3][0] == 'col' and len(tabs) == 2 and tl in tabs:
                    a, b = self.resolve(c[2]), self.resolve(c[3])
                    if key is not None:
                        raise PlanError('LEFT JOIN %s: more than one equality in ON is not supported yet' % tl)
                    key = (b, a) if self.col_table[a] == tl else (a, b)
                elif tabs == {tl}:
                    flt.append(c)
                else:
                    raise PlanError('LEFT JOIN %s: ON must be one equality with an earlier table plus conditions on %s only' % (tl, tl))
            if key is None:
                raise PlanError('LEFT JOIN %s: ON needs an equality with an earlier table' % tl)
            if self.col_table[key[0]] in left_tables:
                raise PlanError('LEFT JOIN %s: joining to another left-joined table is not supported yet' % tl)
            left_joins.append((key[0], tl, key[1], flt))
            left_tables.add(tl)
        for c in conj:
            tabs = self.tables_of(c)
            if tabs & left_tables:
                post.append(c)          # WHERE conditions over the NULL-extended side run after the outer join
                continue
            if (c[0] == 'bin' and c[1] == 'eq' and c[2][0] == 'col' and c[3][0] == 'col' and len(tabs) == 2):
                a, b = self.resolve(c[2]), self.resolve(c[3])
                edges.append((self.col_table[a], a, self.col_table[b], b))
            elif len(tabs) == 1:
                table_filters.setdefault(next(iter(tabs)), []).append(c)
            elif len(tabs) == 0:
                raise PlanError('constant predicates are not supported')
            else:
                post.append(c)
        # ---- root (fact) table
        facts = [t for t in self.tables if t in nds_schema.TABLE_PARTITIONING and t not in left_tables]
        if [t for t in left_tables if t in nds_schema.TABLE_PARTITIONING] and facts:
            raise PlanError('joins between fact tables are not supported yet')
        if len(facts) > 1:
            raise PlanError('joins between fact tables (%s) are not supported yet' % ', '.join(facts))
        root = facts[0] if facts else self.tables[0]
        replicated_root = distributed and root not in (nds_schema.TABLE_PARTITIONING if sharded is None else sharded)
        if replicated_root:
            distributed = False
        # ---- columns each table must produce
        need = {t: set() for t in self.tables}
        used = set()
        for e, _ in q['select']:
            self.columns_of(e, used)
        for e in q['group']:
            self.columns_of(e, used)
        select_aliases = {a for _, a in q['select'] if a}
        for e, _, _ in q['order']:
            if e[0] == 'col' and e[1] is None and e[2] in select_aliases:
                continue
            self.columns_of(e, used)
        for c in conj:
            self.columns_of(c, used)
        for okey, tl, ikey, flt in left_joins:
            used.update([okey, ikey])
            for c in flt:
                self.columns_of(c, used)
        for sj in sub_joins:
            used.add(sj['outer'])
        if q.get('having') is not None:
            self.columns_of(q['having'], used)
        for c in used:
            need[self.col_table[c]].add(c)
        # ---- join tree
        children = {t: [] for t in self.tables}
        visited = {root}
        frontier = [root]
        remaining = list(edges)
        while frontier:
            cur = frontier.pop(0)
            for e in list(remaining):
                ta, ca, tb, cb = e
                if ta == cur and tb not in visited:
                    children[cur].append((ca, tb, cb))
                elif tb == cur and ta not in visited:
                    children[cur].append((cb, ta, ca))
                else:
                    continue
                nxt = children[cur][-1][1]
                visited.add(nxt)
                frontier.append(nxt)
                remaining.remove(e)
        for t in self.tables:
            if t not in visited and t not in left_tables:
                raise PlanError('table %r is not connected by an equi-join (cross joins are not supported)' % t)
        for ta, ca, tb, cb in remaining:       # extra equalities between already-joined tables
            post.append(('bin', 'eq', ('col', None, ca), ('col', None, cb)))

        # columns a table's scan must hand on as data: everything used outside the equi-join conditions that became joins.
        # A probe key used for nothing else is 'key_only': the fused scan probes with it and never materialises it.
        payload = set()
        for e, _ in q['select']:
            self.columns_of(e, payload)
        for e in q['group']:
            self.columns_of(e, payload)
        for e, _, _ in q['order']:
            if not (e[0] == 'col' and e[1] is None and e[2] in select_aliases):
                self.columns_of(e, payload)
        for flt in table_filters.values():
            for c in flt:
                self.columns_of(c, payload)
        for c in post:
            self.columns_of(c, payload)
        for okey, tl, ikey, flt in left_joins:
            payload.add(okey)
        for sj in sub_joins:
            payload.add(sj['outer'])
        if q.get('having') is not None:
            self.columns_of(q['having'], payload)

        def is_primary_key(table, column):
            return nds_schema.SOURCE_TABLES[table][0][0] == column and table not in nds_schema.TABLE_PARTITIONING

        def selectivity(t):
            s = 1.0
            for c in table_filters.get(t, []):
                s *= self.pred_selectivity(c)
            for _, ct, _ in children[t]:
                s *= min(1.0, selectivity(ct) * 1.5)
            return s

        def subtree(t):
            cols = sorted(need[t], key=lambda c: [n for n, _, _ in nds_schema.SOURCE_TABLES[t]].index(c))
            node = {'op': 'scan', 'table': t, 'columns': cols}
            key_only = [c for c in cols if c not in payload and any(pk == c for pk, _, _ in children[t])]
            if key_only:
                node['key_only'] = key_only
            flt = table_filters.get(t, [])
            if flt:
                irs = [self.expr(c)[0] for c in flt]
                node = {'op': 'filter', 'cond': irs[0] if len(irs) == 1 else _fn('and', *irs), 'child': node}
            pcol = nds_schema.TABLE_PARTITIONING.get(t)
            joins = []
            for pk, ct, ck in children[t]:
                prune = pk == pcol
                joins.append((2.0 if prune else selectivity(ct), pk, ct, ck, prune))
            joins.sort(key=lambda j: j[0])
            for _, pk, ct, ck, prune in joins:
                node = {'op': 'join', 'type': 'inner', 'probe': node, 'build': subtree(ct), 'probe_key': pk, 'build_key': ck}
                if prune:
                    node['prune_partitions'] = True
                if not is_primary_key(ct, ck):
                    node['unique_build'] = False
            for okey, tl, ikey, lflt in left_joins:
                if self.col_table[okey] != t:
                    continue
                bcols = sorted(need[tl], key=lambda c: [n for n, _, _ in nds_schema.SOURCE_TABLES[tl]].index(c))
                build = {'op': 'scan', 'table': tl, 'columns': bcols}
                if lflt:
                    irs = [self.expr(c)[0] for c in lflt]
                    build = {'op': 'filter', 'cond': irs[0] if len(irs) == 1 else _fn('and', *irs), 'child': build}
                node = {'op': 'join', 'type': 'left', 'probe': node, 'build': build, 'probe_key': okey, 'build_key': ikey}
                if not is_primary_key(tl, ikey):
                    node['unique_build'] = False
            for sj in sub_joins:
                if self.col_table[sj['outer']] != t:
                    continue
                if sj['drop_null_outer']:
                    node = {'op': 'filter', 'cond': _fn('isnotnull', _col(sj['outer'])), 'child': node}
                node = {'op': 'join', 'type': sj['type'], 'probe': node, 'build': sj['build'], 'probe_key': sj['outer'], 'build_key': sj['key']}
            return node
        node = subtree(root)
        if post:
            irs = [self.expr(c)[0] for c in post]
            node = {'op': 'filter', 'cond': irs[0] if len(irs) == 1 else _fn('and', *irs), 'child': node}

        # ---- aggregation
        def has_agg(a):
            if not isinstance(a, tuple):
                return False
            if a[0] == 'call' and a[1] in AGG_FUNCS:
                return True
            return any(has_agg(x) if isinstance(x, tuple) else any(has_agg(y) for y in x if isinstance(y, tuple))
                       for x in a[1:] if isinstance(x, (tuple, list)))
        is_agg = bool(q['group']) or any(has_agg(e) for e, _ in q['select'])
        group_asts = list(q['group'])
        if q.get('distinct'):
            # SELECT DISTINCT a, b == SELECT a, b GROUP BY a, b (no aggregates, no GROUP BY of its own)
            if is_agg:
                raise PlanError('SELECT DISTINCT together with GROUP BY / aggregates is not supported yet')
            is_agg = True
            group_asts = [e for e, _ in q['select']]
        if q.get('having') is not None and not is_agg:
            raise PlanError('HAVING without GROUP BY / aggregates')
        out_cols = []          # [(name, ir, type)]
        names = []
        for e, alias in q['select']:
            names.append(alias if alias else self.text(e))
        final_partial = None
        if is_agg:
            key_map, keys = {}, []
            for i, g in enumerate(group_asts):
                ir, t = self.expr(g)
                name = ir['col'] if 'col' in ir else '_gk%d' % i
                key_map[self.ast_key(g)] = (name, t)
                keys.append({'name': name, 'expr': ir, '_type': t})
            agg_map = _AggMap()
            self._win = win_map = _WinMap()
            try:
                sel_ir = [self.expr(e, agg_map, key_map) for e, _ in q['select']]
            finally:
                self._win = None
            having_ir = self.expr(q['having'], agg_map, key_map)[0] if q.get('having') is not None else None
            order_extra = []
            for e, desc, nf in q['order']:
                order_extra.append(None if (e[0] == 'col' and e[1] is None and e[2] in select_aliases)
                                   else self.expr(e, agg_map, key_map))
            specs = agg_map.specs()
            dist = [a for a in specs if a['fn'] == 'count_distinct']
            if dist and q.get('rollup'):
                raise PlanError('count(DISTINCT ...) under GROUP BY ROLLUP is not supported yet')
            if dist:
                # count(DISTINCT x) [+ other aggregates]: the inner aggregate groups by (keys, x) and carries the other
                # aggregates as partial states; the outer one counts x per key (NULL x is not counted) and merges the states
                if distributed:
                    raise PlanError('count(DISTINCT ...) is not supported across GPUs yet')
                if len({repr(a['expr']) for a in dist}) != 1:
                    raise PlanError('count(DISTINCT ...) over more than one expression is not supported yet')
                inner_aggs = [a for a in specs if a['fn'] != 'count_distinct']
                inner = {'op': 'aggregate',
                         'keys': [{'name': k['name'], 'expr': k['expr']} for k in keys] + [{'name': '_dx', 'expr': dist[0]['expr']}],
                         'aggs': inner_aggs, 'child': node}
                outer_aggs = []
                for a in specs:
                    if a['fn'] == 'count_distinct':
                        outer_aggs.append({'name': a['name'], 'fn': 'count', 'expr': _col('_dx')})
                    else:
                        outer_aggs.append({'name': a['name'], 'fn': {'count': 'sum', 'count_star': 'sum'}.get(a['fn'], a['fn']),
                                           'expr': _col(a['name']), 'type': agg_map.types[a['name']]})
                agg_node = {'op': 'aggregate', 'keys': [{'name': k['name'], 'expr': _col(k['name'])} for k in keys],
                            'aggs': outer_aggs, 'child': inner}
            elif q.get('rollup'):
                # GROUP BY ROLLUP(k1 .. kn): one engine operator runs the finest aggregate over the child, then every coarser
                # level over the level before it (SUM of sums, SUM of counts, MIN of mins ...), and concatenates the levels
                # with NULL for the rolled-up keys.  The grand total groups by a constant instead of by nothing: like
                # Spark's Expand + Aggregate it yields no row when there is no input row.
                if distributed:
                    raise PlanError('GROUP BY ROLLUP is not supported across GPUs yet')
                merged = [{'name': a['name'], 'fn': {'count': 'sum', 'count_star': 'sum'}.get(a['fn'], a['fn']),
                           'expr': _col(a['name']), 'type': agg_map.types[a['name']]} for a in specs]
                levels = [{'keys': [{'name': k['name'], 'expr': k['expr']} for k in keys], 'aggs': specs}]
                for depth in range(len(keys) - 1, -1, -1):
                    lkeys = [{'name': k['name'], 'expr': _col(k['name'])} for k in keys[:depth]]
                    if not lkeys:
                        lkeys = [{'name': '_rollup_all', 'expr': {'lit': 0, 'type': 'int64'}}]
                    levels.append({'keys': lkeys, 'aggs': merged})
                agg_node = {'op': 'rollup', 'levels': levels, 'child': node}
            else:
                agg_node = {'op': 'aggregate', 'keys': [{'name': k['name'], 'expr': k['expr']} for k in keys],
                            'aggs': specs, 'child': node}
            if distributed:
                final_partial = agg_node
                merge = {'op': 'aggregate',
                         'keys': [{'name': k['name'], 'expr': _col(k['name'])} for k in keys],
                         # 'type': the declared result type, so that the merged SUM of partial sums is not widened twice
                         'aggs': [{'name': a['name'], 'fn': {'count': 'sum', 'count_star': 'sum'}.get(a['fn'], a['fn']),
                                   'expr': _col(a['name']), 'type': agg_map.types[a['name']]} for a in agg_map.specs()],
                         'child': {'op': 'scan', 'table': '__partial__',
                                   'columns': [k['name'] for k in keys] + [a['name'] for a in agg_map.specs()]}}
                node = merge
            else:
                node = agg_node
            if having_ir is not None:        # HAVING: a filter over the (merged) groups
                node = {'op': 'filter', 'cond': having_ir, 'child': node}
            if win_map.parts or win_map.ranks:                # window functions see the groups that are left
                if q.get('rollup'):
                    raise PlanError('window functions over GROUP BY ROLLUP are not supported yet')
                node = win_map.wrap(node)
            for (ir, t), name in zip(sel_ir, names):
                out_cols.append((name, ir, t))
        else:
            self._win = win_map = _WinMap()
            try:
                for (e, _), name in zip(q['select'], names):
                    ir, t = self.expr(e)
                    out_cols.append((name, ir, t))
            finally:
                self._win = None
            if win_map.parts or win_map.ranks:
                if distributed:
                    raise PlanError('window functions are not supported across GPUs yet')
                node = win_map.wrap(node)
            order_extra = [None if (e[0] == 'col' and e[1] is None and e[2] in select_aliases) else self.expr(e)
                           for e, _, _ in q['order']]
        # ---- order by / limit
        hidden = []                # ORDER BY expressions that are no output column: computed beside the outputs, dropped after the sort
        order_names = []
        for (e, desc, nf), extra in zip(q['order'], order_extra):
            if extra is None:
                order_names.append(e[2])
                continue
            ir, t = extra
            match = [n for n, oir, _ in out_cols if oir == ir] + [n for n, oir, _ in hidden if oir == ir]
            if not match:
                if q.get('distinct'):
                    raise PlanError('ORDER BY expression %s is not in the select list of SELECT DISTINCT' % self.text(e))
                match = ['_ord%d' % len(hidden)]
                hidden.append((match[0], ir, t))
            order_names.append(match[0])
        node = {'op': 'project', 'exprs': [{'name': n, 'expr': ir} for n, ir, _ in out_cols + hidden], 'child': node}
        if q['order'] or q['limit'] is not None:
            keys = []
            used_names = []
            for (e, desc, nf), name in zip(q['order'], order_names):
                k = {'expr': _col(name), 'desc': bool(desc)}
                if nf is not None:
                    k['nulls_first'] = bool(nf)
                keys.append(k)
                used_names.append(name)
            # deterministic tie-break: the remaining output columns - unless the ORDER BY keys already contain
            # every GROUP BY key, in which case no two result rows can tie
            total = False
            if is_agg and group_asts and not q.get('rollup'):      # (a rolled-up row ties with the NULL group of its key)
                key_outs = []
                for kname, _ in key_map.values():
                    key_outs.append([n for n, oir, _ in out_cols if oir == _col(kname)])
                total = all(any(n in used_names for n in outs) for outs in key_outs)
            for n, _, _ in out_cols:
                if n not in used_names and not total:
                    keys.append({'expr': _col(n), 'desc': False})
            node = {'op': 'sort', 'keys': keys, 'limit': -1 if q['limit'] is None else q['limit'], 'child': node}
        if hidden:
            if distributed:
                raise PlanError('ORDER BY expressions outside the select list are not supported across GPUs yet')
            node = {'op': 'project', 'exprs': [{'name': n, 'expr': _col(n)} for n, _, _ in out_cols], 'child': node}
        out = {'root': node, 'output': [{'name': n, 'type': t} for n, _, t in out_cols]}
        if replicated_root:
            out['replicated'] = True
        if distributed:
            if final_partial is None:
                raise PlanError('distributed execution needs an aggregating query')
            out['partial'] = final_partial
            # one plan for engines that share a communicator: the scan of the gathered states becomes an exchange node
            out['fused'] = _replace_partial_scan(node, {'op': 'exchange', 'child': final_partial})
            # the same with Spark's own shape - HashAggregate(partial) -> Exchange(hashpartitioning(keys)) -> HashAggregate(final):
            # the partial states are hash-partitioned by the group keys (all-to-all), every rank merges and ranks ITS groups
            # only, and the per-rank results (at most LIMIT rows each) are gathered and ranked once more
            gkeys = [k['name'] for k in final_partial['keys']]
            # (not with window functions: a partition's groups would be spread over the ranks - the gather form has them all)
            if gkeys and not (win_map.parts or win_map.ranks):
                local = _replace_partial_scan(node, {'op': 'exchange', 'mode': 'hash', 'keys': gkeys, 'child': final_partial})
                top = {'op': 'exchange', 'child': local}
                if node['op'] == 'sort':
                    top = {'op': 'sort', 'keys': node['keys'], 'limit': node['limit'], 'child': top}
                out['fused_hash'] = top
        return out

    def plan_subquery(self, c):
        """('exists', q2, negated) | ('in_sub', x, q2, negated) -> {'outer', 'key', 'build', 'type', 'drop_null_outer'}.
        The subquery reads one table that the outer query does not; EXISTS correlates it with one equality."""
        q2 = c[1] if c[0] == 'exists' else c[2]
        negated = c[2] if c[0] == 'exists' else c[3]
        if len(q2['from']) != 1 or q2['group'] or q2.get('having') or q2['order'] or q2['limit'] is not None or q2.get('left_joins'):
            raise PlanError('subqueries are supported over one table, without GROUP BY / ORDER BY / LIMIT')
        t2, a2 = q2['from'][0]
        if t2 not in nds_schema.SOURCE_TABLES:
            raise PlanError('unknown table %r' % t2)
        if t2 in self.tables:
            raise PlanError('a subquery over a table of the outer query (%s) is not supported yet' % t2)
        inner_cols = {n for n, _, _ in nds_schema.SOURCE_TABLES[t2]}

        def res(node):               # column of the subquery's table, or of the outer query
            _, qual, name = node
            if qual is not None and qual == a2:
                if name not in inner_cols:
                    raise PlanError('column %r is not in table %r' % (name, t2))
                return name, True
            if qual is None and name in inner_cols:
                return name, True
            return self.resolve(node), False
        parts = []

        def split(e):
            if e is None:
                return
            if e[0] == 'bin' and e[1] == 'and':
                split(e[2])
                split(e[3])
            else:
                parts.append(e)
        split(q2['where'])
        sub = Planner({'from': [(t2, a2)], 'select': [], 'where': None, 'group': [], 'order': [], 'limit': None})
        key, outer, flt = None, None, []
        for e in parts:
            if e[0] == 'bin' and e[1] == 'eq' and e[2][0] == 'col' and e[3][0] == 'col':
                (n1, in1), (n2, in2) = res(e[2]), res(e[3])
                if in1 != in2:
                    if key is not None:
                        raise PlanError('a subquery correlated by more than one equality is not supported yet')
                    key, outer = (n1, n2) if in1 else (n2, n1)
                    continue
            cols = sub.columns_of(e)          # raises on columns outside the subquery's table
            flt.append(e)
        if c[0] == 'in_sub':
            if key is not None:
                raise PlanError('a correlated IN subquery is not supported yet')
            if len(q2['select']) != 1 or q2['select'][0][0][0] != 'col' or c[1][0] != 'col':
                raise PlanError('IN (subquery): one column on each side is supported')
            key, inner = res(q2['select'][0][0])
            if not inner:
                raise PlanError('IN (subquery): the selected column must belong to the subquery table')
            outer = self.resolve(c[1])
        elif key is None:
            raise PlanError('EXISTS (subquery) needs an equality with a column of the outer query')
        nullable = {n: nu for n, _, nu in nds_schema.SOURCE_TABLES[t2]}[key]
        if c[0] == 'in_sub' and negated and nullable:
            raise PlanError('NOT IN over a nullable subquery column (%s) is not supported: one NULL empties the result' % key)
        need = {key}
        for e in flt:
            need |= sub.columns_of(e)
        order = [n for n, _, _ in nds_schema.SOURCE_TABLES[t2]]
        build = {'op': 'scan', 'table': t2, 'columns': sorted(need, key=order.index)}
        if flt:
            irs = [sub.expr(e)[0] for e in flt]
            build = {'op': 'filter', 'cond': irs[0] if len(irs) == 1 else _fn('and', *irs), 'child': build}
        return {'outer': outer, 'key': key, 'build': build, 'type': 'anti' if negated else 'semi',
                'drop_null_outer': c[0] == 'in_sub' and negated}

    def pred_selectivity(self, c):
        k = c[0]
        if k == 'bin':
            if c[1] == 'eq':
                return 0.02
            if c[1] == 'ne':
                return 0.9
            if c[1] == 'or':
                return min(1.0, self.pred_selectivity(c[2]) + self.pred_selectivity(c[3]) + 0.3)
            if c[1] == 'and':
                return self.pred_selectivity(c[2]) * self.pred_selectivity(c[3])
            return 0.3
        if k == 'between':
            return 0.25
        if k == 'in':
            return min(1.0, 0.02 * len(c[2]))
        return 0.5


class _WinMap:
    """Window aggregates of a select list, grouped by their PARTITION BY list: one engine `window` node per list."""
    def __init__(self):
        self.parts = []          # [(partition irs, [agg specs])]
        self.ranks = []          # [(partition irs, order keys, [{'name', 'fn'}])]
        self.names = {}
        self.count = 0

    def add(self, pirs, fn, ir, rtype):
        key = (repr(pirs), fn, repr(ir))
        if key in self.names:
            return self.names[key]
        name = '_win%d' % self.count
        self.count += 1
        self.names[key] = name
        spec = {'name': name, 'fn': fn}
        if ir is not None:
            spec['expr'] = ir
        for p, specs in self.parts:
            if repr(p) == repr(pirs):
                specs.append(spec)
                break
        else:
            self.parts.append((pirs, [spec]))
        return name

    def add_rank(self, pirs, okeys, fn):
        key = ('rank', repr(pirs), repr(okeys), fn)
        if key in self.names:
            return self.names[key]
        name = '_win%d' % self.count
        self.count += 1
        self.names[key] = name
        for p, o, funcs in self.ranks:
            if repr(p) == repr(pirs) and repr(o) == repr(okeys):
                funcs.append({'name': name, 'fn': fn})
                break
        else:
            self.ranks.append((pirs, okeys, [{'name': name, 'fn': fn}]))
        return name

    def wrap(self, node):
        for i, (pirs, specs) in enumerate(self.parts):
            node = {'op': 'window', 'keys': [{'name': '_wk%d_%d' % (i, j), 'expr': p} for j, p in enumerate(pirs)], 'aggs': specs, 'child': node}
        for pirs, okeys, funcs in self.ranks:
            node = {'op': 'rank', 'keys': [{'expr': p, 'desc': False} for p in pirs] + okeys, 'npart': len(pirs), 'funcs': funcs, 'child': node}
        return node


class _AggMap:
    def __init__(self):
        self.items = {}
        self.order = []
        self.types = {}

    def add(self, fn, ir, rtype):
        key = (fn, repr(ir))
        if key not in self.items:
            name = '_agg%d' % len(self.order)
            self.items[key] = name
            spec = {'name': name, 'fn': fn}
            if ir is not None:
                spec['expr'] = ir
            self.order.append(spec)
            self.types[name] = rtype
        return self.items[key]

    def specs(self):
        return [dict(s) for s in self.order]


def _replace_partial_scan(node, replacement):
    """Copy of a merge plan in which the scan of `__partial__` is replaced (the rest of the tree is shared)."""
    if not isinstance(node, dict):
        return node
    if node.get('op') == 'scan' and node.get('table') == '__partial__':
        return replacement
    out = dict(node)
    for k in ('child', 'probe', 'build'):
        if k in out:
            out[k] = _replace_partial_scan(out[k], replacement)
    return out


class _DerivedPlanner(Planner):
    """The outer query of `select ... from (subquery) alias [where ...] [order by ...] [limit n]`: columns resolve against the
    subquery's output; no joins, no aggregation of its own (NDS query53 / 63 / 89: a filter over window results)."""
    def __init__(self, ast, alias, outputs):
        self.q = ast
        self.col_types = {o['name']: o['type'] for o in outputs}
        self.col_table = {}
        self.alias = {alias: '__derived__'}
        self.tables = ['__derived__']
        self.outputs = outputs

    def resolve(self, node):
        _, qual, name = node
        if qual is not None and qual not in self.alias:
            raise PlanError('unknown table alias %r' % qual)
        if name not in self.col_types:
            raise PlanError('unknown column %r' % name)
        return name

    def plan_over(self, inner_root):
        q = self.q
        if q['group'] or q.get('having') is not None or q.get('distinct') or q.get('left_joins'):
            raise PlanError('GROUP BY / HAVING / DISTINCT / joins over a derived table are not supported yet')
        node = inner_root
        if q['where'] is not None:
            cond, t = self.expr(q['where'])
            if t != 'bool':
                raise PlanError('WHERE condition is not boolean')
            node = {'op': 'filter', 'cond': cond, 'child': node}
        out_cols = []
        self._win = win_map = _WinMap()
        try:
            for e, alias in q['select']:
                if e == ('star',):
                    out_cols += [(o['name'], _col(o['name']), o['type']) for o in self.outputs]
                    continue
                ir, t = self.expr(e)
                out_cols.append((alias if alias else self.text(e), ir, t))
        finally:
            self._win = None
        node = win_map.wrap(node)
        names = [n for n, _, _ in out_cols]
        if len(set(names)) != len(names):
            raise PlanError('duplicate output column names over a derived table')
        node = {'op': 'project', 'exprs': [{'name': n, 'expr': ir} for n, ir, _ in out_cols], 'child': node}
        if q['order'] or q['limit'] is not None:
            keys, used = [], []
            for e, desc, nf in q['order']:
                if e[0] == 'col' and e[1] is None and e[2] in names:
                    name = e[2]
                else:
                    ir, _ = self.expr(e)
                    match = [n for n, oir, _ in out_cols if oir == ir]
                    if not match:
                        raise PlanError('ORDER BY expression %s is not in the select list (not supported yet)' % self.text(e))
                    name = match[0]
                k = {'expr': _col(name), 'desc': bool(desc)}
                if nf is not None:
                    k['nulls_first'] = bool(nf)
                keys.append(k)
                used.append(name)
            keys += [{'expr': _col(n), 'desc': False} for n in names if n not in used]      # deterministic tie-break
            node = {'op': 'sort', 'keys': keys, 'limit': -1 if q['limit'] is None else q['limit'], 'child': node}
        return {'root': node, 'output': [{'name': n, 'type': t} for n, _, t in out_cols]}


def plan_sql(sql_text, distributed=False, sharded=None):
    """SQL text -> plan dict ({'root': ..., 'output': [...]} [+ 'partial', 'fused' | 'replicated'])."""
    body = '\n'.join(l for l in sql_text.split('\n') if not l.strip().startswith('--'))
    try:
        ast = parse(body)
    except SqlError as e:
        raise PlanError(str(e))
    if len(ast['from']) == 1 and isinstance(ast['from'][0][0], tuple):
        if distributed:
            raise PlanError('derived tables are not supported across GPUs yet')
        return _plan_query(ast)
    return Planner(ast).plan(distributed=distributed, sharded=sharded)


def _plan_query(ast):
    """select ... from (subquery) alias ... (also nested, also written as WITH alias AS (subquery)): the subquery is planned as a
    query of its own, the outer query over its output."""
    if len(ast['from']) == 1 and isinstance(ast['from'][0][0], tuple):
        (_, sub), alias = ast['from'][0]
        inner = _plan_query(sub)
        return _DerivedPlanner(ast, alias, inner['output']).plan_over(inner['root'])
    return Planner(ast).plan()
