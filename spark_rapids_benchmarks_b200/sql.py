"""A small SQL front-end for the NDS star-join/aggregate query shape.

In the reference, `spark_session.sql(query)` (/root/reference/nds/nds_power.py:130) parses and
plans the statement inside Spark (Catalyst).  No JVM exists on the build or GPU boxes, so the host
side carries its own parser for the subset of Spark SQL the supported templates use:

    SELECT [DISTINCT] <exprs [AS alias]> FROM <t [alias], ... | t [INNER] JOIN u ON <expr> | t LEFT [OUTER] JOIN u ON <expr> ...>
    [WHERE <expr>] [GROUP BY ...] [HAVING <expr>] [ORDER BY <expr [ASC|DESC] [NULLS FIRST|LAST]>, ...] [LIMIT n]

with `[NOT] EXISTS (SELECT ... FROM t WHERE ...)` and `x [NOT] IN (SELECT c FROM t [WHERE ...])` as WHERE conjuncts
(one-table subqueries; planner.py turns them into semi / anti joins).

Anything outside the subset raises SqlError - the statement is never silently approximated.
The AST is plain tuples:  ('col', qualifier|None, name) | ('num', text) | ('str', value) | ('null',)
 | ('call', fname, [args], distinct) | ('wcall', fname, [args], [partition exprs], [(expr, desc, nulls_first)]) | ('bin', op, l, r) | ('not', x) | ('neg', x)
 | ('isnull', x, negated) | ('between', x, lo, hi, negated) | ('in', x, [items], negated)
 | ('interval', n, 'day') | ('like', x, pattern, negated) | ('case', [(cond, value)...], else|None) | ('cast', x, typename) | ('star',)
 | ('exists', query, negated) | ('in_sub', x, query, negated)
A query is {'select', 'from', 'where', 'group', 'having', 'order', 'limit', 'distinct', 'left_joins': [(alias, on_expr)],
'rollup': GROUP BY ROLLUP(...)}.
"""
import re


class SqlError(ValueError):
    pass


# one alternative per token kind; comments match without a group, the last group catches anything unexpected
_TOKEN = re.compile(r"""--[^\n]*|(\d+\.\d*|\.\d+|\d+)|('(?:[^']|'')*')|(`[^`]*`)|([A-Za-z_][A-Za-z_0-9]*)|(<>|!=|<=|>=|\|\||[=<>+\-*/(),.;])|(\S)""")

KEYWORDS = {'select', 'from', 'where', 'group', 'by', 'order', 'limit', 'as', 'and', 'or', 'not', 'is', 'null',
            'between', 'in', 'asc', 'desc', 'nulls', 'first', 'last', 'case', 'when', 'then', 'else', 'end',
            'cast', 'distinct', 'having', 'union', 'intersect', 'except', 'join', 'on', 'inner', 'left',
            'right', 'full', 'outer', 'with', 'over', 'partition', 'rollup', 'exists', 'interval', 'like'}


def tokenize(text):
    out = []
    add = out.append
    for num, st, bq, ident, op, bad in _TOKEN.findall(text):
        if ident:
            lv = ident.lower()
            add(('kw', lv) if lv in KEYWORDS else ('id', lv))
        elif op:
            add(('op', op))
        elif num:
            add(('num', num))
        elif st:
            add(('str', st[1:-1].replace("''", "'")))
        elif bq:
            add(('id', bq[1:-1]))
        elif bad:
            raise SqlError('cannot tokenize SQL at: %r' % text[text.index(bad):][:30])
    out.extend([('eof', '')] * 4)      # look-ahead never runs off the end
    return out


class Parser:
    def __init__(self, text):
        self.t = tokenize(text)
        self.i = 0
        self.ctes = {}           # WITH name AS (...): name -> [ast, references]

    def peek(self, k=0):
        return self.t[self.i + k]

    def next(self):
        tok = self.t[self.i]
        if tok[0] != 'eof':
            self.i += 1
        return tok

    def accept(self, kind, val=None):
        tok = self.t[self.i]
        if tok[0] == kind and (val is None or tok[1] == val):
            if kind != 'eof':
                self.i += 1
            return tok
        return None

    def expect(self, kind, val=None):
        tok = self.accept(kind, val)
        if not tok:
            raise SqlError('expected %s %s, found %r' % (kind, val or '', self.peek()[1]))
        return tok

    def kw(self, word):
        tok = self.t[self.i]
        if tok[0] == 'kw' and tok[1] == word:
            self.i += 1
            return tok
        return None

    # ---- statement -----------------------------------------------------------------------------
    def parse_query(self, sub=False):
        if self.kw('with'):
            # WITH name AS (subquery) [, ...]: a name that is referenced once is the same as a derived table with that alias
            if sub:
                raise SqlError('WITH inside a subquery is not supported')
            while True:
                name = self.expect('id')[1]
                self.expect('kw', 'as')
                self.expect('op', '(')
                if self.peek() != ('kw', 'select'):
                    raise SqlError('expected a subquery after WITH %s AS (' % name)
                self.ctes[name] = [self.parse_query(sub=True), 0]
                self.expect('op', ')')
                if not self.accept('op', ','):
                    break
        self.expect('kw', 'select')
        distinct = bool(self.kw('distinct'))
        select = [self.select_item()]
        while self.accept('op', ','):
            select.append(self.select_item())
        self.expect('kw', 'from')
        tables = [self.table_ref()]
        on_conds = []
        left_joins = []
        while True:
            if self.accept('op', ','):
                tables.append(self.table_ref())
                continue
            tok = self.peek()
            if tok == ('kw', 'left'):
                # LEFT [OUTER] JOIN t ON cond: t joins the tables before it, rows without a match keep NULLs in t's columns
                self.next()
                self.kw('outer')
                self.expect('kw', 'join')
                ref = self.table_ref()
                tables.append(ref)
                self.expect('kw', 'on')
                left_joins.append((ref[1], self.expr()))
                continue
            if tok[0] == 'kw' and tok[1] in ('right', 'full', 'outer'):
                raise SqlError('%s OUTER JOIN is not supported by the B200 planner yet (inner and left joins only)' % tok[1].upper())
            if tok == ('kw', 'inner') or tok == ('kw', 'join'):
                # [INNER] JOIN t ON cond  ==  , t  ... WHERE cond: an inner join's ON clause is just another conjunct
                self.kw('inner')
                self.expect('kw', 'join')
                tables.append(self.table_ref())
                self.expect('kw', 'on')
                on_conds.append(self.expr())
                continue
            break
        where = self.expr() if self.kw('where') else None
        for c in reversed(on_conds):
            where = c if where is None else ('bin', 'and', c, where)
        group = []
        rollup = False
        if self.kw('group'):
            self.expect('kw', 'by')
            if self.kw('rollup'):
                # GROUP BY ROLLUP(a, b, ...): the grouping sets (a, b, ...), ..., (a), ()
                rollup = True
                self.expect('op', '(')
                group.append(self.expr())
                while self.accept('op', ','):
                    group.append(self.expr())
                self.expect('op', ')')
            else:
                group.append(self.expr())
                while self.accept('op', ','):
                    group.append(self.expr())
        having = self.expr() if self.kw('having') else None
        order = []
        if self.kw('order'):
            self.expect('kw', 'by')
            order.append(self.order_item())
            while self.accept('op', ','):
                order.append(self.order_item())
        limit = None
        if self.kw('limit'):
            limit = int(self.expect('num')[1])
        if not sub:
            self.accept('op', ';')
            if self.peek()[0] != 'eof':
                raise SqlError('unsupported trailing SQL: %r' % self.peek()[1])
        return {'select': select, 'from': tables, 'where': where, 'group': group, 'having': having, 'order': order, 'limit': limit,
                'distinct': distinct, 'left_joins': left_joins, 'rollup': rollup}

    def select_item(self):
        if self.accept('op', '*'):
            return (('star',), None)       # the planner expands it over a derived table; over base tables it raises
        e = self.expr()
        alias = None
        if self.kw('as'):
            alias = self.expect('id')[1]
        elif self.peek()[0] == 'id':
            alias = self.next()[1]
        return (e, alias)

    def table_ref(self):
        if self.accept('op', '('):
            # derived table: ( select ... ) [as] alias
            if self.peek() != ('kw', 'select'):
                raise SqlError('expected a subquery after ( in FROM')
            sub = self.parse_query(sub=True)
            self.expect('op', ')')
            self.kw('as')
            if self.peek()[0] != 'id':
                raise SqlError('a derived table needs an alias')
            return (('subquery', sub), self.next()[1])
        name = self.expect('id')[1]
        alias = None
        if self.kw('as'):
            alias = self.expect('id')[1]
        elif self.peek()[0] == 'id':
            alias = self.next()[1]
        if name in self.ctes:
            cte = self.ctes[name]
            cte[1] += 1
            if cte[1] > 1:
                raise SqlError('common table expression %r is referenced more than once (not supported yet)' % name)
            return (('subquery', cte[0]), alias or name)
        return (name, alias or name)

    def order_item(self):
        e = self.expr()
        desc = False
        if self.kw('desc'):
            desc = True
        else:
            self.kw('asc')
        nulls_first = None
        if self.kw('nulls'):
            if self.kw('first'):
                nulls_first = True
            else:
                self.expect('kw', 'last')
                nulls_first = False
        return (e, desc, nulls_first)

    # ---- expressions -------------------------------------------------------------------------------
    def expr(self):
        e = self.and_expr()
        while self.kw('or'):
            e = ('bin', 'or', e, self.and_expr())
        return e

    def and_expr(self):
        e = self.not_expr()
        while self.kw('and'):
            e = ('bin', 'and', e, self.not_expr())
        return e

    def not_expr(self):
        if self.kw('not'):
            inner = self.not_expr()
            if inner[0] == 'exists':
                return ('exists', inner[1], not inner[2])
            return ('not', inner)
        if self.peek() == ('kw', 'exists'):
            self.next()
            self.expect('op', '(')
            q = self.parse_query(sub=True)
            self.expect('op', ')')
            return ('exists', q, False)
        return self.predicate()

    def predicate(self):
        e = self.additive()
        tok = self.peek()
        if tok[0] == 'op' and tok[1] in ('=', '<>', '!=', '<', '<=', '>', '>='):
            self.next()
            op = {'=': 'eq', '<>': 'ne', '!=': 'ne', '<': 'lt', '<=': 'le', '>': 'gt', '>=': 'ge'}[tok[1]]
            return ('bin', op, e, self.additive())
        if self.kw('is'):
            neg = bool(self.kw('not'))
            self.expect('kw', 'null')
            return ('isnull', e, neg)
        neg = False
        if self.peek() == ('kw', 'not') and self.peek(1)[0] == 'kw' and self.peek(1)[1] in ('between', 'in', 'like'):
            self.next()
            neg = True
        if self.kw('between'):
            lo = self.additive()
            self.expect('kw', 'and')
            hi = self.additive()
            return ('between', e, lo, hi, neg)
        if self.kw('in'):
            self.expect('op', '(')
            if self.peek() == ('kw', 'select'):
                q = self.parse_query(sub=True)
                self.expect('op', ')')
                return ('in_sub', e, q, neg)
            items = [self.expr()]
            while self.accept('op', ','):
                items.append(self.expr())
            self.expect('op', ')')
            return ('in', e, items, neg)
        if self.kw('like'):
            pat = self.next()
            if pat[0] != 'str':
                raise SqlError('LIKE needs a string literal pattern')
            return ('like', e, pat[1], neg)
        return e

    def additive(self):
        e = self.mult()
        while True:
            if self.accept('op', '+'):
                e = ('bin', 'add', e, self.mult())
            elif self.accept('op', '-'):
                e = ('bin', 'sub', e, self.mult())
            else:
                return e

    def mult(self):
        e = self.unary()
        while True:
            if self.accept('op', '*'):
                e = ('bin', 'mul', e, self.unary())
            elif self.accept('op', '/'):
                e = ('bin', 'div', e, self.unary())
            else:
                return e

    def unary(self):
        if self.accept('op', '-'):
            return ('neg', self.unary())
        if self.accept('op', '+'):
            return self.unary()
        return self.primary()

    def primary(self):
        tok = self.next()
        if tok[0] == 'num':
            return ('num', tok[1])
        if tok[0] == 'str':
            return ('str', tok[1])
        if tok == ('kw', 'null'):
            return ('null',)
        if tok == ('kw', 'interval'):
            # INTERVAL n DAY[S] (Spark's form in the NDS templates: cast('1999-02-22' as date) + interval 30 days)
            n = self.next()
            if n[0] == 'str':
                n = ('num', n[1].strip())
            if n[0] != 'num' or not n[1].isdigit():
                raise SqlError('INTERVAL needs an integer count')
            unit = self.next()
            if unit[0] != 'id' or unit[1].lower() not in ('day', 'days'):
                raise SqlError('only INTERVAL n DAYS is supported')
            return ('interval', int(n[1]), 'day')
        if tok == ('op', '('):
            if self.peek() == ('kw', 'select'):
                raise SqlError('scalar subqueries are not supported by the B200 planner yet')
            e = self.expr()
            self.expect('op', ')')
            return e
        if tok == ('kw', 'case'):
            whens = []
            if self.peek() != ('kw', 'when'):
                raise SqlError('CASE <expr> WHEN is not supported; use CASE WHEN <cond>')
            while self.kw('when'):
                c = self.expr()
                self.expect('kw', 'then')
                whens.append((c, self.expr()))
            other = self.expr() if self.kw('else') else None
            self.expect('kw', 'end')
            return ('case', whens, other)
        if tok == ('kw', 'cast'):
            self.expect('op', '(')
            e = self.expr()
            self.expect('kw', 'as')
            tname = self.expect('id')[1]
            if self.accept('op', '('):
                args = [self.expect('num')[1]]
                while self.accept('op', ','):
                    args.append(self.expect('num')[1])
                self.expect('op', ')')
                tname += '(%s)' % ','.join(args)
            self.expect('op', ')')
            return ('cast', e, tname)
        if tok[0] == 'id':
            if self.accept('op', '('):
                distinct = bool(self.kw('distinct'))
                args = []
                if self.accept('op', '*'):
                    args.append(('star',))
                elif self.peek() != ('op', ')'):
                    args.append(self.expr())
                    while self.accept('op', ','):
                        args.append(self.expr())
                self.expect('op', ')')
                if self.kw('over'):
                    # fn(x) OVER (PARTITION BY a, b): aggregate window functions without ORDER BY / frame
                    if distinct:
                        raise SqlError('DISTINCT in a window function is not supported')
                    self.expect('op', '(')
                    part = []
                    if self.kw('partition'):
                        self.expect('kw', 'by')
                        part.append(self.expr())
                        while self.accept('op', ','):
                            part.append(self.expr())
                    order = []
                    if self.kw('order'):
                        self.expect('kw', 'by')
                        order.append(self.order_item())
                        while self.accept('op', ','):
                            order.append(self.order_item())
                    if self.peek() != ('op', ')'):
                        raise SqlError('window frames (ROWS / RANGE BETWEEN ...) are not supported by the B200 planner yet')
                    self.expect('op', ')')
                    return ('wcall', tok[1], args, part, order)
                return ('call', tok[1], args, distinct)
            if self.accept('op', '.'):
                return ('col', tok[1], self.expect('id')[1])
            return ('col', None, tok[1])
        raise SqlError('unexpected token %r' % (tok[1],))


def parse(text):
    return Parser(text).parse_query()
