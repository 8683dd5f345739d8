"""GPU parity for the operator kernels on in-memory tables: fused filter/project with SQL null
semantics, hash join, hash group-by (sum/count/min/max, decimal avg), radix sort + limit.
Checked against pyarrow.compute / plain Python - bit exact for integers, strings, decimals."""
import decimal

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc
import pytest

pytestmark = pytest.mark.gpu


def col(n):
    return {'col': n}


def lit(v, t=None):
    d = {'lit': v}
    if t:
        d['type'] = t
    return d


def fn(name, *args, **kw):
    d = {'fn': name, 'args': list(args)}
    d.update(kw)
    return d


def _mk(n, seed=7):
    rng = np.random.default_rng(seed)
    a = pa.array(rng.integers(-50, 50, n).astype(np.int32), mask=rng.random(n) < 0.1)
    b = pa.array(rng.integers(-10**12, 10**12, n), mask=rng.random(n) < 0.1)
    words = ['alpha', 'beta', 'gamma', 'delta', '', 'epsilon-longer-string', 'b', 'beta2']
    s = pa.array([None if rng.random() < 0.1 else words[i] for i in rng.integers(0, len(words), n)], type=pa.string())
    d = pa.array([None if rng.random() < 0.1 else decimal.Decimal(int(v)).scaleb(-2) for v in rng.integers(-10**6, 10**6, n)],
                 type=pa.decimal128(7, 2))
    g = pa.array(rng.integers(0, 13, n), mask=rng.random(n) < 0.05)
    return pa.table({'a': a, 'b': b, 's': s, 'd': d, 'g': g})


@pytest.fixture(scope='module')
def tbl(gpu_engine):
    t = _mk(10007)
    gpu_engine.register_arrow('t', t)
    return t


def test_roundtrip_arrow(gpu_engine, tbl):
    got = gpu_engine.execute({'op': 'scan', 'table': 't', 'columns': tbl.schema.names})
    assert pa.Table.from_batches([got]).equals(tbl.combine_chunks())


def test_filter_three_valued_logic(gpu_engine, tbl):
    # (a > 3 and s <> 'beta') or b is null or d >= 12.34
    cond = fn('or', fn('and', fn('gt', col('a'), lit(3)), fn('ne', col('s'), lit('beta'))),
              fn('isnull', col('b')), fn('ge', col('d'), lit(1234, 'decimal(7,2)')))
    got = gpu_engine.execute({'op': 'sort', 'keys': [{'expr': col('rid')}],
                              'child': {'op': 'filter', 'cond': cond,
                                        'child': {'op': 'scan', 'table': 't2', 'columns': ['rid', 'a', 'b', 's', 'd']}}}) \
        if False else None
    t = tbl.append_column('rid', pa.array(np.arange(tbl.num_rows)))
    gpu_engine.register_arrow('t2', t)
    got = gpu_engine.execute({'op': 'sort', 'keys': [{'expr': col('rid')}],
                              'child': {'op': 'filter', 'cond': cond,
                                        'child': {'op': 'scan', 'table': 't2', 'columns': ['rid', 'a', 'b', 's', 'd']}}})
    m = pc.or_kleene(pc.or_kleene(pc.and_kleene(pc.greater(t['a'], 3), pc.not_equal(t['s'], 'beta')), pc.is_null(t['b'])),
                     pc.greater_equal(t['d'], pa.scalar(decimal.Decimal('12.34'), pa.decimal128(7, 2))))
    exp = t.filter(m.combine_chunks().fill_null(False)).select(['rid', 'a', 'b', 's', 'd']).combine_chunks()
    assert pa.Table.from_batches([got]).equals(exp)


def test_project_arithmetic_substr(gpu_engine, tbl):
    t = tbl.append_column('rid', pa.array(np.arange(tbl.num_rows)))
    gpu_engine.register_arrow('t2', t)
    plan = {'op': 'project', 'child': {'op': 'scan', 'table': 't2', 'columns': ['rid', 'a', 'b', 's']},
            'exprs': [{'name': 'rid', 'expr': col('rid')},
                      {'name': 'x', 'expr': fn('add', fn('mul', col('a'), lit(3)), col('b'))},
                      {'name': 'p', 'expr': fn('substr', col('s'), start=2, len=3)},
                      {'name': 'c', 'expr': fn('if', fn('lt', col('a'), lit(0)), lit(-1), col('a'))}]}
    got = pa.Table.from_batches([gpu_engine.execute(plan)])
    x = pc.add(pc.multiply(t['a'].cast(pa.int64()), 3), t['b'])
    p = pc.utf8_slice_codeunits(t['s'], 1, 4)
    a64 = t['a'].cast(pa.int64())
    c = pc.if_else(pc.less(a64, 0).fill_null(False), pa.scalar(-1, pa.int64()), a64)
    assert got['x'].combine_chunks().equals(x.combine_chunks())
    assert got['p'].combine_chunks().equals(p.combine_chunks())
    assert got['c'].combine_chunks().equals(c.combine_chunks())


def test_join_inner_unique_build(gpu_engine, tbl):
    n = 5000
    rng = np.random.default_rng(3)
    dim = pa.table({'k': pa.array(np.arange(100, 100 + 300, dtype=np.int32)),
                    'name': pa.array(['n%03d' % (i % 37) for i in range(300)]),
                    'w': pa.array(rng.integers(0, 1000, 300))})
    fact = pa.table({'fk': pa.array(rng.integers(0, 600, n).astype(np.int32), mask=rng.random(n) < 0.05),
                     'v': pa.array(rng.integers(0, 10**6, n)), 'rid': pa.array(np.arange(n))})
    gpu_engine.register_arrow('dim', dim)
    gpu_engine.register_arrow('fact', fact)
    plan = {'op': 'sort', 'keys': [{'expr': col('rid')}],
            'child': {'op': 'join', 'probe': {'op': 'scan', 'table': 'fact', 'columns': ['fk', 'v', 'rid']},
                      'build': {'op': 'filter', 'cond': fn('lt', col('w'), lit(700)),
                                'child': {'op': 'scan', 'table': 'dim', 'columns': ['k', 'name', 'w']}},
                      'probe_key': 'fk', 'build_key': 'k'}}
    got = pa.Table.from_batches([gpu_engine.execute(plan)])
    d2 = dim.filter(pc.less(dim['w'], 700))
    exp = fact.join(d2, keys='fk', right_keys='k', join_type='inner').sort_by('rid')
    assert got.num_rows == exp.num_rows
    for c in ['fk', 'v', 'rid', 'name', 'w']:
        assert got[c].combine_chunks().equals(exp[c].combine_chunks()), c


def test_join_duplicate_build_keys_fail(gpu_engine):
    from spark_rapids_benchmarks_b200.engine import EngineError
    gpu_engine.register_arrow('dup', pa.table({'k': pa.array([1, 2, 2], type=pa.int32())}))
    gpu_engine.register_arrow('pr', pa.table({'fk': pa.array([1, 2, 3], type=pa.int32())}))
    with pytest.raises(EngineError):
        gpu_engine.execute({'op': 'join', 'probe': {'op': 'scan', 'table': 'pr', 'columns': ['fk']},
                            'build': {'op': 'scan', 'table': 'dup', 'columns': ['k']}, 'probe_key': 'fk', 'build_key': 'k'})


def test_group_by_sum_count_min_max_avg(gpu_engine, tbl):
    plan = {'op': 'sort',
            'keys': [{'expr': col('g')}, {'expr': col('s')}],
            'child': {'op': 'project',
                      'exprs': [{'name': 'g', 'expr': col('g')}, {'name': 's', 'expr': col('s')},
                                {'name': 'sd', 'expr': col('sd')}, {'name': 'cb', 'expr': col('cb')},
                                {'name': 'n', 'expr': col('n')}, {'name': 'mn', 'expr': col('mn')}, {'name': 'mx', 'expr': col('mx')},
                                {'name': 'avg_d', 'expr': fn('decavg', col('sd'), col('cd'), up=4, type='decimal(11,6)')},
                                {'name': 'avg_a', 'expr': fn('divi2f', col('sa'), col('ca'))}],
                      'child': {'op': 'aggregate',
                                'keys': [{'name': 'g', 'expr': col('g')}, {'name': 's', 'expr': col('s')}],
                                'aggs': [{'name': 'sd', 'fn': 'sum', 'expr': col('d')}, {'name': 'cd', 'fn': 'count', 'expr': col('d')},
                                         {'name': 'cb', 'fn': 'count', 'expr': col('b')}, {'name': 'n', 'fn': 'count_star'},
                                         {'name': 'mn', 'fn': 'min', 'expr': col('b')}, {'name': 'mx', 'fn': 'max', 'expr': col('a')},
                                         {'name': 'sa', 'fn': 'sum', 'expr': col('a')}, {'name': 'ca', 'fn': 'count', 'expr': col('a')}],
                                'child': {'op': 'scan', 'table': 't', 'columns': ['a', 'b', 's', 'd', 'g']}}}}
    got = pa.Table.from_batches([gpu_engine.execute(plan)]).to_pylist()
    # plain-Python oracle
    groups = {}
    for r in tbl.to_pylist():
        groups.setdefault((r['g'], r['s']), []).append(r)

    def keyf(k):   # NULLS FIRST ascending
        return tuple((0, 0) if v is None else (1, v if not isinstance(v, str) else v.encode()) for v in k)
    exp = []
    for k in sorted(groups, key=keyf):
        rows = groups[k]
        ds = [r['d'] for r in rows if r['d'] is not None]
        bs = [r['b'] for r in rows if r['b'] is not None]
        as_ = [r['a'] for r in rows if r['a'] is not None]
        sd = sum(ds) if ds else None
        avg_d = None
        if ds:
            avg_d = (sd / decimal.Decimal(len(ds))).quantize(decimal.Decimal('0.000001'), rounding=decimal.ROUND_HALF_UP)
        exp.append({'g': k[0], 's': k[1], 'sd': sd, 'cb': len(bs), 'n': len(rows), 'mn': min(bs) if bs else None,
                    'mx': max(as_) if as_ else None, 'avg_d': avg_d, 'avg_a': (sum(as_) / len(as_)) if as_ else None})
    assert len(got) == len(exp)
    for g, e in zip(got, exp):
        assert g == e, (g, e)


def test_sort_multi_key_desc_nulls_limit(gpu_engine, tbl):
    t = tbl.append_column('rid', pa.array(np.arange(tbl.num_rows)))
    gpu_engine.register_arrow('t2', t)
    plan = {'op': 'sort', 'limit': 250,
            'keys': [{'expr': col('s'), 'desc': True}, {'expr': col('d')}, {'expr': col('b'), 'desc': True}, {'expr': col('rid')}],
            'child': {'op': 'scan', 'table': 't2', 'columns': ['rid', 's', 'd', 'b']}}
    got = pa.Table.from_batches([gpu_engine.execute(plan)]).to_pylist()
    rows = t.select(['rid', 's', 'd', 'b']).to_pylist()

    def key(r):
        s, d, b = r['s'], r['d'], r['b']
        ks = (1, b'') if s is None else (0, bytes(255 - c for c in s.encode()) + b'\xff')   # desc, nulls last
        kd = (0, 0) if d is None else (1, d)                                                # asc, nulls first
        kb = (1, 0) if b is None else (0, -b)                                               # desc, nulls last
        return (ks, kd, kb, r['rid'])
    # descending byte order with prefix rule: a longer string sorts BEFORE its prefix when descending
    import functools

    def cmp(x, y):
        for kx, ky, desc in ((x['s'], y['s'], True), (x['d'], y['d'], False), (x['b'], y['b'], True), (x['rid'], y['rid'], False)):
            if kx is None and ky is None:
                continue
            if kx is None:
                return 1 if desc else -1
            if ky is None:
                return -1 if desc else 1
            if isinstance(kx, str):
                kx, ky = kx.encode(), ky.encode()
            if kx == ky:
                continue
            r = -1 if kx < ky else 1
            return -r if desc else r
        return 0
    exp = sorted(rows, key=functools.cmp_to_key(cmp))[:250]
    assert got == exp


def test_empty_inputs(gpu_engine):
    gpu_engine.register_arrow('e', pa.table({'k': pa.array([], type=pa.int32()), 'v': pa.array([], type=pa.int64())}))
    got = gpu_engine.execute({'op': 'sort', 'keys': [{'expr': col('k')}], 'limit': 10,
                              'child': {'op': 'aggregate', 'keys': [{'name': 'k', 'expr': col('k')}],
                                        'aggs': [{'name': 'sv', 'fn': 'sum', 'expr': col('v')}],
                                        'child': {'op': 'filter', 'cond': fn('gt', col('v'), lit(0)),
                                                  'child': {'op': 'scan', 'table': 'e', 'columns': ['k', 'v']}}}})
    assert got.num_rows == 0 and got.schema.names == ['k', 'sv']


def _sort_case(gpu_engine, n, limit, pattern, keep_frac, seed, options=None):
    """Every sort route: rank sort (<= 8192 rows, count on the device), top-k prefilter + rank sort over the
    candidates, candidate overflow -> LSD radix, no LIMIT -> LSD radix; expected order from numpy.lexsort."""
    rng = np.random.default_rng(seed)
    if pattern == 'unique':
        k1 = rng.permutation(n).astype(np.int64) - n // 2
    elif pattern == 'ties':
        k1 = rng.integers(0, 5, n).astype(np.int64)
    else:                                   # 'flat': one value -> every row ties on the first key
        k1 = np.full(n, 42, dtype=np.int64)
    k2 = rng.integers(-1000, 1000, n).astype(np.int32)
    words = np.array(['AAAAAAAA%08d' % v for v in rng.integers(0, 3000, n)])
    v = rng.integers(0, 100, n).astype(np.int32)
    rid = np.arange(n, dtype=np.int64)
    t = pa.table({'k1': k1, 'k2': k2, 'w': pa.array(words.tolist(), type=pa.string()), 'v': v, 'rid': rid})
    gpu_engine.register_arrow('srt', t)
    thr = int(100 * (1 - keep_frac))
    plan = {'op': 'sort',
            'keys': [{'expr': col('k1')}, {'expr': col('w'), 'desc': True}, {'expr': col('k2'), 'desc': True}, {'expr': col('rid')}],
            'child': {'op': 'filter', 'cond': fn('ge', col('v'), lit(thr)),
                      'child': {'op': 'scan', 'table': 'srt', 'columns': ['k1', 'k2', 'w', 'v', 'rid']}}}
    if limit is not None:
        plan['limit'] = limit
    got = gpu_engine.execute(plan, options=options)
    keep = v >= thr
    kk1, kk2, kw, krid = k1[keep], k2[keep], words[keep], rid[keep]
    # fixed-width ASCII strings: descending order = ascending order of the negated code points
    wcodes = np.frombuffer(''.join(kw.tolist()).encode(), dtype=np.uint8).reshape(len(kw), 16).astype(np.int16) if len(kw) else np.zeros((0, 16), np.int16)
    lex = [krid, -kk2.astype(np.int64)] + [-wcodes[:, c] for c in range(15, -1, -1)] + [kk1]
    order = np.lexsort(lex)
    if limit is not None:
        order = order[:limit]
    assert got.num_rows == len(order)
    assert got.column('rid').to_pylist() == krid[order].tolist()
    assert got.column('w').to_pylist() == kw[order].tolist()
    assert got.column('k2').to_pylist() == kk2[order].tolist()
    gpu_engine.drop_table('srt')


@pytest.mark.parametrize('n,limit,pattern,keep', [
    (1, 100, 'unique', 1.0), (100, 100, 'ties', 1.0), (100, None, 'flat', 0.5), (5000, 7, 'ties', 0.5),
    (8192, 100, 'unique', 1.0), (8192, None, 'ties', 1.0), (8193, 100, 'unique', 1.0), (8193, None, 'flat', 0.7),
    (40000, 100, 'unique', 0.5), (40000, 100, 'ties', 1.0), (40000, 1, 'flat', 1.0), (40000, 1024, 'ties', 0.3),
    (40000, 5000, 'unique', 1.0), (40000, None, 'unique', 0.6), (300000, 100, 'flat', 1.0), (300000, 100, 'unique', 0.9),
    (300000, 1000, 'ties', 0.01), (20000, 100, 'unique', 0.0),
])
def test_sort_routes(gpu_engine, n, limit, pattern, keep):
    _sort_case(gpu_engine, n, limit, pattern, keep, seed=n + (limit or 0))
    if limit is not None and limit <= 128:
        # LIMIT <= 128 takes the one-launch streaming top-k by default; the select / rank-sort route must agree
        _sort_case(gpu_engine, n, limit, pattern, keep, seed=n + (limit or 0), options={'topk_stream': False})


@pytest.mark.parametrize('n,limit', [(3, 1), (255, 128), (257, 100), (700, 100), (1024, 100), (1025, 100), (4097, 128), (16384, 100),
                                     (16385, 100), (70000, 128), (1500000, 100), (1500000, 3)])
def test_streaming_topk_sizes(gpu_engine, n, limit):
    """The one-launch top-k at sizes around its tile (256 rows), its candidate area (512), one CTA's minimum share (1024 rows:
    one CTA and no merge up to there), one merge group (16 CTAs: a single merge up to 16384 rows) and its grid, with
    heavy ties on the leading keys and with most rows filtered out on the device."""
    _sort_case(gpu_engine, n, limit, 'ties', 1.0, seed=n * 7 + limit)
    _sort_case(gpu_engine, n, limit, 'flat', 0.35, seed=n * 11 + limit)
    _sort_case(gpu_engine, n, limit, 'unique', 1.0, seed=n * 13 + limit)


def test_sort_topk_candidate_overflow_falls_back(gpu_engine):
    """More than 8192 rows tie on everything the prefilter looks at (4 leading bytes of compacted word 0): the
    candidate list overflows and the sort must take the radix route - same answer."""
    n = 60000
    rng = np.random.default_rng(5)
    hi = np.full(n, 0x0101010100000000, dtype=np.int64)    # the four leading bytes are the same for every row...
    lo = rng.integers(0, 1 << 20, n).astype(np.int64)      # (differences only in the low 20 bits)
    k = hi + lo
    k[0] = 0                                               # ...except one outlier, which makes those bytes "active"
    rid = np.arange(n, dtype=np.int64)
    gpu_engine.register_arrow('ovf', pa.table({'k': k, 'rid': rid}))
    gpu_engine.execute({'op': 'scan', 'table': 'ovf', 'columns': ['rid']})
    retries0 = gpu_engine.last_stats().get('speculation_retries', 0)
    got = gpu_engine.execute({'op': 'sort', 'limit': 50, 'keys': [{'expr': col('k'), 'desc': True}, {'expr': col('rid')}],
                              'child': {'op': 'scan', 'table': 'ovf', 'columns': ['k', 'rid']}}, options={'topk_stream': False})
    order = np.lexsort([rid, -k])[:50]
    assert got.column('rid').to_pylist() == rid[order].tolist()
    # the device-sized first attempt reported the overflow and the engine ran the plan again without speculation
    assert gpu_engine.last_stats()['speculation_retries'] == retries0 + 1
    plan = {'op': 'sort', 'limit': 50, 'keys': [{'expr': col('k'), 'desc': True}, {'expr': col('rid')}],
            'child': {'op': 'scan', 'table': 'ovf', 'columns': ['k', 'rid']}}
    again = gpu_engine.execute(plan, options={'speculate': False, 'topk_stream': False})
    assert again.column('rid').to_pylist() == rid[order].tolist()
    assert gpu_engine.last_stats()['speculation_retries'] == retries0 + 1
    # the default route (streaming top-k) has no candidate list to overflow: same rows, no retry
    dflt = gpu_engine.execute(plan)
    assert dflt.column('rid').to_pylist() == rid[order].tolist()
    assert gpu_engine.last_stats()['speculation_retries'] == retries0 + 1
    gpu_engine.drop_table('ovf')


XROUTES = {'peer': {}, 'single': {'exchange_peer': False, 'exchange_single': True}, 'exact': {'exchange_peer': False}}


@pytest.mark.parametrize('route', sorted(XROUTES))
@pytest.mark.parametrize('k', [2, 3])
def test_exchange_loopback_pack_unpack(gpu_engine, tbl, k, route):
    """The exchange node's pack -> (all-gather) -> unpack path on one GPU: options.exchange_loopback=k stands in
    for k ranks that all hold this rank's rows, so the result is the input k times, in rank order - integers,
    decimals, strings (empty and long) and nulls included.  The filter leaves the row count on the device."""
    plan = {'op': 'exchange',
            'child': {'op': 'filter', 'cond': fn('ge', col('g'), lit(3)),
                      'child': {'op': 'scan', 'table': 't', 'columns': ['a', 'b', 's', 'd', 'g']}}}
    # peer (default): rows packed straight into the peers' windows, row counts in the slot headers, no host round trip;
    # single: one NCCL collective of fixed-capacity segments; exact: the counts are read back first, segments sized exactly
    xopts = dict({'exchange_loopback': k}, **XROUTES[route])
    got = pa.Table.from_batches([gpu_engine.execute(plan, options=xopts)])
    keep = pc.fill_null(pc.greater_equal(tbl.column('g'), 3), False)
    one = tbl.filter(keep).select(['a', 'b', 's', 'd', 'g'])
    exp = pa.concat_tables([one] * k)
    assert got.num_rows == exp.num_rows
    # (the filter compacts with atomics, so row order inside a rank is not the input order: compare as multisets)
    def rows(t):
        return sorted(zip(*[t.column(n).to_pylist() for n in exp.schema.names]), key=repr)
    assert rows(got) == rows(exp)
    # rank order: the second copy repeats the first, row for row
    n1 = one.num_rows
    for name in exp.schema.names:
        c = got.column(name).to_pylist()
        assert c[:n1] == c[n1:2 * n1], name
    # and an aggregate over the gathered rows: every group's sum is k times the single-rank sum
    agg = {'op': 'sort', 'keys': [{'expr': col('g')}],
           'child': {'op': 'aggregate', 'keys': [{'name': 'g', 'expr': col('g')}],
                     'aggs': [{'name': 'sb', 'fn': 'sum', 'expr': col('b')}, {'name': 'c', 'fn': 'count_star'}],
                     'child': plan}}
    g1 = pa.Table.from_batches([gpu_engine.execute(agg)]).to_pylist()                      # world = 1: pass-through
    gk = pa.Table.from_batches([gpu_engine.execute(agg, options=xopts)]).to_pylist()
    assert [(r['g'], r['sb'] * k, r['c'] * k) for r in g1] == [(r['g'], r['sb'], r['c']) for r in gk]


def test_exchange_empty_input(gpu_engine, tbl):
    plan = {'op': 'exchange', 'child': {'op': 'filter', 'cond': fn('gt', col('g'), lit(1000)),
                                        'child': {'op': 'scan', 'table': 't', 'columns': ['a', 's', 'g']}}}
    for route in XROUTES.values():
        got = gpu_engine.execute(plan, options=dict({'exchange_loopback': 2}, **route))
        assert got.num_rows == 0 and got.schema.names == ['a', 's', 'g']


def test_exchange_beyond_the_speculative_capacity_is_retried(gpu_engine):
    """More rows than a speculative segment holds (peer window slot: 256 k rows, single collective: 128 k): the pack kernel
    reports it and the query runs again with the counts read back - same rows."""
    n = 300000
    rid = np.arange(n, dtype=np.int64)
    gpu_engine.register_arrow('big', pa.table({'rid': rid, 'v': (rid * 7 % 1000).astype(np.int32)}))
    gpu_engine.execute({'op': 'scan', 'table': 'big', 'columns': ['rid']})
    retries0 = gpu_engine.last_stats().get('speculation_retries', 0)
    plan = {'op': 'aggregate', 'keys': [], 'aggs': [{'name': 'n', 'fn': 'count_star'}, {'name': 's', 'fn': 'sum', 'expr': col('rid')}],
            'child': {'op': 'exchange', 'child': {'op': 'filter', 'cond': fn('ge', col('v'), lit(0)),
                                                  'child': {'op': 'scan', 'table': 'big', 'columns': ['rid', 'v']}}}}
    for i, route in enumerate(('peer', 'single')):
        got = pa.Table.from_batches([gpu_engine.execute(plan, options=dict({'exchange_loopback': 2}, **XROUTES[route]))]).to_pylist()
        assert got == [{'n': 2 * n, 's': 2 * int(rid.sum())}], route
        assert gpu_engine.last_stats()['speculation_retries'] == retries0 + 1 + i, route
    # the same exchange twice more: the window halves alternate and their arrival counters keep counting
    small = {'op': 'aggregate', 'keys': [], 'aggs': [{'name': 'n', 'fn': 'count_star'}],
             'child': {'op': 'exchange', 'child': {'op': 'filter', 'cond': fn('lt', col('v'), lit(10)),
                                                   'child': {'op': 'scan', 'table': 'big', 'columns': ['rid', 'v']}}}}
    exp_n = 3 * int(((rid * 7 % 1000) < 10).sum())
    for _ in range(5):
        got = pa.Table.from_batches([gpu_engine.execute(small, options={'exchange_loopback': 3})]).to_pylist()
        assert got == [{'n': exp_n}]
        assert any(o['op'] == 'exchange:peer' for o in gpu_engine.last_stats()['ops'])
    gpu_engine.drop_table('big')


def test_expression_deeper_than_the_evaluator_stack_is_refused(gpu_engine, tbl):
    """CASE with many branches nests if() two stack entries per level: the compiler must refuse what the evaluator
    cannot hold instead of writing past its stack."""
    from spark_rapids_benchmarks_b200.engine import EngineError

    def case(depth):
        e = lit(0)
        for k in range(depth):
            e = fn('if', fn('eq', col('g'), lit(k)), lit(100 + k), e)
        return e
    plan = lambda depth: {'op': 'project', 'exprs': [{'name': 'x', 'expr': case(depth)}, {'name': 'g', 'expr': col('g')}],
                          'child': {'op': 'scan', 'table': 't', 'columns': ['g']}}
    got = pa.Table.from_batches([gpu_engine.execute(plan(5))])          # a 5-branch CASE fits
    exp = [None if g is None else (100 + g if g < 5 else 0) for g in tbl['g'].to_pylist()]
    assert [x if g is not None else None for x, g in zip(got['x'].to_pylist(), got['g'].to_pylist())] == exp
    with pytest.raises(EngineError, match='too deep'):
        gpu_engine.execute(plan(10))      # 22 stack entries, 51 instructions: under the code-size limit, over the stack


def test_dates_through_case_and_min_max_export_as_date32(gpu_engine):
    """A DATE that passes through if() or min()/max() is held as int64 on the device; the Arrow export must still be a
    well-formed date32 column."""
    import datetime
    rng = np.random.default_rng(5)
    n = 1000
    days = rng.integers(9000, 12000, n).astype(np.int32)
    d = pa.array(days, type=pa.int32(), mask=rng.random(n) < 0.1).cast(pa.date32())
    k = pa.array(rng.integers(0, 7, n).astype(np.int64))
    gpu_engine.register_arrow('dt', pa.table({'k': k, 'd': d}))
    scan = {'op': 'scan', 'table': 'dt', 'columns': ['k', 'd']}
    got = pa.Table.from_batches([gpu_engine.execute({'op': 'project', 'exprs': [
        {'name': 'c', 'expr': fn('if', fn('lt', col('k'), lit(3)), col('d'), lit(None, 'date'))}, {'name': 'k', 'expr': col('k')},
        {'name': 'd', 'expr': col('d')}], 'child': scan})])
    assert got.schema.field('c').type == pa.date32()
    assert got['c'].to_pylist() == [dv if kv < 3 else None for dv, kv in zip(got['d'].to_pylist(), got['k'].to_pylist())]
    agg = pa.Table.from_batches([gpu_engine.execute({'op': 'sort', 'keys': [{'expr': col('k')}], 'child': {
        'op': 'aggregate', 'keys': [{'name': 'k', 'expr': col('k')}],
        'aggs': [{'name': 'lo', 'fn': 'min', 'expr': col('d')}, {'name': 'hi', 'fn': 'max', 'expr': col('d')}], 'child': scan}})])
    assert agg.schema.field('lo').type == pa.date32() and agg.schema.field('hi').type == pa.date32()
    by = {}
    for kv, dv in zip(k.to_pylist(), d.to_pylist()):
        if dv is not None:
            by.setdefault(kv, []).append(dv)
    assert agg['lo'].to_pylist() == [min(by[kv]) for kv in agg['k'].to_pylist()]
    assert agg['hi'].to_pylist() == [max(by[kv]) for kv in agg['k'].to_pylist()]
    assert isinstance(agg['lo'].to_pylist()[0], datetime.date)
    gpu_engine.drop_table('dt')


def test_keyless_aggregate(gpu_engine, tbl):
    """SELECT sum / count / min / max without GROUP BY: one row - also when no input row survives (COUNT 0, the rest NULL).
    Every warp of the insert kernel sees one group, so this is the warp-reduction path of the group-by."""
    def run(cond):
        child = {'op': 'scan', 'table': 't', 'columns': ['a', 'b', 'd', 'g']}
        if cond is not None:
            child = {'op': 'filter', 'cond': cond, 'child': child}
        return pa.Table.from_batches([gpu_engine.execute({'op': 'aggregate', 'keys': [], 'aggs': [
            {'name': 'n', 'fn': 'count_star'}, {'name': 'ca', 'fn': 'count', 'expr': col('a')}, {'name': 'sa', 'fn': 'sum', 'expr': col('a')},
            {'name': 'sb', 'fn': 'sum', 'expr': col('b')}, {'name': 'sd', 'fn': 'sum', 'expr': col('d')},
            {'name': 'lo', 'fn': 'min', 'expr': col('b')}, {'name': 'hi', 'fn': 'max', 'expr': col('b')}], 'child': child})])
    got = run(None).to_pylist()
    assert got == [{'n': tbl.num_rows, 'ca': pc.count(tbl['a']).as_py(), 'sa': pc.sum(tbl['a']).as_py(), 'sb': pc.sum(tbl['b']).as_py(),
                    'sd': pc.sum(tbl['d']).as_py(), 'lo': pc.min(tbl['b']).as_py(), 'hi': pc.max(tbl['b']).as_py()}]
    keep = pc.fill_null(pc.greater(tbl['g'], 6), False)
    sub = tbl.filter(keep)
    got = run(fn('gt', col('g'), lit(6))).to_pylist()
    assert got == [{'n': sub.num_rows, 'ca': pc.count(sub['a']).as_py(), 'sa': pc.sum(sub['a']).as_py(), 'sb': pc.sum(sub['b']).as_py(),
                    'sd': pc.sum(sub['d']).as_py(), 'lo': pc.min(sub['b']).as_py(), 'hi': pc.max(sub['b']).as_py()}]
    none = run(fn('gt', col('g'), lit(1000))).to_pylist()
    assert none == [{'n': 0, 'ca': 0, 'sa': None, 'sb': None, 'sd': None, 'lo': None, 'hi': None}]


def test_group_by_with_few_groups_per_warp(gpu_engine):
    """Sorted input: every warp of the insert kernel holds one group (the warp-reduced path), group boundaries fall inside
    warps (the per-lane path), and negative values / NULLs go through the 16-bit digit sums."""
    rng = np.random.default_rng(99)
    n = 50000
    k = np.sort(rng.integers(0, 37, n)).astype(np.int64)
    v = pa.array(rng.integers(-2**40, 2**40, n), mask=rng.random(n) < 0.2)
    gpu_engine.register_arrow('fg', pa.table({'k': k, 'v': v}))
    got = pa.Table.from_batches([gpu_engine.execute({'op': 'sort', 'keys': [{'expr': col('k')}], 'child': {
        'op': 'aggregate', 'keys': [{'name': 'k', 'expr': col('k')}],
        'aggs': [{'name': 's', 'fn': 'sum', 'expr': col('v')}, {'name': 'c', 'fn': 'count', 'expr': col('v')}, {'name': 'n', 'fn': 'count_star'}],
        'child': {'op': 'scan', 'table': 'fg', 'columns': ['k', 'v']}}})])
    exp = pa.table({'k': k, 'v': v}).group_by('k').aggregate([('v', 'sum'), ('v', 'count'), ([], 'count_all')]).sort_by('k')
    assert got['k'].to_pylist() == exp['k'].to_pylist()
    assert got['s'].to_pylist() == exp['v_sum'].to_pylist()
    assert got['c'].to_pylist() == exp['v_count'].to_pylist()
    assert got['n'].to_pylist() == exp['count_all'].to_pylist()
    gpu_engine.drop_table('fg')


def _join_tables(seed=17, n=6000, nd=400, dup=False):
    rng = np.random.default_rng(seed)
    keys = rng.integers(50, 50 + nd // (3 if dup else 1), nd).astype(np.int32) if dup else np.arange(50, 50 + nd, dtype=np.int32)
    dim = pa.table({'k': pa.array(keys, mask=rng.random(nd) < (0.03 if dup else 0.0)),
                    'name': pa.array(['n%03d' % (i % 41) for i in range(nd)]),
                    'w': pa.array(rng.integers(0, 1000, nd)), 'brow': pa.array(np.arange(nd, dtype=np.int64))})
    fact = pa.table({'fk': pa.array(rng.integers(0, 700, n).astype(np.int32), mask=rng.random(n) < 0.05),
                     'v': pa.array(rng.integers(0, 10**6, n)), 'rid': pa.array(np.arange(n, dtype=np.int64))})
    return dim, fact


@pytest.mark.parametrize('jtype,pa_type', [('semi', 'left semi'), ('anti', 'left anti'), ('left', 'left outer')])
@pytest.mark.parametrize('dense', [True, False])
def test_join_semi_anti_left(gpu_engine, jtype, pa_type, dense):
    """Left semi / anti / outer joins against pyarrow: NULL probe keys match nothing (semi drops them, anti and outer keep
    them), the unmatched side of an outer join reads NULL in every build column - also through a later filter + sort."""
    dim, fact = _join_tables()
    gpu_engine.register_arrow('jd', dim)
    gpu_engine.register_arrow('jf', fact)
    build = {'op': 'filter', 'cond': fn('lt', col('w'), lit(600)), 'child': {'op': 'scan', 'table': 'jd', 'columns': ['k', 'name', 'w', 'brow']}}
    plan = {'op': 'sort', 'keys': [{'expr': col('rid')}],
            'child': {'op': 'join', 'type': jtype, 'probe': {'op': 'scan', 'table': 'jf', 'columns': ['fk', 'v', 'rid']}, 'build': build,
                      'probe_key': 'fk', 'build_key': 'k'}}
    got = pa.Table.from_batches([gpu_engine.execute(plan, options={'dense_tables': dense})])
    d2 = dim.filter(pc.less(dim['w'], 600))
    exp = fact.join(d2, keys='fk', right_keys='k', join_type=pa_type).sort_by('rid')
    assert got.num_rows == exp.num_rows
    cols = ['fk', 'v', 'rid'] + (['name', 'w', 'brow'] if jtype == 'left' else [])
    assert got.schema.names == ['fk', 'v', 'rid'] + (['k', 'name', 'w', 'brow'] if jtype == 'left' else [])
    for c in cols:
        assert got[c].to_pylist() == exp[c].to_pylist(), c
    if jtype == 'left':       # the build key itself: the probe key where a match exists, NULL elsewhere
        assert got['k'].to_pylist() == [f if b is not None else None for f, b in zip(exp['fk'].to_pylist(), exp['brow'].to_pylist())]
    if jtype == 'left':
        # a predicate over the NULL-extended side and an aggregate above the outer join
        plan2 = {'op': 'aggregate', 'keys': [], 'aggs': [{'name': 'n', 'fn': 'count_star'}, {'name': 'cw', 'fn': 'count', 'expr': col('w')},
                                                          {'name': 'sw', 'fn': 'sum', 'expr': col('w')}],
                 'child': {'op': 'filter', 'cond': fn('or', fn('isnull', col('name')), fn('gt', col('w'), lit(300))), 'child': plan['child']}}
        got2 = pa.Table.from_batches([gpu_engine.execute(plan2, options={'dense_tables': dense})]).to_pylist()
        keep = pc.fill_null(pc.or_kleene(pc.is_null(exp['name']), pc.greater(exp['w'], 300)), False)
        e2 = exp.filter(keep)
        assert got2 == [{'n': e2.num_rows, 'cw': pc.count(e2['w']).as_py(), 'sw': pc.sum(e2['w']).as_py()}]
    gpu_engine.drop_table('jd')
    gpu_engine.drop_table('jf')


@pytest.mark.parametrize('jtype,pa_type', [('inner', 'inner'), ('left', 'left outer'), ('semi', 'left semi'), ('anti', 'left anti')])
@pytest.mark.parametrize('dense', [True, False])
def test_join_repeated_build_keys(gpu_engine, jtype, pa_type, dense):
    """Build keys that repeat (and NULL build keys): inner / left joins expand every probe row by its matches
    (unique_build=false -> chained table, count pass + write pass), semi / anti joins only ask for existence."""
    dim, fact = _join_tables(seed=23, dup=True)
    gpu_engine.register_arrow('jd', dim)
    gpu_engine.register_arrow('jf', fact)
    node = {'op': 'join', 'type': jtype, 'probe': {'op': 'scan', 'table': 'jf', 'columns': ['fk', 'v', 'rid']},
            'build': {'op': 'scan', 'table': 'jd', 'columns': ['k', 'w', 'brow']}, 'probe_key': 'fk', 'build_key': 'k'}
    if jtype in ('inner', 'left'):
        node['unique_build'] = False
    with_build = jtype in ('inner', 'left')
    keys = [{'expr': col('rid')}] + ([{'expr': col('brow')}] if with_build else [])
    got = pa.Table.from_batches([gpu_engine.execute({'op': 'sort', 'keys': keys, 'child': node}, options={'dense_tables': dense})])
    exp = fact.join(dim.filter(pc.is_valid(dim['k'])), keys='fk', right_keys='k', join_type=pa_type)
    exp = exp.sort_by([('rid', 'ascending')] + ([('brow', 'ascending')] if with_build else []))
    assert got.num_rows == exp.num_rows and (jtype != 'inner' or got.num_rows > fact.num_rows // 2)
    for c in ['fk', 'v', 'rid'] + (['w', 'brow'] if with_build else []):
        assert got[c].to_pylist() == exp[c].to_pylist(), c
    assert ('k' in got.schema.names) == with_build
    gpu_engine.drop_table('jd')
    gpu_engine.drop_table('jf')


def test_distinct_is_an_aggregate_without_aggregates(gpu_engine, tbl):
    got = pa.Table.from_batches([gpu_engine.execute({'op': 'sort', 'keys': [{'expr': col('g')}, {'expr': col('s')}], 'child': {
        'op': 'aggregate', 'keys': [{'name': 'g', 'expr': col('g')}, {'name': 's', 'expr': col('s')}], 'aggs': [],
        'child': {'op': 'scan', 'table': 't', 'columns': ['g', 's']}}})])
    exp = sorted(set(zip(tbl['g'].to_pylist(), tbl['s'].to_pylist())), key=lambda r: ((r[0] is not None, r[0] or 0), (r[1] is not None, (r[1] or '').encode())))
    assert list(zip(got['g'].to_pylist(), got['s'].to_pylist())) == exp


@pytest.mark.parametrize('k', [2, 3])
def test_hash_exchange_loopback_keeps_whole_groups(gpu_engine, tbl, k):
    """The hash-partition exchange on one GPU (exchange_loopback=k: k ranks that all hold this rank's rows, so this rank
    receives partition 0 of each): every row arrives k times or not at all, rows with equal keys (NULL keys included) go
    the same way, and about 1/k of the keys land here - integers, decimals, strings and nulls intact."""
    plan = {'op': 'exchange', 'mode': 'hash', 'keys': ['g', 's'],
            'child': {'op': 'filter', 'cond': fn('ge', col('a'), lit(-40)),
                      'child': {'op': 'scan', 'table': 't', 'columns': ['a', 'b', 's', 'd', 'g']}}}
    got = pa.Table.from_batches([gpu_engine.execute(plan, options={'exchange_loopback': k})])
    keep = pc.fill_null(pc.greater_equal(tbl.column('a'), -40), False)
    one = tbl.filter(keep).select(['a', 'b', 's', 'd', 'g'])
    names = one.schema.names
    from collections import Counter
    src = Counter(zip(*[one.column(n).to_pylist() for n in names]))
    out = Counter(zip(*[got.column(n).to_pylist() for n in names]))
    assert all(r in src and c == k * src[r] for r, c in out.items())           # whole rows, k copies each
    here = {(r[4], r[2]) for r in out}                                         # keys (g, s) that landed on this rank
    assert all(((r[4], r[2]) in here) == (r in out) for r in src)              # a key's rows all go the same way
    all_keys = {(r[4], r[2]) for r in src}
    assert 0.2 / k < len(here) / len(all_keys) < min(1.0, 2.5 / k)
    # an aggregate over the partition: groups are complete, so per-group sums are exactly k times the single-rank sums
    agg = lambda child: {'op': 'sort', 'keys': [{'expr': col('g')}, {'expr': col('s')}],
                         'child': {'op': 'aggregate', 'keys': [{'name': 'g', 'expr': col('g')}, {'name': 's', 'expr': col('s')}],
                                   'aggs': [{'name': 'sb', 'fn': 'sum', 'expr': col('b')}, {'name': 'c', 'fn': 'count_star'}], 'child': child}}
    full = {(r['g'], r['s']): (r['sb'], r['c']) for r in pa.Table.from_batches([gpu_engine.execute(agg(plan['child']))]).to_pylist()}
    part = pa.Table.from_batches([gpu_engine.execute(agg(plan), options={'exchange_loopback': k})]).to_pylist()
    assert len(part) == len(here)
    for r in part:
        sb, c = full[(r['g'], r['s'])]
        assert r['c'] == k * c and r['sb'] == (None if sb is None else k * sb)
