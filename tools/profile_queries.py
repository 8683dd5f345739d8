"""Per-query, per-operator CUDA-event timings on the bench warehouse (development aid).
usage: python tools/profile_queries.py [--sf 10] [--reps 3] [--source hbm|host] [--queries query3,query7]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--sf', type=float, default=10)
    ap.add_argument('--sales_days', type=int, default=1823)
    ap.add_argument('--reps', type=int, default=3)
    ap.add_argument('--source', default='hbm')
    ap.add_argument('--queries', default='')
    ap.add_argument('--no_profile', action='store_true')
    ap.add_argument('--decode_timing', action='store_true')
    ap.add_argument('--subset_tables', action='store_true', help='generate / register only the tables the six queries read (large SF)')
    ap.add_argument('--no_keep_host', action='store_true', help='no pinned host copy of the encoded pages (large SF)')
    ap.add_argument('--summary', default='', help='write a per-operator summary (markdown) to this path')
    ap.add_argument('--serial', action='store_true', help='one stream for every operator (kernel times without overlap)')
    ap.add_argument('--opt', action='append', default=[], help='engine option key=int (repeatable), e.g. small_scan_pages=0')
    ap.add_argument('--world', type=int, default=1, help='one rank of N on this GPU: 1/N of the fact files, distributed plans, the exchange '
                                                         'emulated in device memory (exchange_loopback=N)')
    ap.add_argument('--trace', action='store_true', help='print the host timeline (marks and sync waits) of the last rep')
    a = ap.parse_args()
    import bench
    import __graft_entry__ as entry
    entry.build()
    from spark_rapids_benchmarks_b200 import nds_schema, planner
    from spark_rapids_benchmarks_b200.session import Session

    class A:
        pass
    args = A()
    args.sf, args.sales_days, args.data_dir = a.sf, a.sales_days, ''
    needed = ['store_sales', 'date_dim', 'item', 'customer_demographics', 'promotion', 'customer', 'customer_address', 'store']
    wh = bench.data_dir(args)
    if a.subset_tables:
        from spark_rapids_benchmarks_b200 import datagen
        wh = wh + '_subset'
        t0 = time.time()
        rows = datagen.ensure(wh, int(a.sf) if a.sf == int(a.sf) else a.sf, tables=needed, sales_days=a.sales_days)
        print('warehouse %s ready in %.1f s (store_sales rows: %d)' % (wh, time.time() - t0, rows['store_sales']), flush=True)
        from spark_rapids_benchmarks_b200 import nds_power, streams
        sp = os.path.join(wh, 'query_0.sql')
        if not os.path.exists(sp):
            streams.generate_stream(sp, stream=0)
        qd = nds_power.gen_sql_from_stream(sp)
    else:
        bench.ensure_data(args, print)
        qd = bench.make_stream(args)
    sess = Session('profile', device=0, keep_host=not a.no_keep_host, rank=0, world=a.world)
    t0 = time.time()
    for t in (needed if a.subset_tables else nds_schema.get_schemas(True)):
        sess.register_table(t, os.path.join(wh, t))
    print('registered in %.1f s' % (time.time() - t0), flush=True)
    summary = {}
    names = [q for q in qd if not a.queries or q in a.queries.split(',')]
    for name in names:
        plan = planner.plan_sql(qd[name], distributed=a.world > 1)
        for rep in range(a.reps):
            t0 = time.perf_counter()
            sess.engine.execute({'root': plan['fused' if a.world > 1 else 'root']}, source=None if a.source == 'hbm' else a.source, profile=not a.no_profile, decode_timing=a.decode_timing,
                                options=dict(({'trace': True} if a.trace else {}), **({'side_stream': False} if a.serial else {}),
                                             **({'exchange_loopback': a.world} if a.world > 1 else {}),
                                             **{kv.split('=')[0]: int(kv.split('=')[1]) for kv in a.opt}) or None)
            wall = (time.perf_counter() - t0) * 1e3
            st = sess.engine.last_stats()
        print('== %s: wall %.3f ms, device %.3f ms, host(execute) %.3f ms, launches %d, h2d %d' % (name, wall, st['device_ms'], st.get('host_ms', 0), st['launches'], st['h2d_bytes']))
        for o in st['ops']:
            gbs = (o['bytes'] / (o['ms'] * 1e-3) / 1e9) if o['ms'] > 0 and o['bytes'] else 0
            if o['op'].startswith('decode_phases') or o['op'].startswith('decode_routes'):
                print('   ' + o['op'])
                continue
            print('   %-28s %9.3f ms (host %6.3f)  %12d B  %10d rows  %8.1f GB/s   [%7.3f .. %7.3f]' % (o['op'][:28], o['ms'], o.get('host_ms', 0), o['bytes'], o['rows'], gbs,
                                                                                                o.get('start_ms', -1), o.get('start_ms', -1) + o['ms']))
        if a.trace:
            print_trace(st)
        for o in st['ops']:
            if o['ms'] > 0 and o['bytes'] > 0 and not o['op'].startswith('decode:'):
                e = summary.setdefault(o['op'].split(':')[0], [0.0, 0, 0])
                e[0] += o['ms']
                e[1] += o['bytes']
                e[2] += 1
    if a.summary:
        peak = 6549.1
        try:
            peak = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json'))).get('hbm_gbs', peak)
        except Exception:
            pass
        lines = ['| operator | launches | total ms | algorithmic MB | achieved GB/s | fraction of %.1f GB/s |' % peak, '|---|---:|---:|---:|---:|---:|']
        fr = []
        for op, (ms, b, n) in sorted(summary.items(), key=lambda kv: -kv[1][0]):
            g = b / (ms * 1e-3) / 1e9
            fr.append(g / peak)
            lines.append('| %s | %d | %.3f | %.1f | %.1f | %.3f |' % (op, n, ms, b / 1e6, g, g / peak))
        fr.sort()
        lines.append('')
        lines.append('median operator fraction of the measured HBM peak: %.3f' % (fr[len(fr) // 2] if fr else 0.0))
        with open(a.summary, 'w') as f:
            f.write('\n'.join(lines) + '\n')
        print('\n'.join(lines))


def print_trace(st):
    prev = 0.0
    for label, t, wait in st.get('trace', []):
        extra = '   waited %.3f ms' % wait if wait >= 0 else ''
        print('      t=%8.3f ms (+%.3f)  %s%s' % (t, t - prev, label, extra))
        prev = t


if __name__ == '__main__':
    main()
