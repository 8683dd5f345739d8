"""Development aid: why did a fact scan not take the fused scan -> probe route?  Prints the `fused_scan_declined` notes
(profile mode) of the six queries on a small warehouse.  usage: python tools/fused_why.py [--sf 1] [--sales_days 182]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--sf', type=float, default=1)
    ap.add_argument('--sales_days', type=int, default=182)
    ap.add_argument('--month', type=int, default=3)
    ap.add_argument('--loopback', type=int, default=0, help='run the distributed (partial -> exchange -> merge) plan with exchange_loopback=k')
    ap.add_argument('--trace', action='store_true', help='production schedule (side streams) with the host timeline of every query')
    a = ap.parse_args()
    import __graft_entry__ as entry
    entry.build()
    from spark_rapids_benchmarks_b200 import datagen, engine, planner, streams
    sf = int(a.sf) if a.sf == int(a.sf) else a.sf
    wh = '/tmp/nds_why_sf%s_d%d' % (sf, a.sales_days)
    tables = list(datagen.SUBSET_TABLES)
    datagen.ensure(wh, sf, tables=tables, sales_days=a.sales_days)
    eng = engine.Engine(device=0)
    for t in tables:
        eng.register_table(t, os.path.join(wh, t))
    for q in streams.SUBSET:
        plan = planner.plan_sql(streams.render(q, overrides={'MONTH': a.month, 'YEAR': 1998, 'MANUFACT': 606, 'MANAGER': 63}), distributed=a.loopback > 0)
        if a.loopback > 0:
            plan['root'] = plan['fused']
        if a.trace:
            for _ in range(3):
                out = eng.execute({'root': plan['root']}, options=dict({'trace': True}, **({'exchange_loopback': a.loopback} if a.loopback else {})))
            st = eng.last_stats()
            print('== %s: %d rows, device %.3f ms, host %.3f ms, %d launches' % (q, out.num_rows, st['device_ms'], st.get('host_ms', 0), st['launches']))
            prev = 0.0
            for label, t, wait in st.get('trace', []):
                if label in ('end', 'scan:col_alloc'):
                    continue
                print('      t=%8.3f ms (+%.3f)  %s%s' % (t, t - prev, label, '   waited %.3f ms' % wait if wait >= 0 else ''))
                prev = t
            continue
        out = eng.execute({'root': plan['root']}, profile=True, options={'side_stream': False})
        st = eng.last_stats()
        print('== %s: %d rows, device %.3f ms, %d launches' % (q, out.num_rows, st['device_ms'], st['launches']))
        for o in st['ops']:
            if 'declined' in o['op'] or o['op'].split(':')[0] in ('scan_probe', 'gather_decode', 'def_levels', 'decode_kernel', 'join_probe_chain', 'aggregate', 'sort'):
                gbs = (o['bytes'] / (o['ms'] * 1e-3) / 1e9) if o['ms'] > 0 and o['bytes'] else 0
                print('   %-90s %8.3f ms %12d B %10d rows %8.1f GB/s' % (o['op'][:90], o['ms'], o['bytes'], o['rows'], gbs))


if __name__ == '__main__':
    main()
