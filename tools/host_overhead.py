"""Where the time of one engine call goes outside the device (development aid): Python wrapper, ctypes, C++ host work before
the first and after the last event, Arrow import.  usage: python tools/host_overhead.py [--sf 1] [--query query42] [--reps 200]"""
import argparse
import ctypes
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--sf', type=float, default=1)
    ap.add_argument('--query', default='query42')
    ap.add_argument('--reps', type=int, default=200)
    a = ap.parse_args()
    import pyarrow as pa
    import __graft_entry__ as entry
    entry.build()
    from spark_rapids_benchmarks_b200 import datagen, engine as eng_mod, planner, streams
    from spark_rapids_benchmarks_b200.session import Session
    wh = '/tmp/nds_b200_host_overhead_sf%g' % a.sf
    tables = list(datagen.SUBSET_TABLES)
    datagen.ensure(wh, int(a.sf) if a.sf == int(a.sf) else a.sf, tables=tables, sales_days=182)
    sess = Session('overhead', device=0)
    for t in tables:
        sess.register_table(t, os.path.join(wh, t))
    plan = planner.plan_sql(streams.render(a.query))
    e = sess.engine
    payload = e.encode_plan({'root': plan['root']})
    for _ in range(20):
        sess.execute_plan(plan)
    lib, h = e._lib, e._h
    acc = {'execute_plan': 0.0, 'engine.execute': 0.0, 'buffers': 0.0, 'ctypes call': 0.0, 'arrow import': 0.0, 'from_batches': 0.0, 'device_ms': 0.0, 'host_ms(C++)': 0.0}
    for _ in range(a.reps):
        t0 = time.perf_counter()
        sess.execute_plan(plan)
        acc['execute_plan'] += time.perf_counter() - t0
        t0 = time.perf_counter()
        e.execute(payload)
        acc['engine.execute'] += time.perf_counter() - t0
        t0 = time.perf_counter()
        cs = ctypes.create_string_buffer(eng_mod._SCHEMA_BYTES)
        ca = ctypes.create_string_buffer(eng_mod._ARRAY_BYTES)
        t1 = time.perf_counter()
        rc = lib.nds_engine_execute(h, payload, ctypes.addressof(cs), ctypes.addressof(ca))
        t2 = time.perf_counter()
        assert rc == 0
        b = pa.RecordBatch._import_from_c(ctypes.addressof(ca), ctypes.addressof(cs))
        t3 = time.perf_counter()
        pa.Table.from_batches([b])
        t4 = time.perf_counter()
        acc['buffers'] += t1 - t0
        acc['ctypes call'] += t2 - t1
        acc['arrow import'] += t3 - t2
        acc['from_batches'] += t4 - t3
        st = e.last_stats()
        acc['device_ms'] += st['device_ms'] * 1e-3
        acc['host_ms(C++)'] += st['host_ms'] * 1e-3
    for k, v in acc.items():
        print('%-16s %8.1f us' % (k, v / a.reps * 1e6))
    # the engine's own host timeline of one call
    e.execute({'root': plan['root']}, options={'trace': True})
    tr = e.last_stats()['trace']
    print('trace (ms since the executor was set up):')
    for lab, t, w in tr:
        print('   %8.3f  %s%s' % (t, lab, ('  waited %.3f' % w) if w >= 0 else ''))


if __name__ == '__main__':
    main()
