"""Where the end-to-end (host-source, public API) time of each query goes (development aid).
usage: python tools/e2e_breakdown.py [--sf 10] [--reps 5]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--sf', type=float, default=10)
    ap.add_argument('--sales_days', type=int, default=1823)
    ap.add_argument('--reps', type=int, default=5)
    a = ap.parse_args()
    import torch
    import bench
    import __graft_entry__ as entry
    entry.build()
    from spark_rapids_benchmarks_b200 import nds_schema, planner
    from spark_rapids_benchmarks_b200.session import Session, DataFrame

    class A:
        pass
    args = A()
    args.sf, args.sales_days, args.data_dir = a.sf, a.sales_days, ''
    bench.ensure_data(args, print)
    qd = bench.make_stream(args)
    wh = bench.data_dir(args)
    sess = Session('e2e', device=0, keep_host=True)
    for t in nds_schema.get_schemas(True):
        sess.register_table(t, os.path.join(wh, t))
    sess.source = 'host'
    for name, sql in qd.items():
        for _ in range(3):
            sess.sql(sql).collect()
        tp = te = tc = 0.0
        dev = host = h2d = 0.0
        for _ in range(a.reps):
            t0 = time.perf_counter()
            plan = planner.plan_sql(sql)
            t1 = time.perf_counter()
            tbl = sess.execute_plan(plan)
            t2 = time.perf_counter()
            DataFrame(tbl).collect()
            t3 = time.perf_counter()
            st = sess.last_stats
            tp += t1 - t0
            te += t2 - t1
            tc += t3 - t2
            dev += st['device_ms']
            host += st['host_ms']
            h2d = st['h2d_bytes']
        r = a.reps
        print('%-8s plan %.3f ms | execute_plan %.3f ms (engine host %.3f, device %.3f, h2d %.1f MB -> %.1f GB/s if serial) | collect %.3f ms' %
              (name, tp / r * 1e3, te / r * 1e3, host / r, dev / r, h2d / 1e6, h2d / 1e9 / max(dev / r / 1e3, 1e-9), tc / r * 1e3))
    # PCIe: one stream vs two streams
    for nstreams in (1, 2, 4):
        mb = 256
        hs = [torch.empty((mb // nstreams) << 20, dtype=torch.uint8).pin_memory() for _ in range(nstreams)]
        ds = [torch.empty((mb // nstreams) << 20, dtype=torch.uint8, device='cuda') for _ in range(nstreams)]
        ss = [torch.cuda.Stream() for _ in range(nstreams)]
        torch.cuda.synchronize()
        for it in range(2):
            t0 = time.perf_counter()
            for _ in range(5):
                for h, d, s in zip(hs, ds, ss):
                    with torch.cuda.stream(s):
                        d.copy_(h, non_blocking=True)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        print('H2D %d MB over %d stream(s): %.1f GB/s' % (mb, nstreams, 5 * mb * 1.048576 / 1e3 / dt))


if __name__ == '__main__':
    main()
