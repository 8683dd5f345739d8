#!/usr/bin/env python3
"""bench.py - NDS power-run seconds on the SF10 star-join subset (BASELINE.json configs[1]).

One "step" = one power run of the six-query stream (q3, q7, q19, q42, q52, q55) over the seeded
synthetic SF10 warehouse (uncompressed Parquet, reference layout).  Prints ONE JSON line:

  value      seconds per step with the encoded column chunks already resident in HBM, plans prebuilt,
             timed with CUDA events around each step (L2 flushed before every step), max over ranks
  e2e        seconds per step through the public API (Session.sql(text).collect()): SQL planning, host->
             device copy of the needed encoded pages from pinned host memory, execution, result read-back
  roofline   the page decode (def_pages_kernel + rows_pages_kernel, the dominant kernels): algorithmic bytes /
             CUDA-event launch time vs the measured HBM peak in MEASURED_PEAKS.json; timed in extra single-stream
             steps because inside the timed region other streams' kernels share the SMs (that figure: in_step)
  cpu_baseline  the oracle's pyarrow-Acero restatement ("port", NOT Spark) on the same warehouse/stream

`--impl reference` times that CPU restatement as the reference arm (Spark itself cannot run here:
no JVM / pyspark on either box - BASELINE.md section 2).
Under torchrun (N > 1) the fact table is sharded by file across ranks (strong scaling) and the
partial aggregate states are all-gathered over NCCL.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'NDS power-run seconds (SF10 star-join subset q3,q7,q19,q42,q52,q55)'
WORKLOAD = 'NDS SF10 star-join aggregate subset (q3,q7,q19,q42,q52,q55), uncompressed Parquet, reference layout'


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--sf', type=float, default=10)
    ap.add_argument('--data_dir', default=os.environ.get('NDS_B200_DATA', ''))
    ap.add_argument('--sales_days', type=int, default=1823)
    ap.add_argument('--skip_cpu_baseline', action='store_true')
    ap.add_argument('--profile_ops', action='store_true', help='per-operator CUDA-event timings in the output')
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index=0):
        self.proc = None
        self.lines = []
        self.idx = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.idx), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits',
                                          '-lms', '200'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for l in self.lines:
            f = [x.strip() for x in l.split(',')]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), f[5:9]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': statistics.median(sm) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def data_dir(args):
    if args.data_dir:
        return args.data_dir
    sf = int(args.sf) if args.sf == int(args.sf) else args.sf
    return '/tmp/nds_b200_sf%s_d%d' % (sf, args.sales_days)


def ensure_data(args, log):
    from spark_rapids_benchmarks_b200 import datagen
    sf = int(args.sf) if args.sf == int(args.sf) else args.sf
    t0 = time.time()
    rows = datagen.ensure(data_dir(args), sf, sales_days=args.sales_days)
    log('warehouse %s ready in %.1f s (store_sales rows: %d)' % (data_dir(args), time.time() - t0, rows['store_sales']))
    return rows


def make_stream(args):
    from spark_rapids_benchmarks_b200 import nds_power, streams
    path = os.path.join(data_dir(args), 'query_0.sql')
    if int(os.environ.get('RANK', 0)) == 0 and not os.path.exists(path):
        streams.generate_stream(path + '.tmp', stream=0)
        os.replace(path + '.tmp', path)
    while not os.path.exists(path):
        time.sleep(0.1)
    return nds_power.gen_sql_from_stream(path)


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1, which pyarrow takes as its pool size: undo that for the CPU arm."""
    import pyarrow as pa
    n = os.cpu_count() or 1
    pa.set_cpu_count(n)
    pa.set_io_thread_count(max(n, 8))
    return n


def cpu_power_run(query_dict, wh):
    from oracle import nds_queries as oq
    t0 = time.perf_counter()
    rows = 0
    for name, sql in query_dict.items():
        rows += oq.run_query(name, sql, wh).num_rows
    return time.perf_counter() - t0, rows


def measure_cpu(query_dict, wh, warmup, steps):
    import pyarrow as pa
    use_all_host_threads()
    for _ in range(warmup):
        cpu_power_run(query_dict, wh)
    times = [cpu_power_run(query_dict, wh)[0] for _ in range(steps)]
    return statistics.median(times), pa.cpu_count()


def run_reference(args, rank, world):
    """Reference arm: the CPU restatement of the path (oracle, pyarrow-Acero, all host threads)."""
    if rank != 0:
        return
    log = lambda *a: print('[bench:reference]', *a, file=sys.stderr, flush=True)
    ensure_data(args, log)
    qd = make_stream(args)
    import pyarrow as pa
    use_all_host_threads()
    wh = data_dir(args)
    for _ in range(args.warmup):
        cpu_power_run(qd, wh)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_power_run(qd, wh)
    per_step = (time.perf_counter() - t0) / args.steps
    line = {'impl': 'reference', 'metric': METRIC, 'value': per_step, 'unit': 's', 'n_gpus': args.gpus, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': per_step * 1000.0, 'higher_is_better': False, 'scaling': 'strong',
            'vs_baseline': None, 'dtype': 'int64', 'data': 'synthetic',
            'config': {'workload': WORKLOAD, 'engine': 'pyarrow-Acero CPU restatement (oracle port), not Spark: no JVM/pyspark on this box',
                       'sf': args.sf, 'partition_pruning': True},
            'cpu_baseline': {'value': per_step, 'unit': 's', 'cores': pa.cpu_count(), 'kind': 'port',
                             'sample': 'the full six-query stream per step, page cache warm'},
            'e2e': {'value': per_step, 'unit': 's', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))
    if args.impl == 'reference':
        run_reference(args, rank, world)
        return
    log = (lambda *a: print('[bench]', *a, file=sys.stderr, flush=True)) if rank == 0 else (lambda *a: None)

    import torch
    import torch.distributed as tdist
    torch.cuda.set_device(local_rank)
    # stdout carries exactly one JSON line: anything libraries print there meanwhile (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        tdist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))

    def barrier():
        if world > 1:
            tdist.barrier()
        torch.cuda.synchronize()

    import __graft_entry__ as entry
    if rank == 0:
        entry.build()
        ensure_data(args, log)
    barrier()
    from spark_rapids_benchmarks_b200 import nds_schema, planner
    from spark_rapids_benchmarks_b200.session import Session
    qd = make_stream(args)
    wh = data_dir(args)

    t0 = time.time()
    sess = Session('NDS - bench', device=local_rank, rank=rank, world=world, keep_host=True)
    for t in nds_schema.get_schemas(True):
        sess.register_table(t, os.path.join(wh, t))
    setup_s = time.time() - t0
    info = sess.engine.table_info('store_sales')
    log('registered 24 tables in %.1f s; store_sales shard: %d rows, %d files, %.2f GB encoded' %
        (setup_s, info['rows'], info['files'], info['encoded_bytes'] / 1e9))

    plans = {name: planner.plan_sql(sql, distributed=world > 1) for name, sql in qd.items()}
    flush = torch.empty(512 << 20, dtype=torch.uint8, device='cuda')   # > 126 MB L2

    def resident_step(collect):
        launches, ops, eng_ms = 0, [], 0.0
        for name, plan in plans.items():
            sess.source = 'hbm'
            sess.execute_plan(plan)
            if collect:
                st = sess.last_stats
                launches += st['launches']
                eng_ms += st['device_ms']
                ops.extend(st['ops'])
        return launches, ops, eng_ms

    def e2e_step(collect=False):
        h2d = d2h = 0
        for name, sql in qd.items():
            sess.source = 'host'
            sess.sql(sql).collect()
            if collect:
                h2d += sess.last_stats['h2d_bytes']
                d2h += sess.last_stats['d2h_bytes']
        return h2d, d2h

    # ---------------------------------------------------------------- warm-up (both paths)
    for _ in range(max(args.warmup, 3)):
        resident_step(False)
        e2e_step()
    barrier()

    # ---------------------------------------------------------------- value: resident, CUDA events per step
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    total_ms, launches, ops, eng_ms = 0.0, 0, [], 0.0
    for _ in range(args.steps):
        flush.zero_()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        resident_step(False)
        torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
        total_ms += e0.elapsed_time(e1)
    barrier()
    # the same steps once more, untimed, to read the engine's statistics (kernel launches, per-operator events):
    # fetching and parsing them costs ~0.1 ms per query and does not belong in the timed region
    for _ in range(args.steps):
        flush.zero_()
        barrier()
        l, o, m = resident_step(True)
        launches += l
        ops.extend(o)
        eng_ms += m
    barrier()
    # ---------------------------------------------------------------- e2e: host buffers, public API
    t_e2e = 0.0
    h2d = d2h = 0
    for _ in range(args.steps):
        flush.zero_()
        barrier()
        t1 = time.perf_counter()
        e2e_step()
        torch.cuda.synchronize()
        t_e2e += time.perf_counter() - t1
    barrier()
    for _ in range(args.steps):          # bytes moved per step: from an identical untimed repeat (see above)
        a, b = e2e_step(True)
        h2d += a
        d2h += b
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    # ---------------------------------------------------------------- roofline pass: the same steps on ONE stream
    # In the timed region the fact-table decode shares the SMs with the build-side kernels of the other streams, so
    # its CUDA-event bracket there is not the kernel's own duration.  These extra steps (same plans, same inputs,
    # L2 flushed) keep every operator on one stream: the decode kernels run alone, as they do under ncu.
    serial_ops = []
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        for name, plan in plans.items():
            # N > 1: the rank-local part of the plan (scan -> joins -> partial aggregate) holds the decode launches
            sess.engine.execute({'root': plan.get('partial', plan['root'])}, options={'side_stream': False})
            serial_ops.extend(sess.engine.last_stats()['ops'])
    barrier()

    ms_per_step = total_ms / args.steps
    e2e_s = t_e2e / args.steps
    if world > 1:
        v = torch.tensor([ms_per_step, e2e_s], dtype=torch.float64, device='cuda')
        tdist.all_reduce(v, op=tdist.ReduceOp.MAX)
        ms_per_step, e2e_s = float(v[0]), float(v[1])
        c = torch.tensor([launches, h2d, d2h], dtype=torch.int64, device='cuda')
        tdist.all_reduce(c, op=tdist.ReduceOp.SUM)
        launches, h2d, d2h = int(c[0]), int(c[1]), int(c[2])

    # ---------------------------------------------------------------- roofline of the dominant kernel (rank 0's launches)
    dec = [o for o in serial_ops if o['op'] == 'decode_kernel' and o['ms'] > 0]
    dec_in_step = [o for o in ops if o['op'] == 'decode_kernel' and o['ms'] > 0]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    peak = peaks.get('hbm_gbs', 6650.0)
    roof = None
    if dec:
        # the dominant launch: the largest decode launch of the stream (query7's store_sales scan), one per step
        big = max(o['bytes'] for o in dec)
        top = [o for o in dec if o['bytes'] == big]
        ach = big / (sum(o['ms'] for o in top) / len(top) * 1e-3) / 1e9
        all_bytes = sum(o['bytes'] for o in dec)
        all_ms = sum(o['ms'] for o in dec)
        traffic = None
        tp = os.path.join(ROOT, 'profiles', 'decode_traffic.json')
        if os.path.exists(tp) and world == 1:
            try:
                traffic = json.load(open(tp)).get('dram_bytes_per_launch')
            except Exception:
                traffic = None
        top_in = [o for o in dec_in_step if o['bytes'] == big]
        in_us = sum(o['ms'] for o in top_in) / max(len(top_in), 1) * 1000.0
        roof = {'bound': 'hbm', 'kernel': 'def_pages_kernel + rows_pages_kernel (split page decode, one launch each)', 'achieved': ach, 'peak': peak, 'unit': 'GB/s', 'frac': ach / peak,
                'traffic': traffic, 'peak_source': 'MEASURED_PEAKS.json (measured)' if 'hbm_gbs' in peaks else 'fallback 6.65 TB/s',
                'launch': 'largest decode launch of the stream (store_sales scan of query7), %d launches' % len(top),
                'measured': 'CUDA events on the launching stream, in %d extra steps after the timed region that keep every operator on '
                            'one stream (options.side_stream=false) so that the kernels run alone; in_step is the same launch inside the '
                            'timed region, where build-side kernels on other streams share the SMs' % args.steps,
                'in_step': {'avg_launch_us': in_us, 'achieved': (big / (in_us * 1e-6) / 1e9) if in_us > 0 else None},
                'algorithmic_bytes_per_launch': big, 'avg_launch_us': sum(o['ms'] for o in top) / len(top) * 1000.0,
                'all_decode_launches': {'n': len(dec), 'achieved': all_bytes / (all_ms * 1e-3) / 1e9,
                                        'share_of_step': sum(o['ms'] for o in dec_in_step) / (eng_ms if eng_ms > 0 else 1.0)}}

    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    if rank != 0:
        if world > 1:
            tdist.destroy_process_group()
        return
    # ---------------------------------------------------------------- CPU baseline (rank 0, N = 1 only)
    cpu = None
    if world == 1 and not args.skip_cpu_baseline:
        sec, cores = measure_cpu(qd, wh, 1, 2)
        cpu = {'value': sec, 'unit': 's', 'cores': cores, 'kind': 'port',
               'sample': 'the full six-query stream (median of 2 after 1 warm-up), pyarrow-Acero restatement, not Spark'}
    line = {'metric': METRIC, 'value': ms_per_step / 1000.0, 'unit': 's', 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
            'ms_per_step': ms_per_step, 'higher_is_better': False, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'int64',
            'data': 'synthetic',
            'config': {'workload': WORKLOAD, 'sf': args.sf, 'store_sales_rows_per_rank': info['rows'], 'files_per_rank': info['files'],
                       'l2': 'flushed (512 MiB write) before every timed step', 'partition_pruning': True,
                       'table_registration_s': setup_s, 'engine_device_ms_per_step': eng_ms / args.steps,
                       'gpu_launches_counted': 'in an identical untimed repeat of the timed steps',
                       'parallelism': 'fact files sharded over %d GPU(s), dimensions replicated' % world},
            'e2e': {'value': e2e_s, 'unit': 's', 'h2d_bytes_per_step': h2d // args.steps, 'd2h_bytes_per_step': d2h // args.steps},
            'gpu_launches': launches, 'roofline': roof, 'cpu_baseline': cpu, 'clocks': clocks}
    if args.profile_ops:
        agg = {}
        for o in ops:
            a = agg.setdefault(o['op'].split(':')[0], [0.0, 0, 0])
            a[0] += o['ms']
            a[1] += o['bytes']
            a[2] += 1
        line['ops'] = {k: {'ms': v[0], 'bytes': v[1], 'n': v[2]} for k, v in agg.items()}
    print(json.dumps(line), flush=True)
    if world > 1:
        tdist.destroy_process_group()


if __name__ == '__main__':
    main()
