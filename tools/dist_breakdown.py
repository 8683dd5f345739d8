"""Stage timings of the distributed query path (run under torchrun, development aid).
usage: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dist_breakdown.py"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as tdist
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    tdist.init_process_group('nccl', device_id=torch.device('cuda', local))
    import bench
    from spark_rapids_benchmarks_b200 import nds_schema, planner, dist
    from spark_rapids_benchmarks_b200.session import Session

    class A:
        pass
    args = A()
    args.sf, args.sales_days, args.data_dir = 10, 1823, ''
    if rank == 0:
        bench.ensure_data(args, print)
    tdist.barrier()
    qd = bench.make_stream(args)
    wh = bench.data_dir(args)
    sess = Session('dist', device=local, rank=rank, world=world, keep_host=False)
    for t in nds_schema.get_schemas(True):
        sess.register_table(t, os.path.join(wh, t))
    plans = {n: planner.plan_sql(q, distributed=True) for n, q in qd.items()}
    for name, plan in plans.items():
        for rep in range(4):
            tdist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            partial = sess.engine.execute({'root': plan['partial']})
            t1 = time.perf_counter()
            payload = dist.serialize_batch(partial)
            t2 = time.perf_counter()
            parts = dist.all_gather_bytes(payload, None)
            t3 = time.perf_counter()
            import pyarrow as pa
            tables = [dist.deserialize_table(b) for b in parts]
            merged = pa.concat_tables(tables).combine_chunks()
            t4 = time.perf_counter()
            sess.engine.register_arrow('__partial__', merged)
            t5 = time.perf_counter()
            out = sess.engine.execute({'root': plan['root']})
            t6 = time.perf_counter()
            sess.engine.drop_table('__partial__')
            t7 = time.perf_counter()
        if rank == 0:
            print('%-8s partial %.3f | serialize %.3f (%d B, %d rows) | all_gather %.3f | deserialize %.3f | register %.3f | merge %.3f | drop %.3f  = %.3f ms' %
                  (name, (t1 - t0) * 1e3, (t2 - t1) * 1e3, len(payload), partial.num_rows, (t3 - t2) * 1e3, (t4 - t3) * 1e3, (t5 - t4) * 1e3,
                   (t6 - t5) * 1e3, (t7 - t6) * 1e3, (t7 - t0) * 1e3), flush=True)
    tdist.destroy_process_group()


if __name__ == '__main__':
    main()
