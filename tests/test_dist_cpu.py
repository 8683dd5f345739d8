"""CPU (gloo, world_size 2): the host-side exchange of partial aggregate states, and the file-sharding
rule the engine uses (file i of a partitioned table belongs to rank i % world)."""
import os

import pyarrow as pa
import pytest
import torch.multiprocessing as mp


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from spark_rapids_benchmarks_b200 import dist as ndist
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:%d' % port, rank=rank, world_size=world)
    import decimal
    if rank == 0:
        b = pa.record_batch({'k': pa.array(['a', None, 'ccc']), 's': pa.array([decimal.Decimal('1.50'), None, decimal.Decimal('-2.25')],
                                                                                type=pa.decimal128(17, 2)),
                             'c': pa.array([1, 0, 7], type=pa.int64())})
    else:
        b = pa.record_batch({'k': pa.array([], type=pa.string()), 's': pa.array([], type=pa.decimal128(17, 2)),
                             'c': pa.array([], type=pa.int64())})
    t = ndist.all_gather_batches(b)
    assert t.num_rows == 3 and t['k'].to_pylist() == ['a', None, 'ccc']
    b2 = pa.record_batch({'x': pa.array(list(range(rank * 1000, rank * 1000 + 500 * (rank + 1))), type=pa.int64())})
    t2 = ndist.all_gather_batches(b2)
    assert t2['x'].to_pylist() == list(range(0, 500)) + list(range(1000, 2000))
    # payloads beyond the inline size take the second collective (only rank 1 is long here)
    big = bytes(range(256)) * (3 * ndist.INLINE_BYTES // 256) if rank == 1 else b'short-%d' % rank
    parts = ndist.all_gather_bytes(big)
    assert parts[0] == b'short-0' and len(parts[1]) == 3 * ndist.INLINE_BYTES and parts[1][:256] == bytes(range(256))
    assert parts[1][-1] == 255
    open(os.path.join(out_dir, 'ok%d' % rank), 'w').close()
    dist.destroy_process_group()


def test_all_gather_batches_gloo(tmp_path):
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / 'ok0') and os.path.exists(tmp_path / 'ok1')


def test_shards_partition_the_fact_table(warehouse):
    from oracle import nds_queries as oq
    cols = ['ss_sold_date_sk', 'ss_item_sk', 'ss_ticket_number']
    whole = oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', cols)
    parts = [oq.read_fact(warehouse, 'store_sales', 'ss_sold_date_sk', cols, shard=(r, 3)) for r in range(3)]
    assert sum(p.num_rows for p in parts) == whole.num_rows
    assert sorted(pa.concat_tables(parts)['ss_ticket_number'].to_pylist()) == sorted(whole['ss_ticket_number'].to_pylist())


def test_partial_states_merge_to_the_global_answer(warehouse):
    """sum/count partials per shard, merged by key, equal the unsharded oracle (the algebra the
    distributed plan relies on)."""
    from oracle import nds_queries as oq
    from spark_rapids_benchmarks_b200 import streams
    sql = streams.render('query7', overrides={'YEAR': 1998, 'MONTH': 1})
    whole = oq.run_query('query7', sql, warehouse).to_pylist()
    merged = {}
    vals = ['ss_quantity', 'ss_list_price', 'ss_coupon_amt', 'ss_sales_price']
    for r in range(2):
        rows, _, _ = oq.q7(sql, warehouse, shard=(r, 2), partial=True)
        for row in rows:
            m = merged.setdefault(row['i_item_id'], {v: [None, 0] for v in vals})
            for v in vals:
                if row[v + '_sum'] is not None:
                    m[v][0] = row[v + '_sum'] if m[v][0] is None else m[v][0] + row[v + '_sum']
                m[v][1] += row[v + '_count']
    out = []
    for k, m in merged.items():
        q = m['ss_quantity']
        out.append({'i_item_id': k, 'agg1': float(q[0]) / q[1] if q[1] else None, 'agg2': oq._avg_dec(*m['ss_list_price']),
                    'agg3': oq._avg_dec(*m['ss_coupon_amt']), 'agg4': oq._avg_dec(*m['ss_sales_price'])})
    got = oq.order_rows(out, ['i_item_id', 'agg1', 'agg2', 'agg3', 'agg4'], [('i_item_id', False)], 100)
    assert got == whole
