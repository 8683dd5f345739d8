"""Worker of tests/test_gpu_queries.py::test_exchange_over_nccl_two_gpus (launched with torchrun, one rank per GPU)."""
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as tdist
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    tdist.init_process_group('nccl', device_id=torch.device('cuda', local))
    from oracle import nds_queries as oq
    from spark_rapids_benchmarks_b200 import nds_schema, streams
    from spark_rapids_benchmarks_b200.session import Session
    wh = os.environ['NDS_TEST_WAREHOUSE']
    sess = Session('nccl-exchange-test', device=local, rank=rank, world=world, keep_host=True)
    assert sess.engine.has_comm, 'engine communicator was not initialised'
    for t in nds_schema.get_schemas(True):
        sess.register_table(t, os.path.join(wh, t))
    ok = 0
    for query in streams.SUBSET:
        sql = streams.render(query, overrides={'MANUFACT': 606, 'MANAGER': 63, 'MONTH': 1, 'YEAR': 1998})
        for source in ('hbm', 'host'):
            sess.source = source
            got = sess.sql(sql).toArrow().to_pylist()
            # the default exchange between the GPUs of a node: rows pushed into the peers' windows over NVLink (CUDA IPC)
            assert any(o['op'] == 'exchange:peer' for o in sess.last_stats['ops']), 'the peer-push exchange did not run: %s' % [o['op'] for o in sess.last_stats['ops']]
            if rank == 0:
                exp = oq.run_query(query, sql, wh).to_pylist()
                assert len(got) == len(exp), (query, len(got), len(exp))
                for g, e in zip(got, exp):
                    for k in e:
                        if isinstance(e[k], float) and g[k] is not None:
                            assert math.isclose(g[k], e[k], rel_tol=1e-5), (query, k, g[k], e[k])
                        else:
                            assert g[k] == e[k], (query, k, g[k], e[k])
        ok += 1
    # the one-collective form of the exchange (fixed-capacity segments, row counts in their headers, no host round trip)
    # gives the same rows as the default two-step form
    import pyarrow as pa
    for query in ('query3', 'query7'):
        plan = sess.plan(streams.render(query, overrides={'MANUFACT': 606, 'MANAGER': 63, 'MONTH': 1, 'YEAR': 1998}))
        two = pa.Table.from_batches([sess.engine.execute({'root': plan['fused']}, options={'exchange_peer': False})]).to_pylist()
        assert any(o['op'] == 'exchange' for o in sess.engine.last_stats()['ops']), 'the NCCL exchange did not run'
        one = pa.Table.from_batches([sess.engine.execute({'root': plan['fused']}, options={'exchange_peer': False, 'exchange_single': True})]).to_pylist()
        peer = pa.Table.from_batches([sess.engine.execute({'root': plan['fused']})]).to_pylist()
        assert one == two and peer == two and len(one) > 0, query
    # Spark's own plan shape: partial aggregate -> hash-partition all-to-all (grouped ncclSend / ncclRecv) -> every rank merges and
    # ranks its own groups -> gather of the per-rank top rows -> final ranking.  Same rows as the all-gather form.
    for query in streams.SUBSET:
        plan = sess.plan(streams.render(query, overrides={'MANUFACT': 606, 'MANAGER': 63, 'MONTH': 1, 'YEAR': 1998}))
        two = pa.Table.from_batches([sess.engine.execute({'root': plan['fused']})]).to_pylist()
        hashed = pa.Table.from_batches([sess.engine.execute({'root': plan['fused_hash']})]).to_pylist()
        assert any(o['op'] == 'exchange:hash' for o in sess.engine.last_stats()['ops']), 'the all-to-all did not run'
        assert len(hashed) == len(two) and len(two) > 0, query
        for a, b in zip(hashed, two):
            for k in b:
                if isinstance(b[k], float) and a[k] is not None:
                    assert math.isclose(a[k], b[k], rel_tol=1e-5), (query, k)
                else:
                    assert a[k] == b[k], (query, k, a[k], b[k])
    sess.conf['b200.agg_exchange'] = 'hash'
    via_session = sess.sql(streams.render('query42', overrides={'MANAGER': 63, 'MONTH': 1, 'YEAR': 1998})).toArrow().to_pylist()
    assert any(o['op'] == 'exchange:hash' for o in sess.last_stats['ops'])
    sess.conf['b200.agg_exchange'] = 'gather'
    assert via_session == sess.sql(streams.render('query42', overrides={'MANAGER': 63, 'MONTH': 1, 'YEAR': 1998})).toArrow().to_pylist()
    # schemas must not depend on the GPU count: sum(decimal(7,2)) stays decimal(17,2) through the merge
    sess.source = 'hbm'
    t3 = sess.sql(streams.render('query3', overrides={'MANUFACT': 606, 'MONTH': 1})).toArrow()
    assert str(t3.schema.field('sum_agg').type) == 'decimal128(17, 2)', t3.schema
    # a query over a replicated table: every rank computes the whole answer, nothing is exchanged, nothing is multiplied
    import pyarrow.compute as pc
    got = sess.sql('select i_category, count(*) cnt, sum(i_manager_id) s from item group by i_category order by i_category').toArrow()
    item = oq.read_dim(wh, 'item', ['i_category', 'i_manager_id'])
    exp = item.group_by('i_category').aggregate([([], 'count_all'), ('i_manager_id', 'sum')]).sort_by([('i_category', 'ascending')])
    rows = {r['i_category']: (r['count_all'], r['i_manager_id_sum']) for r in exp.to_pylist()}
    assert {r['i_category']: (r['cnt'], r['s']) for r in got.to_pylist()} == rows, (got.to_pylist()[:3], exp.to_pylist()[:3])
    tdist.barrier()
    if rank == 0:
        print('EXCHANGE-OK %d' % ok, flush=True)
    tdist.destroy_process_group()


if __name__ == '__main__':
    main()
