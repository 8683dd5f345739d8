"""Multi-GPU plumbing: one process (and one engine) per GPU under torchrun; the hive-partitioned fact
tables are sharded by file (engine cfg rank/world), dimension tables are replicated, every rank
runs scan -> joins -> PARTIAL aggregate on its shard, the partial group states are all-gathered and
every rank merges them with the final aggregate + project + sort on its own GPU.

This is the one exchange step the star-join subset has (Spark: HashAggregate(partial) ->
ShuffleExchange -> HashAggregate(final)).  torch.distributed is the
transport (NCCL over NVLink on the GPU box, gloo in the CPU tests); group states are a few KB to a
few MB, so the step is latency-bound, not bandwidth-bound.
"""
import io

import pyarrow as pa


def serialize_batch(batch):
    sink = io.BytesIO()
    with pa.ipc.new_stream(sink, batch.schema) as w:
        if batch.num_rows:
            w.write_batch(batch)
    return sink.getvalue()


def deserialize_table(buf):
    return pa.ipc.open_stream(pa.BufferReader(buf)).read_all()


INLINE_BYTES = 512 << 10     # payloads up to this size travel with the length in ONE collective


def all_gather_bytes(payload, group=None):
    """Variable-length all-gather of one bytes object per rank over torch.distributed.

    One collective in the common case: every rank sends an 8-byte length followed by the first
    INLINE_BYTES of its payload; only if some payload is longer does a second collective carry the
    remainders (padded to the longest)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    backend = dist.get_backend(group)
    dev = torch.device('cuda', torch.cuda.current_device()) if backend == 'nccl' else torch.device('cpu')
    n = len(payload)
    head = np.zeros(8 + INLINE_BYTES, dtype=np.uint8)
    head[:8] = np.frombuffer(np.int64(n).tobytes(), dtype=np.uint8)
    k = min(n, INLINE_BYTES)
    if k:
        head[8:8 + k] = np.frombuffer(payload, dtype=np.uint8, count=k)
    send = torch.from_numpy(head).to(dev)
    recv = torch.empty(world * head.size, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(recv, send, group=group)
    host = recv.cpu().numpy().reshape(world, head.size)
    sizes = [int(np.frombuffer(host[r, :8].tobytes(), dtype=np.int64)[0]) for r in range(world)]
    parts = [host[r, 8:8 + min(sizes[r], INLINE_BYTES)].tobytes() for r in range(world)]
    longest = max(sizes)
    if longest > INLINE_BYTES:
        m = longest - INLINE_BYTES
        rest = np.zeros(m, dtype=np.uint8)
        if n > INLINE_BYTES:
            rest[:n - INLINE_BYTES] = np.frombuffer(payload, dtype=np.uint8, count=n - INLINE_BYTES, offset=INLINE_BYTES)
        send2 = torch.from_numpy(rest).to(dev)
        recv2 = torch.empty(world * m, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(recv2, send2, group=group)
        host2 = recv2.cpu().numpy().reshape(world, m)
        for r in range(world):
            if sizes[r] > INLINE_BYTES:
                parts[r] += host2[r, :sizes[r] - INLINE_BYTES].tobytes()
    return parts


def all_gather_batches(batch, group=None):
    """Each rank contributes one record batch; every rank gets the concatenation (rank order)."""
    parts = [deserialize_table(b) for b in all_gather_bytes(serialize_batch(batch), group)]
    return pa.concat_tables(parts).combine_chunks()


def merge_and_finish(engine, plan, tables, source=None):
    """Final step shared by the torchrun path and the one-GPU emulation: register the gathered partial
    states as `__partial__` and run the merge aggregate + projection + sort on the GPU."""
    merged = pa.concat_tables(tables).combine_chunks() if len(tables) != 1 else tables[0].combine_chunks()
    engine.register_arrow('__partial__', merged)
    try:
        return engine.execute({'root': plan['root']})
    finally:
        engine.drop_table('__partial__')


def _add_stats(a, b):
    out = dict(b)
    for k in ('launches', 'device_ms', 'h2d_bytes', 'd2h_bytes'):
        out[k] = a.get(k, 0) + b.get(k, 0)
    out['ops'] = list(a.get('ops', [])) + list(b.get('ops', []))
    return out


def init_engine_comm(session, group=None):
    """Collective over torch.distributed: give every rank's engine the NCCL communicator of the plan IR's
    `exchange` node (rank 0 creates the id, torch.distributed carries its 128 bytes).  Only with the nccl
    backend - under gloo (CPU tests) the host path below stays in use."""
    import torch.distributed as tdist
    if tdist.get_backend(group) != 'nccl':
        return False
    box = [session.engine.comm_unique_id() if session.rank == 0 else None]
    tdist.broadcast_object_list(box, src=0, group=group)
    session.engine.comm_init(box[0], session.rank, session.world)
    return True


def execute_distributed(session, plan):
    """Engines that share a communicator run ONE plan: partial aggregate -> exchange (NCCL all-gather in device
    memory) -> merge aggregate -> project -> sort.  Otherwise (gloo, or no communicator): partial plan on this rank's
    shard -> all-gather of the partial states through the host -> merge on every rank; the summed statistics of
    both engine calls are left in session.dist_stats."""
    src = session.source if session.source != 'hbm' else None
    if plan.get('replicated'):
        # the root table is replicated on every rank: each rank computes the whole result, nothing is exchanged
        out = session.engine.execute({'root': plan['root']}, source=src)
        session.dist_stats = None
        return out
    if 'partial' not in plan:
        raise ValueError('plan was not built with distributed=True')
    if getattr(session.engine, 'has_comm', False) and 'fused' in plan:
        # b200.agg_exchange=hash: the partial states are hash-partitioned by the group keys (all-to-all) and every rank
        # merges its own groups; the default all-gathers them and every rank merges everything (cheaper for few groups)
        form = 'fused_hash' if session.conf.get('b200.agg_exchange', 'gather') == 'hash' and 'fused_hash' in plan else 'fused'
        key = '_json_%s_%s' % (form, session.source)
        payload = plan.get(key)
        if payload is None:
            payload = session.engine.encode_plan({'root': plan[form]}, source=src)
            plan[key] = payload
        out = session.engine.execute(payload)
        session.dist_stats = None          # one engine call: Session.last_stats fetches its statistics on demand
        return out
    partial = session.engine.execute({'root': plan['partial']}, source=src)
    st = session.engine.last_stats()
    gathered = all_gather_batches(partial, session.comm)
    out = merge_and_finish(session.engine, plan, [gathered])
    session.dist_stats = _add_stats(st, session.engine.last_stats())
    return out
