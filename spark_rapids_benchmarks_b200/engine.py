"""ctypes binding over the C ABI in include/nds_b200.h (libndsb200.so).

The product path has no CPU implementation: if the shared library is missing or no CUDA device
is present, constructing an Engine raises.
"""
import ctypes
import json
import os

import pyarrow as pa

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libndsb200.so')

# every symbol include/nds_b200.h declares
SYMBOLS = ['nds_version', 'nds_last_error', 'nds_engine_create', 'nds_engine_destroy',
           'nds_engine_last_error', 'nds_engine_register_table', 'nds_engine_register_arrow',
           'nds_engine_drop_table', 'nds_engine_table_info', 'nds_engine_execute',
           'nds_engine_last_stats', 'nds_join_probe_i64_noalloc', 'nds_engine_comm_unique_id',
           'nds_engine_comm_init']
COMM_ID_BYTES = 128

_lib = None


class EngineError(RuntimeError):
    pass


def load_library(path=None):
    """dlopen libndsb200.so and declare the prototypes. Raises if the library was not built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise EngineError('%s not found: run `python -c "import __graft_entry__ as g; g.build()"` '
                          '(the NDS B200 engine has no CPU fallback)' % p)
    lib = ctypes.CDLL(p)
    vp, cp, i64 = ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64
    lib.nds_version.restype = cp
    lib.nds_last_error.restype = cp
    lib.nds_engine_create.argtypes = [cp, ctypes.POINTER(vp)]
    lib.nds_engine_destroy.argtypes = [vp]
    lib.nds_engine_destroy.restype = None
    lib.nds_engine_last_error.argtypes = [vp]
    lib.nds_engine_last_error.restype = cp
    lib.nds_engine_register_table.argtypes = [vp, cp, cp, cp]
    lib.nds_engine_register_arrow.argtypes = [vp, cp, vp, vp]
    lib.nds_engine_drop_table.argtypes = [vp, cp]
    lib.nds_engine_table_info.argtypes = [vp, cp, ctypes.c_char_p, i64]
    lib.nds_engine_execute.argtypes = [vp, cp, vp, vp]
    lib.nds_engine_last_stats.argtypes = [vp, ctypes.c_char_p, i64]
    lib.nds_join_probe_i64_noalloc.argtypes = [vp, ctypes.c_int32, vp, i64, vp, vp]
    lib.nds_engine_comm_unique_id.argtypes = [vp, ctypes.c_char_p, i64]
    lib.nds_engine_comm_init.argtypes = [vp, ctypes.c_char_p, i64, ctypes.c_int, ctypes.c_int]
    if path is None:
        _lib = lib
    return lib


_SCHEMA_BYTES = 72   # sizeof(struct ArrowSchema)
_ARRAY_BYTES = 80    # sizeof(struct ArrowArray)


class Engine:
    """One engine per process and GPU (the reference runs one SparkSession per stream process,
    /root/reference/nds/nds-throughput:23)."""

    def __init__(self, device=0, resident=True, keep_host=False, rank=0, world=1, slab_mb=None):
        self._lib = load_library()
        self._h = ctypes.c_void_p()
        cfg = {'device': device, 'resident': resident, 'keep_host': keep_host, 'rank': rank, 'world': world}
        if slab_mb is not None:
            cfg['slab_mb'] = int(slab_mb)      # initial size of the per-query device slab (grows when a query overflows it)
        cfg = json.dumps(cfg).encode()
        rc = self._lib.nds_engine_create(cfg, ctypes.byref(self._h))
        if rc != 0:
            self._h = None
            raise EngineError(self._lib.nds_last_error().decode())
        self.rank, self.world = rank, world
        self.has_comm = False

    def _check(self, rc):
        if rc != 0:
            raise EngineError(self._lib.nds_engine_last_error(self._h).decode())

    def close(self):
        if getattr(self, '_h', None):
            self._lib.nds_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def register_table(self, name, path, fmt='parquet'):
        self._check(self._lib.nds_engine_register_table(self._h, name.encode(), path.encode(), fmt.encode()))

    def register_arrow(self, name, table):
        """Upload a pyarrow Table / RecordBatch as an in-memory table."""
        if isinstance(table, pa.Table):
            table = table.combine_chunks()
            batches = table.to_batches()
            if len(batches) == 0:
                batch = pa.RecordBatch.from_arrays([pa.array([], type=f.type) for f in table.schema],
                                                   schema=table.schema)
            elif len(batches) == 1:
                batch = batches[0]
            else:
                raise EngineError('register_arrow: table did not combine into one batch')
        else:
            batch = table
        cs = ctypes.create_string_buffer(_SCHEMA_BYTES)
        ca = ctypes.create_string_buffer(_ARRAY_BYTES)
        batch._export_to_c(ctypes.addressof(ca), ctypes.addressof(cs))
        try:
            self._check(self._lib.nds_engine_register_arrow(self._h, name.encode(), ctypes.addressof(cs),
                                                            ctypes.addressof(ca)))
        finally:
            # we still own the exported structures: re-import to release them
            pa.RecordBatch._import_from_c(ctypes.addressof(ca), ctypes.addressof(cs))

    def drop_table(self, name):
        self._check(self._lib.nds_engine_drop_table(self._h, name.encode()))

    def table_info(self, name):
        buf = ctypes.create_string_buffer(1 << 20)
        self._check(self._lib.nds_engine_table_info(self._h, name.encode(), buf, len(buf)))
        return json.loads(buf.value.decode())

    def encode_plan(self, plan, source=None, profile=False, decode_timing=False, options=None):
        """Plan dict -> the JSON bytes nds_engine_execute takes (cache these when a plan is run repeatedly)."""
        if isinstance(plan, (bytes, bytearray)):
            return bytes(plan)
        if isinstance(plan, str):
            plan = json.loads(plan)
        if 'root' not in plan:
            plan = {'root': plan}
        else:
            plan = {'root': plan['root'], **({'options': plan['options']} if 'options' in plan else {})}
        if source or profile or decode_timing or options:
            opts = dict(plan.get('options', {}))
            if options:
                opts.update(options)
            if source:
                opts['source'] = source
            if profile:
                opts['profile'] = True
            if decode_timing:
                opts['decode_timing'] = True
            plan['options'] = opts
        return json.dumps(plan, separators=(',', ':')).encode()

    def execute(self, plan, source=None, profile=False, decode_timing=False, options=None):
        """Run a plan (dict, JSON string, or bytes from encode_plan) and return a pyarrow.RecordBatch."""
        payload = self.encode_plan(plan, source, profile, decode_timing, options)
        cs = ctypes.create_string_buffer(_SCHEMA_BYTES)
        ca = ctypes.create_string_buffer(_ARRAY_BYTES)
        self._check(self._lib.nds_engine_execute(self._h, payload, ctypes.addressof(cs), ctypes.addressof(ca)))
        return pa.RecordBatch._import_from_c(ctypes.addressof(ca), ctypes.addressof(cs))

    def comm_unique_id(self):
        """128 opaque bytes (an ncclUniqueId) that rank 0 hands to every other rank's comm_init."""
        buf = ctypes.create_string_buffer(COMM_ID_BYTES)
        self._check(self._lib.nds_engine_comm_unique_id(self._h, buf, COMM_ID_BYTES))
        return buf.raw

    def comm_init(self, comm_id, rank, world):
        """Collective: every rank's engine joins the NCCL communicator the plan IR's `exchange` node uses."""
        self._check(self._lib.nds_engine_comm_init(self._h, comm_id, len(comm_id), int(rank), int(world)))
        self.has_comm = world > 1

    def last_stats(self):
        buf = ctypes.create_string_buffer(1 << 20)
        self._check(self._lib.nds_engine_last_stats(self._h, buf, len(buf)))
        return json.loads(buf.value.decode())
