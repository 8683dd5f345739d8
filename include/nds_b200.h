/* nds_b200.h - C ABI of the B200-native NDS query engine (libndsb200.so).
 *
 * The reference (bilalbari/spark-rapids-benchmarks) has no FFI of its own: its hot path is the
 * py4j hand-off to Apache Spark.  Each entry point below cites the reference call it stands in
 * for.  All functions return 0 on success and a non-zero code on failure; the message is
 * available from nds_engine_last_error() (or nds_last_error() when no engine exists yet).
 * Plain pointers and sizes only; results travel through the Arrow C Data Interface.
 * There is no CPU fallback: without a CUDA device nds_engine_create() fails.
 */
#ifndef NDS_B200_H
#define NDS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define NDS_API __attribute__((visibility("default")))
#else
#define NDS_API
#endif

struct ArrowSchema;
struct ArrowArray;
typedef struct nds_engine nds_engine;

/* Library / ABI version string. */
NDS_API const char* nds_version(void);

/* Message of the last failure that happened without an engine handle (e.g. in nds_engine_create). */
NDS_API const char* nds_last_error(void);

/* Create an engine bound to one GPU.  cfg_json: {"device":0, "resident":true, "keep_host":false,
 * "rank":0, "world":1, "slab_mb":512}.  resident: encoded column chunks are kept in HBM; keep_host: a pinned host
 * copy is kept as well so that a plan can run with {"options":{"source":"host"}} (host->device copy
 * of the needed chunks inside the query).  rank/world: hive-partitioned (fact) tables are sharded by
 * file across `world` engines, dimension tables are replicated.  slab_mb: initial size of the per-query
 * device slab (it grows when a query overflows it).
 * Replaces: SparkSession.builder...getOrCreate()   /root/reference/nds/nds_power.py:224-240 */
NDS_API int nds_engine_create(const char* cfg_json, nds_engine** out);

/* Replaces: sparkContext.stop()                      /root/reference/nds/nds_power.py:279-280 */
NDS_API void nds_engine_destroy(nds_engine* e);

NDS_API const char* nds_engine_last_error(nds_engine* e);

/* Register `path` (a directory of uncompressed parquet files, optionally one level of
 * <col>=<value> hive partition directories) under `name`.  Parses footers and page headers and stages
 * the encoded pages in HBM.  format must be "parquet"; anything else fails.
 * Replaces: spark.read.format(fmt).load(path).createOrReplaceTempView(name)
 *                                                    /root/reference/nds/nds_power.py:97-100 */
NDS_API int nds_engine_register_table(nds_engine* e, const char* name, const char* path, const char* format);

/* Register an in-memory record batch (Arrow C Data Interface, struct array) as a table; the engine
 * copies the buffers to the GPU and does not take ownership of the Arrow structures.  Used for
 * operator-level tests and for the gathered partial aggregates of a multi-GPU run.
 * Replaces: spark.createDataFrame(...).createOrReplaceTempView(name) (no call site on the hot path) */
NDS_API int nds_engine_register_arrow(nds_engine* e, const char* name, struct ArrowSchema* schema, struct ArrowArray* array);

NDS_API int nds_engine_drop_table(nds_engine* e, const char* name);

/* JSON description of a registered table (rows, files, encoded bytes, columns) into buf. */
NDS_API int nds_engine_table_info(nds_engine* e, const char* name, char* buf, int64_t buflen);

/* Execute one physical plan (JSON plan IR, DESIGN.md "Plan IR") and return the result as one record
 * batch.  The caller owns out_schema/out_array and must call their release callbacks.
 * An unsupported node or expression fails with a message; nothing is computed on the CPU.
 * Optional top-level "options": {"source":"hbm"|"host", "profile":bool (CUDA events per operator in the statistics),
 * "trace":bool (host timeline), and the test / measurement switches "side_stream", "speculate", "split_decode",
 * "small_scan_pages", "exchange_loopback"} - DESIGN.md, Plan IR.
 * Replaces: df = spark.sql(query); df.collect()     /root/reference/nds/nds_power.py:130-132 */
NDS_API int nds_engine_execute(nds_engine* e, const char* plan_json, struct ArrowSchema* out_schema, struct ArrowArray* out_array);

/* Communicator of the plan IR's "exchange" node (all-gather of partial aggregate states between the per-GPU
 * processes of one node, NCCL over NVLink; libnccl.so.2 is opened at run time on first use).  Rank 0 obtains an id
 * with nds_engine_comm_unique_id, the host distributes its NDS_COMM_ID_BYTES bytes to every process (any channel:
 * torch.distributed, MPI, a file), and every process calls nds_engine_comm_init with its rank - a collective call.
 * Replaces: the shuffle between HashAggregate(partial) and HashAggregate(final) that Spark inserts for
 * `spark.sql.shuffle.partitions`   /root/reference/nds/power_run_gpu.template:29, power_run_cpu.template:27
 * (the reference has no call site of its own - the exchange is inside spark.sql()).
 * nds_engine_comm_init also allocates this engine's exchange window in device memory and maps every peer's window through
 * CUDA IPC (all ranks on one node): the default exchange then packs rows straight into the peers' windows over NVLink and
 * uses NCCL only as the fallback / second attempt.  Every rank must execute the same plans in the same order. */
#define NDS_COMM_ID_BYTES 128
NDS_API int nds_engine_comm_unique_id(nds_engine* e, char* out, int64_t out_len);
NDS_API int nds_engine_comm_init(nds_engine* e, const char* id, int64_t id_len, int rank, int world);

/* JSON statistics of the last nds_engine_execute: kernel launches, device ms, h2d/d2h bytes and
 * per-operator {op, ms, bytes, rows}. */
NDS_API int nds_engine_last_stats(nds_engine* e, char* buf, int64_t buflen);

/* Hash-join lookup that allocates nothing (no counterpart in the reference).
 *
 * For callers that must not allocate GPU memory while a query runs: an engine embedded in a host whose own
 * memory manager already holds the device (spark-rapids gives nearly all HBM to its RMM pool at executor
 * start-up, so a cudaMalloc from inside a library fails or escapes the pool's spill accounting), or a caller
 * that cannot afford the driver's allocator lock on its latency path.  nds_engine_execute draws its hash
 * tables from CUDA's memory pool; this entry point instead keeps the table in a fixed array inside the
 * library, so it calls no cudaMalloc / cudaFree / pool routine at all.
 *
 *   out_build_row_dev[i] = index of the build key equal to probe_keys_dev[i], or -1 if there is none.
 *
 * All three pointers are device pointers supplied by the caller; keys are int64, every key is valid, build
 * keys must be unique and INT64_MIN is reserved (both are reported as errors).  cuda_stream is a cudaStream_t
 * (NULL = default stream); the call returns after the work on that stream has finished.
 * Limits and sharing: at most NDS_NOALLOC_MAX_BUILD_KEYS build keys - more is refused with an error, nothing
 * is truncated.  The table (about 1.5 MB) is part of the loaded module: each GPU that uses the library holds
 * it until the process exits.  There is ONE table per GPU, reused by every call and left holding the last
 * call's keys; each call clears it first.  Two calls whose GPU work overlaps (different streams or threads)
 * would race on it - serialise them.  The first call on a device looks the table's address up
 * (cudaGetSymbolAddress, with that device current), and EVERY call ends by copying its status flags to the host and
 * synchronising cuda_stream: call it outside any stream capture, the first time and every time after.
 * On failure returns non-zero and nds_last_error() holds the message. */
#define NDS_NOALLOC_MAX_BUILD_KEYS 65536
NDS_API int nds_join_probe_i64_noalloc(const int64_t* build_keys_dev, int32_t n_build, const int64_t* probe_keys_dev,
                                       int64_t n_probe, int32_t* out_build_row_dev, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* NDS_B200_H */
