// Engine internals: table registry (encoded Parquet column chunks staged in HBM), plan executor.
#pragma once
#include <map>
#include <memory>
#include <set>
#include "common.h"
#include "json.h"
#include "ops.cuh"
#include "parquet_meta.h"
#include <vector_types.h>

// Arrow C Data Interface (https://arrow.apache.org/docs/format/CDataInterface.html)
extern "C" {
#ifndef ARROW_C_DATA_INTERFACE
#define ARROW_C_DATA_INTERFACE
struct ArrowSchema {
  const char* format;
  const char* name;
  const char* metadata;
  int64_t flags;
  int64_t n_children;
  struct ArrowSchema** children;
  struct ArrowSchema* dictionary;
  void (*release)(struct ArrowSchema*);
  void* private_data;
};
struct ArrowArray {
  int64_t length;
  int64_t null_count;
  int64_t offset;
  int64_t n_buffers;
  int64_t n_children;
  const void** buffers;
  struct ArrowArray** children;
  struct ArrowArray* dictionary;
  void (*release)(struct ArrowArray*);
  void* private_data;
};
#endif
}

namespace nds {

struct LType {                 // logical type of a column / expression
  enum Kind { INT = 0, DATE = 1, DECIMAL = 2, F64 = 3, STR = 4, BOOL = 5 } kind = INT;
  int precision = 0, scale = 0;
  bool is32 = false;           // stored as int32 on the device
  int maxlen = 0;              // strings: upper bound of the byte length when known (0 = unknown)
  DType dtype() const { return kind == STR ? DT_STR : (kind == F64 ? DT_F64 : (is32 ? DT_I32 : DT_I64)); }
  std::string str() const;
};

struct PageInfo {
  int32_t file, dict;          // dict: index into TableColumn::dicts or -1
  int64_t off;                 // payload offset in the column's staged byte buffer
  uint32_t size, num_values, def_len;
  int64_t row_in_file;
  uint8_t enc;                 // 0 PLAIN, 1 dictionary
  uint8_t bw = 0;              // dictionary index bit width (first byte of the values section)
  uint8_t fast = 0;            // 1: eligible for the split def/rows kernels (PLAIN fixed width, or few long index runs)
  // 1: value k of the page can be addressed without walking run headers - PLAIN fixed width, a one-entry dictionary
  // (bit width 0), or an index stream of back-to-back 504-value bit-packed runs (every header byte 0x7F), then at most
  // one shorter bit-packed run, then at most one closing RLE run (tail_k / tail_ix): what the fused scan -> probe
  // kernels and the late-materialising gather need (fused_scan.cu)
  uint8_t direct = 0;
  uint32_t tail_k = 0xFFFFFFFFu, tail_ix = 0;
  uint32_t direct_values = 0;  // values the uniform index stream holds (>= the page's non-null count when well formed)
};
struct DictInfo {
  int64_t off;                 // fixed width: values; BYTE_ARRAY: StrView table (chars follow)
  uint32_t count;
};
struct FileInfo {
  std::string path;
  int64_t rows = 0;
  bool part_null = false;
  int64_t part_val = 0;
};
struct TableColumn {
  std::string name;
  int phys = -1;
  LType lt;
  bool nullable = false;
  bool supported = true;
  std::string why_unsupported;
  std::vector<uint8_t> stage;          // host copy of the encoded pages while registering (file 0, or one file of a worker thread)
  std::vector<std::pair<int64_t, uint32_t>> view_tables;   // while registering: (offset, entries) of the string-view tables in `stage`
  std::vector<int64_t> part_base;      // while registering: where each file's bytes start in the column's buffer
  int64_t total_bytes = 0;
  uint8_t* h_enc = nullptr;            // pinned (cudaHostAlloc) host copy kept when keep_host: source of source=host scans
  bool pinned = false;
  uint8_t* d_enc = nullptr;            // resident copy in HBM
  int64_t enc_bytes = 0;
  std::vector<PageInfo> pages;
  std::vector<DictInfo> dicts;
  std::vector<int64_t> file_page_begin;  // per file: first page index (size nfiles + 1)
  std::vector<int64_t> file_byte_begin;  // per file: first staged byte (16 B aligned), size nfiles + 1
  // One PageTask per page with everything that does not depend on the query filled in at registration: src / dict
  // hold OFFSETS into the encoded buffer (dict: offset + 1, 0 = no dictionary), out_row the page's first row inside its
  // file, vr != null marks a split-route page. A scan copies a file's run of templates and patches five fields.
  std::vector<PageTask> task_tmpl;
  std::vector<int64_t> file_bytes_in;    // per file: page payload bytes + dictionary bytes a scan of it reads
  // Device copies for the fused scan (fused_scan.cu, expand_tasks_kernel): the templates, and per page its offset inside
  // its file's validity / rank scratch (words) and the index of its first 2048-row chunk inside its file; per file the
  // totals a scan needs to place a file's pages without touching them on the host.
  PageTask* d_tmpl = nullptr;
  uint2* d_page_aux = nullptr;           // x: vr words before this page in its file, y: chunks before this page in its file
  std::vector<int32_t> file_vr_words, file_chunks, file_max_def;
  std::vector<int64_t> file_def_bytes;
  std::vector<uint2> aux_stage;          // host copy of d_page_aux while registering
  std::vector<uint8_t> file_direct;      // per file: every page is directly addressable (PageInfo::direct) and def_len <= 8192
  // value range from the Parquet column-chunk statistics (INT32 / INT64 columns; valid only when every chunk had them):
  // bounds the direct-address join table built over this column (ops.cu, dense tables)
  bool has_stats = false;
  int64_t stat_lo = 0, stat_hi = -1;
  int stat_chunks = 0, total_chunks = 0;
};
struct MemColumn {                      // table registered from Arrow buffers
  std::string name;
  LType lt;
  void* d_data = nullptr;
  uint32_t* d_valid = nullptr;
  uint8_t* d_chars = nullptr;
};
struct Table {
  std::string name, path;
  std::vector<FileInfo> files;
  std::vector<TableColumn> cols;
  std::map<std::string, int> col_index;
  std::string part_col;                 // hive partition column ("" if none)
  int64_t rows = 0;
  int64_t encoded_bytes = 0;
  bool is_mem = false;
  std::vector<MemColumn> mem_cols;
};

struct Column {
  std::string name;
  LType lt;
  const void* data = nullptr;
  const uint32_t* valid = nullptr;
  const uint32_t* rowmap = nullptr;
  const uint8_t* strbase = nullptr;
  // value range promised by the Parquet statistics of the table column this one still is (unchanged values, possibly
  // fewer rows); lost by anything computed.  Lets a join build a direct-address table over it (ops.cuh, HashTable).
  bool has_range = false;
  int64_t range_lo = 0, range_hi = -1;
  ColRef ref() const {
    ColRef r;
    r.data = data; r.valid = valid; r.rowmap = rowmap; r.strbase = strbase; r.dtype = lt.dtype(); r.pad = 0;
    return r;
  }
};

struct Batch {
  int64_t n_max = 0;            // upper bound (exact when n_dev == nullptr)
  int* n_dev = nullptr;
  int64_t n_exact = -1;         // cached after a sync
  std::vector<Column> cols;
  RowCount rc() const { return RowCount{n_dev, n_exact >= 0 ? n_exact : n_max}; }
  int find(const std::string& name) const {
    for (size_t i = 0; i < cols.size(); ++i)
      if (cols[i].name == name) return (int)i;
    return -1;
  }
};

struct OpStat {
  std::string op;
  double ms = 0;
  int64_t bytes = 0;            // algorithmic bytes
  int64_t rows = 0;
  double host_ms = 0;           // wall-clock time the host spent inside this operator (launch + bookkeeping)
  int64_t row_bytes = 0, fixed_bytes = 0;   // bytes = fixed_bytes + rows * row_bytes once the device-side count is known
  const int* n_dev = nullptr;
  double t0 = 0;
  double start_ms = -1;         // device time from the query's first event to this operator's (profile mode)
  cudaEvent_t e0 = nullptr, e1 = nullptr;
};

typedef int (*nds_allgather_fn)(void* ctx, const void* send_dev, int64_t send_bytes, void* recv_dev,
                                int64_t max_bytes_per_rank, int64_t* recv_bytes);

constexpr int kDimLanes = 6, kBaseLane = 6, kLowLane = 7;

class Engine {
 public:
  explicit Engine(const Json& cfg);
  ~Engine();
  void register_table(const std::string& name, const std::string& path, const std::string& format);
  void register_arrow(const std::string& name, ArrowSchema* schema, ArrowArray* array);
  void drop_table(const std::string& name);
  void execute(const Json& plan, ArrowSchema* out_schema, ArrowArray* out_array);
  std::string comm_unique_id();
  void comm_init(const std::string& id_bytes, int rank, int world);
  void execute_once(const Json& plan, ArrowSchema* out_schema, ArrowArray* out_array, bool speculate);
  std::string stats_json() const;
  std::string table_info_json(const std::string& name) const;

  std::string last_error;
  int device = 0;
  int rank = 0, world = 1;

 private:
  friend struct Exec;
  cudaStream_t stream_ = nullptr;
  // lanes 0..kDimLanes-1: build sides; kBaseLane: the fact scan; kLowLane: its payload columns' def levels
  std::vector<cudaStream_t> side_streams_;   // build sides and base scan of a join chain run side by side (exec.cu: enter_side / leave_side)
  bool resident_ = true, keep_host_ = false;
  std::map<std::string, Table> tables_;
  std::vector<OpStat> stats_;
  int64_t launches_ = 0;
  int64_t h2d_bytes_ = 0, d2h_bytes_ = 0;
  double exec_ms_ = 0;
  double host_ms_ = 0;
  std::string trace_;             // options.trace: host timeline of the last execute
  uint8_t* pin_ = nullptr;      // pinned host ring for small host->device uploads (plan blobs, page tasks)
  int64_t pin_cap_ = 0;
  uint8_t* slab_ = nullptr;     // device slab behind the per-query bump allocator
  int64_t slab_cap_ = 0;
  int64_t slab_wanted_ = 0;     // > slab_cap_: a query overflowed the slab; grown before the next query runs
  void grow_slab_if_wanted();
  void* comm_ = nullptr;          // ncclComm_t of the exchange operator (nccl_dyn.h); null until comm_init
  int comm_rank_ = 0, comm_world_ = 1;
  // Peer-push exchange (ops.cuh, PeerWindow): this rank's window in HBM, its peers' windows mapped through CUDA IPC at
  // comm_init, and how many arrivals each half has been promised so far (the arrival counters are monotonic).
  uint8_t* xwin_ = nullptr;
  std::vector<uint8_t*> xpeers_;
  bool xpeer_ok_ = false;          // every rank mapped every window: the exchange may push over NVLink
  int64_t xwin_half_bytes_ = 0;
  int64_t xwin_uses_[2] = {0, 0}, xwin_seq_ = 0;
  int64_t xwin_mb_ = 128;          // cfg exchange_window_mb: size of one window half
  void ensure_window();
  void close_windows();
  int64_t spec_retries_ = 0;    // queries re-run because a device-side size assumption failed (stats: speculation_retries)
  int64_t zero_dirty_ = -1;     // bytes of the slab's zeroed prefix the last query used (-1: unknown)
  std::vector<cudaEvent_t> event_pool_;
  cudaEvent_t get_event();
  void put_event(cudaEvent_t e);
};

void set_global_error(const std::string& s);

}  // namespace nds
