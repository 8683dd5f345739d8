// Launcher declarations for the operator kernels (ops.cu, decode.cu).
#pragma once
#include "device_types.h"
#include "expr.cuh"
#include <cuda_runtime.h>

namespace nds {

struct Prog {            // one compiled expression, resident in device memory
  const Instr* code;
  int32_t n;
  int32_t kind;          // 0 integer-like, 1 double, 2 string
};

struct OpCtx {
  const ColRef* cols;
  const uint8_t* lit;
};

struct RowCount {        // rows of a batch: device-side count if known only there
  const int* dev;        // may be null
  int64_t host;          // exact count when dev == null, else an upper bound
};

constexpr int64_t kEmptyKey = INT64_MIN;

// Join table over the build side's keys, one of two layouts:
//  * open addressing (dense == nullptr): 64-bit keys + 32-bit payload (the build-side batch row), linear probing,
//    load factor <= 0.5;
//  * direct addressing (dense != nullptr): the keys are known to lie in [lo, lo + range) - the Parquet statistics of a
//    surrogate-key column - and dense[key - lo] holds build row + 1 (0: no such key).  One 4-byte load per probe and
//    no hashing: the perfect-hash special case every NDS dimension key falls into.
// multi (next != nullptr): build keys may repeat; the slot holds the head of a chain through next[] (entries are
// build row + 1, 0 ends the chain) and a probe walks it (ht_first / ht_next).
struct HashTable {
  int64_t* keys = nullptr;
  uint32_t* vals = nullptr;
  uint32_t mask = 0;
  uint32_t range = 0;
  uint32_t* dense = nullptr;
  int64_t lo = 0;
  uint32_t* next = nullptr;
};

#ifdef __CUDACC__
// build row of `key` (a non-null key), or false
__device__ __forceinline__ bool ht_find(const HashTable& ht, int64_t key, uint32_t& row) {
  if (ht.dense) {
    const uint64_t d = (uint64_t)(key - ht.lo);
    if (d >= (uint64_t)ht.range) return false;
    const uint32_t v = __ldg(ht.dense + d);
    row = v - 1u;
    return v != 0u;
  }
  if (key == kEmptyKey) return false;
  uint32_t s = (uint32_t)mix64((uint64_t)key) & ht.mask;
  for (;;) {
    const int64_t cur = ht.keys[s];
    if (cur == key) { row = ht.next ? ht.vals[s] - 1u : ht.vals[s]; return true; }
    if (cur == kEmptyKey) return false;
    s = (s + 1) & ht.mask;
  }
}
// multi tables: first match of `key` as a chain link (build row + 1, 0 = none); ht_next follows the chain
__device__ __forceinline__ uint32_t ht_first(const HashTable& ht, int64_t key) {
  uint32_t row;
  return ht_find(ht, key, row) ? row + 1u : 0u;
}
__device__ __forceinline__ uint32_t ht_next(const HashTable& ht, uint32_t link) { return ht.next[link - 1u]; }
#endif

// AGG_COUNT_REF: count(x) where another aggregate over the same x already counts the non-null rows (AVG = SUM + COUNT):
// prog holds that aggregate's index, nothing is accumulated, the finalize step reads the other aggregate's counter.
enum AggFn : int32_t { AGG_SUM_I = 0, AGG_COUNT = 1, AGG_COUNT_STAR = 2, AGG_MIN_I = 3, AGG_MAX_I = 4, AGG_SUM_F = 5, AGG_COUNT_REF = 6 };
struct AggSpec {
  int32_t fn;
  int32_t prog;          // index into the program array, -1 for count(*)
};
constexpr int kMaxGroupKeys = 8;

struct SortKey {
  int32_t prog;
  int32_t kind;          // 0 int, 1 double, 2 string
  int32_t desc;
  int32_t nulls_first;
  int32_t width;         // payload bytes (8 for numbers, padded length for strings)
  int32_t offset;        // byte offset of this key (flag byte first) inside the encoded row
};

// decode.cu
int decode_slices(int npages);
// dbg (optional): 8 clock64 checkpoints per CTA, development aid
void launch_decode_pages(const PageTask* d_tasks, const int* d_idx, int npages, int* d_err, cudaStream_t s, long long* dbg = nullptr);
// decode_fast.cu: def kernel over d_def_idx pages, then rows kernel over d_rows_idx pages (indices into d_tasks)
void launch_decode_split(const PageTask* d_tasks, const int* d_def_idx, int ndef, int max_def_len, const int* d_rows_idx, int nrows, int* d_err,
                         cudaStream_t s, int* launches);
void launch_fill_segments(const int64_t* d_seg_start, const int64_t* d_seg_val, int nseg, int32_t* out, cudaStream_t s);

// ops.cu
void launch_filter(Prog p, OpCtx c, RowCount n, uint32_t* out_sel, int* out_count, cudaStream_t s);
void launch_project(Prog p, OpCtx c, RowCount n, const uint32_t* rows, void* out, uint32_t* out_valid, int out_dtype,
                    const uint8_t* str_out_base, int* err, cudaStream_t s);
struct ProjSpec {        // one output column of a multi-column projection
  Prog prog;
  void* out;
  uint32_t* valid;
  const uint8_t* strbase;
  int32_t dtype;
  int32_t pad;
};
void launch_project_multi(const ProjSpec* specs, int nspecs, OpCtx c, RowCount n, const uint32_t* rows, int* err, cudaStream_t s);
void launch_compose(const uint32_t* inner, const uint32_t* idx, RowCount n, uint32_t* out, cudaStream_t s);
void launch_fill_i64(int64_t* p, int64_t v, int64_t n, cudaStream_t s);
// allow_dup: repeated build keys are not an error (semi / anti joins only ask whether a key exists; multi tables chain them)
void launch_join_build(ColRef key, RowCount n, HashTable ht, int* flags, cudaStream_t s, bool allow_dup = false);
// bits: ceil((hi - lo + 1) / 32) pre-zeroed words; bit (k - lo) is set for every non-null build key k in [lo, hi]
void launch_keys_to_bitmap(ColRef key, RowCount n, int64_t lo, int64_t hi, uint32_t* bits, cudaStream_t s);
// filter + direct-address build (+ pruning bitmap) in one pass over the unfiltered build side; table values = unfiltered rows
void launch_join_build_filter(Prog pred, OpCtx c, ColRef key, RowCount n, HashTable ht, int* flags, bool allow_dup, int64_t prune_lo,
                              int64_t prune_hi, uint32_t* prune_bits, cudaStream_t s);
enum JoinKind : int { JOIN_INNER = 0, JOIN_SEMI = 1, JOIN_ANTI = 2, JOIN_LEFT = 3 };
// inner: matching probe rows + their build row; semi / anti: probe rows with / without a match (out_build unused);
// left: every probe row, out_build = the build row or kNoRow.  Unique-key tables.
void launch_join_probe(ColRef key, RowCount n, HashTable ht, uint32_t* out_probe, uint32_t* out_build, int* out_count,
                       cudaStream_t s, int kind = JOIN_INNER);
// multi tables (repeated build keys), inner / left: pass 1 counts the output rows, pass 2 writes them (any order)
void launch_join_probe_multi_count(ColRef key, RowCount n, HashTable ht, int kind, unsigned long long* out_total, cudaStream_t s);
void launch_join_probe_multi_write(ColRef key, RowCount n, HashTable ht, int kind, uint32_t* out_probe, uint32_t* out_build,
                                   unsigned long long* cursor, cudaStream_t s);
constexpr int kMaxChain = 8;
struct ProbeChain {            // a left-deep chain of inner hash joins probed in one pass over the base rows
  ColRef key[kMaxChain];
  HashTable ht[kMaxChain];
  uint32_t* out_build[kMaxChain];
  int32_t n;
  int32_t pad;
};
void launch_join_probe_chain(const ProbeChain& chain, RowCount n, uint32_t* out_probe, int* out_count, cudaStream_t s);

// fused_scan.cu: page decode -> join-chain probe in one kernel over directly addressable pages, then the payload
// columns of the surviving rows only (late materialisation).
constexpr int kFusedThreads = 256;
constexpr int kFusedChunk = 2048;          // rows of one page a CTA of scan_probe_kernel handles
struct FusedChunk {
  uint32_t page;                           // index into every key column's task array (the key columns are aligned page for page)
  uint32_t row0;                           // first row of the chunk inside the page (multiple of kFusedChunk)
  uint32_t file;                           // index of the page's file among the scan's selected files
  uint32_t pad;
};
struct FusedKey {
  const PageTask* tasks;                   // src / dict / vr patched, out_row = global first row of the page
  HashTable ht;
  uint32_t* out_build;                     // matched build row per survivor
};
struct FusedScanArgs {
  FusedKey key[kMaxChain];                 // probe order: most selective join first
  int32_t nkeys;
  int32_t has_part;                        // the chain also joins on the partition column: one lookup per file
  HashTable part_ht;
  uint32_t* part_out_build;
  const int64_t* file_part_val;            // per selected file
  const FusedChunk* chunks;
  uint32_t* out_row;                       // global row of each survivor
  uint32_t* out_file;                      // selected-file index of each survivor
  int* out_count;
  uint32_t cap;                            // capacity of the output arrays; overflow raises err[2]
  int* err;
  // Existence filter of the FIRST join for the persistent variant: bit ((key - lo) >> shift) of `bloom` is set when some
  // build key maps there (shift 0: exact).  Each CTA keeps a copy in shared memory, so that most rows are rejected without
  // touching the join table (a random 4-byte read that costs a 32-byte L2 sector per row).  nullptr: the plain variant.
  const uint32_t* bloom;
  int64_t bloom_lo;
  uint32_t bloom_range, bloom_shift, bloom_words;
  int32_t nchunks;
  // Per selected file: one bit per dictionary entry of the first key column, set when the entry is a key of the first join
  // (dict_match_kernel).  dmatch_stride words per file; nullptr = test the join table per row.
  const uint32_t* dmatch;
  uint32_t dmatch_stride;
  const int32_t* page_file;                // per page of the first key column: index of its file among the selected files
  int32_t npages, split;                   // scan_probe_dm_kernel: `split` CTAs share a page's chunks
};
// dict_match_kernel: kDictMatchEntries dictionary entries per CTA; `coarse` = existence bits of the first join's key range,
// one bit per 2^coarse_shift keys (launch_bloom_from_dense), copied to shared memory; 0 words = test the table directly.
// The match bits of one file must fit the probe kernel's shared memory: kDictMatchMaxWords.
constexpr int kDictMatchThreads = 512;
constexpr int kDictMatchEntries = 16384;
constexpr int kDictMatchCoarseWords = 8192;     // 32 KB
constexpr int kDictMatchMaxWords = 16384;       // 64 KB: dictionaries of up to 524288 entries
// Concatenation of row sets under one column (rollup levels): segment = one source column of one level, copied to the rows
// [*off, *off + n) of the destination; present == 0 writes NULLs.  The destination's validity words start out zero.
struct UnionSeg {
  const void* src;
  const uint32_t* src_valid;
  void* dst;
  uint32_t* dst_valid;
  const int* n_dev;            // row count of the level on the device, or nullptr: n_host is exact
  int64_t n_host;
  const int* off;              // first destination row (device)
  int32_t width;               // 4 or 8 bytes per value
  int32_t present;
};
void launch_union_copy(const UnionSeg* d_segs, int nsegs, int64_t max_rows, cudaStream_t s);
// off[0] = 0, off[i + 1] = off[i] + min(cnt[i], bound[i]) for i < n (one thread); bound: the host's upper bounds (device copy)
void launch_prefix_counts(const int* d_cnt, const int* d_bound, int n, int* d_off, cudaStream_t s);
constexpr int kDmQueue = 6144;                  // queued rows (uint16 row in page) before pass B must run
void launch_dict_match(const PageTask* tasks, const int32_t* dict_task, int nfiles, uint32_t max_dict, int width, const HashTable& ht,
                       const uint32_t* coarse, uint32_t coarse_words, uint32_t coarse_shift, uint32_t* out, uint32_t stride_words,
                       cudaStream_t s);
constexpr int kBloomMaxWords = 36 * 1024;      // 144 KB of shared memory for the filter (+ 25 KB for the queues)
void launch_scan_probe(const FusedScanArgs& args, int nchunks, cudaStream_t s);
// dense join table -> existence filter (bloom_words pre-zeroed words)
void launch_bloom_from_dense(const uint32_t* dense, uint32_t range, uint32_t shift, uint32_t* bloom, cudaStream_t s);
// One (column, selected file) pair of a fused scan: expand_tasks_kernel copies the file's page templates, patches them
// and lists the pages (def kernel) and their 2048-row chunks (probe kernel) - the host only places whole files.
struct FileJob {
  const PageTask* tmpl;                    // the column's templates (device), indexed from page 0 of the table
  const uint2* page_aux;
  const uint8_t* src_base;                 // address of staged byte 0 of the column for this query
  uint32_t* vr_base;                       // this file's validity / rank scratch, or nullptr (column without nulls)
  int64_t row_start;                       // global row of the file's first row
  int32_t page_begin, npages;              // the file's pages within tmpl
  int32_t out_task0;                       // where the file's first task goes in the combined task array
  int32_t col_task0;                       // first task of the column (chunk page numbers are relative to it)
  int32_t def_base;                        // where the file's pages go in the def kernel's page list, or -1
  int32_t chunk_base;                      // where the file's chunks go in the chunk list, or -1 (not the first key column)
  int32_t file;                            // index among the selected files
  int32_t pad;
};
void launch_expand_tasks(const FileJob* d_jobs, int njobs, PageTask* tasks, int* def_idx, FusedChunk* chunks, int32_t* page_file,
                         cudaStream_t s);
struct GatherCol {
  const PageTask* tasks;                   // this column's tasks for the selected files, scan order (null: the partition column)
  const int32_t* file_page0;               // per selected file: first task; [nsel] = task count
  void* out;
  uint32_t* out_valid;
  int32_t width;                           // 4 or 8 bytes per output value
  int32_t is_part;
};
void launch_gather_decode(const GatherCol* d_cols, int ncols, const uint32_t* rows, const uint32_t* files, const int* n_dev, int64_t cap,
                          const int64_t* file_part_val, const uint8_t* file_part_null, int* err, cudaStream_t s);
// Hash group-by.  slots / acc / cnt are allocated for cap_max = agg_capacity_for(upper bound of the input rows); the
// kernels size the table from the row count on the device (<= cap_max), so no host round trip is needed.
// acc / cnt: naggs * cap_max int64 each; out_valid: naggs * valid_stride_words words (cleared by launch_agg_init).
uint32_t agg_capacity_for(int64_t n_bound);
void launch_agg_init(const AggSpec* aggs, int naggs, RowCount n, uint32_t cap_max, uint32_t* slots, int64_t* acc, int64_t* cnt,
                     uint32_t* out_valid, int64_t valid_stride_words, cudaStream_t s);
void launch_agg_insert(const Prog* progs, int nkeys, const AggSpec* aggs, int naggs, OpCtx c, RowCount n, uint32_t* slots,
                       uint32_t cap_max, int64_t* acc, int64_t* cnt, cudaStream_t s);
void launch_agg_finalize(const AggSpec* aggs, int naggs, const uint32_t* slots, uint32_t cap_max, RowCount n, const int64_t* acc,
                         const int64_t* cnt, uint32_t* rep_out, int64_t* out_vals, uint32_t* out_valid, int64_t out_stride,
                         int64_t valid_stride_words, int* ngroups, cudaStream_t s);
// ranking window functions over rows sorted by (npart partition keys, then the order keys): three launches; m1 / m2 / d are
// scratch of one word per row; an output pointer may be null
void launch_rank(const Prog* progs, int nkeys, int npart, OpCtx c, RowCount n, uint32_t* m1, uint32_t* m2, uint32_t* d, int64_t* out_rank,
                 int64_t* out_rown, int64_t* out_dense, cudaStream_t s);
// window aggregates: the partition's finished state per input row (after launch_agg_init + launch_agg_insert)
void launch_win_gather(const Prog* progs, int nkeys, const AggSpec* aggs, int naggs, OpCtx c, RowCount n, const uint32_t* slots,
                       uint32_t cap_max, const int64_t* acc, const int64_t* cnt, int64_t* out_vals, uint32_t* out_valid, int64_t out_stride,
                       int64_t valid_stride_words, cudaStream_t s);
void launch_agg_keyless_fix(const AggSpec* aggs, int naggs, int* ngroups, uint32_t* rep_out, int64_t* out_vals, uint32_t* out_valid,
                            int64_t out_stride, int64_t valid_stride_words, cudaStream_t s);
void launch_str_maxlen(Prog p, OpCtx c, RowCount n, int* out_max, cudaStream_t s);
void launch_encode_keys(const Prog* progs, const SortKey* keys, int nk, OpCtx c, RowCount n, uint8_t* keybuf, int B, int* active,
                        int* err, cudaStream_t s);
// ORDER BY ... LIMIT k (k <= kTopkMaxK) in ONE launch (ops.cu, topk_stream_kernel).  scratch: topk_stream_scratch_words
// 64-bit words; with g = topk_stream_groups(ctas): fin_keys (ctas + g) * k * ceil(B / 8) words, fin_rows (ctas + g) * k,
// fin_count ctas + g ints, tickets 1 + g zeroed ints.  out_rows[0 .. min(k, n)) sorted, out_count[0] = min(k, n).
constexpr int kTopkMaxK = 128;
constexpr int kTopkMaxKeyBytes = 1024;
int topk_stream_ctas(int64_t n_bound);
int topk_stream_groups(int ctas);
int64_t topk_stream_scratch_words(int ctas, int Bw);
void launch_topk_stream(const Prog* progs, const SortKey* keys, int nk, OpCtx c, RowCount n, int B, int k, int ctas, uint64_t* scratch,
                        uint64_t* fin_keys, uint32_t* fin_rows, int* fin_count, int* tickets, uint32_t* out_rows, int* out_count, int* err,
                        cudaStream_t s);
// Stable LSD radix sort of row ids by B-byte big-endian keys (8-bit digits).
//  * launch_radix_sort_small: one CTA, rows <= kSmallSortMax, skips bytes whose active[] flag is 0; the index of
//    the result buffer (0: perm_a, 1: perm_b) is written to which[0].
//  * launch_radix_sort: multi-CTA, one hist/scan/scatter triple per listed byte (passes[] in LSD order);
//    which[] must hold npasses + 1 ints; result buffer index is which[npasses]. scratch: grid*256 + npasses + 16 uint32.
constexpr int kSmallSortMax = 8192;
int radix_sort_grid(int64_t n_max);
void launch_radix_sort_small(const uint8_t* keybuf, int B, const int* active, RowCount n, uint32_t* perm_a, uint32_t* perm_b,
                             int* which, cudaStream_t s);
void launch_radix_sort(const uint8_t* keybuf, int B, const int* passes, int npasses, RowCount n, uint32_t* perm_a,
                       uint32_t* perm_b, int* which, uint32_t* scratch, cudaStream_t s, int* launches);
// Sort without byte passes (see ops.cu): compacted keys (W = words per row in ckeys, w_out[0] = words in use),
// the top-k prefilter, and the rank sort that writes out_rows[0 .. min(limit, n)) and out_count[0] directly.
// rank_tmp: rows_bound uint32 zeros, tile_done: ceil(rows_bound / 128) int zeros (used when the range is sliced).
constexpr int kMaxCompactBytes = 512;      // encoded key bytes per row the compact route accepts
constexpr int kRankSortMax = 8192;         // rows (or top-k candidates) the rank sort is used for
void launch_compact_keys(const uint8_t* keybuf, int B, const int* active, RowCount n, uint64_t* ckeys, int W, int* w_out,
                         uint32_t* zero_rank, int* zero_tiles, int64_t zero_n, cudaStream_t s);   // also zeroes rank scratch
constexpr int kSelectPasses = 4;           // top-k prefilter: leading bytes of word 0 examined; hist = kSelectPasses * 256 zeros
void launch_topk_select(const uint64_t* ckeys, int W, RowCount n, int64_t k, uint32_t* cand, int cap, uint32_t* hist, int* nc_out,
                        cudaStream_t s, int* launches);
int rank_sort_slices(int64_t rows_bound);
void launch_rank_sort(const uint64_t* ckeys, int W, const int* w_dev, const uint32_t* cand, const int* nc_dev, int cand_cap,
                      RowCount n, int64_t rows_bound, int64_t limit, uint32_t* rank_tmp, int* tile_done, uint32_t* out_rows,
                      int* out_count, int* overflow, cudaStream_t s);   // *overflow = 1 when rows/candidates exceed rows_bound
void launch_take_perm(const uint32_t* perm_a, const uint32_t* perm_b, const int* which, int widx, RowCount n, int64_t limit,
                      uint32_t* out_rows, int* out_count, cudaStream_t s);
// Exchange (exec.cu, Exec::exchange): a batch is packed into one segment per rank, all-gathered with NCCL, and
// unpacked into concatenated columns.  kind: 0 int32, 1 int64/double, 2 string (values = int32 lengths, bytes in
// fixed maxlen slots at off_aux).  Error 14: a string longer than the column's declared maximum.
struct XCol {
  const void* data;
  const uint32_t* valid;
  const uint8_t* strbase;
  int64_t off_valid, off_data, off_aux;
  int32_t kind, maxlen;
};
void launch_set_count(RowCount n, int32_t* out, cudaStream_t s, const int* flag_in = nullptr, int32_t* flag_out = nullptr);
void launch_exchange_pack(const XCol* cols, int ncols, int64_t n, uint8_t* out, int* err, cudaStream_t s);
void launch_exchange_unpack(const XCol* srcs, const XCol* dst, int ncols, const int64_t* row_begin, int world, int64_t max_rows,
                            const uint8_t* recv, cudaStream_t s);
// Hash-partition exchange: destination rank of every row (dest[i] = hash(keys) % world; counts[0 .. world) rows per
// destination, counts[world] = rows), the rows ordered by destination, and one packed segment per destination.
void launch_exchange_dest(const Prog* progs, int nkeys, OpCtx c, RowCount n, int world, uint8_t* dest, int32_t* counts, cudaStream_t s);
void launch_exchange_perm(const uint8_t* dest, int64_t n, const int64_t* begin, int32_t* cursor, uint32_t* perm, cudaStream_t s);
void launch_exchange_pack_perm(const XCol* cols, const XCol* segs, int ncols, const uint32_t* perm, const int64_t* begin, int world,
                               int64_t n, int* err, cudaStream_t s);
// The same with the row counts left on the device (one collective, no host round trip): fixed-capacity segments with a
// kXchgHeader-byte header (row count, string slot widths); more rows than cap_rows raise err[2] (speculation failed).
constexpr int kXchgHeader = 64;
void launch_exchange_pack_dev(const XCol* cols, int ncols, RowCount n, int64_t cap_rows, uint8_t* out, int* err, cudaStream_t s);
void launch_exchange_unpack_dev(const XCol* proto, const XCol* dst, int ncols, int world, int64_t cap_rows, int64_t seg_bytes,
                                const uint8_t* recv, int* total_out, int* err, cudaStream_t s);
// Peer-push exchange (all-gather without NCCL and without a host round trip): every rank owns a window in its HBM that
// its peers map through CUDA IPC; a rank's push kernel packs its rows straight into slot `me` of every peer's window
// over NVLink and then raises that peer's arrival counter; the unpack kernel of each rank waits - bounded - until all
// `world` pushes of this use of the window half have arrived and copies the rows out.  The window has two halves used
// alternately, each with `world` slots of slot_bytes (a kXchgHeader header, then the columns laid out for cap_rows rows).
struct PeerWindow {
  uint8_t* win[16];            // base of every rank's window as mapped into this process (win[me] = the local one)
  int32_t world, me;
  int64_t half_bytes, slot_bytes, ctrl_bytes;
};
constexpr int kPeerMaxWorld = 16;
constexpr int64_t kPeerCtrlBytes = 256;      // per window: arrival counter of half 0 at +0, of half 1 at +64 (uint32, monotonic)
// header of a slot: [0] rows (-1: the sender had more rows than a slot holds), [1 + c] string slot width of column c,
// [15] the sender's "a size assumption failed" flag
void launch_peer_push(const XCol* cols, const XCol* proto, int ncols, RowCount n, int64_t cap_rows, PeerWindow w, int half, int loopback,
                      int* ticket, int* err, cudaStream_t s);
void launch_peer_unpack(const XCol* proto, const XCol* dst, int ncols, PeerWindow w, int half, uint32_t arrivals_wanted, int64_t cap_rows,
                        int* total_out, int* err, cudaStream_t s);
// Small-result export: segment kinds 0 = raw bytes (multiple of 4), 1 = string column -> int32 offsets at dst_off
// and characters at aux_off (cap bytes available; error 13 if exceeded), 2 = the row count as one int.
struct ExportSeg {
  const void* src;
  const uint32_t* valid;
  const uint8_t* strbase;
  int64_t dst_off, bytes, aux_off, cap;
  int32_t kind, pad;
};
void launch_export_pack(const ExportSeg* segs, int nsegs, RowCount n, uint8_t* out, int* err, cudaStream_t s);
void launch_copy_str_bytes(const StrView* views, const uint8_t* base, const int64_t* offsets, int64_t n, uint8_t* out,
                           cudaStream_t s);

}  // namespace nds
