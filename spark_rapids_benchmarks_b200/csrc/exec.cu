// Plan executor: walks the JSON physical-plan IR (scan / filter / project / join / aggregate / sort)
// and drives the kernels in decode.cu and ops.cu.  Replaces the engine behind the reference's
// `spark_session.sql(query)` + `collect()` / `write...save()` hand-off
// (/root/reference/nds/nds_power.py:125-135).  An unknown node or expression raises; there is no CPU path.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <exception>
#include <memory>
#include <unordered_set>
#include "engine.h"
#include "nccl_dyn.h"

namespace nds {

namespace {

// Per-query device memory: a bump allocator over a slab the engine owns (stream-ordered reuse: every query runs on
// the engine's one stream), with cudaMallocAsync only for what does not fit.  A query makes ~100 small allocations;
// as driver calls they cost more than the kernels they feed.
struct Arena {
  cudaStream_t s;
  std::vector<void*> ptrs;      // overflow allocations (freed at the end of the query)
  uint8_t* slab = nullptr;
  int64_t slab_cap = 0, slab_off = 0;
  int64_t zero_off = 0, zero_cap = 0;   // pre-zeroed prefix of the slab for counters and flags
  int64_t overflow_bytes = 0;           // what did not fit the slab this query (the engine grows the slab for the next one)
  uint8_t* pin = nullptr;       // pinned staging ring owned by the engine (reset per query)
  int64_t pin_cap = 0, pin_off = 0;
  explicit Arena(cudaStream_t st) : s(st) {}
  // The first zero_cap bytes of the slab are kept zero between queries: a query hands out pieces of it (zeros())
  // instead of issuing a memset per bitmap / counter / hash table, and the next query re-zeroes only the part the
  // previous one used (dirty; negative = unknown, clear everything).
  void attach_slab(uint8_t* p, int64_t cap, int64_t dirty) {
    slab = p;
    slab_cap = cap;
    zero_cap = p ? std::min<int64_t>(cap / 4, int64_t(32) << 20) : 0;
    zero_off = 0;
    slab_off = zero_cap;
    const int64_t clear = dirty < 0 ? zero_cap : std::min(dirty, zero_cap);
    if (clear > 0) NDS_CUDA(cudaMemsetAsync(slab, 0, (size_t)clear, s));
  }
  void* alloc(int64_t bytes) {
    if (bytes < 16) bytes = 16;
    const int64_t need = align_up(bytes + 64, 256);
    if (slab && slab_off + need <= slab_cap) {
      void* p = slab + slab_off;
      slab_off += need;
      return p;
    }
    void* p = nullptr;
    NDS_CUDA(cudaMallocAsync(&p, (size_t)bytes + 64, s));
    ptrs.push_back(p);
    overflow_bytes += need;
    return p;
  }
  template <typename T>
  T* alloc_n(int64_t n) { return reinterpret_cast<T*>(alloc(n * (int64_t)sizeof(T))); }
  template <typename T>
  T* zeros(int64_t n) {
    const int64_t bytes = std::max<int64_t>(n, 1) * (int64_t)sizeof(T);
    const int64_t need = align_up(bytes, 16);
    if (need <= (int64_t(4) << 20) && zero_off + need <= zero_cap) {
      T* p = reinterpret_cast<T*>(slab + zero_off);
      zero_off += need;
      return p;
    }
    T* p = alloc_n<T>(n);
    NDS_CUDA(cudaMemsetAsync(p, 0, (size_t)(bytes < 16 ? 16 : bytes), s));
    return p;
  }
  // host -> device copy that does not need a stream sync when the bytes fit the pinned ring
  void upload(void* dst, const void* src, int64_t bytes) {
    if (bytes <= 0) return;
    int64_t need = align_up(bytes, 64);
    if (pin && pin_off + need <= pin_cap) {
      memcpy(pin + pin_off, src, (size_t)bytes);
      NDS_CUDA(cudaMemcpyAsync(dst, pin + pin_off, (size_t)bytes, cudaMemcpyHostToDevice, s));
      pin_off += need;
    } else {
      NDS_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyHostToDevice, s));
      NDS_CUDA(cudaStreamSynchronize(s));
    }
  }
  // pinned bytes for a device -> host read-back (nullptr when the ring is full: the caller falls back to pageable memory)
  uint8_t* pin_take(int64_t bytes) {
    int64_t need = align_up(std::max<int64_t>(bytes, 1), 64);
    if (!pin || pin_off + need > pin_cap) return nullptr;
    uint8_t* p = pin + pin_off;
    pin_off += need;
    return p;
  }
  ~Arena() {
    for (void* p : ptrs) cudaFreeAsync(p, s);
  }
};

int64_t pow10i(int k) {
  int64_t v = 1;
  for (int i = 0; i < k; ++i) v *= 10;
  return v;
}

uint32_t next_pow2(int64_t v) {
  if (v > (int64_t(1) << 31)) fail("hash table for " + std::to_string(v / 2) + " rows exceeds 2^31 slots; shard the build side across GPUs");
  uint32_t c = 64;
  while ((int64_t)c < v) c <<= 1;
  return c;
}

// Deepest evaluation stack a postfix program reaches (expr.cuh: eval_program keeps kEvalStack values in registers /
// local memory): leaves push, binary operators pop one, if() pops two.
int stack_depth(const std::vector<Instr>& code) {
  int d = 0, mx = 0;
  for (const Instr& in : code) {
    switch (in.op) {
      case OP_COL: case OP_LIT_I: case OP_LIT_F: case OP_LIT_S: case OP_LIT_NULL: ++d; break;
      case OP_NOT: case OP_ISNULL: case OP_ISNOTNULL: case OP_NEG_I: case OP_NEG_F: case OP_I2F: case OP_SUBSTR: break;
      case OP_IF: d -= 2; break;
      default: d -= 1; break;          // every other operator is binary
    }
    mx = std::max(mx, d);
  }
  return mx;
}

LType parse_type(const std::string& s) {
  LType t;
  if (s == "int" || s == "int64" || s == "long" || s == "bigint") { t.kind = LType::INT; return t; }
  if (s == "int32") { t.kind = LType::INT; t.is32 = true; return t; }
  if (s == "double") { t.kind = LType::F64; return t; }
  if (s == "string") { t.kind = LType::STR; return t; }
  if (s == "bool") { t.kind = LType::BOOL; return t; }
  if (s == "date") { t.kind = LType::DATE; t.is32 = true; return t; }
  int p, sc;
  if (sscanf(s.c_str(), "decimal(%d,%d)", &p, &sc) == 2) { t.kind = LType::DECIMAL; t.precision = p; t.scale = sc; return t; }
  fail("plan: unknown type '" + s + "'");
}

// Collects the column slots, literal pool and programs of one operator and uploads them in one copy.
struct Blob {
  std::vector<ColRef> cols;
  std::vector<int> col_src;
  std::string lit;
  std::vector<std::vector<Instr>> progs;
  std::vector<int> kinds;

  int slot(const Batch& b, int ci) {
    for (size_t i = 0; i < col_src.size(); ++i)
      if (col_src[i] == ci) return (int)i;
    col_src.push_back(ci);
    cols.push_back(b.cols[ci].ref());
    return (int)cols.size() - 1;
  }
  int add_prog(std::vector<Instr> code, const LType& t) {
    const int depth = stack_depth(code);
    if (depth > kEvalStack)
      fail("plan: expression too deep (needs " + std::to_string(depth) + " stack entries, the evaluator has " + std::to_string(kEvalStack) +
           "): nest fewer CASE branches / operands");
    progs.push_back(std::move(code));
    kinds.push_back(t.kind == LType::STR ? 2 : (t.kind == LType::F64 ? 1 : 0));
    return (int)progs.size() - 1;
  }
  struct Dev {
    OpCtx ctx;
    std::vector<Prog> host_progs;
    const Prog* progs = nullptr;
    const uint8_t* extra = nullptr;   // caller-provided POD block (AggSpec / SortKey arrays)
  };
  Dev upload(Arena& ar, cudaStream_t s, const void* extra = nullptr, int64_t extra_bytes = 0) {
    int64_t off_cols = 0;
    int64_t off_lit = align_up(off_cols + (int64_t)cols.size() * sizeof(ColRef), 16);
    int64_t off_code = align_up(off_lit + (int64_t)lit.size() + 1, 16);
    int64_t ninstr = 0;
    for (auto& p : progs) ninstr += (int64_t)p.size();
    int64_t off_progs = align_up(off_code + ninstr * (int64_t)sizeof(Instr), 16);
    int64_t off_extra = align_up(off_progs + (int64_t)progs.size() * sizeof(Prog), 16);
    int64_t total = off_extra + extra_bytes + 16;
    std::vector<uint8_t> h(total, 0);
    uint8_t* d = (uint8_t*)ar.alloc(total);
    if (!cols.empty()) memcpy(h.data() + off_cols, cols.data(), cols.size() * sizeof(ColRef));
    if (!lit.empty()) memcpy(h.data() + off_lit, lit.data(), lit.size());
    Dev dev;
    int64_t ci = 0;
    for (size_t k = 0; k < progs.size(); ++k) {
      if (!progs[k].empty()) memcpy(h.data() + off_code + ci * sizeof(Instr), progs[k].data(), progs[k].size() * sizeof(Instr));
      Prog p;
      p.code = reinterpret_cast<const Instr*>(d + off_code + ci * sizeof(Instr));
      p.n = (int32_t)progs[k].size();
      p.kind = kinds[k];
      dev.host_progs.push_back(p);
      ci += (int64_t)progs[k].size();
    }
    if (!progs.empty()) memcpy(h.data() + off_progs, dev.host_progs.data(), progs.size() * sizeof(Prog));
    if (extra_bytes) memcpy(h.data() + off_extra, extra, extra_bytes);
    ar.upload(d, h.data(), total);
    dev.ctx.cols = reinterpret_cast<const ColRef*>(d + off_cols);
    dev.ctx.lit = d + off_lit;
    dev.progs = reinterpret_cast<const Prog*>(d + off_progs);
    dev.extra = d + off_extra;
    return dev;
  }
};

struct RuntimeFilter {           // dynamic partition pruning: build-side join keys
  std::string column;
  // The keys travel to the host asynchronously (pinned memory + an event); the scan that uses the filter waits for
  // the event, so build sides issued in between keep the GPU busy instead of stalling behind a stream sync.
  struct State {
    bool ready = false;
    cudaEvent_t ev = nullptr;
    const uint8_t* hk = nullptr;        // keys of the upper-bound row range
    const uint32_t* hv = nullptr;       // their validity words
    const int* hcount = nullptr;        // exact row count
    int64_t nb = 0;
    bool is32 = false;
    std::vector<uint8_t> pageable;      // backing store when the pinned ring is full
    const uint32_t* hbits = nullptr;    // device-built bitmap over [lo, hi] (32-bit words) when the partition range is known
    int64_t hbits_words = 0;
    // a bitmap over [lo, hi] when the key range is small (surrogate keys are dense), else a sorted list
    int64_t lo = 0, hi = -1;
    std::vector<uint64_t> bits;
    std::vector<int64_t> sorted;
  };
  std::shared_ptr<State> st;
  void assign(std::vector<int64_t>& k) {
    State& t = *st;
    if (k.empty()) return;
    t.lo = t.hi = k[0];
    for (int64_t v : k) { t.lo = std::min(t.lo, v); t.hi = std::max(t.hi, v); }
    if (t.hi - t.lo < (int64_t(1) << 26)) {
      t.bits.assign((size_t)((t.hi - t.lo) / 64 + 1), 0);
      for (int64_t v : k) t.bits[(size_t)((v - t.lo) >> 6)] |= uint64_t(1) << ((v - t.lo) & 63);
    } else {
      std::sort(k.begin(), k.end());
      t.sorted.swap(k);
    }
  }
  bool contains(int64_t v) const {
    const State& t = *st;
    if (v < t.lo || v > t.hi) return false;
    if (!t.bits.empty()) return (t.bits[(size_t)((v - t.lo) >> 6)] >> ((v - t.lo) & 63)) & 1;
    return std::binary_search(t.sorted.begin(), t.sorted.end(), v);
  }
};

}  // namespace

struct SpeculationFailed : std::exception {     // a device-side size assumption did not hold; Engine::execute re-runs the plan without them
  const char* what() const noexcept override { return "device-side size speculation failed twice"; }
};

struct Exec {
  Engine& e;
  Arena ar;
  cudaStream_t s;
  int* err = nullptr;            // err[0]: kernel error code, err[1]: join flags
  bool source_host = false;
  bool profile = false;
  bool decode_timing = false;
  bool split_decode = true;      // options.split_decode=false forces every page through the general kernel
  bool side_stream = true;       // options.side_stream=false keeps every operator on one stream
  int64_t small_scan_pages = 64; // options.small_scan_pages: scans with fewer split-route pages use the general kernel only
  bool speculate = true;         // options.speculate=false: never size a launch on an unverified device-side count
  bool dense_tables = true;      // options.dense_tables=false: every join table is an open-addressing hash table
  // options.exchange_single=true: the exchange is one collective of fixed-capacity segments (options.exchange_rows rows,
  // 128 k by default) with the row counts in their headers and no host round trip; by default the counts are all-gathered
  // and read back first, so that the segments carry exactly the rows there are
  bool exchange_single = false;
  // options.exchange_peer=false: never the peer-push exchange (rows packed straight into the peers' windows over NVLink,
  // row counts in the slot headers, no NCCL call and no host round trip) - the default whenever the windows are mapped
  bool exchange_peer = true;
  bool topk_stream = true;       // options.topk_stream=false: ORDER BY ... LIMIT takes the multi-launch select / rank-sort route
  // options.probe_bloom=true: large fused scans use the persistent probe kernel with the first join's keys as an existence
  // filter in shared memory (scan_probe_persistent_kernel).  OFF by default: measured slower than the plain kernel (SF10
  // query7: 0.49 ms against 0.21 ms) - one 1024-thread CTA per SM hides the load latency worse than six 256-thread CTAs,
  // and the saved join-table reads do not make up for it.  Kept switchable for the next attempt (several chunks in flight
  // per CTA); options.probe_bloom_chunks = 2048-row chunks from which it applies.
  bool probe_bloom = false;
  int64_t probe_bloom_chunks = 4 * 148;
  // options.dict_match=false: the fused scan tests the first join per row (dictionary entry, then join table) instead of
  // once per dictionary entry and file (dict_match_kernel); options.dict_match_chunks = 2048-row chunks from which it applies
  bool dict_match = true;
  int64_t dict_match_chunks = 4096;
  bool build_filter = true;      // options.build_filter=false: a filtered build side is filtered first, then built (two or three launches)
  bool fused_scan = true;        // options.fused_scan=false: fact scans decode every column, then probe (decode_fast.cu + ops.cu)
  bool trace = false;            // options.trace: host timeline marks + time spent waiting in each stream sync
  double t_base = 0;
  std::string trace_out;
  std::vector<OpStat> stats;
  int64_t launches = 0;
  int64_t h2d = 0, d2h = 0;

  const Json* opts = nullptr;
  Exec(Engine& en, const Json* options) : e(en), ar(en.stream_), s(en.stream_), opts(options) {
    ar.pin = en.pin_;
    ar.pin_cap = en.pin_cap_;
    ar.attach_slab(en.slab_, en.slab_cap_, en.zero_dirty_);
    en.zero_dirty_ = -1;          // until this query ends in an orderly way
    err = ar.zeros<int>(4);
    if (options) {
      source_host = options->str_or("source", "hbm") == "host";
      profile = options->bool_or("profile", false);
      decode_timing = options->bool_or("decode_timing", false);
      split_decode = options->bool_or("split_decode", true);
      trace = options->bool_or("trace", false);
      side_stream = options->bool_or("side_stream", true);
      speculate = options->bool_or("speculate", true);
      small_scan_pages = options->int_or("small_scan_pages", 64);
      dense_tables = options->bool_or("dense_tables", true);
      fused_scan = options->bool_or("fused_scan", true);
      probe_bloom = options->bool_or("probe_bloom", false);
      dict_match = options->bool_or("dict_match", true);
      dict_match_chunks = options->int_or("dict_match_chunks", 4096);
      build_filter = options->bool_or("build_filter", true);
      probe_bloom_chunks = options->int_or("probe_bloom_chunks", 4 * 148);
      topk_stream = options->bool_or("topk_stream", true);
      exchange_single = options->bool_or("exchange_single", false);
      exchange_peer = options->bool_or("exchange_peer", true);
    }
    t_base = now_ms();
    if (!e.side_streams_.empty() && side_stream) {
      ev_start = e.get_event();
      NDS_CUDA(cudaEventRecord(ev_start, s));
    }
  }

  void mark(const char* label, double wait_ms = -1) {
    if (!trace) return;
    char buf[160];
    snprintf(buf, sizeof buf, "%s[\"%s\",%.4f,%.4f]", trace_out.empty() ? "" : ",", label, now_ms() - t_base, wait_ms);
    trace_out += buf;
  }
  void sync(const char* label) {
    if (!trace) { NDS_CUDA(cudaStreamSynchronize(s)); return; }
    mark(label);
    const double t0 = now_ms();
    NDS_CUDA(cudaStreamSynchronize(s));
    mark("sync_done", now_ms() - t0);
  }

  ~Exec() {     // a failed query still returns its timing events to the pool
    e.zero_dirty_ = ar.zero_off;
    if (ar.overflow_bytes > 0) e.slab_wanted_ = std::max(e.slab_wanted_, ar.slab_off + ar.overflow_bytes + ar.overflow_bytes / 4);
    if (s_saved) { cudaStreamSynchronize(s); s = s_saved; ar.s = s; }     // failed inside a side-stream section
    if (std::uncaught_exceptions() > 0) {
      // a failed query may leave work queued on the side streams that still writes into the slab: drain them before
      // the next query reuses that memory
      for (cudaStream_t st : e.side_streams_) cudaStreamSynchronize(st);
      cudaStreamSynchronize(e.stream_);
      cudaGetLastError();
    }
    if (ev_start) e.put_event(ev_start);
    for (cudaEvent_t ev : side_events) e.put_event(ev);
    for (auto& st : rf_states)
      if (st->ev) { e.put_event(st->ev); st->ev = nullptr; }     // a filter no scan asked for
    for (OpStat& st : stats) {
      if (st.e0) e.put_event(st.e0);
      if (st.e1) e.put_event(st.e1);
    }
  }

  // ---------------------------------------------------------------- op timing
  static double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
  }
  size_t begin_stat(const std::string& op, bool always = false) {
    OpStat st;
    st.op = op;
    st.t0 = now_ms();
    if (trace) mark(op.c_str());
    if (profile || always) {
      st.e0 = e.get_event();
      st.e1 = e.get_event();
      NDS_CUDA(cudaEventRecord(st.e0, s));
    }
    stats.push_back(st);
    return stats.size() - 1;
  }
  void end_stat(size_t i, int64_t bytes, int64_t rows) {
    stats[i].bytes = bytes;
    stats[i].rows = rows;
    stats[i].host_ms = now_ms() - stats[i].t0;
    if (trace) mark("end");
    if (stats[i].e1) NDS_CUDA(cudaEventRecord(stats[i].e1, s));
  }

  // Algorithmic bytes of a row-at-a-time operator = rows x (bytes of the column slots its programs read: values at
  // their width, 4 B per row-map hop, 1 bit of validity) + what it writes per row.  The row count may only exist on
  // the device; in profile mode it is read back after the query (resolve_stat_rows) and the product is formed then.
  static int64_t row_bytes_of(const Blob& bb) {
    int64_t b = 0;
    for (const ColRef& c : bb.cols) b += (c.dtype == DT_I32 ? 4 : 8) + (c.rowmap ? 4 : 0);
    return b;
  }
  void end_stat_rows(size_t i, int64_t row_bytes, RowCount rows, int64_t fixed_bytes = 0) {
    stats[i].row_bytes = row_bytes;
    stats[i].n_dev = rows.dev;
    end_stat(i, fixed_bytes + rows.host * row_bytes, rows.host);
    stats[i].fixed_bytes = fixed_bytes;
  }
  void resolve_stat_rows() {
    for (OpStat& st : stats) {
      if (!st.n_dev || st.row_bytes <= 0) continue;
      int v = 0;
      if (cudaMemcpy(&v, st.n_dev, 4, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); continue; }
      const int64_t n = std::min<int64_t>(v, st.rows);
      st.rows = n;
      st.bytes = st.fixed_bytes + n * st.row_bytes;
    }
  }

  void check_err() {
    int h[4];
    NDS_CUDA(cudaMemcpyAsync(h, err, sizeof h, cudaMemcpyDeviceToHost, s));
    sync("sync:check_err");
    check_err_values(h);
  }
  void check_err_values(const int* h) {
    if (h[0]) fail("device kernel reported error code " + std::to_string(h[0]) +
                   " (3-6: malformed parquet page, 11: string projection outside its column buffer, 12: sort key wider than its declared maximum, 13: result string "
                   "longer than its column's known maximum, 14: exchanged string longer than its declared maximum, 15: ranks disagree on string widths, "
                   "16: a peer's rows did not arrive within the exchange's time limit - re-create the communicator)");
    if (h[2]) throw SpeculationFailed();
    if (h[1] & 1) fail("hash join: duplicate keys on the build side (only unique build keys are supported)");
    if (h[1] & 2) fail("hash join: key equals the reserved empty marker");
    if (h[1] & 4) fail("hash join: a build key lies outside the range its column's Parquet statistics state (corrupt statistics)");
  }

  // ---- dynamic partition pruning: wait for the build keys (usually long since on the host) and index them
  void resolve(const RuntimeFilter& rf) {
    RuntimeFilter::State& t = *rf.st;
    if (t.ready) return;
    if (t.ev) {
      mark("wait:prune_keys");
      const double t0 = now_ms();
      NDS_CUDA(cudaEventSynchronize(t.ev));
      mark("wait_done", now_ms() - t0);
      e.put_event(t.ev);
      t.ev = nullptr;
    }
    if (t.hbits) {                     // the bitmap was built on the device: repack 32-bit words into the 64-bit ones
      t.bits.assign((size_t)((t.hbits_words + 1) / 2), 0);
      for (int64_t w = 0; w < t.hbits_words; ++w) t.bits[(size_t)(w >> 1)] |= (uint64_t)t.hbits[w] << ((w & 1) * 32);
      t.ready = true;
      return;
    }
    int64_t hcount = std::min<int64_t>(*t.hcount, t.nb);
    std::vector<int64_t> keys;
    keys.reserve((size_t)std::max<int64_t>(hcount, 0));
    for (int64_t i = 0; i < hcount; ++i)
      if ((t.hv[i >> 5] >> (i & 31)) & 1u)
        keys.push_back(t.is32 ? (int64_t) reinterpret_cast<const int32_t*>(t.hk)[i] : reinterpret_cast<const int64_t*>(t.hk)[i]);
    const_cast<RuntimeFilter&>(rf).assign(keys);
    t.ready = true;
  }

  // ---- second stream: the probe-side scan of a join chain runs beside the build sides
  cudaStream_t s_saved = nullptr;
  std::vector<cudaEvent_t> side_events;
  std::vector<std::shared_ptr<RuntimeFilter::State>> rf_states;
  void enter_side(int which) {
    if (e.side_streams_.empty() || s_saved || !ev_start) return;
    // a subtree made of scans/filters/builds reads registered table bytes and its own allocations only: all it must
    // wait for is the arena's zeroed prefix (ev_start, recorded when the query began) - not what the main stream
    // queued since
    // which >= 0: one of the build-side lanes; kBaseLane: the fact scan's lane
    cudaStream_t side = which == kBaseLane ? e.side_streams_[kBaseLane] : e.side_streams_[(size_t)which % kDimLanes];
    NDS_CUDA(cudaStreamWaitEvent(side, ev_start, 0));
    s_saved = s;
    s = side;
    ar.s = s;
  }
  void leave_side() {
    if (!s_saved) return;
    cudaEvent_t ev = e.get_event();
    NDS_CUDA(cudaEventRecord(ev, s));
    side_events.push_back(ev);
    s = s_saved;
    ar.s = s;
    s_saved = nullptr;
    NDS_CUDA(cudaStreamWaitEvent(s, ev, 0));
  }
  cudaEvent_t ev_start = nullptr;

  int64_t count(Batch& b) {
    if (b.n_exact >= 0) return b.n_exact;
    if (!b.n_dev) { b.n_exact = b.n_max; return b.n_exact; }
    int v = 0;
    NDS_CUDA(cudaMemcpyAsync(&v, b.n_dev, 4, cudaMemcpyDeviceToHost, s));
    sync("sync:count");
    d2h += 4;
    b.n_exact = std::min<int64_t>(v, b.n_max);
    return b.n_exact;
  }

  // ---------------------------------------------------------------- expressions
  static bool is_intlike(const LType& t) { return t.kind == LType::INT || t.kind == LType::DATE || t.kind == LType::DECIMAL || t.kind == LType::BOOL; }

  void to_double(std::vector<Instr>& code, const LType& t) {
    if (t.kind == LType::F64) return;
    code.push_back(Instr{OP_I2F, 0, 0});
    if (t.kind == LType::DECIMAL && t.scale > 0) {
      double d = std::pow(10.0, t.scale);
      int64_t bits;
      memcpy(&bits, &d, 8);
      code.push_back(Instr{OP_LIT_F, 0, bits});
      code.push_back(Instr{OP_DIV_F, 0, 0});
    }
  }
  void scale_up(std::vector<Instr>& code, int k) {
    if (k <= 0) return;
    code.push_back(Instr{OP_LIT_I, 0, pow10i(k)});
    code.push_back(Instr{OP_MUL_I, 0, 0});
  }

  // str_origin: batch column whose byte buffer a string result points into (-1: none / literal, -2: mixed)
  LType compile(const Json& x, const Batch& b, Blob& bb, std::vector<Instr>& code, int* str_origin = nullptr) {
    if (str_origin) *str_origin = -1;
    if (x.kind != Json::Obj) fail("plan: expression must be an object");
    if (x.has("col")) {
      const std::string& name = x.str("col");
      int ci = b.find(name);
      if (ci < 0) fail("plan: unknown column '" + name + "'");
      code.push_back(Instr{OP_COL, bb.slot(b, ci), 0});
      if (str_origin && b.cols[ci].lt.kind == LType::STR) *str_origin = ci;
      return b.cols[ci].lt;
    }
    if (x.has("lit")) {
      const Json& v = x.at("lit");
      if (v.kind == Json::Null) { code.push_back(Instr{OP_LIT_NULL, 0, 0}); return x.has("type") ? parse_type(x.str("type")) : LType(); }
      LType t = parse_type(x.str_or("type", v.kind == Json::Str ? "string" : (v.is_int ? "int" : "double")));
      if (t.kind == LType::STR) {
        if (v.kind != Json::Str) fail("plan: string literal expected");
        int64_t off = (int64_t)bb.lit.size();
        bb.lit += v.s;
        t.maxlen = std::max<int>(1, (int)v.s.size());
        code.push_back(Instr{OP_LIT_S, 0, (int64_t)(((uint64_t)v.s.size() << 32) | (uint64_t)off)});
      } else if (t.kind == LType::F64) {
        double d = v.d;
        int64_t bits;
        memcpy(&bits, &d, 8);
        code.push_back(Instr{OP_LIT_F, 0, bits});
      } else {
        if (v.kind != Json::Num || !v.is_int) fail("plan: integer literal expected (decimals are passed unscaled)");
        code.push_back(Instr{OP_LIT_I, 0, v.i});
      }
      return t;
    }
    const std::string& fn = x.str("fn");
    const Json& args = x.at("args");
    auto arg = [&](size_t i) -> const Json& {
      if (i >= args.a.size()) fail("plan: function '" + fn + "' is missing argument " + std::to_string(i));
      return args.a[i];
    };
    LType out;
    static const std::map<std::string, int> cmp = {{"eq", 0}, {"ne", 1}, {"lt", 2}, {"le", 3}, {"gt", 4}, {"ge", 5}};
    auto cit = cmp.find(fn);
    if (cit != cmp.end()) {
      std::vector<Instr> ca, cb;
      LType ta = compile(arg(0), b, bb, ca), tb = compile(arg(1), b, bb, cb);
      int base;
      if (ta.kind == LType::STR || tb.kind == LType::STR) {
        if (ta.kind != tb.kind) fail("plan: comparison between string and non-string");
        base = OP_EQ_S;
      } else if (ta.kind == LType::F64 || tb.kind == LType::F64) {
        to_double(ca, ta);
        to_double(cb, tb);
        base = OP_EQ_F;
      } else {
        int sa = ta.kind == LType::DECIMAL ? ta.scale : 0, sb = tb.kind == LType::DECIMAL ? tb.scale : 0;
        scale_up(ca, sb - sa);
        scale_up(cb, sa - sb);
        base = OP_EQ_I;
      }
      code.insert(code.end(), ca.begin(), ca.end());
      code.insert(code.end(), cb.begin(), cb.end());
      code.push_back(Instr{base + cit->second, 0, 0});
      out.kind = LType::BOOL;
    } else if (fn == "and" || fn == "or") {
      if (args.a.empty()) fail("plan: empty " + fn);
      for (size_t i = 0; i < args.a.size(); ++i) {
        LType t = compile(args.a[i], b, bb, code);
        if (t.kind != LType::BOOL) fail("plan: " + fn + " over a non-boolean");
        if (i) code.push_back(Instr{fn == "and" ? OP_AND : OP_OR, 0, 0});
      }
      out.kind = LType::BOOL;
    } else if (fn == "not") {
      compile(arg(0), b, bb, code);
      code.push_back(Instr{OP_NOT, 0, 0});
      out.kind = LType::BOOL;
    } else if (fn == "isnull" || fn == "isnotnull") {
      compile(arg(0), b, bb, code);
      code.push_back(Instr{fn == "isnull" ? OP_ISNULL : OP_ISNOTNULL, 0, 0});
      out.kind = LType::BOOL;
    } else if (fn == "add" || fn == "sub" || fn == "mul") {
      std::vector<Instr> ca, cb;
      LType ta = compile(arg(0), b, bb, ca), tb = compile(arg(1), b, bb, cb);
      if (ta.kind == LType::STR || tb.kind == LType::STR) fail("plan: arithmetic on strings");
      if (ta.kind == LType::F64 || tb.kind == LType::F64) {
        to_double(ca, ta);
        to_double(cb, tb);
        code.insert(code.end(), ca.begin(), ca.end());
        code.insert(code.end(), cb.begin(), cb.end());
        code.push_back(Instr{fn == "add" ? OP_ADD_F : (fn == "sub" ? OP_SUB_F : OP_MUL_F), 0, 0});
        out.kind = LType::F64;
      } else {
        int sa = ta.kind == LType::DECIMAL ? ta.scale : 0, sb = tb.kind == LType::DECIMAL ? tb.scale : 0;
        bool dec = ta.kind == LType::DECIMAL || tb.kind == LType::DECIMAL;
        int pa = ta.kind == LType::DECIMAL ? ta.precision : (ta.is32 ? 10 : 20), pb = tb.kind == LType::DECIMAL ? tb.precision : (tb.is32 ? 10 : 20);
        if (fn != "mul") { scale_up(ca, sb - sa); scale_up(cb, sa - sb); }
        code.insert(code.end(), ca.begin(), ca.end());
        code.insert(code.end(), cb.begin(), cb.end());
        code.push_back(Instr{fn == "add" ? OP_ADD_I : (fn == "sub" ? OP_SUB_I : OP_MUL_I), 0, 0});
        if (dec) {
          out.kind = LType::DECIMAL;
          if (fn == "mul") { out.scale = sa + sb; out.precision = std::min(38, pa + pb + 1); }
          else { out.scale = std::max(sa, sb); out.precision = std::min(38, std::max(pa - sa, pb - sb) + out.scale + 1); }
        } else out.kind = LType::INT;
      }
    } else if (fn == "neg") {
      LType t = compile(arg(0), b, bb, code);
      code.push_back(Instr{t.kind == LType::F64 ? OP_NEG_F : OP_NEG_I, 0, 0});
      out = t;
      out.is32 = false;
    } else if (fn == "substr") {
      int origin = -1;
      LType t = compile(arg(0), b, bb, code, &origin);
      if (t.kind != LType::STR) fail("plan: substr over a non-string");
      code.push_back(Instr{OP_SUBSTR, (int32_t)x.int_or("start", 1), x.int_or("len", 1 << 30)});
      if (str_origin) *str_origin = origin;
      out.kind = LType::STR;
      {
        const int64_t want = std::max<int64_t>(1, x.int_or("len", 1 << 30));
        out.maxlen = (int)std::min<int64_t>(want, t.maxlen > 0 ? t.maxlen : (want < (1 << 20) ? want : 0));
      }
    } else if (fn == "if") {
      LType tc = compile(arg(0), b, bb, code);
      if (tc.kind != LType::BOOL) fail("plan: if() condition must be boolean");
      int o1 = -1, o2 = -1;
      LType tt = compile(arg(1), b, bb, code, &o1), te = compile(arg(2), b, bb, code, &o2);
      if ((tt.kind == LType::STR) != (te.kind == LType::STR) || (tt.kind == LType::F64) != (te.kind == LType::F64))
        fail("plan: if() branches of different type families");
      code.push_back(Instr{OP_IF, 0, 0});
      if (str_origin) {
        // a string result points into ONE column's bytes: both branches from that column, or the other branch a NULL
        // (CASE WHEN c THEN col END); anything else (a literal beside a column, two columns) is mixed
        const bool n1 = arg(1).has("lit") && arg(1).at("lit").kind == Json::Null;
        const bool n2 = arg(2).has("lit") && arg(2).at("lit").kind == Json::Null;
        *str_origin = n1 ? o2 : (n2 ? o1 : (o1 == o2 ? o1 : -2));
      }
      out = tt;
      out.is32 = false;
      out.maxlen = (tt.maxlen > 0 && te.maxlen > 0) ? std::max(tt.maxlen, te.maxlen) : 0;
    } else if (fn == "decavg") {
      LType ts = compile(arg(0), b, bb, code);
      compile(arg(1), b, bb, code);
      int up = (int)x.int_or("up", 4);
      code.push_back(Instr{OP_DECAVG, up, 0});
      out.kind = LType::DECIMAL;
      out.scale = (ts.kind == LType::DECIMAL ? ts.scale : 0) + up;
      out.precision = std::min(38, (ts.kind == LType::DECIMAL ? ts.precision : 20) - 10 + up);
    } else if (fn == "divi2f") {
      compile(arg(0), b, bb, code);
      compile(arg(1), b, bb, code);
      code.push_back(Instr{OP_DIVI2F, 0, 0});
      out.kind = LType::F64;
    } else if (fn == "div") {     // Spark: `/` on integers and doubles yields double, x / 0 = NULL
      std::vector<Instr> ca, cb;
      LType ta = compile(arg(0), b, bb, ca), tb = compile(arg(1), b, bb, cb);
      if (ta.kind == LType::STR || tb.kind == LType::STR) fail("plan: division on strings");
      to_double(ca, ta);
      to_double(cb, tb);
      code.insert(code.end(), ca.begin(), ca.end());
      code.insert(code.end(), cb.begin(), cb.end());
      code.push_back(Instr{OP_DIV_F, 0, 0});
      out.kind = LType::F64;
    } else if (fn == "cast_double") {
      LType t = compile(arg(0), b, bb, code);
      if (t.kind == LType::STR) fail("plan: cast string -> double is not supported");
      to_double(code, t);
      out.kind = LType::F64;
    } else {
      fail("plan: unsupported expression function '" + fn + "'");
    }
    if (x.has("type")) {
      LType forced = parse_type(x.str("type"));
      if ((forced.kind == LType::STR) != (out.kind == LType::STR) || (forced.kind == LType::F64) != (out.kind == LType::F64))
        fail("plan: type annotation '" + x.str("type") + "' contradicts the expression");
      const int keep_len = out.maxlen;
      out = forced;
      out.is32 = false;
      out.maxlen = keep_len;
    }
    if ((int)code.size() > 200) fail("plan: expression too large");
    return out;
  }

  // rowmap composition with per-operator caching (columns that share a rowmap share the result)
  struct MapCache {
    std::map<const uint32_t*, const uint32_t*> m;
  };
  const uint32_t* remap(const uint32_t* old_map, const uint32_t* sel, RowCount n, MapCache& mc) {
    if (!old_map) return sel;
    auto it = mc.m.find(old_map);
    if (it != mc.m.end()) return it->second;
    uint32_t* out = ar.alloc_n<uint32_t>(n.host);
    launch_compose(old_map, sel, n, out, s);
    ++launches;
    mc.m[old_map] = out;
    return out;
  }

  // ---------------------------------------------------------------- scan
  struct FileSel {                 // files of a table that survive partition pruning, and where their rows start
    std::vector<int> sel;
    std::vector<int64_t> row_start;      // sel.size() + 1 entries
    int64_t N = 0;
    bool pruned = false;
  };
  FileSel select_files(Table& t, std::vector<RuntimeFilter>& filters) {
    const RuntimeFilter* rf = nullptr;
    for (const RuntimeFilter& f : filters)
      if (f.column == t.part_col && !t.part_col.empty()) rf = &f;
    if (rf) resolve(*rf);
    FileSel fs;
    fs.pruned = rf != nullptr;
    for (size_t f = 0; f < t.files.size(); ++f) {
      const FileInfo& fi = t.files[f];
      if (rf && (fi.part_null || !rf->contains(fi.part_val))) continue;
      fs.sel.push_back((int)f);
      fs.row_start.push_back(fs.N);
      fs.N += fi.rows;
    }
    fs.row_start.push_back(fs.N);
    return fs;
  }
  // Address of staged byte 0 of a column for a scan of the selected files: the resident copy in HBM, or - source=host - a
  // per-query device buffer into which the selected files' byte ranges are copied (adjacent ranges coalesced).
  const uint8_t* column_bytes(TableColumn& tc, const std::vector<int>& sel) {
    if (sel.empty()) return nullptr;
    if (!source_host) {
      if (!tc.d_enc) fail("engine was created with resident=false; use source=host");
      return tc.d_enc;
    }
    if (!tc.h_enc) fail("engine was created without keep_host; source=host is unavailable");
    const int64_t span_b = tc.file_byte_begin[sel.front()], span_e = tc.file_byte_begin[sel.back() + 1];
    uint8_t* tr = (uint8_t*)ar.alloc(span_e - span_b + 256);
    size_t k = 0;
    while (k < sel.size()) {
      size_t k2 = k;
      while (k2 + 1 < sel.size() && sel[k2 + 1] == sel[k2] + 1) ++k2;
      const int64_t b0 = tc.file_byte_begin[sel[k]], b1 = tc.file_byte_begin[sel[k2] + 1];
      NDS_CUDA(cudaMemcpyAsync(tr + (b0 - span_b), tc.h_enc + b0, b1 - b0, cudaMemcpyHostToDevice, s));
      h2d += b1 - b0;
      k = k2 + 1;
    }
    return tr - span_b;
  }

  Batch scan(const Json& node, std::vector<RuntimeFilter>& filters) {
    const std::string& tname = node.str("table");
    auto it = e.tables_.find(tname);
    if (it == e.tables_.end()) fail("plan: table '" + tname + "' is not registered");
    Table& t = it->second;
    const Json& jc = node.at("columns");
    Batch out;
    if (t.is_mem) {
      out.n_max = t.rows;
      for (const Json& c : jc.a) {
        auto ci = t.col_index.find(c.s);
        if (ci == t.col_index.end()) fail("plan: table '" + tname + "' has no column '" + c.s + "'");
        const MemColumn& mc = t.mem_cols[ci->second];
        Column col;
        col.name = mc.name; col.lt = mc.lt; col.data = mc.d_data; col.valid = mc.d_valid; col.strbase = mc.d_chars;
        out.cols.push_back(col);
      }
      return out;
    }
    // ---- partition pruning (Spark: dynamic partition pruning on the broadcast side's keys)
    FileSel fs = select_files(t, filters);
    std::vector<int>& sel = fs.sel;
    std::vector<int64_t>& row_start = fs.row_start;
    const int64_t N = fs.N;
    if (N >= (int64_t(1) << 31)) fail("scan of '" + tname + "' selects 2^31 or more rows; shard the table across GPUs");
    out.n_max = N;
    size_t st = begin_stat("decode:" + tname, true);
    // the task list is built in place in the pinned upload ring (one copy less); a std::vector only when the ring is full
    std::vector<PageTask> tasks_fallback;
    PageTask* tasks = nullptr;
    size_t ntasks = 0;
    bool tasks_pinned = false;
    std::vector<int> def_idx, rows_idx, gen_idx;     // routes: split def/rows kernels vs the general kernel
    std::vector<int64_t> vr_off;                      // per task: offset of its validity/rank scratch (words) or -1
    int64_t vr_words = 0;
    int64_t bytes_in = 0, bytes_out = 0;
    {
      // one pass to size the task vectors: growing them page by page costs more than filling them
      size_t npages = 0;
      for (const Json& c : jc.a) {
        auto ci = t.col_index.find(c.s);
        if (ci == t.col_index.end()) continue;
        const TableColumn& tc = t.cols[ci->second];
        for (int f : sel) npages += (size_t)(tc.file_page_begin[f + 1] - tc.file_page_begin[f]);
      }
      tasks = reinterpret_cast<PageTask*>(ar.pin_take((int64_t)(npages * sizeof(PageTask)) + 16));
      tasks_pinned = tasks != nullptr;
      if (!tasks) { tasks_fallback.resize(npages + 1); tasks = tasks_fallback.data(); }
      vr_off.reserve(npages);
      def_idx.reserve(npages);
      rows_idx.reserve(npages);
      gen_idx.reserve(npages);
    }
    for (const Json& c : jc.a) {
      const std::string& cname = c.s;
      Column col;
      col.name = cname;
      if (cname == t.part_col) {
        col.lt.kind = LType::INT;
        col.lt.is32 = true;
        int32_t* d = ar.alloc_n<int32_t>(N);
        uint32_t* valid = nullptr;
        std::vector<int64_t> seg(2 * sel.size() + 1);
        bool any_null = false;
        for (size_t k = 0; k < sel.size(); ++k) any_null |= t.files[sel[k]].part_null;
        if (any_null) {
          // validity: whole files are null; build on host (tiny) and upload
          std::vector<uint32_t> hv((N + 31) / 32 + 1, 0);
          for (size_t k = 0; k < sel.size(); ++k) {
            if (t.files[sel[k]].part_null) continue;
            for (int64_t r = row_start[k]; r < row_start[k + 1];) {
              if ((r & 31) == 0 && r + 32 <= row_start[k + 1]) { hv[r >> 5] = 0xFFFFFFFFu; r += 32; }
              else { hv[r >> 5] |= 1u << (r & 31); ++r; }
            }
          }
          valid = ar.alloc_n<uint32_t>((int64_t)hv.size());
          ar.upload(valid, hv.data(), (int64_t)hv.size() * 4);
          h2d += (int64_t)hv.size() * 4;
        }
        if (!sel.empty()) {
          std::vector<int64_t> hs(sel.size() + 1), hvv(sel.size());
          for (size_t k = 0; k < sel.size(); ++k) { hs[k] = row_start[k]; hvv[k] = t.files[sel[k]].part_null ? 0 : t.files[sel[k]].part_val; }
          hs[sel.size()] = N;
          int64_t* dseg = ar.alloc_n<int64_t>((int64_t)hs.size() + (int64_t)hvv.size());
          ar.upload(dseg, hs.data(), (int64_t)hs.size() * 8);
          ar.upload(dseg + hs.size(), hvv.data(), (int64_t)hvv.size() * 8);
          launch_fill_segments(dseg, dseg + hs.size(), (int)sel.size(), d, s);
          ++launches;
          bytes_out += N * 4;
        }
        col.data = d;
        col.valid = valid;
        out.cols.push_back(col);
        continue;
      }
      auto ci = t.col_index.find(cname);
      if (ci == t.col_index.end()) fail("plan: table '" + tname + "' has no column '" + cname + "'");
      TableColumn& tc = t.cols[ci->second];
      if (!tc.supported) fail("column '" + cname + "' of '" + tname + "' cannot be decoded: " + tc.why_unsupported);
      col.lt = tc.lt;
      col.has_range = tc.has_stats;
      col.range_lo = tc.stat_lo;
      col.range_hi = tc.stat_hi;
      const int width = tc.lt.dtype() == DT_I32 ? 4 : 8;
      void* d_out = ar.alloc(N * width);
      uint32_t* d_valid = tc.nullable ? ar.zeros<uint32_t>((N + 31) / 32 + 1) : nullptr;
      mark("scan:col_alloc");
      col.data = d_out;
      col.valid = d_valid;
      // ---- where do the encoded bytes come from
      const uint8_t* src_base = column_bytes(tc, sel);   // address of staged byte 0
      col.strbase = src_base;
      // the registration-time templates of each selected file, copied in one piece and patched (engine.h, task_tmpl)
      for (size_t k = 0; k < sel.size(); ++k) {
        const int f = sel[k];
        const int64_t p0 = tc.file_page_begin[f], p1 = tc.file_page_begin[f + 1];
        const size_t base = ntasks;
        if (p1 > p0) memcpy(tasks + ntasks, tc.task_tmpl.data() + p0, (size_t)(p1 - p0) * sizeof(PageTask));
        ntasks += (size_t)(p1 - p0);
        bytes_in += tc.file_bytes_in[f];
        const int64_t r0 = row_start[k];
        for (size_t q = base; q < ntasks; ++q) {
          PageTask& task = tasks[q];
          const bool fast = task.vr != nullptr;
          task.vr = nullptr;
          task.src = src_base + reinterpret_cast<uintptr_t>(task.src);
          if (task.dict) task.dict = src_base + (reinterpret_cast<uintptr_t>(task.dict) - 1);
          task.out = d_out;
          task.valid = d_valid;
          task.out_row += r0;
          if (fast && split_decode) {
            rows_idx.push_back((int)q);
            if (d_valid) {
              def_idx.push_back((int)q);
              vr_off.push_back(vr_words);
              vr_words += 2 * (((int64_t)task.num_values + 31) / 32) + 1;
            } else vr_off.push_back(-1);
          } else {
            gen_idx.push_back((int)q);
            vr_off.push_back(-1);
          }
        }
      }
      bytes_out += N * width + (tc.nullable ? N / 8 : 0);
      out.cols.push_back(col);
    }
    mark("scan:tasks_built");
    if (ntasks > 0) {
      // a scan of a few pages is latency-bound: one general launch beats a def + rows (+ general) sequence
      if ((int64_t)rows_idx.size() < small_scan_pages) {
        gen_idx.clear();
        for (size_t k = 0; k < ntasks; ++k) gen_idx.push_back((int)k);
        def_idx.clear();
        rows_idx.clear();
        vr_words = 0;
        std::fill(vr_off.begin(), vr_off.end(), (int64_t)-1);
      }
      // scratch of the split route: per nullable page its validity words and their rank prefix
      uint32_t* d_vr = vr_words ? ar.alloc_n<uint32_t>(vr_words) : nullptr;
      for (size_t k = 0; k < ntasks; ++k) tasks[k].vr = vr_off[k] >= 0 ? d_vr + vr_off[k] : nullptr;
      PageTask* d_tasks = ar.alloc_n<PageTask>((int64_t)ntasks);
      if (tasks_pinned) NDS_CUDA(cudaMemcpyAsync(d_tasks, tasks, ntasks * sizeof(PageTask), cudaMemcpyHostToDevice, s));
      else ar.upload(d_tasks, tasks, (int64_t)(ntasks * sizeof(PageTask)));
      std::vector<int> idx_all;
      idx_all.reserve(def_idx.size() + rows_idx.size() + gen_idx.size());
      idx_all.insert(idx_all.end(), def_idx.begin(), def_idx.end());
      idx_all.insert(idx_all.end(), rows_idx.begin(), rows_idx.end());
      idx_all.insert(idx_all.end(), gen_idx.begin(), gen_idx.end());
      int* d_idx = ar.alloc_n<int>((int64_t)idx_all.size());
      ar.upload(d_idx, idx_all.data(), (int64_t)(idx_all.size() * sizeof(int)));
      // a dedicated pair of events brackets only the decode launches (roofline.achieved in bench.py)
      mark("scan:uploaded");
      size_t ks = begin_stat("decode_kernel", true);
      int l = 0;
      int max_def_len = 0;
      for (int ti : def_idx) max_def_len = std::max<int>(max_def_len, (int)tasks[ti].def_len);
      launch_decode_split(d_tasks, d_idx, (int)def_idx.size(), max_def_len, d_idx + def_idx.size(), (int)rows_idx.size(), err, s, &l);
      launches += l;
      long long* dbg = nullptr;
      const int ngen = (int)gen_idx.size();
      const int nctas = ngen * decode_slices(ngen);
      if (decode_timing && ngen) dbg = ar.zeros<long long>((int64_t)nctas * 8);
      if (ngen) {
        launch_decode_pages(d_tasks, d_idx + def_idx.size() + rows_idx.size(), ngen, err, s, dbg);
        ++launches;
      }
      end_stat(ks, bytes_in + bytes_out, N);
      if (dbg) {
        std::vector<long long> h((size_t)nctas * 8);
        NDS_CUDA(cudaMemcpyAsync(h.data(), dbg, h.size() * 8, cudaMemcpyDeviceToHost, s));
        NDS_CUDA(cudaStreamSynchronize(s));
        double acc[6] = {0, 0, 0, 0, 0, 0};
        double defA = 0, defB = 0, defC = 0;
        for (int c = 0; c < nctas; ++c) {
          long long prev = h[c * 8];
          for (int j = 1; j < 6; ++j) {
            long long v = h[c * 8 + j];
            if (v) { acc[j] += (double)(v - prev); prev = v; }
          }
          if (h[c * 8 + 6] && h[c * 8 + 7] && h[c * 8 + 1] && h[c * 8 + 2]) {
            defA += (double)(h[c * 8 + 6] - h[c * 8 + 1]);
            defB += (double)(h[c * 8 + 7] - h[c * 8 + 6]);
            defC += (double)(h[c * 8 + 2] - h[c * 8 + 7]);
          }
        }
        char buf[320];
        snprintf(buf, sizeof buf, "decode_phases(general kernel, avg cycles/CTA over %d CTAs): stage=%.0f def=%.0f [A=%.0f B=%.0f C+=%.0f] scan+valid=%.0f headers=%.0f rows=%.0f",
                 nctas, acc[1] / nctas, acc[2] / nctas, defA / nctas, defB / nctas, defC / nctas, acc[3] / nctas, acc[4] / nctas, acc[5] / nctas);
        size_t di = begin_stat(buf);
        end_stat(di, 0, 0);
      }
      {
        char buf[160];
        snprintf(buf, sizeof buf, "decode_routes: def=%zu rows=%zu general=%zu pages", def_idx.size(), rows_idx.size(), gen_idx.size());
        size_t di = begin_stat(buf);
        end_stat(di, 0, 0);
      }
    }
    end_stat(st, bytes_in + bytes_out, N);
    return out;
  }

  // ---------------------------------------------------------------- filter
  Batch filter(const Json& node, std::vector<RuntimeFilter>& filters) {
    Batch in = exec(node.at("child"), filters);
    return filter_batch(node, in);
  }
  Batch filter_batch(const Json& node, Batch& in) {
    size_t st = begin_stat("filter");
    Blob bb;
    std::vector<Instr> code;
    LType t = compile(node.at("cond"), in, bb, code);
    if (t.kind != LType::BOOL) fail("plan: filter condition is not boolean");
    bb.add_prog(code, t);
    Blob::Dev dev = bb.upload(ar, s);
    const int64_t filter_row_bytes = row_bytes_of(bb);
    RowCount n = in.rc();
    uint32_t* sel = ar.alloc_n<uint32_t>(n.host);
    int* cnt = ar.zeros<int>(1);
    launch_filter(dev.host_progs[0], dev.ctx, n, sel, cnt, s);
    ++launches;
    Batch out;
    out.n_max = n.host;
    out.n_dev = cnt;
    MapCache mc;
    RowCount nout{cnt, n.host};
    for (const Column& c : in.cols) {
      Column o = c;
      o.rowmap = remap(c.rowmap, sel, nout, mc);
      out.cols.push_back(o);
    }
    end_stat_rows(st, filter_row_bytes + 4, n);
    return out;
  }

  // ---------------------------------------------------------------- project
  Column eval_to_column(const Json& expr, const std::string& name, Batch& in, const uint32_t* rows, RowCount n) {
    Blob bb;
    std::vector<Instr> code;
    int origin = -1;
    LType t = compile(expr, in, bb, code, &origin);
    bb.add_prog(code, t);
    Blob::Dev dev = bb.upload(ar, s);
    Column o;
    o.name = name;
    o.lt = t;
    if (t.kind == LType::BOOL) { o.lt.kind = LType::INT; o.lt.is32 = false; }
    const DType dt = o.lt.dtype();
    const uint8_t* sbase = nullptr;
    if (dt == DT_STR) {
      if (origin < 0) fail("plan: a string-valued projection must be a column or substr(column)");
      sbase = in.cols[origin].strbase;
      o.strbase = sbase;
    }
    void* d = ar.alloc(n.host * (dt == DT_I32 ? 4 : 8));
    uint32_t* v = ar.alloc_n<uint32_t>((n.host + 31) / 32 + 1);
    launch_project(dev.host_progs[0], dev.ctx, n, rows, d, v, dt, sbase, err, s);
    ++launches;
    o.data = d;
    o.valid = v;
    return o;
  }

  // Several expressions -> flat columns with ONE kernel launch (rows: optional gather indices).
  std::vector<Column> eval_columns(const std::vector<std::pair<std::string, const Json*>>& exprs, Batch& in, const uint32_t* rows,
                                   RowCount n) {
    std::vector<Column> outc;
    if (exprs.empty()) return outc;
    Blob bb;
    std::vector<ProjSpec> specs(exprs.size());
    for (size_t k = 0; k < exprs.size(); ++k) {
      std::vector<Instr> code;
      int origin = -1;
      LType t = compile(*exprs[k].second, in, bb, code, &origin);
      bb.add_prog(code, t);
      Column o;
      o.name = exprs[k].first;
      o.lt = t;
      if (t.kind == LType::BOOL) { o.lt.kind = LType::INT; o.lt.is32 = false; }
      const DType dt = o.lt.dtype();
      memset(&specs[k], 0, sizeof(ProjSpec));
      if (dt == DT_STR) {
        if (origin < 0) fail("plan: a string-valued projection must be a column or substr(column)");
        o.strbase = in.cols[origin].strbase;
        specs[k].strbase = o.strbase;
      }
      void* d = ar.alloc(n.host * (dt == DT_I32 ? 4 : 8));
      uint32_t* v = ar.alloc_n<uint32_t>((n.host + 31) / 32 + 1);
      o.data = d;
      o.valid = v;
      specs[k].out = d;
      specs[k].valid = v;
      specs[k].dtype = dt;
      outc.push_back(o);
    }
    // program pointers are only known after the upload layout is fixed: upload once to learn them, patch, upload specs
    Blob::Dev dev = bb.upload(ar, s);
    for (size_t k = 0; k < exprs.size(); ++k) specs[k].prog = dev.host_progs[k];
    ProjSpec* d_specs = ar.alloc_n<ProjSpec>((int64_t)specs.size());
    ar.upload(d_specs, specs.data(), (int64_t)(specs.size() * sizeof(ProjSpec)));
    launch_project_multi(d_specs, (int)specs.size(), dev.ctx, n, rows, err, s);
    ++launches;
    return outc;
  }

  static Json col_expr(const std::string& name) {
    Json ce;
    ce.kind = Json::Obj;
    Json cn;
    cn.kind = Json::Str;
    cn.s = name;
    ce.o.emplace_back("col", cn);
    return ce;
  }

  Batch project(const Json& node, std::vector<RuntimeFilter>& filters) {
    Batch in = exec(node.at("child"), filters);
    size_t st = begin_stat("project");
    Batch out;
    out.n_max = in.n_max;
    out.n_dev = in.n_dev;
    out.n_exact = in.n_exact;
    std::vector<std::pair<std::string, const Json*>> computed;
    std::vector<int> slot;       // position in out.cols of each computed column
    for (const Json& pe : node.at("exprs").a) {
      const std::string& name = pe.str("name");
      const Json& ex = pe.at("expr");
      if (ex.has("col")) {
        int ci = in.find(ex.str("col"));
        if (ci < 0) fail("plan: unknown column '" + ex.str("col") + "'");
        Column o = in.cols[ci];
        o.name = name;
        out.cols.push_back(o);
      } else {
        computed.emplace_back(name, &ex);
        slot.push_back((int)out.cols.size());
        out.cols.push_back(Column());
      }
    }
    if (!computed.empty()) {
      std::vector<Column> cc = eval_columns(computed, in, nullptr, in.rc());
      for (size_t k = 0; k < cc.size(); ++k) out.cols[slot[k]] = cc[k];
    }
    end_stat(st, 0, in.n_max);
    return out;
  }

  // ---------------------------------------------------------------- join
  struct BuiltSide {
    Batch build;
    HashTable ht;
    std::string pk, bk;
    bool multi = false;          // repeated build keys are chained (HashTable::next)
  };

  // The value range of the partitions of the table(s) partitioned by `pk` (dynamic partition pruning), or false.
  bool partition_range(const std::string& pk, int64_t& plo, int64_t& phi) const {
    plo = INT64_MAX;
    phi = INT64_MIN;
    for (const auto& kv : e.tables_)
      if (kv.second.part_col == pk)
        for (const FileInfo& fi : kv.second.files)
          if (!fi.part_null) { plo = std::min(plo, fi.part_val); phi = std::max(phi, fi.part_val); }
    return plo <= phi && phi - plo < (int64_t(1) << 22);
  }

  // Filter + build of a dimension in ONE launch (join_build_filter_kernel): the predicate of `fnode` is evaluated over the
  // unfiltered scan `in` and the rows that pass are entered into a direct-address table under their row in `in`, which then
  // is the build side (no selection vector, no row map between the probe's matched rows and the decoded columns); a pruning
  // join's key bitmap is filled in the same pass.  False (nothing launched) when the table cannot be a direct-address one
  // or the pruning keys would have to travel as a list.
  bool build_filtered(const Json& node, const Json& fnode, Batch& in, bool existence, BuiltSide& bs, std::vector<RuntimeFilter>& down) {
    const int kb = in.find(bs.bk);
    if (kb < 0 || !is_intlike(in.cols[kb].lt)) return false;
    const Column& kc = in.cols[kb];
    const int64_t nb_bound = in.rc().host;
    const int64_t range = kc.has_range ? kc.range_hi - kc.range_lo + 1 : 0;
    if (!dense_tables || !kc.has_range || range <= 0 || range > (int64_t(1) << 26) || range > 16 * std::max<int64_t>(nb_bound, 1) + 65536 ||
        nb_bound <= 0 || nb_bound >= (int64_t(1) << 31))
      return false;
    const bool prune = node.bool_or("prune_partitions", false);
    int64_t plo = 0, phi = -1;
    if (prune && !partition_range(bs.pk, plo, phi)) return false;
    size_t stb = begin_stat("join_build");
    Blob bb;
    std::vector<Instr> code;
    LType t = compile(fnode.at("cond"), in, bb, code);
    if (t.kind != LType::BOOL) fail("plan: filter condition is not boolean");
    bb.add_prog(code, t);
    Blob::Dev dev = bb.upload(ar, s);
    HashTable& ht = bs.ht;
    ht.dense = ar.zeros<uint32_t>(range);
    ht.lo = kc.range_lo;
    ht.range = (uint32_t)range;
    uint32_t* dbits = nullptr;
    std::shared_ptr<RuntimeFilter::State> st;
    if (prune) {
      st = std::make_shared<RuntimeFilter::State>();
      st->lo = plo;
      st->hi = phi;
      st->hbits_words = (phi - plo) / 32 + 1;
      dbits = ar.zeros<uint32_t>(st->hbits_words);
    }
    launch_join_build_filter(dev.host_progs[0], dev.ctx, kc.ref(), in.rc(), ht, err + 1, existence, plo, phi, dbits, s);
    ++launches;
    if (prune) {
      RuntimeFilter rf;
      rf.column = bs.pk;
      rf.st = st;
      uint8_t* hbuf = ar.pin_take(st->hbits_words * 4);
      if (!hbuf) { st->pageable.resize((size_t)st->hbits_words * 4); hbuf = st->pageable.data(); }
      st->hbits = reinterpret_cast<const uint32_t*>(hbuf);
      NDS_CUDA(cudaMemcpyAsync(hbuf, dbits, (size_t)st->hbits_words * 4, cudaMemcpyDeviceToHost, s));
      st->ev = e.get_event();
      NDS_CUDA(cudaEventRecord(st->ev, s));
      rf_states.push_back(st);
      d2h += st->hbits_words * 4;
      down.push_back(std::move(rf));
    }
    end_stat(stb, nb_bound * (row_bytes_of(bb) + 8) + range * 4, nb_bound);
    bs.build = in;
    return true;
  }

  // Executes the build child of a join node, builds its hash table and, for a partition-pruning join, turns the build
  // keys into a runtime filter for the scan below.
  BuiltSide build_side(const Json& node, std::vector<RuntimeFilter>& down) {
    const std::string type = node.str_or("type", "inner");
    if (type != "inner" && type != "semi" && type != "anti" && type != "left")
      fail("plan: join type '" + type + "' is not supported (inner, left, semi, anti)");
    // semi / anti joins only ask whether a key exists: repeated build keys are fine.  inner / left joins over a build side
    // that is not declared unique chain the repeated keys (multi table) and expand the probe rows.
    const bool existence = type == "semi" || type == "anti";
    const bool multi = !existence && !node.bool_or("unique_build", true);
    BuiltSide bs;
    bs.multi = multi;
    std::vector<RuntimeFilter> none;
    bs.bk = node.str("build_key");
    bs.pk = node.str("probe_key");
    const Json& bnode = node.at("build");
    if (build_filter && !multi && bnode.str("op") == "filter" && bnode.at("child").str("op") == "scan") {
      // a filtered dimension scan: filter + build in one pass when the table can be a direct-address one (build_filtered)
      Batch in = exec(bnode.at("child"), none);
      if (build_filtered(node, bnode, in, existence, bs, down)) return bs;
      bs.build = filter_batch(bnode, in);
    } else {
      bs.build = exec(bnode, none);
    }
    Batch& build = bs.build;
    const std::string& bk = bs.bk;
    int kb = build.find(bk);
    if (kb < 0) fail("plan: build key '" + bk + "' not found");
    if (!is_intlike(build.cols[kb].lt)) fail("plan: join keys must be integer columns");
    // small build sides (every dimension table): size the table from the upper bound, no host round trip
    const int64_t nb_bound = build.rc().host;
    const int64_t nb = nb_bound <= (int64_t(4) << 20) ? nb_bound : count(build);
    auto build_table = [&]() {
      size_t stb = begin_stat("join_build");
      HashTable& ht = bs.ht;
      const Column& kc = build.cols[kb];
      // surrogate keys are dense: when the column statistics bound the keys to a range not much larger than the table,
      // the join table is a direct-address array over that range (ops.cuh) instead of an open-addressing hash table
      const int64_t range = kc.has_range ? kc.range_hi - kc.range_lo + 1 : 0;
      if (dense_tables && kc.has_range && range > 0 && range <= (int64_t(1) << 26) && range <= 16 * std::max<int64_t>(nb_bound, 1) + 65536) {
        ht.dense = ar.zeros<uint32_t>(range);
        ht.lo = kc.range_lo;
        ht.range = (uint32_t)range;
        if (multi) ht.next = ar.alloc_n<uint32_t>(std::max<int64_t>(nb_bound, 1));
        launch_join_build(kc.ref(), build.rc(), ht, err + 1, s, existence);
        ++launches;
        end_stat(stb, nb * 8 + range * 4, nb);
        return;
      }
      uint32_t cap = next_pow2(2 * nb);
      ht.mask = cap - 1;
      ht.keys = ar.alloc_n<int64_t>(cap);
      ht.vals = multi ? ar.zeros<uint32_t>(cap) : ar.alloc_n<uint32_t>(cap);
      if (multi) ht.next = ar.alloc_n<uint32_t>(std::max<int64_t>(nb_bound, 1));
      launch_fill_i64(ht.keys, kEmptyKey, cap, s);
      launch_join_build(kc.ref(), build.rc(), ht, err + 1, s, existence);
      launches += 2;
      end_stat(stb, nb * 8 + (int64_t)cap * 12, nb);
    };
    const bool prune = node.bool_or("prune_partitions", false);
    if (!prune) build_table();
    // dynamic partition pruning: ship the (small) build key set to the host and let the scan skip files. The keys are
    // queued for the host BEFORE the hash table is built: the scan that waits for them is the head of the critical path.
    if (prune) {
      RuntimeFilter rf;
      rf.column = bs.pk;
      rf.st = std::make_shared<RuntimeFilter::State>();
      RuntimeFilter::State& t = *rf.st;
      // the partition values of the table(s) partitioned by the probe key bound the keys that matter
      int64_t plo = 0, phi = -1;
      if (partition_range(bs.pk, plo, phi) && nb > 0) {
        // few partitions: a bitmap over their value range is filled on the device and only that travels to the host
        t.lo = plo;
        t.hi = phi;
        t.hbits_words = (phi - plo) / 32 + 1;
        uint32_t* dbits = ar.zeros<uint32_t>(t.hbits_words);
        launch_keys_to_bitmap(build.cols[kb].ref(), build.rc(), plo, phi, dbits, s);
        ++launches;
        uint8_t* hbuf = ar.pin_take(t.hbits_words * 4);
        if (!hbuf) { t.pageable.resize((size_t)t.hbits_words * 4); hbuf = t.pageable.data(); }
        t.hbits = reinterpret_cast<const uint32_t*>(hbuf);
        NDS_CUDA(cudaMemcpyAsync(hbuf, dbits, (size_t)t.hbits_words * 4, cudaMemcpyDeviceToHost, s));
        t.ev = e.get_event();
        NDS_CUDA(cudaEventRecord(t.ev, s));
        rf_states.push_back(rf.st);
        d2h += t.hbits_words * 4;
        build_table();
        down.push_back(std::move(rf));
        return bs;
      }
      build_table();
      Json kexpr = col_expr(bk);
      Column flat = eval_to_column(kexpr, bk, build, nullptr, build.rc());
      if (nb > 0) {
        // keys of the upper-bound row range and the exact count: copies queued now, read when the scan needs them
        t.is32 = flat.lt.dtype() == DT_I32;
        t.nb = nb;
        const int64_t kbytes = nb * (t.is32 ? 4 : 8), vbytes = ((nb + 31) / 32) * 4;
        uint8_t* hbuf = ar.pin_take(kbytes + vbytes + 16);
        if (!hbuf) { t.pageable.resize((size_t)(kbytes + vbytes + 16)); hbuf = t.pageable.data(); }
        const int64_t off_v = align_up(kbytes, 8), off_c = align_up(off_v + vbytes, 4);
        t.hk = hbuf;
        t.hv = reinterpret_cast<const uint32_t*>(hbuf + off_v);
        int* hcount_p = reinterpret_cast<int*>(hbuf + off_c);
        *hcount_p = build.n_exact >= 0 ? (int)build.n_exact : (int)nb;
        t.hcount = hcount_p;
        NDS_CUDA(cudaMemcpyAsync(hbuf, flat.data, (size_t)kbytes, cudaMemcpyDeviceToHost, s));
        NDS_CUDA(cudaMemcpyAsync(hbuf + off_v, flat.valid, (size_t)vbytes, cudaMemcpyDeviceToHost, s));
        if (build.n_dev && build.n_exact < 0) NDS_CUDA(cudaMemcpyAsync(hcount_p, build.n_dev, 4, cudaMemcpyDeviceToHost, s));
        t.ev = e.get_event();
        NDS_CUDA(cudaEventRecord(t.ev, s));
        rf_states.push_back(rf.st);
        d2h += kbytes + vbytes + 4;
      } else {
        t.ready = true;
      }
      down.push_back(std::move(rf));
    }
    return bs;
  }


  // ---------------------------------------------------------------- fused scan -> probe (fused_scan.cu)
  struct FusedCol {                 // one column of the base table the fused scan touches
    std::string name;
    TableColumn* tc = nullptr;      // null: the partition column
    bool is_key = false, is_payload = false;
    int64_t task0 = 0;              // first task of this column in the combined task array
    const int32_t* d_file_page0 = nullptr;
  };
  struct FusedPlan {
    Table* t = nullptr;
    FileSel fs;
    std::vector<FusedCol> cols;                 // key columns first (probe order), then payload-only columns
    std::vector<int> key_side;                  // per key column: index into `sides` / `chain`
    int part_side = -1;                         // chain index of the join on the partition column, or -1
    PageTask* d_tasks = nullptr;
    int64_t pages_per_col = 0;                  // pages of one key column (all key columns have the same)
    FusedChunk* d_chunks = nullptr;
    int32_t* d_page_file = nullptr;             // per page of the first key column: its file among the selected files
    int nchunks = 0;
    int64_t* d_part_val = nullptr;
    uint8_t* d_part_null = nullptr;
    int64_t key_bytes_in = 0;
    cudaEvent_t ev_payload_def = nullptr;       // the payload columns' def kernel ran on another stream: the gather waits for it
  };

  // Decides whether the probe side of a join chain can run as the fused scan and, if so, selects the files, builds and
  // uploads the page tasks of every column involved and launches the def-level kernel.  Returns false (nothing launched)
  // when the shape does not fit; the caller then decodes and probes the ordinary way.
  bool fused_prepare(const Json& base_node, const std::vector<const Json*>& chain, std::vector<RuntimeFilter>& down, FusedPlan& fp) {
    if (!fused_scan || !split_decode || base_node.str("op") != "scan") return false;
    auto it = e.tables_.find(base_node.str("table"));
    if (it == e.tables_.end()) return false;
    Table& t = it->second;
    if (t.is_mem) return false;
    // why a Parquet scan under a join chain did not take the fused route shows up in the statistics (profile mode)
    auto decline = [&](const std::string& why) {
      if (profile || trace) { size_t di = begin_stat("fused_scan_declined(" + t.name + "): " + why); end_stat(di, 0, 0); }
      return false;
    };
    std::set<std::string> key_only;
    if (const Json* ko = base_node.find("key_only"))
      for (const Json& c : ko->a) key_only.insert(c.s);
    // ---- columns: probe keys in probe order (innermost join first), then the payload
    auto add_col = [&](const std::string& name) -> int {
      for (size_t i = 0; i < fp.cols.size(); ++i)
        if (fp.cols[i].name == name) return (int)i;
      FusedCol fc;
      fc.name = name;
      if (name != t.part_col) {
        auto ci = t.col_index.find(name);
        if (ci == t.col_index.end()) return -1;
        fc.tc = &t.cols[ci->second];
        if (!fc.tc->supported || fc.tc->lt.kind == LType::STR) return -1;     // strings are not gathered (yet): ordinary route
      }
      fp.cols.push_back(fc);
      return (int)fp.cols.size() - 1;
    };
    for (size_t q = 0; q < chain.size(); ++q) {
      const size_t j = chain.size() - 1 - q;
      const std::string& pk = chain[j]->str("probe_key");
      if (pk == t.part_col && !t.part_col.empty()) {
        if (fp.part_side >= 0) return decline("two joins on the partition column");
        fp.part_side = (int)j;
        continue;
      }
      const int ci = add_col(pk);
      if (ci < 0 || fp.cols[ci].is_key) return decline("probe key '" + pk + "' is not a fixed-width column, or is probed twice");
      fp.cols[ci].is_key = true;
      fp.key_side.push_back((int)j);
    }
    const int nkeys = (int)fp.key_side.size();
    if (nkeys < 1) return decline("no probe key besides the partition column");
    for (const Json& c : base_node.at("columns").a) {
      if (key_only.count(c.s)) continue;
      const int ci = add_col(c.s);
      if (ci < 0) return decline("payload column '" + c.s + "' is a string / unsupported column");
      fp.cols[ci].is_payload = true;
    }
    // ---- files
    fp.t = &t;
    fp.fs = select_files(t, down);
    const std::vector<int>& sel = fp.fs.sel;
    if (sel.empty() || fp.fs.N <= 0) return decline("no rows selected");
    if (fp.fs.N >= (int64_t(1) << 31)) return decline("2^31 or more rows selected");
    for (const FusedCol& fc : fp.cols) {
      if (!fc.tc) continue;
      for (int f : sel)
        if (!fc.tc->file_direct[f]) return decline("column '" + fc.name + "' has a page that needs a run-header walk (file " + std::to_string(f) + ")");
    }
    // the key columns must be cut into pages at the same rows
    const TableColumn& k0 = *fp.cols[0].tc;
    for (int c = 1; c < nkeys; ++c) {
      const TableColumn& kc = *fp.cols[c].tc;
      for (int f : sel) {
        const int64_t a0 = k0.file_page_begin[f], a1 = k0.file_page_begin[f + 1], b0 = kc.file_page_begin[f];
        if (kc.file_page_begin[f + 1] - b0 != a1 - a0) return decline("key columns are paged differently");
        for (int64_t p = 0; p < a1 - a0; ++p)
          if (k0.pages[a0 + p].num_values != kc.pages[b0 + p].num_values) return decline("key columns are paged differently");
      }
    }
    mark("fused:checked");
    // ---- place whole files: the page tasks themselves are expanded on the device from the registration-time templates
    // (expand_tasks_kernel); the host only adds up per-file page / chunk / scratch totals
    size_t sdef = begin_stat("def_levels", true);
    std::vector<FileJob> jobs;
    jobs.reserve(fp.cols.size() * sel.size());
    std::vector<int64_t> job_vr;                 // per job: offset (words) of the file's scratch, -1 = none
    job_vr.reserve(jobs.capacity());
    std::vector<int32_t> file_page0;             // per column with pages: sel.size() + 1 entries
    std::vector<size_t> fp0_off(fp.cols.size(), 0);
    int64_t ntasks = 0, vr_words = 0, def_bytes = 0;
    int ndef_key = 0, ndef_all = 0, nchunks = 0, max_def_key = 0, max_def_pay = 0;
    // def page list: the key columns' pages first (the probe kernel waits for those), then the payload-only columns'
    for (int pass = 0; pass < 2; ++pass) {
      for (size_t c = 0; c < fp.cols.size(); ++c) {
        FusedCol& fc = fp.cols[c];
        if (!fc.tc || (pass == 0) != fc.is_key) continue;
        TableColumn& tc = *fc.tc;
        const uint8_t* src_base = column_bytes(tc, sel);
        fc.task0 = ntasks;
        fp0_off[c] = file_page0.size();
        for (size_t k = 0; k < sel.size(); ++k) {
          const int f = sel[k];
          FileJob j;
          memset(&j, 0, sizeof j);
          j.tmpl = tc.d_tmpl;
          j.page_aux = tc.d_page_aux;
          j.src_base = src_base;
          j.row_start = fp.fs.row_start[k];
          j.page_begin = (int32_t)tc.file_page_begin[f];
          j.npages = (int32_t)(tc.file_page_begin[f + 1] - tc.file_page_begin[f]);
          j.out_task0 = (int32_t)ntasks;
          j.col_task0 = (int32_t)fc.task0;
          j.def_base = -1;
          j.chunk_base = -1;
          j.file = (int32_t)k;
          file_page0.push_back((int32_t)(ntasks - fc.task0));
          ntasks += j.npages;
          if (fc.is_key) fp.key_bytes_in += tc.file_bytes_in[f];
          int64_t vo = -1;
          if (tc.nullable) {
            j.def_base = ndef_all;
            ndef_all += j.npages;
            vo = vr_words;
            vr_words += tc.file_vr_words[f];
            def_bytes += tc.file_def_bytes[f];
            if (fc.is_key) { fp.key_bytes_in -= tc.file_def_bytes[f]; max_def_key = std::max(max_def_key, tc.file_max_def[f]); }
            else max_def_pay = std::max(max_def_pay, tc.file_max_def[f]);
          }
          if (c == 0) { j.chunk_base = nchunks; nchunks += tc.file_chunks[f]; }
          jobs.push_back(j);
          job_vr.push_back(vo);
        }
        file_page0.push_back((int32_t)(ntasks - fc.task0));
      }
      if (pass == 0) ndef_key = ndef_all;
    }
    fp.pages_per_col = (int64_t)file_page0[fp0_off[0] + sel.size()];
    fp.nchunks = nchunks;
    uint32_t* d_vr = vr_words ? ar.alloc_n<uint32_t>(vr_words) : nullptr;
    for (size_t q = 0; q < jobs.size(); ++q) jobs[q].vr_base = job_vr[q] >= 0 ? d_vr + job_vr[q] : nullptr;
    fp.d_tasks = ar.alloc_n<PageTask>(std::max<int64_t>(ntasks, 1));
    int* d_def_idx = ar.alloc_n<int>(std::max(ndef_all, 1));
    fp.d_chunks = ar.alloc_n<FusedChunk>(std::max(nchunks, 1));
    fp.d_page_file = ar.alloc_n<int32_t>(std::max<int64_t>(fp.pages_per_col, 1));
    // ---- jobs, per-file partition values, per-column page directories: one upload
    std::vector<int64_t> pv(sel.size());
    std::vector<uint8_t> pn(sel.size());
    for (size_t k = 0; k < sel.size(); ++k) { pv[k] = t.files[sel[k]].part_null ? 0 : t.files[sel[k]].part_val; pn[k] = t.files[sel[k]].part_null ? 1 : 0; }
    const int64_t off_jobs = 0;
    const int64_t off_pv = align_up(off_jobs + (int64_t)(jobs.size() * sizeof(FileJob)), 16);
    const int64_t off_pn = align_up(off_pv + (int64_t)pv.size() * 8, 16);
    const int64_t off_fp0 = align_up(off_pn + (int64_t)pn.size(), 16);
    const int64_t total = off_fp0 + (int64_t)file_page0.size() * 4 + 16;
    uint8_t* hbuf = ar.pin_take(total);
    std::vector<uint8_t> pageable;
    const bool pinned = hbuf != nullptr;
    if (!hbuf) { pageable.resize((size_t)total); hbuf = pageable.data(); }
    memcpy(hbuf + off_jobs, jobs.data(), jobs.size() * sizeof(FileJob));
    memcpy(hbuf + off_pv, pv.data(), pv.size() * 8);
    memcpy(hbuf + off_pn, pn.data(), pn.size());
    if (!file_page0.empty()) memcpy(hbuf + off_fp0, file_page0.data(), file_page0.size() * 4);
    uint8_t* d = (uint8_t*)ar.alloc(total);
    if (pinned) NDS_CUDA(cudaMemcpyAsync(d, hbuf, (size_t)total, cudaMemcpyHostToDevice, s));
    else ar.upload(d, hbuf, total);
    fp.d_part_val = reinterpret_cast<int64_t*>(d + off_pv);
    fp.d_part_null = d + off_pn;
    for (size_t c = 0; c < fp.cols.size(); ++c)
      if (fp.cols[c].tc) fp.cols[c].d_file_page0 = reinterpret_cast<const int32_t*>(d + off_fp0) + fp0_off[c];
    launch_expand_tasks(reinterpret_cast<const FileJob*>(d + off_jobs), (int)jobs.size(), fp.d_tasks, d_def_idx, fp.d_chunks, fp.d_page_file, s);
    ++launches;
    mark("fused:uploaded");
    // ---- def levels -> page-local validity + rank words: the key columns' on this stream (the probe kernel needs them),
    // the payload-only columns' on another stream, beside the probe kernel (only the gather needs those)
    int l = 0;
    launch_decode_split(fp.d_tasks, d_def_idx, ndef_key, max_def_key, nullptr, 0, err, s, &l);
    if (ndef_all > ndef_key) {
      cudaStream_t s2 = s;
      if (side_stream && (int)e.side_streams_.size() > kLowLane) s2 = e.side_streams_[kLowLane];
      if (s2 != s) {
        cudaEvent_t ev_up = e.get_event();
        NDS_CUDA(cudaEventRecord(ev_up, s));
        side_events.push_back(ev_up);
        NDS_CUDA(cudaStreamWaitEvent(s2, ev_up, 0));
      }
      launch_decode_split(fp.d_tasks, d_def_idx + ndef_key, ndef_all - ndef_key, max_def_pay, nullptr, 0, err, s2, &l);
      if (s2 != s) {
        fp.ev_payload_def = e.get_event();
        NDS_CUDA(cudaEventRecord(fp.ev_payload_def, s2));
        side_events.push_back(fp.ev_payload_def);
      }
    }
    launches += l;
    end_stat(sdef, def_bytes + vr_words * 4, fp.fs.N);
    return true;
  }

  // The probe kernel over the prepared scan (after every build side is done), then the payload of the survivors.
  Batch fused_probe(FusedPlan& fp, const std::vector<const Json*>& chain, std::vector<BuiltSide>& sides) {
    const int64_t N = fp.fs.N;
    const int nkeys = (int)fp.key_side.size();
    // survivors: a star join over a large fact scan keeps a small part of the rows; beyond 16 M selected rows there is room
    // for an eighth of them, and if that ever overflows the kernel raises err[2] and the query is run again with room for all
    const int64_t cap = (speculate && N > (int64_t(16) << 20)) ? N / 8 : N;
    size_t stp = begin_stat("scan_probe", true);
    FusedScanArgs a;
    memset(&a, 0, sizeof a);
    a.nkeys = nkeys;
    for (int c = 0; c < nkeys; ++c) {
      BuiltSide& bs = sides[(size_t)fp.key_side[c]];
      if (!is_intlike(fp.cols[c].tc->lt)) fail("plan: join keys must be integer columns");
      a.key[c].tasks = fp.d_tasks + fp.cols[c].task0;
      a.key[c].ht = bs.ht;
      a.key[c].out_build = ar.alloc_n<uint32_t>(cap);
    }
    if (fp.part_side >= 0) {
      a.has_part = 1;
      a.part_ht = sides[(size_t)fp.part_side].ht;
      a.part_out_build = ar.alloc_n<uint32_t>(cap);
      a.file_part_val = fp.d_part_val;
    }
    a.chunks = fp.d_chunks;
    a.out_row = ar.alloc_n<uint32_t>(cap);
    a.out_file = ar.alloc_n<uint32_t>(cap);
    a.out_count = ar.zeros<int>(1);
    a.cap = (uint32_t)cap;
    a.err = err;
    a.nchunks = fp.nchunks;
    // a large scan whose first join has a direct-address table: existence filter of that join's keys in shared memory
    // (exact when the key range fits 144 KB of bits, else several keys per bit), persistent CTAs
    if (probe_bloom && a.key[0].ht.dense && !a.key[0].ht.next && fp.nchunks >= probe_bloom_chunks) {
      const uint32_t range = a.key[0].ht.range;
      uint32_t shift = 0;
      while (((((uint64_t)range >> shift) + 31) / 32) > (uint64_t)kBloomMaxWords) ++shift;
      a.bloom_words = (uint32_t)((((uint64_t)range >> shift) + 31) / 32 + 1);
      a.bloom_shift = shift;
      a.bloom_range = range;
      a.bloom_lo = a.key[0].ht.lo;
      uint32_t* bits = ar.zeros<uint32_t>(a.bloom_words);
      launch_bloom_from_dense(a.key[0].ht.dense, range, shift, bits, s);
      ++launches;
      a.bloom = bits;
    }
    // the first join tested once per dictionary entry and file instead of once per row: every selected file's pages of
    // the first key column share one dictionary (one row group per file), whose entries the match kernel reads in order
    if (dict_match && !a.bloom && fp.nchunks >= dict_match_chunks) {
      const TableColumn& tc = *fp.cols[0].tc;
      const std::vector<int>& sel = fp.fs.sel;
      std::vector<int32_t> dict_task(sel.size(), -1);
      uint32_t max_dict = 0;
      bool ok = true;
      int64_t task = 0, entries = 0;
      for (size_t k = 0; k < sel.size() && ok; ++k) {
        int32_t dict = -1;
        for (int64_t p = tc.file_page_begin[sel[k]]; p < tc.file_page_begin[sel[k] + 1]; ++p, ++task) {
          const PageInfo& pg = tc.pages[p];
          if (pg.num_values > 65535u) { ok = false; break; }   // the probe kernel queues 16-bit row numbers
          if (pg.enc != 1 || pg.dict < 0) continue;            // PLAIN page: tested per row
          if (dict >= 0 && pg.dict != dict) { ok = false; break; }    // several dictionaries in one file (several row groups)
          if (dict < 0) { dict = pg.dict; dict_task[k] = (int32_t)task; }
        }
        if (dict >= 0) { max_dict = std::max(max_dict, tc.dicts[dict].count); entries += tc.dicts[dict].count; }
      }
      // worth it when the dictionaries are smaller than the rows they stand for
      const uint32_t stride = ((max_dict + 31) / 32 + 3) & ~3u;
      if (ok && max_dict > 0 && entries < N && stride <= (uint32_t)kDictMatchMaxWords) {
        size_t sdm = begin_stat("dict_match");       // inside the scan_probe statistic: part of the same operator
        uint32_t* bits = ar.alloc_n<uint32_t>((int64_t)stride * (int64_t)sel.size());
        int32_t* d_dt = ar.alloc_n<int32_t>((int64_t)sel.size());
        ar.upload(d_dt, dict_task.data(), (int64_t)sel.size() * 4);
        // coarse existence bits of a direct-address table's key range: 32 KB of shared memory in the match kernel
        const HashTable& h0 = a.key[0].ht;
        const uint32_t* coarse = nullptr;
        uint32_t cwords = 0, cshift = 0;
        if (h0.dense && h0.range > 0) {
          while (((((uint64_t)h0.range >> cshift) + 31) / 32 + 4) > (uint64_t)kDictMatchCoarseWords) ++cshift;
          cwords = (uint32_t)(((((uint64_t)h0.range >> cshift) + 31) / 32 + 4) & ~uint64_t(3));
          uint32_t* cb = ar.zeros<uint32_t>(cwords);
          launch_bloom_from_dense(h0.dense, h0.range, cshift, cb, s);
          ++launches;
          coarse = cb;
        }
        launch_dict_match(a.key[0].tasks, d_dt, (int)sel.size(), max_dict, tc.lt.dtype() == DT_I32 ? 4 : 8, h0, coarse, cwords, cshift, bits, stride, s);
        ++launches;
        a.dmatch = bits;
        a.dmatch_stride = stride;
        a.page_file = fp.d_page_file;
        a.npages = (int32_t)fp.pages_per_col;
        // a few CTAs per page when there are not many pages: several waves of CTAs keep the tail short
        a.split = fp.pages_per_col >= 148 * 8 * 4 ? 1 : (fp.pages_per_col >= 148 * 8 ? 2 : 5);
        end_stat(sdm, entries * (tc.lt.dtype() == DT_I32 ? 4 : 8) + entries / 8, entries);
      }
    }
    launch_scan_probe(a, fp.nchunks, s);
    ++launches;
    RowCount nout{a.out_count, cap};
    end_stat_rows(stp, 4 * ((int64_t)chain.size() + 2), nout, fp.key_bytes_in);
    Batch out;
    out.n_max = cap;
    out.n_dev = a.out_count;
    // ---- payload columns of the survivors
    std::vector<GatherCol> gcs;
    int64_t gather_row_bytes = 0;
    for (const FusedCol& fc : fp.cols) {
      if (!fc.is_payload) continue;
      Column col;
      col.name = fc.name;
      GatherCol g;
      memset(&g, 0, sizeof g);
      if (fc.tc) {
        col.lt = fc.tc->lt;
        col.has_range = fc.tc->has_stats;
        col.range_lo = fc.tc->stat_lo;
        col.range_hi = fc.tc->stat_hi;
        g.tasks = fp.d_tasks + fc.task0;
        g.file_page0 = fc.d_file_page0;
      } else {
        col.lt.kind = LType::INT;
        col.lt.is32 = true;
        g.is_part = 1;
      }
      g.width = col.lt.dtype() == DT_I32 ? 4 : 8;
      void* d = ar.alloc(cap * g.width);
      uint32_t* v = ar.alloc_n<uint32_t>((cap + 31) / 32 + 1);
      g.out = d;
      g.out_valid = v;
      col.data = d;
      col.valid = v;
      out.cols.push_back(col);
      gcs.push_back(g);
      gather_row_bytes += 3 * g.width + 8 + 4;       // index word + dictionary entry + value written, rank words, row id
    }
    if (!gcs.empty()) {
      if (fp.ev_payload_def) NDS_CUDA(cudaStreamWaitEvent(s, fp.ev_payload_def, 0));
      size_t stg = begin_stat("gather_decode", true);
      GatherCol* d_gcs = ar.alloc_n<GatherCol>((int64_t)gcs.size());
      ar.upload(d_gcs, gcs.data(), (int64_t)(gcs.size() * sizeof(GatherCol)));
      launch_gather_decode(d_gcs, (int)gcs.size(), a.out_row, a.out_file, a.out_count, cap, fp.d_part_val, fp.d_part_null, err, s);
      ++launches;
      end_stat_rows(stg, gather_row_bytes, nout);
    }
    // ---- build-side columns in the order the unfused plan produces them (innermost join first)
    for (size_t q = 0; q < chain.size(); ++q) {
      const size_t j = chain.size() - 1 - q;
      const BuiltSide& bs = sides[j];
      const uint32_t* map = nullptr;
      if ((int)j == fp.part_side) map = a.part_out_build;
      for (int c = 0; c < nkeys; ++c)
        if (fp.key_side[c] == (int)j) map = a.key[c].out_build;
      MapCache mb;
      for (const Column& c : bs.build.cols) {
        if (out.find(c.name) >= 0) continue;
        Column o = c;
        o.rowmap = remap(c.rowmap, map, nout, mb);
        out.cols.push_back(o);
      }
    }
    return out;
  }

  // scans, filters, projections and joins between them: such a subtree reads nothing another operator of the query
  // produces (a join inside it stays on the stream the subtree was given)
  static bool independent_subtree(const Json& node) {
    const std::string& op = node.str("op");
    if (op == "scan") return true;
    if (op == "filter" || op == "project") return independent_subtree(node.at("child"));
    if (op == "join") return !node.bool_or("prune_partitions", false) && independent_subtree(node.at("probe")) && independent_subtree(node.at("build"));
    return false;
  }

  // rows of the table a scan/filter/project subtree reads (0 when it is something else): an estimate of its cost
  int64_t subtree_rows(const Json& node) const {
    const std::string& op = node.str("op");
    if (op == "scan") {
      auto it = e.tables_.find(node.str("table"));
      return it == e.tables_.end() ? 0 : it->second.rows;
    }
    if (op == "filter" || op == "project") return subtree_rows(node.at("child"));
    if (op == "join") return subtree_rows(node.at("probe")) + subtree_rows(node.at("build"));
    return 0;
  }

  // column names a plan subtree produces (static; used to decide whether a join chain can be fused)
  static void plan_columns(const Json& node, std::set<std::string>& out) {
    const std::string& op = node.str("op");
    if (op == "scan") { for (const Json& c : node.at("columns").a) out.insert(c.s); }
    else if (op == "filter" || op == "sort") plan_columns(node.at("child"), out);
    else if (op == "join") { plan_columns(node.at("probe"), out); plan_columns(node.at("build"), out); }
    else if (op == "project") { for (const Json& e : node.at("exprs").a) out.insert(e.str("name")); }
    else if (op == "aggregate") {
      for (const Json& k : node.at("keys").a) out.insert(k.str("name"));
      for (const Json& a : node.at("aggs").a) out.insert(a.str("name"));
    }
  }

  Batch join(const Json& node, std::vector<RuntimeFilter>& filters) {
    // ---- a left-deep chain of inner joins whose probe keys all come from the innermost probe side is probed in ONE pass
    std::vector<const Json*> chain;          // outermost first
    const Json* cur = &node;
    while (cur->str("op") == "join" && (int)chain.size() < kMaxChain) {
      chain.push_back(cur);
      cur = &cur->at("probe");
    }
    const Json& base_node = *cur;
    // (a single join over a Parquet scan also takes this route: the fused scan needs no decoded key column)
    bool fuse = (chain.size() >= 2 || (base_node.str("op") == "scan" && fused_scan && split_decode)) && base_node.str("op") != "join";
    if (fuse) {
      std::set<std::string> base_cols;
      plan_columns(base_node, base_cols);
      for (const Json* j : chain)
        if (!base_cols.count(j->str("probe_key")) || j->str_or("type", "inner") != "inner" || !j->bool_or("unique_build", true)) fuse = false;
    }
    if (!fuse) {
      std::vector<RuntimeFilter> down = filters;
      BuiltSide bs = build_side(node, down);
      Batch probe = exec(node.at("probe"), down);
      int kp = probe.find(bs.pk);
      if (kp < 0) fail("plan: probe key '" + bs.pk + "' not found");
      if (!is_intlike(probe.cols[kp].lt)) fail("plan: join keys must be integer columns");
      size_t stp = begin_stat("join_probe");
      RowCount np = probe.rc();
      const std::string jtype = node.str_or("type", "inner");
      const int kind = jtype == "semi" ? JOIN_SEMI : (jtype == "anti" ? JOIN_ANTI : (jtype == "left" ? JOIN_LEFT : JOIN_INNER));
      const bool with_build = kind == JOIN_INNER || kind == JOIN_LEFT;
      uint32_t* out_p = nullptr;
      uint32_t* out_b = nullptr;
      Batch out;
      if (bs.multi) {
        // repeated build keys: count the output rows first (one read-back), then write them
        unsigned long long* tot = ar.zeros<unsigned long long>(2);
        launch_join_probe_multi_count(probe.cols[kp].ref(), np, bs.ht, kind, tot, s);
        ++launches;
        unsigned long long h_tot = 0;
        NDS_CUDA(cudaMemcpyAsync(&h_tot, tot, 8, cudaMemcpyDeviceToHost, s));
        sync("sync:join_multi_count");
        d2h += 8;
        if (h_tot >= (1ull << 31)) fail("hash join: 2^31 or more output rows; shard the tables across GPUs");
        const int64_t total = (int64_t)h_tot;
        out_p = ar.alloc_n<uint32_t>(std::max<int64_t>(total, 1));
        out_b = ar.alloc_n<uint32_t>(std::max<int64_t>(total, 1));
        launch_join_probe_multi_write(probe.cols[kp].ref(), np, bs.ht, kind, out_p, out_b, tot + 1, s);
        ++launches;
        out.n_max = total;
        out.n_exact = total;
      } else {
        out_p = ar.alloc_n<uint32_t>(np.host);
        out_b = with_build ? ar.alloc_n<uint32_t>(np.host) : nullptr;
        int* cnt = ar.zeros<int>(1);
        launch_join_probe(probe.cols[kp].ref(), np, bs.ht, out_p, out_b, cnt, s, kind);
        ++launches;
        out.n_max = np.host;
        out.n_dev = cnt;
      }
      RowCount nout = out.rc();
      MapCache mp, mb;
      for (const Column& c : probe.cols) {
        Column o = c;
        o.rowmap = remap(c.rowmap, out_p, nout, mp);
        out.cols.push_back(o);
      }
      if (with_build)
        for (const Column& c : bs.build.cols) {
          if (out.find(c.name) >= 0) continue;
          Column o = c;
          o.rowmap = remap(c.rowmap, out_b, nout, mb);
          out.cols.push_back(o);
        }
      end_stat(stp, np.host * (probe.cols[kp].lt.dtype() == DT_I32 ? 4 : 8), np.host);
      return out;
    }
    // ---- fused: all build sides first (their partition filters reach the base scan), then one probe pass
    std::vector<RuntimeFilter> down = filters;
    std::vector<BuiltSide> sides;
    // build sides that prune the base scan's partitions go first: their keys are on the way to the host while the
    // other build sides are issued
    // Neither the build sides nor the base scan depend on one another: each runs on a stream of its own (round
    // robin over the engine's side streams) and the main stream waits for all of them before the probe pass. Their
    // kernels are small-grid and latency-bound, so they overlap almost completely.
    // Issue order: the pruning build sides first (their keys start travelling to the host), then the other build sides
    // from the largest table down - and the base scan as soon as the pruning keys have arrived (polled after every
    // build side; issuing it earlier would make the host wait for them), so that the scan of the fact table, usually
    // the longest piece of the query, is not queued behind the issue time of every dimension.
    sides.resize(chain.size());
    int lane_no = 0;
    std::vector<size_t> order;
    for (size_t j = 0; j < chain.size(); ++j)
      if (chain[j]->bool_or("prune_partitions", false)) order.push_back(j);
    const size_t npruning = order.size();
    for (size_t j = 0; j < chain.size(); ++j)
      if (!chain[j]->bool_or("prune_partitions", false)) order.push_back(j);
    std::stable_sort(order.begin() + npruning, order.end(), [&](size_t a, size_t b) {
      return subtree_rows(chain[a]->at("build")) > subtree_rows(chain[b]->at("build"));
    });
    Batch base;
    bool base_done = false;
    FusedPlan fplan;
    bool fused_ready = false;
    auto issue_base = [&]() {
      const bool side = side_stream && independent_subtree(base_node);
      if (side) enter_side(kBaseLane);
      // a Parquet fact scan under the chain: decode the probe keys inside the probe kernel instead of materialising columns
      fused_ready = fused_prepare(base_node, chain, down, fplan);
      if (!fused_ready) base = exec(base_node, down);
      if (side) leave_side();
      base_done = true;
    };
    auto pruning_keys_arrived = [&]() {
      for (const RuntimeFilter& f : down) {
        const RuntimeFilter::State& t = *f.st;
        if (!t.ready && t.ev && cudaEventQuery(t.ev) != cudaSuccess) { cudaGetLastError(); return false; }
      }
      return true;
    };
    for (size_t k = 0; k < order.size(); ++k) {
      const size_t j = order[k];
      if (!base_done && k >= npruning && side_stream && pruning_keys_arrived()) issue_base();
      const bool side = side_stream && independent_subtree(chain[j]->at("build"));
      if (side) enter_side(lane_no++ % kDimLanes);
      sides[j] = build_side(*chain[j], down);
      if (side) leave_side();
    }
    if (!base_done) issue_base();
    if (fused_ready) return fused_probe(fplan, chain, sides);
    size_t stp = begin_stat("join_probe_chain");
    RowCount np = base.rc();
    ProbeChain pc;
    memset(&pc, 0, sizeof pc);
    pc.n = (int32_t)chain.size();
    int64_t key_bytes = 0;
    // probe order: innermost join first (the planner put the most selective join there)
    for (size_t q = 0; q < chain.size(); ++q) {
      const BuiltSide& bs = sides[chain.size() - 1 - q];
      int kp = base.find(bs.pk);
      if (kp < 0) fail("plan: probe key '" + bs.pk + "' not found");
      if (!is_intlike(base.cols[kp].lt)) fail("plan: join keys must be integer columns");
      pc.key[q] = base.cols[kp].ref();
      pc.ht[q] = bs.ht;
      pc.out_build[q] = ar.alloc_n<uint32_t>(np.host);
      key_bytes += np.host * (base.cols[kp].lt.dtype() == DT_I32 ? 4 : 8);
    }
    uint32_t* out_p = ar.alloc_n<uint32_t>(np.host);
    int* cnt = ar.zeros<int>(1);
    launch_join_probe_chain(pc, np, out_p, cnt, s);
    ++launches;
    Batch out;
    out.n_max = np.host;
    out.n_dev = cnt;
    RowCount nout{cnt, np.host};
    MapCache mp;
    for (const Column& c : base.cols) {
      Column o = c;
      o.rowmap = remap(c.rowmap, out_p, nout, mp);
      out.cols.push_back(o);
    }
    // build-side columns in the order the unfused plan would produce them (innermost join first)
    for (size_t q = 0; q < chain.size(); ++q) {
      const BuiltSide& bs = sides[chain.size() - 1 - q];
      MapCache mb;
      for (const Column& c : bs.build.cols) {
        if (out.find(c.name) >= 0) continue;
        Column o = c;
        o.rowmap = remap(c.rowmap, pc.out_build[q], nout, mb);
        out.cols.push_back(o);
      }
    }
    end_stat_rows(stp, (np.host > 0 ? key_bytes / np.host : 0) + 4 * ((int64_t)chain.size() + 1), np);
    return out;
  }

  // ---------------------------------------------------------------- aggregate
  Batch aggregate(const Json& node, std::vector<RuntimeFilter>& filters) {
    Batch in = exec(node.at("child"), filters);
    return aggregate_batch(node, in);
  }

  // GROUP BY ROLLUP (Spark: Expand + HashAggregate over the grouping sets): levels[0] is the finest
  // aggregate over the child, every further level aggregates the level before it (the planner wrote its keys as column
  // references and its aggregates in their merge form), and the result is the concatenation of all levels under the
  // columns of the first one - a key a level does not group by is NULL there.  Row counts stay on the device: the levels'
  // offsets in the output are a prefix sum computed by a one-thread kernel.
  Batch rollup(const Json& node, std::vector<RuntimeFilter>& filters) {
    Batch in = exec(node.at("child"), filters);
    const Json& jlevels = node.at("levels");
    const int nl = (int)jlevels.a.size();
    if (nl < 1 || nl > 16) fail("plan: rollup needs 1..16 levels");
    std::vector<Batch> lv;
    lv.reserve(nl);
    lv.push_back(aggregate_batch(jlevels.a[0], in));
    for (int i = 1; i < nl; ++i) lv.push_back(aggregate_batch(jlevels.a[i], lv[i - 1]));
    size_t st = begin_stat("rollup_union");
    // every level's columns as flat columns (row maps resolved), one launch per level
    const size_t nc = lv[0].cols.size();
    std::vector<std::vector<Column>> flat(nl);
    std::vector<std::vector<int>> present(nl, std::vector<int>(nc, 0));
    std::vector<Json> ces(nc);
    for (size_t c = 0; c < nc; ++c) ces[c] = col_expr(lv[0].cols[c].name);
    int64_t N = 0;
    for (int i = 0; i < nl; ++i) {
      std::vector<std::pair<std::string, const Json*>> exs;
      for (size_t c = 0; c < nc; ++c)
        if (lv[i].find(lv[0].cols[c].name) >= 0) { present[i][c] = 1; exs.emplace_back(lv[0].cols[c].name, &ces[c]); }
      if (lv[i].rc().host > 0) flat[i] = eval_columns(exs, lv[i], nullptr, lv[i].rc());
      N += lv[i].rc().host;
    }
    if (N >= (int64_t(1) << 31)) fail("rollup: 2^31 or more result rows");
    // per-level row counts and their prefix sum on the device
    std::vector<int> hcnt(nl, 0);
    int* d_cnt = ar.alloc_n<int>(nl);
    int* d_bound = ar.alloc_n<int>(nl);
    int* d_off = ar.alloc_n<int>(nl + 1);
    for (int i = 0; i < nl; ++i) hcnt[i] = (int)lv[i].rc().host;
    ar.upload(d_cnt, hcnt.data(), (int64_t)nl * 4);
    ar.upload(d_bound, hcnt.data(), (int64_t)nl * 4);
    for (int i = 0; i < nl; ++i)
      if (lv[i].n_dev && lv[i].n_exact < 0) NDS_CUDA(cudaMemcpyAsync(d_cnt + i, lv[i].n_dev, 4, cudaMemcpyDeviceToDevice, s));
    launch_prefix_counts(d_cnt, d_bound, nl, d_off, s);
    ++launches;
    Batch out;
    out.n_max = N;
    out.n_dev = d_off + nl;
    std::vector<UnionSeg> segs;
    const int64_t vwords = (N + 31) / 32 + 1;
    for (size_t c = 0; c < nc; ++c) {
      Column oc;
      bool have = false;
      for (int i = 0; i < nl && !have; ++i) {
        if (!present[i][c] || flat[i].empty()) continue;
        int k = 0;
        for (size_t q = 0; q < c; ++q) k += present[i][q];
        oc = flat[i][(size_t)k];
        have = true;
      }
      if (!have) {                 // no level has a row: an empty column of the declared type
        oc = lv[0].cols[c];
        oc.rowmap = nullptr;
        if (oc.lt.kind == LType::BOOL) { oc.lt.kind = LType::INT; oc.lt.is32 = false; }
      }
      const int w = oc.lt.dtype() == DT_I32 ? 4 : 8;
      void* d = ar.alloc(std::max<int64_t>(N, 1) * w);
      uint32_t* v = ar.zeros<uint32_t>(vwords);
      for (int i = 0; i < nl; ++i) {
        if (lv[i].rc().host <= 0) continue;
        UnionSeg g;
        memset(&g, 0, sizeof g);
        g.dst = d;
        g.dst_valid = v;
        g.n_dev = lv[i].n_dev && lv[i].n_exact < 0 ? lv[i].n_dev : nullptr;
        g.n_host = lv[i].rc().host;
        g.off = d_off + i;
        g.width = w;
        if (present[i][c] && !flat[i].empty()) {
          int k = 0;
          for (size_t q = 0; q < c; ++q) k += present[i][q];
          const Column& fc = flat[i][(size_t)k];
          if (fc.lt.dtype() != oc.lt.dtype()) fail("rollup: column '" + oc.name + "' changes its type between levels");
          if (fc.lt.kind == LType::STR && fc.strbase != oc.strbase) fail("rollup: string column '" + oc.name + "' has different sources");
          g.src = fc.data;
          g.src_valid = fc.valid;
          g.present = 1;
        }
        segs.push_back(g);
      }
      oc.data = d;
      oc.valid = v;
      oc.rowmap = nullptr;
      oc.has_range = false;
      out.cols.push_back(oc);
    }
    if (!segs.empty()) {
      UnionSeg* d_segs = ar.alloc_n<UnionSeg>((int64_t)segs.size());
      ar.upload(d_segs, segs.data(), (int64_t)(segs.size() * sizeof(UnionSeg)));
      int64_t biggest = 1;
      for (int i = 0; i < nl; ++i) biggest = std::max<int64_t>(biggest, lv[i].rc().host);
      launch_union_copy(d_segs, (int)segs.size(), biggest, s);
      ++launches;
    }
    end_stat(st, 0, N);
    return out;
  }

  // Window aggregates `fn(x) OVER (PARTITION BY keys)` (no ORDER BY, no frame; Spark: WindowExec): the
  // child's rows with one more column per aggregate.  The partitions are accumulated like groups (agg_insert_kernel); then
  // every row reads its partition's state (win_gather_kernel) - nothing is sorted and the groups are never compacted.
  Batch window(const Json& node, std::vector<RuntimeFilter>& filters) {
    Batch in = exec(node.at("child"), filters);
    return aggregate_batch(node, in, true);
  }

  Batch aggregate_batch(const Json& node, Batch& in, bool per_row = false) {
    const Json& jkeys = node.at("keys");
    const Json& jaggs = node.at("aggs");
    const int nkeys = (int)jkeys.a.size(), naggs = (int)jaggs.a.size();
    if (nkeys > kMaxGroupKeys) fail("plan: more than " + std::to_string(kMaxGroupKeys) + " group-by keys");
    // Row count of the input: exact if already known; otherwise its upper bound, with the table sized on the device
    // (ops.cu, agg_capacity) - unless that would reserve more than kAggReserveMax bytes, then one host round trip.
    constexpr int64_t kAggReserveMax = int64_t(6) << 30;
    int64_t n = in.rc().host;
    if (in.n_dev && in.n_exact < 0 &&
        (int64_t)agg_capacity_for(n) * (4 + 16 * (int64_t)std::max(naggs, 1)) + n * 8 * (int64_t)std::max(naggs, 1) > kAggReserveMax)
      n = count(in);
    if (per_row && nkeys == 0) fail("plan: a window aggregate needs PARTITION BY keys");
    size_t st = begin_stat(per_row ? "window" : "aggregate");
    Blob bb;
    std::vector<LType> ktypes, atypes;
    std::vector<int> korigin;
    for (const Json& k : jkeys.a) {
      std::vector<Instr> code;
      int origin = -1;
      LType t = compile(k.at("expr"), in, bb, code, &origin);
      bb.add_prog(code, t);
      ktypes.push_back(t);
      korigin.push_back(origin);
    }
    std::vector<AggSpec> specs;
    std::vector<std::vector<Instr>> agg_codes;      // compiled argument of each aggregate (empty: count(*) / count-ref)
    for (const Json& a : jaggs.a) {
      const std::string& fn = a.str("fn");
      AggSpec sp;
      LType ot;
      std::vector<Instr> agg_code_of_this;
      if (fn == "count_star") { sp.fn = AGG_COUNT_STAR; sp.prog = -1; ot.kind = LType::INT; }
      else {
        std::vector<Instr> code;
        LType t = compile(a.at("expr"), in, bb, code);
        int same_as = -1;          // an earlier aggregate over the very same compiled expression
        for (size_t j = 0; j < agg_codes.size() && same_as < 0; ++j)
          if (specs[j].fn != AGG_COUNT_STAR && specs[j].fn != AGG_COUNT_REF && agg_codes[j].size() == code.size() &&
              (code.empty() || memcmp(agg_codes[j].data(), code.data(), code.size() * sizeof(Instr)) == 0))
            same_as = (int)j;
        if (fn == "count" && same_as >= 0) {
          sp.fn = AGG_COUNT_REF;
          sp.prog = same_as;
          ot.kind = LType::INT;
          specs.push_back(sp);
          atypes.push_back(ot);
          agg_codes.emplace_back();
          continue;
        }
        sp.prog = bb.add_prog(code, t);
        agg_code_of_this = code;
        if (fn == "count") { sp.fn = AGG_COUNT; ot.kind = LType::INT; }
        else if (t.kind == LType::STR) fail("plan: " + fn + " over strings is not supported");
        else if (fn == "sum") {
          if (t.kind == LType::F64) { sp.fn = AGG_SUM_F; ot.kind = LType::F64; }
          else {
            sp.fn = AGG_SUM_I;
            ot = t;
            ot.is32 = false;
            if (t.kind == LType::DECIMAL) ot.precision = std::min(38, t.precision + 10);   // Spark: sum(dec(p,s)) -> dec(p+10,s)
            else ot.kind = LType::INT;
          }
        } else if (fn == "min" || fn == "max") {
          if (t.kind == LType::F64) fail("plan: min/max over doubles is not supported yet");
          sp.fn = fn == "min" ? AGG_MIN_I : AGG_MAX_I;
          ot = t;
          ot.is32 = false;
        } else fail("plan: unsupported aggregate function '" + fn + "'");
      }
      if (a.has("type")) {
        // the planner's declared result type (the merge side of a two-phase aggregate re-applies SUM to partial sums and
        // must not widen the decimal precision a second time)
        LType forced = parse_type(a.str("type"));
        if ((forced.kind == LType::F64) != (ot.kind == LType::F64) || forced.kind == LType::STR)
          fail("plan: type annotation '" + a.str("type") + "' contradicts aggregate '" + a.str("name") + "'");
        ot = forced;
        ot.is32 = false;
      }
      specs.push_back(sp);
      atypes.push_back(ot);
      agg_codes.push_back(std::move(agg_code_of_this));
    }
    Batch out;
    if (per_row && n == 0) {
      out = in;
      for (int a = 0; a < naggs; ++a) { Column c; c.name = jaggs.a[a].str("name"); c.lt = atypes[a]; c.data = ar.alloc(16); c.valid = ar.zeros<uint32_t>(1); out.cols.push_back(c); }
      end_stat(st, 0, 0);
      return out;
    }
    if (n == 0 && nkeys > 0) {
      out.n_max = 0;
      out.n_exact = 0;
      for (int k = 0; k < nkeys; ++k) { Column c; c.name = jkeys.a[k].str("name"); c.lt = ktypes[k]; if (c.lt.kind == LType::BOOL) c.lt.kind = LType::INT; c.data = ar.alloc(16); out.cols.push_back(c); }
      for (int a = 0; a < naggs; ++a) { Column c; c.name = jaggs.a[a].str("name"); c.lt = atypes[a]; c.data = ar.alloc(16); out.cols.push_back(c); }
      end_stat(st, 0, 0);
      return out;
    }
    // a keyless aggregate has one group whatever the input (also none at all: agg_keyless_fix_kernel): one slot, one row out
    if (nkeys == 0) n = 1;
    Blob::Dev dev = bb.upload(ar, s, specs.data(), (int64_t)specs.size() * sizeof(AggSpec));
    const AggSpec* d_specs = reinterpret_cast<const AggSpec*>(dev.extra);
    const uint32_t cap = agg_capacity_for(n);
    uint32_t* slots = ar.alloc_n<uint32_t>(cap);
    int64_t* acc = ar.alloc_n<int64_t>((int64_t)cap * std::max(naggs, 1));
    int64_t* cnt = ar.alloc_n<int64_t>((int64_t)cap * std::max(naggs, 1));
    const int64_t vwords = (n + 31) / 32 + 1;
    uint32_t* rep = ar.alloc_n<uint32_t>(n);
    int64_t* ovals = ar.alloc_n<int64_t>(n * std::max(naggs, 1));
    uint32_t* ovalid = ar.alloc_n<uint32_t>(vwords * std::max(naggs, 1));
    int* ng = ar.zeros<int>(1);
    launch_agg_init(d_specs, naggs, in.rc(), cap, slots, acc, cnt, ovalid, vwords, s);
    launch_agg_insert(dev.progs, nkeys, d_specs, naggs, dev.ctx, in.rc(), slots, cap, acc, cnt, s);
    if (per_row) {
      launch_win_gather(dev.progs, nkeys, d_specs, naggs, dev.ctx, in.rc(), slots, cap, acc, cnt, ovals, ovalid, n, vwords, s);
      launches += 3;
      out = in;
      for (int a = 0; a < naggs; ++a) {
        Column c;
        c.name = jaggs.a[a].str("name");
        c.lt = atypes[a];
        c.data = ovals + (int64_t)a * n;
        c.valid = ovalid + (int64_t)a * vwords;
        out.cols.push_back(c);
      }
      end_stat_rows(st, row_bytes_of(bb) * 2 + 4 + 24 * (int64_t)std::max(naggs, 1), in.rc());
      return out;
    }
    launch_agg_finalize(d_specs, naggs, slots, cap, in.rc(), acc, cnt, rep, ovals, ovalid, n, vwords, ng, s);
    launches += 3;
    if (nkeys == 0) { launch_agg_keyless_fix(d_specs, naggs, ng, rep, ovals, ovalid, n, vwords, s); ++launches; }
    out.n_max = n;
    out.n_dev = ng;
    RowCount nout{ng, n};
    {
      std::vector<std::pair<std::string, const Json*>> kx;
      for (int k = 0; k < nkeys; ++k) kx.emplace_back(jkeys.a[k].str("name"), &jkeys.a[k].at("expr"));
      for (Column& c : eval_columns(kx, in, rep, nout)) out.cols.push_back(c);
    }
    for (int a = 0; a < naggs; ++a) {
      Column c;
      c.name = jaggs.a[a].str("name");
      c.lt = atypes[a];
      c.data = ovals + (int64_t)a * n;
      c.valid = ovalid + (int64_t)a * vwords;
      out.cols.push_back(c);
    }
    end_stat_rows(st, row_bytes_of(bb) + 4 + 16 * (int64_t)std::max(naggs, 1), in.rc());   // + slot + aggregate states touched
    return out;
  }

  // ---------------------------------------------------------------- exchange (all-gather between the GPUs of a node)
  // Spark: HashAggregate(partial) -> ShuffleExchange -> HashAggregate(final).  Every rank
  // contributes the rows of its child (the partial aggregate states of its fact-table shard) and receives the
  // concatenation in rank order, all in device memory: row counts are all-gathered first (one host read-back, the
  // only synchronisation), then the packed segments.  With one rank the operator passes its child through;
  // options.exchange_loopback = k emulates k ranks that all hold this rank's rows (device copies instead of NCCL),
  // which exercises the pack / layout / unpack path on a single GPU.
  struct SegLayout { std::vector<int64_t> off_valid, off_data, off_aux; int64_t bytes = 0; };
  static SegLayout seg_layout(const std::vector<XCol>& proto, int64_t n) {
    SegLayout L;
    int64_t off = 0;
    for (const XCol& x : proto) {
      L.off_valid.push_back(off);
      off = align_up(off + n, 16);
      L.off_data.push_back(off);
      off = align_up(off + n * (x.kind == 1 ? 8 : 4), 16);
      L.off_aux.push_back(off);
      if (x.kind == 2) off = align_up(off + n * (int64_t)x.maxlen, 16);
    }
    L.bytes = align_up(std::max<int64_t>(off, 16), 256);
    return L;
  }

  // Hash-partition exchange (all-to-all): row i goes to rank hash(keys) % world.  Spark: ShuffleExchange(hashpartitioning);
  // here the per-destination row counts of every rank are all-gathered first (one host read-back), then every rank packs
  // one segment per destination and a grouped ncclSend / ncclRecv pair per peer moves them over NVLink.
  Batch exchange_hash(const Json& node, Batch& in, std::vector<Column>& flat, int world, int loop, size_t st) {
    const int ncols = (int)flat.size();
    if (world > 255) fail("exchange: more than 255 ranks");
    Batch fb;                       // the flattened child as a batch (the key programs read it)
    fb.n_max = in.n_max;
    fb.n_dev = in.n_dev;
    fb.n_exact = in.n_exact;
    fb.cols = flat;
    Blob bb;
    const Json& jkeys = node.at("keys");
    if (jkeys.a.empty()) fail("plan: hash exchange without keys");
    for (const Json& k : jkeys.a) {
      std::vector<Instr> code;
      Json ce = col_expr(k.s);
      LType t = compile(ce, fb, bb, code);
      bb.add_prog(code, t);
    }
    Blob::Dev dev = bb.upload(ar, s);
    const int64_t nb = std::max<int64_t>(in.rc().host, 1);
    uint8_t* dest = (uint8_t*)ar.alloc(nb);
    // meta of one rank: rows per destination [world], rows, "size assumption failed" flag, string slot width per column
    const int mlen = world + 2 + ncols;
    std::vector<int32_t> meta(mlen, 0);
    for (int c = 0; c < ncols; ++c) {
      const LType& t = flat[c].lt;
      if (t.kind == LType::STR && t.maxlen <= 0) fail("plan: exchange of string column '" + flat[c].name + "' without a known maximum length");
      meta[world + 2 + c] = t.kind == LType::STR ? t.maxlen : 0;
    }
    int32_t* d_meta = ar.alloc_n<int32_t>(mlen);
    int32_t* d_all = ar.alloc_n<int32_t>((int64_t)mlen * world);
    ar.upload(d_meta, meta.data(), mlen * 4);
    launch_exchange_dest(dev.progs, (int)jkeys.a.size(), dev.ctx, in.rc(), world, dest, d_meta, s);
    launch_set_count(in.rc(), d_meta + world, s, err + 2, d_meta + world + 1);
    launches += 2;
    if (loop > 0) {
      for (int r = 0; r < world; ++r)
        NDS_CUDA(cudaMemcpyAsync(d_all + (int64_t)r * mlen, d_meta, mlen * 4, cudaMemcpyDeviceToDevice, s));
    } else {
      nccl_check(nccl_api().AllGather(d_meta, d_all, (size_t)mlen * 4, kNcclChar, e.comm_, s), "ncclAllGather(partition counts)");
    }
    std::vector<int32_t> all((size_t)mlen * world);
    uint8_t* pin = ar.pin_take((int64_t)all.size() * 4);
    int32_t* hall = pin ? reinterpret_cast<int32_t*>(pin) : all.data();
    NDS_CUDA(cudaMemcpyAsync(hall, d_all, all.size() * 4, cudaMemcpyDeviceToHost, s));
    sync("sync:exchange_counts");
    d2h += (int64_t)all.size() * 4;
    for (int r = 0; r < world; ++r)
      if (hall[(int64_t)r * mlen + world + 1]) throw SpeculationFailed();
    const int my = loop > 0 ? 0 : e.comm_rank_;
    std::vector<XCol> proto(ncols);
    for (int c = 0; c < ncols; ++c) {
      XCol& x = proto[c];
      memset(&x, 0, sizeof x);
      const LType& t = flat[c].lt;
      x.kind = t.kind == LType::STR ? 2 : (t.dtype() == DT_I32 ? 0 : 1);
      for (int r = 0; r < world; ++r) x.maxlen = std::max(x.maxlen, hall[(int64_t)r * mlen + world + 2 + c]);
    }
    const int64_t n_local = std::max<int32_t>(hall[(int64_t)my * mlen + world], 0);
    in.n_exact = n_local;
    std::vector<int64_t> begin(world + 1, 0), send_off(world + 1, 0), recv_off(world + 1, 0), row_begin(world + 1, 0);
    std::vector<SegLayout> Ls(world), Lr(world);
    int64_t max_from = 0;
    for (int d = 0; d < world; ++d) {
      const int64_t to = std::max<int32_t>(hall[(int64_t)my * mlen + d], 0), from = std::max<int32_t>(hall[(int64_t)d * mlen + my], 0);
      begin[d + 1] = begin[d] + to;
      Ls[d] = seg_layout(proto, to);
      Lr[d] = seg_layout(proto, from);
      send_off[d + 1] = send_off[d] + Ls[d].bytes;
      recv_off[d + 1] = recv_off[d] + Lr[d].bytes;
      row_begin[d + 1] = row_begin[d] + from;
      max_from = std::max(max_from, from);
    }
    if (begin[world] != n_local) fail("exchange: partition counts do not add up");
    const int64_t N = row_begin[world];
    if (N >= (int64_t(1) << 31)) fail("exchange: 2^31 or more rows arrive at one rank");
    uint8_t* d_send = (uint8_t*)ar.alloc(std::max<int64_t>(send_off[world], 16));
    uint8_t* d_recv = (uint8_t*)ar.alloc(std::max<int64_t>(recv_off[world], 16));
    // ---- rows ordered by destination, one packed segment per destination
    if (n_local > 0) {
      int64_t* d_begin = ar.alloc_n<int64_t>(world + 1);
      ar.upload(d_begin, begin.data(), (world + 1) * 8);
      int32_t* cursor = ar.zeros<int32_t>(world);
      uint32_t* perm = ar.alloc_n<uint32_t>(n_local);
      launch_exchange_perm(dest, n_local, d_begin, cursor, perm, s);
      std::vector<XCol> desc((size_t)ncols * (world + 1));
      for (int c = 0; c < ncols; ++c) {
        XCol x = proto[c];
        x.data = flat[c].data;
        x.valid = flat[c].valid;
        x.strbase = flat[c].strbase;
        desc[c] = x;
      }
      for (int d = 0; d < world; ++d)
        for (int c = 0; c < ncols; ++c) {
          XCol x = proto[c];
          x.data = d_send + send_off[d];
          x.off_valid = Ls[d].off_valid[c];
          x.off_data = Ls[d].off_data[c];
          x.off_aux = Ls[d].off_aux[c];
          desc[(size_t)ncols * (d + 1) + c] = x;
        }
      XCol* d_desc = ar.alloc_n<XCol>((int64_t)desc.size());
      ar.upload(d_desc, desc.data(), (int64_t)(desc.size() * sizeof(XCol)));
      launch_exchange_pack_perm(d_desc, d_desc + ncols, ncols, perm, d_begin, world, n_local, err, s);
      launches += 2;
    }
    // ---- all-to-all
    int64_t bytes_out = 0;
    if (loop > 0) {
      for (int r = 0; r < world; ++r)        // every emulated rank holds my rows: each sends me its partition 0
        if (Lr[r].bytes > 0 && hall[(int64_t)r * mlen + my] > 0)
          NDS_CUDA(cudaMemcpyAsync(d_recv + recv_off[r], d_send + send_off[my], (size_t)Ls[my].bytes, cudaMemcpyDeviceToDevice, s));
    } else {
      const NcclApi& nc = nccl_api();
      nccl_check(nc.GroupStart(), "ncclGroupStart");
      for (int p = 0; p < world; ++p) {
        const int64_t to = hall[(int64_t)my * mlen + p], from = hall[(int64_t)p * mlen + my];
        if (p == my) {
          if (to > 0) NDS_CUDA(cudaMemcpyAsync(d_recv + recv_off[p], d_send + send_off[p], (size_t)Ls[p].bytes, cudaMemcpyDeviceToDevice, s));
          continue;
        }
        if (to > 0) { nccl_check(nc.Send(d_send + send_off[p], (size_t)Ls[p].bytes, kNcclChar, p, e.comm_, s), "ncclSend"); bytes_out += Ls[p].bytes; }
        if (from > 0) nccl_check(nc.Recv(d_recv + recv_off[p], (size_t)Lr[p].bytes, kNcclChar, p, e.comm_, s), "ncclRecv");
      }
      nccl_check(nc.GroupEnd(), "ncclGroupEnd");
    }
    // ---- unpack the received segments into concatenated columns
    Batch out;
    out.n_max = N;
    out.n_exact = N;
    std::vector<XCol> srcs((size_t)world * ncols), dst(ncols);
    const int64_t vwords = (N + 31) / 32 + 1;
    for (int c = 0; c < ncols; ++c) {
      Column oc;
      oc.name = flat[c].name;
      oc.lt = flat[c].lt;
      if (oc.lt.kind == LType::STR) oc.lt.maxlen = proto[c].maxlen;
      const int w = proto[c].kind == 0 ? 4 : 8;
      oc.data = ar.alloc(std::max<int64_t>(N, 1) * w);
      oc.valid = ar.zeros<uint32_t>(vwords);
      oc.strbase = proto[c].kind == 2 ? d_recv : nullptr;
      out.cols.push_back(oc);
      memset(&dst[c], 0, sizeof(XCol));
      dst[c].data = oc.data;
      dst[c].valid = oc.valid;
      dst[c].kind = proto[c].kind;
    }
    for (int r = 0; r < world; ++r)
      for (int c = 0; c < ncols; ++c) {
        XCol x = proto[c];
        x.data = d_recv + recv_off[r];
        x.off_valid = Lr[r].off_valid[c];
        x.off_data = Lr[r].off_data[c];
        x.off_aux = Lr[r].off_aux[c];
        srcs[(size_t)r * ncols + c] = x;
      }
    if (N > 0) {
      const int64_t bytes_desc = (int64_t)(srcs.size() + dst.size()) * (int64_t)sizeof(XCol) + (world + 1) * 8;
      uint8_t* d_desc = (uint8_t*)ar.alloc(bytes_desc);
      std::vector<uint8_t> h((size_t)bytes_desc);
      memcpy(h.data(), srcs.data(), srcs.size() * sizeof(XCol));
      memcpy(h.data() + srcs.size() * sizeof(XCol), dst.data(), dst.size() * sizeof(XCol));
      memcpy(h.data() + (srcs.size() + dst.size()) * sizeof(XCol), row_begin.data(), (world + 1) * 8);
      ar.upload(d_desc, h.data(), bytes_desc);
      const XCol* d_srcs = reinterpret_cast<const XCol*>(d_desc);
      launch_exchange_unpack(d_srcs, d_srcs + srcs.size(), ncols, reinterpret_cast<const int64_t*>(d_desc + (srcs.size() + dst.size()) * sizeof(XCol)),
                             world, max_from, d_recv, s);
      ++launches;
    }
    stats[st].op = "exchange:hash";
    end_stat(st, bytes_out, N);       // bytes that leave this GPU over NVLink
    return out;
  }

  Batch exchange(const Json& node, std::vector<RuntimeFilter>& filters) {
    Batch in = exec(node.at("child"), filters);
    const int loop = (int)(opts ? opts->int_or("exchange_loopback", 0) : 0);
    const int world = loop > 0 ? loop : e.comm_world_;
    if (world <= 1) return in;
    if (loop <= 0 && !e.comm_) fail("plan: exchange needs a communicator (nds_engine_comm_init) when the engine's world size is > 1");
    size_t st = begin_stat("exchange");
    const int ncols = (int)in.cols.size();
    // flatten the columns (row maps resolved) - one launch
    std::vector<Column> flat = in.cols;
    {
      std::vector<Json> ces;
      std::vector<size_t> which;
      for (size_t c = 0; c < in.cols.size(); ++c)
        if (in.cols[c].rowmap) { which.push_back(c); ces.push_back(col_expr(in.cols[c].name)); }
      std::vector<std::pair<std::string, const Json*>> exs;
      for (size_t k = 0; k < which.size(); ++k) exs.emplace_back(in.cols[which[k]].name, &ces[k]);
      if (!which.empty() && in.rc().host > 0) {
        std::vector<Column> fc = eval_columns(exs, in, nullptr, in.rc());
        for (size_t k = 0; k < which.size(); ++k) flat[which[k]] = fc[k];
      }
    }
    if (node.str_or("mode", "gather") == "hash") return exchange_hash(node, in, flat, world, loop, st);
    // ---- peer-push route (default between the GPUs of one node): no NCCL call, no host round trip.  The push kernel packs
    // this rank's rows straight into its slot of every peer's window over NVLink and raises the peers' arrival counters;
    // the unpack kernel waits for all `world` arrivals and copies the rows out.  More rows than a slot holds, or a failed
    // size assumption on any rank, sends every rank to the second attempt (exact NCCL route).
    if (speculate && exchange_peer && ncols <= kXchgHeader / 4 - 2 && world <= kPeerMaxWorld && (loop > 0 || e.xpeer_ok_)) {
      std::vector<XCol> proto(ncols);
      bool ok_types = true;
      int64_t row_bytes = 0;
      for (int c = 0; c < ncols; ++c) {
        XCol& x = proto[c];
        memset(&x, 0, sizeof x);
        const LType& t = flat[c].lt;
        x.kind = t.kind == LType::STR ? 2 : (t.dtype() == DT_I32 ? 0 : 1);
        x.maxlen = t.kind == LType::STR ? t.maxlen : 0;
        if (t.kind == LType::STR && t.maxlen <= 0) ok_types = false;
        row_bytes += 1 + (x.kind == 1 ? 8 : 4) + x.maxlen;
      }
      if (ok_types) {
        e.ensure_window();
        PeerWindow w;
        memset(&w, 0, sizeof w);
        w.world = world;
        w.me = loop > 0 ? 0 : e.comm_rank_;
        w.half_bytes = e.xwin_half_bytes_;
        w.slot_bytes = (e.xwin_half_bytes_ / world) & ~int64_t(255);
        w.ctrl_bytes = kPeerCtrlBytes;
        for (int p = 0; p < world; ++p) w.win[p] = loop > 0 ? e.xwin_ : e.xpeers_[p];
        // rows a slot holds: the layout is per column, so solve for it (and keep the bound modest: 256 k rows)
        int64_t cap_rows = std::min<int64_t>((w.slot_bytes - kXchgHeader - 256 - 48 * ncols) / std::max<int64_t>(row_bytes, 1), int64_t(256) << 10);
        while (cap_rows > 0 && kXchgHeader + seg_layout(proto, cap_rows).bytes > w.slot_bytes) --cap_rows;
        if (cap_rows >= 1) {
          SegLayout L = seg_layout(proto, cap_rows);
          for (int c = 0; c < ncols; ++c) { proto[c].off_valid = L.off_valid[c]; proto[c].off_data = L.off_data[c]; proto[c].off_aux = L.off_aux[c]; }
          std::vector<XCol> mine = proto, dst(ncols);
          for (int c = 0; c < ncols; ++c) { mine[c].data = flat[c].data; mine[c].valid = flat[c].valid; mine[c].strbase = flat[c].strbase; }
          const int64_t Nmax = cap_rows * world;
          const int64_t vwords = (Nmax + 31) / 32 + 1;
          Batch out;
          out.n_max = Nmax;
          out.n_dev = ar.zeros<int>(1);
          for (int c = 0; c < ncols; ++c) {
            Column oc;
            oc.name = flat[c].name;
            oc.lt = flat[c].lt;
            const int wd = proto[c].kind == 0 ? 4 : 8;
            oc.data = ar.alloc(Nmax * wd);
            oc.valid = ar.zeros<uint32_t>(vwords);
            oc.strbase = proto[c].kind == 2 ? e.xwin_ : nullptr;
            out.cols.push_back(oc);
            memset(&dst[c], 0, sizeof(XCol));
            dst[c].data = oc.data;
            dst[c].valid = oc.valid;
            dst[c].kind = proto[c].kind;
          }
          std::vector<XCol> all3;
          all3.insert(all3.end(), mine.begin(), mine.end());
          all3.insert(all3.end(), proto.begin(), proto.end());
          all3.insert(all3.end(), dst.begin(), dst.end());
          XCol* d_desc = ar.alloc_n<XCol>((int64_t)all3.size());
          ar.upload(d_desc, all3.data(), (int64_t)(all3.size() * sizeof(XCol)));
          const int half = (int)(e.xwin_seq_++ & 1);
          const uint32_t wanted = (uint32_t)(e.xwin_uses_[half] += world);      // arrivals this half has seen, this use included
          int* ticket = ar.zeros<int>(1);
          launch_peer_push(d_desc, d_desc + ncols, ncols, in.rc(), cap_rows, w, half, loop > 0 ? 1 : 0, ticket, err, s);
          launch_peer_unpack(d_desc + ncols, d_desc + 2 * ncols, ncols, w, half, wanted, cap_rows, out.n_dev, err, s);
          launches += 2;
          stats[st].op = "exchange:peer";
          end_stat_rows(st, row_bytes, RowCount{out.n_dev, Nmax});
          return out;
        }
      }
    }
    // ---- speculative route: ONE collective and no host round trip.  Every rank sends a fixed-capacity segment whose
    // header holds its row count; the unpack kernel reads the headers on the device.  More rows than the capacity raise
    // err[2] and the query runs again on the exact two-step route below.
    constexpr int64_t kXchgSpecRows = int64_t(128) << 10;
    if (speculate && exchange_single && ncols <= kXchgHeader / 4 - 2) {
      std::vector<XCol> proto(ncols);
      bool ok_types = true;
      for (int c = 0; c < ncols; ++c) {
        XCol& x = proto[c];
        memset(&x, 0, sizeof x);
        const LType& t = flat[c].lt;
        x.kind = t.kind == LType::STR ? 2 : (t.dtype() == DT_I32 ? 0 : 1);
        x.maxlen = t.kind == LType::STR ? t.maxlen : 0;
        if (t.kind == LType::STR && t.maxlen <= 0) ok_types = false;
      }
      // (the capacity must be the same on every rank: nothing rank-local - like this rank's row bound - may enter it)
      const int64_t cap_rows = opts ? std::max<int64_t>(opts->int_or("exchange_rows", kXchgSpecRows), 1) : kXchgSpecRows;
      SegLayout L = seg_layout(proto, cap_rows);
      const int64_t seg_bytes = kXchgHeader + L.bytes;
      if (ok_types && seg_bytes * world < (int64_t(1) << 32)) {
        for (int c = 0; c < ncols; ++c) { proto[c].off_valid = L.off_valid[c]; proto[c].off_data = L.off_data[c]; proto[c].off_aux = L.off_aux[c]; }
        std::vector<XCol> mine = proto, dst(ncols);
        for (int c = 0; c < ncols; ++c) { mine[c].data = flat[c].data; mine[c].valid = flat[c].valid; mine[c].strbase = flat[c].strbase; }
        uint8_t* d_send = (uint8_t*)ar.alloc(seg_bytes);
        uint8_t* d_recv = (uint8_t*)ar.alloc(seg_bytes * world);
        const int64_t Nmax = cap_rows * world;
        const int64_t vwords = (Nmax + 31) / 32 + 1;
        Batch out;
        out.n_max = Nmax;
        out.n_dev = ar.zeros<int>(1);
        for (int c = 0; c < ncols; ++c) {
          Column oc;
          oc.name = flat[c].name;
          oc.lt = flat[c].lt;
          const int w = proto[c].kind == 0 ? 4 : 8;
          oc.data = ar.alloc(Nmax * w);
          oc.valid = ar.zeros<uint32_t>(vwords);
          oc.strbase = proto[c].kind == 2 ? d_recv : nullptr;
          out.cols.push_back(oc);
          memset(&dst[c], 0, sizeof(XCol));
          dst[c].data = oc.data;
          dst[c].valid = oc.valid;
          dst[c].kind = proto[c].kind;
        }
        // descriptors: [mine][proto][dst] in one upload
        std::vector<XCol> all3;
        all3.insert(all3.end(), mine.begin(), mine.end());
        all3.insert(all3.end(), proto.begin(), proto.end());
        all3.insert(all3.end(), dst.begin(), dst.end());
        XCol* d_desc = ar.alloc_n<XCol>((int64_t)all3.size());
        ar.upload(d_desc, all3.data(), (int64_t)(all3.size() * sizeof(XCol)));
        launch_exchange_pack_dev(d_desc, ncols, in.rc(), cap_rows, d_send, err, s);
        ++launches;
        if (loop > 0) {
          for (int r = 0; r < world; ++r)
            NDS_CUDA(cudaMemcpyAsync(d_recv + (int64_t)r * seg_bytes, d_send, (size_t)seg_bytes, cudaMemcpyDeviceToDevice, s));
        } else {
          nccl_check(nccl_api().AllGather(d_send, d_recv, (size_t)seg_bytes, kNcclChar, e.comm_, s), "ncclAllGather(segments)");
        }
        launch_exchange_unpack_dev(d_desc + ncols, d_desc + 2 * ncols, ncols, world, cap_rows, seg_bytes, d_recv, out.n_dev, err, s);
        ++launches;
        end_stat(st, seg_bytes * (world - 1), Nmax);
        return out;
      }
    }
    // ---- counts (and string slot widths: the widest any rank declares) travel first
    std::vector<int32_t> meta(2 + ncols, 0);
    for (int c = 0; c < ncols; ++c) {
      const LType& t = flat[c].lt;
      if (t.kind == LType::STR && t.maxlen <= 0) fail("plan: exchange of string column '" + flat[c].name + "' without a known maximum length");
      meta[1 + c] = t.kind == LType::STR ? t.maxlen : 0;
    }
    // the last word carries this rank's "a device-side size assumption failed" flag (err[2]): a rank that has to run the
    // query again must take every other rank with it, or the collectives of the second attempt would have no partners
    const int mlen = 2 + ncols;
    int32_t* d_meta = ar.alloc_n<int32_t>(mlen);
    int32_t* d_all = ar.alloc_n<int32_t>((int64_t)mlen * world);
    ar.upload(d_meta, meta.data(), mlen * 4);
    launch_set_count(in.rc(), d_meta, s, err + 2, d_meta + 1 + ncols);
    ++launches;
    if (loop > 0) {
      for (int r = 0; r < world; ++r)
        NDS_CUDA(cudaMemcpyAsync(d_all + (int64_t)r * mlen, d_meta, mlen * 4, cudaMemcpyDeviceToDevice, s));
    } else {
      nccl_check(nccl_api().AllGather(d_meta, d_all, (size_t)mlen * 4, kNcclChar, e.comm_, s), "ncclAllGather(counts)");
    }
    std::vector<int32_t> all((size_t)mlen * world);
    uint8_t* pin = ar.pin_take((int64_t)all.size() * 4);
    int32_t* hall = pin ? reinterpret_cast<int32_t*>(pin) : all.data();
    NDS_CUDA(cudaMemcpyAsync(hall, d_all, all.size() * 4, cudaMemcpyDeviceToHost, s));
    sync("sync:exchange_counts");
    d2h += (int64_t)all.size() * 4;
    for (int r = 0; r < world; ++r)
      if (hall[(int64_t)r * mlen + 1 + ncols]) throw SpeculationFailed();
    std::vector<int64_t> row_begin(world + 1, 0);
    int64_t max_rows = 0;
    std::vector<XCol> proto(ncols);
    for (int c = 0; c < ncols; ++c) {
      XCol& x = proto[c];
      memset(&x, 0, sizeof x);
      const LType& t = flat[c].lt;
      x.kind = t.kind == LType::STR ? 2 : (t.dtype() == DT_I32 ? 0 : 1);
      for (int r = 0; r < world; ++r) x.maxlen = std::max(x.maxlen, hall[(int64_t)r * mlen + 1 + c]);
    }
    for (int r = 0; r < world; ++r) {
      const int64_t nr = std::max<int32_t>(hall[(int64_t)r * mlen], 0);
      row_begin[r + 1] = row_begin[r] + nr;
      max_rows = std::max(max_rows, nr);
    }
    const int my = loop > 0 ? 0 : e.comm_rank_;
    const int64_t n_mine = row_begin[my + 1] - row_begin[my];
    in.n_exact = n_mine;
    const int64_t N = row_begin[world];
    const int64_t seg_bytes = seg_layout(proto, max_rows).bytes;        // every rank sends this many bytes
    if (seg_bytes * world >= (int64_t(1) << 32)) fail("exchange: more than 4 GiB of partial state; repartition instead of gathering");
    // ---- pack my segment
    uint8_t* d_send = (uint8_t*)ar.alloc(seg_bytes);
    uint8_t* d_recv = (uint8_t*)ar.alloc(seg_bytes * world);
    {
      SegLayout L = seg_layout(proto, n_mine);
      std::vector<XCol> mine = proto;
      for (int c = 0; c < ncols; ++c) {
        mine[c].data = flat[c].data;
        mine[c].valid = flat[c].valid;
        mine[c].strbase = flat[c].strbase;
        mine[c].off_valid = L.off_valid[c];
        mine[c].off_data = L.off_data[c];
        mine[c].off_aux = L.off_aux[c];
      }
      XCol* d_cols = ar.alloc_n<XCol>(ncols);
      ar.upload(d_cols, mine.data(), (int64_t)ncols * (int64_t)sizeof(XCol));
      launch_exchange_pack(d_cols, ncols, n_mine, d_send, err, s);
      ++launches;
    }
    if (loop > 0) {
      for (int r = 0; r < world; ++r)
        NDS_CUDA(cudaMemcpyAsync(d_recv + (int64_t)r * seg_bytes, d_send, (size_t)seg_bytes, cudaMemcpyDeviceToDevice, s));
    } else {
      nccl_check(nccl_api().AllGather(d_send, d_recv, (size_t)seg_bytes, kNcclChar, e.comm_, s), "ncclAllGather(segments)");
    }
    // ---- unpack into concatenated columns
    Batch out;
    out.n_max = N;
    out.n_exact = N;
    std::vector<XCol> srcs((size_t)world * ncols), dst(ncols);
    const int64_t vwords = (N + 31) / 32 + 1;
    for (int c = 0; c < ncols; ++c) {
      Column oc;
      oc.name = flat[c].name;
      oc.lt = flat[c].lt;
      if (oc.lt.kind == LType::STR) oc.lt.maxlen = proto[c].maxlen;
      const int w = proto[c].kind == 0 ? 4 : 8;
      oc.data = ar.alloc(std::max<int64_t>(N, 1) * w);
      oc.valid = ar.zeros<uint32_t>(vwords);
      oc.strbase = proto[c].kind == 2 ? d_recv : nullptr;
      out.cols.push_back(oc);
      memset(&dst[c], 0, sizeof(XCol));
      dst[c].data = oc.data;
      dst[c].valid = oc.valid;
      dst[c].kind = proto[c].kind;
    }
    for (int r = 0; r < world; ++r) {
      SegLayout L = seg_layout(proto, row_begin[r + 1] - row_begin[r]);
      for (int c = 0; c < ncols; ++c) {
        XCol x = proto[c];
        x.data = d_recv + (int64_t)r * seg_bytes;
        x.off_valid = L.off_valid[c];
        x.off_data = L.off_data[c];
        x.off_aux = L.off_aux[c];
        srcs[(size_t)r * ncols + c] = x;
      }
    }
    if (N > 0) {
      const int64_t bytes_desc = (int64_t)(srcs.size() + dst.size()) * (int64_t)sizeof(XCol) + (world + 1) * 8;
      uint8_t* d_desc = (uint8_t*)ar.alloc(bytes_desc);
      std::vector<uint8_t> h((size_t)bytes_desc);
      memcpy(h.data(), srcs.data(), srcs.size() * sizeof(XCol));
      memcpy(h.data() + srcs.size() * sizeof(XCol), dst.data(), dst.size() * sizeof(XCol));
      memcpy(h.data() + (srcs.size() + dst.size()) * sizeof(XCol), row_begin.data(), (world + 1) * 8);
      ar.upload(d_desc, h.data(), bytes_desc);
      const XCol* d_srcs = reinterpret_cast<const XCol*>(d_desc);
      const XCol* d_dst = d_srcs + srcs.size();
      const int64_t* d_rb = reinterpret_cast<const int64_t*>(d_desc + (srcs.size() + dst.size()) * sizeof(XCol));
      launch_exchange_unpack(d_srcs, d_dst, ncols, d_rb, world, max_rows, d_recv, s);
      ++launches;
    }
    end_stat(st, seg_bytes * (world - 1), N);     // bytes that cross NVLink into this GPU
    return out;
  }

  // ---------------------------------------------------------------- sort / top-k
  Batch sort(const Json& node, std::vector<RuntimeFilter>& filters) {
    Batch in = exec(node.at("child"), filters);
    return sort_batch(node, in);
  }

  // rank() / row_number() / dense_rank() OVER (PARTITION BY p ORDER BY o) (Spark: WindowExec over a sort):
  // the child's rows sorted by (p, o) - the node's `keys`, the first `npart` of them the partition keys - plus one int64 column
  // per function.  Partition and tie-group starts come from comparing neighbouring rows and a max-scan (ops.cu, launch_rank).
  Batch rank(const Json& node, std::vector<RuntimeFilter>& filters) {
    Batch in = exec(node.at("child"), filters);
    Batch out = sort_batch(node, in);
    size_t st = begin_stat("rank");
    const Json& jkeys = node.at("keys");
    const int nkeys = (int)jkeys.a.size(), npart = (int)node.int_or("npart", 0);
    if (npart < 0 || npart > nkeys || nkeys > 16) fail("plan: rank node with bad key counts");
    const int64_t n = out.rc().host;
    const Json& jf = node.at("funcs");
    int64_t* cols3[3] = {nullptr, nullptr, nullptr};
    std::vector<std::pair<std::string, int>> names;
    for (const Json& f : jf.a) {
      const std::string& fn = f.str("fn");
      const int which = fn == "rank" ? 0 : (fn == "row_number" ? 1 : (fn == "dense_rank" ? 2 : -1));
      if (which < 0) fail("plan: unsupported ranking function '" + fn + "'");
      if (!cols3[which]) cols3[which] = ar.alloc_n<int64_t>(std::max<int64_t>(n, 1));
      names.emplace_back(f.str("name"), which);
    }
    if (n > 0) {
      Blob bb;
      for (const Json& k : jkeys.a) {
        std::vector<Instr> code;
        LType t = compile(k.at("expr"), out, bb, code);
        bb.add_prog(code, t);
      }
      Blob::Dev dev = bb.upload(ar, s);
      uint32_t* scratch = ar.alloc_n<uint32_t>(3 * n);
      launch_rank(dev.progs, nkeys, npart, dev.ctx, out.rc(), scratch, scratch + n, scratch + 2 * n, cols3[0], cols3[1], cols3[2], s);
      launches += 3;
    }
    const int64_t vwords = (std::max<int64_t>(n, 1) + 31) / 32 + 1;
    uint32_t* ones = ar.alloc_n<uint32_t>(vwords);
    NDS_CUDA(cudaMemsetAsync(ones, 0xFF, (size_t)vwords * 4, s));
    for (const auto& nm : names) {
      Column c;
      c.name = nm.first;
      c.lt.kind = LType::INT;
      c.data = cols3[nm.second];
      c.valid = ones;
      out.cols.push_back(c);
    }
    end_stat(st, n * 16, n);
    return out;
  }

  Batch sort_batch(const Json& node, Batch& in) {
    const int64_t limit = node.int_or("limit", -1);
    size_t st = begin_stat("sort");
    // compile the keys first: the string keys' maximum lengths are measured on the device with the row count still
    // there, so that count and lengths come back in ONE synchronisation
    Blob bb;
    std::vector<SortKey> keys;
    std::vector<LType> key_types;
    int B = 0;
    for (const Json& k : node.at("keys").a) {
      std::vector<Instr> code;
      LType t = compile(k.at("expr"), in, bb, code);
      key_types.push_back(t);
      SortKey sk;
      sk.prog = bb.add_prog(code, t);
      sk.kind = t.kind == LType::STR ? 2 : (t.kind == LType::F64 ? 1 : 0);
      sk.desc = k.bool_or("desc", false) ? 1 : 0;
      sk.nulls_first = k.bool_or("nulls_first", !sk.desc) ? 1 : 0;   // Spark default null ordering
      sk.width = 8;
      sk.offset = 0;
      keys.push_back(sk);
    }
    // string key widths: the column's known maximum length (dictionary / page scan at registration, substr length),
    // else measured on the device - with the row count still there, so that both come back in ONE synchronisation
    bool need_measure = false;
    for (SortKey& sk : keys)
      if (sk.kind == 2) {
        const int ml = key_types[&sk - keys.data()].maxlen;
        if (ml > 0) sk.width = ml; else need_measure = true;
      }
    if (need_measure && in.rc().host > 0) {
      Blob::Dev dev0 = bb.upload(ar, s);
      int* dmax = ar.zeros<int>((int64_t)keys.size());
      for (size_t k = 0; k < keys.size(); ++k)
        if (keys[k].kind == 2) { launch_str_maxlen(dev0.host_progs[keys[k].prog], dev0.ctx, in.rc(), dmax + k, s); ++launches; }
      std::vector<int> hmax(keys.size());
      int hcount = 0;
      NDS_CUDA(cudaMemcpyAsync(hmax.data(), dmax, keys.size() * 4, cudaMemcpyDeviceToHost, s));
      if (in.n_dev && in.n_exact < 0) NDS_CUDA(cudaMemcpyAsync(&hcount, in.n_dev, 4, cudaMemcpyDeviceToHost, s));
      sync("sync:sort_widths");
      d2h += (int64_t)keys.size() * 4 + 4;
      if (in.n_dev && in.n_exact < 0) in.n_exact = std::min<int64_t>(hcount, in.n_max);
      for (size_t k = 0; k < keys.size(); ++k)
        if (keys[k].kind == 2) keys[k].width = std::max(hmax[k], 1);
    }
    for (SortKey& sk : keys) { sk.offset = B; B += 1 + sk.width; }
    const int64_t nb = in.rc().host;                   // exact when known, else the upper bound
    Batch out;
    int64_t m = (limit >= 0 && limit < nb) ? limit : nb;
    uint32_t* rows = ar.alloc_n<uint32_t>(std::max<int64_t>(m, 1));
    int* out_count = nullptr;
    if (nb > 0) {
      RowCount rc = in.rc();
      Blob::Dev dev = bb.upload(ar, s, keys.data(), (int64_t)keys.size() * sizeof(SortKey));
      if (topk_stream && limit >= 1 && limit <= kTopkMaxK && B <= kTopkMaxKeyBytes) {
        // ORDER BY ... LIMIT k: one launch, whatever the row count turns out to be on the device
        const int ctas = topk_stream_ctas(nb);
        const int Bw = (B + 7) / 8;
        uint64_t* scratch = ar.alloc_n<uint64_t>(topk_stream_scratch_words(ctas, Bw));
        const int groups = topk_stream_groups(ctas);
        uint64_t* fin_keys = ar.alloc_n<uint64_t>((int64_t)(ctas + groups) * limit * Bw);
        uint32_t* fin_rows = ar.alloc_n<uint32_t>((int64_t)(ctas + groups) * limit);
        int* fin_count = ar.alloc_n<int>(ctas + groups);
        int* meta = ar.zeros<int>(groups + 3);              // tickets (1 + groups), then the result count
        out_count = meta + groups + 2;
        launch_topk_stream(dev.progs, reinterpret_cast<const SortKey*>(dev.extra), (int)keys.size(), dev.ctx, rc, B, (int)limit, ctas, scratch,
                           fin_keys, fin_rows, fin_count, meta, rows, out_count, err, s);
        ++launches;
      } else {
      uint8_t* keybuf = (uint8_t*)ar.alloc(nb * B);
      int* active = ar.zeros<int>(B + 1);
      launch_encode_keys(dev.progs, reinterpret_cast<const SortKey*>(dev.extra), (int)keys.size(), dev.ctx, rc, keybuf, B, active, err, s);
      ++launches;
      out_count = ar.zeros<int>(1);
      bool done = false;
      const bool compact_ok = B <= kMaxCompactBytes;
      const bool topk = limit >= 1 && limit <= kRankSortMax / 8;
      if (compact_ok && (nb <= kRankSortMax || topk)) {
        const int W = (B + 7) / 8;
        uint64_t* ckeys = ar.alloc_n<uint64_t>(nb * W);
        int* wact = ar.zeros<int>(1);
        const bool direct = nb <= kRankSortMax;
        uint32_t* rank_tmp = direct ? ar.alloc_n<uint32_t>(nb) : nullptr;
        int* tile_done = direct ? ar.alloc_n<int>((nb + 127) / 128 + 1) : nullptr;
        launch_compact_keys(keybuf, B, active, rc, ckeys, W, wact, rank_tmp, tile_done, direct ? nb : 0, s);
        ++launches;
        if (direct) {
          // no host round trip at all: the row count stays on the device
          launch_rank_sort(ckeys, W, wact, nullptr, nullptr, 0, rc, nb, limit, rank_tmp, tile_done, rows, out_count, err + 2, s);
          ++launches;
          done = true;
        } else {
          uint32_t* cand = ar.alloc_n<uint32_t>(kRankSortMax);
          int* nc_dev = ar.zeros<int>(1);
          uint32_t* hist = ar.zeros<uint32_t>(kSelectPasses * 256);
          int l = 0;
          launch_topk_select(ckeys, W, rc, limit, cand, kRankSortMax, hist, nc_dev, s, &l);
          launches += l;
          if (speculate) {
            // no host round trip: the rank sort is sized for the candidate list's capacity; should the list overflow
            // (more than kRankSortMax rows tie on the selected prefix) err[2] is raised and the query runs again
            // without speculation
            uint32_t* rt = ar.zeros<uint32_t>(kRankSortMax);
            int* td = ar.zeros<int>(kRankSortMax / 128);
            launch_rank_sort(ckeys, W, wact, cand, nc_dev, kRankSortMax, rc, kRankSortMax, limit, rt, td, rows, out_count, err + 2, s);
            ++launches;
            done = true;
          } else {
            int hv[2] = {0, (int)nb};
            NDS_CUDA(cudaMemcpyAsync(&hv[0], nc_dev, 4, cudaMemcpyDeviceToHost, s));
            if (in.n_dev && in.n_exact < 0) NDS_CUDA(cudaMemcpyAsync(&hv[1], in.n_dev, 4, cudaMemcpyDeviceToHost, s));
            sync("sync:sort_topk");
            d2h += 8;
            if (in.n_exact < 0) in.n_exact = std::min<int64_t>(hv[1], in.n_max);
            if (hv[0] <= kRankSortMax) {
              const int64_t ncand = std::max(hv[0], 1);
              uint32_t* rt = ar.zeros<uint32_t>(ncand);
              int* td = ar.zeros<int>((ncand + 127) / 128);
              launch_rank_sort(ckeys, W, wact, cand, nc_dev, kRankSortMax, in.rc(), ncand, limit, rt, td, rows, out_count, nullptr, s);
              ++launches;
              done = true;
            }
          }
        }
      }
      if (!done) {
        // LSD radix sort over the key bytes that differ between rows (one small read-back decides the passes)
        const int64_t n = count(in);
        RowCount rn{nullptr, n};
        uint32_t* pa = ar.alloc_n<uint32_t>(std::max<int64_t>(n, 1));
        uint32_t* pb = ar.alloc_n<uint32_t>(std::max<int64_t>(n, 1));
        if (n > 0 && n <= kSmallSortMax) {
          int* which = ar.zeros<int>(2);
          launch_radix_sort_small(keybuf, B, active, rn, pa, pb, which, s);
          launch_take_perm(pa, pb, which, 0, rn, limit, rows, out_count, s);
          launches += 2;
        } else if (n > 0) {
          std::vector<int> hact(B);
          NDS_CUDA(cudaMemcpyAsync(hact.data(), active, B * sizeof(int), cudaMemcpyDeviceToHost, s));
          sync("sync:sort_active");
          d2h += B * 4;
          std::vector<int> passes;
          for (int b = B - 1; b >= 0; --b)
            if (hact[b]) passes.push_back(b);
          const int np = (int)passes.size();
          int* which = ar.zeros<int>(np + 2);
          uint32_t* scratch = ar.alloc_n<uint32_t>((int64_t)radix_sort_grid(n) * 256 + np + 16);
          int l = 0;
          launch_radix_sort(keybuf, B, passes.data(), np, rn, pa, pb, which, scratch, s, &l);
          launches += l;
          launch_take_perm(pa, pb, which, np, rn, limit, rows, out_count, s);
          ++launches;
        }
      }
      }
    }
    if (in.n_exact >= 0 || !in.n_dev) {
      const int64_t n = in.n_exact >= 0 ? in.n_exact : in.n_max;
      m = (limit >= 0 && limit < n) ? limit : n;
      out.n_max = m;
      out.n_exact = m;
    } else {
      out.n_max = m;              // upper bound; the exact min(limit, n) is on the device
      out.n_dev = out_count;
    }
    const int64_t n = nb;
    RowCount rm = out.rc();
    if (m > 0) {
      std::vector<Json> ces;
      ces.reserve(in.cols.size());
      for (size_t c = 0; c < in.cols.size(); ++c) ces.push_back(col_expr(in.cols[c].name));
      std::vector<std::pair<std::string, const Json*>> ex;
      for (size_t c = 0; c < in.cols.size(); ++c) ex.emplace_back(in.cols[c].name, &ces[c]);
      out.cols = eval_columns(ex, in, rows, rm);
    } else {
      for (size_t c = 0; c < in.cols.size(); ++c) { Column e0 = in.cols[c]; e0.rowmap = nullptr; e0.valid = nullptr; e0.data = ar.alloc(16); out.cols.push_back(e0); }
    }
    (void)n;
    end_stat_rows(st, row_bytes_of(bb) + 2 * (int64_t)B, in.rc());    // key columns read, encoded key written and read once
    return out;
  }

  Batch exec(const Json& node, std::vector<RuntimeFilter>& filters) {
    const std::string& op = node.str("op");
    if (op == "scan") return scan(node, filters);
    if (op == "filter") return filter(node, filters);
    if (op == "project") return project(node, filters);
    if (op == "join") return join(node, filters);
    if (op == "aggregate") return aggregate(node, filters);
    if (op == "rollup") return rollup(node, filters);
    if (op == "window") return window(node, filters);
    if (op == "rank") return rank(node, filters);
    if (op == "sort") return sort(node, filters);
    if (op == "exchange") return exchange(node, filters);
    fail("plan: unsupported plan node '" + op + "' (no CPU fallback exists)");
  }
};

// ------------------------------------------------------------------------------------------ result export
namespace {

struct ExportPriv {
  std::vector<void*> bufs;           // malloc'd buffers
  std::vector<std::string> strings;  // format / name storage
  std::vector<ArrowSchema*> schema_children;
  std::vector<ArrowArray*> array_children;
  std::vector<const void*> bufptrs;
  std::vector<ArrowSchema> schema_store;
  std::vector<ArrowArray> array_store;
  std::vector<std::vector<const void*>> child_bufptrs;
};

void release_schema(ArrowSchema* s) {
  if (!s || !s->release) return;
  for (int64_t i = 0; i < s->n_children; ++i)
    if (s->children[i] && s->children[i]->release) s->children[i]->release(s->children[i]);
  if (s->private_data) delete reinterpret_cast<ExportPriv*>(s->private_data);
  s->release = nullptr;
}
void release_schema_child(ArrowSchema* s) { if (s) s->release = nullptr; }
void release_array(ArrowArray* a) {
  if (!a || !a->release) return;
  for (int64_t i = 0; i < a->n_children; ++i)
    if (a->children[i] && a->children[i]->release) a->children[i]->release(a->children[i]);
  if (a->private_data) {
    ExportPriv* p = reinterpret_cast<ExportPriv*>(a->private_data);
    for (void* b : p->bufs) free(b);
    delete p;
  }
  a->release = nullptr;
}
void release_array_child(ArrowArray* a) { if (a) a->release = nullptr; }

}  // namespace

void Engine::execute(const Json& plan, ArrowSchema* out_schema, ArrowArray* out_array) {
  try {
    execute_once(plan, out_schema, out_array, true);
  } catch (const SpeculationFailed&) {
    ++spec_retries_;
    execute_once(plan, out_schema, out_array, false);
  }
}

void Engine::execute_once(const Json& plan, ArrowSchema* out_schema, ArrowArray* out_array, bool speculate) {
  const double t_exec0 = Exec::now_ms();
  NDS_CUDA(cudaSetDevice(device));
  grow_slab_if_wanted();
  const Json* options = plan.find("options");
  const Json& root = plan.has("root") ? plan.at("root") : plan;
  cudaEvent_t ev0 = get_event(), ev1 = get_event();
  struct EventGuard {          // a failed or retried query returns its timing events to the pool too
    Engine& en; cudaEvent_t a, b; bool armed = true;
    ~EventGuard() { if (armed) { en.put_event(a); en.put_event(b); } }
  } ev_guard{*this, ev0, ev1};
  NDS_CUDA(cudaEventRecord(ev0, stream_));
  bool ev1_recorded = false;
  std::vector<OpStat> op_stats;
  {
    Exec ex(*this, options);
    if (!speculate) ex.speculate = false;
    std::vector<RuntimeFilter> none;
    Batch res = ex.exec(root, none);
    ex.mark("plan_issued");
    const size_t nc = res.cols.size();
    // Small results (every NDS top-k) leave the device in one packed buffer, row count included: one copy, one
    // synchronisation.  Anything larger takes the column-at-a-time route below.
    constexpr int64_t kSmallRows = 8192, kSmallBytes = 2 << 20;
    int64_t pack_bytes = 64;
    bool small = res.rc().host <= kSmallRows && ex.ar.pin != nullptr;
    for (size_t c = 0; c < nc && small; ++c) {
      const Column& col = res.cols[c];
      const int64_t nb = std::max<int64_t>(res.rc().host, 1);
      pack_bytes += ((nb + 31) / 32) * 4 + 64;
      if (col.lt.kind == LType::STR) {
        if (col.lt.maxlen <= 0) small = false;
        pack_bytes += (nb + 1) * 4 + nb * (int64_t)col.lt.maxlen + 128;
      } else {
        pack_bytes += nb * 8 + 64;
      }
      if (pack_bytes > kSmallBytes) small = false;
    }
    int64_t n = small ? -1 : ex.count(res);
    // flatten every result column that still goes through a row map (one launch for all of them)
    std::vector<Column> flat = res.cols;
    const RowCount rn = small ? res.rc() : RowCount{nullptr, n};
    if (rn.host > 0) {
      std::vector<Json> ces;
      std::vector<size_t> which;
      for (size_t c = 0; c < nc; ++c)
        if (res.cols[c].rowmap) { which.push_back(c); ces.push_back(Exec::col_expr(res.cols[c].name)); }
      std::vector<std::pair<std::string, const Json*>> exs;
      for (size_t k = 0; k < which.size(); ++k) exs.emplace_back(res.cols[which[k]].name, &ces[k]);
      std::vector<Column> fc = ex.eval_columns(exs, res, nullptr, rn);
      for (size_t k = 0; k < which.size(); ++k) flat[which[k]] = fc[k];
    }
    // ---- schema
    ExportPriv* sp = new ExportPriv();
    ExportPriv* apv = new ExportPriv();
    sp->schema_store.resize(nc);
    apv->array_store.resize(nc);
    apv->child_bufptrs.resize(nc);
    sp->strings.reserve(2 * nc + 2);
    for (size_t c = 0; c < nc; ++c) {
      const Column& col = flat[c];
      std::string fmt;
      switch (col.lt.kind) {
        case LType::INT: case LType::BOOL: fmt = col.lt.dtype() == DT_I32 ? "i" : "l"; break;
        case LType::DATE: fmt = "tdD"; break;
        case LType::DECIMAL: fmt = "d:" + std::to_string(col.lt.precision) + "," + std::to_string(col.lt.scale); break;
        case LType::F64: fmt = "g"; break;
        case LType::STR: fmt = "u"; break;
      }
      sp->strings.push_back(fmt);
      const char* fmt_c = sp->strings.back().c_str();
      sp->strings.push_back(col.name);
      const char* name_c = sp->strings.back().c_str();
      ArrowSchema& cs = sp->schema_store[c];
      memset(&cs, 0, sizeof cs);
      cs.format = fmt_c;
      cs.name = name_c;
      cs.flags = 2;  // nullable
      cs.release = release_schema_child;
      sp->schema_children.push_back(&cs);
    }
    auto drop_export = [&]() {
      for (void* b : apv->bufs) free(b);
      delete apv;
      delete sp;
    };
    // ---- device -> host.  Per column afterwards: hvalid (or null), hdata (fixed width values or StrViews),
    // and for strings either (offs32, chars) already final (small route) or pending a device gather (large route).
    std::vector<uint8_t*> hvalid_v(nc, nullptr), hdata_v(nc, nullptr), hchars_v(nc, nullptr);
    std::vector<int32_t*> hoffs_v(nc, nullptr);
    int h_err[4] = {0, 0, 0, 0};
    try {
      if (small) {
        const int64_t nb = std::max<int64_t>(rn.host, 1);
        const int64_t vbytes_b = ((nb + 31) / 32) * 4;
        std::vector<ExportSeg> segs;
        struct Place { int64_t valid = -1, data = -1, offs = -1, chars = -1; };
        std::vector<Place> place(nc);
        int64_t off = 0;
        auto take = [&](int64_t bytes) { int64_t o = off; off = align_up(off + bytes, 16); return o; };
        const int64_t off_count = take(4), off_err = take(16);
        segs.push_back(ExportSeg{nullptr, nullptr, nullptr, off_count, 4, 0, 0, 2, 0});
        for (size_t c = 0; c < nc; ++c) {
          const Column& col = flat[c];
          if (col.valid) {
            place[c].valid = take(vbytes_b);
            segs.push_back(ExportSeg{col.valid, nullptr, nullptr, place[c].valid, vbytes_b, 0, 0, 0, 0});
          }
          if (col.lt.kind == LType::STR) {
            const int64_t cap = nb * (int64_t)col.lt.maxlen;
            place[c].offs = take((nb + 1) * 4);
            place[c].chars = take(cap);
            segs.push_back(ExportSeg{col.data, col.valid, col.strbase, place[c].offs, 0, place[c].chars, cap, 1, 0});
          } else {
            const int64_t w = col.lt.dtype() == DT_I32 ? 4 : 8;
            place[c].data = take(nb * w);
            segs.push_back(ExportSeg{col.data, nullptr, nullptr, place[c].data, nb * w, 0, 0, 0, 0});
          }
        }
        const int64_t total = off;
        uint8_t* hpack = ex.ar.pin_take(total);
        std::vector<uint8_t> pageable;
        if (!hpack) { pageable.resize((size_t)total); hpack = pageable.data(); }
        uint8_t* dpack = (uint8_t*)ex.ar.alloc(total);
        ExportSeg* dsegs = ex.ar.alloc_n<ExportSeg>((int64_t)segs.size());
        ex.ar.upload(dsegs, segs.data(), (int64_t)segs.size() * (int64_t)sizeof(ExportSeg));
        launch_export_pack(dsegs, (int)segs.size(), rn, dpack, ex.err, stream_);
        ++ex.launches;
        NDS_CUDA(cudaMemcpyAsync(dpack + off_err, ex.err, sizeof h_err, cudaMemcpyDeviceToDevice, stream_));
        NDS_CUDA(cudaMemcpyAsync(hpack, dpack, (size_t)total, cudaMemcpyDeviceToHost, stream_));
        NDS_CUDA(cudaEventRecord(ev1, stream_));       // the query's last device operation: no second wait for the timing event
        ev1_recorded = true;
        ex.sync("sync:export");
        ex.d2h += total;
        memcpy(h_err, hpack + off_err, sizeof h_err);
        ex.check_err_values(h_err);
        n = std::min<int64_t>(*reinterpret_cast<const int*>(hpack + off_count), rn.host);
        if (n < 0) n = 0;
        const int64_t vbytes = ((n + 31) / 32) * 4;
        for (size_t c = 0; c < nc; ++c) {
          const Column& col = flat[c];
          if (place[c].valid >= 0 && n > 0) {
            hvalid_v[c] = (uint8_t*)malloc(vbytes + 8);
            apv->bufs.push_back(hvalid_v[c]);
            memcpy(hvalid_v[c], hpack + place[c].valid, (size_t)vbytes);
          }
          if (col.lt.kind == LType::STR) {
            const int32_t* po = reinterpret_cast<const int32_t*>(hpack + place[c].offs);
            hoffs_v[c] = (int32_t*)malloc((n + 1) * 4 + 8);
            apv->bufs.push_back(hoffs_v[c]);
            if (n > 0) memcpy(hoffs_v[c], po, (size_t)(n + 1) * 4); else hoffs_v[c][0] = 0;
            const int64_t cb = n > 0 ? po[n] : 0;
            hchars_v[c] = (uint8_t*)malloc(std::max<int64_t>(cb, 8));
            apv->bufs.push_back(hchars_v[c]);
            if (cb > 0) memcpy(hchars_v[c], hpack + place[c].chars, (size_t)cb);
          } else {
            const int64_t w = col.lt.dtype() == DT_I32 ? 4 : 8;
            hdata_v[c] = (uint8_t*)malloc(std::max<int64_t>(n * w, 8));
            apv->bufs.push_back(hdata_v[c]);
            if (n > 0) memcpy(hdata_v[c], hpack + place[c].data, (size_t)(n * w));
          }
        }
      } else {
        NDS_CUDA(cudaMemcpyAsync(h_err, ex.err, sizeof h_err, cudaMemcpyDeviceToHost, stream_));   // checked after the export sync
        for (size_t c = 0; c < nc; ++c) {
          const Column& col = flat[c];
          const int64_t vbytes = ((n + 31) / 32) * 4;
          if (col.valid && n > 0) {
            hvalid_v[c] = (uint8_t*)malloc(vbytes + 8);
            apv->bufs.push_back(hvalid_v[c]);
            NDS_CUDA(cudaMemcpyAsync(hvalid_v[c], col.valid, vbytes, cudaMemcpyDeviceToHost, stream_));
            ex.d2h += vbytes;
          }
          const int w = col.lt.dtype() == DT_I32 ? 4 : 8;
          hdata_v[c] = (uint8_t*)malloc(std::max<int64_t>(n * w, 8));
          apv->bufs.push_back(hdata_v[c]);
          if (n > 0) { NDS_CUDA(cudaMemcpyAsync(hdata_v[c], col.data, n * w, cudaMemcpyDeviceToHost, stream_)); ex.d2h += n * w; }
        }
        ex.sync("sync:export");
        ex.check_err_values(h_err);
        // string characters are gathered on the device for all string columns, one more sync
        bool pending = false;
        for (size_t c = 0; c < nc; ++c) {
          const Column& col = flat[c];
          if (col.lt.kind != LType::STR) continue;
          const uint8_t* hvalid = hvalid_v[c];
          const StrView* views = reinterpret_cast<const StrView*>(hdata_v[c]);
          int32_t* offs = (int32_t*)malloc((n + 1) * 4 + 8);
          apv->bufs.push_back(offs);
          hoffs_v[c] = offs;
          std::vector<int64_t> offs64(n + 1, 0);
          for (int64_t i = 0; i < n; ++i) {
            bool ok = !hvalid || ((hvalid[i >> 3] >> (i & 7)) & 1);
            offs64[i + 1] = offs64[i] + (ok ? (int64_t)(views[i] >> 32) : 0);
          }
          if (offs64[n] > 0x7FFFFFFFll) fail("result string column exceeds 2 GiB");
          for (int64_t i = 0; i <= n; ++i) offs[i] = (int32_t)offs64[i];
          uint8_t* chars = (uint8_t*)malloc(std::max<int64_t>(offs64[n], 8));
          apv->bufs.push_back(chars);
          hchars_v[c] = chars;
          if (offs64[n] > 0) {
            Arena& ar = ex.ar;
            int64_t* doffs = ar.alloc_n<int64_t>(n + 1);
            uint8_t* dchars = (uint8_t*)ar.alloc(offs64[n]);
            ar.upload(doffs, offs64.data(), (n + 1) * 8);
            launch_copy_str_bytes(reinterpret_cast<const StrView*>(col.data), col.strbase, doffs, n, dchars, stream_);
            ++ex.launches;
            NDS_CUDA(cudaMemcpyAsync(chars, dchars, offs64[n], cudaMemcpyDeviceToHost, stream_));
            ex.d2h += offs64[n];
            pending = true;
          }
        }
        if (pending) ex.sync("sync:export_strings");
      }
    } catch (...) {
      drop_export();
      throw;
    }
    // ---- Arrow arrays over the host buffers
    for (size_t c = 0; c < nc; ++c) {
      const Column& col = flat[c];
      const DType dt = col.lt.dtype();
      uint8_t* hvalid = hvalid_v[c];
      uint8_t* hdata = hdata_v[c];
      int64_t null_count = 0;
      if (hvalid) {
        // clear padding bits, count nulls
        int64_t valid_cnt = 0;
        for (int64_t i = 0; i < n; ++i) valid_cnt += (hvalid[i >> 3] >> (i & 7)) & 1;
        null_count = n - valid_cnt;
        if (n & 7) hvalid[n >> 3] &= (uint8_t)((1u << (n & 7)) - 1u);
      }
      std::vector<const void*>& bp = apv->child_bufptrs[c];
      ArrowArray& ca = apv->array_store[c];
      memset(&ca, 0, sizeof ca);
      ca.length = n;
      ca.null_count = null_count;
      ca.release = release_array_child;
      if (col.lt.kind == LType::DECIMAL) {
        int64_t* wide = (int64_t*)malloc(std::max<int64_t>(n * 16, 16));
        apv->bufs.push_back(wide);
        for (int64_t i = 0; i < n; ++i) {
          int64_t v = dt == DT_I32 ? (int64_t) reinterpret_cast<int32_t*>(hdata)[i] : reinterpret_cast<int64_t*>(hdata)[i];
          wide[2 * i] = v;
          wide[2 * i + 1] = v >> 63;
        }
        bp = {hvalid, wide};
      } else if (col.lt.kind == LType::STR) {
        bp = {hvalid, hoffs_v[c], hchars_v[c]};
      } else if (col.lt.kind == LType::DATE && dt != DT_I32) {
        // a date that went through if() / min / max is held as int64 on the device; Arrow's date32 ("tdD") is 4 bytes wide
        int32_t* narrow = (int32_t*)malloc(std::max<int64_t>(n * 4, 8));
        apv->bufs.push_back(narrow);
        for (int64_t i = 0; i < n; ++i) narrow[i] = (int32_t) reinterpret_cast<int64_t*>(hdata)[i];
        bp = {hvalid, narrow};
      } else {
        bp = {hvalid, hdata};
      }
      ca.n_buffers = (int64_t)bp.size();
      ca.buffers = bp.data();
      apv->array_children.push_back(&ca);
    }
    memset(out_schema, 0, sizeof *out_schema);
    out_schema->format = "+s";
    out_schema->name = "";
    out_schema->n_children = (int64_t)nc;
    out_schema->children = sp->schema_children.data();
    out_schema->release = release_schema;
    out_schema->private_data = sp;
    memset(out_array, 0, sizeof *out_array);
    apv->bufptrs = {nullptr};
    out_array->length = n;
    out_array->n_buffers = 1;
    out_array->buffers = apv->bufptrs.data();
    out_array->n_children = (int64_t)nc;
    out_array->children = apv->array_children.data();
    out_array->release = release_array;
    out_array->private_data = apv;

    if (!ev1_recorded) {
      NDS_CUDA(cudaEventRecord(ev1, stream_));
      ex.sync("sync:final");
    }
    if (ex.profile) ex.resolve_stat_rows();
    trace_ = ex.trace_out;
    launches_ = ex.launches;
    h2d_bytes_ = ex.h2d;
    d2h_bytes_ = ex.d2h;
    op_stats = std::move(ex.stats);
  }
  float ms = 0;
  cudaEventElapsedTime(&ms, ev0, ev1);
  exec_ms_ = ms;
  ev_guard.armed = false;
  for (OpStat& st : op_stats) {
    if (st.e0 && st.e1) {
      float t = 0;
      if (cudaEventElapsedTime(&t, st.e0, st.e1) == cudaSuccess) st.ms = t;
      else cudaGetLastError();
      if (cudaEventElapsedTime(&t, ev0, st.e0) == cudaSuccess) st.start_ms = t;
      else cudaGetLastError();
    }
    if (st.e0) put_event(st.e0);
    if (st.e1) put_event(st.e1);
    st.e0 = st.e1 = nullptr;
  }
  put_event(ev0);
  put_event(ev1);
  stats_ = std::move(op_stats);
  host_ms_ = Exec::now_ms() - t_exec0;
}

std::string Engine::stats_json() const {
  std::string s = "{\"launches\":" + std::to_string(launches_) + ",\"device_ms\":" + std::to_string(exec_ms_) + ",\"host_ms\":" + std::to_string(host_ms_) +
                  ",\"h2d_bytes\":" + std::to_string(h2d_bytes_) + ",\"d2h_bytes\":" + std::to_string(d2h_bytes_) +
                  ",\"speculation_retries\":" + std::to_string(spec_retries_) + ",\"trace\":[" + trace_ + "],\"ops\":[";
  for (size_t i = 0; i < stats_.size(); ++i) {
    if (i) s += ",";
    s += "{\"op\":\"" + json_escape(stats_[i].op) + "\",\"ms\":" + std::to_string(stats_[i].ms) + ",\"bytes\":" +
         std::to_string(stats_[i].bytes) + ",\"rows\":" + std::to_string(stats_[i].rows) + ",\"host_ms\":" +
         std::to_string(stats_[i].host_ms) + ",\"start_ms\":" + std::to_string(stats_[i].start_ms) + "}";
  }
  s += "]}";
  return s;
}

Engine::Engine(const Json& cfg) {
  device = (int)cfg.int_or("device", 0);
  resident_ = cfg.bool_or("resident", true);
  keep_host_ = cfg.bool_or("keep_host", false);
  rank = (int)cfg.int_or("rank", 0);
  world = (int)cfg.int_or("world", 1);
  xwin_mb_ = std::max<int64_t>(cfg.int_or("exchange_window_mb", 128), 1);
  if (world < 1 || rank < 0 || rank >= world) fail("engine config: bad rank/world");
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev == 0)
    fail(std::string("no CUDA device is available (") + cudaGetErrorString(ce) + "); the NDS B200 engine has no CPU path");
  if (device >= ndev) fail("engine config: device " + std::to_string(device) + " does not exist");
  NDS_CUDA(cudaSetDevice(device));
  // Lanes: the build sides of a join chain, the fact scan's key columns, and the def-level kernel over the payload columns
  // (which only the gather after the probe pass needs) each get a stream.  Stream priorities (build sides first, payload
  // last) were measured at SF100 and changed nothing: query7 is bound by the sum of the kernels' SM time, so what the build
  // sides gain the gather loses waiting for the deprioritised payload kernel.  All lanes share one priority.
  NDS_CUDA(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
  for (int k = 0; k <= kLowLane; ++k) {
    cudaStream_t st;
    NDS_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    side_streams_.push_back(st);
  }
  cudaMemPool_t pool;
  NDS_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
  uint64_t thr = UINT64_MAX;
  NDS_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
  slab_cap_ = cfg.int_or("slab_mb", 512) << 20;
  if (slab_cap_ > 0 && cudaMalloc(reinterpret_cast<void**>(&slab_), (size_t)slab_cap_) != cudaSuccess) {
    cudaGetLastError();
    slab_ = nullptr;
    slab_cap_ = 0;
  }
  pin_cap_ = 8 << 20;
  if (cudaHostAlloc(reinterpret_cast<void**>(&pin_), (size_t)pin_cap_, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    pin_ = nullptr;
    pin_cap_ = 0;
  }
}

// The previous query did not fit the per-query slab and fell back to cudaMallocAsync (tens of microseconds per column
// at SF100): make the slab large enough for a query of that size, within a quarter of the memory that is free now.
void Engine::grow_slab_if_wanted() {
  if (slab_wanted_ <= slab_cap_) return;
  size_t free_b = 0, total_b = 0;
  NDS_CUDA(cudaStreamSynchronize(stream_));
  NDS_CUDA(cudaMemGetInfo(&free_b, &total_b));
  const int64_t limit = slab_cap_ + (int64_t)(free_b / 4);
  const int64_t want = std::min<int64_t>(align_up(slab_wanted_, 64 << 20), limit);
  slab_wanted_ = 0;
  if (want <= slab_cap_) return;
  uint8_t* fresh = nullptr;
  if (cudaMalloc(reinterpret_cast<void**>(&fresh), (size_t)want) != cudaSuccess) { cudaGetLastError(); return; }
  if (slab_) cudaFree(slab_);
  slab_ = fresh;
  slab_cap_ = want;
  zero_dirty_ = -1;
}

cudaEvent_t Engine::get_event() {
  if (!event_pool_.empty()) {
    cudaEvent_t e = event_pool_.back();
    event_pool_.pop_back();
    return e;
  }
  cudaEvent_t e;
  NDS_CUDA(cudaEventCreate(&e));
  return e;
}
void Engine::put_event(cudaEvent_t e) { event_pool_.push_back(e); }

// ---- communicator for the exchange operator (one rank per process / GPU, NCCL over NVLink)
std::string Engine::comm_unique_id() {
  NcclUniqueId id;
  memset(&id, 0, sizeof id);
  nccl_check(nccl_api().GetUniqueId(&id), "ncclGetUniqueId");
  return std::string(id.internal, sizeof id.internal);
}

void Engine::comm_init(const std::string& id_bytes, int rank_in, int world_in) {
  if (id_bytes.size() != sizeof(NcclUniqueId)) fail("comm_init: the unique id must be 128 bytes");
  if (world_in < 1 || rank_in < 0 || rank_in >= world_in) fail("comm_init: bad rank/world");
  NDS_CUDA(cudaSetDevice(device));
  if (comm_) { nccl_api().CommDestroy(comm_); comm_ = nullptr; }
  comm_rank_ = rank_in;
  comm_world_ = world_in;
  if (world_in == 1) return;
  NcclUniqueId id;
  memcpy(id.internal, id_bytes.data(), sizeof id.internal);
  nccl_check(nccl_api().CommInitRank(&comm_, world_in, id, rank_in), "ncclCommInitRank");
  // ---- windows of the peer-push exchange: allocate mine, all-gather the IPC handles, map the peers'.  Any failure
  // (no peer access, a refused mapping) leaves the NCCL exchange in charge - on every rank, or on none.
  close_windows();
  xpeer_ok_ = false;
  if (world_in > kPeerMaxWorld) return;
  ensure_window();
  cudaIpcMemHandle_t mine;
  int ok = cudaIpcGetMemHandle(&mine, xwin_) == cudaSuccess ? 1 : 0;
  if (!ok) cudaGetLastError();
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
  const size_t rec = 64 + 8;
  std::vector<uint8_t> host(rec * world_in, 0);
  memcpy(host.data() + rec * rank_in, &mine, 64);
  memcpy(host.data() + rec * rank_in + 64, &ok, sizeof ok);
  uint8_t* d_all = nullptr;
  NDS_CUDA(cudaMalloc(&d_all, rec * world_in));
  NDS_CUDA(cudaMemcpy(d_all, host.data(), rec * world_in, cudaMemcpyHostToDevice));
  nccl_check(nccl_api().AllGather(d_all + rec * rank_in, d_all, rec, kNcclChar, comm_, stream_), "ncclAllGather(window handles)");
  NDS_CUDA(cudaStreamSynchronize(stream_));
  NDS_CUDA(cudaMemcpy(host.data(), d_all, rec * world_in, cudaMemcpyDeviceToHost));
  xpeers_.assign(world_in, nullptr);
  xpeers_[rank_in] = xwin_;
  int all_ok = 1;
  for (int p = 0; p < world_in; ++p) {
    int pok = 0;
    memcpy(&pok, host.data() + rec * p + 64, sizeof pok);
    if (!pok) all_ok = 0;
  }
  if (all_ok)
    for (int p = 0; p < world_in && all_ok; ++p) {
      if (p == rank_in) continue;
      cudaIpcMemHandle_t h;
      memcpy(&h, host.data() + rec * p, 64);
      void* ptr = nullptr;
      if (cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); all_ok = 0; break; }
      xpeers_[p] = (uint8_t*)ptr;
    }
  // second round: a rank that could not map a window must switch the others off too
  int* d_flag = reinterpret_cast<int*>(d_all);
  std::vector<int> flags(world_in, 0);
  flags[rank_in] = all_ok;
  NDS_CUDA(cudaMemcpy(d_flag, flags.data(), sizeof(int) * world_in, cudaMemcpyHostToDevice));
  nccl_check(nccl_api().AllGather(d_flag + rank_in, d_flag, sizeof(int), kNcclChar, comm_, stream_), "ncclAllGather(window status)");
  NDS_CUDA(cudaStreamSynchronize(stream_));
  NDS_CUDA(cudaMemcpy(flags.data(), d_flag, sizeof(int) * world_in, cudaMemcpyDeviceToHost));
  cudaFree(d_all);
  xpeer_ok_ = true;
  for (int f : flags) xpeer_ok_ = xpeer_ok_ && f != 0;
  if (!xpeer_ok_) {
    for (int p = 0; p < world_in; ++p)
      if (p != rank_in && xpeers_[p]) { cudaIpcCloseMemHandle(xpeers_[p]); xpeers_[p] = nullptr; }
  }
}

// The local window: control block (arrival counters, zeroed once) + two halves.
void Engine::ensure_window() {
  if (xwin_) return;
  xwin_half_bytes_ = xwin_mb_ << 20;
  NDS_CUDA(cudaMalloc(reinterpret_cast<void**>(&xwin_), (size_t)(kPeerCtrlBytes + 2 * xwin_half_bytes_)));
  NDS_CUDA(cudaMemset(xwin_, 0, (size_t)kPeerCtrlBytes));
  xwin_uses_[0] = xwin_uses_[1] = 0;
  xwin_seq_ = 0;
}

void Engine::close_windows() {
  for (size_t p = 0; p < xpeers_.size(); ++p)
    if (xpeers_[p] && xpeers_[p] != xwin_) cudaIpcCloseMemHandle(xpeers_[p]);
  xpeers_.clear();
  if (xwin_) { cudaFree(xwin_); xwin_ = nullptr; }
  xpeer_ok_ = false;
}

Engine::~Engine() {
  cudaSetDevice(device);
  close_windows();
  if (comm_) { nccl_api().CommDestroy(comm_); comm_ = nullptr; }
  for (cudaEvent_t e : event_pool_) cudaEventDestroy(e);
  std::vector<std::string> names;
  for (auto& kv : tables_) names.push_back(kv.first);
  for (auto& n : names) drop_table(n);
  for (cudaStream_t st : side_streams_) cudaStreamDestroy(st);
  if (stream_) cudaStreamDestroy(stream_);
  if (pin_) cudaFreeHost(pin_);
  if (slab_) cudaFree(slab_);
}

}  // namespace nds
