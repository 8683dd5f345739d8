// Parquet page decode: RLE/bit-packed hybrid definition levels -> Arrow validity bitmap,
// RLE_DICTIONARY indices / PLAIN values -> dense device columns.  One CTA (512 threads) per data page,
// or several CTAs per page when a launch has too few pages to fill the GPU.
//
// Reference boundary: the scan the reference triggers with
//   spark_session.read.format(input_format).load(table_path)   (/root/reference/nds/nds_power.py:97-100)
// and executes inside Spark's vectorised Parquet reader (not vendored).  Format facts follow the
// public Parquet "Encodings" spec: hybrid run header = varint, LSB 1 -> bit-packed run of
// (header>>1) groups of 8 values, LSB 0 -> RLE run of (header>>1) copies of a ceil(bw/8)-byte value;
// data page v1 def levels carry a 4-byte little-endian length prefix; dictionary-encoded values
// start with a 1-byte bit width.
//
// Structure of one page:
//   0. one warp starts chasing the dictionary-index run headers (32 predicted header positions per
//      round, verified with a ballot) while
//   1. the def-level bytes are staged in shared memory by a bulk async copy (cp.async.bulk + mbarrier,
//      SASS UBLKCP);
//   2. the def-level run headers - a sequential format - are resolved in parallel: every byte position is
//      parsed speculatively as a header by one thread per 32-byte chunk (phase A: where does a walk that
//      enters the chunk at this offset leave it, and how many values does it emit), one thread hops chunk
//      to chunk (phase B), the entered chunks expand their runs into the validity bitmap (phase C);
//   3. a block scan of popcounts gives rank(row) = index into the non-null value stream;
//   4. small index streams (<= 16 KB, where thousands of short runs can hide) are staged and resolved like
//      the def levels;
//   5. each warp expands a contiguous block of 32-row tiles with a warp-uniform run cursor, four tiles per
//      iteration with the loads of all four issued before the first use: bits -> dictionary -> store.
//
// Roofline: HBM.  Algorithmic bytes per page = payload bytes read + width*num_values written
// + num_values/8 validity written.
#include "decode_common.cuh"
#include "ops.cuh"

namespace nds {

using namespace dec;
namespace {
constexpr int kT = kDecodeThreads;
}  // namespace

#define NDS_CKPT(j) do { if (dbg && threadIdx.x == 0) dbg[(int64_t)blockIdx.x * 8 + (j)] = clock64(); } while (0)

__global__ void __launch_bounds__(kT, 2)
decode_pages_kernel(const PageTask* __restrict__ tasks, const int* __restrict__ task_idx, int nslices, int* __restrict__ err,
                    long long* __restrict__ dbg) {
  __shared__ uint32_t s_valid[kWords];
  __shared__ uint32_t s_rank[kWords];
  __shared__ uint32_t ent_pos[kEntCap + 1];
  __shared__ uint32_t ent_k[kEntCap + 1];
  __shared__ uint32_t early_k[kEarlyCap + 1];
  __shared__ uint32_t early_info[kEarlyCap];
  __shared__ __align__(16) uint8_t s_stage[kStage + 96];
  __shared__ uint32_t s_hop[kHopWords];                       // (exit position << 16) | values emitted
  __shared__ uint32_t s_long[64][2];                        // long RLE runs of 1s, filled cooperatively
  __shared__ __align__(8) uint64_t s_mbar;
  __shared__ __align__(8) uint64_t s_mbar_dict;
  extern __shared__ __align__(16) uint8_t s_dict[];         // kDictStage bytes: the page's dictionary
  __shared__ int s_nlong;
  __shared__ int s_nent;
  __shared__ int s_done;
  __shared__ int s_early_n;                                 // -1: early chase not done / did not finish
  __shared__ uint32_t s_warp_tot[kT / 32];
  __shared__ Walker s_walk;

  NDS_CKPT(0);
  // a page may be shared by `nslices` CTAs when a launch has too few pages to fill the GPU: every slice resolves
  // the page's run headers, then expands only its own share of the rows; slice 0 writes the validity bitmap
  const PageTask t = tasks[task_idx ? task_idx[blockIdx.x / nslices] : (int)(blockIdx.x / nslices)];
  const int slice = blockIdx.x % nslices;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t n = t.num_values;
  const uint32_t nw = (n + 31) >> 5;
  const bool has_def = t.valid != nullptr && t.def_len > 0;
  const uint32_t values_off = has_def ? 4 + t.def_len : 0;
  const bool def_staged = has_def && t.def_len <= (uint32_t)kStage;
  const bool is_dict = t.enc == 1;
  const uint32_t bw = t.bw;
  const uint32_t VL = is_dict ? t.size - (values_off + 1) : 0;     // index stream bytes
  const uint8_t* vstream = t.src + values_off + 1;                 // index stream in global memory
  const bool v_small = is_dict && bw > 0 && VL <= (uint32_t)kStage;   // may be staged if it has many short runs
  uint32_t nn = n;          // non-null count
  uint32_t mbar_phase = 0;

  const uint32_t dict_bytes = is_dict ? t.dict_count * (uint32_t)t.width : 0u;
  const bool dict_staged = is_dict && dict_bytes > 0 && dict_bytes <= (uint32_t)kDictStage;
  if (tid == 0) {
    s_early_n = -1;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_mbar)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_mbar_dict)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // the dictionary streams into shared memory while the def levels are resolved: the per-value gathers then hit
  // shared memory instead of 32 different L2 sectors per warp
  if (dict_staged && tid == 0) bulk_load_issue(s_dict, t.dict, (dict_bytes + 15u) & ~15u, &s_mbar_dict);

  // ---- 1. def-level bytes -> shared memory (async) ---------------------------------------------------------------
  uint32_t def_sd = 0;
  if (def_staged) {
    const uint8_t* g0 = t.src + 4;
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(g0) & ~uintptr_t(15);
    const uintptr_t a1 = (reinterpret_cast<uintptr_t>(g0) + t.def_len + 15) & ~uintptr_t(15);
    if (tid == 0) bulk_load_issue(s_stage, reinterpret_cast<const void*>(a0), (uint32_t)(a1 - a0), &s_mbar);
    def_sd = smem_u32(s_stage) + (uint32_t)(reinterpret_cast<uintptr_t>(g0) - a0);
  }
  // ---- 0. meanwhile the last warp chases the index stream's headers (large streams = long bit-packed runs) ----------
  if (is_dict && bw > 0 && bw <= 32 && warp == kT / 32 - 1) {
    Walker w;
    w.pos = 0;
    w.end = VL;
    w.k = 0;
    int nr = walk_runs_warp(vstream, w, bw, n, kEarlyCap, 4, early_k, early_info);
    __syncwarp();
    if (lane == 0 && (w.pos >= VL || w.k >= n)) s_early_n = nr;     // finished: the runs cover every value
  }
  // the chase warp stays out of the def-level phases (they synchronise on a named barrier of the other warps)
  const bool chasing = is_dict && bw > 0 && bw <= 32;
  const int defN = chasing ? kT - 32 : kT;            // threads taking part in the def-level phases
  const int defBar = chasing ? 1 : 0;
  const bool in_def = tid < defN;
  if (in_def)
    for (uint32_t w = tid; w < nw; w += defN) s_valid[w] = 0;

  if (has_def) {
    const uint32_t L = t.def_len;
    if (def_staged) {
      if (in_def) {
        mbar_wait(&s_mbar, mbar_phase);
        NDS_CKPT(1);
        if (tid == 0) { s_walk.pos = 0; s_walk.k = 0; s_walk.end = L; }
        bar_sync(defBar, defN);
        // ---- 2. resolve the def-level run headers in parallel, segment by segment --------------------------------
        for (;;) {
          if (s_walk.pos >= L || s_walk.k >= n) break;
          if (tid == 0) { s_nent = 0; s_nlong = 0; }
          resolve_segment(def_sd, L, 1, n, &s_walk, s_hop, ent_pos, ent_k, &s_nent, tid, defBar, defN, dbg);
          // phase C: the entered chunks list their runs (count, scan, write - the hop table is free again and holds
          // the list), then every run is expanded into the bitmap by its own thread; long runs of valid rows (a
          // column without NULLs is one run) go to the whole CTA.
          const int ne = s_nent;
          const uint32_t seg_end = s_walk.pos;
          uint32_t my_runs = 0;
          if (tid < ne) {
            const uint32_t pe = tid + 1 < ne ? ent_pos[tid + 1] : seg_end;
            uint32_t p = ent_pos[tid], k = ent_k[tid];
            while (p < pe && k < n) {
              uint32_t cnt, info;
              p += parse_run_s(def_sd, p, L, 1, cnt, info);
              k += cnt;
              ++my_runs;
            }
          }
          // exclusive scan of my_runs over the (<= kDefSeg / kChunk = 64) entry threads: warps 0 and 1
          uint32_t run0 = 0;
          if (tid < 64) {
            uint32_t incl = my_runs;
            for (int d = 1; d < 32; d <<= 1) {
              uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
              if (lane >= d) incl += v;
            }
            if (lane == 31) s_warp_tot[warp] = incl;
            run0 = incl - my_runs;
          }
          bar_sync(defBar, defN);
          const uint32_t nruns = min(s_warp_tot[0] + s_warp_tot[1], (uint32_t)kDefSeg / 2);
          if (tid < ne) {
            if (warp == 1) run0 += s_warp_tot[0];
            const uint32_t pe = tid + 1 < ne ? ent_pos[tid + 1] : seg_end;
            uint32_t p = ent_pos[tid], k = ent_k[tid], r = run0;
            while (p < pe && k < n) {
              uint32_t cnt, info;
              p += parse_run_s(def_sd, p, L, 1, cnt, info);
              if (r < (uint32_t)kDefSeg / 2) {
                s_hop[2 * r] = k;
                s_hop[2 * r + 1] = (min(cnt, 0xFFFFu) << 16) | (info & 0xFFFFu);
              } else {            // more runs than the list holds (degenerate zero-length runs): expand in place
                const uint32_t k1 = min(k + cnt, n);
                if (info & 1) copy_bits(s_valid, def_sd + (info >> 1), nullptr, k, k1);
                else if (info >> 1) set_bits(s_valid, k, k1);
              }
              k += cnt;
              ++r;
            }
          }
          bar_sync(defBar, defN);
          for (uint32_t r = tid; r < nruns; r += defN) {
            const uint32_t k = s_hop[2 * r], ci = s_hop[2 * r + 1];
            const uint32_t info = ci & 0xFFFFu;
            const uint32_t k1 = min(k + (ci >> 16), n);
            if (info & 1) copy_bits(s_valid, def_sd + (info >> 1), nullptr, k, k1);
            else if (info >> 1) {
              if (k1 - k >= 256) {
                int j = atomicAdd(&s_nlong, 1);
                if (j < 64) { s_long[j][0] = k; s_long[j][1] = k1; }
                else set_bits(s_valid, k, k1);
              } else set_bits(s_valid, k, k1);
            }
          }
          bar_sync(defBar, defN);
          const int nl = min(s_nlong, 64);
          for (int j = 0; j < nl; ++j) {
            const uint32_t b0 = s_long[j][0], b1 = s_long[j][1];
            for (uint32_t w = (b0 >> 5) + tid; w <= ((b1 - 1) >> 5); w += defN) {
              const uint32_t lo = max(b0, w << 5), hi = min(b1, (w + 1) << 5);
              const uint32_t cntb = hi - lo;
              atomicOr(&s_valid[w], (cntb == 32 ? 0xFFFFFFFFu : ((1u << cntb) - 1u)) << (lo & 31));
            }
          }
          bar_sync(defBar, defN);
        }
      }
      mbar_phase ^= 1;
    } else {
      // ---- oversized def stream: sequential walk straight from global memory ---------------------------------
      if (tid == 0) { s_walk.pos = 4; s_walk.end = 4 + L; s_walk.k = 0; s_done = 0; }
      __syncthreads();
      for (;;) {
        if (tid == 0) {
          Walker w = s_walk;
          s_nent = walk_runs_seq(t.src, w, 1, n, ent_k, ent_pos);
          s_walk = w;
          s_done = (w.k >= n || w.pos >= w.end) ? 1 : 0;
        }
        __syncthreads();
        const int nr = s_nent;
        for (int j = tid; j < nr; j += kT) {
          const uint32_t info = ent_pos[j], k = ent_k[j], k1 = min(ent_k[j + 1], n);
          if (info & 1) copy_bits(s_valid, 0, t.src + (info >> 1), k, k1);
          else if (info >> 1) set_bits(s_valid, k, k1);
        }
        __syncthreads();
        if (s_done) break;
      }
    }
    __syncthreads();          // the chase warp rejoins; s_valid is complete
    NDS_CKPT(2);
    // ---- 3. exclusive prefix of popcounts per word -> s_rank ----------------------------------------------
    {
      const uint32_t per = (nw + kT - 1) / kT;
      const uint32_t w0 = min((uint32_t)tid * per, nw), w1 = min(w0 + per, nw);
      uint32_t sum = 0;
      for (uint32_t w = w0; w < w1; ++w) sum += __popc(s_valid[w]);
      uint32_t incl = sum;
      for (int d = 1; d < 32; d <<= 1) {
        uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += v;
      }
      if (lane == 31) s_warp_tot[warp] = incl;
      __syncthreads();
      uint32_t base = 0;
      for (int wv = 0; wv < warp; ++wv) base += s_warp_tot[wv];
      uint32_t run = base + incl - sum;
      for (uint32_t w = w0; w < w1; ++w) {
        s_rank[w] = run;
        run += __popc(s_valid[w]);
      }
      nn = 0;
      for (int wv = 0; wv < kT / 32; ++wv) nn += s_warp_tot[wv];
    }
    __syncthreads();
    // ---- validity bitmap out (the page may start at any bit of the column bitmap) ----------------------------
    if (slice == 0) {
      const uint32_t sh = (uint32_t)(t.out_row & 31);
      uint32_t* vout = t.valid + (t.out_row >> 5);
      const uint32_t nwo = (n + sh + 31) >> 5;
      for (uint32_t j = tid; j < nwo; j += kT) {
        uint32_t lo = j < nw ? s_valid[j] : 0u;
        uint32_t hi = (j > 0 && j - 1 < nw) ? s_valid[j - 1] : 0u;
        uint32_t v = sh ? ((lo << sh) | (hi >> (32 - sh))) : lo;
        if (j == 0 || j == nwo - 1) { if (v) atomicOr(&vout[j], v); }
        else vout[j] = v;
      }
    }
  } else if (t.valid != nullptr && slice == 0) {
    // page without def levels (REQUIRED in this file) of a column that is nullable elsewhere
    const uint32_t sh = (uint32_t)(t.out_row & 31);
    uint32_t* vout = t.valid + (t.out_row >> 5);
    const uint32_t nwo = (n + sh + 31) >> 5;
    for (uint32_t j = tid; j < nwo; j += kT) {
      uint32_t v = 0xFFFFFFFFu;
      if (j == 0) v &= 0xFFFFFFFFu << sh;
      uint32_t endbit = n + sh - j * 32;
      if (endbit < 32) v &= (1u << endbit) - 1u;
      if (j == 0 || j == nwo - 1) atomicOr(&vout[j], v);
      else vout[j] = v;
    }
  }
  NDS_CKPT(3);

  const uint8_t* vsrc = t.src + values_off;
  const int64_t orow = t.out_row;

  // rank / validity of row i
  auto row_state = [&](uint32_t i, uint32_t& k) -> bool {
    if (!has_def) { k = i; return true; }
    uint32_t wv = s_valid[i >> 5];
    k = s_rank[i >> 5] + __popc(wv & ((1u << (i & 31)) - 1u));
    return (wv >> (i & 31)) & 1u;
  };

  // this slice's share of the page, in 32-row tiles
  const uint32_t ntiles = (n + 31) >> 5;
  const uint32_t tiles_per_slice = (ntiles + nslices - 1) / nslices;
  const uint32_t stile0 = min((uint32_t)slice * tiles_per_slice, ntiles), stile1 = min(stile0 + tiles_per_slice, ntiles);
  const uint32_t row0 = stile0 * 32, row1 = min(stile1 * 32, n);

  // ---- 4a. PLAIN fixed width -------------------------------------------------------------------------------------
  if (t.enc == 0 && t.phys == 0) {
    for (uint32_t base = row0; base < row1; base += kT * kUnroll) {
      uint32_t k[kUnroll];
      bool ok[kUnroll];
      int64_t v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        uint32_t i = base + u * kT + tid;
        ok[u] = i < row1 && row_state(i, k[u]);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        v[u] = 0;
        if (ok[u]) v[u] = t.width == 4 ? (int64_t)__ldg(reinterpret_cast<const int32_t*>(vsrc) + k[u])
                                       : __ldg(reinterpret_cast<const int64_t*>(vsrc) + k[u]);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        uint32_t i = base + u * kT + tid;
        if (i < row1) {
          if (t.width == 4) reinterpret_cast<int32_t*>(t.out)[orow + i] = (int32_t)v[u];
          else reinterpret_cast<int64_t*>(t.out)[orow + i] = v[u];
        }
      }
    }
    return;
  }

  // ---- 4b. PLAIN BYTE_ARRAY -> string views --------------------------------------------------------------------------
  if (t.enc == 0 && t.phys == 1) {
    // fast path: every value has the same length L (TPC-DS *_id columns are char(16))
    const uint32_t vbytes = t.size - values_off;
    uint32_t L = nn ? load_u32_unaligned(vsrc) : 0;
    bool uniform = nn == 0 || ((uint64_t)(L + 4) * nn == vbytes);
    if (uniform) {
      for (uint32_t k = tid; k < nn && uniform; k += kT) uniform = load_u32_unaligned(vsrc + (uint64_t)k * (L + 4)) == L;
    }
    uniform = __syncthreads_and(uniform);
    if (uniform) {
      for (uint32_t i = row0 + tid; i < row1; i += kT) {
        uint32_t k;
        bool ok = row_state(i, k);
        StrView v = ok ? ((uint64_t)L << 32) | (uint64_t)(t.src_off + values_off + k * (L + 4) + 4) : 0ull;
        reinterpret_cast<StrView*>(t.out)[orow + i] = v;
      }
    } else if (tid == 0 && slice == 0) {
      // slow path: one thread follows the length prefixes (variable-length free text)
      uint32_t pos = 0;
      for (uint32_t i = 0; i < n; ++i) {
        bool ok = !has_def || ((s_valid[i >> 5] >> (i & 31)) & 1u);
        StrView v = 0;
        if (ok) {
          if (pos + 4 > vbytes) { if (err) atomicExch(err, 4); break; }
          uint32_t len = load_u32_unaligned(vsrc + pos);
          v = ((uint64_t)len << 32) | (uint64_t)(t.src_off + values_off + pos + 4);
          pos += 4 + len;
        }
        reinterpret_cast<StrView*>(t.out)[orow + i] = v;
      }
    }
    return;
  }

  // ---- 4c. RLE_DICTIONARY -------------------------------------------------------------------------------------------
  if (dict_staged) mbar_wait(&s_mbar_dict, 0);       // the dictionary has landed in shared memory
  if (bw > 32) { if (tid == 0 && err) atomicExch(err, 5); return; }
  const uint32_t L = VL;
  uint32_t vsd = 0;                                     // shared-window address of the staged index stream
  __syncthreads();                                      // s_early_n is final; everyone is done with the def bytes
  // long bit-packed runs (the early chase covered the stream in a few runs): read the bits straight from global
  // memory; otherwise a small stream is staged and its many short runs are resolved in parallel
  const bool v_staged = v_small && s_early_n < 0;
  if (v_staged) {
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(vstream) & ~uintptr_t(15);
    const uintptr_t a1 = (reinterpret_cast<uintptr_t>(vstream) + L + 15) & ~uintptr_t(15);
    if (tid == 0) bulk_load_issue(s_stage, reinterpret_cast<const void*>(a0), (uint32_t)(a1 - a0), &s_mbar);
    mbar_wait(&s_mbar, mbar_phase);
    mbar_phase ^= 1;
    vsd = smem_u32(s_stage) + (uint32_t)(reinterpret_cast<uintptr_t>(vstream) - a0);
  }
  if (tid == 0) { s_walk.pos = 0; s_walk.end = L; s_walk.k = 0; s_done = 0; }
  __syncthreads();
  const bool use_early = !v_staged && bw > 0 && s_early_n >= 0;
  const uint32_t dict_s = smem_u32(s_dict);
  const uint32_t bwmask = bw >= 32 ? 0xFFFFFFFFu : ((1u << bw) - 1u);
  const uint32_t vsd_al = vsd & ~3u, vsd_mis = (vsd & 3u) * 8u;
  // contiguous block of 32-row tiles per warp, so that a warp's run cursor only moves forward
  const uint32_t tiles_per_warp = (stile1 - stile0 + kT / 32 - 1) / (kT / 32);
  const uint32_t tile0 = min(stile0 + (uint32_t)warp * tiles_per_warp, stile1), tile1 = min(tile0 + tiles_per_warp, stile1);
  for (;;) {
    // ---- entries: (header position | run info, first value index) of run headers, in stream order --------------
    const uint32_t* ek = ent_k;
    const uint32_t* ep = ent_pos;
    int ne;
    uint32_t k_lo, k_hi;
    if (use_early) {
      ek = early_k;
      ep = early_info;
      ne = s_early_n;
      k_lo = 0;
      k_hi = nn;
      if (tid == 0) s_done = 1;
    } else {
      if (tid == 0) s_nent = 0;
      __syncthreads();
      if (bw == 0) {
        if (tid == 0) { s_walk.k = nn; s_walk.pos = L; }
        __syncthreads();
      } else if (v_staged) {
        // small stream (many short runs fit here): resolve every header in parallel, segment by segment
        while (s_walk.pos < L && s_walk.k < nn && s_nent + kDefSeg / kChunk <= kEntCap)
          resolve_segment(vsd, L, bw, nn, &s_walk, s_hop, ent_pos, ent_k, &s_nent, tid, 0, kT);
      } else {
        // large stream with more runs than the early chase could record: batches of kEntCap runs
        if (tid < 32) {
          Walker w = s_walk;
          int nr = walk_runs_warp(vstream, w, bw, nn, kEntCap, 1 << 30, ent_k, ent_pos);
          __syncwarp();
          if (tid == 0) { s_nent = nr; s_walk = w; }
        }
        __syncthreads();
      }
      ne = s_nent;
      k_lo = (bw && ne) ? ent_k[0] : 0u;
      k_hi = s_walk.k;
      if (tid == 0) s_done = (s_walk.k >= nn || s_walk.pos >= L) ? 1 : 0;
    }
    NDS_CKPT(4);
    // ---- rows: 4 tiles of 32 rows per warp iteration; a warp-uniform cursor tracks the current run --------------
    int ce = 0;                                   // entry index at or before the cursor
    uint32_t cp = 0, ck = 0, ccnt = 0, cinfo = 0; // current run: header position (staged), first k, values, info
    bool cur_ok = false;
    // rows are expanded in blocks of 8 tiles (256 rows). Block fast path: when all values of the block lie in the
    // cursor's run or the one after it (runs of 504 values: almost always), the run of a value is a compare+select
    // and the per-tile cursor bookkeeping below is skipped.
    const uintptr_t vs_al = reinterpret_cast<uintptr_t>(vstream) & ~uintptr_t(3);
    const uint32_t vs_mis = (uint32_t)(reinterpret_cast<uintptr_t>(vstream) & 3) * 8;
    const uint32_t* vwords = reinterpret_cast<const uint32_t*>(vs_al);
    const uint32_t lt_mask = (1u << lane) - 1u;
    for (uint32_t bb = tile0; bb < tile1; bb += 8) {
      const uint32_t be = min(bb + 8u, tile1);
      const uint32_t kA = has_def ? s_rank[bb] : bb * 32;
      const uint32_t kB = has_def ? (be < nw ? s_rank[be] : nn) : min(be * 32, n);
      bool fast = kA >= k_lo && kB <= k_hi;
      uint32_t rk1 = 0, info1 = 0, rk2 = 0xFFFFFFFFu, info2 = 0;
      if (fast && bw && kB > kA) {
        // uniform cursor move to the run holding kA (same rules as the per-tile path below)
        const uint32_t kmin = kA;
        if (!cur_ok || kmin < ck || kmin >= ck + ccnt) {
          if (!cur_ok || kmin < ck || (ce + 1 < ne && ek[ce + 1] <= kmin)) {
            int lo = (cur_ok && kmin >= ck) ? ce : 0, hi = ne - 1;
            while (lo < hi) {
              int mid = (lo + hi + 1) >> 1;
              if (ek[mid] <= kmin) lo = mid; else hi = mid - 1;
            }
            ce = lo;
            cur_ok = false;
          }
          if (!v_staged) {
            ck = ek[ce];
            ccnt = ek[ce + 1] - ck;
            cinfo = ep[ce];
          } else {
            uint32_t len;
            if (!cur_ok) { cp = ep[ce]; ck = ek[ce]; }
            else {
              uint32_t c0, i0;
              len = parse_run_s(vsd, cp, L, bw, c0, i0);
              cp += len;
              ck += c0;
            }
            for (;;) {
              len = parse_run_s(vsd, cp, L, bw, ccnt, cinfo);
              if (ck + ccnt > kmin || cp + len >= L) break;
              cp += len;
              ck += ccnt;
            }
          }
          cur_ok = true;
        }
        rk1 = ck;
        info1 = cinfo;
        uint32_t cnt2 = 0;
        if (kB > ck + ccnt) {               // the block reaches into the next run
          if (!v_staged) {
            if (ce + 1 < ne) { rk2 = ek[ce + 1]; cnt2 = ek[ce + 2] - rk2; info2 = ep[ce + 1]; }
          } else {
            uint32_t c1, i1;
            const uint32_t len1 = parse_run_s(vsd, cp, L, bw, c1, i1);
            if (cp + len1 < L) { parse_run_s(vsd, cp + len1, L, bw, cnt2, info2); rk2 = ck + ccnt; }
          }
          fast = rk2 != 0xFFFFFFFFu && kB <= rk2 + cnt2;
        }
      }
      if (fast) {
        uint32_t kk[8], word0[8], word1[8], sh[8];
        int stt[8];                 // 0 skip, 1 zero, 2 RLE / bw == 0 value in word0, 3 bit-packed
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const uint32_t tcur = bb + u;
          const uint32_t i = tcur * 32 + lane;
          stt[u] = 0;
          word0[u] = 0;
          word1[u] = 0;
          sh[u] = 0;
          kk[u] = 0;
          if (tcur < be && i < n) {
            bool ok = true;
            if (has_def) {
              const uint32_t wv = s_valid[tcur];
              kk[u] = s_rank[tcur] + __popc(wv & lt_mask);
              ok = (wv >> lane) & 1u;
            } else kk[u] = i;
            stt[u] = 1;
            if (ok) {
              const bool in2 = kk[u] >= rk2;
              const uint32_t rk = in2 ? rk2 : rk1, info = in2 ? info2 : info1;
              if (!bw || !(info & 1)) { word0[u] = bw ? (info >> 1) : 0u; stt[u] = 2; }
              else if (!v_staged) {
                const uint32_t bit = (info >> 1) * 8u + (kk[u] - rk) * bw + vs_mis;
                word0[u] = __ldg(vwords + (bit >> 5));
                word1[u] = __ldg(vwords + (bit >> 5) + 1);
                sh[u] = bit & 31;
                stt[u] = 3;
              } else {
                word0[u] = smem_bits(vsd_al, vsd_mis + (info >> 1) * 8u + (kk[u] - rk) * bw, bwmask);
                stt[u] = 2;
              }
            }
          }
        }
        int64_t vv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          vv[u] = 0;
          if (stt[u] >= 2) {
            uint32_t ix = stt[u] == 3 ? (__funnelshift_r(word0[u], word1[u], sh[u]) & bwmask) : word0[u];
            if (ix >= t.dict_count) { if (err) atomicExch(err, 6); continue; }
            if (dict_staged) vv[u] = t.width == 4 ? (int64_t)(int32_t)lds32(dict_s + ix * 4u) : lds64(dict_s + ix * 8u);
            else vv[u] = t.width == 4 ? (int64_t)__ldg(reinterpret_cast<const int32_t*>(t.dict) + ix)
                                      : __ldg(reinterpret_cast<const int64_t*>(t.dict) + ix);
          }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if (stt[u]) {
            const uint32_t i = (bb + u) * 32 + lane;
            if (t.width == 4) reinterpret_cast<int32_t*>(t.out)[orow + i] = (int32_t)vv[u];
            else reinterpret_cast<int64_t*>(t.out)[orow + i] = vv[u];
          }
        }
        continue;
      }
      // ---- general path: per-tile cursor, any mix of runs, batch boundaries ---------------------------------
      const uint32_t sb = bb, se = be;
        for (uint32_t tb = sb; tb < se; tb += kUnroll) {
        uint32_t k[kUnroll], idx[kUnroll];
        uint64_t gbit[kUnroll];                     // global bit offset of a bit-packed value not yet fetched
        int st[kUnroll];                            // 0 skip, 1 write null/zero, 2 write value, 3 value needs a global fetch
  #pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          uint32_t i = (tb + u) * 32 + lane;
          st[u] = 0;
          idx[u] = 0;
          k[u] = 0;
          gbit[u] = 0;
          if (tb + u < se && i < n) {
            bool ok = row_state(i, k[u]);
            if (!ok) st[u] = k_lo == 0 ? 1 : 0;        // nulls are written once, with the first batch
            else if (k[u] >= k_lo && k[u] < k_hi) st[u] = 2;
          }
        }
        if (bw) {
          // pass 1: locate every value's run (no global loads for bit-packed values yet)
  #pragma unroll
          for (int u = 0; u < kUnroll; ++u) {
            const uint32_t want = __ballot_sync(0xFFFFFFFFu, st[u] == 2);
            if (!want) continue;
            // k grows with the lane: first / last wanted lanes give the tile's k range
            const uint32_t kmin = __shfl_sync(0xFFFFFFFFu, k[u], __ffs(want) - 1);
            const uint32_t kmax = __shfl_sync(0xFFFFFFFFu, k[u], 31 - __clz(want));
            if (!cur_ok || kmin < ck || kmin >= ck + ccnt) {
              // move the warp cursor to the run containing kmin (uniform across the warp)
              if (!cur_ok || kmin < ck || (ce + 1 < ne && ek[ce + 1] <= kmin)) {
                int lo = (cur_ok && kmin >= ck) ? ce : 0, hi = ne - 1;     // last entry with ek <= kmin
                while (lo < hi) {
                  int mid = (lo + hi + 1) >> 1;
                  if (ek[mid] <= kmin) lo = mid; else hi = mid - 1;
                }
                ce = lo;
                cur_ok = false;
              }
              if (!v_staged) {
                ck = ek[ce];
                ccnt = ek[ce + 1] - ck;
                cinfo = ep[ce];
              } else {
                uint32_t len;
                if (!cur_ok) { cp = ep[ce]; ck = ek[ce]; }
                else {                      // current run is known and too early: step past it first
                  uint32_t c0, i0;
                  len = parse_run_s(vsd, cp, L, bw, c0, i0);
                  cp += len;
                  ck += c0;
                }
                for (;;) {
                  len = parse_run_s(vsd, cp, L, bw, ccnt, cinfo);
                  if (ck + ccnt > kmin || cp + len >= L) break;
                  cp += len;
                  ck += ccnt;
                }
              }
              cur_ok = true;
            }
            if (st[u] != 2) continue;
            uint32_t rk = ck, info = cinfo;
            if (kmax >= ck + ccnt && k[u] >= ck + ccnt) {
              // the tile straddles runs and my value lies in a later one: walk on privately
              if (!v_staged) {
                int e = ce;
                while (e + 1 < ne && ek[e + 1] <= k[u]) ++e;
                rk = ek[e];
                info = ep[e];
              } else {
                uint32_t p = cp, cnt = ccnt;
                for (;;) {
                  uint32_t len = parse_run_s(vsd, p, L, bw, cnt, info);
                  if (rk + cnt > k[u] || p + len >= L) break;
                  p += len;
                  rk += cnt;
                }
              }
            }
            if (!(info & 1)) idx[u] = info >> 1;
            else if (!v_staged) {
              gbit[u] = (uint64_t)(info >> 1) * 8u + (uint64_t)(k[u] - rk) * bw;
              st[u] = 3;
            } else {
              idx[u] = smem_bits(vsd_al, vsd_mis + (info >> 1) * 8u + (k[u] - rk) * bw, bwmask);
            }
          }
          // pass 2: all global fetches of the four tiles back to back, pass 3: combine
          uint32_t w0[kUnroll], w1[kUnroll];
  #pragma unroll
          for (int u = 0; u < kUnroll; ++u) {
            w0[u] = 0;
            w1[u] = 0;
            if (st[u] == 3) {
              const uintptr_t a = reinterpret_cast<uintptr_t>(vstream) + (uintptr_t)(gbit[u] >> 3);
              const uint32_t* wp = reinterpret_cast<const uint32_t*>(a & ~uintptr_t(3));
              w0[u] = __ldg(wp);
              w1[u] = __ldg(wp + 1);
            }
          }
  #pragma unroll
          for (int u = 0; u < kUnroll; ++u) {
            if (st[u] == 3) {
              const uintptr_t a = reinterpret_cast<uintptr_t>(vstream) + (uintptr_t)(gbit[u] >> 3);
              uint64_t v = ((uint64_t)w1[u] << 32) | w0[u];
              v >>= (uint32_t)(a & 3) * 8 + (uint32_t)(gbit[u] & 7);
              idx[u] = bw == 32 ? (uint32_t)v : ((uint32_t)v & ((1u << bw) - 1u));
              st[u] = 2;
            }
          }
        }
        int64_t v[kUnroll];
  #pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          v[u] = 0;
          if (st[u] == 2) {
            if (idx[u] >= t.dict_count) { if (err) atomicExch(err, 6); st[u] = 1; continue; }
            if (dict_staged) v[u] = t.width == 4 ? (int64_t)(int32_t)lds32(dict_s + idx[u] * 4u) : lds64(dict_s + idx[u] * 8u);
            else v[u] = t.width == 4 ? (int64_t)__ldg(reinterpret_cast<const int32_t*>(t.dict) + idx[u])
                                     : __ldg(reinterpret_cast<const int64_t*>(t.dict) + idx[u]);
          }
        }
  #pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          uint32_t i = (tb + u) * 32 + lane;
          if (st[u]) {
            if (t.width == 4) reinterpret_cast<int32_t*>(t.out)[orow + i] = (int32_t)v[u];
            else reinterpret_cast<int64_t*>(t.out)[orow + i] = v[u];
          }
        }
      }
    }
    __syncthreads();
    NDS_CKPT(5);
    if (s_done) break;
  }
}

// Fill a partition-injected column: row range [seg_start[s], seg_start[s+1]) gets seg_val[s].
__global__ void fill_segments_kernel(const int64_t* __restrict__ seg_start, const int64_t* __restrict__ seg_val,
                                     int nseg, int32_t* __restrict__ out) {
  for (int s = blockIdx.y; s < nseg; s += gridDim.y) {
    int64_t a = seg_start[s], b = seg_start[s + 1];
    int32_t v = (int32_t)seg_val[s];
    for (int64_t i = a + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < b; i += (int64_t)gridDim.x * blockDim.x) out[i] = v;
  }
}

int decode_slices(int npages) {
  // few pages: share each page between several CTAs so that the launch still covers the 148 SMs twice
  int nslices = 1;
  if (npages > 0 && npages < 2 * 148) {
    nslices = (2 * 148 + npages - 1) / npages;
    if (nslices > 8) nslices = 8;
  }
  return nslices;
}

void launch_decode_pages(const PageTask* d_tasks, const int* d_idx, int npages, int* d_err, cudaStream_t stream, long long* dbg) {
  if (npages <= 0) return;
  const int nslices = decode_slices(npages);
  static bool attr_set[64] = {false};
  int dev = 0;
  NDS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    NDS_CUDA(cudaFuncSetAttribute(decode_pages_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDictStage));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  decode_pages_kernel<<<npages * nslices, kDecodeThreads, kDictStage, stream>>>(d_tasks, d_idx, nslices, d_err, dbg);
  NDS_CUDA(cudaGetLastError());
}

void launch_fill_segments(const int64_t* d_seg_start, const int64_t* d_seg_val, int nseg, int32_t* out,
                          cudaStream_t stream) {
  if (nseg <= 0) return;
  dim3 grid(8, nseg < 4096 ? nseg : 4096);
  fill_segments_kernel<<<grid, 256, 0, stream>>>(d_seg_start, d_seg_val, nseg, out);
  NDS_CUDA(cudaGetLastError());
}

}  // namespace nds
