// Fused fact-table scan: page decode -> join-chain probe in one kernel, and late materialisation of the surviving rows.
//
// A star-join query reads a few columns of a fact table, probes the join tables of its dimensions with the key columns
// and keeps a small fraction of the rows (NDS query7: ~1 %).  Decoding every column of every row into HBM first
// (decode_fast.cu) and probing the decoded keys afterwards (ops.cu, join_probe_chain_kernel) writes and re-reads the
// decoded columns and decodes measures of rows no join keeps.  Here instead:
//
//   scan_probe_kernel    one CTA per 2048-row chunk of a page.  Pass A: every thread decodes four consecutive values of
//                        the FIRST probe key straight from the page (validity / rank words of the def kernel -> value
//                        index -> bit-packed dictionary index -> dictionary -> key), probes that join's table and queues
//                        the hits in shared memory.  Pass B: one thread per queued row decodes the remaining probe keys
//                        of that row only and walks the rest of the chain; survivors are appended to the output with the
//                        matched build row of every join.  No decoded key column ever exists in memory.
//   gather_decode_kernel one thread per survivor and payload column: the value of that row only (page lookup by row,
//                        rank from the def kernel's words, index, dictionary) into a flat column + validity bitmap.
//
// Both rely on pages whose value k can be addressed without walking run headers (engine.h, PageInfo::direct): PLAIN
// fixed-width pages, one-entry dictionaries, and index streams of back-to-back 504-value bit-packed runs, where value
// k sits at bit k * bw + 8 * (k / 504 + 1) of the stream.  Exec::join falls back to decode + probe for anything else.
//
// Spark operators these replace: FileSourceScanExec + BroadcastHashJoinExec chains under whole-stage codegen; the
// reference hands SQL text to Spark at /root/reference/nds/nds_power.py:130.
#include "decode_common.cuh"
#include "ops.cuh"
#include <climits>
#include <type_traits>

namespace nds {

namespace {

struct PageView {                 // what decoding one value of a page needs, hoisted out of the row loop
  const uint32_t* words;          // 4 B aligned base of the index stream (dictionary pages)
  const uint8_t* plain;           // PLAIN values (16 B aligned)
  const void* dict;
  const uint32_t* vr_valid;       // validity words of the page (def kernel scratch), nullptr = no nulls
  const uint32_t* vr_rank;        // rank prefix per word; vr_rank[nw] = non-null values in the page
  int64_t out_row;                // global row of the page's first row
  uint32_t n;                     // rows in the page
  uint32_t dict_count;
  uint32_t bw, bwmask;
  uint32_t off8;                  // bit offset of value 0 from `words`: stream misalignment + the first header byte
  uint32_t kmax;                  // last valid value index (non-null count - 1; 0 when the page holds no value)
  uint32_t is_dict, width;
  uint32_t tail_k, tail_ix;       // values from tail_k on: one closing RLE run of index tail_ix (PageTask)
};

__device__ __forceinline__ PageView make_view(const PageTask& t) {
  PageView v;
  const uint32_t values_off = t.def_len ? 4u + t.def_len : 0u;
  const uint32_t nw = (t.num_values + 31u) >> 5;
  v.n = t.num_values;
  v.out_row = t.out_row;
  v.is_dict = t.enc == 1;
  v.width = t.width;
  v.dict = t.dict;
  v.dict_count = t.dict_count;
  v.bw = t.bw;
  v.bwmask = t.bw >= 32 ? 0xFFFFFFFFu : ((1u << t.bw) - 1u);
  v.plain = t.src + values_off;
  const uint8_t* vstream = t.src + values_off + 1;
  v.words = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(vstream) & ~uintptr_t(3));
  v.off8 = (uint32_t)(reinterpret_cast<uintptr_t>(vstream) & 3) * 8u + 8u;
  v.vr_valid = t.vr;
  v.vr_rank = t.vr ? t.vr + nw : nullptr;
  const uint32_t nn = t.vr ? t.vr[2 * nw] : t.num_values;
  v.kmax = nn ? nn - 1u : 0u;
  v.tail_k = t.tail_k;
  v.tail_ix = t.tail_ix;
  return v;
}

// value index k of a dictionary page -> dictionary index; false + error 6 when it is out of range
__device__ __forceinline__ bool page_index(const PageView& v, uint32_t k, uint32_t& ix, int* err) {
  ix = 0;
  if (k >= v.tail_k) ix = v.tail_ix;
  else if (v.bw) {
    const uint32_t bit = k * v.bw + (k / 504u) * 8u + v.off8;
    const uint32_t* wp = v.words + (bit >> 5);
    ix = __funnelshift_r(__ldg(wp), __ldg(wp + 1), bit & 31) & v.bwmask;
  }
  if (ix >= v.dict_count) { if (err) atomicExch(err, 6); return false; }
  return true;
}
// value index k (k-th non-null value of the page) -> value
__device__ __forceinline__ bool page_value(const PageView& v, uint32_t k, int64_t& out, int* err) {
  if (!v.is_dict) {
    out = v.width == 4 ? (int64_t)__ldg(reinterpret_cast<const int32_t*>(v.plain) + k)
                       : (int64_t)__ldg(reinterpret_cast<const long long*>(v.plain) + k);
    return true;
  }
  uint32_t ix;
  if (!page_index(v, k, ix, err)) return false;
  out = v.width == 4 ? (int64_t)__ldg(reinterpret_cast<const int32_t*>(v.dict) + ix)
                     : (int64_t)__ldg(reinterpret_cast<const long long*>(v.dict) + ix);
  return true;
}

// row r of the page -> (is it non-null, its value index)
__device__ __forceinline__ bool row_rank(const PageView& v, uint32_t r, uint32_t& k) {
  if (!v.vr_valid) { k = r; return true; }
  const uint32_t w = __ldg(v.vr_valid + (r >> 5));
  k = __ldg(v.vr_rank + (r >> 5)) + __popc(w & ((1u << (r & 31)) - 1u));
  return (w >> (r & 31)) & 1u;
}

}  // namespace

// Page tasks of a scan, made on the device: one warp per (column, selected file).
__global__ void __launch_bounds__(128)
expand_tasks_kernel(const FileJob* __restrict__ jobs, int njobs, PageTask* __restrict__ tasks, int* __restrict__ def_idx,
                    FusedChunk* __restrict__ chunks, int32_t* __restrict__ page_file) {
  const int w = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (w >= njobs) return;
  const FileJob j = jobs[w];
  for (int p = lane; p < j.npages; p += 32) {
    PageTask t = j.tmpl[j.page_begin + p];
    const uint2 aux = j.page_aux[j.page_begin + p];
    t.src = j.src_base + reinterpret_cast<uintptr_t>(t.src);
    if (t.dict) t.dict = j.src_base + (reinterpret_cast<uintptr_t>(t.dict) - 1);
    t.out = nullptr;
    t.valid = nullptr;
    t.out_row += j.row_start;
    t.vr = j.vr_base ? j.vr_base + aux.x : nullptr;
    tasks[j.out_task0 + p] = t;
    if (j.def_base >= 0) def_idx[j.def_base + p] = j.out_task0 + p;
    if (j.chunk_base >= 0) {
      page_file[j.out_task0 + p - j.col_task0] = j.file;
      uint32_t c = (uint32_t)j.chunk_base + aux.y;
      for (uint32_t r0 = 0; r0 < t.num_values; r0 += kFusedChunk, ++c)
        chunks[c] = FusedChunk{(uint32_t)(j.out_task0 + p - j.col_task0), r0, (uint32_t)j.file, 0u};
    }
  }
}

void launch_expand_tasks(const FileJob* d_jobs, int njobs, PageTask* tasks, int* def_idx, FusedChunk* chunks, int32_t* page_file,
                         cudaStream_t s) {
  if (njobs <= 0) return;
  expand_tasks_kernel<<<(njobs * 32 + 127) / 128, 128, 0, s>>>(d_jobs, njobs, tasks, def_idx, chunks, page_file);
  NDS_CUDA(cudaGetLastError());
}

__global__ void __launch_bounds__(kFusedThreads)
scan_probe_kernel(const FusedScanArgs a) {
  __shared__ PageView s_view[kMaxChain];
  __shared__ uint16_t q_row[kFusedChunk];          // pass A hits: row inside the chunk
  __shared__ uint32_t q_b0[kFusedChunk];           //              matched build row of the first join
  __shared__ int s_nq;
  __shared__ int s_part_ok;
  __shared__ uint32_t s_part_row;

  const FusedChunk ch = a.chunks[blockIdx.x];
  const int tid = threadIdx.x;
  if (tid < a.nkeys) s_view[tid] = make_view(a.key[tid].tasks[ch.page]);
  if (tid == 32) {
    s_nq = 0;
    s_part_ok = 1;
    s_part_row = 0;
    if (a.has_part) {
      // the join on the partition column: one lookup per file (the value is the directory name's)
      uint32_t r = 0;
      s_part_ok = ht_find(a.part_ht, a.file_part_val[ch.file], r) ? 1 : 0;
      s_part_row = r;
    }
  }
  __syncthreads();
  if (!s_part_ok) return;

  // ---- pass A: the first probe key of every row of the chunk, four consecutive rows per thread
  {
    const PageView v = s_view[0];
    const HashTable ht = a.key[0].ht;
    const uint32_t rend = min(ch.row0 + (uint32_t)kFusedChunk, v.n);
    for (uint32_t base = ch.row0 + (uint32_t)tid * 4u; base < rend; base += kFusedThreads * 4u) {
      uint32_t vbits = 0xFu, k0 = base;
      if (v.vr_valid) {
        const uint32_t w = __ldg(v.vr_valid + (base >> 5));
        const uint32_t sh = base & 31u;       // multiple of 4: the four rows share one validity word
        k0 = __ldg(v.vr_rank + (base >> 5)) + __popc(w & ((1u << sh) - 1u));
        vbits = (w >> sh) & 0xFu;
      }
      if (base + 4u > rend) vbits &= (1u << (rend - base)) - 1u;
      if (!vbits) continue;
      // straight-line: every slot fetches a value (null slots the neighbouring one's, clamped to the page's last), the
      // validity bits decide afterwards - four independent load chains in flight per thread
      int64_t key[4];
      bool okv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t k = min(k0 + __popc(vbits & ((1u << u) - 1u)), v.kmax);
        okv[u] = page_value(v, k, key[u], a.err) && ((vbits >> u) & 1u);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        uint32_t b = 0;
        if (okv[u] && ht_find(ht, key[u], b)) {
          const int pos = atomicAdd(&s_nq, 1);
          q_row[pos] = (uint16_t)(base + u - ch.row0);
          q_b0[pos] = b;
        }
      }
    }
  }
  __syncthreads();

  // ---- pass B: the remaining joins for the queued rows only; survivors go to the output
  const int nq = s_nq;
  const int nq_round = (nq + 31) & ~31;
  const uint32_t row_base = (uint32_t)s_view[0].out_row;
  for (int i = tid; i < nq_round; i += kFusedThreads) {
    bool alive = i < nq;
    uint32_t r = 0;
    uint32_t b[kMaxChain];
#pragma unroll
    for (int j = 0; j < kMaxChain; ++j) b[j] = 0;
    if (alive) { r = ch.row0 + q_row[i]; b[0] = q_b0[i]; }
#pragma unroll
    for (int j = 1; j < kMaxChain; ++j) {
      if (j < a.nkeys && alive) {
        uint32_t k;
        int64_t key;
        alive = row_rank(s_view[j], r, k) && page_value(s_view[j], k, key, a.err) && ht_find(a.key[j].ht, key, b[j]);
      }
    }
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, alive);
    if (!m) continue;
    const int lane = tid & 31, leader = __ffs(m) - 1;
    int pos0 = 0;
    if (lane == leader) pos0 = atomicAdd(a.out_count, __popc(m));
    pos0 = __shfl_sync(0xFFFFFFFFu, pos0, leader);
    if (alive) {
      const uint32_t pos = (uint32_t)pos0 + __popc(m & ((1u << lane) - 1u));
      if (pos < a.cap) {
        a.out_row[pos] = row_base + r;
        a.out_file[pos] = ch.file;
#pragma unroll
        for (int j = 0; j < kMaxChain; ++j)
          if (j < a.nkeys) a.key[j].out_build[pos] = b[j];
        if (a.has_part) a.part_out_build[pos] = s_part_row;
      } else {
        atomicExch(a.err + 2, 1);      // more survivors than the speculative capacity: the query is run again with room for all
      }
    }
  }
}

// ---- the first join tested per dictionary entry instead of per row
//
// A warp load whose 32 lanes hit 32 different cache lines occupies the L1 for 32 cycles, and scan_probe_kernel issues two
// of those per row (dictionary entry, join-table entry): that, not DRAM, bounds it (NDS query7 at SF100: 55 M rows x 2
// scattered reads / 148 SMs = 0.39 ms).  A partition file's pages share one dictionary, so:
//
//   dict_match_kernel      reads every selected file's dictionary ONCE, in order (coalesced), tests each entry against the
//                          first join - a coarse existence bitmap of the key range in shared memory rejects most entries,
//                          the rest look the join table up - and writes one bit per entry.
//   scan_probe_dm_kernel   keeps the current file's match bits in shared memory; pass A costs a row its index bits and one
//                          shared-memory bit.  Rows whose bit is set (a few percent) go through pass B, which now also
//                          does the first join's dictionary + table lookup.
// MODE 0: any join table (ht_find per entry); 1: direct-address table; 2: direct-address table, 4-byte keys and a key range
// inside int32 - all arithmetic in 32 bits, no branches besides the (rare) table read behind a coarse hit
template <int WIDTH, int MODE>
__global__ void __launch_bounds__(kDictMatchThreads)
dict_match_kernel(const PageTask* __restrict__ tasks, const int32_t* __restrict__ dict_task, HashTable ht,
                  const uint32_t* __restrict__ coarse, uint32_t coarse_words, uint32_t coarse_shift,
                  uint32_t* __restrict__ out, uint32_t stride_words) {
  extern __shared__ __align__(16) uint32_t s_coarse[];
  const int32_t dt = dict_task[blockIdx.y];
  if (dt < 0) return;
  const uint32_t cnt = tasks[dt].dict_count;
  const uint32_t e0 = blockIdx.x * kDictMatchEntries;
  if (e0 >= cnt) return;
  using T = typename std::conditional<WIDTH == 4, int32_t, long long>::type;
  const T* __restrict__ dict = reinterpret_cast<const T*>(tasks[dt].dict);
  if (MODE != 0) {
    // (coarse_words is a multiple of 4 and the bitmap 16-byte aligned: launch_dict_match's callers round up)
    for (uint32_t w = threadIdx.x; w < coarse_words / 4u; w += kDictMatchThreads)
      reinterpret_cast<uint4*>(s_coarse)[w] = __ldg(reinterpret_cast<const uint4*>(coarse) + w);
    __syncthreads();
  }
  const uint32_t e1 = min(e0 + (uint32_t)kDictMatchEntries, (cnt + 31u) & ~31u);
  const uint32_t lo32 = (uint32_t)ht.lo, range = ht.range;
  const uint32_t* __restrict__ dense = ht.dense;
  constexpr int U = 8;
  for (uint32_t i0 = e0 + threadIdx.x; i0 < e1; i0 += kDictMatchThreads * U) {
    T key[U];
    const T* __restrict__ dp = dict + i0;
    uint32_t* __restrict__ op = out + (size_t)blockIdx.y * stride_words + (i0 >> 5);
#pragma unroll
    for (int u = 0; u < U; ++u)                      // U independent loads in flight per thread
      key[u] = i0 + u * kDictMatchThreads < cnt ? __ldg(dp + u * kDictMatchThreads) : (T)0;
    bool hit[U];
    if (MODE != 0) {
      // coarse test of all U entries, then the table reads of the coarse hits together (several keys per coarse bit: the
      // table decides), so that their latencies overlap
      uint32_t dv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t i = i0 + u * kDictMatchThreads;
        uint32_t d32;
        bool in;
        if (MODE == 2) {
          d32 = (uint32_t)key[u] - lo32;
          in = (i < cnt) & (d32 < range);
        } else {
          const uint64_t d = (uint64_t)((int64_t)key[u] - ht.lo);
          in = (i < cnt) & (d < (uint64_t)range);
          d32 = (uint32_t)d;
        }
        const uint32_t b = in ? d32 >> coarse_shift : 0u;
        hit[u] = in & (bool)((s_coarse[b >> 5] >> (b & 31u)) & 1u);
        dv[u] = 1u;
        if (coarse_shift && hit[u]) dv[u] = __ldg(dense + d32);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) hit[u] = hit[u] & (dv[u] != 0u);
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        uint32_t row;
        hit[u] = i0 + u * kDictMatchThreads < cnt && ht_find(ht, (int64_t)key[u], row);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (i0 + u * kDictMatchThreads >= e1) break;   // warp-uniform: e1 and the strides are multiples of 32
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, hit[u]);
      if ((threadIdx.x & 31) == 0) op[u * (kDictMatchThreads / 32)] = m;
    }
  }
}
void launch_dict_match(const PageTask* tasks, const int32_t* dict_task, int nfiles, uint32_t max_dict, int width, const HashTable& ht,
                       const uint32_t* coarse, uint32_t coarse_words, uint32_t coarse_shift, uint32_t* out, uint32_t stride_words,
                       cudaStream_t s) {
  if (nfiles <= 0 || max_dict == 0) return;
  const dim3 grid((max_dict + kDictMatchEntries - 1) / kDictMatchEntries, nfiles);
  const size_t sm = (size_t)coarse_words * 4;
  const bool dense = ht.dense && coarse_words > 0;
  const bool narrow = dense && ht.lo >= INT32_MIN && ht.lo + (int64_t)ht.range <= INT32_MAX;
#define NDS_DM(W, M, SM) dict_match_kernel<W, M><<<grid, kDictMatchThreads, SM, s>>>(tasks, dict_task, ht, coarse, dense ? coarse_words : 0u, \
                                                                                coarse_shift, out, stride_words)
  if (width == 4) {
    if (narrow) NDS_DM(4, 2, sm);
    else if (dense) NDS_DM(4, 1, sm);
    else NDS_DM(4, 0, 0);
  } else {
    if (dense) NDS_DM(8, 1, sm);
    else NDS_DM(8, 0, 0);
  }
#undef NDS_DM
  NDS_CUDA(cudaGetLastError());
}

// One CTA per (page of the first key column, part): the page's 2048-row chunks c = part, part + split, ...  The views and
// the file's match bits are set up once; pass A queues rows chunk after chunk and pass B runs when the queue may not hold
// another chunk (an unselective join) or after the last chunk (the usual case: one pass B with all warps busy).
__global__ void __launch_bounds__(kFusedThreads)
scan_probe_dm_kernel(const FusedScanArgs a) {
  extern __shared__ __align__(16) uint32_t s_bits[];    // match bits of the file's dictionary
  __shared__ PageView s_view[kMaxChain];
  __shared__ uint16_t q_row[kDmQueue];             // pass A hits: row inside the page (pages of this route hold < 65536 rows)
  __shared__ int s_nq;
  __shared__ int s_part_ok;
  __shared__ uint32_t s_part_row;
  const int tid = threadIdx.x;
  const uint32_t page = blockIdx.x / (uint32_t)a.split, part = blockIdx.x % (uint32_t)a.split;
  const uint32_t file = (uint32_t)a.page_file[page];
  if (tid < a.nkeys) s_view[tid] = make_view(a.key[tid].tasks[page]);
  if (tid == 32) {
    s_nq = 0;
    s_part_ok = 1;
    s_part_row = 0;
    if (a.has_part) {
      uint32_t r = 0;
      s_part_ok = ht_find(a.part_ht, a.file_part_val[file], r) ? 1 : 0;
      s_part_row = r;
    }
  }
  {
    const uint32_t words4 = min(a.dmatch_stride / 4u, (a.key[0].tasks[page].dict_count + 127u) / 128u);
    const uint4* src = reinterpret_cast<const uint4*>(a.dmatch + (size_t)file * a.dmatch_stride);   // stride: multiple of 4 words
    for (uint32_t w = tid; w < words4; w += kFusedThreads) reinterpret_cast<uint4*>(s_bits)[w] = __ldg(src + w);
  }
  __syncthreads();
  if (!s_part_ok) return;
  const PageView v = s_view[0];
  const uint32_t nchunk = (v.n + (uint32_t)kFusedChunk - 1u) / (uint32_t)kFusedChunk;
  const uint32_t row_base = (uint32_t)v.out_row;
  const uint32_t dmax = v.dict_count ? v.dict_count - 1u : 0u;
  for (uint32_t c = part; c < nchunk; c += (uint32_t)a.split) {
    // ---- pass A: four consecutive rows per thread; a dictionary page's row is kept when its entry's match bit is set
    const uint32_t row0 = c * (uint32_t)kFusedChunk;
    const uint32_t rend = min(row0 + (uint32_t)kFusedChunk, v.n);
    for (uint32_t base = row0 + (uint32_t)tid * 4u; base < rend; base += kFusedThreads * 4u) {
      uint32_t vbits = 0xFu, k0 = base;
      if (v.vr_valid) {
        const uint32_t w = __ldg(v.vr_valid + (base >> 5));
        const uint32_t sh = base & 31u;
        k0 = __ldg(v.vr_rank + (base >> 5)) + __popc(w & ((1u << sh) - 1u));
        vbits = (w >> sh) & 0xFu;
      }
      if (base + 4u > rend) vbits &= (1u << (rend - base)) - 1u;
      if (!vbits) continue;
      uint32_t hits = 0;
      if (v.is_dict) {
        uint32_t bad = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t k = min(k0 + __popc(vbits & ((1u << u) - 1u)), v.kmax);
          uint32_t ix = 0;
          if (v.bw) {
            const uint32_t bit = k * v.bw + (k / 504u) * 8u + v.off8;
            const uint32_t* wp = v.words + (bit >> 5);
            ix = __funnelshift_r(__ldg(wp), __ldg(wp + 1), bit & 31) & v.bwmask;
          }
          ix = k >= v.tail_k ? v.tail_ix : ix;
          bad |= ix > dmax;
          ix = min(ix, dmax);
          hits |= ((s_bits[ix >> 5] >> (ix & 31u)) & 1u) << u;
        }
        hits &= vbits;
        if (bad) atomicExch(a.err, 6);
      } else {
        const HashTable ht = a.key[0].ht;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t k = min(k0 + __popc(vbits & ((1u << u) - 1u)), v.kmax);
          int64_t key;
          uint32_t b;
          if (page_value(v, k, key, a.err) && ((vbits >> u) & 1u) && ht_find(ht, key, b)) hits |= 1u << u;
        }
      }
      if (hits) {
        int pos = atomicAdd(&s_nq, __popc(hits));
        do {
          q_row[pos++] = (uint16_t)(base + __ffs(hits) - 1);
          hits &= hits - 1u;
        } while (hits);
      }
    }
    __syncthreads();
    const int nq = s_nq;
    if (nq + kFusedChunk <= kDmQueue && c + (uint32_t)a.split < nchunk) continue;    // room for another chunk's hits
    // ---- pass B: every join of the chain for the queued rows; survivors go to the output
    const int nq_round = (nq + 31) & ~31;
    for (int i = tid; i < nq_round; i += kFusedThreads) {
      bool alive = i < nq;
      uint32_t r = 0;
      uint32_t b[kMaxChain];
#pragma unroll
      for (int j = 0; j < kMaxChain; ++j) b[j] = 0;
      if (alive) r = q_row[i];
#pragma unroll
      for (int j = 0; j < kMaxChain; ++j) {
        if (j < a.nkeys && alive) {
          uint32_t k;
          int64_t key;
          alive = row_rank(s_view[j], r, k) && page_value(s_view[j], k, key, a.err) && ht_find(a.key[j].ht, key, b[j]);
        }
      }
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, alive);
      if (!m) continue;
      const int lane = tid & 31, leader = __ffs(m) - 1;
      int pos0 = 0;
      if (lane == leader) pos0 = atomicAdd(a.out_count, __popc(m));
      pos0 = __shfl_sync(0xFFFFFFFFu, pos0, leader);
      if (alive) {
        const uint32_t pos = (uint32_t)pos0 + __popc(m & ((1u << lane) - 1u));
        if (pos < a.cap) {
          a.out_row[pos] = row_base + r;
          a.out_file[pos] = file;
#pragma unroll
          for (int j = 0; j < kMaxChain; ++j)
            if (j < a.nkeys) a.key[j].out_build[pos] = b[j];
          if (a.has_part) a.part_out_build[pos] = s_part_row;
        } else {
          atomicExch(a.err + 2, 1);
        }
      }
    }
    if (c + (uint32_t)a.split < nchunk) {
      __syncthreads();
      if (tid == 0) s_nq = 0;
      __syncthreads();
    }
  }
}

// Persistent variant for large scans whose first join has a direct-address table: one 1024-thread CTA per SM loops over
// the chunks with an existence filter of the first join's keys in shared memory (FusedScanArgs::bloom).  A row whose key
// misses the filter - most rows of a selective star join - never touches the join table; the rest is the plain kernel.
constexpr int kPersistThreads = 1024;
__global__ void __launch_bounds__(kPersistThreads, 1)
scan_probe_persistent_kernel(const FusedScanArgs a) {
  extern __shared__ __align__(16) uint32_t s_bloom[];
  __shared__ PageView s_view[kMaxChain];
  __shared__ uint16_t q_row[kFusedChunk];
  __shared__ uint32_t q_b0[kFusedChunk];
  __shared__ int s_nq;
  __shared__ int s_part_ok;
  __shared__ uint32_t s_part_row;
  const int tid = threadIdx.x;
  for (uint32_t w = tid; w < a.bloom_words; w += kPersistThreads) s_bloom[w] = __ldg(a.bloom + w);
  const HashTable ht0 = a.key[0].ht;
  for (int cidx = blockIdx.x; cidx < a.nchunks; cidx += gridDim.x) {
    const FusedChunk ch = a.chunks[cidx];
    __syncthreads();                       // the previous chunk's pass B is done with the views and the queues
    if (tid < a.nkeys) s_view[tid] = make_view(a.key[tid].tasks[ch.page]);
    if (tid == 32) {
      s_nq = 0;
      s_part_ok = 1;
      s_part_row = 0;
      if (a.has_part) {
        uint32_t r = 0;
        s_part_ok = ht_find(a.part_ht, a.file_part_val[ch.file], r) ? 1 : 0;
        s_part_row = r;
      }
    }
    __syncthreads();
    if (!s_part_ok) continue;
    // ---- pass A: 1024 threads x 4 consecutive rows = the whole chunk in one step
    {
      const PageView& v = s_view[0];
      const uint32_t rend = min(ch.row0 + (uint32_t)kFusedChunk, v.n);
      const uint32_t base = ch.row0 + (uint32_t)tid * 4u;
      if (base < rend) {
        uint32_t vbits = 0xFu, k0 = base;
        if (v.vr_valid) {
          const uint32_t w = __ldg(v.vr_valid + (base >> 5));
          const uint32_t sh = base & 31u;
          k0 = __ldg(v.vr_rank + (base >> 5)) + __popc(w & ((1u << sh) - 1u));
          vbits = (w >> sh) & 0xFu;
        }
        if (base + 4u > rend) vbits &= (1u << (rend - base)) - 1u;
        if (vbits) {
          int64_t key[4];
          bool okv[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const uint32_t k = min(k0 + __popc(vbits & ((1u << u) - 1u)), v.kmax);
            okv[u] = page_value(v, k, key[u], a.err) && ((vbits >> u) & 1u);
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const uint64_t d = (uint64_t)(key[u] - a.bloom_lo);
            bool cand = okv[u] && d < (uint64_t)a.bloom_range;
            if (cand) {
              const uint32_t bit = (uint32_t)d >> a.bloom_shift;
              cand = (s_bloom[bit >> 5] >> (bit & 31)) & 1u;
            }
            uint32_t b = 0;
            if (cand && ht_find(ht0, key[u], b)) {
              const int pos = atomicAdd(&s_nq, 1);
              q_row[pos] = (uint16_t)(base + u - ch.row0);
              q_b0[pos] = b;
            }
          }
        }
      }
    }
    __syncthreads();
    // ---- pass B: as in the plain kernel
    const int nq = s_nq;
    const int nq_round = (nq + 31) & ~31;
    const uint32_t row_base = (uint32_t)s_view[0].out_row;
    for (int i = tid; i < nq_round; i += kPersistThreads) {
      bool alive = i < nq;
      uint32_t r = 0;
      uint32_t b[kMaxChain];
#pragma unroll
      for (int j = 0; j < kMaxChain; ++j) b[j] = 0;
      if (alive) { r = ch.row0 + q_row[i]; b[0] = q_b0[i]; }
#pragma unroll
      for (int j = 1; j < kMaxChain; ++j) {
        if (j < a.nkeys && alive) {
          uint32_t k;
          int64_t key;
          alive = row_rank(s_view[j], r, k) && page_value(s_view[j], k, key, a.err) && ht_find(a.key[j].ht, key, b[j]);
        }
      }
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, alive);
      if (!m) continue;
      const int lane = tid & 31, leader = __ffs(m) - 1;
      int pos0 = 0;
      if (lane == leader) pos0 = atomicAdd(a.out_count, __popc(m));
      pos0 = __shfl_sync(0xFFFFFFFFu, pos0, leader);
      if (alive) {
        const uint32_t pos = (uint32_t)pos0 + __popc(m & ((1u << lane) - 1u));
        if (pos < a.cap) {
          a.out_row[pos] = row_base + r;
          a.out_file[pos] = ch.file;
#pragma unroll
          for (int j = 0; j < kMaxChain; ++j)
            if (j < a.nkeys) a.key[j].out_build[pos] = b[j];
          if (a.has_part) a.part_out_build[pos] = s_part_row;
        } else {
          atomicExch(a.err + 2, 1);
        }
      }
    }
  }
}

__global__ void bloom_from_dense_kernel(const uint32_t* __restrict__ dense, uint32_t range, uint32_t shift, uint32_t* bloom) {
  for (uint32_t d = blockIdx.x * blockDim.x + threadIdx.x; d < range; d += gridDim.x * blockDim.x)
    if (__ldg(dense + d)) {
      const uint32_t bit = d >> shift;
      atomicOr(&bloom[bit >> 5], 1u << (bit & 31));
    }
}
void launch_bloom_from_dense(const uint32_t* dense, uint32_t range, uint32_t shift, uint32_t* bloom, cudaStream_t s) {
  if (!range) return;
  int g = (int)((range + 255) / 256);
  if (g > 148 * 8) g = 148 * 8;
  bloom_from_dense_kernel<<<g, 256, 0, s>>>(dense, range, shift, bloom);
  NDS_CUDA(cudaGetLastError());
}

void launch_scan_probe(const FusedScanArgs& args, int nchunks, cudaStream_t s) {
  if (nchunks <= 0) return;
  if (args.bloom) {
    static bool attr_set[64] = {false};
    int dev = 0;
    NDS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
      NDS_CUDA(cudaFuncSetAttribute(scan_probe_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBloomMaxWords * 4));
      if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int grid = nchunks < 148 ? nchunks : 148;
    scan_probe_persistent_kernel<<<grid, kPersistThreads, (size_t)args.bloom_words * 4, s>>>(args);
  } else if (args.dmatch) {
    static bool attr_set[64] = {false};
    int dev = 0;
    NDS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
      NDS_CUDA(cudaFuncSetAttribute(scan_probe_dm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDictMatchMaxWords * 4));
      if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    scan_probe_dm_kernel<<<args.npages * args.split, kFusedThreads, (size_t)args.dmatch_stride * 4, s>>>(args);
  } else {
    scan_probe_kernel<<<nchunks, kFusedThreads, 0, s>>>(args);
  }
  NDS_CUDA(cudaGetLastError());
}

// One survivor per thread, every payload column in turn (a handful of columns, a few percent of the rows).
__global__ void __launch_bounds__(256)
gather_decode_kernel(const GatherCol* __restrict__ cols, int ncols, const uint32_t* __restrict__ rows, const uint32_t* __restrict__ files,
                     const int* n_dev, int64_t cap, const int64_t* __restrict__ file_part_val, const uint8_t* __restrict__ file_part_null,
                     int* err) {
  int64_t n = *n_dev;
  if (n > cap) n = cap;
  const int64_t nround = (n + 31) & ~int64_t(31);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
    const bool live = i < n;
    const uint32_t g = live ? rows[i] : 0u;
    const uint32_t f = live ? files[i] : 0u;
    for (int c = 0; c < ncols; ++c) {
      const GatherCol gc = cols[c];
      bool ok = false;
      int64_t val = 0;
      if (live) {
        if (gc.is_part) {
          ok = !file_part_null[f];
          val = file_part_val[f];
        } else {
          // the page of this column that holds global row g: the file's pages are in row order
          int lo = gc.file_page0[f], hi = gc.file_page0[f + 1] - 1;
          while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (gc.tasks[mid].out_row <= (int64_t)g) lo = mid; else hi = mid - 1;
          }
          const PageView v = make_view(gc.tasks[lo]);
          const uint32_t r = (uint32_t)((int64_t)g - v.out_row);
          uint32_t k;
          if (r < v.n && row_rank(v, r, k)) ok = page_value(v, k, val, err);
        }
      }
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, ok);
      if ((threadIdx.x & 31) == 0) gc.out_valid[i >> 5] = m;
      if (live) {
        if (gc.width == 4) reinterpret_cast<int32_t*>(gc.out)[i] = ok ? (int32_t)val : 0;
        else reinterpret_cast<int64_t*>(gc.out)[i] = ok ? val : 0;
      }
    }
  }
}

void launch_gather_decode(const GatherCol* d_cols, int ncols, const uint32_t* rows, const uint32_t* files, const int* n_dev, int64_t cap,
                          const int64_t* file_part_val, const uint8_t* file_part_null, int* err, cudaStream_t s) {
  if (ncols <= 0 || cap <= 0) return;
  int64_t g = (cap + 255) / 256;
  if (g > 148 * 8) g = 148 * 8;
  gather_decode_kernel<<<(int)g, 256, 0, s>>>(d_cols, ncols, rows, files, n_dev, cap, file_part_val, file_part_null, err);
  NDS_CUDA(cudaGetLastError());
}

}  // namespace nds
