// Split page decode for the common page shapes: PLAIN fixed-width pages and dictionary pages whose index
// stream is a few long bit-packed runs - what parquet-mr and Arrow write for high-cardinality columns such as every
// surrogate key and measure of the NDS fact tables.
//
//   def_pages_kernel   one small CTA (64 threads) per nullable page: def levels -> validity bitmap in the output column
//                      plus a page-local copy of the validity words and their rank prefix in scratch memory. The
//                      serial, barrier-separated header resolution lives here, ~10 pages per SM in flight.
//   rows_pages_kernel  one or more CTAs (384 threads) per page: every warp expands rows from the first cycle on -
//                      index bits -> dictionary (staged in shared memory by a bulk async copy) -> store.
//
// Pages that do not fit (RLE-heavy index streams, PLAIN strings, oversized def streams) take decode_pages_kernel.
// Same reference boundary as decode.cu: Spark's Parquet scan behind /root/reference/nds/nds_power.py:97-100.
#include "decode_common.cuh"
#include "ops.cuh"

namespace nds {
using namespace dec;

#ifndef NDS_UNIFORM_TB
#define NDS_UNIFORM_TB 8
#endif
#ifndef NDS_DEF_SEG
#define NDS_DEF_SEG 2048
#endif
#ifndef NDS_DEF_TD
#define NDS_DEF_TD 64
#endif
#ifndef NDS_DEF_MINB
#define NDS_DEF_MINB 13
#endif
#ifndef NDS_ROWS_MINB
#define NDS_ROWS_MINB 3
#endif
namespace {
constexpr int kTD = 64;                  // def kernel threads
constexpr int kTR = 384;                 // rows kernel threads
}  // namespace

// STAGE = bytes of def-level stream a page may have in this instantiation: the 4 KB variant (every page a 20000-row
// writer produces) leaves room for 13 CTAs per SM instead of 10, which is what this latency-bound kernel needs.
template <int STAGE, int MINB, int SEG, int TD>
__global__ void __launch_bounds__(TD, MINB)
def_pages_kernel(const PageTask* __restrict__ tasks, const int* __restrict__ task_idx, int* __restrict__ err) {
  __shared__ uint32_t s_valid[kMaxPageValues / 32];
  constexpr int kEntD = SEG / kChunk;      // entries per resolved segment
  __shared__ uint32_t ent_pos[kEntD + 1];
  __shared__ uint32_t ent_k[kEntD + 1];
  __shared__ __align__(16) uint8_t s_stage[STAGE + 96];
  __shared__ uint32_t s_hop[SEG + SEG / 32 + 2];
  __shared__ uint32_t s_long[64][2];
  __shared__ __align__(8) uint64_t s_mbar;
  __shared__ int s_nlong;
  __shared__ int s_nent;
  __shared__ uint32_t s_warp_tot[TD / 32];
  __shared__ Walker s_walk;

  const PageTask t = tasks[task_idx[blockIdx.x]];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t n = t.num_values;
  const uint32_t nw = (n + 31) >> 5;
  uint32_t* vr_valid = t.vr;
  uint32_t* vr_rank = t.vr + nw;
  const uint32_t sh = (uint32_t)(t.out_row & 31);
  uint32_t* vout = t.valid ? t.valid + (t.out_row >> 5) : nullptr;   // null: only the page-local words are wanted (fused scan)
  const uint32_t nwo = vout ? (n + sh + 31) >> 5 : 0u;

  if (t.def_len == 0) {
    // page without def levels (REQUIRED in this file) of a column that is nullable elsewhere: everything valid
    for (uint32_t j = tid; j < nwo; j += TD) {
      uint32_t v = 0xFFFFFFFFu;
      if (j == 0) v &= 0xFFFFFFFFu << sh;
      uint32_t endbit = n + sh - j * 32;
      if (endbit < 32) v &= (1u << endbit) - 1u;
      if (j == 0 || j == nwo - 1) atomicOr(&vout[j], v);
      else vout[j] = v;
    }
    for (uint32_t w = tid; w < nw; w += TD) {
      const uint32_t rem = n - w * 32;
      vr_valid[w] = rem >= 32 ? 0xFFFFFFFFu : ((1u << rem) - 1u);
      vr_rank[w] = w * 32;
    }
    if (tid == 0) vr_rank[nw] = n;
    return;
  }

  const uint32_t L = t.def_len;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (uint32_t w = tid; w < nw; w += TD) s_valid[w] = 0;
  __syncthreads();
  const uint8_t* g0 = t.src + 4;
  const uintptr_t a0 = reinterpret_cast<uintptr_t>(g0) & ~uintptr_t(15);
  const uintptr_t a1 = (reinterpret_cast<uintptr_t>(g0) + L + 15) & ~uintptr_t(15);
  if (tid == 0) {
    bulk_load_issue(s_stage, reinterpret_cast<const void*>(a0), (uint32_t)(a1 - a0), &s_mbar);
    s_walk.pos = 0;
    s_walk.k = 0;
    s_walk.end = L;
  }
  const uint32_t def_sd = smem_u32(s_stage) + (uint32_t)(reinterpret_cast<uintptr_t>(g0) - a0);
  mbar_wait(&s_mbar, 0);
  __syncthreads();
  for (;;) {
    if (s_walk.pos >= L || s_walk.k >= n) break;
    if (tid == 0) { s_nent = 0; s_nlong = 0; }
    resolve_segment<SEG>(def_sd, L, 1, n, &s_walk, s_hop, ent_pos, ent_k, &s_nent, tid, 0, TD);
    // the entered chunks list the runs that start in them into their own 16 slots of the (now free) hop table -
    // a 32-byte chunk holds at most 16 runs of 2+ bytes - then the listed runs are expanded one thread per run
    const int ne = s_nent;
    const uint32_t seg_end = s_walk.pos;
    if (tid < ne) {
      const uint32_t pe = tid + 1 < ne ? ent_pos[tid + 1] : seg_end;
      uint32_t p = ent_pos[tid], k = ent_k[tid], r = 0;
      while (p < pe && k < n) {
        // def-level header: one varint byte unless the run is longer than 63 values / groups
        uint32_t b0 = lds8(def_sd + p), h = b0 & 0x7F, hl = 1;
        if (b0 & 0x80) {
          const uint32_t b1 = lds8(def_sd + p + 1);
          h |= (b1 & 0x7F) << 7;
          hl = 2;
          if (b1 & 0x80) { h |= (lds8(def_sd + p + 2) & 0x7F) << 14; hl = 3; }
        }
        const bool packed = h & 1;
        const uint32_t cnt = packed ? (h >> 1) * 8 : (h >> 1);
        const uint32_t info = packed ? (((p + hl) << 1) | 1u) : ((lds8(def_sd + p + hl) & 1u) << 1);
        if (r < 16) {
          s_hop[tid * 33 + r * 2] = k;                 // 33-word rows: lanes write different banks
          s_hop[tid * 33 + r * 2 + 1] = (min(cnt, 0xFFFFu) << 16) | (info & 0xFFFFu);
        } else {              // degenerate zero-length runs overflowed the chunk's slots: expand in place
          const uint32_t k1 = min(k + cnt, n);
          if (info & 1) copy_bits(s_valid, def_sd + (info >> 1), nullptr, k, k1);
          else if (info >> 1) set_bits(s_valid, k, k1);
        }
        p += hl + (packed ? (h >> 1) : 1u);
        k += cnt;
        ++r;
      }
      ent_k[tid] = min(r, 16u);       // ent_k is free after phase B: number of listed runs of this entry
    }
    __syncthreads();
    for (uint32_t idx = tid; idx < (uint32_t)ne * 16; idx += TD) {
      if ((idx & 15) >= ent_k[idx >> 4]) continue;
      const uint32_t hx = (idx >> 4) * 33 + (idx & 15) * 2;
      const uint32_t k = s_hop[hx], ci = s_hop[hx + 1];
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
    __syncthreads();
    const int nl = min(s_nlong, 64);
    for (int j = 0; j < nl; ++j) {
      const uint32_t b0 = s_long[j][0], b1 = s_long[j][1];
      for (uint32_t w = (b0 >> 5) + tid; w <= ((b1 - 1) >> 5); w += TD) {
        const uint32_t lo = max(b0, w << 5), hi = min(b1, (w + 1) << 5);
        const uint32_t cntb = hi - lo;
        atomicOr(&s_valid[w], (cntb == 32 ? 0xFFFFFFFFu : ((1u << cntb) - 1u)) << (lo & 31));
      }
    }
    __syncthreads();
  }
  // rank prefix per word
  {
    const uint32_t per = (nw + TD - 1) / TD;
    const uint32_t w0 = min((uint32_t)tid * per, nw), w1 = min(w0 + per, nw);
    uint32_t sum = 0;
    for (uint32_t w = w0; w < w1; ++w) sum += __popc(s_valid[w]);
    uint32_t incl = sum;
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if (lane >= d) incl += v;
    }
    __syncthreads();          // s_warp_tot is reused
    if (lane == 31) s_warp_tot[warp] = incl;
    __syncthreads();
    uint32_t run = incl - sum;
    for (int q = 0; q < warp; ++q) run += s_warp_tot[q];
    for (uint32_t w = w0; w < w1; ++w) {
      vr_rank[w] = run;
      vr_valid[w] = s_valid[w];
      run += __popc(s_valid[w]);
    }
    if (tid == 0) {
      uint32_t tot = 0;
      for (int q = 0; q < TD / 32; ++q) tot += s_warp_tot[q];
      vr_rank[nw] = tot;
    }
  }
  // validity bitmap of the output column (the page may start at any bit)
  for (uint32_t j = tid; j < nwo; j += TD) {
    uint32_t lo = j < nw ? s_valid[j] : 0u;
    uint32_t hi = (j > 0 && j - 1 < nw) ? s_valid[j - 1] : 0u;
    uint32_t v = sh ? ((lo << sh) | (hi >> (32 - sh))) : lo;
    if (j == 0 || j == nwo - 1) { if (v) atomicOr(&vout[j], v); }
    else vout[j] = v;
  }
}

namespace {

struct RowsCtx {
  const uint32_t* s_v;      // validity words (shared)
  const uint32_t* s_r;      // rank prefix words, s_r[nw] = non-null count (shared)
  const uint32_t* ek;       // run start value indices (shared), ek[ne] = end
  const uint32_t* ep;       // run infos
  int ne;
  uint32_t n, nw;
  uint32_t bw, bwmask;
  const uint32_t* vwords;   // 4 B aligned base of the index stream
  uint32_t vs_mis;          // misalignment of the stream start, in bits
  uint32_t dict_s;          // shared-window address of the staged dictionary (0: gather from global)
  const void* dict;
  uint32_t dict_count;
  void* out;
  int64_t orow;
  int* err;
};

template <int W>
__device__ __forceinline__ void store_out(void* out, int64_t row, int64_t v) {
  if (W == 4) reinterpret_cast<int32_t*>(out)[row] = (int32_t)v;
  else reinterpret_cast<int64_t*>(out)[row] = v;
}

// One warp expands tiles [tile0, tile1) of a dictionary page, 8 tiles (256 rows) per iteration.
template <int W, bool DEF, bool SDICT>
__device__ __forceinline__ void expand_dict(const RowsCtx& c, uint32_t tile0, uint32_t tile1, int lane) {
  const uint32_t lt_mask = (1u << lane) - 1u;
  int ce = 0;
  for (uint32_t bb = tile0; bb < tile1; bb += 8) {
    const uint32_t be = min(bb + 8u, tile1);
    // runs: cursor at the run holding the block's first value; next1/next2 = starts of the following two runs
    uint32_t rk1 = 0, info1 = 0, next1 = 0xFFFFFFFFu, info2 = 0, next2 = 0xFFFFFFFFu;
    if (c.bw && c.ne) {
      const uint32_t kA = DEF ? c.s_r[bb] : bb * 32;
      if (bb == tile0) {                 // first block of this warp: binary search, afterwards the cursor creeps forward
        int lo = 0, hi = c.ne - 1;
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if (c.ek[mid] <= kA) lo = mid; else hi = mid - 1;
        }
        ce = lo;
      }
      // advance the cursor: all lanes look at the next 32 run starts at once
      for (;;) {
        const int e = ce + 1 + lane;
        const int adv = __popc(__ballot_sync(0xFFFFFFFFu, e < c.ne && c.ek[e] <= kA));
        ce += adv;
        if (adv < 32) break;
      }
      rk1 = c.ek[ce];
      info1 = c.ep[ce];
      if (ce + 1 < c.ne) { next1 = c.ek[ce + 1]; info2 = c.ep[ce + 1]; }
      if (ce + 2 < c.ne) next2 = c.ek[ce + 2];
    }
    const uint32_t kB = DEF ? c.s_r[be] : min(be * 32, c.n);
    const bool two_packed = c.bw && c.ne && (info1 & 1) && (next1 == 0xFFFFFFFFu || (info2 & 1)) && kB <= next2;
    if (two_packed && be == bb + 8 && be * 32 <= c.n && c.dict_count) {
      // The common block: 256 rows, every value in one of two bit-packed runs (bit position = base(run) + k * bw).
      // Straight-line code, no per-row state: null lanes fetch the index of the next non-null value (in range, the
      // page is padded) and store 0.
      const uint32_t base1 = (info1 >> 1) * 8u + c.vs_mis - rk1 * c.bw;
      const uint32_t base2 = (info2 >> 1) * 8u + c.vs_mis - next1 * c.bw;
      const uint32_t kmax = (DEF ? c.s_r[c.nw] : c.n) - 1u;      // last value index of the page (block is non-empty)
      uint32_t ix[8];
      bool okv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t tcur = bb + u;
        uint32_t k = tcur * 32 + lane;
        okv[u] = true;
        if (DEF) {
          const uint32_t wv = c.s_v[tcur];
          k = min(c.s_r[tcur] + __popc(wv & lt_mask), kmax);
          okv[u] = (wv >> lane) & 1u;
        }
        const uint32_t bit = (k >= next1 ? base2 : base1) + k * c.bw;
        const uint32_t* wp = c.vwords + (bit >> 5);
        ix[u] = __funnelshift_r(__ldg(wp), __ldg(wp + 1), bit & 31) & c.bwmask;
      }
      bool bad = false;
      const int64_t row_base = c.orow + (int64_t)bb * 32 + lane;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        bad |= okv[u] && ix[u] >= c.dict_count;
        const uint32_t j = min(ix[u], c.dict_count - 1u);
        int64_t v;
        if (SDICT) v = W == 4 ? (int64_t)(int32_t)lds32(c.dict_s + j * 4u) : lds64(c.dict_s + j * 8u);
        else v = W == 4 ? (int64_t)__ldg(reinterpret_cast<const int32_t*>(c.dict) + j) : __ldg(reinterpret_cast<const int64_t*>(c.dict) + j);
        store_out<W>(c.out, row_base + u * 32, okv[u] ? v : 0);
      }
      if (bad && c.err) atomicExch(c.err, 6);
      continue;
    }
    uint32_t w0[8], w1[8], shv[8];
    int stt[8];          // 0 skip, 1 zero, 2 index in w0, 3 bit-packed words in w0/w1
    if (two_packed) {
      // partial block (page tail): same addressing, bounds checked per row
      const uint32_t base1 = (info1 >> 1) * 8u + c.vs_mis - rk1 * c.bw;
      const uint32_t base2 = (info2 >> 1) * 8u + c.vs_mis - next1 * c.bw;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t tcur = bb + u;
        stt[u] = 0;
        w0[u] = 0;
        w1[u] = 0;
        shv[u] = 0;
        if (tcur < be && tcur * 32 + lane < c.n) {
          uint32_t k = tcur * 32 + lane;
          bool ok = true;
          if (DEF) {
            const uint32_t wv = c.s_v[tcur];
            k = c.s_r[tcur] + __popc(wv & lt_mask);
            ok = (wv >> lane) & 1u;
          }
          stt[u] = 1;
          if (ok) {
            const uint32_t bit = (k >= next1 ? base2 : base1) + k * c.bw;
            w0[u] = __ldg(c.vwords + (bit >> 5));
            w1[u] = __ldg(c.vwords + (bit >> 5) + 1);
            shv[u] = bit & 31;
            stt[u] = 3;
          }
        }
      }
    } else {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const uint32_t tcur = bb + u;
      const uint32_t i = tcur * 32 + lane;
      stt[u] = 0;
      w0[u] = 0;
      w1[u] = 0;
      shv[u] = 0;
      if (tcur < be && i < c.n) {
        uint32_t k = i;
        bool ok = true;
        if (DEF) {
          const uint32_t wv = c.s_v[tcur];
          k = c.s_r[tcur] + __popc(wv & lt_mask);
          ok = (wv >> lane) & 1u;
        }
        stt[u] = 1;
        if (ok) {
          uint32_t rk = rk1, info = info1;
          if (k >= next1) {
            rk = next1;
            info = info2;
            if (k >= next2) {            // rare: more than two runs under one block
              int e = ce + 2;
              while (e + 1 < c.ne && c.ek[e + 1] <= k) ++e;
              rk = c.ek[e];
              info = c.ep[e];
            }
          }
          if (!c.bw || !(info & 1)) { w0[u] = c.bw ? (info >> 1) : 0u; stt[u] = 2; }
          else {
            const uint32_t bit = (info >> 1) * 8u + (k - rk) * c.bw + c.vs_mis;
            w0[u] = __ldg(c.vwords + (bit >> 5));
            w1[u] = __ldg(c.vwords + (bit >> 5) + 1);
            shv[u] = bit & 31;
            stt[u] = 3;
          }
        }
      }
    }
    }
    int64_t vv[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      vv[u] = 0;
      if (stt[u] >= 2) {
        const uint32_t ix = stt[u] == 3 ? (__funnelshift_r(w0[u], w1[u], shv[u]) & c.bwmask) : w0[u];
        if (ix >= c.dict_count) { if (c.err) atomicExch(c.err, 6); continue; }
        if (SDICT) vv[u] = W == 4 ? (int64_t)(int32_t)lds32(c.dict_s + ix * 4u) : lds64(c.dict_s + ix * 8u);
        else vv[u] = W == 4 ? (int64_t)__ldg(reinterpret_cast<const int32_t*>(c.dict) + ix) : __ldg(reinterpret_cast<const int64_t*>(c.dict) + ix);
      }
    }
    const int64_t row_base = c.orow + (int64_t)bb * 32 + lane;
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (stt[u]) store_out<W>(c.out, row_base + u * 32, vv[u]);
  }
}

// The index stream every writer produces for a high-cardinality column is UNIFORM: bit-packed runs of 63 groups (504
// values, one header byte 0x7F) back to back, the last one shorter. Then value k sits at bit k * bw + 8 * (k / 504 + 1)
// of the stream and no run cursor is needed at all: full 256-row blocks are straight-line code (null lanes fetch the
// index of a neighbouring value, so the range check needs no validity), the page tail takes the general path above.
template <int W, bool DEF, bool SDICT>
__device__ __forceinline__ void expand_dict_uniform(const RowsCtx& c, uint32_t tile0, uint32_t tile1, int lane) {
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t kmax = (DEF ? c.s_r[c.nw] : c.n) - 1u;
  const uint32_t dmax = c.dict_count - 1u;
  const uint32_t off8 = c.vs_mis + 8u;
  uint32_t mx = 0;
  constexpr int TB = NDS_UNIFORM_TB;      // tiles per iteration: 2 * TB index loads in flight per thread
  uint32_t bb = tile0;
  if (kmax != 0xFFFFFFFFu) {
    for (; bb + TB <= tile1 && (bb + TB) * 32 <= c.n; bb += TB) {
      uint32_t ix[TB];
      bool okv[TB];
#pragma unroll
      for (int u = 0; u < TB; ++u) {
        const uint32_t tcur = bb + u;
        uint32_t k = tcur * 32 + lane;
        okv[u] = true;
        if (DEF) {
          const uint32_t wv = c.s_v[tcur];
          k = min(c.s_r[tcur] + __popc(wv & lt_mask), kmax);
          okv[u] = (wv >> lane) & 1u;
        }
        const uint32_t bit = k * c.bw + (k / 504u) * 8u + off8;
        const uint32_t* wp = c.vwords + (bit >> 5);
        ix[u] = __funnelshift_r(__ldg(wp), __ldg(wp + 1), bit & 31) & c.bwmask;
      }
      const int64_t row_base = c.orow + (int64_t)bb * 32 + lane;
#pragma unroll
      for (int u = 0; u < TB; ++u) {
        mx = max(mx, ix[u]);
        const uint32_t j = min(ix[u], dmax);
        int64_t v;
        if (SDICT) v = W == 4 ? (int64_t)(int32_t)lds32(c.dict_s + j * 4u) : lds64(c.dict_s + j * 8u);
        else v = W == 4 ? (int64_t)__ldg(reinterpret_cast<const int32_t*>(c.dict) + j) : __ldg(reinterpret_cast<const int64_t*>(c.dict) + j);
        store_out<W>(c.out, row_base + u * 32, okv[u] ? v : 0);
      }
    }
    if (mx > dmax && c.err) atomicExch(c.err, 6);
  }
  if (bb < tile1) expand_dict<W, DEF, SDICT>(c, bb, tile1, lane);
}

template <int W, bool DEF>
__device__ __forceinline__ void expand_plain(const RowsCtx& c, const uint8_t* vsrc, uint32_t tile0, uint32_t tile1, int lane) {
  const uint32_t lt_mask = (1u << lane) - 1u;
  for (uint32_t bb = tile0; bb < tile1; bb += 8) {
    const uint32_t be = min(bb + 8u, tile1);
    int64_t vv[8];
    bool wr[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const uint32_t tcur = bb + u;
      const uint32_t i = tcur * 32 + lane;
      vv[u] = 0;
      wr[u] = tcur < be && i < c.n;
      if (wr[u]) {
        uint32_t k = i;
        bool ok = true;
        if (DEF) {
          const uint32_t wv = c.s_v[tcur];
          k = c.s_r[tcur] + __popc(wv & lt_mask);
          ok = (wv >> lane) & 1u;
        }
        if (ok) vv[u] = W == 4 ? (int64_t)__ldg(reinterpret_cast<const int32_t*>(vsrc) + k) : __ldg(reinterpret_cast<const int64_t*>(vsrc) + k);
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (wr[u]) store_out<W>(c.out, c.orow + (bb + u) * 32 + lane, vv[u]);
  }
}

}  // namespace

__global__ void __launch_bounds__(kTR, NDS_ROWS_MINB)
rows_pages_kernel(const PageTask* __restrict__ tasks, const int* __restrict__ task_idx, int nslices, int* __restrict__ err) {
  __shared__ uint32_t s_v[kMaxPageValues / 32];
  __shared__ uint32_t s_r[kMaxPageValues / 32 + 1];
  __shared__ uint32_t early_k[kEarlyCap + 1];
  __shared__ uint32_t early_info[kEarlyCap];
  __shared__ __align__(8) uint64_t s_mbar_dict;
  __shared__ int s_early_n;
  __shared__ int s_uniform;
  extern __shared__ __align__(16) uint8_t s_dict[];

  const PageTask t = tasks[task_idx[blockIdx.x / nslices]];
  const int slice = blockIdx.x % nslices;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t n = t.num_values;
  const uint32_t nw = (n + 31) >> 5;
  const bool has_def = t.vr != nullptr;
  const uint32_t values_off = t.def_len ? 4 + t.def_len : 0;
  const bool is_dict = t.enc == 1;
  const uint32_t bw = t.bw;
  const uint32_t VL = is_dict ? t.size - (values_off + 1) : 0;
  const uint8_t* vstream = t.src + values_off + 1;
  const uint32_t dict_bytes = is_dict ? t.dict_count * (uint32_t)t.width : 0u;
  const bool dict_staged = is_dict && dict_bytes > 0 && dict_bytes <= (uint32_t)kDictStage;

  if (tid == 0) {
    s_early_n = 0;
    s_uniform = 0;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_mbar_dict)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (dict_staged && tid == 0) bulk_load_issue(s_dict, t.dict, (dict_bytes + 15u) & ~15u, &s_mbar_dict);
  // this slice's share of the page, in 32-row tiles
  const uint32_t ntiles = nw;
  const uint32_t tiles_per_slice = (ntiles + nslices - 1) / nslices;
  const uint32_t stile0 = min((uint32_t)slice * tiles_per_slice, ntiles), stile1 = min(stile0 + tiles_per_slice, ntiles);
  if (is_dict && bw > 0 && bw <= 32 && warp == kTR / 32 - 1) {
    Walker w;
    w.pos = 0;
    w.end = VL;
    w.k = 0;
    int nr = walk_runs_warp(vstream, w, bw, n, kEarlyCap, 1 << 30, early_k, early_info);
    __syncwarp();
    bool uni = nr > 0;
    for (int e = lane; e < nr; e += 32)
      uni = uni && early_k[e] == 504u * (uint32_t)e && early_info[e] == ((((uint32_t)e * (1u + 63u * bw) + 1u) << 1) | 1u);
    uni = __all_sync(0xFFFFFFFFu, uni);
    if (lane == 0) { s_early_n = (w.pos >= VL || w.k >= n) ? nr : -1; s_uniform = uni ? 1 : 0; }
  } else if (has_def) {
    // validity + rank words of my tiles (and the sentinel) from the def kernel's scratch
    const int nthr = (is_dict && bw > 0 && bw <= 32) ? kTR - 32 : kTR;
    for (uint32_t w = stile0 + tid; w < stile1; w += nthr) { s_v[w] = t.vr[w]; s_r[w] = t.vr[nw + w]; }
    if (tid == 0) { s_r[stile1] = t.vr[nw + stile1]; s_r[nw] = t.vr[nw + nw]; }
  }
  __syncthreads();
  if (dict_staged) mbar_wait(&s_mbar_dict, 0);
  if (bw > 32 || s_early_n < 0) { if (tid == 0 && err) atomicExch(err, s_early_n < 0 ? 7 : 5); return; }

  RowsCtx c;
  c.s_v = s_v;
  c.s_r = s_r;
  c.ek = early_k;
  c.ep = early_info;
  c.ne = s_early_n;
  c.n = n;
  c.nw = nw;
  c.bw = bw;
  c.bwmask = bw >= 32 ? 0xFFFFFFFFu : ((1u << bw) - 1u);
  c.vwords = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(vstream) & ~uintptr_t(3));
  c.vs_mis = (uint32_t)(reinterpret_cast<uintptr_t>(vstream) & 3) * 8;
  c.dict_s = smem_u32(s_dict);
  c.dict = t.dict;
  c.dict_count = t.dict_count;
  c.out = t.out;
  c.orow = t.out_row;
  c.err = err;
  // contiguous block of tiles per warp
  const uint32_t tiles_per_warp = (stile1 - stile0 + kTR / 32 - 1) / (kTR / 32);
  const uint32_t tile0 = min(stile0 + (uint32_t)warp * tiles_per_warp, stile1), tile1 = min(tile0 + tiles_per_warp, stile1);
  if (!is_dict) {
    const uint8_t* vsrc = t.src + values_off;
    if (t.width == 4) { if (has_def) expand_plain<4, true>(c, vsrc, tile0, tile1, lane); else expand_plain<4, false>(c, vsrc, tile0, tile1, lane); }
    else { if (has_def) expand_plain<8, true>(c, vsrc, tile0, tile1, lane); else expand_plain<8, false>(c, vsrc, tile0, tile1, lane); }
    return;
  }
  if (s_uniform && t.dict_count && dict_staged) {        // the common case gets the cursor-free expansion
    if (t.width == 4) { if (has_def) expand_dict_uniform<4, true, true>(c, tile0, tile1, lane); else expand_dict_uniform<4, false, true>(c, tile0, tile1, lane); }
    else { if (has_def) expand_dict_uniform<8, true, true>(c, tile0, tile1, lane); else expand_dict_uniform<8, false, true>(c, tile0, tile1, lane); }
    return;
  }
  if (t.width == 4) {
    if (has_def) { if (dict_staged) expand_dict<4, true, true>(c, tile0, tile1, lane); else expand_dict<4, true, false>(c, tile0, tile1, lane); }
    else { if (dict_staged) expand_dict<4, false, true>(c, tile0, tile1, lane); else expand_dict<4, false, false>(c, tile0, tile1, lane); }
  } else {
    if (has_def) { if (dict_staged) expand_dict<8, true, true>(c, tile0, tile1, lane); else expand_dict<8, true, false>(c, tile0, tile1, lane); }
    else { if (dict_staged) expand_dict<8, false, true>(c, tile0, tile1, lane); else expand_dict<8, false, false>(c, tile0, tile1, lane); }
  }
}

int rows_slices(int npages) {
  // few pages: several CTAs per page so that the launch still covers the 148 SMs three times
  int ns = 1;
  if (npages > 0 && npages < 3 * 148) {
    ns = (3 * 148 + npages - 1) / npages;
    if (ns > 16) ns = 16;
  }
  return ns;
}

void launch_decode_split(const PageTask* d_tasks, const int* d_def_idx, int ndef, int max_def_len, const int* d_rows_idx, int nrows,
                         int* d_err, cudaStream_t stream, int* launches) {
  if (ndef > 0) {
    static_assert(NDS_DEF_SEG / 32 <= NDS_DEF_TD, "one thread per 32-byte chunk of a segment");
    if (max_def_len <= 4096) def_pages_kernel<4096, NDS_DEF_MINB, NDS_DEF_SEG, NDS_DEF_TD><<<ndef, NDS_DEF_TD, 0, stream>>>(d_tasks, d_def_idx, d_err);
    else def_pages_kernel<8192, 8, kDefSeg, kTD><<<ndef, kTD, 0, stream>>>(d_tasks, d_def_idx, d_err);
    NDS_CUDA(cudaGetLastError());
    ++*launches;
  }
  if (nrows > 0) {
    static bool attr_set[64] = {false};
    int dev = 0;
    NDS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
      NDS_CUDA(cudaFuncSetAttribute(rows_pages_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDictStage));
      if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int ns = rows_slices(nrows);
    rows_pages_kernel<<<nrows * ns, kTR, kDictStage, stream>>>(d_tasks, d_rows_idx, ns, d_err);
    NDS_CUDA(cudaGetLastError());
    ++*launches;
  }
}

}  // namespace nds
