// Device helpers shared by the page-decode kernels (decode.cu: general kernel; decode_fast.cu: split def/rows kernels).
#pragma once
#include "device_types.h"
#include "common.h"

namespace nds {
namespace dec {

constexpr int kWords = kMaxPageValues / 32;
constexpr int kStage = 16384;            // bytes of one hybrid stream staged in shared memory
constexpr int kEntCap = 1024;            // header entries (position, first value index) per batch
constexpr int kDefSeg = 2048;            // byte positions resolved per segment of the parallel header walk
// The hop table has one word per byte position and is written by one thread per 32-position chunk: stored plainly,
// the 32 lanes of a warp would hit one bank on every access. One pad word per 32 positions spreads them over all banks.
constexpr int kHopWords = kDefSeg + kDefSeg / 32 + 2;
__device__ __forceinline__ uint32_t hop_ix(uint32_t x) { return x + (x >> 5); }
constexpr int kChunk = 32;               // bytes per chunk of the parallel header walk
constexpr int kUnroll = 4;
constexpr int kDictStage = 65536;        // dictionary bytes staged in (dynamic) shared memory; larger ones are gathered from L2
constexpr int kEarlyCap = 80;            // runs the early index-header chase may record (32768 / 504 = 65 full runs)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t lds8(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}

__device__ __forceinline__ void lds128(uint32_t a, uint32_t& x, uint32_t& y, uint32_t& z, uint32_t& w) {
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "r"(a));
}
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ int64_t lds64(uint32_t a) {
  long long v;
  asm volatile("ld.shared.s64 %0, [%1];" : "=l"(v) : "r"(a));
  return (int64_t)v;
}
// bw bits starting `bit` bits after the 4 B aligned shared address base (two aligned word loads + funnel shift)
__device__ __forceinline__ uint32_t smem_bits(uint32_t base_al, uint32_t bit, uint32_t mask) {
  const uint32_t a = base_al + ((bit >> 5) << 2);
  return __funnelshift_r(lds32(a), lds32(a + 4), bit & 31) & mask;
}

__device__ __forceinline__ uint32_t load_u32_unaligned(const uint8_t* p) {
  return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}

struct Walker {       // header walker state
  uint32_t pos;       // byte position inside the stream
  uint32_t end;       // end of the stream
  uint32_t k;         // values covered so far
};

// One header at global src[pos]: returns the run's byte length (header included); cnt = values,
// info = (payload position << 1) | 1 for bit-packed, (value << 1) for RLE.
__device__ __forceinline__ uint32_t parse_run_g(const uint8_t* src, uint32_t pos, uint32_t end, uint32_t bw, uint32_t& cnt,
                                                uint32_t& info) {
  uint32_t h = 0, s = 0, c, p = pos;
  do {
    c = src[p++];
    h |= (c & 0x7F) << s;
    s += 7;
  } while ((c & 0x80) && s < 35 && p < end);
  if (h & 1) {
    info = (p << 1) | 1u;
    cnt = (h >> 1) * 8;
    return (p - pos) + (h >> 1) * bw;
  }
  const uint32_t bytew = (bw + 7) >> 3;
  uint32_t v = 0;
  for (uint32_t b = 0; b < bytew; ++b) v |= (uint32_t)src[p + b] << (8 * b);
  info = v << 1;
  cnt = h >> 1;
  return (p - pos) + bytew;
}

// Same, for a stream staged in shared memory (sd = shared-window address of stream byte 0).
__device__ __forceinline__ uint32_t parse_run_s(uint32_t sd, uint32_t pos, uint32_t end, uint32_t bw, uint32_t& cnt, uint32_t& info) {
  uint32_t h = 0, s = 0, c, p = pos;
  do {
    c = lds8(sd + p);
    ++p;
    h |= (c & 0x7F) << s;
    s += 7;
  } while ((c & 0x80) && s < 35 && p < end);
  if (h & 1) {
    info = (p << 1) | 1u;
    cnt = (h >> 1) * 8;
    return (p - pos) + (h >> 1) * bw;
  }
  const uint32_t bytew = (bw + 7) >> 3;
  uint32_t v = 0;
  for (uint32_t b = 0; b < bytew; ++b) v |= lds8(sd + p + b) << (8 * b);
  info = v << 1;
  cnt = h >> 1;
  return (p - pos) + bytew;
}

// Sequential walk of up to kEntCap headers from global memory (fallback for oversized def streams; one thread).
__device__ inline int walk_runs_seq(const uint8_t* src, Walker& w, uint32_t bw, uint32_t limit, uint32_t* r_start, uint32_t* r_info) {
  int n = 0;
  while (n < kEntCap && w.k < limit && w.pos < w.end) {
    uint32_t cnt, info;
    uint32_t len = parse_run_g(src, w.pos, w.end, bw, cnt, info);
    r_start[n] = w.k;
    r_info[n] = info;
    w.pos += len;
    w.k += cnt;
    ++n;
  }
  if (w.k > limit) w.k = limit;
  r_start[n] = w.k;
  return n;
}

// Index-stream headers, one warp: lanes load the 32 positions where the next headers sit if the runs are full
// 63-group bit-packed runs (what parquet-mr and Arrow emit for long stretches) and verify them with a ballot;
// the run after the verified ones is parsed from the byte its lane already holds when it is a short bit-packed
// run, otherwise by one ordinary parse.  All lanes return the same run count; r_start[n] is the end sentinel.
__device__ inline int walk_runs_warp(const uint8_t* src, Walker& w, uint32_t bw, uint32_t limit, int cap, int max_single,
                              uint32_t* r_start, uint32_t* r_info) {
  const int lane = threadIdx.x & 31;
  const uint32_t S = 1 + 63 * bw;
  int n = 0, singles = 0;
  uint32_t pos = w.pos, k = w.k;
  while (n < cap && k < limit && pos < w.end && singles <= max_single) {
    const uint32_t cand = pos + (uint32_t)lane * S;
    const uint32_t byte = cand < w.end ? (uint32_t)src[cand] : 0x100u;
    const bool full = byte == 0x7Fu && cand + S <= w.end;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, full);
    uint32_t lead = (m == 0xFFFFFFFFu) ? 32u : (uint32_t)(__ffs(~m) - 1);
    lead = min(lead, min((uint32_t)(cap - n), (limit - k + 503u) / 504u));
    if (lead > 0) {
      if ((uint32_t)lane < lead) {
        r_start[n + lane] = k + (uint32_t)lane * 504u;
        r_info[n + lane] = ((cand + 1) << 1) | 1u;
      }
      n += (int)lead;
      k += lead * 504u;
      pos += lead * S;
      if (n >= cap || k >= limit || pos >= w.end || lead >= 32) continue;
      // the lane after the verified runs already holds the next header byte: short bit-packed run?
      const uint32_t nb = __shfl_sync(0xFFFFFFFFu, byte, (int)lead);
      if (nb < 0x80u && (nb & 1u)) {
        const uint32_t groups = nb >> 1;
        if (lane == 0) {
          r_start[n] = k;
          r_info[n] = ((pos + 1) << 1) | 1u;
        }
        ++n;
        k += groups * 8;
        pos += 1 + groups * bw;
      }
    } else {
      uint32_t cnt = 0, info = 0, len = 0;
      if (lane == 0) {
        len = parse_run_g(src, pos, w.end, bw, cnt, info);
        r_start[n] = k;
        r_info[n] = info;
      }
      len = __shfl_sync(0xFFFFFFFFu, len, 0);
      cnt = __shfl_sync(0xFFFFFFFFu, cnt, 0);
      pos += len;
      k += cnt;
      ++n;
      ++singles;     // each of these costs dependent global loads: a stream of short runs is given up on early
    }
  }
  if (k > limit) k = limit;
  if (lane == 0) r_start[n] = k;
  w.pos = pos;
  w.k = k;
  return n;
}

__device__ __forceinline__ void set_bits(uint32_t* bm, uint32_t b0, uint32_t b1) {   // [b0, b1) <- 1
  uint32_t b = b0;
  while (b < b1) {
    uint32_t wd = b >> 5, lo = b & 31;
    uint32_t cnt = min(32u - lo, b1 - b);
    uint32_t m = (cnt == 32 ? 0xFFFFFFFFu : ((1u << cnt) - 1u)) << lo;
    atomicOr(&bm[wd], m);
    b += cnt;
  }
}

// bit-packed bw=1 payload bytes are validity bits for rows [b0, b1); payload in shared (sp) or global (gp) memory
__device__ __forceinline__ void copy_bits(uint32_t* bm, uint32_t sp, const uint8_t* gp, uint32_t b0, uint32_t b1) {
  for (uint32_t b = b0; b < b1; b += 8) {
    uint32_t byte = gp ? (uint32_t)gp[(b - b0) >> 3] : lds8(sp + ((b - b0) >> 3));
    uint32_t rem = b1 - b;
    if (rem < 8) byte &= (1u << rem) - 1u;
    if (!byte) continue;
    uint32_t lo = b & 31;
    atomicOr(&bm[b >> 5], byte << lo);
    if (lo > 24 && (byte >> (32 - lo))) atomicOr(&bm[(b >> 5) + 1], byte >> (32 - lo));
  }
}

// ---- bulk async copy (TMA 1-D) of a 16 B aligned window into shared memory -----------------------------
__device__ __forceinline__ void bulk_load_issue(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* mbar) {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // order earlier generic accesses to smem_dst first
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mbar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(mbar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* mbar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(mbar)), "r"(parity)
        : "memory");
  }
}

// CTA-wide barrier over the first `nthr` threads (nthr multiple of 32); id 0 with nthr == blockDim is __syncthreads()
__device__ __forceinline__ void bar_sync(int id, int nthr) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthr) : "memory"); }

// ---- parallel resolution of a staged hybrid stream's run headers -------------------------------------------------
// sd: shared-window address of stream byte 0, L bytes long. Starting at header position w->pos with w->k values
// covered, resolves up to kDefSeg byte positions: appends (position, first value index) of the first header of every
// entered kChunk-byte chunk to ent_pos/ent_k (at most kDefSeg/kChunk new entries) and advances *w.
// SEG (default kDefSeg) = positions per call; s_hop needs SEG + SEG / 32 + 2 words (kHopWords for the default).
// Called by the first bar_n threads of the CTA (they synchronise on named barrier bar_id); s_hop is kHopWords words of scratch;
// the staging buffer behind sd must be readable up to 64 bytes past the stream's end.
template <int SEG = kDefSeg>
__device__ __forceinline__ void resolve_segment(uint32_t sd, uint32_t L, uint32_t bw, uint32_t limit, Walker* w, uint32_t* s_hop,
                                                uint32_t* ent_pos, uint32_t* ent_k, int* n_ent, int tid, int bar_id, int bar_n,
                                                long long* dbg = nullptr) {
  const uint32_t seg0 = w->pos, k0 = w->k;
  const uint32_t seg1 = min(seg0 + (uint32_t)SEG, L);
  const uint32_t nchunks = (seg1 - seg0 + kChunk - 1) / kChunk;
  const uint32_t bytew = (bw + 7) >> 3;
  bar_sync(bar_id, bar_n);
  // phase A: speculative parse of every byte position of my chunk, last to first. The chunk's bytes (+2 of look-ahead)
  // are loaded up front so that the per-position chain is ALU + one hop-table load.
  if ((uint32_t)tid < nchunks) {
    const uint32_t cb = seg0 + tid * kChunk, ce = min(cb + (uint32_t)kChunk, seg1);
    // The chunk's 34 bytes come in as four aligned 16-byte loads (byte loads at a 32-byte stride between lanes are
    // 8-way bank conflicts, 34 times over); the stream's misalignment m - the same for every lane - is removed with a
    // barrel shift over the 16 words, after which every byte sits at a compile-time register and shift.
    uint32_t A[17];
    {
      const uint32_t a0 = (sd + cb) & ~15u, m = (sd + cb) & 15u;
      lds128(a0, A[0], A[1], A[2], A[3]);
      lds128(a0 + 16, A[4], A[5], A[6], A[7]);
      lds128(a0 + 32, A[8], A[9], A[10], A[11]);
      lds128(a0 + 48, A[12], A[13], A[14], A[15]);
      A[16] = 0;
      if (m & 8) {
#pragma unroll
        for (int i = 0; i < 15; ++i) A[i] = A[i + 2];
      }
      if (m & 4) {
#pragma unroll
        for (int i = 0; i < 14; ++i) A[i] = A[i + 1];
      }
      const uint32_t sh = (m & 3u) * 8u;
#pragma unroll
      for (int i = 0; i < 9; ++i) A[i] = __funnelshift_r(A[i], A[i + 1], sh);
    }
    uint32_t by[kChunk + 2];
#pragma unroll
    for (int j = 0; j < kChunk + 2; ++j) by[j] = cb + j < L ? ((A[j >> 2] >> ((j & 3) * 8)) & 0xFFu) : 0u;
#pragma unroll
    for (int j = kChunk - 1; j >= 0; --j) {
      const uint32_t p = cb + j;
      if (p < ce) {
        const uint32_t b0 = by[j], b1 = by[j + 1], b2 = by[j + 2];
        uint32_t h = b0 & 0x7F, hl = 1;
        if (b0 & 0x80) {
          h |= (b1 & 0x7F) << 7;
          hl = 2;
          if (b1 & 0x80) {
            h |= (b2 & 0x7F) << 14;
            hl = 3;     // a run of one page never needs more than 3 varint bytes (<= 32768 values)
          }
        }
        uint32_t cnt = (h & 1) ? (h >> 1) * 8 : (h >> 1);
        uint32_t nxt = p + hl + ((h & 1) ? (h >> 1) * bw : bytew);
        if (nxt > L || nxt <= p) nxt = L;
        if (nxt < ce) {
          uint32_t f = s_hop[hop_ix(nxt - seg0)];
          cnt += f & 0xFFFFu;
          nxt = f >> 16;
        }
        if (cnt > 0xFFFFu) cnt = 0xFFFFu;
        s_hop[hop_ix(p - seg0)] = (nxt << 16) | cnt;
      }
    }
  }
  bar_sync(bar_id, bar_n);
  if (dbg && tid == 0 && seg0 == 0) dbg[(int64_t)blockIdx.x * 8 + 6] = clock64();
  // phase B: hop from chunk to chunk, one shared-memory load per chunk
  if (tid == 0) {
    uint32_t p = seg0, k = k0;
    int ne = *n_ent;
    while (p < seg1 && k < limit) {
      ent_pos[ne] = p;
      ent_k[ne] = k;
      ++ne;
      uint32_t f = s_hop[hop_ix(p - seg0)];
      k += f & 0xFFFFu;
      p = f >> 16;
    }
    *n_ent = ne;
    w->pos = p;
    w->k = k > limit ? limit : k;
    if (dbg && seg0 == 0) dbg[(int64_t)blockIdx.x * 8 + 7] = clock64();
  }
  bar_sync(bar_id, bar_n);
}


}  // namespace dec
}  // namespace nds
