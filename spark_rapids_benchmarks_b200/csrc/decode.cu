// Parquet page decode: RLE/bit-packed hybrid definition levels -> Arrow validity bitmap,
// RLE_DICTIONARY indices / PLAIN values -> dense device columns.  One CTA per data page.
//
// Reference boundary: the scan the reference triggers with
//   spark_session.read.format(input_format).load(table_path)   (/root/reference/nds/nds_power.py:97-100)
// and executes inside Spark's vectorised Parquet reader (not vendored).  Format facts follow the
// public Parquet "Encodings" spec: hybrid run header = varint, LSB 1 -> bit-packed run of
// (header>>1) groups of 8 values, LSB 0 -> RLE run of (header>>1) copies of a ceil(bw/8)-byte value;
// data page v1 def levels carry a 4-byte little-endian length prefix; dictionary-encoded values
// start with a 1-byte bit width.
//
// Roofline: HBM.  Algorithmic bytes per page = payload bytes read + width*num_values written
// + num_values/8 validity written (DESIGN.md "decode").
#include "device_types.h"
#include "common.h"

namespace nds {

namespace {

// Read `bw` (<= 32) bits starting `bit` bits after byte pointer p. Reads two aligned words;
// the staging layout guarantees kPagePad readable bytes after each payload.
__device__ __forceinline__ uint32_t extract_bits(const uint8_t* p, uint64_t bit, uint32_t bw) {
  const uint8_t* a = p + (bit >> 3);
  uintptr_t ai = reinterpret_cast<uintptr_t>(a);
  const uint32_t* w = reinterpret_cast<const uint32_t*>(ai & ~uintptr_t(3));
  uint64_t v = (uint64_t)__ldg(w) | ((uint64_t)__ldg(w + 1) << 32);
  uint32_t sh = (uint32_t)(ai & 3) * 8 + (uint32_t)(bit & 7);
  v >>= sh;
  // sh <= 31, bw <= 32: when sh + bw > 64 is impossible (31 + 32 = 63)
  return bw == 32 ? (uint32_t)v : ((uint32_t)v & ((1u << bw) - 1u));
}

__device__ __forceinline__ uint32_t load_u32_unaligned(const uint8_t* p) {
  return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}

struct Walker {       // sequential hybrid-run header walker state (thread 0 only)
  uint32_t pos;       // byte position inside the page payload
  uint32_t end;       // end of the stream
  uint32_t k;         // values covered so far
};

// Walk up to kRunCap run headers. Writes r_start[j] (first value index of run j) and
// r_info[j] ((payload byte pos << 1) | 1 for bit-packed, (value << 1) | 0 for RLE).
// Returns number of runs recorded; r_start[n] holds the end index (sentinel).
__device__ int walk_runs(const uint8_t* src, Walker& w, uint32_t bw, uint32_t limit,
                         uint32_t* r_start, uint32_t* r_info) {
  int n = 0;
  const uint32_t bytew = (bw + 7) >> 3;
  while (n < kRunCap && w.k < limit && w.pos < w.end) {
    uint32_t h = 0, s = 0, c;
    do {  // varint
      c = src[w.pos++];
      h |= (c & 0x7F) << s;
      s += 7;
    } while ((c & 0x80) && s < 35 && w.pos < w.end);
    r_start[n] = w.k;
    if (h & 1) {
      uint32_t groups = h >> 1;
      r_info[n] = (w.pos << 1) | 1u;
      w.pos += groups * bw;
      w.k += groups * 8;
    } else {
      uint32_t v = 0;
      for (uint32_t b = 0; b < bytew; ++b) v |= (uint32_t)src[w.pos + b] << (8 * b);
      r_info[n] = v << 1;
      w.pos += bytew;
      w.k += h >> 1;
    }
    ++n;
  }
  if (w.k > limit) w.k = limit;  // padding values of the last bit-packed group
  r_start[n] = w.k;
  return n;
}

template <typename T>
__device__ __forceinline__ void store_val(void* out, int64_t row, T v) {
  reinterpret_cast<T*>(out)[row] = v;
}

}  // namespace

__global__ void __launch_bounds__(kDecodeThreads)
decode_pages_kernel(const PageTask* __restrict__ tasks, int* __restrict__ err) {
  __shared__ uint32_t s_valid[kMaxPageValues / 32];
  __shared__ uint32_t s_rank[kMaxPageValues / 32];
  __shared__ uint32_t r_start[kRunCap + 1];
  __shared__ uint32_t r_info[kRunCap];
  __shared__ int s_nruns;
  __shared__ int s_done;
  __shared__ uint32_t s_warp_tot[kDecodeThreads / 32];
  __shared__ Walker s_walk;

  const PageTask t = tasks[blockIdx.x];
  const int tid = threadIdx.x;
  const uint32_t n = t.num_values;
  const uint32_t nw = (n + 31) >> 5;
  const bool nullable = t.valid != nullptr && t.def_len > 0;
  uint32_t values_off = 0;
  uint32_t nn = n;  // non-null count

  if (nullable) {
    // ---- 1. definition levels (bit width 1) -> s_valid ------------------------------------
    for (uint32_t w = tid; w < nw; w += kDecodeThreads) s_valid[w] = 0;
    if (tid == 0) {
      s_walk.pos = 4;
      s_walk.end = 4 + t.def_len;
      s_walk.k = 0;
      s_done = 0;
    }
    __syncthreads();
    for (;;) {
      if (tid == 0) {
        Walker w = s_walk;
        s_nruns = walk_runs(t.src, w, 1, n, r_start, r_info);
        s_walk = w;
        s_done = (w.k >= n || w.pos >= w.end) ? 1 : 0;
      }
      __syncthreads();
      const int nr = s_nruns;
      for (int j = tid; j < nr; j += kDecodeThreads) {
        uint32_t b0 = r_start[j], b1 = r_start[j + 1];
        if (b1 > n) b1 = n;
        uint32_t info = r_info[j];
        if (info & 1) {                       // bit-packed: payload bits are validity bits
          const uint8_t* p = t.src + (info >> 1);
          for (uint32_t b = b0; b < b1; b += 8) {
            uint32_t byte = p[(b - b0) >> 3];
            uint32_t rem = b1 - b;
            if (rem < 8) byte &= (1u << rem) - 1u;
            if (byte) atomicOr(&s_valid[b >> 5], byte << (b & 31));
            if ((b & 31) > 24 && (byte >> (32 - (b & 31)))) atomicOr(&s_valid[(b >> 5) + 1], byte >> (32 - (b & 31)));
          }
        } else if (info >> 1) {               // RLE run of 1s
          uint32_t b = b0;
          while (b < b1) {
            uint32_t w = b >> 5, lo = b & 31;
            uint32_t cnt = min(32u - lo, b1 - b);
            uint32_t m = (cnt == 32 ? 0xFFFFFFFFu : ((1u << cnt) - 1u)) << lo;
            atomicOr(&s_valid[w], m);
            b += cnt;
          }
        }
      }
      __syncthreads();
      if (s_done) break;
    }
    // ---- 2. exclusive prefix of popcounts per word -> s_rank ----------------------------------
    {
      const uint32_t per = (nw + kDecodeThreads - 1) / kDecodeThreads;
      uint32_t w0 = tid * per, w1 = min(w0 + per, nw);
      uint32_t sum = 0;
      for (uint32_t w = w0; w < w1; ++w) sum += __popc(s_valid[w]);
      uint32_t incl = sum;
      for (int d = 1; d < 32; d <<= 1) {
        uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if ((tid & 31) >= d) incl += v;
      }
      if ((tid & 31) == 31) s_warp_tot[tid >> 5] = incl;
      __syncthreads();
      uint32_t base = 0;
      for (int wv = 0; wv < (tid >> 5); ++wv) base += s_warp_tot[wv];
      uint32_t run = base + incl - sum;
      for (uint32_t w = w0; w < w1; ++w) {
        s_rank[w] = run;
        run += __popc(s_valid[w]);
      }
      nn = 0;
      for (int wv = 0; wv < kDecodeThreads / 32; ++wv) nn += s_warp_tot[wv];
    }
    __syncthreads();
    // ---- 3. validity bitmap out (page may start at any bit of the column bitmap) ---------------
    {
      const uint32_t sh = (uint32_t)(t.out_row & 31);
      uint32_t* vout = t.valid + (t.out_row >> 5);
      // output word j (0..nwo) = bits from s_valid[j-1] (high part) and s_valid[j] (low part)
      const uint32_t nwo = (n + sh + 31) >> 5;
      for (uint32_t j = tid; j < nwo; j += kDecodeThreads) {
        uint32_t lo = j < nw ? s_valid[j] : 0u;
        uint32_t hi = (j > 0 && j - 1 < nw) ? s_valid[j - 1] : 0u;
        uint32_t v = sh ? ((lo << sh) | (hi >> (32 - sh))) : lo;
        if (j == 0 || j == nwo - 1) { if (v) atomicOr(&vout[j], v); }
        else vout[j] = v;
      }
    }
    values_off = 4 + t.def_len;
  } else if (t.valid != nullptr) {
    // page without def levels (REQUIRED in this file) of a column that is nullable elsewhere
    const uint32_t sh = (uint32_t)(t.out_row & 31);
    uint32_t* vout = t.valid + (t.out_row >> 5);
    const uint32_t nwo = (n + sh + 31) >> 5;
    for (uint32_t j = tid; j < nwo; j += kDecodeThreads) {
      uint32_t v = 0xFFFFFFFFu;
      if (j == 0) v &= 0xFFFFFFFFu << sh;
      uint32_t endbit = n + sh - j * 32;
      if (endbit < 32) v &= (1u << endbit) - 1u;
      if (j == 0 || j == nwo - 1) atomicOr(&vout[j], v);
      else vout[j] = v;
    }
  }

  const uint8_t* vsrc = t.src + values_off;

  // ---- 4. values --------------------------------------------------------------------------
  if (t.enc == 0 && t.phys == 0) {                       // PLAIN fixed width
    for (uint32_t i = tid; i < n; i += kDecodeThreads) {
      bool ok = true;
      uint32_t k = i;
      if (nullable) {
        uint32_t wv = s_valid[i >> 5];
        ok = (wv >> (i & 31)) & 1u;
        k = s_rank[i >> 5] + __popc(wv & ((1u << (i & 31)) - 1u));
      }
      if (t.width == 4) {
        int32_t v = ok ? __ldg(reinterpret_cast<const int32_t*>(vsrc) + k) : 0;
        store_val<int32_t>(t.out, t.out_row + i, v);
      } else {
        int64_t v = ok ? __ldg(reinterpret_cast<const int64_t*>(vsrc) + k) : 0;
        store_val<int64_t>(t.out, t.out_row + i, v);
      }
    }
    return;
  }

  if (t.enc == 0 && t.phys == 1) {                       // PLAIN BYTE_ARRAY -> string views
    // fast path: every value has the same length L (TPC-DS *_id columns are char(16))
    const uint32_t vbytes = t.size - values_off;
    uint32_t L = nn ? load_u32_unaligned(vsrc) : 0;
    bool uniform = nn == 0 || ((uint64_t)(L + 4) * nn == vbytes);
    if (uniform) {
      for (uint32_t k = tid; k < nn && uniform; k += kDecodeThreads)
        uniform = load_u32_unaligned(vsrc + (uint64_t)k * (L + 4)) == L;
    }
    uniform = __syncthreads_and(uniform);
    if (uniform) {
      for (uint32_t i = tid; i < n; i += kDecodeThreads) {
        bool ok = true;
        uint32_t k = i;
        if (nullable) {
          uint32_t wv = s_valid[i >> 5];
          ok = (wv >> (i & 31)) & 1u;
          k = s_rank[i >> 5] + __popc(wv & ((1u << (i & 31)) - 1u));
        }
        StrView v = ok ? ((uint64_t)L << 32) | (uint64_t)(t.src_off + values_off + k * (L + 4) + 4) : 0ull;
        store_val<StrView>(t.out, t.out_row + i, v);
      }
    } else if (tid == 0) {
      // slow path: one thread follows the length prefixes (variable-length free text)
      uint32_t pos = 0;
      for (uint32_t i = 0; i < n; ++i) {
        bool ok = !nullable || ((s_valid[i >> 5] >> (i & 31)) & 1u);
        StrView v = 0;
        if (ok) {
          if (pos + 4 > vbytes) { if (err) atomicExch(err, 4); break; }
          uint32_t len = load_u32_unaligned(vsrc + pos);
          v = ((uint64_t)len << 32) | (uint64_t)(t.src_off + values_off + pos + 4);
          pos += 4 + len;
        }
        store_val<StrView>(t.out, t.out_row + i, v);
      }
    }
    return;
  }

  // RLE_DICTIONARY: bit width byte, then hybrid runs of dictionary indices over the nn values
  const uint32_t bw = vsrc[0];
  if (bw > 32) { if (tid == 0 && err) atomicExch(err, 5); return; }
  if (tid == 0) {
    s_walk.pos = values_off + 1;
    s_walk.end = t.size;
    s_walk.k = 0;
    s_done = 0;
  }
  __syncthreads();
  for (;;) {
    if (tid == 0) {
      Walker w = s_walk;
      s_nruns = bw ? walk_runs(t.src, w, bw, nn, r_start, r_info) : 0;
      if (!bw) { r_start[0] = 0; w.k = nn; }
      s_walk = w;
      s_done = (w.k >= nn || w.pos >= w.end) ? 1 : 0;
    }
    __syncthreads();
    const int nr = s_nruns;
    const uint32_t k_lo = r_start[0], k_hi = bw ? r_start[nr] : nn;
    for (uint32_t i = tid; i < n; i += kDecodeThreads) {
      bool ok = true;
      uint32_t k = i;
      if (nullable) {
        uint32_t wv = s_valid[i >> 5];
        ok = (wv >> (i & 31)) & 1u;
        k = s_rank[i >> 5] + __popc(wv & ((1u << (i & 31)) - 1u));
      }
      if (!ok) {
        if (k_lo == 0) {  // nulls are written once, in the first batch
          if (t.width == 4) store_val<int32_t>(t.out, t.out_row + i, 0);
          else store_val<int64_t>(t.out, t.out_row + i, 0);
        }
        continue;
      }
      if (k < k_lo || k >= k_hi) continue;
      uint32_t idx = 0;
      if (bw) {
        int lo = 0, hi = nr - 1;          // last run with r_start <= k
        while (lo < hi) {
          int mid = (lo + hi + 1) >> 1;
          if (r_start[mid] <= k) lo = mid; else hi = mid - 1;
        }
        uint32_t info = r_info[lo];
        idx = (info & 1) ? extract_bits(t.src + (info >> 1), (uint64_t)(k - r_start[lo]) * bw, bw)
                         : (info >> 1);
      }
      if (idx >= t.dict_count) { if (err) atomicExch(err, 6); idx = 0; if (!t.dict_count) continue; }
      if (t.width == 4) store_val<int32_t>(t.out, t.out_row + i, __ldg(reinterpret_cast<const int32_t*>(t.dict) + idx));
      else store_val<int64_t>(t.out, t.out_row + i, __ldg(reinterpret_cast<const int64_t*>(t.dict) + idx));
    }
    __syncthreads();
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

void launch_decode_pages(const PageTask* d_tasks, int npages, int* d_err, cudaStream_t stream) {
  if (npages <= 0) return;
  decode_pages_kernel<<<npages, kDecodeThreads, 0, stream>>>(d_tasks, d_err);
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
