// Operator kernels: fused filter/project, hash-join build/probe, hash group-by, key encoding and
// LSD radix sort.  All are HBM-bound integer/byte kernels (no tensor-core work on this path).
// Row counts may live on the device (RowCount::dev) so that a chain of operators needs no host sync.
//
// Spark operators these replace (none of them is in /root/reference; the reference hands SQL text to
// Spark at /root/reference/nds/nds_power.py:130): FilterExec/ProjectExec, BroadcastHashJoinExec,
// HashAggregateExec, TakeOrderedAndProjectExec.
#include "ops.cuh"
#include "common.h"

namespace nds {

namespace {

constexpr int kThreads = 256;

inline int grid_for(int64_t n, int per_block = kThreads, int max_blocks = 148 * 8) {
  int64_t g = (n + per_block - 1) / per_block;
  if (g < 1) g = 1;
  if (g > max_blocks) g = max_blocks;
  return (int)g;
}

__device__ __forceinline__ int64_t row_count(const int* dev, int64_t host) {
  if (dev) { int64_t d = *dev; return d < host ? d : host; }
  return host;
}

__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// warp-aggregated append: returns the output position for lanes with keep == true
__device__ __forceinline__ int warp_append(bool keep, int* counter) {
  uint32_t m = __ballot_sync(0xFFFFFFFFu, keep);
  int base = 0;
  int lane = threadIdx.x & 31;
  if (m) {
    if (lane == (__ffs(m) - 1)) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, __ffs(m) - 1);
  }
  return base + __popc(m & lanemask_lt());
}

}  // namespace

// ------------------------------------------------------------------ filter
__global__ void __launch_bounds__(kThreads)
filter_kernel(Prog p, OpCtx c, const int* n_dev, int64_t n_host, uint32_t* __restrict__ out_sel, int* out_count) {
  const int64_t n = row_count(n_dev, n_host);
  const int64_t nround = (n + 31) & ~int64_t(31);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
    bool keep = false;
    if (i < n) {
      Val v = eval_program(p.code, p.n, c.cols, c.lit, i);
      keep = v.ok && v.v != 0;
    }
    int pos = warp_append(keep, out_count);
    if (keep) out_sel[pos] = (uint32_t)i;
  }
}

void launch_filter(Prog p, OpCtx c, RowCount n, uint32_t* out_sel, int* out_count, cudaStream_t s) {
  filter_kernel<<<grid_for(n.host), kThreads, 0, s>>>(p, c, n.dev, n.host, out_sel, out_count);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ project
// Evaluates one expression for every row (or for rows[i] when given) into a flat column.
__global__ void __launch_bounds__(kThreads)
project_kernel(Prog p, OpCtx c, const int* n_dev, int64_t n_host, const uint32_t* __restrict__ rows, void* out,
               uint32_t* __restrict__ out_valid, int out_dtype, const uint8_t* str_out_base, int* err) {
  const int64_t n = row_count(n_dev, n_host);
  const int64_t nround = (n + 31) & ~int64_t(31);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
    Val v;
    v.ok = false; v.v = 0; v.len = 0;
    if (i < n) v = eval_program(p.code, p.n, c.cols, c.lit, rows ? (int64_t)rows[i] : i);
    uint32_t m = __ballot_sync(0xFFFFFFFFu, v.ok);
    if (out_valid && (threadIdx.x & 31) == 0) out_valid[i >> 5] = m;
    if (i < n) {
      if (out_dtype == DT_I32) reinterpret_cast<int32_t*>(out)[i] = v.ok ? (int32_t)v.v : 0;
      else if (out_dtype == DT_STR) {
        StrView sv = 0;
        if (v.ok) {
          int64_t off = (int64_t)((const uint8_t*)(uintptr_t)v.v - str_out_base);
          if (off < 0 || off > 0xFFFFFFFFll) { if (err) atomicExch(err, 11); off = 0; }
          sv = ((uint64_t)v.len << 32) | (uint64_t)(uint32_t)off;
        }
        reinterpret_cast<StrView*>(out)[i] = sv;
      } else reinterpret_cast<int64_t*>(out)[i] = v.ok ? v.v : 0;
    }
  }
}

void launch_project(Prog p, OpCtx c, RowCount n, const uint32_t* rows, void* out, uint32_t* out_valid, int out_dtype,
                    const uint8_t* str_out_base, int* err, cudaStream_t s) {
  project_kernel<<<grid_for(n.host), kThreads, 0, s>>>(p, c, n.dev, n.host, rows, out, out_valid, out_dtype, str_out_base, err);
  NDS_CUDA(cudaGetLastError());
}

// Several expressions in one launch (sort output, group keys, final result columns): every thread evaluates all of
// them for its row; one launch instead of one per column.
__global__ void __launch_bounds__(kThreads)
project_multi_kernel(const ProjSpec* __restrict__ specs, int nspecs, OpCtx c, const int* n_dev, int64_t n_host,
                     const uint32_t* __restrict__ rows, int* err) {
  const int64_t n = row_count(n_dev, n_host);
  const int64_t nround = (n + 31) & ~int64_t(31);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = (i < n && rows) ? (int64_t)rows[i] : i;
    for (int sidx = 0; sidx < nspecs; ++sidx) {
      const ProjSpec sp = specs[sidx];
      Val v;
      v.ok = false; v.v = 0; v.len = 0;
      if (i < n) v = eval_program(sp.prog.code, sp.prog.n, c.cols, c.lit, r);
      uint32_t m = __ballot_sync(0xFFFFFFFFu, v.ok);
      if (sp.valid && (threadIdx.x & 31) == 0) sp.valid[i >> 5] = m;
      if (i < n) {
        if (sp.dtype == DT_I32) reinterpret_cast<int32_t*>(sp.out)[i] = v.ok ? (int32_t)v.v : 0;
        else if (sp.dtype == DT_STR) {
          StrView sv = 0;
          if (v.ok) {
            int64_t off = (int64_t)((const uint8_t*)(uintptr_t)v.v - sp.strbase);
            if (off < 0 || off > 0xFFFFFFFFll) { if (err) atomicExch(err, 11); off = 0; }
            sv = ((uint64_t)v.len << 32) | (uint64_t)(uint32_t)off;
          }
          reinterpret_cast<StrView*>(sp.out)[i] = sv;
        } else reinterpret_cast<int64_t*>(sp.out)[i] = v.ok ? v.v : 0;
      }
    }
  }
}
void launch_project_multi(const ProjSpec* specs, int nspecs, OpCtx c, RowCount n, const uint32_t* rows, int* err, cudaStream_t s) {
  if (nspecs <= 0) return;
  project_multi_kernel<<<grid_for(n.host), kThreads, 0, s>>>(specs, nspecs, c, n.dev, n.host, rows, err);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ small utilities
__global__ void compose_kernel(const uint32_t* __restrict__ inner, const uint32_t* __restrict__ idx, const int* n_dev,
                               int64_t n_host, uint32_t* __restrict__ out) {
  const int64_t n = row_count(n_dev, n_host);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
  {
    const uint32_t j = idx[i];
    out[i] = j == kNoRow ? kNoRow : inner[j];      // an outer join's unmatched row stays unmatched through every composition
  }
}
void launch_compose(const uint32_t* inner, const uint32_t* idx, RowCount n, uint32_t* out, cudaStream_t s) {
  compose_kernel<<<grid_for(n.host), kThreads, 0, s>>>(inner, idx, n.dev, n.host, out);
  NDS_CUDA(cudaGetLastError());
}

__global__ void fill_i64_kernel(int64_t* p, int64_t v, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
void launch_fill_i64(int64_t* p, int64_t v, int64_t n, cudaStream_t s) {
  if (n <= 0) return;
  fill_i64_kernel<<<grid_for(n), kThreads, 0, s>>>(p, v, n);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ hash join
// Open addressing, linear probing, 64-bit keys + 32-bit payload (the build-side batch row).
// Build keys must be unique (FK -> PK joins); a duplicate raises flag bit 0.
__global__ void __launch_bounds__(kThreads)
join_build_kernel(ColRef key, const int* n_dev, int64_t n_host, HashTable ht, int* flags, int allow_dup) {
  const int64_t n = row_count(n_dev, n_host);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    Val k = load_col(key, i);
    if (!k.ok) continue;                       // NULL keys never match an equi-join
    if (ht.dense) {
      const uint64_t d = (uint64_t)(k.v - ht.lo);
      if (d >= (uint64_t)ht.range) { atomicOr(flags, 4); continue; }      // outside the range the column statistics promised
      if (ht.next) ht.next[i] = atomicExch(&ht.dense[d], (uint32_t)i + 1u);   // push onto the key's chain
      else if (atomicCAS(&ht.dense[d], 0u, (uint32_t)i + 1u) != 0u && !allow_dup) atomicOr(flags, 1);
      continue;
    }
    if (k.v == kEmptyKey) { atomicOr(flags, 2); continue; }
    uint32_t s = (uint32_t)mix64((uint64_t)k.v) & ht.mask;
    for (;;) {
      unsigned long long prev = atomicCAS(reinterpret_cast<unsigned long long*>(&ht.keys[s]),
                                          (unsigned long long)kEmptyKey, (unsigned long long)k.v);
      if (ht.next) {
        if ((int64_t)prev == kEmptyKey || (int64_t)prev == k.v) { ht.next[i] = atomicExch(&ht.vals[s], (uint32_t)i + 1u); break; }
      } else {
        if ((int64_t)prev == kEmptyKey) { ht.vals[s] = (uint32_t)i; break; }
        if ((int64_t)prev == k.v) { if (!allow_dup) atomicOr(flags, 1); break; }
      }
      s = (s + 1) & ht.mask;
    }
  }
}

// Dynamic partition pruning: the build keys that fall into [lo, hi] (the range of the probe table's partition values)
// become bits of a small bitmap, so that a few hundred bytes travel to the host instead of every build key.
__global__ void __launch_bounds__(kThreads)
keys_to_bitmap_kernel(ColRef key, const int* n_dev, int64_t n_host, int64_t lo, int64_t hi, uint32_t* bits) {
  const int64_t n = row_count(n_dev, n_host);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    Val k = load_col(key, i);
    if (!k.ok || k.v < lo || k.v > hi) continue;
    const uint64_t d = (uint64_t)(k.v - lo);
    atomicOr(&bits[d >> 5], 1u << (d & 31));
  }
}
void launch_keys_to_bitmap(ColRef key, RowCount n, int64_t lo, int64_t hi, uint32_t* bits, cudaStream_t s) {
  keys_to_bitmap_kernel<<<grid_for(n.host), kThreads, 0, s>>>(key, n.dev, n.host, lo, hi, bits);
  NDS_CUDA(cudaGetLastError());
}

__global__ void __launch_bounds__(kThreads)
join_probe_kernel(ColRef key, const int* n_dev, int64_t n_host, HashTable ht, uint32_t* __restrict__ out_probe,
                  uint32_t* __restrict__ out_build, int* out_count, int kind) {
  const int64_t n = row_count(n_dev, n_host);
  const int64_t nround = (n + 31) & ~int64_t(31);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
    bool hit = false;
    uint32_t b = 0;
    if (i < n) {
      Val k = load_col(key, i);
      if (k.ok) hit = ht_find(ht, k.v, b);
    }
    // inner / semi keep the matching rows, anti the others (a NULL key matches nothing), left keeps every row
    const bool keep = i < n && (kind == JOIN_LEFT || (kind == JOIN_ANTI ? !hit : hit));
    int pos = warp_append(keep, out_count);
    if (keep) {
      out_probe[pos] = (uint32_t)i;
      if (kind == JOIN_INNER) out_build[pos] = b;
      else if (kind == JOIN_LEFT) out_build[pos] = hit ? b : kNoRow;
    }
  }
}

// Fused star-join probe: every base row walks the chain of hash tables (most selective first) and leaves at its first
// miss; survivors are appended once with the matched build row of every join. One kernel instead of one probe plus
// the row-map compositions per join, and no intermediate selection vectors.
__global__ void __launch_bounds__(kThreads)
join_probe_chain_kernel(const ProbeChain c, const int* n_dev, int64_t n_host, uint32_t* __restrict__ out_probe, int* out_count) {
  const int64_t n = row_count(n_dev, n_host);
  const int64_t nround = (n + 31) & ~int64_t(31);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
    bool hit = i < n;
    uint32_t b[kMaxChain];
#pragma unroll
    for (int j = 0; j < kMaxChain; ++j) {
      b[j] = 0;
      if (j < c.n && hit) {
        Val k = load_col(c.key[j], i);
        hit = k.ok && ht_find(c.ht[j], k.v, b[j]);
      }
    }
    int pos = warp_append(hit, out_count);
    if (hit) {
      out_probe[pos] = (uint32_t)i;
#pragma unroll
      for (int j = 0; j < kMaxChain; ++j)
        if (j < c.n) c.out_build[j][pos] = b[j];
    }
  }
}
void launch_join_probe_chain(const ProbeChain& chain, RowCount n, uint32_t* out_probe, int* out_count, cudaStream_t s) {
  join_probe_chain_kernel<<<grid_for(n.host), kThreads, 0, s>>>(chain, n.dev, n.host, out_probe, out_count);
  NDS_CUDA(cudaGetLastError());
}

// Filter + build in one pass over an unfiltered build side (a dimension scan under a filter): the predicate is evaluated per
// row and the rows that pass go straight into a direct-address table with their row in the UNFILTERED batch as the value,
// so no selection vector is written and the probe's matched rows address the decoded columns directly.  prune_bits != null:
// the keys inside [prune_lo, prune_hi] of the rows that pass also set a bit (dynamic partition pruning, keys_to_bitmap_kernel).
__global__ void __launch_bounds__(kThreads)
join_build_filter_kernel(Prog pred, OpCtx c, ColRef key, const int* n_dev, int64_t n_host, HashTable ht, int* flags, int allow_dup,
                         int64_t prune_lo, int64_t prune_hi, uint32_t* prune_bits) {
  const int64_t n = row_count(n_dev, n_host);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const Val pv = eval_program(pred.code, pred.n, c.cols, c.lit, i);
    if (!pv.ok || pv.v == 0) continue;
    const Val k = load_col(key, i);
    if (!k.ok) continue;                       // NULL keys never match an equi-join
    const uint64_t d = (uint64_t)(k.v - ht.lo);
    if (d >= (uint64_t)ht.range) { atomicOr(flags, 4); continue; }
    if (atomicCAS(&ht.dense[d], 0u, (uint32_t)i + 1u) != 0u && !allow_dup) atomicOr(flags, 1);
    if (prune_bits && k.v >= prune_lo && k.v <= prune_hi) {
      const uint64_t b = (uint64_t)(k.v - prune_lo);
      atomicOr(&prune_bits[b >> 5], 1u << (b & 31));
    }
  }
}
void launch_join_build_filter(Prog pred, OpCtx c, ColRef key, RowCount n, HashTable ht, int* flags, bool allow_dup, int64_t prune_lo,
                              int64_t prune_hi, uint32_t* prune_bits, cudaStream_t s) {
  join_build_filter_kernel<<<grid_for(n.host), kThreads, 0, s>>>(pred, c, key, n.dev, n.host, ht, flags, allow_dup ? 1 : 0, prune_lo, prune_hi,
                                                                 prune_bits);
  NDS_CUDA(cudaGetLastError());
}

void launch_join_build(ColRef key, RowCount n, HashTable ht, int* flags, cudaStream_t s, bool allow_dup) {
  join_build_kernel<<<grid_for(n.host), kThreads, 0, s>>>(key, n.dev, n.host, ht, flags, allow_dup ? 1 : 0);
  NDS_CUDA(cudaGetLastError());
}
void launch_join_probe(ColRef key, RowCount n, HashTable ht, uint32_t* out_probe, uint32_t* out_build, int* out_count,
                       cudaStream_t s, int kind) {
  join_probe_kernel<<<grid_for(n.host), kThreads, 0, s>>>(key, n.dev, n.host, ht, out_probe, out_build, out_count, kind);
  NDS_CUDA(cudaGetLastError());
}

// Repeated build keys: a probe row has as many output rows as its key's chain is long (left: at least one).
__global__ void __launch_bounds__(kThreads)
join_probe_multi_count_kernel(ColRef key, const int* n_dev, int64_t n_host, HashTable ht, int kind, unsigned long long* out_total) {
  const int64_t n = row_count(n_dev, n_host);
  unsigned long long mine = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    Val k = load_col(key, i);
    uint32_t c = 0;
    if (k.ok)
      for (uint32_t l = ht_first(ht, k.v); l; l = ht_next(ht, l)) ++c;
    mine += (c == 0 && kind == JOIN_LEFT) ? 1u : c;
  }
  for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(out_total, mine);
}
__global__ void __launch_bounds__(kThreads)
join_probe_multi_write_kernel(ColRef key, const int* n_dev, int64_t n_host, HashTable ht, int kind, uint32_t* __restrict__ out_probe,
                              uint32_t* __restrict__ out_build, unsigned long long* cursor) {
  const int64_t n = row_count(n_dev, n_host);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    Val k = load_col(key, i);
    uint32_t c = 0;
    if (k.ok)
      for (uint32_t l = ht_first(ht, k.v); l; l = ht_next(ht, l)) ++c;
    const uint32_t rows = (c == 0 && kind == JOIN_LEFT) ? 1u : c;
    if (!rows) continue;
    unsigned long long pos = atomicAdd(cursor, (unsigned long long)rows);
    if (c == 0) { out_probe[pos] = (uint32_t)i; out_build[pos] = kNoRow; continue; }
    for (uint32_t l = ht_first(ht, k.v); l; l = ht_next(ht, l), ++pos) { out_probe[pos] = (uint32_t)i; out_build[pos] = l - 1u; }
  }
}
void launch_join_probe_multi_count(ColRef key, RowCount n, HashTable ht, int kind, unsigned long long* out_total, cudaStream_t s) {
  join_probe_multi_count_kernel<<<grid_for(n.host), kThreads, 0, s>>>(key, n.dev, n.host, ht, kind, out_total);
  NDS_CUDA(cudaGetLastError());
}
void launch_join_probe_multi_write(ColRef key, RowCount n, HashTable ht, int kind, uint32_t* out_probe, uint32_t* out_build,
                                   unsigned long long* cursor, cudaStream_t s) {
  join_probe_multi_write_kernel<<<grid_for(n.host), kThreads, 0, s>>>(key, n.dev, n.host, ht, kind, out_probe, out_build, cursor);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ concatenation of row sets (rollup levels)
__global__ void prefix_counts_kernel(const int* __restrict__ cnt, const int* __restrict__ bound, int n, int* __restrict__ off) {
  if (blockIdx.x || threadIdx.x) return;
  int run = 0;
  off[0] = 0;
  for (int i = 0; i < n; ++i) {
    int c = cnt[i];
    if (c > bound[i]) c = bound[i];
    if (c < 0) c = 0;
    run += c;
    off[i + 1] = run;
  }
}
void launch_prefix_counts(const int* d_cnt, const int* d_bound, int n, int* d_off, cudaStream_t s) {
  prefix_counts_kernel<<<1, 32, 0, s>>>(d_cnt, d_bound, n, d_off);
  NDS_CUDA(cudaGetLastError());
}

__global__ void __launch_bounds__(kThreads)
union_copy_kernel(const UnionSeg* __restrict__ segs) {
  const UnionSeg g = segs[blockIdx.y];
  const int64_t n = row_count(g.n_dev, g.n_host);
  const int64_t off = *g.off;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = off + i;
    const bool ok = g.present && ((g.src_valid[i >> 5] >> (i & 31)) & 1u);
    if (g.width == 4) reinterpret_cast<int32_t*>(g.dst)[o] = ok ? reinterpret_cast<const int32_t*>(g.src)[i] : 0;
    else reinterpret_cast<int64_t*>(g.dst)[o] = ok ? reinterpret_cast<const int64_t*>(g.src)[i] : 0;
    if (ok) atomicOr(&g.dst_valid[o >> 5], 1u << (o & 31));
  }
}
void launch_union_copy(const UnionSeg* d_segs, int nsegs, int64_t max_rows, cudaStream_t s) {
  if (nsegs <= 0) return;
  int64_t gx = (max_rows + kThreads - 1) / kThreads;
  if (gx < 1) gx = 1;
  if (gx > 148 * 4) gx = 148 * 4;
  union_copy_kernel<<<dim3((unsigned)gx, (unsigned)nsegs), kThreads, 0, s>>>(d_segs);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ hash group-by
// The table is ALLOCATED for the upper bound of the input rows (cap_max slots) but SIZED on the device from the actual
// row count: every kernel derives the same power-of-two capacity from n, so only that part is cleared, probed and
// compacted, and the host never waits for the count of the join below.
// (a power of two of at least 4/3 of the rows: the load factor stays below 0.75 even if every row is a group of its own,
// and the table - cleared and compacted in full - is half the size "twice the rows" would make it for most row counts)
__host__ __device__ __forceinline__ uint32_t agg_capacity(int64_t n, uint32_t cap_max) {
  uint32_t c = 64;
  while ((int64_t)c < n + n / 3 + 1 && c < cap_max) c <<= 1;
  return c < cap_max ? c : cap_max;
}
uint32_t agg_capacity_for(int64_t n_bound) {
  uint32_t c = 64;
  while ((int64_t)c < n_bound + n_bound / 3 + 1 && c < (1u << 31)) c <<= 1;
  return c;
}

__global__ void __launch_bounds__(kThreads)
agg_init_kernel(const AggSpec* __restrict__ aggs, int naggs, const int* n_dev, int64_t n_host, uint32_t cap_max, uint32_t* slots,
                int64_t* acc, int64_t* cnt, uint32_t* out_valid, int64_t valid_stride_words) {
  const int64_t n = row_count(n_dev, n_host);
  const uint64_t cap = agg_capacity(n, cap_max);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (int64_t i = t0; i < (int64_t)cap; i += stride) slots[i] = 0;
  for (int a = 0; a < naggs; ++a) {
    const int fn = aggs[a].fn;
    const int64_t init = fn == AGG_MIN_I ? INT64_MAX : (fn == AGG_MAX_I ? INT64_MIN : 0);
    // a count that shares another aggregate's counter owns no state of its own; plain counts never touch acc
    if (fn != AGG_COUNT_REF) {
      const bool counting = fn == AGG_COUNT || fn == AGG_COUNT_STAR;
      for (int64_t i = t0; i < (int64_t)cap; i += stride) { if (!counting) acc[a * cap + i] = init; cnt[a * cap + i] = 0; }
    }
    const int64_t vw = min((n + 31) / 32 + 1, valid_stride_words);
    for (int64_t i = t0; i < vw; i += stride) out_valid[a * valid_stride_words + i] = 0;
  }
}
// slots[s] = representative batch row + 1 (0 = empty). Key equality is decided by re-evaluating the key
// expressions at the representative row, so no key columns are stored in the table.
__device__ __forceinline__ bool val_equal(const Val& a, const Val& b, int kind) {
  if (a.ok != b.ok) return false;
  if (!a.ok) return true;
  if (kind == 2) {
    if (a.len != b.len) return false;
    if (a.v == b.v) return true;               // the same dictionary entry (or the same bytes): no compare needed
    const uint8_t* x = (const uint8_t*)(uintptr_t)a.v;
    const uint8_t* y = (const uint8_t*)(uintptr_t)b.v;
    for (uint32_t i = 0; i < a.len; ++i)
      if (x[i] != y[i]) return false;
    return true;
  }
  return a.v == b.v;
}

// Accumulation goes through a per-CTA cache in shared memory (pre-aggregation): a direct-mapped table of kAggCache
// entries tagged with the global slot. A row whose slot owns (or can claim) its entry accumulates with shared-memory
// atomics; the others go straight to the global table. At the end every CTA flushes its entries with one global atomic
// per aggregate. Few groups (the NDS top-k queries: tens to thousands) -> almost every row stays in shared memory and
// the same-address contention in L2 disappears; many groups -> the CTA notices that most rows miss and turns the
// cache off for its remaining rows.
constexpr int kAggCache = 256;
__device__ __forceinline__ void agg_apply(int fn, int64_t* acc, int64_t* cnt, int64_t v) {
  atomicAdd(reinterpret_cast<unsigned long long*>(cnt), 1ull);
  if (fn == AGG_SUM_I) atomicAdd(reinterpret_cast<unsigned long long*>(acc), (unsigned long long)v);
  else if (fn == AGG_MIN_I) atomicMin(reinterpret_cast<long long*>(acc), (long long)v);
  else if (fn == AGG_MAX_I) atomicMax(reinterpret_cast<long long*>(acc), (long long)v);
  else if (fn == AGG_SUM_F) atomicAdd(reinterpret_cast<double*>(acc), __longlong_as_double(v));
}

__global__ void __launch_bounds__(kThreads)
agg_insert_kernel(const Prog* __restrict__ progs, int nkeys, const AggSpec* __restrict__ aggs, int naggs, OpCtx c,
                  const int* n_dev, int64_t n_host, uint32_t* slots, uint32_t cap_max, int64_t* acc, int64_t* cnt, int use_cache) {
  extern __shared__ __align__(16) uint8_t agg_smem[];
  uint32_t* c_tag = reinterpret_cast<uint32_t*>(agg_smem);                       // global slot + 1, 0 = free
  int64_t* c_acc = reinterpret_cast<int64_t*>(agg_smem + kAggCache * 4);         // [naggs][kAggCache]
  int64_t* c_cnt = c_acc + (int64_t)naggs * kAggCache;
  __shared__ int s_seen, s_miss, s_off;
  const int64_t n = row_count(n_dev, n_host);
  const uint64_t cap = agg_capacity(n, cap_max);
  const uint32_t mask = (uint32_t)cap - 1u;
  if (use_cache) {
    for (int e = threadIdx.x; e < kAggCache; e += blockDim.x) c_tag[e] = 0;
    for (int k = threadIdx.x; k < naggs * kAggCache; k += blockDim.x) {
      const int fn = aggs[k / kAggCache].fn;
      c_acc[k] = fn == AGG_MIN_I ? INT64_MAX : (fn == AGG_MAX_I ? INT64_MIN : 0);
      c_cnt[k] = 0;
    }
    if (threadIdx.x == 0) { s_seen = 0; s_miss = 0; s_off = 0; }
    __syncthreads();
  }
  for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x; i0 < n; i0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = i0 + threadIdx.x;
    const uint32_t active = __ballot_sync(0xFFFFFFFFu, i < n);      // the lanes that have a row: they run the body together
    if (i < n) {
      Val mine[kMaxGroupKeys];
      uint64_t h = 0x9E3779B97F4A7C15ULL;
      for (int k = 0; k < nkeys; ++k) {
        mine[k] = eval_program(progs[k].code, progs[k].n, c.cols, c.lit, i);
        uint64_t hv = !mine[k].ok ? 0x7F4A7C15ULL
                      : (progs[k].kind == 2 ? hash_bytes((const uint8_t*)(uintptr_t)mine[k].v, mine[k].len) : (uint64_t)mine[k].v);
        h = mix64(h ^ hv) + 0x9E3779B97F4A7C15ULL * (k + 1);
      }
      uint32_t s = (uint32_t)mix64(h) & mask;
      for (;;) {
        uint32_t prev = atomicCAS(&slots[s], 0u, (uint32_t)i + 1u);
        if (prev == 0) break;
        const int64_t rep = (int64_t)prev - 1;
        bool same = true;
        for (int k = 0; k < nkeys && same; ++k) {
          Val o = eval_program(progs[k].code, progs[k].n, c.cols, c.lit, rep);
          same = val_equal(mine[k], o, progs[k].kind);
        }
        if (same) break;
        s = (s + 1) & mask;
      }
      // Every row of the warp in the same group (few groups, or none at all - a keyless aggregate): SUM / COUNT are reduced
      // across the warp (redux) and one lane applies them, instead of 32 atomics on one address.
      const int lane = threadIdx.x & 31, leader = __ffs(active) - 1;
      const bool uniform = __popc(active) > 1 && __all_sync(active, s == __shfl_sync(active, s, leader));
      bool hit = false;
      uint32_t e = 0;
      if (use_cache && !s_off && (!uniform || lane == leader)) {
        e = (s ^ (s >> 8)) & (kAggCache - 1);
        const uint32_t tag = atomicCAS(&c_tag[e], 0u, s + 1u);
        hit = tag == 0u || tag == s + 1u;
        if (!hit) atomicAdd(&s_miss, 1);
      }
      for (int a = 0; a < naggs; ++a) {
        const AggSpec sp = aggs[a];
        int64_t* pa = hit ? &c_acc[a * kAggCache + e] : &acc[a * cap + s];
        int64_t* pc = hit ? &c_cnt[a * kAggCache + e] : &cnt[a * cap + s];
        if (sp.fn == AGG_COUNT_REF) continue;
        Val v;
        v.ok = true; v.v = 0; v.len = 0;
        if (sp.fn != AGG_COUNT_STAR) v = eval_program(progs[sp.prog].code, progs[sp.prog].n, c.cols, c.lit, i);
        if (uniform && (sp.fn == AGG_SUM_I || sp.fn == AGG_COUNT || sp.fn == AGG_COUNT_STAR)) {
          const unsigned cntv = __reduce_add_sync(active, v.ok ? 1u : 0u);
          unsigned long long sum = 0;
          if (sp.fn == AGG_SUM_I) {
            // 64-bit sum from four 16-bit digit sums (each fits 21 bits for 32 lanes); wraps like the atomics do
            const unsigned long long x = v.ok ? (unsigned long long)v.v : 0ull;
#pragma unroll
            for (int q = 0; q < 4; ++q) sum += (unsigned long long)__reduce_add_sync(active, (unsigned)((x >> (16 * q)) & 0xFFFFull)) << (16 * q);
          }
          if (lane != leader || cntv == 0) continue;
          atomicAdd(reinterpret_cast<unsigned long long*>(pc), (unsigned long long)cntv);
          if (sp.fn == AGG_SUM_I) atomicAdd(reinterpret_cast<unsigned long long*>(pa), sum);
          continue;
        }
        if (sp.fn == AGG_COUNT_STAR) { atomicAdd(reinterpret_cast<unsigned long long*>(pc), 1ull); continue; }
        if (!v.ok) continue;
        agg_apply(sp.fn, pa, pc, v.v);
      }
    }
    if (use_cache && !s_off) {
      // after the first rounds: if most rows missed the cache this input has too many groups for it
      __syncthreads();
      if (threadIdx.x == 0) {
        s_seen += blockDim.x;
        if (s_seen >= 4 * kAggCache && s_miss * 2 > s_seen) s_off = 1;
      }
      __syncthreads();
    }
  }
  if (!use_cache) return;
  __syncthreads();
  for (int e = threadIdx.x; e < kAggCache; e += blockDim.x) {
    const uint32_t tag = c_tag[e];
    if (!tag) continue;
    const uint64_t s = tag - 1u;
    for (int a = 0; a < naggs; ++a) {
      const int fn = aggs[a].fn;
      const int64_t cv = c_cnt[a * kAggCache + e];
      if (cv == 0) continue;
      atomicAdd(reinterpret_cast<unsigned long long*>(&cnt[a * cap + s]), (unsigned long long)cv);
      const int64_t av = c_acc[a * kAggCache + e];
      if (fn == AGG_SUM_I) atomicAdd(reinterpret_cast<unsigned long long*>(&acc[a * cap + s]), (unsigned long long)av);
      else if (fn == AGG_MIN_I) atomicMin(reinterpret_cast<long long*>(&acc[a * cap + s]), (long long)av);
      else if (fn == AGG_MAX_I) atomicMax(reinterpret_cast<long long*>(&acc[a * cap + s]), (long long)av);
      else if (fn == AGG_SUM_F) atomicAdd(reinterpret_cast<double*>(&acc[a * cap + s]), __longlong_as_double(av));
    }
  }
}

// Compacts occupied slots into group rows. out_vals[a * out_stride + g], out_valid[a * valid_stride_words + g/32].
__global__ void __launch_bounds__(kThreads)
agg_finalize_kernel(const AggSpec* __restrict__ aggs, int naggs, const uint32_t* __restrict__ slots, uint32_t cap_max,
                    const int* n_dev, int64_t n_host, const int64_t* __restrict__ acc, const int64_t* __restrict__ cnt,
                    uint32_t* __restrict__ rep_out, int64_t* __restrict__ out_vals, uint32_t* out_valid, int64_t out_stride,
                    int64_t valid_stride_words, int* ngroups) {
  const uint32_t cap = agg_capacity(row_count(n_dev, n_host), cap_max);
  const uint32_t capround = (cap + 31u) & ~31u;
  for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < capround; s += gridDim.x * blockDim.x) {
    uint32_t rep = s < cap ? slots[s] : 0u;
    int g = warp_append(rep != 0, ngroups);
    if (!rep) continue;
    rep_out[g] = rep - 1;
    for (int a = 0; a < naggs; ++a) {
      const int fn = aggs[a].fn;
      const bool counting = fn == AGG_COUNT || fn == AGG_COUNT_STAR || fn == AGG_COUNT_REF;
      int64_t c = cnt[(uint64_t)(fn == AGG_COUNT_REF ? aggs[a].prog : a) * cap + s];
      int64_t v = counting ? c : acc[(uint64_t)a * cap + s];
      bool ok = counting ? true : c > 0;
      out_vals[a * out_stride + g] = ok ? v : 0;
      if (ok) atomicOr(&out_valid[a * valid_stride_words + (g >> 5)], 1u << (g & 31));
    }
  }
}

// Window aggregates over PARTITION BY keys (no ORDER BY, no frame): after agg_insert_kernel has accumulated the partitions'
// states, every ROW looks its partition's slot up again (same hash, same probe sequence, keys compared with the slot's
// representative row) and reads the finished value - the groups are never compacted.  out_vals[a * out_stride + row].
__global__ void __launch_bounds__(kThreads)
win_gather_kernel(const Prog* __restrict__ progs, int nkeys, const AggSpec* __restrict__ aggs, int naggs, OpCtx c, const int* n_dev,
                  int64_t n_host, const uint32_t* __restrict__ slots, uint32_t cap_max, const int64_t* __restrict__ acc,
                  const int64_t* __restrict__ cnt, int64_t* __restrict__ out_vals, uint32_t* __restrict__ out_valid, int64_t out_stride,
                  int64_t valid_stride_words) {
  const int64_t n = row_count(n_dev, n_host);
  const uint64_t cap = agg_capacity(n, cap_max);
  const uint32_t mask = (uint32_t)cap - 1u;
  const int64_t nround = (n + 31) & ~int64_t(31);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nround; i += (int64_t)gridDim.x * blockDim.x) {
    const bool live = i < n;
    bool found = false;
    uint32_t s = 0;
    if (live) {
      Val mine[kMaxGroupKeys];
      uint64_t h = 0x9E3779B97F4A7C15ULL;
      for (int k = 0; k < nkeys; ++k) {
        mine[k] = eval_program(progs[k].code, progs[k].n, c.cols, c.lit, i);
        uint64_t hv = !mine[k].ok ? 0x7F4A7C15ULL
                      : (progs[k].kind == 2 ? hash_bytes((const uint8_t*)(uintptr_t)mine[k].v, mine[k].len) : (uint64_t)mine[k].v);
        h = mix64(h ^ hv) + 0x9E3779B97F4A7C15ULL * (k + 1);
      }
      s = (uint32_t)mix64(h) & mask;
      for (;;) {
        const uint32_t prev = slots[s];
        if (prev == 0) break;
        const int64_t rep = (int64_t)prev - 1;
        bool same = true;
        for (int k = 0; k < nkeys && same; ++k) {
          Val o = eval_program(progs[k].code, progs[k].n, c.cols, c.lit, rep);
          same = val_equal(mine[k], o, progs[k].kind);
        }
        if (same) { found = true; break; }
        s = (s + 1) & mask;
      }
    }
    for (int a = 0; a < naggs; ++a) {
      const int fn = aggs[a].fn;
      const bool counting = fn == AGG_COUNT || fn == AGG_COUNT_STAR || fn == AGG_COUNT_REF;
      int64_t v = 0;
      bool ok = false;
      if (found) {
        const int64_t cv = cnt[(uint64_t)(fn == AGG_COUNT_REF ? aggs[a].prog : a) * cap + s];
        v = counting ? cv : acc[(uint64_t)a * cap + s];
        ok = counting ? true : cv > 0;
      }
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, ok);
      if ((threadIdx.x & 31) == 0) out_valid[a * valid_stride_words + (i >> 5)] = m;
      if (live) out_vals[a * out_stride + i] = ok ? v : 0;
    }
  }
}
void launch_win_gather(const Prog* progs, int nkeys, const AggSpec* aggs, int naggs, OpCtx c, RowCount n, const uint32_t* slots,
                       uint32_t cap_max, const int64_t* acc, const int64_t* cnt, int64_t* out_vals, uint32_t* out_valid, int64_t out_stride,
                       int64_t valid_stride_words, cudaStream_t s) {
  win_gather_kernel<<<grid_for(n.host), kThreads, 0, s>>>(progs, nkeys, aggs, naggs, c, n.dev, n.host, slots, cap_max, acc, cnt, out_vals,
                                                          out_valid, out_stride, valid_stride_words);
  NDS_CUDA(cudaGetLastError());
}

// Ranking window functions over rows SORTED by (partition keys, order keys): row j starts a partition when a partition key
// differs from row j - 1's, a tie group when an order key differs too (NULL equals NULL, as in GROUP BY).
//   rank_flags_kernel  m1[j] = j if row j starts a partition else 0;  m2[j] = j if it starts a tie group else 0;  d[j] = 1 / 0
//   rank_scan_kernel   in place: m1, m2 -> inclusive max (the row's partition / tie-group start), d -> inclusive sum
//   rank_out_kernel    rank = tie start - partition start + 1, row_number = j - partition start + 1,
//                      dense_rank = d[j] - d[partition start] + 1
__global__ void __launch_bounds__(kThreads)
rank_flags_kernel(const Prog* __restrict__ progs, int nkeys, int npart, OpCtx c, const int* n_dev, int64_t n_host,
                  uint32_t* __restrict__ m1, uint32_t* __restrict__ m2, uint32_t* __restrict__ d) {
  const int64_t n = row_count(n_dev, n_host);
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    bool newp = j == 0, newo = j == 0;
    for (int k = 0; k < nkeys && !newp && j > 0; ++k) {
      if (newo && k >= npart) break;
      const Val a = eval_program(progs[k].code, progs[k].n, c.cols, c.lit, j);
      const Val b = eval_program(progs[k].code, progs[k].n, c.cols, c.lit, j - 1);
      if (!val_equal(a, b, progs[k].kind)) { if (k < npart) newp = true; else newo = true; }
    }
    newo = newo || newp;
    m1[j] = newp ? (uint32_t)j : 0u;
    m2[j] = newo ? (uint32_t)j : 0u;
    d[j] = newo ? 1u : 0u;
  }
}
constexpr int kRankScanThreads = 1024;
__global__ void __launch_bounds__(kRankScanThreads)
rank_scan_kernel(const int* n_dev, int64_t n_host, uint32_t* __restrict__ m1, uint32_t* __restrict__ m2, uint32_t* __restrict__ d) {
  __shared__ uint32_t w1[32], w2[32], w3[32];
  __shared__ uint32_t c1, c2, c3;
  const int64_t n = row_count(n_dev, n_host);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) { c1 = 0; c2 = 0; c3 = 0; }
  __syncthreads();
  for (int64_t base = 0; base < n; base += kRankScanThreads) {
    const int64_t j = base + tid;
    uint32_t v1 = j < n ? m1[j] : 0u, v2 = j < n ? m2[j] : 0u, v3 = j < n ? d[j] : 0u;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t a = __shfl_up_sync(0xFFFFFFFFu, v1, o), b = __shfl_up_sync(0xFFFFFFFFu, v2, o), e = __shfl_up_sync(0xFFFFFFFFu, v3, o);
      if (lane >= o) { v1 = max(v1, a); v2 = max(v2, b); v3 += e; }
    }
    if (lane == 31) { w1[warp] = v1; w2[warp] = v2; w3[warp] = v3; }
    __syncthreads();
    uint32_t p1 = c1, p2 = c2, p3 = c3;
    for (int q = 0; q < warp; ++q) { p1 = max(p1, w1[q]); p2 = max(p2, w2[q]); p3 += w3[q]; }
    v1 = max(v1, p1);
    v2 = max(v2, p2);
    v3 += p3;
    if (j < n) { m1[j] = v1; m2[j] = v2; d[j] = v3; }
    __syncthreads();
    if (tid == kRankScanThreads - 1) { c1 = v1; c2 = v2; c3 = v3; }     // (rows past n carry zeros / nothing new: harmless)
    __syncthreads();
  }
}
__global__ void __launch_bounds__(kThreads)
rank_out_kernel(const int* n_dev, int64_t n_host, const uint32_t* __restrict__ m1, const uint32_t* __restrict__ m2,
                const uint32_t* __restrict__ d, int64_t* __restrict__ out_rank, int64_t* __restrict__ out_rown, int64_t* __restrict__ out_dense) {
  const int64_t n = row_count(n_dev, n_host);
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t ps = m1[j];
    if (out_rank) out_rank[j] = (int64_t)m2[j] - ps + 1;
    if (out_rown) out_rown[j] = j - ps + 1;
    if (out_dense) out_dense[j] = (int64_t)d[j] - d[ps] + 1;
  }
}
void launch_rank(const Prog* progs, int nkeys, int npart, OpCtx c, RowCount n, uint32_t* m1, uint32_t* m2, uint32_t* d, int64_t* out_rank,
                 int64_t* out_rown, int64_t* out_dense, cudaStream_t s) {
  rank_flags_kernel<<<grid_for(n.host), kThreads, 0, s>>>(progs, nkeys, npart, c, n.dev, n.host, m1, m2, d);
  NDS_CUDA(cudaGetLastError());
  rank_scan_kernel<<<1, kRankScanThreads, 0, s>>>(n.dev, n.host, m1, m2, d);
  NDS_CUDA(cudaGetLastError());
  rank_out_kernel<<<grid_for(n.host), kThreads, 0, s>>>(n.dev, n.host, m1, m2, d, out_rank, out_rown, out_dense);
  NDS_CUDA(cudaGetLastError());
}

// A keyless aggregate (no GROUP BY) yields exactly one row, also over an empty input: COUNT = 0, everything else NULL.
__global__ void agg_keyless_fix_kernel(const AggSpec* __restrict__ aggs, int naggs, int* ngroups, uint32_t* rep_out, int64_t* out_vals,
                                       uint32_t* out_valid, int64_t out_stride, int64_t valid_stride_words) {
  if (threadIdx.x != 0 || blockIdx.x != 0 || *ngroups != 0) return;
  *ngroups = 1;
  rep_out[0] = 0;
  for (int a = 0; a < naggs; ++a) {
    const int fn = aggs[a].fn;
    out_vals[a * out_stride] = 0;
    out_valid[a * valid_stride_words] = (fn == AGG_COUNT || fn == AGG_COUNT_STAR || fn == AGG_COUNT_REF) ? 1u : 0u;
  }
}
void launch_agg_keyless_fix(const AggSpec* aggs, int naggs, int* ngroups, uint32_t* rep_out, int64_t* out_vals, uint32_t* out_valid,
                            int64_t out_stride, int64_t valid_stride_words, cudaStream_t s) {
  agg_keyless_fix_kernel<<<1, 32, 0, s>>>(aggs, naggs, ngroups, rep_out, out_vals, out_valid, out_stride, valid_stride_words);
  NDS_CUDA(cudaGetLastError());
}

void launch_agg_init(const AggSpec* aggs, int naggs, RowCount n, uint32_t cap_max, uint32_t* slots, int64_t* acc, int64_t* cnt,
                     uint32_t* out_valid, int64_t valid_stride_words, cudaStream_t s) {
  const int64_t work = (int64_t)agg_capacity(n.host, cap_max) * std::max(naggs, 1);
  agg_init_kernel<<<grid_for(work), kThreads, 0, s>>>(aggs, naggs, n.dev, n.host, cap_max, slots, acc, cnt, out_valid,
                                                      valid_stride_words);
  NDS_CUDA(cudaGetLastError());
}
void launch_agg_insert(const Prog* progs, int nkeys, const AggSpec* aggs, int naggs, OpCtx c, RowCount n, uint32_t* slots,
                       uint32_t cap_max, int64_t* acc, int64_t* cnt, cudaStream_t s) {
  // shared-memory pre-aggregation needs 16 B per aggregate and cache entry; beyond 48 KB it is simply not used
  const size_t smem = (size_t)kAggCache * 4 + (size_t)std::max(naggs, 1) * kAggCache * 16;
  const int use_cache = smem <= 48 * 1024 ? 1 : 0;
  agg_insert_kernel<<<grid_for(n.host), kThreads, use_cache ? smem : 0, s>>>(progs, nkeys, aggs, naggs, c, n.dev, n.host, slots, cap_max,
                                                                            acc, cnt, use_cache);
  NDS_CUDA(cudaGetLastError());
}
void launch_agg_finalize(const AggSpec* aggs, int naggs, const uint32_t* slots, uint32_t cap_max, RowCount n, const int64_t* acc,
                         const int64_t* cnt, uint32_t* rep_out, int64_t* out_vals, uint32_t* out_valid, int64_t out_stride,
                         int64_t valid_stride_words, int* ngroups, cudaStream_t s) {
  agg_finalize_kernel<<<grid_for(agg_capacity(n.host, cap_max)), kThreads, 0, s>>>(aggs, naggs, slots, cap_max, n.dev, n.host, acc, cnt,
                                                                                  rep_out, out_vals, out_valid, out_stride,
                                                                                  valid_stride_words, ngroups);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ sort: key encoding
__global__ void str_maxlen_kernel(Prog p, OpCtx c, const int* n_dev, int64_t n_host, int* out_max) {
  const int64_t n = row_count(n_dev, n_host);
  int m = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    Val v = eval_program(p.code, p.n, c.cols, c.lit, i);
    if (v.ok && (int)v.len > m) m = (int)v.len;
  }
  for (int d = 16; d; d >>= 1) m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, d));
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out_max, m);
}
void launch_str_maxlen(Prog p, OpCtx c, RowCount n, int* out_max, cudaStream_t s) {
  str_maxlen_kernel<<<grid_for(n.host), kThreads, 0, s>>>(p, c, n.dev, n.host, out_max);
  NDS_CUDA(cudaGetLastError());
}

// Order-preserving big-endian byte strings: per key [flag][payload]. Spark's default null ordering
// (NULLS FIRST ascending, NULLS LAST descending) arrives as nulls_first; desc inverts the payload.
// While encoding, every byte is compared with row 0's: active[b] != 0 afterwards means byte b differs between
// rows, so the radix sort only needs passes over those bytes.
__device__ __forceinline__ void encode_row(const Prog* __restrict__ progs, const SortKey* __restrict__ keys, int nk, const OpCtx& c,
                                           int64_t i, uint8_t* row, const uint8_t* first, uint8_t* flag, int* err) {
  for (int k = 0; k < nk; ++k) {
    const SortKey sk = keys[k];
    Val v = eval_program(progs[sk.prog].code, progs[sk.prog].n, c.cols, c.lit, i);
    const int o = sk.offset;
    const uint8_t inv = sk.desc ? 0xFF : 0x00;
    auto put = [&](int j, uint8_t b) {
      row[o + j] = b;
      if (first && b != first[o + j]) flag[o + j] = 1;
    };
    put(0, sk.nulls_first ? (v.ok ? 1 : 0) : (v.ok ? 0 : 1));
    if (!v.ok) {
      for (int b = 0; b < sk.width; ++b) put(1 + b, 0);
    } else if (sk.kind == 2) {
      const uint8_t* p = (const uint8_t*)(uintptr_t)v.v;
      if ((int)v.len > sk.width) { if (err) atomicExch(err, 12); }
      for (int b = 0; b < sk.width; ++b) put(1 + b, (uint8_t)((b < (int)v.len ? p[b] : 0) ^ inv));
    } else {
      uint64_t u = (uint64_t)v.v;
      if (sk.kind == 1) u = (u >> 63) ? ~u : (u | 0x8000000000000000ULL);
      else u ^= 0x8000000000000000ULL;
      for (int b = 0; b < 8; ++b) put(1 + b, (uint8_t)((uint8_t)(u >> (56 - 8 * b)) ^ inv));
    }
  }
}

__global__ void __launch_bounds__(kThreads)
encode_keys_kernel(const Prog* __restrict__ progs, const SortKey* __restrict__ keys, int nk, OpCtx c, const int* n_dev,
                   int64_t n_host, uint8_t* __restrict__ keybuf, int B, int* active, int* err) {
  extern __shared__ uint8_t sm_keys[];
  uint8_t* first = sm_keys;
  uint8_t* flag = sm_keys + B;
  const int64_t n = row_count(n_dev, n_host);
  if (blockIdx.x * (int64_t)blockDim.x >= n) return;     // the grid is sized for the upper bound of the row count
  for (int b = threadIdx.x; b < B; b += blockDim.x) flag[b] = 0;
  if (threadIdx.x == 0) encode_row(progs, keys, nk, c, 0, first, nullptr, nullptr, nullptr);
  __syncthreads();
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    encode_row(progs, keys, nk, c, i, keybuf + i * B, first, flag, err);
  __syncthreads();
  for (int b = threadIdx.x; b < B; b += blockDim.x)
    if (flag[b]) atomicOr(&active[b], 1);
}
void launch_encode_keys(const Prog* progs, const SortKey* keys, int nk, OpCtx c, RowCount n, uint8_t* keybuf, int B, int* active,
                        int* err, cudaStream_t s) {
  encode_keys_kernel<<<grid_for(n.host), kThreads, 2 * B, s>>>(progs, keys, nk, c, n.dev, n.host, keybuf, B, active, err);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ sort: ORDER BY ... LIMIT k in one launch
// Streaming top-k: every CTA walks its share of the rows in 256-row tiles; a thread encodes its row's key (the same
// order-preserving byte string the radix route sorts) and keeps it only if it beats the CTA's current k-th best key.
// Kept keys collect in a 512-slot candidate area in shared memory; when it fills, a bitonic sort of the slot indices
// (45 compare-exchange steps, keys stay put) finds the k best, which move, sorted, into the other half of a ping-pong
// pair, and the k-th becomes the new threshold - after the first tiles almost every row fails the threshold compare
// and costs nothing more.  The last CTA to finish (atomic ticket) streams every CTA's <= k survivors through the same
// routine and writes the result rows.  One launch replaces encode + compact + four select passes + collect + rank
// sort; ties are broken by row id, so the order is total and the outcome deterministic.
constexpr int kTopkThreads = 256;
constexpr int kTopkCand = 512;
constexpr int kTopkSmemMax = 200 * 1024;      // candidate areas beyond this stay in global memory (very wide keys)

__device__ __forceinline__ uint64_t bswap64(uint64_t v) {
  const uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
  return ((uint64_t)__byte_perm(lo, 0, 0x0123) << 32) | (uint64_t)__byte_perm(hi, 0, 0x0123);
}
// (key a, row ra) ordered before (key b, row rb)?  keys: Bw 8-byte words holding big-endian byte strings.
// CG: a was written by another CTA - read it through L2
template <bool CG>
__device__ __forceinline__ bool topk_less(const uint64_t* a, uint32_t ra, const uint64_t* b, uint32_t rb, int Bw) {
  for (int w = 0; w < Bw; ++w) {
    const uint64_t x = CG ? __ldcg(a + w) : a[w], y = b[w];
    if (x != y) return bswap64(x) < bswap64(y);
  }
  return ra < rb;
}

struct TopkState {
  uint64_t* area[2];        // ping-pong candidate key areas of this CTA, kTopkCand slots of Bw words (shared or global)
  uint32_t* row[2];         // row ids of the candidates (shared memory)
  uint16_t* perm;           // kTopkCand slot indices (shared memory)
  int cur;                  // which area collects
  int Bw, k;
  bool have_thresh;         // area[cur][0 .. k) is the sorted k best of everything offered so far
  bool sorted;              // nothing was appended since the last shrink: area[cur][0 .. count) is sorted
};

// CTA-wide: keep the k best of the m collected candidates, sorted, in the other area.  Returns the new count.
__device__ __forceinline__ int topk_shrink(TopkState& st, int m, int* s_m) {
  const int tid = threadIdx.x;
  const uint64_t* A = st.area[st.cur];
  const uint32_t* rA = st.row[st.cur];
  for (int i = tid; i < kTopkCand; i += kTopkThreads) st.perm[i] = (uint16_t)i;
  __syncthreads();
  // bitonic sort of the slot indices over the next power of two >= m; slots >= m rank after everything
  int span = 2;
  while (span < m) span <<= 1;
  for (int size = 2; size <= span; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      // strides below 32 pair elements inside the 64-element chunk a warp owns (thread t <-> elements 2t - t % stride and
      // + stride): only the warp itself wrote them in the previous step
      if (stride >= 32) __syncthreads(); else __syncwarp();
      const int i = 2 * tid - (tid & (stride - 1)), j = i + stride;
      if (j < span) {
        const bool up = (i & size) == 0;
        const int a = st.perm[i], b = st.perm[j];
        bool b_before_a;
        if (b >= m) b_before_a = (a >= m) && b < a;
        else if (a >= m) b_before_a = true;
        else b_before_a = topk_less<false>(A + (int64_t)b * st.Bw, rA[b], A + (int64_t)a * st.Bw, rA[a], st.Bw);
        if (b_before_a == up) { st.perm[i] = (uint16_t)b; st.perm[j] = (uint16_t)a; }
      }
    }
  }
  __syncthreads();
  const int kept = m < st.k ? m : st.k;
  uint64_t* B = st.area[st.cur ^ 1];
  uint32_t* rB = st.row[st.cur ^ 1];
  for (int idx = tid; idx < kept * st.Bw; idx += kTopkThreads) {
    const int r = idx / st.Bw, w = idx - r * st.Bw;
    B[(int64_t)r * st.Bw + w] = A[(int64_t)st.perm[r] * st.Bw + w];
  }
  for (int r = tid; r < kept; r += kTopkThreads) rB[r] = rA[st.perm[r]];
  if (tid == 0) *s_m = kept;
  __syncthreads();
  st.cur ^= 1;
  st.have_thresh = kept == st.k;
  st.sorted = true;
  return kept;
}

// CTA-wide: offer one candidate per thread (have = this thread has one; key = its Bw words in global memory).
template <bool CG>
__device__ __forceinline__ void topk_offer(TopkState& st, bool have, uint32_t row, const uint64_t* key, int* s_m) {
  bool keep = have;
  if (keep && st.have_thresh)
    keep = topk_less<CG>(key, row, st.area[st.cur] + (int64_t)(st.k - 1) * st.Bw, st.row[st.cur][st.k - 1], st.Bw);
  if (keep) {
    const int pos = atomicAdd(s_m, 1);            // at most kTopkThreads appends per call: the area cannot overflow
    uint64_t* dst = st.area[st.cur] + (int64_t)pos * st.Bw;
    for (int w = 0; w < st.Bw; ++w) dst[w] = CG ? __ldcg(key + w) : key[w];
    st.row[st.cur][pos] = row;
  }
  if (__syncthreads_or(keep)) st.sorted = false;
  const int m = *s_m;
  __syncthreads();
  if (m > kTopkCand - kTopkThreads) topk_shrink(st, m, s_m);
}

// CTA-wide: stream the survivor lists of sources [s0, s1) (at most kTopkMaxSrc of them, <= k entries each) through the
// top-k routine, started afresh: the lists are taken as one flat sequence, kTopkThreads candidates per round.
// sort_result: the result must come out sorted (the final merge); otherwise <= k unsorted survivors are fine.
constexpr int kTopkMaxSrc = 64;
__device__ __forceinline__ int topk_merge(TopkState& st, int s0, int s1, const uint64_t* keys_base, const uint32_t* rows_base,
                                          const int* counts, int* s_m, int* s_pre, bool sort_result) {
  const int tid = threadIdx.x;
  const int ns = s1 - s0;
  st.cur = 0;
  st.have_thresh = false;
  st.sorted = false;
  if (tid == 0) *s_m = 0;
  if (tid < ns) s_pre[tid + 1] = __ldcg(counts + s0 + tid);
  __syncthreads();
  if (tid == 0) {
    s_pre[0] = 0;
    for (int q = 0; q < ns; ++q) s_pre[q + 1] += s_pre[q];
  }
  __syncthreads();
  const int total = s_pre[ns];
  for (int c0 = 0; c0 < total; c0 += kTopkThreads) {
    const int cidx = c0 + tid;
    const bool have = cidx < total;
    int src = 0, e = 0;
    if (have) {
      int lo = 0, hi = ns - 1;                     // the source whose range holds candidate cidx
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (s_pre[mid] <= cidx) lo = mid; else hi = mid - 1;
      }
      src = s0 + lo;
      e = cidx - s_pre[lo];
    }
    const uint32_t row = have ? __ldcg(rows_base + (int64_t)src * st.k + e) : 0u;
    topk_offer<true>(st, have, row, keys_base + ((int64_t)src * st.k + e) * st.Bw, s_m);
  }
  const int m = *s_m;
  if (m <= st.k && (!sort_result || st.sorted)) return m;
  return topk_shrink(st, m, s_m);
}

// sorted survivors of this CTA -> slot `slot` of a (keys, rows, counts) triple in global memory
__device__ __forceinline__ void topk_publish(const TopkState& st, int kept, int slot, uint64_t* keys_base, uint32_t* rows_base, int* counts) {
  const int tid = threadIdx.x;
  const uint64_t* A = st.area[st.cur];
  uint64_t* fin = keys_base + (int64_t)slot * st.k * st.Bw;
  for (int idx = tid; idx < kept * st.Bw; idx += kTopkThreads) fin[idx] = A[idx];
  for (int i = tid; i < kept; i += kTopkThreads) rows_base[(int64_t)slot * st.k + i] = st.row[st.cur][i];
  if (tid == 0) counts[slot] = kept;
}

static_assert((148 + 15) / 16 <= kTopkMaxSrc, "final merge fan-in");
constexpr int kTopkGroup = 16;      // CTAs whose survivors one CTA merges before the final merge (two-level last-CTA pattern)

__global__ void __launch_bounds__(kTopkThreads)
topk_stream_kernel(const Prog* __restrict__ progs, const SortKey* __restrict__ keys, int nk, OpCtx c, const int* n_dev, int64_t n_host,
                   int B, int Bw, int k, int smem_areas, uint64_t* scratch, uint64_t* fin_keys, uint32_t* fin_rows, int* fin_count,
                   int* tickets, uint32_t* out_rows, int* out_count, int* err) {
  extern __shared__ __align__(16) uint64_t topk_smem[];
  __shared__ uint32_t s_row[2][kTopkCand];
  __shared__ uint16_t s_perm[kTopkCand];
  __shared__ int s_m;
  __shared__ int s_last;
  __shared__ int s_pre[kTopkMaxSrc + 1];
  const int tid = threadIdx.x;
  const int64_t n = row_count(n_dev, n_host);
  // The grid is sized for the host's bound of the row count; the count itself is only known here.  A CTA gets at least two
  // tiles of rows (one candidate area), the CTAs beyond that leave at once and take no part in the merges: a thousand rows
  // are two CTAs and one merge, instead of 148 publishes, ten group merges and a final merge of mostly empty lists.
  const int64_t want = (n + 2 * kTopkThreads - 1) / (2 * kTopkThreads);
  const int nctas = (int)(want < 1 ? 1 : (want > (int64_t)gridDim.x ? (int64_t)gridDim.x : want));
  if ((int)blockIdx.x >= nctas) return;
  const int ngroups = (nctas + kTopkGroup - 1) / kTopkGroup;
  // global scratch per CTA: 256 staging slots (+ the two candidate areas when they do not fit shared memory)
  const int64_t per_cta = (int64_t)(kTopkThreads + (smem_areas ? 0 : 2 * kTopkCand)) * Bw;
  uint64_t* my = scratch + (int64_t)blockIdx.x * per_cta;
  TopkState st;
  st.area[0] = smem_areas ? topk_smem : my + (int64_t)kTopkThreads * Bw;
  st.area[1] = st.area[0] + (int64_t)kTopkCand * Bw;
  st.row[0] = s_row[0];
  st.row[1] = s_row[1];
  st.perm = s_perm;
  st.cur = 0;
  st.Bw = Bw;
  st.k = k;
  st.have_thresh = false;
  st.sorted = false;
  if (tid == 0) s_m = 0;
  __syncthreads();
  uint64_t* stage = my + (int64_t)tid * Bw;
  uint8_t* mykey = reinterpret_cast<uint8_t*>(stage);
  for (int64_t t0 = (int64_t)blockIdx.x * kTopkThreads; t0 < n; t0 += (int64_t)nctas * kTopkThreads) {
    const int64_t i = t0 + tid;
    const bool have = i < n;
    if (have) {
      encode_row(progs, keys, nk, c, i, mykey, nullptr, nullptr, err);
      for (int b = B; b < Bw * 8; ++b) mykey[b] = 0;
    }
    topk_offer<false>(st, have, (uint32_t)i, stage, &s_m);
  }
  // (<= k collected candidates are handed on as they are: only the result has to be sorted)
  if (nctas == 1) {                     // the only CTA: its sorted candidates are the result
    const int res = (st.sorted && s_m <= k) ? s_m : topk_shrink(st, s_m, &s_m);
    for (int i = tid; i < res; i += kTopkThreads) out_rows[i] = st.row[st.cur][i];
    if (tid == 0) *out_count = res;
    return;
  }
  int kept = s_m > k ? topk_shrink(st, s_m, &s_m) : s_m;
  // level 0 -> level 1: this CTA's survivors; the last CTA of each group of kTopkGroup merges the group's lists
  uint64_t* keys1 = fin_keys;
  uint32_t* rows1 = fin_rows;
  int* count1 = fin_count;
  uint64_t* keys2 = fin_keys + (int64_t)nctas * k * Bw;
  uint32_t* rows2 = fin_rows + (int64_t)nctas * k;
  int* count2 = fin_count + nctas;
  topk_publish(st, kept, (int)blockIdx.x, keys1, rows1, count1);
  const int g = (int)blockIdx.x / kTopkGroup;
  const int g0 = g * kTopkGroup, g1 = min(g0 + kTopkGroup, nctas);
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = atomicAdd(tickets + 1 + g, 1) == (g1 - g0) - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  kept = topk_merge(st, g0, g1, keys1, rows1, count1, &s_m, s_pre, ngroups == 1);
  if (ngroups == 1) {                   // one group: its merge is the final one
    for (int i = tid; i < kept; i += kTopkThreads) out_rows[i] = st.row[st.cur][i];
    if (tid == 0) *out_count = kept;
    return;
  }
  // level 1 -> result: the last group to finish merges the groups' lists and writes the rows
  topk_publish(st, kept, g, keys2, rows2, count2);
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = atomicAdd(tickets, 1) == ngroups - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  kept = topk_merge(st, 0, ngroups, keys2, rows2, count2, &s_m, s_pre, true);
  for (int i = tid; i < kept; i += kTopkThreads) out_rows[i] = st.row[st.cur][i];
  if (tid == 0) *out_count = kept;
}

int topk_stream_ctas(int64_t n_bound) {
  int64_t g = (n_bound + kTopkThreads - 1) / kTopkThreads;
  // at most one CTA per SM: the candidate areas can take most of an SM's shared memory, and a CTA without rows still costs a
  // scheduling slot, a publish and a ticket
  return (int)(g < 1 ? 1 : (g > 148 ? 148 : g));
}
int topk_stream_groups(int ctas) { return (ctas + kTopkGroup - 1) / kTopkGroup; }
static bool topk_areas_in_smem(int Bw) { return (int64_t)2 * kTopkCand * Bw * 8 <= kTopkSmemMax; }
int64_t topk_stream_scratch_words(int ctas, int Bw) {
  return (int64_t)ctas * (kTopkThreads + (topk_areas_in_smem(Bw) ? 0 : 2 * kTopkCand)) * Bw;
}

void launch_topk_stream(const Prog* progs, const SortKey* keys, int nk, OpCtx c, RowCount n, int B, int k, int ctas, uint64_t* scratch,
                        uint64_t* fin_keys, uint32_t* fin_rows, int* fin_count, int* tickets, uint32_t* out_rows, int* out_count, int* err,
                        cudaStream_t s) {
  const int Bw = (B + 7) / 8;
  const bool in_smem = topk_areas_in_smem(Bw);
  const size_t smem = in_smem ? (size_t)2 * kTopkCand * Bw * 8 : 0;
  static bool attr_set[64] = {false};
  int dev = 0;
  NDS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    NDS_CUDA(cudaFuncSetAttribute(topk_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTopkSmemMax));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  topk_stream_kernel<<<ctas, kTopkThreads, smem, s>>>(progs, keys, nk, c, n.dev, n.host, B, Bw, k, in_smem ? 1 : 0, scratch, fin_keys, fin_rows,
                                                      fin_count, tickets, out_rows, out_count, err);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ sort: LSD radix, 8-bit digits
// Pass p (p = 0 is the least significant byte B-1) reads the live permutation buffer which[p], and, unless all rows
// share the digit (skip), scatters into the other buffer; which[p+1] records the result. Three kernels per
// pass: per-block histograms, one-block scan, stable scatter.
constexpr int kSortThreads = 256;
constexpr int kSortTile = 4096;     // rows per block

int radix_sort_grid(int64_t n_max) {
  int64_t g = (n_max + kSortTile - 1) / kSortTile;
  return (int)(g < 1 ? 1 : g);
}

__global__ void radix_init_kernel(const int* n_dev, int64_t n_host, uint32_t* perm_a, int* which) {
  const int64_t n = row_count(n_dev, n_host);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) perm_a[i] = (uint32_t)i;
  if (blockIdx.x == 0 && threadIdx.x == 0) which[0] = 0;
}

__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const uint8_t* __restrict__ keybuf, int B, int byte, const int* n_dev, int64_t n_host,
                  const uint32_t* perm_a, const uint32_t* perm_b, const int* which, int pass, uint32_t* __restrict__ block_hist) {
  __shared__ uint32_t h[256];
  const int64_t n = row_count(n_dev, n_host);
  const uint32_t* perm = which[pass] ? perm_b : perm_a;
  h[threadIdx.x] = 0;
  __syncthreads();
  const int64_t lo = (int64_t)blockIdx.x * kSortTile;
  const int64_t hi = min(lo + (int64_t)kSortTile, n);
  for (int64_t i = lo + threadIdx.x; i < hi; i += kSortThreads) atomicAdd(&h[keybuf[(int64_t)perm[i] * B + byte]], 1u);
  __syncthreads();
  block_hist[(int64_t)blockIdx.x * 256 + threadIdx.x] = h[threadIdx.x];
}

// one block of 256 threads: thread d owns digit d
__global__ void __launch_bounds__(256)
radix_scan_kernel(uint32_t* block_hist, int nblocks, const int* n_dev, int64_t n_host, int* which, int pass, uint32_t* skip_flag) {
  __shared__ uint32_t tot[256];
  __shared__ int s_skip;
  const int64_t n = row_count(n_dev, n_host);
  const int d = threadIdx.x;
  if (d == 0) s_skip = 0;
  __syncthreads();
  uint32_t sum = 0;
  for (int g = 0; g < nblocks; ++g) sum += block_hist[(int64_t)g * 256 + d];
  tot[d] = sum;
  if ((int64_t)sum == n) s_skip = 1;      // every row has this digit (covers n == 0 via d == any)
  __syncthreads();
  // exclusive scan over digits (256 values, serial per thread is fine: 256 adds)
  uint32_t base = 0;
  for (int k = 0; k < d; ++k) base += tot[k];
  uint32_t run = base;
  for (int g = 0; g < nblocks; ++g) {
    uint32_t c = block_hist[(int64_t)g * 256 + d];
    block_hist[(int64_t)g * 256 + d] = run;
    run += c;
  }
  if (d == 0) {
    int skip = s_skip || n <= 1;
    skip_flag[pass] = skip;
    which[pass + 1] = which[pass] ^ (skip ? 0 : 1);
  }
}

__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const uint8_t* __restrict__ keybuf, int B, int byte, const int* n_dev, int64_t n_host, uint32_t* perm_a,
                     uint32_t* perm_b, const int* which, int pass, const uint32_t* __restrict__ block_off,
                     const uint32_t* __restrict__ skip_flag) {
  if (skip_flag[pass]) return;
  __shared__ uint32_t base[256];
  __shared__ uint16_t wcnt[kSortThreads / 32][256];
  const int64_t n = row_count(n_dev, n_host);
  const uint32_t* src = which[pass] ? perm_b : perm_a;
  uint32_t* dst = which[pass] ? perm_a : perm_b;
  const int tid = threadIdx.x, warp = tid >> 5;
  base[tid] = block_off[(int64_t)blockIdx.x * 256 + tid];
  const int64_t lo = (int64_t)blockIdx.x * kSortTile;
  const int64_t hi = min(lo + (int64_t)kSortTile, n);
  for (int64_t c0 = lo; c0 < hi; c0 += kSortThreads) {
    for (int k = tid; k < (kSortThreads / 32) * 256; k += kSortThreads) (&wcnt[0][0])[k] = 0;
    __syncthreads();
    const int64_t i = c0 + tid;
    const bool act = i < hi;
    uint32_t row = act ? src[i] : 0u;
    int d = act ? keybuf[(int64_t)row * B + byte] : 256 + (tid & 31);   // inactive lanes get unique pseudo-digits
    uint32_t m = __match_any_sync(0xFFFFFFFFu, d);
    uint32_t rank = __popc(m & lanemask_lt());
    if (act && rank == 0) wcnt[warp][d] = (uint16_t)__popc(m);
    __syncthreads();
    if (act) {
      uint32_t pos = base[d] + rank;
      for (int w = 0; w < warp; ++w) pos += wcnt[w][d];
      dst[pos] = row;
    }
    __syncthreads();
    uint32_t add = 0;
    for (int w = 0; w < kSortThreads / 32; ++w) add += wcnt[w][tid];
    base[tid] += add;
    __syncthreads();
  }
}

void launch_radix_sort(const uint8_t* keybuf, int B, const int* passes, int npasses, RowCount n, uint32_t* perm_a,
                       uint32_t* perm_b, int* which, uint32_t* scratch, cudaStream_t s, int* launches) {
  const int G = radix_sort_grid(n.host);
  uint32_t* block_hist = scratch;
  uint32_t* skip_flag = scratch + (int64_t)G * 256;
  radix_init_kernel<<<grid_for(n.host), kThreads, 0, s>>>(n.dev, n.host, perm_a, which);
  NDS_CUDA(cudaGetLastError());
  *launches += 1;
  for (int pass = 0; pass < npasses; ++pass) {
    const int byte = passes[pass];
    radix_hist_kernel<<<G, kSortThreads, 0, s>>>(keybuf, B, byte, n.dev, n.host, perm_a, perm_b, which, pass, block_hist);
    radix_scan_kernel<<<1, 256, 0, s>>>(block_hist, G, n.dev, n.host, which, pass, skip_flag);
    radix_scatter_kernel<<<G, kSortThreads, 0, s>>>(keybuf, B, byte, n.dev, n.host, perm_a, perm_b, which, pass, block_hist,
                                                    skip_flag);
    *launches += 3;
  }
  NDS_CUDA(cudaGetLastError());
}

// One CTA sorts up to kSmallSortMax rows: all passes in one launch, bytes whose active flag is 0 are skipped.
// which[0] receives the index of the buffer holding the result.
constexpr int kSmallThreads = 1024;
__global__ void __launch_bounds__(kSmallThreads)
radix_sort_small_kernel(const uint8_t* __restrict__ keybuf, int B, const int* __restrict__ active, const int* n_dev, int64_t n_host,
                        uint32_t* perm_a, uint32_t* perm_b, int* which) {
  __shared__ uint32_t hist[256];
  __shared__ uint32_t base[256];
  __shared__ uint32_t wtot[8];
  __shared__ uint16_t wcnt[kSmallThreads / 32][256];
  __shared__ uint8_t s_active[512];
  const int n = (int)row_count(n_dev, n_host);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < n; i += kSmallThreads) perm_a[i] = (uint32_t)i;
  for (int b = tid; b < B && b < 512; b += kSmallThreads) s_active[b] = active[b] ? 1 : 0;
  int cur = 0;
  __syncthreads();
  for (int byte = B - 1; byte >= 0; --byte) {
    if (byte < 512 ? !s_active[byte] : !active[byte]) continue;
    const uint32_t* src = cur ? perm_b : perm_a;
    uint32_t* dst = cur ? perm_a : perm_b;
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += kSmallThreads) atomicAdd(&hist[keybuf[(int64_t)src[i] * B + byte]], 1u);
    __syncthreads();
    // exclusive scan of the 256 bins: warp shuffles + 8 warp totals
    uint32_t hv = 0, incl = 0;
    if (tid < 256) {
      hv = hist[tid];
      incl = hv;
      for (int d = 1; d < 32; d <<= 1) {
        uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += v;
      }
      if (lane == 31) wtot[warp] = incl;
    }
    __syncthreads();
    if (tid < 256) {
      uint32_t b = 0;
      for (int w = 0; w < warp; ++w) b += wtot[w];
      base[tid] = b + incl - hv;
    }
    __syncthreads();
    for (int c0 = 0; c0 < n; c0 += kSmallThreads) {
      for (int k = tid; k < (kSmallThreads / 32) * 256; k += kSmallThreads) (&wcnt[0][0])[k] = 0;
      __syncthreads();
      const int i = c0 + tid;
      const bool act = i < n;
      uint32_t row = act ? src[i] : 0u;
      int d = act ? keybuf[(int64_t)row * B + byte] : 256 + (tid & 31);
      uint32_t m = __match_any_sync(0xFFFFFFFFu, d);
      uint32_t rank = __popc(m & lanemask_lt());
      if (act && rank == 0) wcnt[warp][d] = (uint16_t)__popc(m);
      __syncthreads();
      if (act) {
        uint32_t pos = base[d] + rank;
        for (int w = 0; w < warp; ++w) pos += wcnt[w][d];
        dst[pos] = row;
      }
      __syncthreads();
      if (tid < 256) {
        uint32_t add = 0;
        for (int w = 0; w < kSmallThreads / 32; ++w) add += wcnt[w][tid];
        base[tid] += add;
      }
      __syncthreads();
    }
    cur ^= 1;
    __syncthreads();
  }
  if (tid == 0) which[0] = cur;
}

void launch_radix_sort_small(const uint8_t* keybuf, int B, const int* active, RowCount n, uint32_t* perm_a, uint32_t* perm_b,
                             int* which, cudaStream_t s) {
  radix_sort_small_kernel<<<1, kSmallThreads, 0, s>>>(keybuf, B, active, n.dev, n.host, perm_a, perm_b, which);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ sort: compacted keys, top-k select, rank sort
// The encoded keys are mostly constant bytes (flags, high bytes of small integers, string padding, shared prefixes).
// compact_keys_kernel keeps only the bytes that differ between rows (active[]), packed big-endian into 64-bit words,
// so that an unsigned integer compare of word 0, then word 1, ... is the lexicographic compare of the full key.
constexpr int kCompactThreads = 256;
__global__ void __launch_bounds__(kCompactThreads)
compact_keys_kernel(const uint8_t* __restrict__ keybuf, int B, const int* __restrict__ active, const int* n_dev, int64_t n_host,
                    uint64_t* __restrict__ ckeys, int W, int* w_out, uint32_t* zero_rank, int* zero_tiles, int64_t zero_n) {
  __shared__ uint16_t list[kMaxCompactBytes];
  __shared__ int wsum[kCompactThreads / 32];
  __shared__ int s_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int b0 = 0; b0 < B; b0 += kCompactThreads) {
    const int b = b0 + tid;
    const bool f = b < B && active[b] != 0;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, f);
    if (lane == 0) wsum[warp] = __popc(bal);
    __syncthreads();
    int pre = s_base;
    for (int w = 0; w < warp; ++w) pre += wsum[w];
    if (f) list[pre + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)b;
    __syncthreads();
    if (tid == 0) { int t = 0; for (int w = 0; w < kCompactThreads / 32; ++w) t += wsum[w]; s_base += t; }
    __syncthreads();
  }
  const int nact = s_base;
  const int wact = nact ? (nact + 7) >> 3 : 1;
  if (blockIdx.x == 0 && tid == 0) w_out[0] = wact;
  const int64_t n = row_count(n_dev, n_host);
  if (zero_rank)      // the rank sort that follows accumulates into these
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + tid; i < zero_n; i += (int64_t)gridDim.x * blockDim.x) {
      zero_rank[i] = 0;
      if ((i & 127) == 0) zero_tiles[i >> 7] = 0;
    }
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + tid; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint8_t* row = keybuf + i * B;
    for (int w = 0; w < wact; ++w) {
      uint64_t v = 0;
      for (int b = 0; b < 8; ++b) {
        const int k = 8 * w + b;
        v = (v << 8) | (k < nact ? (uint64_t)row[list[k]] : 0ull);
      }
      ckeys[i * W + w] = v;
    }
  }
}
void launch_compact_keys(const uint8_t* keybuf, int B, const int* active, RowCount n, uint64_t* ckeys, int W, int* w_out,
                         uint32_t* zero_rank, int* zero_tiles, int64_t zero_n, cudaStream_t s) {
  compact_keys_kernel<<<grid_for(n.host), kCompactThreads, 0, s>>>(keybuf, B, active, n.dev, n.host, ckeys, W, w_out, zero_rank,
                                                                    zero_tiles, zero_n);
  NDS_CUDA(cudaGetLastError());
}

// Top-k prefilter (ORDER BY ... LIMIT k over many rows).  Word 0 of the compacted keys is radix-selected most
// significant byte first: pass q histograms byte q of the rows that match the prefix chosen so far (multi-CTA, one
// launch per pass, at most kSelectPasses), and every CTA re-derives the prefix from the earlier histograms, so no
// state other than hist[][] passes between launches.  As soon as (rows below the prefix) + (rows in the chosen
// bucket) fits the candidate list, later passes are no-ops.  topk_collect_kernel then lists every row whose masked
// word 0 is <= the prefix: the k smallest keys are all in that list, ties only make it longer.  nc_out[0] = list
// length; when it exceeds cap only cap entries were stored and the caller must take another route.
constexpr int kSelectThreads = 256;
constexpr int kSelectTile = 8192;       // rows per CTA and pass
__device__ __forceinline__ void topk_state(const uint32_t* __restrict__ hist, int npass, int64_t k, int cap, int64_t n,
                                           uint64_t* s_prefix, uint64_t* s_mask, int* s_done) {
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    uint64_t prefix = 0, mask = 0;
    int64_t kk = k, less = 0;
    int done = n <= (int64_t)cap;        // everything fits the candidate list: empty mask, every row is taken
    for (int q = 0; q < npass && !done; ++q) {
      uint32_t h[8];
      uint32_t sum = 0;
      for (int b = 0; b < 8; ++b) { h[b] = hist[q * 256 + lane * 8 + b]; sum += h[b]; }
      uint32_t incl = sum;
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += t;
      }
      const uint32_t excl = incl - sum;
      const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
      if (total == 0) { done = 1; break; }
      const int64_t ke = kk < (int64_t)total ? kk : (int64_t)total;
      const bool mine = (int64_t)excl < ke && ke <= (int64_t)incl;
      const uint32_t bal = __ballot_sync(0xFFFFFFFFu, mine);
      const int src = __ffs(bal) - 1;
      uint32_t dsel = 0, before = 0, bucket = 0;
      if (lane == src) {
        uint32_t c = excl;
        for (int b = 0; b < 8; ++b) {
          if ((int64_t)c + h[b] >= ke) { dsel = lane * 8 + b; before = c; bucket = h[b]; break; }
          c += h[b];
        }
      }
      dsel = __shfl_sync(0xFFFFFFFFu, dsel, src);
      before = __shfl_sync(0xFFFFFFFFu, before, src);
      bucket = __shfl_sync(0xFFFFFFFFu, bucket, src);
      const int shift = 56 - 8 * q;
      prefix |= (uint64_t)dsel << shift;
      mask |= 0xFFull << shift;
      less += before;
      kk = ke - before;
      if (less + (int64_t)bucket <= (int64_t)cap) done = 1;
    }
    if (lane == 0) { *s_prefix = prefix; *s_mask = mask; *s_done = done; }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kSelectThreads)
topk_hist_kernel(const uint64_t* __restrict__ ckeys, int W, const int* n_dev, int64_t n_host, int64_t k, int cap, uint32_t* hist,
                 int pass) {
  __shared__ uint32_t h[256];
  __shared__ uint64_t s_prefix, s_mask;
  __shared__ int s_done;
  const int tid = threadIdx.x, lane = tid & 31;
  const int64_t n = row_count(n_dev, n_host);
  const int64_t lo = (int64_t)blockIdx.x * kSelectTile, hi = min(n, lo + (int64_t)kSelectTile);
  if (lo >= n) return;
  h[tid] = 0;
  topk_state(hist, pass, k, cap, n, &s_prefix, &s_mask, &s_done);
  if (s_done) return;
  const uint64_t prefix = s_prefix, mask = s_mask;
  const int shift = 56 - 8 * pass;
  for (int64_t i0 = lo; i0 < hi; i0 += kSelectThreads) {
    const int64_t i = i0 + tid;
    bool in = false;
    uint32_t d = 256 + lane;
    if (i < hi) {
      const uint64_t v = ckeys[i * W];
      in = (v & mask) == prefix;
      if (in) d = (uint32_t)(v >> shift) & 255u;
    }
    const uint32_t m = __match_any_sync(0xFFFFFFFFu, d);
    if (in && (m & ((1u << lane) - 1u)) == 0) atomicAdd(&h[d], (uint32_t)__popc(m));
  }
  __syncthreads();
  if (h[tid]) atomicAdd(&hist[pass * 256 + tid], h[tid]);
}

__global__ void __launch_bounds__(kSelectThreads)
topk_collect_kernel(const uint64_t* __restrict__ ckeys, int W, const int* n_dev, int64_t n_host, int64_t k, int cap,
                    const uint32_t* hist, int npass, uint32_t* __restrict__ cand, int* nc_out) {
  __shared__ uint64_t s_prefix, s_mask;
  __shared__ int s_done;
  const int tid = threadIdx.x, lane = tid & 31;
  const int64_t n = row_count(n_dev, n_host);
  const int64_t lo = (int64_t)blockIdx.x * kSelectTile, hi = min(n, lo + (int64_t)kSelectTile);
  if (lo >= n) return;
  topk_state(hist, npass, k, cap, n, &s_prefix, &s_mask, &s_done);
  const uint64_t prefix = s_prefix, mask = s_mask;
  for (int64_t i0 = lo; i0 < hi; i0 += kSelectThreads) {
    const int64_t i = i0 + tid;
    const bool take = i < hi && (ckeys[i * W] & mask) <= prefix;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, take);
    int base = 0;
    if (lane == 0 && bal) base = atomicAdd(nc_out, __popc(bal));
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (take) {
      const int pos = base + __popc(bal & ((1u << lane) - 1u));
      if (pos < cap) cand[pos] = (uint32_t)i;
    }
  }
}
void launch_topk_select(const uint64_t* ckeys, int W, RowCount n, int64_t k, uint32_t* cand, int cap, uint32_t* hist, int* nc_out,
                        cudaStream_t s, int* launches) {
  const int G = (int)std::max<int64_t>(1, (n.host + kSelectTile - 1) / kSelectTile);
  for (int pass = 0; pass < kSelectPasses; ++pass)
    topk_hist_kernel<<<G, kSelectThreads, 0, s>>>(ckeys, W, n.dev, n.host, k, cap, hist, pass);
  topk_collect_kernel<<<G, kSelectThreads, 0, s>>>(ckeys, W, n.dev, n.host, k, cap, hist, kSelectPasses, cand, nc_out);
  NDS_CUDA(cudaGetLastError());
  *launches += kSelectPasses + 1;
}

// Rank sort: row i's position is the number of rows that order before it (ties broken by row id, which makes it
// the same permutation a stable sort yields).  n^2 compares, but they are independent, mostly decided on word 0,
// and spread over gridDim.x row tiles x gridDim.y slices of the compared range - for the <= kRankSortMax rows an
// NDS top-k sees, this is a few microseconds where 8-bit LSD passes cost one launch or several block barriers per
// key byte.  Rows whose rank is below the limit write their id to out_rows[rank]: sort and LIMIT in one step.
constexpr int kRankThreads = 128;
__global__ void __launch_bounds__(kRankThreads)
rank_sort_kernel(const uint64_t* __restrict__ ckeys, int W, const int* __restrict__ w_dev, const uint32_t* __restrict__ cand,
                 const int* nc_dev, int cand_cap, const int* n_dev, int64_t n_host, int64_t limit, uint32_t* rank_tmp,
                 int* tile_done, uint32_t* __restrict__ out_rows, int* out_count, int* overflow) {
  __shared__ uint64_t s_k0[kRankThreads], s_k1[kRankThreads];
  __shared__ uint32_t s_row[kRankThreads];
  __shared__ int s_last;
  const int tid = threadIdx.x;
  const int64_t n_all = row_count(n_dev, n_host);
  if (cand ? nc_dev[0] > cand_cap : n_all > (int64_t)gridDim.x * kRankThreads) {
    // more candidates / rows than this launch was sized for: tell the host (it re-runs the sort on the radix route)
    // the rows the caller will gather must still be valid ids: the result is thrown away, but it is computed
    if (blockIdx.x == 0 && blockIdx.y == 0) {
      if (tid == 0 && overflow) atomicExch(overflow, 1);
      const int64_t mo = (limit >= 0 && limit < n_all) ? limit : n_all;
      for (int64_t i = tid; i < mo; i += kRankThreads) out_rows[i] = 0;
    }
    return;
  }
  const int64_t n = cand ? min((int64_t)nc_dev[0], n_all) : n_all;
  const int64_t m = (limit >= 0 && limit < n_all) ? limit : n_all;
  if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) out_count[0] = (int)m;
  if ((int64_t)blockIdx.x * kRankThreads >= n) return;
  const int wact = w_dev[0];
  const int64_t gi = (int64_t)blockIdx.x * kRankThreads + tid;
  const bool valid = gi < n;
  const uint32_t ri = valid ? (cand ? cand[gi] : (uint32_t)gi) : 0u;
  const uint64_t* mine = ckeys + (int64_t)ri * W;
  const uint64_t k0 = valid ? mine[0] : 0ull;
  const uint64_t k1 = (valid && wact > 1) ? mine[1] : 0ull;
  const int J = gridDim.y;
  int64_t chunk = (n + J - 1) / J;
  chunk = (chunk + kRankThreads - 1) / kRankThreads * kRankThreads;
  const int64_t j0 = (int64_t)blockIdx.y * chunk;
  const int64_t j1 = min(n, j0 + chunk);
  uint32_t r = 0;
  for (int64_t t = j0; t < j1; t += kRankThreads) {
    __syncthreads();
    const int64_t gj = t + tid;
    if (gj < j1) {
      const uint32_t rj = cand ? cand[gj] : (uint32_t)gj;
      const uint64_t* other = ckeys + (int64_t)rj * W;
      s_row[tid] = rj;
      s_k0[tid] = other[0];
      s_k1[tid] = wact > 1 ? other[1] : 0ull;
    }
    __syncthreads();
    const int cnt = (int)min((int64_t)kRankThreads, j1 - t);
    if (!valid) continue;
#pragma unroll 4
    for (int jj = 0; jj < cnt; ++jj) {
      const uint64_t a0 = s_k0[jj];
      if (a0 < k0) { ++r; continue; }
      if (a0 != k0) continue;
      const uint64_t a1 = s_k1[jj];
      if (a1 < k1) { ++r; continue; }
      if (a1 != k1) continue;
      const uint32_t rj = s_row[jj];
      int c = 0;                                   // -1: other < mine, +1: other > mine
      const uint64_t* other = ckeys + (int64_t)rj * W;
      for (int w = 2; w < wact && c == 0; ++w) {
        const uint64_t a = other[w], b = mine[w];
        c = a < b ? -1 : (a > b ? 1 : 0);
      }
      if (c < 0 || (c == 0 && rj < ri)) ++r;
    }
  }
  if (J > 1) {
    if (valid) atomicAdd(&rank_tmp[gi], r);
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(&tile_done[blockIdx.x], 1) == J - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (valid) r = atomicAdd(&rank_tmp[gi], 0u);
  }
  if (valid && (int64_t)r < m) out_rows[r] = ri;
}
void launch_rank_sort(const uint64_t* ckeys, int W, const int* w_dev, const uint32_t* cand, const int* nc_dev, int cand_cap,
                      RowCount n, int64_t rows_bound, int64_t limit, uint32_t* rank_tmp, int* tile_done, uint32_t* out_rows,
                      int* out_count, int* overflow, cudaStream_t s) {
  const int tiles = (int)std::max<int64_t>(1, (rows_bound + kRankThreads - 1) / kRankThreads);
  const int J = rank_sort_slices(rows_bound);
  rank_sort_kernel<<<dim3(tiles, J), kRankThreads, 0, s>>>(ckeys, W, w_dev, cand, nc_dev, cand_cap, n.dev, n.host, limit, rank_tmp,
                                                           tile_done, out_rows, out_count, overflow);
  NDS_CUDA(cudaGetLastError());
}
int rank_sort_slices(int64_t rows_bound) {
  const int64_t tiles = std::max<int64_t>(1, (rows_bound + kRankThreads - 1) / kRankThreads);
  int64_t J = 296 / tiles;                      // about two CTAs per SM in flight
  if (J > tiles) J = tiles;
  if (J < 1) J = 1;
  return (int)J;
}

__global__ void take_perm_kernel(const uint32_t* perm_a, const uint32_t* perm_b, const int* which, int widx, const int* n_dev,
                                 int64_t n_host, int64_t limit, uint32_t* out_rows, int* out_count) {
  const int64_t n = row_count(n_dev, n_host);
  const int64_t m = (limit >= 0 && limit < n) ? limit : n;
  const uint32_t* perm = which[widx] ? perm_b : perm_a;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) out_rows[i] = perm[i];
  if (blockIdx.x == 0 && threadIdx.x == 0) *out_count = (int)m;
}
void launch_take_perm(const uint32_t* perm_a, const uint32_t* perm_b, const int* which, int widx, RowCount n, int64_t limit,
                      uint32_t* out_rows, int* out_count, cudaStream_t s) {
  int64_t m = (limit >= 0 && limit < n.host) ? limit : n.host;
  take_perm_kernel<<<grid_for(m), kThreads, 0, s>>>(perm_a, perm_b, which, widx, n.dev, n.host, limit, out_rows, out_count);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ result strings
// ------------------------------------------------------------------ exchange (all-gather of a batch between GPUs)
// A rank's rows travel as one segment: per column [validity, one byte per row][values]; string columns send their
// lengths as the values and their bytes in fixed slots of the column's maximum length (no prefix sum, no second size
// exchange - NVLink has the bandwidth to spare).  XCol::off_* are byte offsets inside the segment; they depend on
// the row count only, so every rank can lay out every other rank's segment once the counts are known.
// flag_out (optional) receives *flag_in != 0: this rank's "a device-side size assumption failed" state travels with the count
__global__ void set_count_kernel(const int* n_dev, int64_t n_host, int32_t* out, const int* flag_in, int32_t* flag_out) {
  out[0] = (int32_t)row_count(n_dev, n_host);
  if (flag_out) flag_out[0] = (flag_in && *flag_in) ? 1 : 0;
}
void launch_set_count(RowCount n, int32_t* out, cudaStream_t s, const int* flag_in, int32_t* flag_out) {
  set_count_kernel<<<1, 1, 0, s>>>(n.dev, n.host, out, flag_in, flag_out);
  NDS_CUDA(cudaGetLastError());
}

__global__ void __launch_bounds__(kThreads)
exchange_pack_kernel(const XCol* __restrict__ cols, int ncols, int64_t n, uint8_t* __restrict__ out, int* err) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    for (int c = 0; c < ncols; ++c) {
      const XCol x = cols[c];
      const bool ok = !x.valid || ((x.valid[i >> 5] >> (i & 31)) & 1u);
      out[x.off_valid + i] = ok ? 1 : 0;
      if (x.kind == 0) reinterpret_cast<int32_t*>(out + x.off_data)[i] = ok ? reinterpret_cast<const int32_t*>(x.data)[i] : 0;
      else if (x.kind == 1) reinterpret_cast<int64_t*>(out + x.off_data)[i] = ok ? reinterpret_cast<const int64_t*>(x.data)[i] : 0;
      else {
        const StrView v = ok ? reinterpret_cast<const StrView*>(x.data)[i] : 0;
        int len = (int)(v >> 32);
        if (len > x.maxlen) { if (err) atomicExch(err, 14); len = x.maxlen; }
        reinterpret_cast<int32_t*>(out + x.off_data)[i] = len;
        const uint8_t* p = x.strbase + (uint32_t)v;
        uint8_t* q = out + x.off_aux + i * (int64_t)x.maxlen;
        for (int b = 0; b < len; ++b) q[b] = p[b];
      }
    }
  }
}
void launch_exchange_pack(const XCol* cols, int ncols, int64_t n, uint8_t* out, int* err, cudaStream_t s) {
  if (n <= 0) return;
  exchange_pack_kernel<<<grid_for(n), kThreads, 0, s>>>(cols, ncols, n, out, err);
  NDS_CUDA(cudaGetLastError());
}

// blockIdx.y = source rank; srcs[r * ncols + c] describes column c inside rank r's segment (data = segment base),
// dst[c] the concatenated output column (data, valid bitmap - pre-zeroed -, strbase = the whole receive buffer).
__global__ void __launch_bounds__(kThreads)
exchange_unpack_kernel(const XCol* __restrict__ srcs, const XCol* __restrict__ dst, int ncols, const int64_t* __restrict__ row_begin,
                       const uint8_t* __restrict__ recv) {
  const int r = blockIdx.y;
  const int64_t n = row_begin[r + 1] - row_begin[r];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = row_begin[r] + i;
    for (int c = 0; c < ncols; ++c) {
      const XCol x = srcs[r * ncols + c];
      const XCol d = dst[c];
      const uint8_t* seg = reinterpret_cast<const uint8_t*>(x.data);
      const bool ok = seg[x.off_valid + i] != 0;
      if (ok) atomicOr(const_cast<uint32_t*>(d.valid) + (o >> 5), 1u << (o & 31));
      if (x.kind == 0) reinterpret_cast<int32_t*>(const_cast<void*>(d.data))[o] = reinterpret_cast<const int32_t*>(seg + x.off_data)[i];
      else if (x.kind == 1) reinterpret_cast<int64_t*>(const_cast<void*>(d.data))[o] = reinterpret_cast<const int64_t*>(seg + x.off_data)[i];
      else {
        const uint32_t len = (uint32_t) reinterpret_cast<const int32_t*>(seg + x.off_data)[i];
        const uint64_t off = (uint64_t)(seg + x.off_aux + i * (int64_t)x.maxlen - recv);
        reinterpret_cast<StrView*>(const_cast<void*>(d.data))[o] = ((uint64_t)len << 32) | (uint64_t)(uint32_t)off;
      }
    }
  }
}
void launch_exchange_unpack(const XCol* srcs, const XCol* dst, int ncols, const int64_t* row_begin, int world, int64_t max_rows,
                            const uint8_t* recv, cudaStream_t s) {
  if (max_rows <= 0) return;
  exchange_unpack_kernel<<<dim3(grid_for(max_rows, kThreads, 148 * 2), world), kThreads, 0, s>>>(srcs, dst, ncols, row_begin, recv);
  NDS_CUDA(cudaGetLastError());
}

// ---- hash-partition exchange (Spark: ShuffleExchange(hashpartitioning(keys))): every row goes to rank hash(keys) % world.
// exchange_dest_kernel: the destination of every row (the group-by's hash, over values - never pointers - so that all
// ranks agree) and the rows per destination; exchange_perm_kernel: rows ordered by destination; exchange_pack_perm_kernel:
// one segment per destination, laid out for that destination's row count.
__global__ void __launch_bounds__(kThreads)
exchange_dest_kernel(const Prog* __restrict__ progs, int nkeys, OpCtx c, const int* n_dev, int64_t n_host, int world,
                     uint8_t* __restrict__ dest, int32_t* __restrict__ counts) {
  const int64_t n = row_count(n_dev, n_host);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t h = 0x9E3779B97F4A7C15ULL;
    for (int k = 0; k < nkeys; ++k) {
      const Val v = eval_program(progs[k].code, progs[k].n, c.cols, c.lit, i);
      const uint64_t hv = !v.ok ? 0x7F4A7C15ULL : (progs[k].kind == 2 ? hash_bytes((const uint8_t*)(uintptr_t)v.v, v.len) : (uint64_t)v.v);
      h = mix64(h ^ hv) + 0x9E3779B97F4A7C15ULL * (k + 1);
    }
    const int d = (int)(mix64(h) % (uint64_t)world);
    dest[i] = (uint8_t)d;
    atomicAdd(&counts[d], 1);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) counts[world] = (int32_t)n;
}
void launch_exchange_dest(const Prog* progs, int nkeys, OpCtx c, RowCount n, int world, uint8_t* dest, int32_t* counts, cudaStream_t s) {
  exchange_dest_kernel<<<grid_for(n.host), kThreads, 0, s>>>(progs, nkeys, c, n.dev, n.host, world, dest, counts);
  NDS_CUDA(cudaGetLastError());
}
__global__ void __launch_bounds__(kThreads)
exchange_perm_kernel(const uint8_t* __restrict__ dest, int64_t n, const int64_t* __restrict__ begin, int32_t* cursor, uint32_t* __restrict__ perm) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int d = dest[i];
    perm[begin[d] + atomicAdd(&cursor[d], 1)] = (uint32_t)i;
  }
}
void launch_exchange_perm(const uint8_t* dest, int64_t n, const int64_t* begin, int32_t* cursor, uint32_t* perm, cudaStream_t s) {
  if (n <= 0) return;
  exchange_perm_kernel<<<grid_for(n), kThreads, 0, s>>>(dest, n, begin, cursor, perm);
  NDS_CUDA(cudaGetLastError());
}
// segs[d * ncols + c]: column c inside destination d's segment (data = segment base, offsets for that segment's rows)
__global__ void __launch_bounds__(kThreads)
exchange_pack_perm_kernel(const XCol* __restrict__ cols, const XCol* __restrict__ segs, int ncols, const uint32_t* __restrict__ perm,
                          const int64_t* __restrict__ begin, int world, int64_t n, int* err) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    int d = 0;
    while (d + 1 < world && begin[d + 1] <= p) ++d;
    const int64_t j = p - begin[d];
    const int64_t i = perm[p];
    for (int c = 0; c < ncols; ++c) {
      const XCol x = cols[c];
      const XCol sg = segs[d * ncols + c];
      uint8_t* out = const_cast<uint8_t*>(reinterpret_cast<const uint8_t*>(sg.data));
      const bool ok = !x.valid || ((x.valid[i >> 5] >> (i & 31)) & 1u);
      out[sg.off_valid + j] = ok ? 1 : 0;
      if (x.kind == 0) reinterpret_cast<int32_t*>(out + sg.off_data)[j] = ok ? reinterpret_cast<const int32_t*>(x.data)[i] : 0;
      else if (x.kind == 1) reinterpret_cast<int64_t*>(out + sg.off_data)[j] = ok ? reinterpret_cast<const int64_t*>(x.data)[i] : 0;
      else {
        const StrView v = ok ? reinterpret_cast<const StrView*>(x.data)[i] : 0;
        int len = (int)(v >> 32);
        if (len > sg.maxlen) { if (err) atomicExch(err, 14); len = sg.maxlen; }
        reinterpret_cast<int32_t*>(out + sg.off_data)[j] = len;
        const uint8_t* q = x.strbase + (uint32_t)v;
        uint8_t* w = out + sg.off_aux + j * (int64_t)sg.maxlen;
        for (int b = 0; b < len; ++b) w[b] = q[b];
      }
    }
  }
}
void launch_exchange_pack_perm(const XCol* cols, const XCol* segs, int ncols, const uint32_t* perm, const int64_t* begin, int world,
                               int64_t n, int* err, cudaStream_t s) {
  if (n <= 0) return;
  exchange_pack_perm_kernel<<<grid_for(n), kThreads, 0, s>>>(cols, segs, ncols, perm, begin, world, n, err);
  NDS_CUDA(cudaGetLastError());
}

// ---- the same exchange with the row counts left on the device: ONE collective, no host round trip.
// A segment has a fixed capacity (cap_rows rows, laid out for that capacity) and starts with a 64-byte header:
// int32 row count, then the string slot width of every column.  The unpack kernel reads every rank's header itself.
__global__ void __launch_bounds__(kThreads)
exchange_pack_dev_kernel(const XCol* __restrict__ cols, int ncols, const int* n_dev, int64_t n_host, int64_t cap_rows,
                         uint8_t* __restrict__ out, int* err) {
  int64_t n = row_count(n_dev, n_host);
  const bool over = n > cap_rows;   // more rows than the speculative capacity: the query runs again with exact sizes
  if (over) n = cap_rows;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    int32_t* h = reinterpret_cast<int32_t*>(out);
    h[0] = (int32_t)n;
    for (int c = 0; c < ncols && c < kXchgHeader / 4 - 2; ++c) h[1 + c] = cols[c].kind == 2 ? cols[c].maxlen : 0;
    h[kXchgHeader / 4 - 1] = (err[2] || over) ? 1 : 0;      // a failed size assumption on ANY rank must make EVERY rank run the query again
  }
  uint8_t* body = out + kXchgHeader;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    for (int c = 0; c < ncols; ++c) {
      const XCol x = cols[c];
      const bool ok = !x.valid || ((x.valid[i >> 5] >> (i & 31)) & 1u);
      body[x.off_valid + i] = ok ? 1 : 0;
      if (x.kind == 0) reinterpret_cast<int32_t*>(body + x.off_data)[i] = ok ? reinterpret_cast<const int32_t*>(x.data)[i] : 0;
      else if (x.kind == 1) reinterpret_cast<int64_t*>(body + x.off_data)[i] = ok ? reinterpret_cast<const int64_t*>(x.data)[i] : 0;
      else {
        const StrView v = ok ? reinterpret_cast<const StrView*>(x.data)[i] : 0;
        int len = (int)(v >> 32);
        if (len > x.maxlen) { if (err) atomicExch(err, 14); len = x.maxlen; }
        reinterpret_cast<int32_t*>(body + x.off_data)[i] = len;
        const uint8_t* p = x.strbase + (uint32_t)v;
        uint8_t* q = body + x.off_aux + i * (int64_t)x.maxlen;
        for (int b = 0; b < len; ++b) q[b] = p[b];
      }
    }
  }
}
void launch_exchange_pack_dev(const XCol* cols, int ncols, RowCount n, int64_t cap_rows, uint8_t* out, int* err, cudaStream_t s) {
  exchange_pack_dev_kernel<<<grid_for(std::min<int64_t>(n.host, cap_rows)), kThreads, 0, s>>>(cols, ncols, n.dev, n.host, cap_rows, out, err);
  NDS_CUDA(cudaGetLastError());
}

// blockIdx.y = source rank.  proto[c]: column c's layout inside a segment (offsets for cap_rows; the same for every
// rank), dst[c]: the concatenated output column (valid bitmap pre-zeroed).  total_out[0] = rows of all ranks.
__global__ void __launch_bounds__(kThreads)
exchange_unpack_dev_kernel(const XCol* __restrict__ proto, const XCol* __restrict__ dst, int ncols, int world, int64_t seg_bytes,
                           const uint8_t* __restrict__ recv, int* total_out, int* err) {
  const int r = blockIdx.y;
  int64_t before = 0, total = 0;
  for (int q = 0; q < world; ++q) {
    const int32_t* h = reinterpret_cast<const int32_t*>(recv + (int64_t)q * seg_bytes);
    const int64_t nq = h[0];
    if (q < r) before += nq;
    total += nq;
    // every rank must have laid its strings out with the same slot widths (they come from table metadata)
    if (blockIdx.x == 0 && threadIdx.x == 0 && r == 0)
      for (int c = 0; c < ncols && c < kXchgHeader / 4 - 2; ++c)
        if (proto[c].kind == 2 && h[1 + c] != proto[c].maxlen) atomicExch(err, 15);
    if (blockIdx.x == 0 && threadIdx.x == 0 && r == 0 && h[kXchgHeader / 4 - 1]) atomicExch(err + 2, 1);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && r == 0) *total_out = (int)total;
  const uint8_t* seg = recv + (int64_t)r * seg_bytes;
  const int64_t n = reinterpret_cast<const int32_t*>(seg)[0];
  const uint8_t* body = seg + kXchgHeader;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = before + i;
    for (int c = 0; c < ncols; ++c) {
      const XCol x = proto[c];
      const XCol d = dst[c];
      const bool ok = body[x.off_valid + i] != 0;
      if (ok) atomicOr(const_cast<uint32_t*>(d.valid) + (o >> 5), 1u << (o & 31));
      if (x.kind == 0) reinterpret_cast<int32_t*>(const_cast<void*>(d.data))[o] = reinterpret_cast<const int32_t*>(body + x.off_data)[i];
      else if (x.kind == 1) reinterpret_cast<int64_t*>(const_cast<void*>(d.data))[o] = reinterpret_cast<const int64_t*>(body + x.off_data)[i];
      else {
        const uint32_t len = (uint32_t) reinterpret_cast<const int32_t*>(body + x.off_data)[i];
        const uint64_t off = (uint64_t)(body + x.off_aux + i * (int64_t)x.maxlen - recv);
        reinterpret_cast<StrView*>(const_cast<void*>(d.data))[o] = ((uint64_t)len << 32) | (uint64_t)(uint32_t)off;
      }
    }
  }
}
void launch_exchange_unpack_dev(const XCol* proto, const XCol* dst, int ncols, int world, int64_t cap_rows, int64_t seg_bytes,
                                const uint8_t* recv, int* total_out, int* err, cudaStream_t s) {
  exchange_unpack_dev_kernel<<<dim3(grid_for(std::max<int64_t>(cap_rows, 1), kThreads, 148 * 2), world), kThreads, 0, s>>>(
      proto, dst, ncols, world, seg_bytes, recv, total_out, err);
  NDS_CUDA(cudaGetLastError());
}

// ---- peer-push exchange: pack into every peer's window over NVLink, arrive, wait, unpack (ops.cuh, PeerWindow)
__device__ __forceinline__ uint8_t* peer_slot(const PeerWindow& w, int rank, int half, int slot) {
  return w.win[rank] + w.ctrl_bytes + (int64_t)half * w.half_bytes + (int64_t)slot * w.slot_bytes;
}
__device__ __forceinline__ uint32_t* peer_arrivals(const PeerWindow& w, int rank, int half) {
  return reinterpret_cast<uint32_t*>(w.win[rank] + half * 64);
}

__global__ void __launch_bounds__(kThreads)
peer_push_kernel(const XCol* __restrict__ cols, const XCol* __restrict__ proto, int ncols, const int* n_dev, int64_t n_host, int64_t cap_rows,
                 const PeerWindow w, int half, int loopback, int* ticket, int* err) {
  __shared__ int s_last;
  int64_t n = row_count(n_dev, n_host);
  const bool over = n > cap_rows;
  if (over) n = 0;                                   // nothing is sent: the header says why
  // loopback (one GPU, tests): this rank plays every sender - it fills all `world` slots of its own window
  const int ntargets = w.world;
  for (int t = 0; t < ntargets; ++t) {
    uint8_t* slot = loopback ? peer_slot(w, w.me, half, t) : peer_slot(w, t, half, w.me);
    uint8_t* body = slot + kXchgHeader;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
      for (int c = 0; c < ncols; ++c) {
        const XCol x = cols[c];
        const XCol pr = proto[c];
        const bool ok = !x.valid || ((x.valid[i >> 5] >> (i & 31)) & 1u);
        body[pr.off_valid + i] = ok ? 1 : 0;
        if (x.kind == 0) reinterpret_cast<int32_t*>(body + pr.off_data)[i] = ok ? reinterpret_cast<const int32_t*>(x.data)[i] : 0;
        else if (x.kind == 1) reinterpret_cast<int64_t*>(body + pr.off_data)[i] = ok ? reinterpret_cast<const int64_t*>(x.data)[i] : 0;
        else {
          const StrView v = ok ? reinterpret_cast<const StrView*>(x.data)[i] : 0;
          int len = (int)(v >> 32);
          if (len > pr.maxlen) { if (err) atomicExch(err, 14); len = pr.maxlen; }
          reinterpret_cast<int32_t*>(body + pr.off_data)[i] = len;
          const uint8_t* q = x.strbase + (uint32_t)v;
          uint8_t* o = body + pr.off_aux + i * (int64_t)pr.maxlen;
          for (int b = 0; b < len; ++b) o[b] = q[b];
        }
      }
    }
  }
  // every CTA's rows are out (system-wide) before the last CTA raises the arrival counters
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(ticket, 1) == (int)gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence_system();
  for (int t = threadIdx.x; t < ntargets; t += blockDim.x) {
    uint8_t* slot = loopback ? peer_slot(w, w.me, half, t) : peer_slot(w, t, half, w.me);
    int32_t* h = reinterpret_cast<int32_t*>(slot);
    for (int c = 0; c < ncols && c < kXchgHeader / 4 - 2; ++c) h[1 + c] = proto[c].kind == 2 ? proto[c].maxlen : 0;
    h[kXchgHeader / 4 - 1] = (err[2] || over) ? 1 : 0;
    h[0] = over ? -1 : (int32_t)n;
    __threadfence_system();
    atomicAdd_system(peer_arrivals(w, loopback ? w.me : t, half), 1u);
  }
}
void launch_peer_push(const XCol* cols, const XCol* proto, int ncols, RowCount n, int64_t cap_rows, PeerWindow w, int half, int loopback,
                      int* ticket, int* err, cudaStream_t s) {
  peer_push_kernel<<<grid_for(std::min<int64_t>(n.host, cap_rows), kThreads, 148 * 2), kThreads, 0, s>>>(cols, proto, ncols, n.dev, n.host, cap_rows, w,
                                                                                                       half, loopback, ticket, err);
  NDS_CUDA(cudaGetLastError());
}

__global__ void __launch_bounds__(kThreads)
peer_unpack_kernel(const XCol* __restrict__ proto, const XCol* __restrict__ dst, int ncols, const PeerWindow w, int half, uint32_t arrivals_wanted,
                   int* total_out, int* err) {
  __shared__ int s_ok;
  // wait (bounded: ~4 s of SM clocks) until every rank's push into this half has arrived
  if (threadIdx.x == 0) {
    volatile uint32_t* arr = peer_arrivals(w, w.me, half);
    const long long t0 = clock64();
    int ok = 1;
    while ((int32_t)(*arr - arrivals_wanted) < 0) {
      if (clock64() - t0 > 8000000000LL) { ok = 0; break; }
      __nanosleep(100);
    }
    __threadfence_system();
    s_ok = ok;
    if (!ok) atomicExch(err, 16);
  }
  __syncthreads();
  if (!s_ok) return;
  const int r = blockIdx.y;
  int64_t before = 0, total = 0;
  bool failed = false;
  for (int q = 0; q < w.world; ++q) {
    const volatile int32_t* h = reinterpret_cast<const volatile int32_t*>(peer_slot(w, w.me, half, q));
    const int64_t nq = h[0];
    if (nq < 0 || h[kXchgHeader / 4 - 1]) failed = true;
    if (q < r && nq > 0) before += nq;
    if (nq > 0) total += nq;
    if (blockIdx.x == 0 && threadIdx.x == 0 && r == 0)
      for (int c = 0; c < ncols && c < kXchgHeader / 4 - 2; ++c)
        if (proto[c].kind == 2 && h[1 + c] != proto[c].maxlen) atomicExch(err, 15);
  }
  if (failed) {      // some rank could not send (too many rows) or had a failed size assumption: every rank runs the query again
    if (blockIdx.x == 0 && threadIdx.x == 0 && r == 0) { atomicExch(err + 2, 1); *total_out = 0; }
    return;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && r == 0) *total_out = (int)total;
  const uint8_t* slot = peer_slot(w, w.me, half, r);
  const int64_t n = reinterpret_cast<const volatile int32_t*>(slot)[0];
  const uint8_t* body = slot + kXchgHeader;
  const uint8_t* base = w.win[w.me];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = before + i;
    for (int c = 0; c < ncols; ++c) {
      const XCol x = proto[c];
      const XCol d = dst[c];
      const bool ok = __ldcg(body + x.off_valid + i) != 0;
      if (ok) atomicOr(const_cast<uint32_t*>(d.valid) + (o >> 5), 1u << (o & 31));
      if (x.kind == 0) reinterpret_cast<int32_t*>(const_cast<void*>(d.data))[o] = __ldcg(reinterpret_cast<const int32_t*>(body + x.off_data) + i);
      else if (x.kind == 1) reinterpret_cast<int64_t*>(const_cast<void*>(d.data))[o] = __ldcg(reinterpret_cast<const long long*>(body + x.off_data) + i);
      else {
        const uint32_t len = (uint32_t)__ldcg(reinterpret_cast<const int32_t*>(body + x.off_data) + i);
        const uint64_t off = (uint64_t)(body + x.off_aux + i * (int64_t)x.maxlen - base);
        reinterpret_cast<StrView*>(const_cast<void*>(d.data))[o] = ((uint64_t)len << 32) | (uint64_t)(uint32_t)off;
      }
    }
  }
}
void launch_peer_unpack(const XCol* proto, const XCol* dst, int ncols, PeerWindow w, int half, uint32_t arrivals_wanted, int64_t cap_rows,
                        int* total_out, int* err, cudaStream_t s) {
  peer_unpack_kernel<<<dim3(grid_for(std::max<int64_t>(cap_rows, 1), kThreads, 16), w.world), kThreads, 0, s>>>(proto, dst, ncols, w, half,
                                                                                                               arrivals_wanted, total_out, err);
  NDS_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ result export, small results
// One launch packs a whole (small) result into one contiguous buffer, so that it reaches the host with ONE copy and
// ONE synchronisation, row count included: block b handles segment b.
__global__ void __launch_bounds__(256)
export_pack_kernel(const ExportSeg* __restrict__ segs, const int* n_dev, int64_t n_host, uint8_t* __restrict__ out, int* err) {
  const ExportSeg sg = segs[blockIdx.x];
  const int tid = threadIdx.x;
  const int64_t n = row_count(n_dev, n_host);
  if (sg.kind == 2) {
    if (tid == 0) *reinterpret_cast<int*>(out + sg.dst_off) = (int)n;
    return;
  }
  if (sg.kind == 0) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(sg.src);
    uint32_t* dst = reinterpret_cast<uint32_t*>(out + sg.dst_off);
    for (int64_t i = tid; i < sg.bytes / 4; i += 256) dst[i] = src[i];
    return;
  }
  // strings: Arrow offsets (null rows get length 0) by a block scan, then the bytes
  __shared__ int wsum[8];
  __shared__ int s_base;
  const StrView* views = reinterpret_cast<const StrView*>(sg.src);
  int32_t* offs = reinterpret_cast<int32_t*>(out + sg.dst_off);
  uint8_t* chars = out + sg.aux_off;
  const int warp = tid >> 5, lane = tid & 31;
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int64_t i0 = 0; i0 < n; i0 += 256) {
    const int64_t i = i0 + tid;
    int len = 0;
    StrView v = 0;
    if (i < n && (!sg.valid || ((sg.valid[i >> 5] >> (i & 31)) & 1u))) { v = views[i]; len = (int)(v >> 32); }
    int incl = len;
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int pre = s_base;
    for (int w = 0; w < warp; ++w) pre += wsum[w];
    const int start = pre + incl - len;
    if (i < n) {
      offs[i] = start;
      if ((int64_t)start + len > sg.cap) { if (len) atomicExch(err, 13); }
      else {
        const uint8_t* p = sg.strbase + (uint32_t)v;
        for (int b = 0; b < len; ++b) chars[start + b] = p[b];
      }
    }
    __syncthreads();
    if (tid == 255) s_base = pre + incl;
    __syncthreads();
  }
  if (tid == 0) offs[n] = s_base;
}
void launch_export_pack(const ExportSeg* segs, int nsegs, RowCount n, uint8_t* out, int* err, cudaStream_t s) {
  export_pack_kernel<<<nsegs, 256, 0, s>>>(segs, n.dev, n.host, out, err);
  NDS_CUDA(cudaGetLastError());
}

__global__ void copy_str_bytes_kernel(const StrView* __restrict__ views, const uint8_t* __restrict__ base,
                                      const int64_t* __restrict__ offsets, int64_t n, uint8_t* __restrict__ out) {
  // one warp per string
  const int lane = threadIdx.x & 31;
  for (int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5; i < n; i += ((int64_t)gridDim.x * blockDim.x) >> 5) {
    StrView v = views[i];
    const uint8_t* p = base + (uint32_t)v;
    uint32_t len = (uint32_t)(v >> 32);
    if ((int64_t)len != offsets[i + 1] - offsets[i]) continue;   // null rows have length 0 in offsets
    for (uint32_t b = lane; b < len; b += 32) out[offsets[i] + b] = p[b];
  }
}
void launch_copy_str_bytes(const StrView* views, const uint8_t* base, const int64_t* offsets, int64_t n, uint8_t* out,
                           cudaStream_t s) {
  if (n <= 0) return;
  copy_str_bytes_kernel<<<grid_for(n * 32), kThreads, 0, s>>>(views, base, offsets, n, out);
  NDS_CUDA(cudaGetLastError());
}

}  // namespace nds
