// Fused filter/project expression evaluation: a per-row postfix interpreter with SQL
// three-valued logic over Arrow validity bitmaps.  One kernel launch evaluates a whole
// predicate / projection tree (Spark: FilterExec / ProjectExec under whole-stage codegen;
// the reference only ever submits SQL text - /root/reference/nds/nds_power.py:125-135).
#pragma once
#include "device_types.h"

namespace nds {

enum Op : int32_t {
  OP_COL = 0,   // a = column slot
  OP_LIT_I,     // imm
  OP_LIT_F,     // imm = bits of a double
  OP_LIT_S,     // imm = offset | len<<32 into the literal pool
  OP_LIT_NULL,
  OP_EQ_I, OP_NE_I, OP_LT_I, OP_LE_I, OP_GT_I, OP_GE_I,
  OP_EQ_F, OP_NE_F, OP_LT_F, OP_LE_F, OP_GT_F, OP_GE_F,
  OP_EQ_S, OP_NE_S, OP_LT_S, OP_LE_S, OP_GT_S, OP_GE_S,
  OP_AND, OP_OR, OP_NOT, OP_ISNULL, OP_ISNOTNULL,
  OP_ADD_I, OP_SUB_I, OP_MUL_I, OP_NEG_I,
  OP_ADD_F, OP_SUB_F, OP_MUL_F, OP_DIV_F, OP_NEG_F,
  OP_I2F,
  OP_SUBSTR,    // a = 1-based start, imm = length (UTF-8 code points)
  OP_IF,        // stack: cond, then, else
  OP_DECAVG,    // stack: sum(unscaled), count -> round_half_up(sum * 10^a / count)
  OP_DIVI2F,    // stack: sum(int), count -> (double)sum / (double)count, null when count = 0
  OP_COUNT_
};

struct Instr {
  int32_t op;
  int32_t a;
  int64_t imm;
};

struct Val {
  int64_t v;      // integer, double bits, or char pointer
  uint32_t len;   // string length
  bool ok;        // not NULL
};

constexpr int kEvalStack = 16;
constexpr uint32_t kNoRow = 0xFFFFFFFFu;      // row-map entry of an outer join's unmatched row: every column it maps reads NULL

__device__ __forceinline__ int str_cmp(const uint8_t* a, uint32_t la, const uint8_t* b, uint32_t lb) {
  uint32_t m = la < lb ? la : lb;
  for (uint32_t i = 0; i < m; ++i) {
    int d = (int)a[i] - (int)b[i];
    if (d) return d;
  }
  return la < lb ? -1 : (la > lb ? 1 : 0);
}

__device__ __forceinline__ Val load_col(const ColRef& c, int64_t row) {
  Val r;
  int64_t b = c.rowmap ? (int64_t)c.rowmap[row] : row;
  r.len = 0;
  r.v = 0;
  if (b == (int64_t)kNoRow) { r.ok = false; return r; }       // the unmatched side of an outer join
  r.ok = !c.valid || ((c.valid[b >> 5] >> (b & 31)) & 1u);
  if (c.dtype == DT_I32) r.v = reinterpret_cast<const int32_t*>(c.data)[b];
  else if (c.dtype == DT_STR) {
    StrView s = reinterpret_cast<const StrView*>(c.data)[b];
    r.v = (int64_t)(uintptr_t)(c.strbase + (uint32_t)s);
    r.len = (uint32_t)(s >> 32);
  } else r.v = reinterpret_cast<const int64_t*>(c.data)[b];
  return r;
}

__device__ inline Val eval_program(const Instr* __restrict__ code, int ncode, const ColRef* __restrict__ cols,
                                   const uint8_t* __restrict__ lit, int64_t row) {
  // a bare column reference (group keys, aggregate arguments, sort keys, gathers: most programs): no interpreter loop
  if (ncode == 1 && code[0].op == OP_COL) return load_col(cols[code[0].a], row);
  Val st[kEvalStack];
  int sp = 0;
  for (int pc = 0; pc < ncode; ++pc) {
    const Instr in = code[pc];
    switch (in.op) {
      case OP_COL: st[sp++] = load_col(cols[in.a], row); break;
      case OP_LIT_I:
      case OP_LIT_F: st[sp].v = in.imm; st[sp].len = 0; st[sp].ok = true; ++sp; break;
      case OP_LIT_S:
        st[sp].v = (int64_t)(uintptr_t)(lit + (uint32_t)in.imm);
        st[sp].len = (uint32_t)((uint64_t)in.imm >> 32);
        st[sp].ok = true;
        ++sp;
        break;
      case OP_LIT_NULL: st[sp].v = 0; st[sp].len = 0; st[sp].ok = false; ++sp; break;
      case OP_EQ_I: case OP_NE_I: case OP_LT_I: case OP_LE_I: case OP_GT_I: case OP_GE_I: {
        Val b = st[--sp]; Val& a = st[sp - 1];
        bool r;
        switch (in.op) {
          case OP_EQ_I: r = a.v == b.v; break;
          case OP_NE_I: r = a.v != b.v; break;
          case OP_LT_I: r = a.v < b.v; break;
          case OP_LE_I: r = a.v <= b.v; break;
          case OP_GT_I: r = a.v > b.v; break;
          default: r = a.v >= b.v;
        }
        a.ok = a.ok && b.ok; a.v = r; break;
      }
      case OP_EQ_F: case OP_NE_F: case OP_LT_F: case OP_LE_F: case OP_GT_F: case OP_GE_F: {
        Val b = st[--sp]; Val& a = st[sp - 1];
        double x = __longlong_as_double(a.v), y = __longlong_as_double(b.v);
        bool r;
        switch (in.op) {
          case OP_EQ_F: r = x == y; break;
          case OP_NE_F: r = x != y; break;
          case OP_LT_F: r = x < y; break;
          case OP_LE_F: r = x <= y; break;
          case OP_GT_F: r = x > y; break;
          default: r = x >= y;
        }
        a.ok = a.ok && b.ok; a.v = r; break;
      }
      case OP_EQ_S: case OP_NE_S: case OP_LT_S: case OP_LE_S: case OP_GT_S: case OP_GE_S: {
        Val b = st[--sp]; Val& a = st[sp - 1];
        bool ok = a.ok && b.ok;
        int c = ok ? str_cmp((const uint8_t*)(uintptr_t)a.v, a.len, (const uint8_t*)(uintptr_t)b.v, b.len) : 0;
        bool r;
        switch (in.op) {
          case OP_EQ_S: r = c == 0; break;
          case OP_NE_S: r = c != 0; break;
          case OP_LT_S: r = c < 0; break;
          case OP_LE_S: r = c <= 0; break;
          case OP_GT_S: r = c > 0; break;
          default: r = c >= 0;
        }
        a.ok = ok; a.v = r; a.len = 0; break;
      }
      case OP_AND: {  // Kleene: false AND x = false
        Val b = st[--sp]; Val& a = st[sp - 1];
        bool af = a.ok && !a.v, bf = b.ok && !b.v;
        if (af || bf) { a.ok = true; a.v = 0; }
        else { a.ok = a.ok && b.ok; a.v = 1; }
        break;
      }
      case OP_OR: {   // true OR x = true
        Val b = st[--sp]; Val& a = st[sp - 1];
        bool at = a.ok && a.v, bt = b.ok && b.v;
        if (at || bt) { a.ok = true; a.v = 1; }
        else { a.ok = a.ok && b.ok; a.v = 0; }
        break;
      }
      case OP_NOT: st[sp - 1].v = !st[sp - 1].v; break;
      case OP_ISNULL: st[sp - 1].v = !st[sp - 1].ok; st[sp - 1].ok = true; st[sp - 1].len = 0; break;
      case OP_ISNOTNULL: st[sp - 1].v = st[sp - 1].ok; st[sp - 1].ok = true; st[sp - 1].len = 0; break;
      case OP_ADD_I: { Val b = st[--sp]; Val& a = st[sp - 1]; a.v += b.v; a.ok = a.ok && b.ok; break; }
      case OP_SUB_I: { Val b = st[--sp]; Val& a = st[sp - 1]; a.v -= b.v; a.ok = a.ok && b.ok; break; }
      case OP_MUL_I: { Val b = st[--sp]; Val& a = st[sp - 1]; a.v *= b.v; a.ok = a.ok && b.ok; break; }
      case OP_NEG_I: st[sp - 1].v = -st[sp - 1].v; break;
      case OP_ADD_F: case OP_SUB_F: case OP_MUL_F: case OP_DIV_F: {
        Val b = st[--sp]; Val& a = st[sp - 1];
        double x = __longlong_as_double(a.v), y = __longlong_as_double(b.v), r;
        a.ok = a.ok && b.ok;
        if (in.op == OP_ADD_F) r = x + y;
        else if (in.op == OP_SUB_F) r = x - y;
        else if (in.op == OP_MUL_F) r = x * y;
        else { if (y == 0.0) a.ok = false; r = y == 0.0 ? 0.0 : x / y; }   // Spark: x / 0 = NULL
        a.v = __double_as_longlong(r);
        break;
      }
      case OP_NEG_F: st[sp - 1].v = __double_as_longlong(-__longlong_as_double(st[sp - 1].v)); break;
      case OP_I2F: st[sp - 1].v = __double_as_longlong((double)st[sp - 1].v); break;
      case OP_SUBSTR: {
        Val& a = st[sp - 1];
        if (a.ok) {
          const uint8_t* p = (const uint8_t*)(uintptr_t)a.v;
          uint32_t pos = 0, cp = 0;
          const uint32_t start = in.a > 0 ? (uint32_t)(in.a - 1) : 0u;
          while (pos < a.len && cp < start) { ++pos; while (pos < a.len && (p[pos] & 0xC0) == 0x80) ++pos; ++cp; }
          uint32_t b0 = pos;
          cp = 0;
          while (pos < a.len && cp < (uint32_t)in.imm) { ++pos; while (pos < a.len && (p[pos] & 0xC0) == 0x80) ++pos; ++cp; }
          a.v = (int64_t)(uintptr_t)(p + b0);
          a.len = pos - b0;
        }
        break;
      }
      case OP_IF: {
        Val e = st[--sp]; Val t = st[--sp]; Val& c = st[sp - 1];
        c = (c.ok && c.v) ? t : e;
        break;
      }
      case OP_DECAVG: {
        Val cnt = st[--sp]; Val& s = st[sp - 1];
        if (!s.ok || !cnt.ok || cnt.v == 0) { s.ok = false; s.v = 0; break; }
        __int128 num = (__int128)s.v;
        for (int k = 0; k < in.a; ++k) num *= 10;
        __int128 den = (__int128)cnt.v;
        bool neg = num < 0;
        if (neg) num = -num;
        __int128 q = (num * 2 + den) / (den * 2);   // HALF_UP on the magnitude
        s.v = (int64_t)(neg ? -q : q);
        break;
      }
      case OP_DIVI2F: {
        Val cnt = st[--sp]; Val& s = st[sp - 1];
        if (!s.ok || !cnt.ok || cnt.v == 0) { s.ok = false; s.v = 0; break; }
        s.v = __double_as_longlong((double)s.v / (double)cnt.v);
        break;
      }
      default: break;
    }
  }
  return st[0];
}

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

__device__ __forceinline__ uint64_t hash_bytes(const uint8_t* p, uint32_t n) {
  uint64_t h = 0xcbf29ce484222325ULL;
  for (uint32_t i = 0; i < n; ++i) { h ^= p[i]; h *= 0x100000001b3ULL; }
  return mix64(h);
}

}  // namespace nds
