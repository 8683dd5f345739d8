#include "parquet_meta.h"

namespace nds {
namespace pq {
namespace {

// Thrift compact protocol cursor.
struct TC {
  const uint8_t* b;
  int64_t p, end;
  uint8_t byte() {
    if (p >= end) fail("parquet: truncated thrift data");
    return b[p++];
  }
  uint64_t varint() {
    uint64_t r = 0;
    int s = 0;
    for (;;) {
      uint8_t c = byte();
      r |= (uint64_t)(c & 0x7F) << s;
      if (!(c & 0x80)) return r;
      s += 7;
      if (s > 63) fail("parquet: varint too long");
    }
  }
  int64_t zigzag() {
    uint64_t v = varint();
    return (int64_t)(v >> 1) ^ -(int64_t)(v & 1);
  }
  std::string binary() {
    uint64_t n = varint();
    if (p + (int64_t)n > end) fail("parquet: truncated thrift binary");
    std::string s((const char*)b + p, n);
    p += n;
    return s;
  }
  void skip(int t) {
    switch (t) {
      case 1: case 2: return;
      case 3: byte(); return;
      case 4: case 5: case 6: varint(); return;
      case 7: p += 8; return;
      case 8: { uint64_t n = varint(); p += n; return; }
      case 9: case 10: {
        uint8_t h = byte();
        uint64_t n = h >> 4;
        int et = h & 0xF;
        if (n == 15) n = varint();
        for (uint64_t i = 0; i < n; ++i) {
          if (et == 1 || et == 2) byte(); else skip(et);
        }
        return;
      }
      case 11: {
        uint64_t n = varint();
        if (n) { uint8_t kv = byte(); for (uint64_t i = 0; i < n; ++i) { skip(kv >> 4); skip(kv & 0xF); } }
        return;
      }
      case 12: skip_struct(); return;
      default: fail("parquet: unknown thrift type " + std::to_string(t));
    }
  }
  void skip_struct() {
    int16_t fid = 0;
    int t;
    while (next_field(fid, t)) skip(t);
  }
  // returns false at STOP
  bool next_field(int16_t& fid, int& t) {
    uint8_t h = byte();
    if (h == 0) return false;
    int d = h >> 4;
    t = h & 0xF;
    fid = d ? (int16_t)(fid + d) : (int16_t)zigzag();
    return true;
  }
  // list header -> (count, elem type)
  uint64_t list_begin(int& et) {
    uint8_t h = byte();
    uint64_t n = h >> 4;
    et = h & 0xF;
    if (n == 15) n = varint();
    return n;
  }
};

void parse_logical_type(TC& c, SchemaElement& e) {
  int16_t fid = 0; int t;
  while (c.next_field(fid, t)) {
    if (fid == 1) e.logical_string = true;       // STRING
    else if (fid == 5) e.logical_decimal = true; // DECIMAL (scale/precision also in element)
    else if (fid == 6) e.logical_date = true;    // DATE
    c.skip(t);
  }
}

SchemaElement parse_schema_element(TC& c) {
  SchemaElement e;
  int16_t fid = 0; int t;
  while (c.next_field(fid, t)) {
    switch (fid) {
      case 1: e.type = (int)c.zigzag(); break;
      case 3: e.repetition = (int)c.zigzag(); break;
      case 4: e.name = c.binary(); break;
      case 5: e.num_children = (int)c.zigzag(); break;
      case 6: e.converted = (int)c.zigzag(); break;
      case 7: e.scale = (int)c.zigzag(); break;
      case 8: e.precision = (int)c.zigzag(); break;
      case 10: parse_logical_type(c, e); break;
      default: c.skip(t);
    }
  }
  return e;
}

// Statistics struct of a column chunk: keeps the raw min / max bytes (new-style fields win over the deprecated ones)
void parse_statistics(TC& c, std::string& mn, std::string& mx) {
  std::string old_min, old_max, new_min, new_max;
  bool have_new_min = false, have_new_max = false, have_old_min = false, have_old_max = false;
  int16_t fid = 0; int t;
  while (c.next_field(fid, t)) {
    if (t == 8 && fid == 1) { old_max = c.binary(); have_old_max = true; }
    else if (t == 8 && fid == 2) { old_min = c.binary(); have_old_min = true; }
    else if (t == 8 && fid == 5) { new_max = c.binary(); have_new_max = true; }
    else if (t == 8 && fid == 6) { new_min = c.binary(); have_new_min = true; }
    else c.skip(t);
  }
  if (have_new_min && have_new_max) { mn = new_min; mx = new_max; }
  else if (have_old_min && have_old_max) { mn = old_min; mx = old_max; }
}

ColumnChunk parse_column_meta(TC& c) {
  ColumnChunk m;
  std::string smin, smax;
  int16_t fid = 0; int t;
  while (c.next_field(fid, t)) {
    switch (fid) {
      case 1: m.type = (int)c.zigzag(); break;
      case 3: {
        int et; uint64_t n = c.list_begin(et);
        for (uint64_t i = 0; i < n; ++i) { std::string s = c.binary(); m.path += (i ? "." : "") + s; }
        break;
      }
      case 4: m.codec = (int)c.zigzag(); break;
      case 5: m.num_values = c.zigzag(); break;
      case 6: m.total_uncompressed = c.zigzag(); break;
      case 7: m.total_compressed = c.zigzag(); break;
      case 9: m.data_page_offset = c.zigzag(); break;
      case 11: m.dictionary_page_offset = c.zigzag(); break;
      case 12: if (t == 12) parse_statistics(c, smin, smax); else c.skip(t); break;
      default: c.skip(t);
    }
  }
  const size_t w = m.type == INT32 ? 4 : (m.type == INT64 ? 8 : 0);
  if (w && smin.size() == w && smax.size() == w) {
    if (w == 4) { int32_t a, b; memcpy(&a, smin.data(), 4); memcpy(&b, smax.data(), 4); m.min_i = a; m.max_i = b; }
    else { memcpy(&m.min_i, smin.data(), 8); memcpy(&m.max_i, smax.data(), 8); }
    m.has_minmax = m.min_i <= m.max_i;
  }
  return m;
}

ColumnChunk parse_column_chunk(TC& c) {
  ColumnChunk m;
  bool have = false;
  int16_t fid = 0; int t;
  while (c.next_field(fid, t)) {
    if (fid == 3 && t == 12) { m = parse_column_meta(c); have = true; }
    else c.skip(t);
  }
  if (!have) fail("parquet: column chunk without inline metadata");
  return m;
}

RowGroup parse_row_group(TC& c) {
  RowGroup rg;
  int16_t fid = 0; int t;
  while (c.next_field(fid, t)) {
    if (fid == 1) {
      int et; uint64_t n = c.list_begin(et);
      for (uint64_t i = 0; i < n; ++i) rg.columns.push_back(parse_column_chunk(c));
    } else if (fid == 3) rg.num_rows = c.zigzag();
    else c.skip(t);
  }
  return rg;
}

}  // namespace

FileMeta parse_footer(const uint8_t* buf, int64_t len) {
  if (len < 12 || memcmp(buf, "PAR1", 4) || memcmp(buf + len - 4, "PAR1", 4))
    fail("parquet: bad magic (encrypted or not a parquet file)");
  uint32_t flen;
  memcpy(&flen, buf + len - 8, 4);
  if ((int64_t)flen + 12 > len) fail("parquet: bad footer length");
  TC c{buf, len - 8 - (int64_t)flen, len - 8};
  FileMeta fm;
  int16_t fid = 0; int t;
  while (c.next_field(fid, t)) {
    switch (fid) {
      case 2: {
        int et; uint64_t n = c.list_begin(et);
        for (uint64_t i = 0; i < n; ++i) fm.schema.push_back(parse_schema_element(c));
        break;
      }
      case 3: fm.num_rows = c.zigzag(); break;
      case 4: {
        int et; uint64_t n = c.list_begin(et);
        for (uint64_t i = 0; i < n; ++i) fm.row_groups.push_back(parse_row_group(c));
        break;
      }
      case 6: fm.created_by = c.binary(); break;
      default: c.skip(t);
    }
  }
  return fm;
}

PageHeader parse_page_header(const uint8_t* buf, int64_t pos, int64_t len) {
  TC c{buf, pos, len};
  PageHeader h;
  int16_t fid = 0; int t;
  while (c.next_field(fid, t)) {
    switch (fid) {
      case 1: h.type = (int)c.zigzag(); break;
      case 2: h.uncompressed_size = (int32_t)c.zigzag(); break;
      case 3: h.compressed_size = (int32_t)c.zigzag(); break;
      case 5: {  // DataPageHeader
        int16_t f2 = 0; int t2;
        while (c.next_field(f2, t2)) {
          if (f2 == 1) h.num_values = (int32_t)c.zigzag();
          else if (f2 == 2) h.encoding = (int)c.zigzag();
          else if (f2 == 3) h.def_encoding = (int)c.zigzag();
          else if (f2 == 4) h.rep_encoding = (int)c.zigzag();
          else c.skip(t2);
        }
        break;
      }
      case 7: {  // DictionaryPageHeader
        int16_t f2 = 0; int t2;
        while (c.next_field(f2, t2)) {
          if (f2 == 1) h.num_values = (int32_t)c.zigzag();
          else if (f2 == 2) h.encoding = (int)c.zigzag();
          else c.skip(t2);
        }
        break;
      }
      default: c.skip(t);
    }
  }
  h.header_len = (int)(c.p - pos);
  return h;
}

}  // namespace pq
}  // namespace nds
