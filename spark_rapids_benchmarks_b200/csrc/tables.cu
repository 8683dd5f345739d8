// Table registration: directory walk (hive partitions), footer + page-header parse, staging of the
// encoded column chunks, upload to HBM.
//
// Mirrors what `setup_tables` makes Spark do per table (/root/reference/nds/nds_power.py:79-106:
// read.format(fmt).load(prefix/table).createOrReplaceTempView(table)) for the layout written by
// `store()` (/root/reference/nds/nds_transcode.py:131-150): fact tables as <col>=<value>/ directories with
// the partition column only in the directory name, dimension tables as one file.
#include <algorithm>
#include <atomic>
#include <mutex>
#include <thread>
#include <dirent.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
#include "engine.h"

namespace nds {

std::string LType::str() const {
  switch (kind) {
    case INT: return is32 ? "int32" : "int64";
    case DATE: return "date";
    case DECIMAL: return "decimal(" + std::to_string(precision) + "," + std::to_string(scale) + ")";
    case F64: return "double";
    case STR: return "string";
    case BOOL: return "bool";
  }
  return "?";
}

namespace {

const char* kHiveNull = "__HIVE_DEFAULT_PARTITION__";

bool ends_with(const std::string& s, const std::string& suf) {
  return s.size() >= suf.size() && !s.compare(s.size() - suf.size(), suf.size(), suf);
}

std::vector<std::string> list_dir(const std::string& path) {
  std::vector<std::string> out;
  DIR* d = opendir(path.c_str());
  if (!d) fail("cannot open directory '" + path + "'");
  while (dirent* e = readdir(d)) {
    std::string n = e->d_name;
    if (n == "." || n == "..") continue;
    out.push_back(n);
  }
  closedir(d);
  std::sort(out.begin(), out.end());
  return out;
}

bool is_dir(const std::string& p) {
  struct stat st;
  return stat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode);
}

void read_file(const std::string& path, std::vector<uint8_t>& buf) {
  int fd = open(path.c_str(), O_RDONLY);
  if (fd < 0) fail("cannot open '" + path + "'");
  struct stat st;
  fstat(fd, &st);
  buf.resize(st.st_size);
  int64_t got = 0;
  while (got < st.st_size) {
    ssize_t r = read(fd, buf.data() + got, st.st_size - got);
    if (r <= 0) { close(fd); fail("short read on '" + path + "'"); }
    got += r;
  }
  close(fd);
}

LType logical_of(const pq::SchemaElement& e, bool& supported, std::string& why) {
  LType t;
  supported = true;
  bool dec = e.converted == pq::CT_DECIMAL || e.logical_decimal;
  switch (e.type) {
    case pq::INT32:
      t.is32 = true;
      if (dec) { t.kind = LType::DECIMAL; t.precision = e.precision; t.scale = e.scale; }
      else if (e.converted == pq::CT_DATE || e.logical_date) t.kind = LType::DATE;
      else t.kind = LType::INT;
      break;
    case pq::INT64:
      if (dec) { t.kind = LType::DECIMAL; t.precision = e.precision; t.scale = e.scale; }
      else t.kind = LType::INT;
      break;
    case pq::DOUBLE: t.kind = LType::F64; break;
    case pq::BYTE_ARRAY:
      if (dec) { supported = false; why = "decimal stored as BYTE_ARRAY"; }
      t.kind = LType::STR;
      break;
    default:
      supported = false;
      why = "parquet physical type " + std::to_string(e.type) + " has no device decoder";
  }
  return t;
}

void stage_align(std::vector<uint8_t>& s, int64_t a) { s.resize(align_up((int64_t)s.size(), a), 0); }

}  // namespace

void Engine::register_table(const std::string& name, const std::string& path, const std::string& format) {
  if (format != "parquet")
    fail("input format '" + format + "' is not supported by the B200 engine (only uncompressed parquet)");
  if (!is_dir(path)) fail("table path '" + path + "' is not a directory");
  Table t;
  t.name = name;
  t.path = path;

  // ---- enumerate files (one level of hive partition directories) ---------------------------------
  std::vector<FileInfo> all;
  for (const std::string& e : list_dir(path)) {
    std::string p = path + "/" + e;
    if (e[0] == '_' || e[0] == '.') continue;  // _SUCCESS, _b200gen.json, .crc
    if (is_dir(p)) {
      size_t eq = e.find('=');
      if (eq == std::string::npos) fail("unexpected directory '" + p + "' (not <col>=<value>)");
      std::string col = e.substr(0, eq), val = e.substr(eq + 1);
      if (t.part_col.empty()) t.part_col = col;
      else if (t.part_col != col) fail("mixed partition columns under '" + path + "'");
      for (const std::string& f : list_dir(p)) {
        if (!ends_with(f, ".parquet")) continue;
        FileInfo fi;
        fi.path = p + "/" + f;
        if (val == kHiveNull) fi.part_null = true;
        else {
          char* endp = nullptr;
          fi.part_val = strtoll(val.c_str(), &endp, 10);
          if (!endp || *endp) fail("non-integer partition value '" + val + "' under '" + path + "'");
        }
        all.push_back(fi);
      }
    } else if (ends_with(e, ".parquet")) {
      FileInfo fi;
      fi.path = p;
      all.push_back(fi);
    }
  }
  if (all.empty()) fail("no parquet files under '" + path + "'");
  // multi-GPU: partitioned (fact) tables are sharded by file, dimension tables replicated
  for (size_t i = 0; i < all.size(); ++i)
    if (t.part_col.empty() || world <= 1 || (int)(i % world) == rank) t.files.push_back(all[i]);

  // ---- parse + stage: the first file establishes the schema; the others are parsed and staged by a pool of threads, each
  // into buffers of its own (offsets relative to the file), and placed one after the other afterwards
  auto stage_one = [&](size_t fidx, std::vector<TableColumn>& cols, bool first) {
    FileInfo& fi = t.files[fidx];
    std::vector<uint8_t> buf;
    read_file(fi.path, buf);
    pq::FileMeta fm = pq::parse_footer(buf.data(), (int64_t)buf.size());
    if (fm.schema.empty()) fail("empty schema in '" + fi.path + "'");
    const int ncol = (int)fm.schema.size() - 1;
    if ((int)fm.schema[0].num_children != ncol) fail("nested schema in '" + fi.path + "' is not supported");
    if (first) {
      for (int c = 0; c < ncol; ++c) {
        const pq::SchemaElement& e = fm.schema[c + 1];
        TableColumn tc;
        tc.name = e.name;
        tc.phys = e.type;
        tc.lt = logical_of(e, tc.supported, tc.why_unsupported);
        if (e.repetition == 2) { tc.supported = false; tc.why_unsupported = "repeated field"; }
        t.col_index[tc.name] = (int)cols.size();
        cols.push_back(std::move(tc));
      }
    } else if (ncol != (int)cols.size()) {
      fail("schema mismatch between files of table '" + name + "'");
    }
    std::vector<int> max_def(ncol, 0);
    for (int c = 0; c < ncol; ++c) {
      const pq::SchemaElement& e = fm.schema[c + 1];
      if (e.name != cols[c].name || e.type != cols[c].phys) fail("schema mismatch in '" + fi.path + "'");
      max_def[c] = e.repetition == 1 ? 1 : 0;
      if (max_def[c]) cols[c].nullable = true;
    }
    fi.rows = fm.num_rows;
    for (int c = 0; c < ncol; ++c) {
      TableColumn& tc = cols[c];
      stage_align(tc.stage, 16);
      tc.file_page_begin.push_back((int64_t)tc.pages.size());
      tc.file_byte_begin.push_back((int64_t)tc.stage.size());
    }
    int64_t row_base = 0;
    for (const pq::RowGroup& rg : fm.row_groups) {
      if ((int)rg.columns.size() != ncol) fail("row group column count mismatch in '" + fi.path + "'");
      for (int c = 0; c < ncol; ++c) {
        TableColumn& tc = cols[c];
        if (!tc.supported) continue;
        const pq::ColumnChunk& cc = rg.columns[c];
        ++tc.total_chunks;
        if (cc.has_minmax) {
          if (tc.stat_chunks == 0) { tc.stat_lo = cc.min_i; tc.stat_hi = cc.max_i; }
          else { tc.stat_lo = std::min(tc.stat_lo, cc.min_i); tc.stat_hi = std::max(tc.stat_hi, cc.max_i); }
          ++tc.stat_chunks;
        }
        if (cc.codec != 0)
          fail("column '" + tc.name + "' in '" + fi.path + "' is compressed (codec " + std::to_string(cc.codec) +
               "); this engine reads uncompressed parquet only (nds_transcode --compression none)");
        int64_t pos = cc.data_page_offset;
        if (cc.dictionary_page_offset > 0 && cc.dictionary_page_offset < pos) pos = cc.dictionary_page_offset;
        int64_t seen = 0, row_in_file = row_base;
        int cur_dict = -1;
        while (seen < cc.num_values) {
          pq::PageHeader ph = pq::parse_page_header(buf.data(), pos, (int64_t)buf.size());
          const uint8_t* payload = buf.data() + pos + ph.header_len;
          const int64_t psize = ph.compressed_size;
          if (ph.compressed_size != ph.uncompressed_size) fail("compressed page in '" + fi.path + "'");
          if (pos + ph.header_len + psize > (int64_t)buf.size()) fail("page overruns file '" + fi.path + "'");
          if (ph.type == pq::DICTIONARY_PAGE) {
            if (ph.encoding != pq::PLAIN && ph.encoding != pq::PLAIN_DICTIONARY) fail("unsupported dictionary encoding");
            DictInfo di;
            di.count = (uint32_t)ph.num_values;
            stage_align(tc.stage, 16);
            if (tc.phys == pq::BYTE_ARRAY) {
              // index the dictionary entries once at load time: StrView table, then the raw page
              int64_t views_off = (int64_t)tc.stage.size();
              tc.stage.resize(views_off + 8 * (int64_t)di.count);
              int64_t chars_off = (int64_t)tc.stage.size();
              tc.stage.insert(tc.stage.end(), payload, payload + psize);
              int64_t p = 0;
              for (uint32_t k = 0; k < di.count; ++k) {
                if (p + 4 > psize) fail("truncated string dictionary in '" + fi.path + "'");
                uint32_t len;
                memcpy(&len, payload + p, 4);
                int64_t abs_off = chars_off + p + 4;
                if (abs_off + len > 0xFFFFFFFFll) fail("string column '" + tc.name + "' exceeds 4 GiB of encoded bytes");
                uint64_t v = ((uint64_t)len << 32) | (uint64_t)abs_off;
                memcpy(tc.stage.data() + views_off + 8 * (int64_t)k, &v, 8);
                tc.lt.maxlen = std::max<int>(tc.lt.maxlen, (int)std::min<uint32_t>(len, 1u << 30));
                p += 4 + len;
              }
              di.off = views_off;
              tc.view_tables.emplace_back(views_off, di.count);
            } else {
              di.off = (int64_t)tc.stage.size();
              tc.stage.insert(tc.stage.end(), payload, payload + psize);
            }
            tc.stage.resize(tc.stage.size() + kPagePad, 0);
            cur_dict = (int)tc.dicts.size();
            tc.dicts.push_back(di);
          } else if (ph.type == pq::DATA_PAGE) {
            PageInfo pi;
            pi.file = (int32_t)fidx;
            pi.num_values = (uint32_t)ph.num_values;
            pi.size = (uint32_t)psize;
            pi.row_in_file = row_in_file;
            pi.def_len = 0;
            if (ph.num_values > kMaxPageValues)
              fail("data page with " + std::to_string(ph.num_values) + " values in '" + fi.path +
                   "' exceeds the decode kernel limit of " + std::to_string(kMaxPageValues));
            int64_t values_off = 0;
            if (max_def[c] > 0) {
              if (ph.def_encoding != pq::RLE) fail("definition levels not RLE-encoded in '" + fi.path + "'");
              uint32_t dl;
              memcpy(&dl, payload, 4);
              pi.def_len = dl;
              values_off = 4 + (int64_t)dl;
              if (values_off > psize) fail("bad definition level length in '" + fi.path + "'");
              if (dl == 0) fail("empty definition level stream in '" + fi.path + "'");
            }
            if (ph.encoding == pq::PLAIN) {
              pi.enc = 0;
              pi.dict = -1;
              if (tc.phys == pq::BYTE_ARRAY) {       // longest value of the page bounds sort-key widths later
                int64_t p = values_off;
                while (p + 4 <= psize) {
                  uint32_t len;
                  memcpy(&len, payload + p, 4);
                  tc.lt.maxlen = std::max<int>(tc.lt.maxlen, (int)std::min<uint32_t>(len, 1u << 30));
                  p += 4 + (int64_t)len;
                }
              }
            } else if (ph.encoding == pq::RLE_DICTIONARY || ph.encoding == pq::PLAIN_DICTIONARY) {
              pi.enc = 1;
              pi.dict = cur_dict;
              if (values_off >= psize) fail("dictionary-encoded page without index data in '" + fi.path + "'");
              pi.bw = payload[values_off];
              if (pi.bw > 32) fail("dictionary index bit width " + std::to_string(pi.bw) + " in '" + fi.path + "'");
              if (cur_dict < 0) fail("dictionary-encoded page without dictionary in '" + fi.path + "'");
            } else fail("unsupported data page encoding " + std::to_string(ph.encoding) + " in '" + fi.path + "'");
            // route: the split def/rows kernels take PLAIN fixed-width pages and dictionary pages whose index stream
            // is a few long runs (counted here, up to 64); everything else goes to the general kernel
            pi.fast = 0;
            if (pi.def_len <= 8192 && ph.num_values > 0) {
              if (pi.enc == 0) pi.fast = tc.phys != pq::BYTE_ARRAY;
              else if (pi.bw == 0) pi.fast = 1;
              else {
                int64_t p = values_off + 1, runs = 0;
                const uint32_t bytew = (pi.bw + 7) / 8;
                while (p < psize && runs <= 64) {
                  uint64_t h = 0;
                  int sft = 0;
                  uint8_t c;
                  do { c = payload[p++]; h |= (uint64_t)(c & 0x7F) << sft; sft += 7; } while ((c & 0x80) && p < psize && sft < 35);
                  p += (h & 1) ? (int64_t)(h >> 1) * pi.bw : bytew;
                  ++runs;
                }
                pi.fast = runs <= 64;
              }
            }
            // directly addressable pages (engine.h, PageInfo::direct): the whole index stream is checked here, once
            pi.direct = 0;
            if (pi.def_len <= 8192 && ph.num_values > 0) {
              if (pi.enc == 0) pi.direct = tc.phys != pq::BYTE_ARRAY;
              else if (pi.bw == 0) pi.direct = 1;
              else {
                int64_t p = values_off + 1;
                const int64_t full = 1 + 63 * (int64_t)pi.bw;
                int64_t covered = 0;
                bool ok = true;
                while (p < psize && ok) {
                  const uint8_t h = payload[p];
                  if (h == 0x7F && p + full <= psize) { p += full; covered += 504; continue; }
                  // after the full runs: at most one shorter bit-packed run (one-byte header) ...
                  if ((h & 1) && h < 0x80) {
                    const int64_t groups = h >> 1;
                    if (p + 1 + groups * pi.bw > psize) { ok = false; break; }
                    covered += groups * 8;
                    p += 1 + groups * pi.bw;
                  }
                  if (p == psize) break;
                  // ... and at most one closing RLE run (what the writer makes of a last value or two)
                  uint64_t hv = 0;
                  int sft = 0;
                  uint8_t c;
                  do { c = payload[p++]; hv |= (uint64_t)(c & 0x7F) << sft; sft += 7; } while ((c & 0x80) && p < psize && sft < 35);
                  const int64_t bytew = (pi.bw + 7) / 8;
                  if ((hv & 1) || p + bytew != psize) { ok = false; break; }
                  uint32_t v = 0;
                  for (int64_t b = 0; b < bytew; ++b) v |= (uint32_t)payload[p + b] << (8 * b);
                  pi.tail_k = (uint32_t)covered;
                  pi.tail_ix = v;
                  covered += (int64_t)(hv >> 1);
                  p += bytew;
                }
                pi.direct = ok && covered >= 0;
                pi.direct_values = (uint32_t)std::min<int64_t>(covered, 0x7FFFFFFF);
              }
            }
            // place the payload so that its values section is 16 B aligned
            int64_t cur = (int64_t)tc.stage.size();
            int64_t off = align_up(cur + values_off, 16) - values_off;
            tc.stage.resize(off, 0);
            tc.stage.insert(tc.stage.end(), payload, payload + psize);
            tc.stage.resize(tc.stage.size() + kPagePad, 0);
            pi.off = off;
            tc.pages.push_back(pi);
            seen += ph.num_values;
            row_in_file += ph.num_values;
          } else if (ph.type == pq::DATA_PAGE_V2) {
            fail("data page v2 in '" + fi.path + "' is not supported (write with data page version 1.0)");
          }  // index pages are skipped
          pos += ph.header_len + psize;
        }
      }
      row_base += rg.num_rows;
    }
    if (row_base != fi.rows) fail("row group rows do not add up in '" + fi.path + "'");
    };
  if (!t.files.empty()) stage_one(0, t.cols, true);
  const size_t nfiles = t.files.size();
  std::vector<std::vector<TableColumn>> parts(nfiles);        // parts[f]: file f's columns (f >= 1), local offsets
  if (nfiles > 1) {
    auto clone_schema = [&]() {
      std::vector<TableColumn> c(t.cols.size());
      for (size_t k = 0; k < t.cols.size(); ++k) {
        c[k].name = t.cols[k].name;
        c[k].phys = t.cols[k].phys;
        c[k].lt = t.cols[k].lt;
        c[k].lt.maxlen = 0;
        c[k].supported = t.cols[k].supported;
        c[k].why_unsupported = t.cols[k].why_unsupported;
      }
      return c;
    };
    unsigned hw = std::thread::hardware_concurrency();
    const size_t nthreads = std::max<size_t>(1, std::min<size_t>({(size_t)16, (size_t)std::max(1u, hw / (unsigned)std::max(world, 1)), nfiles - 1}));
    std::atomic<size_t> next{1};
    std::mutex err_mu;
    std::string first_error;
    auto worker = [&]() {
      for (;;) {
        const size_t f = next.fetch_add(1);
        if (f >= nfiles) return;
        try {
          parts[f] = clone_schema();
          stage_one(f, parts[f], false);
        } catch (const std::exception& ex) {
          std::lock_guard<std::mutex> l(err_mu);
          if (first_error.empty()) first_error = ex.what();
          next.store(nfiles);
          return;
        }
      }
    };
    std::vector<std::thread> pool;
    for (size_t k = 0; k < nthreads; ++k) pool.emplace_back(worker);
    for (std::thread& th : pool) th.join();
    if (!first_error.empty()) fail(first_error);
  }
  // ---- place the files: global byte / page / dictionary offsets, string-view tables rebased
  for (size_t f = 0; f < nfiles; ++f) t.rows += t.files[f].rows;
  for (size_t c = 0; c < t.cols.size(); ++c) {
    TableColumn& tc = t.cols[c];
    tc.part_base.assign(nfiles, 0);
    int64_t cur = (int64_t)tc.stage.size();          // file 0 is already in place (base 0)
    for (size_t f = 1; f < nfiles; ++f) {
      TableColumn& pc = parts[f][c];
      const int64_t base = align_up(cur, 16);
      tc.part_base[f] = base;
      tc.file_page_begin.push_back((int64_t)tc.pages.size());
      tc.file_byte_begin.push_back(base);
      const int dict_base = (int)tc.dicts.size();
      for (DictInfo di : pc.dicts) { di.off += base; tc.dicts.push_back(di); }
      for (PageInfo pi : pc.pages) { pi.off += base; if (pi.dict >= 0) pi.dict += dict_base; tc.pages.push_back(pi); }
      for (const auto& vt : pc.view_tables)
        for (uint32_t k = 0; k < vt.second; ++k) {
          uint64_t v;
          memcpy(&v, pc.stage.data() + vt.first + 8 * (int64_t)k, 8);
          const uint64_t off = (v & 0xFFFFFFFFull) + (uint64_t)base;
          if (off > 0xFFFFFFFFull) fail("string column '" + tc.name + "' exceeds 4 GiB of encoded bytes");
          v = (v & 0xFFFFFFFF00000000ull) | off;
          memcpy(pc.stage.data() + vt.first + 8 * (int64_t)k, &v, 8);
        }
      tc.nullable = tc.nullable || pc.nullable;
      tc.lt.maxlen = std::max(tc.lt.maxlen, pc.lt.maxlen);
      if (pc.stat_chunks) {
        if (tc.stat_chunks == 0) { tc.stat_lo = pc.stat_lo; tc.stat_hi = pc.stat_hi; }
        else { tc.stat_lo = std::min(tc.stat_lo, pc.stat_lo); tc.stat_hi = std::max(tc.stat_hi, pc.stat_hi); }
      }
      tc.stat_chunks += pc.stat_chunks;
      tc.total_chunks += pc.total_chunks;
      cur = base + (int64_t)pc.stage.size();
      pc.pages.clear();
      pc.pages.shrink_to_fit();
    }
    tc.total_bytes = align_up(cur, 16);
  }
  for (TableColumn& tc : t.cols) {
    tc.file_page_begin.push_back((int64_t)tc.pages.size());
    tc.file_byte_begin.push_back(tc.total_bytes);
    tc.enc_bytes = tc.total_bytes;
    t.encoded_bytes += tc.enc_bytes;
    if (tc.phys == pq::BYTE_ARRAY && tc.enc_bytes > 0xFFFFFFFFll)
      fail("string column '" + tc.name + "' exceeds 4 GiB of encoded bytes");
    // scan-time task templates (engine.h)
    const int width = tc.lt.dtype() == DT_I32 ? 4 : 8;
    tc.task_tmpl.resize(tc.pages.size());
    tc.file_bytes_in.assign(tc.file_page_begin.size() - 1, 0);
    tc.file_direct.assign(tc.file_page_begin.size() - 1, 1);
    tc.file_vr_words.assign(tc.file_page_begin.size() - 1, 0);
    tc.file_chunks.assign(tc.file_page_begin.size() - 1, 0);
    tc.file_max_def.assign(tc.file_page_begin.size() - 1, 0);
    tc.file_def_bytes.assign(tc.file_page_begin.size() - 1, 0);
    std::vector<uint2> page_aux(tc.pages.size());
    for (size_t f = 0; f + 1 < tc.file_page_begin.size(); ++f) {
      uint32_t vw = 0, ch = 0;
      for (int64_t p = tc.file_page_begin[f]; p < tc.file_page_begin[f + 1]; ++p) {
        page_aux[p] = make_uint2(vw, ch);
        vw += 2 * ((tc.pages[p].num_values + 31) / 32) + 1;
        ch += (tc.pages[p].num_values + kFusedChunk - 1) / kFusedChunk;
        tc.file_max_def[f] = std::max<int32_t>(tc.file_max_def[f], (int32_t)tc.pages[p].def_len);
        tc.file_def_bytes[f] += tc.pages[p].def_len;
      }
      tc.file_vr_words[f] = (int32_t)vw;
      tc.file_chunks[f] = (int32_t)ch;
    }
    tc.aux_stage.swap(page_aux);
    tc.has_stats = tc.total_chunks > 0 && tc.stat_chunks == tc.total_chunks && tc.lt.kind != LType::STR && tc.lt.kind != LType::F64;
    for (size_t f = 0; f + 1 < tc.file_page_begin.size(); ++f) {
      int last_dict = -1;
      for (int64_t p = tc.file_page_begin[f]; p < tc.file_page_begin[f + 1]; ++p) {
        const PageInfo& pi = tc.pages[p];
        if (!pi.direct) tc.file_direct[f] = 0;
        PageTask& task = tc.task_tmpl[p];
        memset(&task, 0, sizeof task);
        task.src = reinterpret_cast<const uint8_t*>((uintptr_t)pi.off);
        task.out_row = pi.row_in_file;
        task.size = pi.size;
        task.num_values = pi.num_values;
        task.def_len = pi.def_len;
        task.src_off = (uint32_t)pi.off;
        task.enc = pi.enc;
        task.bw = pi.bw;
        task.width = (uint8_t)width;
        task.phys = tc.phys == pq::BYTE_ARRAY ? 1 : 0;
        task.vr = pi.fast ? reinterpret_cast<uint32_t*>((uintptr_t)1) : nullptr;
        task.tail_k = pi.direct ? pi.tail_k : 0xFFFFFFFFu;
        task.tail_ix = pi.tail_ix;
        tc.file_bytes_in[f] += pi.size;
        if (pi.dict >= 0) {
          task.dict = reinterpret_cast<const void*>((uintptr_t)(tc.dicts[pi.dict].off + 1));
          task.dict_count = tc.dicts[pi.dict].count;
          if (pi.dict != last_dict) {          // each dictionary page is read once per scan
            last_dict = pi.dict;
            tc.file_bytes_in[f] += (int64_t)tc.dicts[pi.dict].count * (tc.phys == pq::BYTE_ARRAY ? 8 : width);
          }
        }
      }
    }
  }
  if (t.part_col.size() && t.col_index.count(t.part_col))
    fail("partition column '" + t.part_col + "' is also stored in the files of '" + name + "'");

  // ---- upload -------------------------------------------------------------------------------------------
  NDS_CUDA(cudaSetDevice(device));
  for (TableColumn& tc : t.cols) {
    if (!tc.supported || tc.enc_bytes == 0) {
      tc.stage.clear();
      tc.stage.shrink_to_fit();
      for (size_t f = 1; f < nfiles; ++f) { parts[f][(size_t)(&tc - t.cols.data())].stage.clear(); }
      continue;
    }
    if (!tc.task_tmpl.empty()) {
      NDS_CUDA(cudaMalloc(&tc.d_tmpl, tc.task_tmpl.size() * sizeof(PageTask)));
      NDS_CUDA(cudaMemcpyAsync(tc.d_tmpl, tc.task_tmpl.data(), tc.task_tmpl.size() * sizeof(PageTask), cudaMemcpyHostToDevice, stream_));
      NDS_CUDA(cudaMalloc(&tc.d_page_aux, tc.aux_stage.size() * sizeof(uint2)));
      NDS_CUDA(cudaMemcpyAsync(tc.d_page_aux, tc.aux_stage.data(), tc.aux_stage.size() * sizeof(uint2), cudaMemcpyHostToDevice, stream_));
      NDS_CUDA(cudaStreamSynchronize(stream_));
      tc.aux_stage.clear();
      tc.aux_stage.shrink_to_fit();
    }
    const size_t cidx = (size_t)(&tc - t.cols.data());
    auto part_bytes = [&](size_t f) -> const std::vector<uint8_t>& { return f == 0 ? tc.stage : parts[f][cidx].stage; };
    if (resident_) {
      NDS_CUDA(cudaMalloc(&tc.d_enc, tc.enc_bytes + 256));
      NDS_CUDA(cudaMemsetAsync(tc.d_enc, 0, tc.enc_bytes + 256, stream_));        // the alignment gaps between files
      for (size_t f = 0; f < nfiles; ++f) {
        const std::vector<uint8_t>& b = part_bytes(f);
        if (!b.empty()) NDS_CUDA(cudaMemcpyAsync(tc.d_enc + tc.part_base[f], b.data(), b.size(), cudaMemcpyHostToDevice, stream_));
      }
      NDS_CUDA(cudaStreamSynchronize(stream_));
    }
    if (keep_host_) {
      // page-locked memory from the driver (not a registered std::vector): host->device copies out of it run at the
      // full PCIe rate
      NDS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&tc.h_enc), (size_t)tc.enc_bytes + 256, cudaHostAllocDefault));
      memset(tc.h_enc, 0, (size_t)tc.enc_bytes + 256);
      for (size_t f = 0; f < nfiles; ++f) {
        const std::vector<uint8_t>& b = part_bytes(f);
        if (!b.empty()) memcpy(tc.h_enc + tc.part_base[f], b.data(), b.size());
      }
      tc.pinned = true;
    }
    tc.stage.clear();
    tc.stage.shrink_to_fit();
    for (size_t f = 1; f < nfiles; ++f) { parts[f][cidx].stage.clear(); parts[f][cidx].stage.shrink_to_fit(); }
  }
  drop_table(name);
  tables_[name] = std::move(t);
}

void Engine::drop_table(const std::string& name) {
  auto it = tables_.find(name);
  if (it == tables_.end()) return;
  for (TableColumn& tc : it->second.cols) {
    if (tc.d_enc) cudaFree(tc.d_enc);
    if (tc.d_tmpl) cudaFree(tc.d_tmpl);
    if (tc.d_page_aux) cudaFree(tc.d_page_aux);
    if (tc.h_enc) cudaFreeHost(tc.h_enc);
  }
  for (MemColumn& mc : it->second.mem_cols) {
    if (mc.d_data) cudaFreeAsync(mc.d_data, stream_);
    if (mc.d_valid) cudaFreeAsync(mc.d_valid, stream_);
    if (mc.d_chars) cudaFreeAsync(mc.d_chars, stream_);
  }
  tables_.erase(it);
}

// ---- tables from Arrow buffers (operator-level tests, gathered partial aggregates) ----------------------
void Engine::register_arrow(const std::string& name, ArrowSchema* schema, ArrowArray* array) {
  NDS_REQUIRE(schema && array, "register_arrow: null schema/array");
  NDS_REQUIRE(std::string(schema->format) == "+s", "register_arrow: expected a struct (record batch)");
  NDS_REQUIRE(array->offset == 0, "register_arrow: sliced record batches are not supported");
  NDS_CUDA(cudaSetDevice(device));
  Table t;
  t.name = name;
  t.is_mem = true;
  t.rows = array->length;
  const int64_t n = array->length;
  for (int64_t c = 0; c < schema->n_children; ++c) {
    ArrowSchema* cs = schema->children[c];
    ArrowArray* ca = array->children[c];
    NDS_REQUIRE(ca->offset == 0, "register_arrow: sliced columns are not supported");
    MemColumn mc;
    mc.name = cs->name ? cs->name : "";
    std::string f = cs->format;
    const uint8_t* validity = (const uint8_t*)ca->buffers[0];
    int64_t vwords = (n + 31) / 32 + 1;
    // stream-ordered pool allocations and copies: registering the gathered partial aggregates of a multi-GPU
    // query sits on the query's critical path, so no cudaMalloc / cudaFree device synchronisation here
    auto upload = [&](const void* src, int64_t bytes) {
      void* d = nullptr;
      NDS_CUDA(cudaMallocAsync(&d, (size_t)bytes + 256, stream_));
      if (bytes) NDS_CUDA(cudaMemcpyAsync(d, src, (size_t)bytes, cudaMemcpyHostToDevice, stream_));
      NDS_CUDA(cudaStreamSynchronize(stream_));     // src may be a temporary
      return d;
    };
    if (validity && ca->null_count != 0) {
      std::vector<uint32_t> v(vwords, 0);
      memcpy(v.data(), validity, (n + 7) / 8);
      mc.d_valid = (uint32_t*)upload(v.data(), vwords * 4);
    }
    if (f == "i") { mc.lt.kind = LType::INT; mc.lt.is32 = true; mc.d_data = upload(ca->buffers[1], n * 4); }
    else if (f == "tdD") { mc.lt.kind = LType::DATE; mc.lt.is32 = true; mc.d_data = upload(ca->buffers[1], n * 4); }
    else if (f == "l") { mc.lt.kind = LType::INT; mc.d_data = upload(ca->buffers[1], n * 8); }
    else if (f == "g") { mc.lt.kind = LType::F64; mc.d_data = upload(ca->buffers[1], n * 8); }
    else if (f.rfind("d:", 0) == 0) {
      int p = 0, s = 0, bits = 128;
      int got = sscanf(f.c_str(), "d:%d,%d,%d", &p, &s, &bits);
      NDS_REQUIRE(got >= 2 && bits == 128 && p <= 18, "register_arrow: only decimal128 with precision <= 18 is supported");
      mc.lt.kind = LType::DECIMAL; mc.lt.precision = p; mc.lt.scale = s;
      std::vector<int64_t> lo(n);
      const int64_t* src = (const int64_t*)ca->buffers[1];
      for (int64_t i = 0; i < n; ++i) lo[i] = src[2 * i];
      mc.d_data = upload(lo.data(), n * 8);
    } else if (f == "u") {
      mc.lt.kind = LType::STR;
      const int32_t* offs = (const int32_t*)ca->buffers[1];
      std::vector<uint64_t> views(n);
      for (int64_t i = 0; i < n; ++i) {
        views[i] = ((uint64_t)(uint32_t)(offs[i + 1] - offs[i]) << 32) | (uint32_t)offs[i];
        mc.lt.maxlen = std::max<int>(mc.lt.maxlen, offs[i + 1] - offs[i]);
      }
      mc.lt.maxlen = std::max(mc.lt.maxlen, 1);
      mc.d_data = upload(views.data(), n * 8);
      mc.d_chars = (uint8_t*)upload(ca->buffers[2], n ? offs[n] : 0);
    } else fail("register_arrow: unsupported column format '" + f + "' for column '" + mc.name + "'");
    t.col_index[mc.name] = (int)t.mem_cols.size();
    t.mem_cols.push_back(mc);
  }
  drop_table(name);
  tables_[name] = std::move(t);
}

std::string Engine::table_info_json(const std::string& name) const {
  auto it = tables_.find(name);
  if (it == tables_.end()) fail("unknown table '" + name + "'");
  const Table& t = it->second;
  std::string s = "{\"name\":\"" + json_escape(name) + "\",\"rows\":" + std::to_string(t.rows) + ",\"files\":" +
                  std::to_string(t.files.size()) + ",\"encoded_bytes\":" + std::to_string(t.encoded_bytes) +
                  ",\"partition_column\":\"" + json_escape(t.part_col) + "\",\"columns\":[";
  bool firstc = true;
  auto add = [&](const std::string& n, const LType& lt, bool nullable, int64_t pages, int64_t bytes, bool ok) {
    if (!firstc) s += ",";
    firstc = false;
    s += "{\"name\":\"" + json_escape(n) + "\",\"type\":\"" + lt.str() + "\",\"nullable\":" + (nullable ? "true" : "false") +
         ",\"pages\":" + std::to_string(pages) + ",\"encoded_bytes\":" + std::to_string(bytes) + ",\"supported\":" +
         (ok ? "true" : "false") + "}";
  };
  for (const TableColumn& c : t.cols) add(c.name, c.lt, c.nullable, (int64_t)c.pages.size(), c.enc_bytes, c.supported);
  for (const MemColumn& c : t.mem_cols) add(c.name, c.lt, c.d_valid != nullptr, 0, 0, true);
  s += "]}";
  return s;
}

}  // namespace nds
