// Parquet footer / page-header reader (Thrift compact protocol) for UNCOMPRESSED files.
//
// Replaces, for this path, what Spark's FileSourceScanExec does when the reference calls
// spark.read.format("parquet").load(path) (/root/reference/nds/nds_power.py:97-100): footer
// parse, schema discovery, column-chunk and page enumeration. Field ids follow the public
// parquet.thrift (format 2.x).
#pragma once
#include "common.h"

namespace nds {
namespace pq {

enum PhysType { BOOLEAN = 0, INT32 = 1, INT64 = 2, INT96 = 3, FLOAT = 4, DOUBLE = 5, BYTE_ARRAY = 6, FIXED_LEN_BYTE_ARRAY = 7 };
enum Encoding { PLAIN = 0, PLAIN_DICTIONARY = 2, RLE = 3, BIT_PACKED = 4, RLE_DICTIONARY = 8 };
enum PageType { DATA_PAGE = 0, INDEX_PAGE = 1, DICTIONARY_PAGE = 2, DATA_PAGE_V2 = 3 };
// converted types we care about
enum Converted { CT_NONE = -1, CT_UTF8 = 0, CT_DECIMAL = 5, CT_DATE = 6 };

struct SchemaElement {
  std::string name;
  int type = -1;          // PhysType, -1 for groups
  int repetition = 0;     // 0 required, 1 optional, 2 repeated
  int num_children = 0;
  int converted = CT_NONE;
  int scale = 0, precision = 0;
  bool logical_decimal = false, logical_date = false, logical_string = false;
};

struct ColumnChunk {
  int type = -1;
  int codec = 0;
  int64_t num_values = 0;
  int64_t total_uncompressed = 0, total_compressed = 0;
  int64_t data_page_offset = 0;
  int64_t dictionary_page_offset = -1;
  std::string path;
  // Statistics.min_value / max_value (fields 6 / 5; the deprecated min / max, fields 2 / 1, as a fallback) of
  // INT32 / INT64 chunks, little-endian PLAIN bytes; has_minmax only when both were present and well-formed
  bool has_minmax = false;
  int64_t min_i = 0, max_i = 0;
};

struct RowGroup {
  int64_t num_rows = 0;
  std::vector<ColumnChunk> columns;
};

struct FileMeta {
  int64_t num_rows = 0;
  std::vector<SchemaElement> schema;  // flattened, schema[0] is the root
  std::vector<RowGroup> row_groups;
  std::string created_by;
};

struct PageHeader {
  int type = -1;
  int32_t uncompressed_size = 0, compressed_size = 0;
  int32_t num_values = 0;
  int encoding = -1;
  int def_encoding = -1, rep_encoding = -1;
  int header_len = 0;  // bytes of the thrift header itself
};

// buf/len: the whole file. Throws nds::Error on malformed input.
FileMeta parse_footer(const uint8_t* buf, int64_t len);
// parse one page header starting at buf+pos (bounded by len)
PageHeader parse_page_header(const uint8_t* buf, int64_t pos, int64_t len);

}  // namespace pq
}  // namespace nds
