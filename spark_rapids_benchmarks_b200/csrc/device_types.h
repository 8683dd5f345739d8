// POD structures shared between host code and kernels.
#pragma once
#include <cstdint>

namespace nds {

enum DType : int { DT_I32 = 0, DT_I64 = 1, DT_F64 = 2, DT_STR = 3 };

// A string value on the device is a view into the column's encoded byte buffer:
// low 32 bits = byte offset from ColRef::strbase, high 32 bits = length in bytes.
typedef uint64_t StrView;

constexpr int kMaxPageValues = 32768;     // rows per data page the decode kernel accepts (writers default to 20000)
constexpr int kDecodeThreads = 512;
constexpr int kRunCap = 512;              // run-table entries per batch (shared memory)
constexpr int kPagePad = 16;              // readable slack after every staged page payload

// One data page = one CTA of decode_pages_kernel.
struct PageTask {
  const uint8_t* src;    // page payload: [def levels][bit-width byte][values]; values 16 B aligned
  void* out;             // column output base (int32 / int64 / StrView)
  uint32_t* valid;       // column validity bitmap base; nullptr when the column is REQUIRED
  const void* dict;      // dictionary values (int32/int64 array, or StrView table); may be null
  int64_t out_row;       // first output row of this page
  uint32_t size;         // payload bytes
  uint32_t num_values;   // rows in the page, nulls included
  uint32_t def_len;      // bytes of RLE def-level data after the 4-byte length (0: no def levels)
  uint32_t dict_count;
  uint32_t src_off;      // offset of src inside the column's encoded buffer (string views)
  uint8_t enc;           // 0 PLAIN, 1 RLE_DICTIONARY
  uint8_t width;         // 4 or 8 bytes per output value
  uint8_t phys;          // 0 fixed-width, 1 BYTE_ARRAY
  uint8_t bw;            // dictionary index bit width (enc == 1)
  uint32_t* vr;          // split decode only: page-local scratch [nw validity words][nw + 1 rank words] (nw = ceil(n/32))
  // directly addressable dictionary pages (engine.h, PageInfo::direct): values from index tail_k on are one RLE run of
  // dictionary index tail_ix (writers flush a trailing value or two that way); 0xFFFFFFFF = the page has no such tail
  uint32_t tail_k;
  uint32_t tail_ix;
};

// Column reference as kernels see it.
struct ColRef {
  const void* data;         // typed values (base rows)
  const uint32_t* valid;    // validity bitmap over base rows, or nullptr = all valid
  const uint32_t* rowmap;   // optional indirection: base row = rowmap[i]
  const uint8_t* strbase;   // base of the byte buffer StrViews point into
  int32_t dtype;
  int32_t pad;
};

}  // namespace nds
