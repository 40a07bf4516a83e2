// tape.hpp -- the flat, device-friendly form of a document ("line tape") and the packed
// module-key index.  Shared by the host flattener and the kernels.
//
// One ROW per line of json.MarshalIndent(doc, "", "\t") output (state/state.go:88-90):
//
//     <depth tabs> [ "<key>": ] <value text> [,] \n
//
// Rows are stored as five parallel u32 arrays (struct of arrays, so a warp reads each field
// with one coalesced request) plus one byte arena that holds the UNESCAPED key and string
// bytes.  Escaping, quoting, indentation and separators are produced on the GPU.
#pragma once
#include <cstdint>

namespace tk {

enum RowKind : uint32_t {
    ROW_STR = 0,        // value is a string: quoted + escaped from arena[voff, voff+vlen)
    ROW_RAW = 1,        // value text copied verbatim (number / true / false / null)
    ROW_OPEN_OBJ = 2,   // {
    ROW_OPEN_ARR = 3,   // [
    ROW_CLOSE_OBJ = 4,  // }
    ROW_CLOSE_ARR = 5,  // ]
    ROW_EMPTY_OBJ = 6,  // {}
    ROW_EMPTY_ARR = 7,  // []
};

// meta word: bits 0-3 kind, 4 has_key, 5 comma after the value, 6 last row (no newline),
// 8 key is "plain", 9 string value is "plain" (every byte is printable ASCII other than " \ < > &,
// so the escaped form IS the byte string and serialising it is a copy),
// bits 16-31 indentation depth (<= TK_MAX_DEPTH = 10000)
constexpr uint32_t META_HAS_KEY = 1u << 4;
constexpr uint32_t META_COMMA = 1u << 5;
constexpr uint32_t META_LAST = 1u << 6;
constexpr uint32_t META_KEY_PLAIN = 1u << 8;
constexpr uint32_t META_VAL_PLAIN = 1u << 9;
constexpr uint32_t META_DEPTH_SHIFT = 16;

struct TapeView {        // pointers may be host (pinned) or device
    uint32_t n_rows;
    uint32_t arena_bytes;
    const uint32_t *meta, *koff, *klen, *voff, *vlen;
    const uint8_t *arena;  // allocation padded so that 16 bytes past arena_bytes are readable
};

// Packed keys of the children of "module", in bytewise key order: key i is
// keys[koffs[i], koffs[i+1]).  This is the array Clusters()/Nodes() classify.
struct KeyIndexView {
    uint32_t n_keys;
    uint32_t key_bytes;
    const uint32_t *koffs;  // n_keys + 1 entries
    const uint8_t *keys;
};

}  // namespace tk
