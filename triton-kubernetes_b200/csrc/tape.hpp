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
//
// A tape the parser built (TapeView::text_len != 0) refers to the input text in place and uses the
// fields a container row has no use for: voff of an opening row is the index of its bracket
// (BracketBuffers), koff of a closing row the bracket of its opener, vlen of any container row the
// text position of the bracket.  A tape flattened from the host tree has zeros there.
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
// 10 / 11 (tapes the parser built only) the key reads "name" / "hostname" and the row is a string
// member three levels down: what the module index looks for (index.cu),
// bits 16-31 indentation depth (<= TK_MAX_DEPTH = 10000)
constexpr uint32_t META_HAS_KEY = 1u << 4;
constexpr uint32_t META_COMMA = 1u << 5;
constexpr uint32_t META_LAST = 1u << 6;
constexpr uint32_t META_KEY_PLAIN = 1u << 8;
constexpr uint32_t META_VAL_PLAIN = 1u << 9;
constexpr uint32_t META_KEY_IS_NAME = 1u << 10;
constexpr uint32_t META_KEY_IS_HOSTNAME = 1u << 11;
constexpr uint32_t META_DEPTH_SHIFT = 16;

struct TapeView {        // pointers may be host (pinned) or device
    uint32_t n_rows;
    uint32_t arena_bytes;
    const uint32_t *meta, *koff, *klen, *voff, *vlen;
    const uint8_t *arena;  // allocation padded so that 16 bytes past arena_bytes are readable
    // != 0: the tape came from the parser -- arena[0, text_len) IS the input text, keys and plain
    // strings are referred to in place, and the rows of containers ({ [ } ] {} []) carry the text
    // position of their bracket in vlen (otherwise 0).  emit_kernel uses it to recognise input that
    // is already in output form.  0 for a tape flattened from the host tree.
    uint32_t text_len;
    // true: the tape came from the pass for text in indented form (lines.cu) -- voff of an opening row
    // is the ROW that closes it, koff of a closing row the ROW that opened it.  false (with text_len
    // != 0): they are bracket numbers (BracketBuffers), as the general parse leaves them.
    bool rows_direct;
};

#ifdef __CUDACC__
// Bytes base[off .. off+3] as one big-endian number, so that comparing numbers compares byte
// strings.  base must be 4-byte aligned and readable 7 bytes past off (every arena is padded).
__device__ __forceinline__ uint32_t load4_be(const uint8_t *base, uint32_t off) {
    const uint32_t *w = reinterpret_cast<const uint32_t *>(base + (off & ~3u));
    return __byte_perm(__funnelshift_r(w[0], w[1], (off & 3u) * 8), 0, 0x0123);
}
// memcmp-style order of two byte strings of the arena (bytewise, the shorter first on a tie): -1, 0, 1
__device__ __forceinline__ int compare_bytes(const uint8_t *arena, uint32_t ao, uint32_t al, uint32_t bo, uint32_t bl) {
    const uint32_t m = al < bl ? al : bl;
    for (uint32_t i = 0; i < m; i += 4) {
        uint32_t x = load4_be(arena, ao + i), y = load4_be(arena, bo + i);
        if (m - i < 4) { const uint32_t mask = 0xffffffffu << ((4 - (m - i)) * 8); x &= mask; y &= mask; }
        if (x != y) return x < y ? -1 : 1;
    }
    return al < bl ? -1 : (al > bl ? 1 : 0);
}
#endif

// Packed keys of the children of "module", in bytewise key order: key i is
// keys[koffs[i], koffs[i+1]).  This is the array Clusters()/Nodes() classify.
struct KeyIndexView {
    uint32_t n_keys;
    uint32_t key_bytes;
    const uint32_t *koffs;  // n_keys + 1 entries
    const uint8_t *keys;
};

}  // namespace tk
