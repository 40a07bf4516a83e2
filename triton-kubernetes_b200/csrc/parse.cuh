// parse.cuh -- GPU side of state.New: raw JSON text -> validated line tape (tape.hpp), all on the
// device.  Implemented in parse.cu; driven by Engine::parse (engine.cu).
//
// Pipeline (byte/integer kernels; see DESIGN.md section 3 for what bounds each):
//   P1 scan            bytes -> token stream, with no block waiting on another (scan_fast.cuh):
//                      scan_count  per 8 KiB tile (quote parity, tokens if entered outside a string,
//                                  tokens if entered inside);
//                      scan_prefix one block folds those in order: string state and token count
//                                  at the start of every tile;
//                      scan_emit   tokens (position, kind) written in order at their final places;
//                                  string bodies checked (control bytes, escape codes); strings that
//                                  need decoding (escapes, non-ASCII bytes, < > &) marked.
//                      A backslash run longer than 512 bytes sends the text through the older
//                      single-pass kernel (scan_tokens_kernel: two-stage decoupled look-back).
//   P2a token_prefix   per 2,048-token block: nesting depth, row count and bracket count before it
//                      (per-block aggregates, then one block folds them in order)
//   P2b brackets       bracket tokens compacted with their nesting level
//   P2c radix sort     brackets stably sorted by level (3 bits per pass: histogram, run starts,
//                      scatter), which puts every opener next to its closer
//   P2d match          opener<->closer pairs, enclosing opener of every opener
//   P3  decode         strings marked in P1 are unescaped / UTF-8-repaired into a side arena
//   P2e rows           grammar check (Go's scanner rules restated as local token-pair rules) and
//                      one tape row per value / closer, referring to the raw text in place
//   P4  order          keys strictly ascending inside every object? (then the tape is already in
//                      encoding/json's output order and Bytes() can run straight from it)
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "tape.hpp"

namespace tk {

enum TokKind : uint8_t {
    T_LBRACE = 0, T_RBRACE = 1, T_LBRACKET = 2, T_RBRACKET = 3, T_COLON = 4, T_COMMA = 5,
    T_STR_OPEN = 6, T_STR_CLOSE = 7, T_SCALAR = 8, T_SCALAR_END = 9
};

#ifndef TK_P1_THREADS
#define TK_P1_THREADS 128
#endif
constexpr int P1_THREADS = TK_P1_THREADS;
constexpr int P1_BYTES_PER_THREAD = 64;
constexpr int P1_TILE = P1_THREADS * P1_BYTES_PER_THREAD;  // 8 KiB of text per block
constexpr int TOK_TILE = 2048;                             // tokens per TilePrefix entry (what the token passes index by)

// result block the kernels fill in device memory (one per parse)
struct ParseResult {
    unsigned long long n_tokens;
    unsigned int invalid;         // != 0: json.Unmarshal would fail (the host phrases Go's message)
    unsigned int end_in_string;   // text ended inside a string
    unsigned int n_decode;        // strings that need decoding
    unsigned int n_rows;
    unsigned int n_brackets;
    unsigned int max_level;
    int final_depth;
    unsigned int not_canonical;   // keys out of order / duplicated, or numbers not in canonical form
    unsigned int need_host_numbers;
    unsigned long long side_bytes; // decoded bytes written to the side arena
    unsigned int slow_scan;       // P1 met a backslash run longer than its local look-behind: run the carry protocol
    unsigned int root_kind;       // RowKind of row 0
    unsigned int module_row;      // 1 + the row of the root object's "module" member; 0: there is none
    unsigned int module_kind;     // its RowKind
    unsigned int module_end;      // the row that closes it (an object), else 0
    unsigned int fast_reject;     // the pass for text in indented form (lines.cu) declined: run the general pipeline
    // module index (field_find_kernel, index.cu): per field -- 0 "name", 1 "hostname" -- how many
    // children of "module" carry it as a string member, and the bytes of their packed result records;
    // index_slow: a child had more members than the kernel walks through (Clusters / Nodes then use the host tree)
    unsigned int field_count[2];
    unsigned long long field_bytes[2];
    unsigned int index_slow;
    unsigned int reject_early;    // lines.cu: fast_reject as it stood before the last launch of the pass
};

inline uint32_t p1_tiles(size_t n) { return (uint32_t)((n + P1_TILE - 1) / P1_TILE); }
inline uint32_t tok_tiles(size_t t) { return (uint32_t)((t + TOK_TILE - 1) / TOK_TILE); }

// scratch sizes in 64-bit words
// (two tile tables, then per 64-byte chunk four mask words and one "rare" bit that count hands to emit)
inline size_t p1_scratch_words(size_t n) {
    const size_t tiles = p1_tiles(n), chunks = tiles * P1_THREADS;
    return 4 + 2 * tiles + 4 * chunks + chunks / 64 + 2;
}
inline size_t tok_scratch_words(size_t t) { return 4 + 3 * (size_t)tok_tiles(t); }   // 20 bytes per token block

struct TokenBuffers {
    uint32_t *pos;        // token byte offsets
    uint8_t *kind;        // TokKind
    uint32_t *decode_bits; // 1 bit per token: string needs decoding
    uint32_t cap;         // capacity in tokens
};

struct TilePrefix { int depth; uint32_t rows; uint32_t brackets; uint32_t pad; };

struct BracketBuffers {
    uint32_t *tok;      // token index of bracket b (position order)
    uint32_t *lvkind;   // level << 8 | TokKind
    uint32_t *match;    // partner bracket  (match / parent / row / ct: four slices of one block, cleared to 0xff together)
    uint32_t *parent;   // enclosing opener (0xffffffff: none)
    uint32_t *row;      // row index of an opener's row / a closer's row
    uint32_t *ct;       // container kinds around the bracket (see match_kernel)
    uint32_t *sort_a, *sort_b;  // sorted permutations (ping-pong)
    uint32_t *key_a, *key_b;    // level keys travelling with the permutation
};

struct DecodeBuffers {
    uint32_t *tok;      // token index of the i-th string that needs decoding (ascending)
    uint32_t *off;      // its offset in the arena after decoding
    uint32_t *len;      // its decoded length
};

// raw must be readable (and space-padded) up to p1_tiles(n) * P1_TILE + 16 bytes.
// scratch: p1_scratch_words(n) words.  Clears scratch, *res and tb.decode_bits on the stream first.
// carry_protocol false: every thread finds the backslash-run carry into its 64 bytes by looking at
// most 512 bytes behind it, and sets res->slow_scan if that was not enough (the caller then runs
// the kernel again with carry_protocol true: the carry travels by decoupled look-back).
// stage: 0 clears and counts (or, with carry_protocol, runs the whole single-pass kernel), 1 folds the
// tile counts, 2 emits the tokens -- three calls, so that the caller can time the kernels apart.
cudaError_t launch_scan_tokens(const uint8_t *raw, size_t n, TokenBuffers tb, unsigned long long *scratch,
                               ParseResult *res, bool carry_protocol, int stage, cudaStream_t s);
// scratch: tok_scratch_words(n_tok) words
cudaError_t launch_token_prefix(TokenBuffers tb, uint32_t n_tok, TilePrefix *tiles, unsigned long long *scratch,
                                ParseResult *res, cudaStream_t s);
cudaError_t launch_brackets(TokenBuffers tb, uint32_t n_tok, const TilePrefix *tiles, BracketBuffers bb, uint32_t n_br, cudaStream_t s);
// stable sort of the identity permutation (sort_a/key_a) by level, 3 bits per pass; *in_a says which
// buffer holds the result.  scratch: parse_scan_scratch_words(n_br) words.
cudaError_t launch_sort_brackets(BracketBuffers bb, uint32_t n_br, uint32_t max_level, unsigned long long *scratch, bool *in_a,
                                 cudaStream_t s);
cudaError_t launch_match(BracketBuffers bb, const uint32_t *sorted, const uint32_t *sorted_key, uint32_t n_br,
                         ParseResult *res, cudaStream_t s);
// strings that need decoding: ascending token list (count -> scratch[1]), then decoded lengths and
// their exclusive sum (total -> scratch[1]), then the bytes at arena[side_base + off]
cudaError_t launch_decode_list(TokenBuffers tb, uint32_t n_tok, DecodeBuffers db, uint32_t *counts, uint32_t *offs,
                               unsigned long long *scratch, cudaStream_t s);
cudaError_t launch_decode_lengths(const uint8_t *raw, TokenBuffers tb, DecodeBuffers db, uint32_t n_dec,
                                  unsigned long long *scratch, cudaStream_t s);
cudaError_t launch_decode_write(uint8_t *arena, uint32_t side_base, TokenBuffers tb, DecodeBuffers db, uint32_t n_dec,
                                cudaStream_t s);
// rows: the tape arrays are written in input order; arena = raw text (+ side arena)
struct RowBuffers {
    uint32_t *meta, *koff, *klen, *voff, *vlen;
    // lines.cu only: the open-container words of its last stage; the arena, where its side arena starts and how many
    // bytes of it there are (strings that have to be decoded go there, as in the general parse)
    uint32_t *open = nullptr;
    uint8_t *side = nullptr;
    uint32_t side_base = 0, side_cap = 0;
};
cudaError_t launch_rows(const uint8_t *raw, size_t n, TokenBuffers tb, uint32_t n_tok, const TilePrefix *tiles,
                        BracketBuffers bb, DecodeBuffers db, uint32_t n_dec, RowBuffers rb, uint32_t n_rows,
                        ParseResult *res, cudaStream_t s);
cudaError_t launch_key_order(const uint8_t *arena, RowBuffers rb, uint32_t n_rows, BracketBuffers bb, ParseResult *res,
                             cudaStream_t s);
size_t parse_scan_scratch_words(uint32_t n);
// exclusive scan of n u32 values (single pass, decoupled look-back); the total lands in scratch[1].
// scratch: parse_scan_scratch_words(n) words, cleared on the stream first.
cudaError_t scan_u32(const uint32_t *in, uint32_t *out, uint32_t n, unsigned long long *scratch, cudaStream_t s);
// same for 64-bit words holding two packed sums (high half must stay below 2^30)
cudaError_t scan_u64(const unsigned long long *in, unsigned long long *out, uint32_t n, unsigned long long *scratch, cudaStream_t s);
void count_parse_launch();
unsigned long long parse_launch_count();

}  // namespace tk
