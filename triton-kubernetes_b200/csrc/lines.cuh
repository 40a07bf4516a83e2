// lines.cuh -- state.New for text that is already in encoding/json's indented form (lines.cu).
//
// Every document this library or the reference writes is json.MarshalIndent(doc, "", "\t") output,
// and loading such a file back is what backend.State does all day (backend/local/backend.go:43-66).
// In that form one line of text is one row of the line tape, so the general pipeline's detour
// through seven million tokens is not needed: the lines are found (a raw newline cannot occur
// inside a valid string), each line is parsed and checked on its own against the exact grammar
// of the indented form, neighbouring lines are checked against each other (indentation, commas,
// key order), and one pass over the rows pairs closers with openers and checks what only the
// enclosing container can tell (keys in objects, none in arrays, bracket kinds, order across a
// closed container).  Strings with text beyond ASCII are kept where they stand if they are valid
// UTF-8 as json.Marshal writes it; string VALUES that read differently once decoded (escape
// sequences, bytes Marshal would have escaped, invalid UTF-8) are decoded into a side arena behind
// the text, as the general parser does.  Anything that is not that form -- a key that would have
// to be decoded, a number that would be re-formatted, a key out of order, nesting deeper than LINES_MAX_DEPTH, a
// space too many -- makes the pass say "not for me" (ParseResult::fast_reject) and the caller
// runs the general pipeline on the same text, which also owns every error message.  A text the
// pass accepts is valid JSON in canonical form, and the tape it leaves is row for row the one the
// general pipeline builds, except that an opening row's voff is the ROW that closes it and a closing
// row's koff the ROW that opened it (TapeView::rows_direct), not a bracket number.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "index.cuh"
#include "parse.cuh"
#include "tape.hpp"

namespace tk {

constexpr uint32_t LINES_MAX_DEPTH = 7;   // openers at depths 0 .. 7 are tracked, so rows down to depth 8 can be placed

constexpr int LN_TILE = 8192;             // bytes of text per block (128 threads x 64 bytes)
inline uint32_t lines_tiles(size_t n) { return (uint32_t)((n + 1 + LN_TILE - 1) / LN_TILE); }   // position n (the end) counts as a terminator
// scratch in 32-bit words: per tile the number of line ends, then the position of its last one
inline size_t lines_scratch_words(size_t n) { return 2 * (size_t)lines_tiles(n) + 8; }
constexpr int LN_ROWS_PER_BLOCK = 1024;   // rows per block of the structure pass
inline uint32_t lines_row_blocks(uint32_t rows) { return (rows + LN_ROWS_PER_BLOCK - 1) / LN_ROWS_PER_BLOCK; }
// open_state in 32-bit words: per block of rows 8 open-container words, then (as 64-bit words) a ticket counter,
// a spare, and per field and block one look-back word of the module index's lists
inline size_t lines_open_state_words(uint32_t row_cap) { const size_t b = lines_row_blocks(row_cap); return b * 8 + 2 * (2 + 2 * b); }

// raw: the text at the start of the arena, space padded to a whole tile and 64 bytes more.
// Stage 0: line ends counted per tile, folded (res->n_rows known on the device afterwards);
// open_state (lines_open_state_words(row_cap) words) cleared.
// src != nullptr: the text is not in the arena yet but at src in device memory (16-byte aligned); the count pass
// copies its n bytes to raw as it reads them (the padding behind them must already be there).
cudaError_t launch_lines_count(uint8_t *raw, const uint8_t *src, size_t n, uint32_t *scratch, uint32_t *open_state, uint32_t row_cap,
                               ParseResult *res, cudaStream_t s);
// Stage 1: one row per line into rb (capacity row_cap rows; more lines than that: fast_reject);
// rb.open = open_state, where every opening row leaves its mark.
cudaError_t launch_lines_rows(const uint8_t *raw, size_t n, const uint32_t *scratch, RowBuffers rb, uint32_t row_cap, ParseResult *res,
                              cudaStream_t s);
// Stage 2: closers paired with openers, containers checked; the module index's field lists (fx.refs, fx.off,
// fx.cap given; res->field_count / field_bytes as after launch_field_find) made on the way.
cudaError_t launch_lines_structure(const uint8_t *raw, RowBuffers rb, uint32_t row_cap, uint32_t *open_state, FieldIndexBuffers fx, ParseResult *res,
                                   cudaStream_t s);

}  // namespace tk
