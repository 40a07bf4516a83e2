// kernels.cuh -- launch wrappers of the sm_100a kernels (implemented in kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "tape.hpp"

namespace tk {

// geometry shared by the two serialisation kernels
constexpr int EMIT_THREADS = 256;                       // 8 warps, one 32-row warp tile each per step
constexpr int EMIT_ROWS_PER_BLOCK = EMIT_THREADS;       // rows handled by one block
#ifndef TK_WARP_IN_BYTES
#define TK_WARP_IN_BYTES 2048
#endif
#ifndef TK_WARP_OUT_BYTES
#define TK_WARP_OUT_BYTES 3072
#endif
constexpr int WARP_IN_BYTES = TK_WARP_IN_BYTES;         // per-warp staging of arena bytes
constexpr int WARP_OUT_BYTES = TK_WARP_OUT_BYTES;       // per-warp staging of output text

inline uint32_t emit_tiles(uint32_t n_rows) { return (n_rows + EMIT_ROWS_PER_BLOCK - 1) / EMIT_ROWS_PER_BLOCK; }

// row_offsets works on tiles of RO_TILE rows; its scratch is 2 + tiles 64-bit words (tile counter,
// total, one look-back status word per tile)
constexpr int RO_ITEMS = 8;                             // consecutive rows per thread
constexpr int RO_TILE = EMIT_THREADS * RO_ITEMS;        // 2,048 rows per block: one look-back per 2,048 rows
inline uint32_t row_offsets_tiles(uint32_t n_rows) { return (n_rows + RO_TILE - 1) / RO_TILE; }
inline size_t scan_scratch_words(uint32_t n_rows) { return 2 + (size_t)row_offsets_tiles(n_rows); }

// K1: per-row output length (escaping included) and its exclusive prefix sum.
//     off[r] = byte offset of line r (n_rows + 1 entries); scratch[1] = serialised length as a
//     64-bit number.  Offsets are 32-bit: if the length is 4 GiB or more they have wrapped, and
//     off[n_rows] is set to 0xffffffff to say so.
//     scratch must hold scan_scratch_words(n) words; it is cleared on the stream first.
cudaError_t launch_row_offsets(const TapeView &t, uint32_t *off, unsigned long long *scratch, cudaStream_t s);

// K2: write every line at its offset.
cudaError_t launch_emit(const TapeView &t, const uint32_t *off, uint8_t *out, cudaStream_t s);

// K3: indices of the keys that start with prefix, appended (unordered) to hits[1..];
//     hits[0] = count.  hits must hold n_keys + 1 words; hits[0] is cleared on the stream first.
//     prefix is a device pointer.
cudaError_t launch_prefix_match(const KeyIndexView &k, const uint8_t *prefix, uint32_t prefix_len, uint32_t *hits,
                                cudaStream_t s);

unsigned long long kernel_launch_count();

}  // namespace tk
