// edit.cuh -- device side of edits to a document that lives on the device (edit.cu): the list of
// module's children a first edit needs, and the rebuild of the line tape from row segments.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "parse.cuh"
#include "tape.hpp"

namespace tk {

// flags[r] = 1 for every row that starts a child of the object "module" spans (rows module_row < r < module_end
// at depth 2 that are not closers), else 0; n_rows entries
cudaError_t launch_child_flags(const TapeView &t, uint32_t module_row, uint32_t module_end, uint32_t *flags, cudaStream_t s);
// per child (rank[r] = number of child rows before r): its first and last row and its key reference
struct ChildList { uint32_t *row, *end, *koff, *klen; };
cudaError_t launch_child_entries(const TapeView &t, const uint32_t *flags, const uint32_t *rank, const uint32_t *br_match, const uint32_t *br_row,
                                 ChildList out, cudaStream_t s);
// keys copied one after another: key j at blob + key_off[j]
cudaError_t launch_child_keys(const uint8_t *arena, ChildList c, const uint32_t *key_off, uint32_t n_children, uint8_t *blob, cudaStream_t s);

// The tape rebuilt from segments of the old one: new rows [dst, dst + n) are old rows [src, src + n).
// Segments are sorted by dst and cover [0, n_new) without gaps.
struct RowSegment { uint32_t dst, src, n; };
struct RowArrays { uint32_t *meta, *koff, *klen, *voff, *vlen; };
cudaError_t launch_gather_rows(RowArrays from, RowArrays to, const RowSegment *segs, uint32_t n_segs, uint32_t n_new, cudaStream_t s);
// after the rebuild: the comma after a child's last row set (rows_set) or cleared (rows_clear); 0xffffffff = none
cudaError_t launch_patch_commas(uint32_t *meta, uint32_t row_set, uint32_t row_clear, cudaStream_t s);

}  // namespace tk
