// index.cuh -- launch wrappers of the module-index kernels (index.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "parse.cuh"
#include "tape.hpp"

namespace tk {

// What State.Clusters() / State.Nodes() read (state/state.go:94-147): every string member called
// "name" (field 0) or "hostname" (field 1) of an object that is a child of the root's "module", as
// (field value, child key) references into the arena, in document order.
struct FieldRef { uint32_t foff, flen, koff, klen; };
struct FieldIndexBuffers {
    FieldRef *refs[2];
    uint32_t *off[2];             // cap + 1 entries: where record j starts in the packed blob; off[m] = its size
    uint32_t cap[2];              // a list that outgrows its buffer is counted, not stored: the caller runs the kernel again
    unsigned long long *status;   // field_status_words(n_rows) words, cleared by the launch wrapper
};
size_t field_status_words(uint32_t n_rows);
// One pass over the rows of a tape the parser built; reads where "module" is from *res (rows_kernel and
// key_order_kernel left it there) and leaves the list sizes in res->field_count / field_bytes.
cudaError_t launch_field_find(const TapeView &t, FieldIndexBuffers fx, ParseResult *res, cudaStream_t s);

// Both lists packed into result records { u32 len, field, u32 len, key } at the offsets field_find
// left (fx.off), in one launch; bit j of viol_bits[f] (m[f] bits, cleared beforehand) marks a record
// whose field does not sort strictly after the record before.
cudaError_t launch_field_pack(const uint8_t *arena, const FieldIndexBuffers &fx, const uint32_t m[2], uint32_t *const viol_bits[2],
                              uint8_t *const blob[2], cudaStream_t s);

}  // namespace tk
