// index.cuh -- launch wrappers of the module-index kernels (index.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "parse.cuh"
#include "tape.hpp"

namespace tk {

struct ChildRef { uint32_t koff, klen, noff, nlen, hoff, hlen; };  // nlen / hlen == 0xffffffff: no such string field

// device-side views of what child_pack_kernel reads and writes (field 0 = "name", 1 = "hostname")
struct IndexTables {
    const unsigned long long *rec[2];         // per child: (has field) << 32 | record bytes
    const unsigned long long *rec_excl[2];    // its exclusive scan: (records before) << 32 | bytes before
    const uint32_t *viol[2];                  // per child: 1 if its field does not sort strictly after the previous child's
    uint8_t *blob[2];                         // packed result records of the children that have the field
    uint32_t *off[2];                         // m + 1 offsets into blob, m = number of such children
    uint32_t *viol_bits[2];                   // m bits, cleared beforehand
    unsigned long long *totals;               // [2]: m << 32 | blob bytes
};

// refs must be pre-filled with 0xff bytes
cudaError_t launch_child_refs(const TapeView &t, uint32_t module_row, const uint32_t *br_match, const uint32_t *br_row, const uint32_t *rank,
                              ChildRef *refs, cudaStream_t s);
cudaError_t launch_child_sizes(const uint8_t *arena, const ChildRef *refs, uint32_t n, unsigned long long *rec_n,
                               unsigned long long *rec_h, uint32_t *viol_n, uint32_t *viol_h, cudaStream_t s);
cudaError_t launch_child_pack(const uint8_t *arena, const ChildRef *refs, uint32_t n, const IndexTables &tb, cudaStream_t s);

}  // namespace tk
