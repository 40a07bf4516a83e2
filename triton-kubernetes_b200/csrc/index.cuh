// index.cuh -- launch wrappers of the module-index kernels (index.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "parse.cuh"
#include "tape.hpp"

namespace tk {

struct ModuleInfo {
    uint32_t root_kind;    // RowKind of row 0
    uint32_t module_kind;  // RowKind of root's "module" member; 0xff if there is none
    uint32_t module_row, module_end;
    uint32_t n_children, pad;
};
struct ChildRef { uint32_t koff, klen, noff, nlen, hoff, hlen; };  // nlen / hlen == 0xffffffff: no such string field

// device-side views of everything child_pack_kernel writes (field 0 = "name", 1 = "hostname")
struct IndexTables {
    const uint32_t *key_excl;                 // exclusive scan of key lengths
    const unsigned long long *rec[2];         // (has field) << 32 | record bytes, per child
    const unsigned long long *rec_excl[2];    // its exclusive scan
    uint8_t *keys;                            // packed keys
    uint32_t *key_off;                        // n + 1 offsets into keys
    uint8_t *blob[2];                         // packed result records per field
    uint32_t *off[2], *cnt[2];                // n + 1 entries each: record bytes / records before child i
};

cudaError_t launch_find_module(const TapeView &t, const uint32_t *br_match, const uint32_t *br_row, ModuleInfo *info, cudaStream_t s);
cudaError_t launch_child_flags(const TapeView &t, const ModuleInfo *info, uint32_t *flags, cudaStream_t s);
// refs must be pre-filled with 0xff bytes
cudaError_t launch_child_refs(const TapeView &t, const ModuleInfo *info, const uint32_t *flags, const uint32_t *rank, ChildRef *refs,
                              cudaStream_t s);
cudaError_t launch_child_sizes(const uint8_t *arena, const ChildRef *refs, uint32_t n, uint32_t *key_bytes, unsigned long long *rec_n,
                               unsigned long long *rec_h, uint32_t *viol_n, uint32_t *viol_h, cudaStream_t s);
cudaError_t launch_child_pack(const uint8_t *arena, const ChildRef *refs, uint32_t n, const IndexTables &tb, cudaStream_t s);

}  // namespace tk
