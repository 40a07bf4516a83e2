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
struct ChildEntry { uint32_t off, klen, nlen, hlen; };             // strings packed at off: key, name, hostname

cudaError_t launch_find_module(const TapeView &t, const uint32_t *br_match, const uint32_t *br_row, ModuleInfo *info, cudaStream_t s);
cudaError_t launch_child_flags(const TapeView &t, const ModuleInfo *info, uint32_t *flags, cudaStream_t s);
cudaError_t launch_child_entries(const TapeView &t, const uint32_t *flags, const uint32_t *rank, const uint32_t *br_match,
                                 const uint32_t *br_row, ChildRef *refs, uint32_t *sizes, cudaStream_t s);
cudaError_t launch_child_pack(const uint8_t *arena, const ChildRef *refs, const uint32_t *off, uint32_t n, ChildEntry *entries,
                              uint8_t *packed, cudaStream_t s);

// which: 0 = "name" (Clusters), 1 = "hostname" (Nodes).  cs[i] = has << 32 | record bytes; viol[i] = 1
// where child i's field value does not strictly exceed its predecessor's (or the predecessor has none)
cudaError_t launch_field_sizes(const uint8_t *arena, const ChildRef *refs, uint32_t n, int which, unsigned long long *cs, uint32_t *viol,
                               cudaStream_t s);
// cs_excl = exclusive scan of cs.  Writes the records into blob and, per child, the bytes and the
// records before it (off, cnt: n + 1 entries each)
cudaError_t launch_field_pack(const uint8_t *arena, const ChildRef *refs, uint32_t n, int which, const unsigned long long *cs_excl,
                              const unsigned long long *cs, uint8_t *blob, uint32_t *off, uint32_t *cnt, cudaStream_t s);

}  // namespace tk
