// edit.cu -- kernels behind State.Delete / AddNode / AddCluster / SetManager on a document that
// lives on the device (state/state.go:36-86 over gabs SetP / DeleteP, gabs.go:180-208,276-301).
//
// destroy/node.go:110-180 and destroy/cluster.go:126-161 do Nodes -> Delete ... -> Bytes, and
// create/cluster.go:146-284 does AddCluster -> Bytes -> New -> Nodes -> AddNode ... -> Bytes, all on
// children of "module".  The host keeps the list of those children (key, rows, where they are in
// the tape) and the edits made to it; when the text or the tree is next needed, the tape is
// rebuilt on the device from runs of surviving rows plus the rows of added children -- no host
// tree, no second upload of the document.  See Engine::apply_edits (engine.cu).
#include "edit.cuh"

namespace tk {
namespace {

__global__ void child_flags_kernel(TapeView t, uint32_t module_row, uint32_t module_end, uint32_t *__restrict__ flags) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= t.n_rows) return;
    uint32_t f = 0;
    if (r > module_row && r < module_end) {
        const uint32_t meta = __ldg(t.meta + r), kind = meta & 15u;
        f = (meta >> META_DEPTH_SHIFT) == 2 && kind != ROW_CLOSE_OBJ && kind != ROW_CLOSE_ARR ? 1u : 0u;
    }
    flags[r] = f;
}

__global__ void child_entries_kernel(TapeView t, const uint32_t *__restrict__ flags, const uint32_t *__restrict__ rank,
                                     const uint32_t *__restrict__ br_match, const uint32_t *__restrict__ br_row, ChildList out) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= t.n_rows || !flags[r]) return;
    const uint32_t j = rank[r], kind = __ldg(t.meta + r) & 15u;
    uint32_t end = r;
    if (kind == ROW_OPEN_OBJ || kind == ROW_OPEN_ARR) {   // voff of an opening row: the row that closes it, or its bracket
        const uint32_t v = __ldg(t.voff + r);
        if (t.rows_direct) end = v;
        else { const uint32_t cb = br_match[v]; end = cb != 0xffffffffu ? br_row[cb] : r; }
    }
    out.row[j] = r;
    out.end[j] = end;
    out.koff[j] = __ldg(t.koff + r);
    out.klen[j] = __ldg(t.klen + r);
}

// a warp per key, a lane per byte
__global__ void child_keys_kernel(const uint8_t *__restrict__ arena, ChildList c, const uint32_t *__restrict__ key_off, uint32_t n, uint8_t *blob) {
    const uint32_t j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (j >= n) return;
    const uint32_t ko = c.koff[j], kl = c.klen[j], at = key_off[j];
    for (uint32_t x = lane; x < kl; x += 32) blob[at + x] = arena[ko + x];
}

__global__ void gather_rows_kernel(RowArrays from, RowArrays to, const RowSegment *__restrict__ segs, uint32_t n_segs, uint32_t n_new) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_new) return;
    uint32_t lo = 0, hi = n_segs;   // the last segment whose dst <= i
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(&segs[mid].dst) <= i) lo = mid; else hi = mid;
    }
    const uint32_t src = __ldg(&segs[lo].src) + (i - __ldg(&segs[lo].dst));
    to.meta[i] = from.meta[src]; to.koff[i] = from.koff[src]; to.klen[i] = from.klen[src];
    to.voff[i] = from.voff[src]; to.vlen[i] = from.vlen[src];
}

__global__ void patch_commas_kernel(uint32_t *meta, uint32_t row_set, uint32_t row_clear) {
    if (row_set != 0xffffffffu) meta[row_set] |= META_COMMA;
    if (row_clear != 0xffffffffu) meta[row_clear] &= ~META_COMMA;
}

}  // namespace

cudaError_t launch_child_flags(const TapeView &t, uint32_t module_row, uint32_t module_end, uint32_t *flags, cudaStream_t s) {
    if (!t.n_rows) return cudaSuccess;
    child_flags_kernel<<<(t.n_rows + 255) / 256, 256, 0, s>>>(t, module_row, module_end, flags);
    count_parse_launch();
    return cudaGetLastError();
}
cudaError_t launch_child_entries(const TapeView &t, const uint32_t *flags, const uint32_t *rank, const uint32_t *br_match, const uint32_t *br_row,
                                 ChildList out, cudaStream_t s) {
    if (!t.n_rows) return cudaSuccess;
    child_entries_kernel<<<(t.n_rows + 255) / 256, 256, 0, s>>>(t, flags, rank, br_match, br_row, out);
    count_parse_launch();
    return cudaGetLastError();
}
cudaError_t launch_child_keys(const uint8_t *arena, ChildList c, const uint32_t *key_off, uint32_t n_children, uint8_t *blob, cudaStream_t s) {
    if (!n_children) return cudaSuccess;
    child_keys_kernel<<<(n_children + 3) / 4, 128, 0, s>>>(arena, c, key_off, n_children, blob);
    count_parse_launch();
    return cudaGetLastError();
}
cudaError_t launch_gather_rows(RowArrays from, RowArrays to, const RowSegment *segs, uint32_t n_segs, uint32_t n_new, cudaStream_t s) {
    if (!n_new || !n_segs) return cudaSuccess;
    gather_rows_kernel<<<(n_new + 255) / 256, 256, 0, s>>>(from, to, segs, n_segs, n_new);
    count_parse_launch();
    return cudaGetLastError();
}
cudaError_t launch_patch_commas(uint32_t *meta, uint32_t row_set, uint32_t row_clear, cudaStream_t s) {
    if (row_set == 0xffffffffu && row_clear == 0xffffffffu) return cudaSuccess;
    patch_commas_kernel<<<1, 1, 0, s>>>(meta, row_set, row_clear);
    count_parse_launch();
    return cudaGetLastError();
}

}  // namespace tk
