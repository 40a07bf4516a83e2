// index.cu -- the module index: what State.Clusters() and State.Nodes() need from a document that
// lives on the device as the parser's line tape (rows in encoding/json's output order).
//
// state/state.go:94-147 scan every child of "module" for a key prefix and read one string field
// of each hit ("name" for clusters, "hostname" for nodes).  Here one bulk pass per document
// version classifies all module children at once and packs, per field, the result records of the
// children that have it (about 6.4 MB for the 101,001 children of config #4).  The records go to
// the host once; because the children are in bytewise key order, each Clusters()/Nodes() call is
// then a binary search for the prefix range over them instead of a 3.5 MB scan and a kernel
// launch per call.
#include "index.cuh"

namespace tk {
namespace {

__device__ __forceinline__ uint32_t row_depth(uint32_t meta) { return meta >> META_DEPTH_SHIFT; }
__device__ __forceinline__ bool key_is(const uint8_t *arena, uint32_t off, uint32_t len, const char *s, uint32_t n) {
    if (len != n) return false;
    for (uint32_t i = 0; i < n; i++)
        if (arena[off + i] != (uint8_t)s[i]) return false;
    return true;
}

// Where every child's key, "name" and "hostname" strings are.  One thread per ROW: a child's own
// row records the key, and each depth-3 string member called name / hostname records itself in
// the slot of the child it belongs to -- the last child row before it, i.e. slot rank[r] - 1
// (rank comes from scan_child_rank, parse.cu).  refs must have been filled with 0xff bytes (absent
// fields read as length 0xffffffff).
__global__ void child_refs_kernel(TapeView t, uint32_t module_row, const uint32_t *__restrict__ br_match, const uint32_t *__restrict__ br_row,
                                  const uint32_t *__restrict__ rank, ChildRef *refs) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= t.n_rows || r <= module_row) return;
    // independent loads first, so that they are in flight together (the kernel is a chain of
    // dependent look-ups otherwise: meta -> key length -> key bytes -> rank)
    const uint32_t meta = t.meta[r], kl = t.klen[r], rk = rank[r];
    const uint32_t module_end = closing_row(t, module_row, br_match, br_row);
    if (r >= module_end) return;
    if (starts_child(meta, r, module_row, module_end)) {
        ChildRef *c = refs + rk;
        c->koff = t.koff[r];
        c->klen = kl;
        return;
    }
    if (row_depth(meta) != 3 || !(meta & META_HAS_KEY) || (meta & 15u) != ROW_STR || (kl != 4 && kl != 8)) return;
    const uint32_t ko = t.koff[r];
    ChildRef *c = refs + (rk - 1);
    if (key_is(t.arena, ko, kl, "name", 4)) { c->noff = t.voff[r]; c->nlen = t.vlen[r]; }
    else if (key_is(t.arena, ko, kl, "hostname", 8)) { c->hoff = t.voff[r]; c->hlen = t.vlen[r]; }
}

// ---- what goes to the host -------------------------------------------------------------------
// Per field ("name" for Clusters, "hostname" for Nodes) the finished C-ABI result records of the
// children that have it -- { u32 field_len, field, u32 key_len, key } -- packed in child (= bytewise
// key) order, with one offset per record.  The host finds the run of records whose key starts with
// the prefix by binary search over the records themselves, and the answer to the query is the byte
// range of that run provided the field values in it ascend strictly (the ABI returns pairs sorted
// by field, one per value); a bitmap marks the records where they do not.

__device__ __forceinline__ bool field_of(const ChildRef &c, int which, uint32_t *off, uint32_t *len) {
    *off = which ? c.hoff : c.noff;
    *len = which ? c.hlen : c.nlen;
    return *len != 0xffffffffu;
}

__device__ __forceinline__ uint32_t order_violation(const uint8_t *arena, const ChildRef &p, const ChildRef &c, int which) {
    uint32_t fo, fl, po, pl;
    if (!field_of(c, which, &fo, &fl)) return 0;
    if (!field_of(p, which, &po, &pl)) return 1;  // a predecessor without the field: not a clean run
    return compare_bytes(arena, po, pl, fo, fl) < 0 ? 0u : 1u;
}

// per child: (has field) << 32 | record bytes for both fields, and the violation flags
__global__ void child_sizes_kernel(const uint8_t *__restrict__ arena, const ChildRef *__restrict__ refs, uint32_t n,
                                   unsigned long long *rec_n, unsigned long long *rec_h, uint32_t *viol_n, uint32_t *viol_h) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const ChildRef c = refs[i];
    rec_n[i] = c.nlen != 0xffffffffu ? ((1ull << 32) | (8ull + c.nlen + c.klen)) : 0ull;
    rec_h[i] = c.hlen != 0xffffffffu ? ((1ull << 32) | (8ull + c.hlen + c.klen)) : 0ull;
    uint32_t vn = 0, vh = 0;
    if (i > 0) {
        const ChildRef p = refs[i - 1];
        vn = order_violation(arena, p, c, 0);
        vh = order_violation(arena, p, c, 1);
    }
    viol_n[i] = vn;
    viol_h[i] = vh;
}

// Records written at their offsets.  Per field the children that have it form a compact list (in
// child = key order): off[j] is where record j starts (m + 1 entries), bit j of viol_bits says that
// record j does not sort strictly after the child before it.  The last child also leaves the
// totals (records << 32 | bytes) for the host.  viol_bits must have been cleared.
// One warp per PACK_CHILDREN children: the first lanes fetch one child's references each, then the
// warp writes the records one after another, a lane per byte, so that the stores of a record are
// contiguous.  (Few children per warp: the copy loop is a chain of L2 round trips, and it is the
// number of warps in flight that hides them.)
constexpr uint32_t PACK_CHILDREN = 8;
__global__ void __launch_bounds__(128) child_pack_kernel(const uint8_t *__restrict__ arena, const ChildRef *__restrict__ refs, uint32_t n, IndexTables tb) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t i = lane < PACK_CHILDREN ? warp * PACK_CHILDREN + lane : n;   // this lane's child (none for the other lanes)
    ChildRef c;
    c.koff = 0; c.klen = 0; c.noff = 0; c.nlen = 0xffffffffu; c.hoff = 0; c.hlen = 0xffffffffu;
    if (i < n) c = refs[i];
#pragma unroll
    for (int f = 0; f < 2; f++) {
        uint32_t fo, fl;
        const bool has = i < n && field_of(c, f, &fo, &fl);
        uint32_t byte_off = 0;
        if (i < n) {
            const unsigned long long ex = tb.rec_excl[f][i];
            byte_off = (uint32_t)ex;
            const uint32_t j = (uint32_t)(ex >> 32);
            if (has) {
                tb.off[f][j] = byte_off;
                if (tb.viol[f][i]) atomicOr(tb.viol_bits[f] + (j >> 5), 1u << (j & 31));
            }
            if (i + 1 == n) {
                const unsigned long long total = ex + tb.rec[f][i];
                tb.off[f][(uint32_t)(total >> 32)] = (uint32_t)total;
                tb.totals[f] = total;
            }
        }
        const unsigned who = __ballot_sync(0xffffffffu, has);
        for (unsigned m = who; m; m &= m - 1) {
            const int src = __ffs(m) - 1;
            const uint32_t r_fo = __shfl_sync(0xffffffffu, fo, src), r_fl = __shfl_sync(0xffffffffu, fl, src);
            const uint32_t r_ko = __shfl_sync(0xffffffffu, c.koff, src), r_kl = __shfl_sync(0xffffffffu, c.klen, src);
            uint8_t *w = tb.blob[f] + __shfl_sync(0xffffffffu, byte_off, src);
            // { u32 field_len, field, u32 key_len, key }
            const uint32_t total = 8 + r_fl + r_kl;
            for (uint32_t x = lane; x < total; x += 32) {
                uint32_t v;
                if (x < 4) v = r_fl >> (8 * x);
                else if (x < 4 + r_fl) v = arena[r_fo + x - 4];
                else if (x < 8 + r_fl) v = r_kl >> (8 * (x - 4 - r_fl));
                else v = arena[r_ko + x - 8 - r_fl];
                w[x] = (uint8_t)v;
            }
        }
    }
}

}  // namespace

cudaError_t launch_child_refs(const TapeView &t, uint32_t module_row, const uint32_t *br_match, const uint32_t *br_row, const uint32_t *rank,
                              ChildRef *refs, cudaStream_t s) {
    child_refs_kernel<<<(t.n_rows + 255) / 256, 256, 0, s>>>(t, module_row, br_match, br_row, rank, refs);
    count_parse_launch();
    return cudaGetLastError();
}
cudaError_t launch_child_sizes(const uint8_t *arena, const ChildRef *refs, uint32_t n, unsigned long long *rec_n,
                               unsigned long long *rec_h, uint32_t *viol_n, uint32_t *viol_h, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    child_sizes_kernel<<<(n + 127) / 128, 128, 0, s>>>(arena, refs, n, rec_n, rec_h, viol_n, viol_h);
    count_parse_launch();
    return cudaGetLastError();
}
cudaError_t launch_child_pack(const uint8_t *arena, const ChildRef *refs, uint32_t n, const IndexTables &tb, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    const uint32_t warps = (n + PACK_CHILDREN - 1) / PACK_CHILDREN;
    child_pack_kernel<<<(warps + 3) / 4, 128, 0, s>>>(arena, refs, n, tb);
    count_parse_launch();
    return cudaGetLastError();
}

}  // namespace tk
