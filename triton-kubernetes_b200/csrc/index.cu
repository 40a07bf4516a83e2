// index.cu -- the module index: what State.Clusters() and State.Nodes() need from a document that
// lives on the device as the parser's line tape (rows in encoding/json's output order).
//
// state/state.go:94-147 scan every child of "module" for a key prefix and read one string field
// of each hit ("name" for clusters, "hostname" for nodes).  Here one bulk pass per document
// version classifies all module children at once and packs, per child, its key and those two
// fields into a small arena (about 5 MB for the 101,001 children of config #4).  The arena goes
// to the host once; because the children are in bytewise key order, each Clusters()/Nodes() call
// is then a binary search for the prefix range over that packed index instead of a 3.5 MB scan
// and a kernel launch per call.
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

// the "module" member of the root object and the row that closes it
__global__ void find_module_kernel(TapeView t, const uint32_t *__restrict__ br_match, const uint32_t *__restrict__ br_row, ModuleInfo *info) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= t.n_rows) return;
    const uint32_t meta = t.meta[r];
    if (r == 0) info->root_kind = meta & 15u;
    if (row_depth(meta) != 1 || !(meta & META_HAS_KEY)) return;
    if (!key_is(t.arena, t.koff[r], t.klen[r], "module", 6)) return;
    const uint32_t kind = meta & 15u;
    info->module_kind = kind;
    info->module_row = r;
    if (kind == ROW_OPEN_OBJ) info->module_end = br_row[br_match[t.voff[r]]];  // voff of an opening row: its bracket
    else info->module_end = r;
}

// 1 for every row that starts a child of "module"
__global__ void child_flags_kernel(TapeView t, const ModuleInfo *__restrict__ info, uint32_t *__restrict__ flags) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= t.n_rows) return;
    uint32_t f = 0;
    if (info->module_kind == ROW_OPEN_OBJ && r > info->module_row && r < info->module_end) {
        const uint32_t meta = t.meta[r], kind = meta & 15u;
        f = (row_depth(meta) == 2 && kind != ROW_CLOSE_OBJ && kind != ROW_CLOSE_ARR) ? 1u : 0u;
    }
    flags[r] = f;
}

// Where every child's key, "name" and "hostname" strings are.  One thread per ROW: a child's own
// row records the key, and each depth-3 string member called name / hostname records itself in
// the slot of the child it belongs to -- the last child row before it, i.e. slot rank[r] - 1.
// refs must have been filled with 0xff bytes (absent fields read as length 0xffffffff).
__global__ void child_refs_kernel(TapeView t, const ModuleInfo *__restrict__ info, const uint32_t *__restrict__ flags,
                                  const uint32_t *__restrict__ rank, ChildRef *refs) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= t.n_rows || info->module_kind != ROW_OPEN_OBJ || r <= info->module_row || r >= info->module_end) return;
    const uint32_t meta = t.meta[r];
    if (flags[r]) {
        ChildRef *c = refs + rank[r];
        c->koff = t.koff[r];
        c->klen = t.klen[r];
        return;
    }
    if (row_depth(meta) != 3 || !(meta & META_HAS_KEY) || (meta & 15u) != ROW_STR) return;
    const uint32_t ko = t.koff[r], kl = t.klen[r];
    ChildRef *c = refs + (rank[r] - 1);
    if (key_is(t.arena, ko, kl, "name", 4)) { c->noff = t.voff[r]; c->nlen = t.vlen[r]; }
    else if (key_is(t.arena, ko, kl, "hostname", 8)) { c->hoff = t.voff[r]; c->hlen = t.vlen[r]; }
}

// ---- what goes to the host -------------------------------------------------------------------
// (1) the keys, packed in child (= bytewise key) order, for the prefix search;
// (2) per field ("name" for Clusters, "hostname" for Nodes) the finished C-ABI result records of
//     the children that have it -- { u32 field_len, field, u32 key_len, key } -- packed in child
//     order, so that the answer to a prefix query is the byte range of a run of children provided
//     the field values in the run ascend strictly (the ABI returns pairs sorted by field, one per
//     value); `viol` marks the places where they do not, so the host can tell in O(1).

__device__ __forceinline__ bool field_of(const ChildRef &c, int which, uint32_t *off, uint32_t *len) {
    *off = which ? c.hoff : c.noff;
    *len = which ? c.hlen : c.nlen;
    return *len != 0xffffffffu;
}

__device__ __forceinline__ uint32_t order_violation(const uint8_t *arena, const ChildRef &p, const ChildRef &c, int which) {
    uint32_t fo, fl, po, pl;
    if (!field_of(c, which, &fo, &fl)) return 0;
    if (!field_of(p, which, &po, &pl)) return 1;  // a predecessor without the field: not a clean run
    const uint32_t m = min(pl, fl);
    for (uint32_t k = 0; k < m; k++) {
        const int x = arena[po + k], y = arena[fo + k];
        if (x != y) return x < y ? 0u : 1u;
    }
    return pl < fl ? 0u : 1u;
}

// per child: key bytes, (has field) << 32 | record bytes for both fields, and the violation flags
__global__ void child_sizes_kernel(const uint8_t *__restrict__ arena, const ChildRef *__restrict__ refs, uint32_t n, uint32_t *key_bytes,
                                   unsigned long long *rec_n, unsigned long long *rec_h, uint32_t *viol_n, uint32_t *viol_h) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const ChildRef c = refs[i];
    key_bytes[i] = c.klen;
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

__device__ __forceinline__ void put_u32(uint8_t *p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); }

__device__ __forceinline__ void put_record(uint8_t *w, const uint8_t *arena, uint32_t fo, uint32_t fl, const ChildRef &c) {
    put_u32(w, fl);
    for (uint32_t k = 0; k < fl; k++) w[4 + k] = arena[fo + k];
    put_u32(w + 4 + fl, c.klen);
    for (uint32_t k = 0; k < c.klen; k++) w[8 + fl + k] = arena[c.koff + k];
}

// keys and records written at their offsets; tables get their n + 1 entries
__global__ void child_pack_kernel(const uint8_t *__restrict__ arena, const ChildRef *__restrict__ refs, uint32_t n, IndexTables tb) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const ChildRef c = refs[i];
    const uint32_t ko = tb.key_excl[i];
    const unsigned long long en = tb.rec_excl[0][i], eh = tb.rec_excl[1][i];
    tb.key_off[i] = ko;
    tb.off[0][i] = (uint32_t)en; tb.cnt[0][i] = (uint32_t)(en >> 32);
    tb.off[1][i] = (uint32_t)eh; tb.cnt[1][i] = (uint32_t)(eh >> 32);
    if (i + 1 == n) {
        tb.key_off[n] = ko + c.klen;
        const unsigned long long tn = en + tb.rec[0][i], th = eh + tb.rec[1][i];
        tb.off[0][n] = (uint32_t)tn; tb.cnt[0][n] = (uint32_t)(tn >> 32);
        tb.off[1][n] = (uint32_t)th; tb.cnt[1][n] = (uint32_t)(th >> 32);
    }
    for (uint32_t k = 0; k < c.klen; k++) tb.keys[ko + k] = arena[c.koff + k];
    if (c.nlen != 0xffffffffu) put_record(tb.blob[0] + (uint32_t)en, arena, c.noff, c.nlen, c);
    if (c.hlen != 0xffffffffu) put_record(tb.blob[1] + (uint32_t)eh, arena, c.hoff, c.hlen, c);
}

}  // namespace

cudaError_t launch_child_refs(const TapeView &t, const ModuleInfo *info, const uint32_t *flags, const uint32_t *rank, ChildRef *refs,
                              cudaStream_t s) {
    child_refs_kernel<<<(t.n_rows + 255) / 256, 256, 0, s>>>(t, info, flags, rank, refs);
    count_parse_launch();
    return cudaGetLastError();
}
cudaError_t launch_child_sizes(const uint8_t *arena, const ChildRef *refs, uint32_t n, uint32_t *key_bytes, unsigned long long *rec_n,
                               unsigned long long *rec_h, uint32_t *viol_n, uint32_t *viol_h, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    child_sizes_kernel<<<(n + 127) / 128, 128, 0, s>>>(arena, refs, n, key_bytes, rec_n, rec_h, viol_n, viol_h);
    count_parse_launch();
    return cudaGetLastError();
}
cudaError_t launch_child_pack(const uint8_t *arena, const ChildRef *refs, uint32_t n, const IndexTables &tb, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    child_pack_kernel<<<(n + 127) / 128, 128, 0, s>>>(arena, refs, n, tb);
    count_parse_launch();
    return cudaGetLastError();
}

cudaError_t launch_find_module(const TapeView &t, const uint32_t *br_match, const uint32_t *br_row, ModuleInfo *info, cudaStream_t s) {
    ModuleInfo init;
    init.root_kind = 0xff; init.module_kind = 0xff; init.module_row = 0; init.module_end = 0; init.n_children = 0; init.pad = 0;
    cudaError_t e = cudaMemcpyAsync(info, &init, sizeof init, cudaMemcpyHostToDevice, s);  // 24 bytes from pageable memory: staged by the driver
    if (e != cudaSuccess) return e;
    find_module_kernel<<<(t.n_rows + 255) / 256, 256, 0, s>>>(t, br_match, br_row, info);
    count_parse_launch();
    return cudaGetLastError();
}

cudaError_t launch_child_flags(const TapeView &t, const ModuleInfo *info, uint32_t *flags, cudaStream_t s) {
    child_flags_kernel<<<(t.n_rows + 255) / 256, 256, 0, s>>>(t, info, flags);
    count_parse_launch();
    return cudaGetLastError();
}

}  // namespace tk
