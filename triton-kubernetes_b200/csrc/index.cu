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

// entry of every child: where its key, "name" and "hostname" strings are, and their total size
__global__ void child_entries_kernel(TapeView t, const uint32_t *__restrict__ flags, const uint32_t *__restrict__ rank,
                                     const uint32_t *__restrict__ br_match, const uint32_t *__restrict__ br_row, ChildRef *refs,
                                     uint32_t *sizes) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= t.n_rows || !flags[r]) return;
    ChildRef c;
    c.koff = t.koff[r]; c.klen = t.klen[r];
    c.noff = c.hoff = 0;
    c.nlen = c.hlen = 0xffffffffu;  // absent, or not a string
    if ((t.meta[r] & 15u) == ROW_OPEN_OBJ) {
        const uint32_t end = br_row[br_match[t.voff[r]]];
        for (uint32_t m = r + 1; m < end; m++) {
            const uint32_t meta = t.meta[m];
            if (row_depth(meta) != 3 || !(meta & META_HAS_KEY) || (meta & 15u) != ROW_STR) continue;
            const uint32_t ko = t.koff[m], kl = t.klen[m];
            if (key_is(t.arena, ko, kl, "name", 4)) { c.noff = t.voff[m]; c.nlen = t.vlen[m]; }
            else if (key_is(t.arena, ko, kl, "hostname", 8)) { c.hoff = t.voff[m]; c.hlen = t.vlen[m]; }
        }
    }
    const uint32_t i = rank[r];
    refs[i] = c;
    sizes[i] = c.klen + (c.nlen == 0xffffffffu ? 0 : c.nlen) + (c.hlen == 0xffffffffu ? 0 : c.hlen);
}

// strings of child i copied to packed[off[i] ...): key, then name, then hostname
__global__ void child_pack_kernel(const uint8_t *__restrict__ arena, const ChildRef *__restrict__ refs, const uint32_t *__restrict__ off,
                                  uint32_t n, ChildEntry *entries, uint8_t *packed) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const ChildRef c = refs[i];
    uint32_t o = off[i];
    ChildEntry e;
    e.off = o; e.klen = c.klen; e.nlen = c.nlen; e.hlen = c.hlen;
    entries[i] = e;
    for (uint32_t k = 0; k < c.klen; k++) packed[o + k] = arena[c.koff + k];
    o += c.klen;
    if (c.nlen != 0xffffffffu) { for (uint32_t k = 0; k < c.nlen; k++) packed[o + k] = arena[c.noff + k]; o += c.nlen; }
    if (c.hlen != 0xffffffffu) { for (uint32_t k = 0; k < c.hlen; k++) packed[o + k] = arena[c.hoff + k]; }
}

// ---- result blobs: Clusters()/Nodes() answers laid out ahead of time -------------------------------
// For one field ("name" or "hostname"), every child that has it contributes the record the C ABI
// returns for it -- { u32 field_len, field, u32 key_len, key } -- and the records are packed in
// child (= key) order.  The answer to a prefix query is then the byte range of a run of children,
// provided the field values in that run ascend strictly (the ABI returns pairs sorted by field,
// one per value); `viol` counts the places where they do not, so the host can tell in O(1).

__device__ __forceinline__ bool field_of(const ChildRef &c, int which, uint32_t *off, uint32_t *len) {
    *off = which ? c.hoff : c.noff;
    *len = which ? c.hlen : c.nlen;
    return *len != 0xffffffffu;
}

// per child: (has field) << 32 | record bytes, and the order-violation flag
__global__ void field_sizes_kernel(const uint8_t *__restrict__ arena, const ChildRef *__restrict__ refs, uint32_t n, int which,
                                   unsigned long long *cs, uint32_t *viol) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t fo, fl;
    const ChildRef c = refs[i];
    const bool has = field_of(c, which, &fo, &fl);
    cs[i] = has ? ((1ull << 32) | (8ull + fl + c.klen)) : 0ull;
    uint32_t v = 0;
    if (has && i > 0) {
        uint32_t po, pl;
        const ChildRef p = refs[i - 1];
        v = 1;  // a predecessor without the field, or with a field value that is not smaller
        if (field_of(p, which, &po, &pl)) {
            const uint32_t m = min(pl, fl);
            int cmp = 0;
            for (uint32_t k = 0; k < m; k++) {
                const int x = arena[po + k], y = arena[fo + k];
                if (x != y) { cmp = x < y ? -1 : 1; break; }
            }
            if (cmp == 0) cmp = pl < fl ? -1 : (pl > fl ? 1 : 0);
            if (cmp < 0) v = 0;
        }
    }
    viol[i] = v;
}

__device__ __forceinline__ void put_u32(uint8_t *p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); }

// records written at their offsets; off[i] / cnt[i] = bytes / records before child i (n + 1 entries)
__global__ void field_pack_kernel(const uint8_t *__restrict__ arena, const ChildRef *__restrict__ refs, uint32_t n, int which,
                                  const unsigned long long *__restrict__ cs_excl, const unsigned long long *__restrict__ cs, uint8_t *blob,
                                  uint32_t *off, uint32_t *cnt) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long e = cs_excl[i];
    const uint32_t o = (uint32_t)e;
    off[i] = o;
    cnt[i] = (uint32_t)(e >> 32);
    if (i + 1 == n) {
        const unsigned long long t = e + cs[i];
        off[n] = (uint32_t)t;
        cnt[n] = (uint32_t)(t >> 32);
    }
    uint32_t fo, fl;
    const ChildRef c = refs[i];
    if (!field_of(c, which, &fo, &fl)) return;
    uint8_t *w = blob + o;
    put_u32(w, fl);
    for (uint32_t k = 0; k < fl; k++) w[4 + k] = arena[fo + k];
    put_u32(w + 4 + fl, c.klen);
    for (uint32_t k = 0; k < c.klen; k++) w[8 + fl + k] = arena[c.koff + k];
}

}  // namespace

cudaError_t launch_field_sizes(const uint8_t *arena, const ChildRef *refs, uint32_t n, int which, unsigned long long *cs, uint32_t *viol,
                               cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    field_sizes_kernel<<<(n + 127) / 128, 128, 0, s>>>(arena, refs, n, which, cs, viol);
    count_parse_launch();
    return cudaGetLastError();
}
cudaError_t launch_field_pack(const uint8_t *arena, const ChildRef *refs, uint32_t n, int which, const unsigned long long *cs_excl,
                              const unsigned long long *cs, uint8_t *blob, uint32_t *off, uint32_t *cnt, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    field_pack_kernel<<<(n + 127) / 128, 128, 0, s>>>(arena, refs, n, which, cs_excl, cs, blob, off, cnt);
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

cudaError_t launch_child_entries(const TapeView &t, const uint32_t *flags, const uint32_t *rank, const uint32_t *br_match,
                                 const uint32_t *br_row, ChildRef *refs, uint32_t *sizes, cudaStream_t s) {
    child_entries_kernel<<<(t.n_rows + 255) / 256, 256, 0, s>>>(t, flags, rank, br_match, br_row, refs, sizes);
    count_parse_launch();
    return cudaGetLastError();
}

cudaError_t launch_child_pack(const uint8_t *arena, const ChildRef *refs, const uint32_t *off, uint32_t n, ChildEntry *entries,
                              uint8_t *packed, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    child_pack_kernel<<<(n + 127) / 128, 128, 0, s>>>(arena, refs, off, n, entries, packed);
    count_parse_launch();
    return cudaGetLastError();
}

}  // namespace tk
