// index.cu -- the module index: what State.Clusters() and State.Nodes() need from a document that
// lives on the device as the parser's line tape (rows in encoding/json's output order).
//
// state/state.go:94-147 scan every child of "module" for a key prefix and read one string field
// of each hit ("name" for clusters, "hostname" for nodes).  Here one pass over the rows at the end
// of the parse (field_find_kernel) lists, per field, the children that carry it -- (field value,
// child key) references in document order, which for a tape that stays on the device is bytewise
// key order -- and two more kernels turn a list into the finished C-ABI result records
//     { u32 field_len, field, u32 key_len, key }
// packed one after another, with one offset per record and a bitmap of the records whose field
// does not sort strictly after the record before.  The records go to the host once per document
// version (about 6.4 MB for the 101,001 children of config #4); each Clusters()/Nodes() call is
// then a binary search for the prefix range over them, and its answer a byte range of the blob
// provided the field values in it ascend strictly (the ABI returns pairs sorted by field, one per
// value).
#include "index.cuh"

#include <cstddef>

namespace tk {
namespace {

constexpr unsigned FULL = 0xffffffffu;
typedef unsigned long long u64;
constexpr u64 ST_AGG = 1ull << 62, ST_PREFIX = 2ull << 62, ST_MASK = (1ull << 62) - 1;
__device__ __forceinline__ u64 ld_relaxed(const u64 *p) {
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(u64 *p, u64 v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }

// ---- the field lists ---------------------------------------------------------------------------
// A row qualifies if it is a string member three levels down whose key reads "name" or "hostname";
// the object it sits in is the first row above it that is no deeper than level 2, and that row
// must open an object strictly inside "module".  A block takes 1,024 consecutive rows (and the 64
// before them, for the search upwards) and keeps their meta words in shared memory; every thread
// looks at four rows, a phase at a time -- key lengths, key offsets, key text, the child's and the
// value's references -- so that the loads of a phase are in flight together.  The place of a
// reference in its list, and of its record in the packed blob, is the number / the record bytes of
// the qualifying rows before it: summed per block and handed from block to block by decoupled
// look-back (one status word per field: count in the low 30 bits, bytes above), blocks taking
// their tile from a ticket counter so that a block only ever waits for blocks that have started.
constexpr int FF_THREADS = 256, FF_ITEMS = 4, FF_TILE = FF_THREADS * FF_ITEMS, FF_HALO = 64;
constexpr uint32_t FF_MAX_WALK = 4096;   // rows looked through for the child before giving up (index_slow)

__global__ void __launch_bounds__(FF_THREADS) field_find_kernel(TapeView t, const uint32_t *__restrict__ br_match, const uint32_t *__restrict__ br_row,
                                                               FieldIndexBuffers fx, ParseResult *res) {
    __shared__ uint32_t s_tile, s_module_row, s_module_end;
    __shared__ uint32_t s_meta[FF_HALO + FF_TILE];
    __shared__ u64 s_wc[FF_THREADS / 32], s_wb0[FF_THREADS / 32], s_wb1[FF_THREADS / 32];
    __shared__ u64 s_prefix[2];   // per field: records before this block | record bytes before it << 32
    const uint32_t n_rows = res->n_rows;
    if (threadIdx.x == 0) {
        s_tile = (uint32_t)atomicAdd(&fx.status[0], 1ull);
        // "module" must be an object member of a root object; its closing row through the bracket tables
        uint32_t first = 0xffffffffu, last = 0;
        if (res->invalid == 0 && res->root_kind == ROW_OPEN_OBJ && res->module_row != 0 && res->module_kind == ROW_OPEN_OBJ) {
            first = res->module_row - 1;
            const uint32_t ob = t.voff[first];               // voff of an opening row: its bracket
            const uint32_t cb = br_match[ob];
            last = cb != 0xffffffffu ? br_row[cb] : 0;
        }
        s_module_row = first;
        s_module_end = last;
    }
    __syncthreads();
    const uint32_t tile = s_tile, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t n_tiles = (n_rows + FF_TILE - 1) / FF_TILE;
    if (tile >= n_tiles) return;                             // (the grid covers a bound on the row count)
    const uint32_t module_row = s_module_row, module_end = s_module_end;
    const uint32_t tile0 = tile * FF_TILE;
    for (uint32_t x = threadIdx.x; x < FF_HALO + FF_TILE; x += FF_THREADS) {
        const long long r = (long long)tile0 - FF_HALO + x;
        s_meta[x] = r >= 0 && r < (long long)n_rows ? __ldg(t.meta + r) : 0u;   // (depth 0: stops any search upwards)
    }
    __syncthreads();
    const uint32_t *meta_at = s_meta + FF_HALO - tile0;       // meta_at[r] for rows in the window
    const uint32_t base = tile0 + threadIdx.x * FF_ITEMS;

    // phase 1: which rows can be field rows at all
    uint32_t kl[FF_ITEMS], ko[FF_ITEMS], fld[FF_ITEMS];
    bool cand[FF_ITEMS];
#pragma unroll
    for (int q = 0; q < FF_ITEMS; q++) {
        const uint32_t r = base + q;
        const bool in = r < n_rows && r > module_row && r < module_end;
        const uint32_t meta = in ? meta_at[r] : 0u;
        cand[q] = in && (meta >> META_DEPTH_SHIFT) == 3 && (meta & (15u | META_HAS_KEY)) == (ROW_STR | META_HAS_KEY);
        kl[q] = cand[q] ? __ldg(t.klen + r) : 0u;
    }
#pragma unroll
    for (int q = 0; q < FF_ITEMS; q++) {
        cand[q] = cand[q] && (kl[q] == 4 || kl[q] == 8);
        ko[q] = cand[q] ? __ldg(t.koff + base + q) : 0u;
    }
    // phase 2: the key text, as three aligned words that cover its 8 bytes
    uint32_t w[FF_ITEMS][3];
#pragma unroll
    for (int q = 0; q < FF_ITEMS; q++) {
        const uint32_t *p = reinterpret_cast<const uint32_t *>(t.arena + (ko[q] & ~3u));
#pragma unroll
        for (int k = 0; k < 3; k++) w[q][k] = cand[q] ? __ldg(p + k) : 0u;
    }
    // phase 3: the child (searched in shared memory; past the window, in global memory) and the references
    FieldRef ref[FF_ITEMS];
    uint32_t child[FF_ITEMS];
#pragma unroll
    for (int q = 0; q < FF_ITEMS; q++) {
        fld[q] = 2;
        child[q] = 0;
        if (!cand[q]) continue;
        const uint32_t sh = (ko[q] & 3u) * 8u;
        const uint32_t a = __byte_perm(__funnelshift_r(w[q][0], w[q][1], sh), 0, 0x0123), b = __byte_perm(__funnelshift_r(w[q][1], w[q][2], sh), 0, 0x0123);
        uint32_t f = 2;
        if (kl[q] == 4) { if (a == 0x6e616d65u) f = 0; }                          // "name"
        else if (a == 0x686f7374u && b == 0x6e616d65u) f = 1;                      // "hostname"
        if (f == 2) continue;
        const uint32_t r = base + q;
        uint32_t c = r, cm = 0;
        bool found = false;
        for (uint32_t step = 0; step < FF_MAX_WALK && c > module_row + 1; step++) {
            c--;
            cm = c + FF_HALO >= tile0 ? meta_at[c] : __ldg(t.meta + c);
            if ((cm >> META_DEPTH_SHIFT) <= 2) { found = true; break; }
        }
        if (!found) { if (c > module_row + 1) res->index_slow = 1; continue; }
        if ((cm >> META_DEPTH_SHIFT) != 2 || (cm & 15u) != ROW_OPEN_OBJ || !(cm & META_HAS_KEY)) continue;
        fld[q] = f;
        child[q] = c;
    }
    u64 cnt = 0, b0 = 0, b1 = 0;                               // count of field 0 | count of field 1 << 32; record bytes per field
#pragma unroll
    for (int q = 0; q < FF_ITEMS; q++) {
        const bool is = fld[q] != 2;
        const uint32_t r = base + q;
        ref[q].foff = is ? __ldg(t.voff + r) : 0u;
        ref[q].flen = is ? __ldg(t.vlen + r) : 0u;
        ref[q].koff = is ? __ldg(t.koff + child[q]) : 0u;
        ref[q].klen = is ? __ldg(t.klen + child[q]) : 0u;
    }
#pragma unroll
    for (int q = 0; q < FF_ITEMS; q++) {
        if (fld[q] == 2) continue;
        const u64 rec = 8ull + ref[q].flen + ref[q].klen;      // { u32 len, field, u32 len, key }
        if (fld[q]) { cnt += 1ull << 32; b1 += rec; } else { cnt += 1ull; b0 += rec; }
    }

    // exclusive scans over the block
    u64 ic = cnt, i0 = b0, i1 = b1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u64 xc = __shfl_up_sync(FULL, ic, d), x0 = __shfl_up_sync(FULL, i0, d), x1 = __shfl_up_sync(FULL, i1, d);
        if (lane >= (uint32_t)d) { ic += xc; i0 += x0; i1 += x1; }
    }
    if (lane == 31) { s_wc[warp] = ic; s_wb0[warp] = i0; s_wb1[warp] = i1; }
    __syncthreads();
    u64 wc = 0, w0 = 0, w1 = 0, tc = 0, t0 = 0, t1 = 0;
    for (uint32_t i = 0; i < FF_THREADS / 32; i++) {
        if (i < warp) { wc += s_wc[i]; w0 += s_wb0[i]; w1 += s_wb1[i]; }
        tc += s_wc[i]; t0 += s_wb0[i]; t1 += s_wb1[i];
    }
    if (warp < 2) {
        // warp 0 carries field 0, warp 1 field 1.  Payload: records (30 bits) | record bytes (32 bits).
        const uint32_t f = warp;
        const u64 agg = ((f ? tc >> 32 : tc) & 0x3fffffffull) | (((f ? t1 : t0) & 0xffffffffull) << 30);
        u64 *status = fx.status + 2 + (size_t)f * n_tiles;
        u64 prefix = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed(status, ST_PREFIX | agg);
        } else {
            if (lane == 0) st_relaxed(status + tile, ST_AGG | agg);
            int look = (int)tile - 1;
            for (;;) {
                const int idx = look - (int)lane;
                u64 st = idx >= 0 ? ld_relaxed(status + idx) : ST_PREFIX;
                while (__any_sync(FULL, (st >> 62) == 0)) { if ((st >> 62) == 0) st = ld_relaxed(status + idx); }
                const unsigned pm = __ballot_sync(FULL, (st >> 62) == 2);
                const int first = pm ? __ffs(pm) - 1 : 32;
                u64 x = ((int)lane <= first) ? (st & ST_MASK) : 0ull;
#pragma unroll
                for (int d = 16; d; d >>= 1) x += __shfl_xor_sync(FULL, x, d);
                prefix += x;
                if (pm) break;
                look -= 32;
            }
            if (lane == 0) st_relaxed(status + tile, ST_PREFIX | ((prefix + agg) & ST_MASK));
        }
        if (lane == 0) {
            const u64 records = prefix & 0x3fffffffull, bytes = (prefix >> 30) & 0xffffffffull;
            s_prefix[f] = records | (bytes << 32);
            const u64 block_bytes = f ? t1 : t0;
            if (block_bytes) atomicAdd(&res->field_bytes[f], block_bytes);     // (the 64-bit total: a blob of 4 GiB or more is refused)
            if (tile + 1 == n_tiles) {
                const uint32_t m = (uint32_t)((prefix + agg) & 0x3fffffffull);
                res->field_count[f] = m;
                if (m <= fx.cap[f]) fx.off[f][m] = (uint32_t)(((prefix + agg) >> 30) & 0xffffffffull);
            }
        }
    }
    __syncthreads();
    uint32_t rank[2] = {(uint32_t)s_prefix[0] + (uint32_t)(wc + ic - cnt), (uint32_t)s_prefix[1] + (uint32_t)((wc + ic - cnt) >> 32)};
    uint32_t at_byte[2] = {(uint32_t)(s_prefix[0] >> 32) + (uint32_t)(w0 + i0 - b0), (uint32_t)(s_prefix[1] >> 32) + (uint32_t)(w1 + i1 - b1)};
#pragma unroll
    for (int q = 0; q < FF_ITEMS; q++) {
        if (fld[q] == 2) continue;
        const uint32_t f = fld[q], at = rank[f]++;
        if (at < fx.cap[f]) { fx.refs[f][at] = ref[q]; fx.off[f][at] = at_byte[f]; }
        at_byte[f] += 8u + ref[q].flen + ref[q].klen;
    }
}

// ---- the packed records ------------------------------------------------------------------------
// One warp per PACK_RECORDS records of one field (both fields in one launch): the first lanes
// fetch one reference each and say whether their field sorts strictly after the record before
// (viol_bits: m bits per field, cleared beforehand), then the warp writes the records one after
// another, a lane per byte, so that the stores of a record are contiguous.  (Few records per warp:
// the copy loop is a chain of L2 round trips, and it is the number of warps in flight that hides
// them.)
constexpr uint32_t PACK_RECORDS = 8;
struct PackArgs {
    const FieldRef *refs[2];
    const uint32_t *off[2];
    uint32_t *viol_bits[2];
    uint8_t *blob[2];
    uint32_t m[2];
    uint32_t warps0;   // warps that work on field 0; the rest on field 1
};
__global__ void __launch_bounds__(128) field_pack_kernel(const uint8_t *__restrict__ arena, PackArgs a) {
    const uint32_t lane = threadIdx.x & 31;
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t f = warp >= a.warps0 ? 1u : 0u;
    if (f) warp -= a.warps0;
    const uint32_t m = a.m[f];
    const FieldRef *refs = a.refs[f];
    const uint32_t j = lane < PACK_RECORDS ? warp * PACK_RECORDS + lane : m;   // this lane's record (none for the other lanes)
    FieldRef c = {0, 0, 0, 0};
    uint32_t at = 0;
    if (j < m) { c = refs[j]; at = a.off[f][j]; }
    // field order against the record before: the lane below holds it, lane 0 fetches it
    {
        uint32_t po = __shfl_up_sync(FULL, c.foff, 1), pl = __shfl_up_sync(FULL, c.flen, 1);
        if (lane == 0 && j > 0 && j < m) { const FieldRef p = refs[j - 1]; po = p.foff; pl = p.flen; }
        if (j > 0 && j < m && compare_bytes(arena, po, pl, c.foff, c.flen) >= 0) atomicOr(a.viol_bits[f] + (j >> 5), 1u << (j & 31));
    }
    uint8_t *blob = a.blob[f];
    for (unsigned who = __ballot_sync(FULL, j < m); who; who &= who - 1) {
        const int src = __ffs(who) - 1;
        const uint32_t r_fo = __shfl_sync(FULL, c.foff, src), r_fl = __shfl_sync(FULL, c.flen, src);
        const uint32_t r_ko = __shfl_sync(FULL, c.koff, src), r_kl = __shfl_sync(FULL, c.klen, src);
        uint8_t *w = blob + __shfl_sync(FULL, at, src);
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

}  // namespace

size_t field_status_words(uint32_t n_rows) { return 2 + 2 * (size_t)((n_rows + FF_TILE - 1) / FF_TILE); }

cudaError_t launch_field_find(const TapeView &t, uint32_t n_rows_bound, const uint32_t *br_match, const uint32_t *br_row, FieldIndexBuffers fx,
                              ParseResult *res, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(fx.status, 0, field_status_words(n_rows_bound) * 8, s);   // ticket counter and look-back status words
    if (e != cudaSuccess) return e;
    // (a second run -- with larger lists -- starts the totals again)
    if ((e = cudaMemsetAsync(reinterpret_cast<char *>(res) + offsetof(ParseResult, field_count), 0,
                             sizeof(ParseResult) - offsetof(ParseResult, field_count), s)) != cudaSuccess) return e;
    if (n_rows_bound == 0) return cudaSuccess;
    field_find_kernel<<<(n_rows_bound + FF_TILE - 1) / FF_TILE, FF_THREADS, 0, s>>>(t, br_match, br_row, fx, res);
    count_parse_launch();
    return cudaGetLastError();
}

cudaError_t launch_field_pack(const uint8_t *arena, const FieldIndexBuffers &fx, const uint32_t m[2], uint32_t *const viol_bits[2],
                              uint8_t *const blob[2], cudaStream_t s) {
    PackArgs a;
    uint32_t warps[2];
    for (int f = 0; f < 2; f++) {
        a.refs[f] = fx.refs[f]; a.off[f] = fx.off[f]; a.viol_bits[f] = viol_bits[f]; a.blob[f] = blob[f]; a.m[f] = m[f];
        warps[f] = (m[f] + PACK_RECORDS - 1) / PACK_RECORDS;
    }
    a.warps0 = warps[0];
    if (warps[0] + warps[1] == 0) return cudaSuccess;
    field_pack_kernel<<<(warps[0] + warps[1] + 3) / 4, 128, 0, s>>>(arena, a);
    count_parse_launch();
    return cudaGetLastError();
}

}  // namespace tk
