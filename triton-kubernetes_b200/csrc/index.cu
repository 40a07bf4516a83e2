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
// A row qualifies if key_order_kernel marked it (a string member three levels down whose key reads
// "name" or "hostname"); the object it sits in is the first row above it that is no deeper than
// level 2, and that row must open an object strictly inside "module".  A block takes 4,096
// consecutive rows (and the 64 before them, for the search upwards) and brings their meta words
// into shared memory; the few qualifying rows then fetch their value's and their child's key
// reference.  The place of a reference in its list, and of its record in the packed blob, is the
// number / the record bytes of the qualifying rows before it: summed per block and handed from
// block to block by decoupled look-back (one status word per field: count in the low 30 bits,
// bytes above), blocks taking their tile from a ticket counter so that a block only ever waits
// for blocks that have started.  (4,096 rows per block: the 344 blocks of config #4 are all
// resident at once, so the kernel is one wave.)
constexpr int FF_THREADS = 512, FF_ITEMS = 8, FF_TILE = FF_THREADS * FF_ITEMS, FF_HALO = 64;
constexpr uint32_t FF_MAX_WALK = 4096;   // rows looked through for the child before giving up (index_slow)

__global__ void __launch_bounds__(FF_THREADS, 3) field_find_kernel(TapeView t, uint32_t n_rows_bound, FieldIndexBuffers fx, ParseResult *res) {
    const uint32_t n_rows = min(n_rows_bound, res->n_rows);   // (the real count is on the device; the launch may only have a bound)
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_meta[FF_HALO + FF_TILE];
    __shared__ uint16_t s_up[FF_TILE];            // field rows: rows between the row and its child (at most FF_MAX_WALK)
    __shared__ u64 s_wc[FF_THREADS / 32], s_wb0[FF_THREADS / 32], s_wb1[FF_THREADS / 32];
    __shared__ u64 s_prefix[2];   // per field: records before this block | record bytes before it << 32
    if (threadIdx.x == 0) s_tile = (uint32_t)atomicAdd(&fx.status[0], 1ull);
    // where "module" is: an object member of a root object (rows_kernel and key_order_kernel left it in *res)
    uint32_t module_row = 0xffffffffu, module_end = 0;
    if (res->invalid == 0 && res->root_kind == ROW_OPEN_OBJ && res->module_row != 0 && res->module_kind == ROW_OPEN_OBJ) {
        module_row = res->module_row - 1;
        module_end = res->module_end;
    }
    __syncthreads();
    const uint32_t tile = s_tile, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t n_tiles = (n_rows + FF_TILE - 1) / FF_TILE;
    if (tile >= n_tiles) return;                             // (the grid may cover a bound on the row count)
    const uint32_t tile0 = tile * FF_TILE;
    for (uint32_t x = threadIdx.x; x < FF_HALO + FF_TILE; x += FF_THREADS) {
        const long long r = (long long)tile0 - FF_HALO + x;
        s_meta[x] = r >= 0 && r < (long long)n_rows ? __ldg(t.meta + r) : 0u;   // (depth 0: stops any search upwards)
    }
    __syncthreads();
    const uint32_t base = tile0 + threadIdx.x * FF_ITEMS, sbase = FF_HALO + threadIdx.x * FF_ITEMS;   // rows, and their place in the window

    // per row: 0 / 1 the field, 2 not a field row (two bits each in fbits); how far up the child's row is
    uint32_t fbits = 0;
#pragma unroll
    for (int i = 0; i < FF_ITEMS; i++) {
        const uint32_t r = base + i, meta = s_meta[sbase + i];
        uint32_t f = 2;
        if (r < n_rows && r > module_row && r < module_end && (meta & (META_KEY_IS_NAME | META_KEY_IS_HOSTNAME))) {
            f = (meta & META_KEY_IS_HOSTNAME) ? 1u : 0u;
            // the child: searched upwards in the window; past it, in global memory
            uint32_t c = r, cm = 0;
            bool found = false;
            for (uint32_t step = 0; step < FF_MAX_WALK && c > module_row + 1; step++) {
                c--;
                cm = c + FF_HALO >= tile0 ? s_meta[c + FF_HALO - tile0] : __ldg(t.meta + c);
                if ((cm >> META_DEPTH_SHIFT) <= 2) { found = true; break; }
            }
            if (!found) { if (c > module_row + 1) res->index_slow = 1; f = 2; }
            else if ((cm >> META_DEPTH_SHIFT) != 2 || (cm & 15u) != ROW_OPEN_OBJ || !(cm & META_HAS_KEY)) f = 2;
            s_up[threadIdx.x * FF_ITEMS + i] = (uint16_t)(r - c);
        }
        fbits |= f << (2 * i);
    }
    // (a thread has a field row or two at most: their four loads go out together, row after row)
    auto load_ref = [&](int i) {
        const uint32_t r = base + i, c = r - s_up[threadIdx.x * FF_ITEMS + i];
        return FieldRef{__ldg(t.voff + r), __ldg(t.vlen + r), __ldg(t.koff + c), __ldg(t.klen + c)};
    };
    u64 cnt = 0, b0 = 0, b1 = 0;                               // count of field 0 | count of field 1 << 32; record bytes per field
#pragma unroll
    for (int i = 0; i < FF_ITEMS; i++) {
        const uint32_t f = (fbits >> (2 * i)) & 3u;
        if (f == 2) continue;
        const FieldRef ref = load_ref(i);
        const u64 rec = 8ull + ref.flen + ref.klen;            // { u32 len, field, u32 len, key }
        if (f) { cnt += 1ull << 32; b1 += rec; } else { cnt += 1ull; b0 += rec; }
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
#pragma unroll
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
    uint32_t rank0 = (uint32_t)s_prefix[0] + (uint32_t)(wc + ic - cnt), rank1 = (uint32_t)s_prefix[1] + (uint32_t)((wc + ic - cnt) >> 32);
    uint32_t byte0 = (uint32_t)(s_prefix[0] >> 32) + (uint32_t)(w0 + i0 - b0), byte1 = (uint32_t)(s_prefix[1] >> 32) + (uint32_t)(w1 + i1 - b1);
#pragma unroll
    for (int i = 0; i < FF_ITEMS; i++) {
        const uint32_t f = (fbits >> (2 * i)) & 3u;
        if (f == 2) continue;
        const FieldRef ref = load_ref(i);
        const uint32_t rec = 8u + ref.flen + ref.klen;
        const uint32_t at = f ? rank1 : rank0, at_byte = f ? byte1 : byte0;
        if (f) { rank1++; byte1 += rec; } else { rank0++; byte0 += rec; }
        if (at < fx.cap[f]) { fx.refs[f][at] = ref; fx.off[f][at] = at_byte; }
    }
}

// ---- the packed records ------------------------------------------------------------------------
// A block takes PACK_RECORDS consecutive records of one field (both fields in one launch).  Every
// thread builds its record -- { u32 len, field, u32 len, key } -- in shared memory at the place it
// has among the block's records, reading the two strings as aligned words, and says whether its
// field sorts strictly after the record before (viol_bits: m bits per field, cleared beforehand);
// the block then copies the finished stretch to the blob with coalesced stores.  Records too long
// for the staging area (PACK_STAGE bytes for the block) are written straight to the blob.
constexpr uint32_t PACK_RECORDS = 64, PACK_STAGE = 8192;
struct PackArgs {
    const FieldRef *refs[2];
    const uint32_t *off[2];
    uint32_t *viol_bits[2];
    uint8_t *blob[2];
    uint32_t m[2];
    uint32_t blocks0;   // blocks that work on field 0; the rest on field 1
};
__device__ __forceinline__ void put_u32(uint8_t *dst, uint32_t v) { dst[0] = (uint8_t)v; dst[1] = (uint8_t)(v >> 8); dst[2] = (uint8_t)(v >> 16); dst[3] = (uint8_t)(v >> 24); }
// n bytes of the arena from offset off to dst (any alignment on both sides), read as aligned words,
// COPY_BATCH words at a time so that the loads of a batch are in flight together
constexpr int COPY_BATCH = 12;
__device__ __forceinline__ void copy_string(uint8_t *dst, const uint8_t *__restrict__ arena, uint32_t off, uint32_t n) {
    const uint32_t *wp = reinterpret_cast<const uint32_t *>(arena + (off & ~3u));
    const uint32_t sh = (off & 3u) * 8u;
    for (uint32_t i0 = 0; i0 < n; i0 += 4 * COPY_BATCH) {
        uint32_t v[COPY_BATCH + 1];
#pragma unroll
        for (int k = 0; k <= COPY_BATCH; k++) v[k] = i0 + 4 * k < n + 4 ? __ldg(wp + (i0 >> 2) + k) : 0u;   // (every arena is readable 7 bytes past a string)
#pragma unroll
        for (int k = 0; k < COPY_BATCH; k++) {
            const uint32_t i = i0 + 4 * k;
            if (i >= n) break;
            const uint32_t x = __funnelshift_r(v[k], v[k + 1], sh), left = n - i;
            dst[i] = (uint8_t)x;
            if (left > 1) dst[i + 1] = (uint8_t)(x >> 8);
            if (left > 2) dst[i + 2] = (uint8_t)(x >> 16);
            if (left > 3) dst[i + 3] = (uint8_t)(x >> 24);
        }
    }
}
// memcmp-style order of two byte strings in shared or global memory, byte by byte (short strings)
__device__ __forceinline__ int compare_plain(const uint8_t *a, uint32_t al, const uint8_t *b, uint32_t bl) {
    const uint32_t m = al < bl ? al : bl;
    for (uint32_t i = 0; i < m; i++)
        if (a[i] != b[i]) return a[i] < b[i] ? -1 : 1;
    return al < bl ? -1 : (al > bl ? 1 : 0);
}
__global__ void __launch_bounds__(PACK_RECORDS) field_pack_kernel(const uint8_t *__restrict__ arena, PackArgs a) {
    __shared__ __align__(16) uint8_t s_stage[PACK_STAGE + 16];
    uint32_t block = blockIdx.x;
    const uint32_t f = block >= a.blocks0 ? 1u : 0u;
    if (f) block -= a.blocks0;
    const uint32_t m = a.m[f];
    const FieldRef *refs = a.refs[f];
    const uint32_t *off = a.off[f];
    const uint32_t j0 = block * PACK_RECORDS, j = j0 + threadIdx.x, j1 = min(j0 + PACK_RECORDS, m);
    const uint32_t out0 = off[j0], out1 = off[j1];       // the block's stretch of the blob (off has m + 1 entries)
    const uint32_t shift = out0 & 15u;                     // staged at the blob's own 16-byte phase
    const bool staged = shift + (out1 - out0) <= PACK_STAGE;
    uint8_t *blob = a.blob[f];
    FieldRef c = {0, 0, 0, 0};
    uint32_t at = 0;
    if (j < m) {
        c = refs[j];
        at = off[j];
        uint8_t *w = staged ? s_stage + shift + (at - out0) : blob + at;
        put_u32(w, c.flen);
        copy_string(w + 4, arena, c.foff, c.flen);
        put_u32(w + 4 + c.flen, c.klen);
        copy_string(w + 8 + c.flen, arena, c.koff, c.klen);
        // field order against the record before: the block's first record (and any record of a block that is
        // not staged) fetches its predecessor's field from the arena, the others find it in the staging area
        if (j > 0 && (!staged || j == j0)) {
            const FieldRef p = refs[j - 1];
            if (compare_bytes(arena, p.foff, p.flen, c.foff, c.flen) >= 0) atomicOr(a.viol_bits[f] + (j >> 5), 1u << (j & 31));
        }
    }
    if (!staged) return;
    __syncthreads();
    if (j < m && j > j0) {
        const uint8_t *mine = s_stage + shift + (at - out0), *prev = s_stage + shift + (off[j - 1] - out0);
        uint32_t pl = (uint32_t)prev[0] | ((uint32_t)prev[1] << 8) | ((uint32_t)prev[2] << 16) | ((uint32_t)prev[3] << 24);
        if (compare_plain(prev + 4, pl, mine + 4, c.flen) >= 0) atomicOr(a.viol_bits[f] + (j >> 5), 1u << (j & 31));
    }
    // whole 16-byte chunks as vectors, the ragged first and last chunk (shared with the neighbouring blocks) byte by byte
    uint8_t *g = blob + (out0 - shift);
    const uint32_t end = shift + (out1 - out0), chunks = (end + 15u) >> 4;
    for (uint32_t c = threadIdx.x; c < chunks; c += PACK_RECORDS) {
        const uint32_t b0 = c << 4;
        if (b0 >= shift && b0 + 16 <= end) reinterpret_cast<uint4 *>(g)[c] = reinterpret_cast<const uint4 *>(s_stage)[c];
        else for (uint32_t b = max(b0, shift); b < min(b0 + 16, end); b++) g[b] = s_stage[b];
    }
}

}  // namespace

size_t field_status_words(uint32_t n_rows) { return 2 + 2 * (size_t)((n_rows + FF_TILE - 1) / FF_TILE); }

cudaError_t launch_field_find(const TapeView &t, FieldIndexBuffers fx, ParseResult *res, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(fx.status, 0, field_status_words(t.n_rows) * 8, s);   // ticket counter and look-back status words
    if (e != cudaSuccess) return e;
    // (a second run -- with larger lists -- starts the totals again)
    if ((e = cudaMemsetAsync(reinterpret_cast<char *>(res) + offsetof(ParseResult, field_count), 0,
                             sizeof(ParseResult) - offsetof(ParseResult, field_count), s)) != cudaSuccess) return e;
    if (t.n_rows == 0) return cudaSuccess;
    field_find_kernel<<<(t.n_rows + FF_TILE - 1) / FF_TILE, FF_THREADS, 0, s>>>(t, t.n_rows, fx, res);
    count_parse_launch();
    return cudaGetLastError();
}

cudaError_t launch_field_pack(const uint8_t *arena, const FieldIndexBuffers &fx, const uint32_t m[2], uint32_t *const viol_bits[2],
                              uint8_t *const blob[2], cudaStream_t s) {
    PackArgs a;
    uint32_t blocks[2];
    for (int f = 0; f < 2; f++) {
        a.refs[f] = fx.refs[f]; a.off[f] = fx.off[f]; a.viol_bits[f] = viol_bits[f]; a.blob[f] = blob[f]; a.m[f] = m[f];
        blocks[f] = (m[f] + PACK_RECORDS - 1) / PACK_RECORDS;
    }
    a.blocks0 = blocks[0];
    if (blocks[0] + blocks[1] == 0) return cudaSuccess;
    field_pack_kernel<<<blocks[0] + blocks[1], PACK_RECORDS, 0, s>>>(arena, a);
    count_parse_launch();
    return cudaGetLastError();
}

}  // namespace tk
