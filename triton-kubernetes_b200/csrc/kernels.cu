// kernels.cu -- sm_100a kernels of the state.State bulk passes.
//
//   row_offsets_kernel   State.Bytes(), pass 1: escaped length of every output line and the
//                        exclusive prefix sum of those lengths (single pass, decoupled look-back)
//   emit_kernel          State.Bytes(), pass 2: every line written at its offset
//   prefix_match_kernel  State.Clusters()/Nodes(): module keys that start with a prefix
//
// All three are HBM/L2-bound byte work (no flops): the design rules that matter are coalesced
// 16-byte global accesses, shared-memory staging of the unaligned string bytes, and grids that
// are many multiples of the 148 SMs.  Reference semantics restated here:
//   encoding/json encodeState.string (HTML-safe escaping, \\ufffd for invalid UTF-8, U+2028/9)
//   encoding/json Indent("", "\t")   (line layout, see tape.hpp)
//   state/state.go:102-110,134-144   strings.Index(key, prefix) == 0
#include "kernels.cuh"

#include "bulkcopy.cuh"

#include <atomic>

namespace tk {

static std::atomic<unsigned long long> g_launches{0};
unsigned long long kernel_launch_count() { return g_launches.load(); }

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int WARPS = EMIT_THREADS / 32;

// ---- byte classification (encoding/json: htmlSafeSet) -----------------------------------

// 1 bit per ASCII code: the byte is copied as is
__device__ __forceinline__ bool ascii_plain(uint32_t c) {
    // 0x00-0x1f escape; '"' 0x22, '&' 0x26, '<' 0x3c, '>' 0x3e, '\\' 0x5c escape; 0x7f is plain
    const uint32_t w = (c < 32) ? 0u : (c < 64) ? 0xAFFFFFBBu : (c < 96) ? 0xEFFFFFFFu : 0xFFFFFFFFu;
    return (w >> (c & 31)) & 1u;
}

// escaped size of one ASCII byte that is not plain
__device__ __forceinline__ uint32_t ascii_esc_len(uint32_t c) {
    return (c == '"' || c == '\\' || c == '\n' || c == '\r' || c == '\t') ? 2u : 6u;
}

// unicode/utf8.DecodeRune validity: length of the sequence starting at s[i], 0 if invalid
__device__ __forceinline__ uint32_t utf8_seq_len(const uint8_t *s, uint32_t i, uint32_t n) {
    const uint32_t c = s[i];
    uint32_t lo = 0x80, hi = 0xBF, len;
    if (c >= 0xC2 && c <= 0xDF) len = 2;
    else if (c >= 0xE0 && c <= 0xEF) { len = 3; if (c == 0xE0) lo = 0xA0; if (c == 0xED) hi = 0x9F; }
    else if (c >= 0xF0 && c <= 0xF4) { len = 4; if (c == 0xF0) lo = 0x90; if (c == 0xF4) hi = 0x8F; }
    else return 0;
    if (i + len > n) return 0;
    const uint32_t c1 = s[i + 1];
    if (c1 < lo || c1 > hi) return 0;
    for (uint32_t k = 2; k < len; k++)
        if ((s[i + k] & 0xC0) != 0x80) return 0;
    return len;
}

// length of the JSON-escaped form of s[0, n) (without the surrounding quotes)
__device__ __forceinline__ uint32_t escaped_len(const uint8_t *s, uint32_t n) {
    uint32_t len = 0;
    for (uint32_t i = 0; i < n;) {
        const uint32_t c = s[i];
        if (c < 0x80) {
            len += ascii_plain(c) ? 1u : ascii_esc_len(c);
            i++;
            continue;
        }
        const uint32_t l = utf8_seq_len(s, i, n);
        if (l == 0) { len += 6; i += 1; continue; }                                   // \\ufffd
        if (l == 3 && c == 0xE2 && s[i + 1] == 0x80 && (s[i + 2] & 0xFE) == 0xA8) {    // U+2028 / U+2029
            len += 6; i += 3; continue;
        }
        len += l;
        i += l;
    }
    return len;
}

__device__ __forceinline__ uint8_t hex_digit(uint32_t v) { return (uint8_t)(v < 10 ? '0' + v : 'a' + (v - 10)); }

// writes the escaped form of s[0, n) at dst[pos...]; returns the new pos
__device__ __forceinline__ uint32_t put_escaped(uint8_t *dst, uint32_t pos, const uint8_t *s, uint32_t n) {
    for (uint32_t i = 0; i < n;) {
        const uint32_t c = s[i];
        if (c < 0x80) {
            i++;
            if (ascii_plain(c)) { dst[pos++] = (uint8_t)c; continue; }
            dst[pos++] = '\\';
            switch (c) {
            case '"': case '\\': dst[pos++] = (uint8_t)c; break;
            case '\n': dst[pos++] = 'n'; break;
            case '\r': dst[pos++] = 'r'; break;
            case '\t': dst[pos++] = 't'; break;
            default:
                dst[pos++] = 'u'; dst[pos++] = '0'; dst[pos++] = '0';
                dst[pos++] = hex_digit(c >> 4); dst[pos++] = hex_digit(c & 15);
            }
            continue;
        }
        const uint32_t l = utf8_seq_len(s, i, n);
        if (l == 0) {
            dst[pos++] = '\\'; dst[pos++] = 'u'; dst[pos++] = 'f'; dst[pos++] = 'f'; dst[pos++] = 'f'; dst[pos++] = 'd';
            i += 1;
            continue;
        }
        if (l == 3 && c == 0xE2 && s[i + 1] == 0x80 && (s[i + 2] & 0xFE) == 0xA8) {
            dst[pos++] = '\\'; dst[pos++] = 'u'; dst[pos++] = '2'; dst[pos++] = '0'; dst[pos++] = '2';
            dst[pos++] = (uint8_t)('8' + (s[i + 2] & 1));
            i += 3;
            continue;
        }
        for (uint32_t k = 0; k < l; k++) dst[pos++] = s[i + k];
        i += l;
    }
    return pos;
}

struct Row { uint32_t meta, koff, klen, voff, vlen; };

__device__ __forceinline__ Row load_row(const TapeView &t, uint32_t r) {
    Row w;
    w.meta = __ldg(t.meta + r); w.koff = __ldg(t.koff + r); w.klen = __ldg(t.klen + r);
    w.voff = __ldg(t.voff + r); w.vlen = __ldg(t.vlen + r);
    return w;
}
__device__ __forceinline__ bool row_has_value_bytes(const Row &w) { return (w.meta & 15u) <= ROW_RAW && w.vlen; }

// Stages the arena bytes the 32 rows of a warp refer to into that warp's shared-memory slice with
// coalesced 16-byte loads.  Returns true and sets *base (arena offset of slice byte 0) when the
// span fits; the caller then reads arena byte x at slice[x - *base].
__device__ __forceinline__ bool stage_warp_input(const TapeView &t, const Row &w, bool valid, uint4 *slice, uint32_t *base, uint64_t *bar,
                                                 uint32_t *parity) {
    uint32_t lo = 0xffffffffu, hi = 0;
    if (valid) {
        if ((w.meta & META_HAS_KEY) && w.klen) { lo = min(lo, w.koff); hi = max(hi, w.koff + w.klen); }
        if (row_has_value_bytes(w)) { lo = min(lo, w.voff); hi = max(hi, w.voff + w.vlen); }
    }
    lo = __reduce_min_sync(FULL, lo);
    hi = __reduce_max_sync(FULL, hi);
    if (hi <= lo) { *base = 0; return true; }  // no string bytes at all
    const uint32_t b = lo & ~15u;
    if (hi - b > (uint32_t)WARP_IN_BYTES) return false;
    const uint32_t chunks = (hi - b + 15u) >> 4;
    // one bulk copy for the whole span (the arena is readable 16 bytes past its end)
    if ((threadIdx.x & 31) == 0) bulk_load(slice, t.arena + b, chunks << 4, bar);
    mbar_wait(bar, *parity);
    *parity ^= 1u;
    *base = b;
    return true;
}

// Output length of one line.  A plain string (tape.hpp) is as long escaped as raw, so only the
// rare non-plain ones have their bytes examined.
__device__ __forceinline__ uint32_t row_out_len(const Row &w, const uint8_t *src) {
    uint32_t len = w.meta >> META_DEPTH_SHIFT;
    if (w.meta & META_HAS_KEY) len += ((w.meta & META_KEY_PLAIN) ? w.klen : escaped_len(src + w.koff, w.klen)) + 4;  // "key":<space>
    switch (w.meta & 15u) {
    case ROW_STR: len += ((w.meta & META_VAL_PLAIN) ? w.vlen : escaped_len(src + w.voff, w.vlen)) + 2; break;
    case ROW_RAW: len += w.vlen; break;
    case ROW_EMPTY_OBJ: case ROW_EMPTY_ARR: len += 2; break;
    default: len += 1;
    }
    if (w.meta & META_COMMA) len += 1;
    if (!(w.meta & META_LAST)) len += 1;
    return len;
}

__device__ __forceinline__ bool row_needs_bytes(const Row &w) {
    return ((w.meta & META_HAS_KEY) && !(w.meta & META_KEY_PLAIN) && w.klen) ||
           ((w.meta & 15u) == ROW_STR && !(w.meta & META_VAL_PLAIN) && w.vlen);
}

// ---- K1: line lengths + exclusive scan ------------------------------------------------------

constexpr unsigned long long ST_AGG = 1ull << 62, ST_PREFIX = 2ull << 62, ST_MASK = (1ull << 62) - 1;

__device__ __forceinline__ unsigned long long ld_status(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_status(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__global__ void __launch_bounds__(EMIT_THREADS) row_offsets_kernel(TapeView t, uint32_t *__restrict__ off,
                                                                  unsigned long long *scratch) {
    __shared__ unsigned long long s_warp_sum[WARPS];
    __shared__ uint32_t s_tile;
    __shared__ unsigned long long s_prefix;

    // tiles are handed out in launch order so that a tile only ever waits on tiles already running
    if (threadIdx.x == 0) s_tile = (uint32_t)atomicAdd(&scratch[0], 1ull);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t r0 = tile * RO_TILE + threadIdx.x * RO_ITEMS;

    // a plain string is as long escaped as raw, so the common row needs nothing but its descriptor;
    // the rare non-plain string is read straight from the arena
    uint32_t len[RO_ITEMS];
    unsigned long long sum = 0;
    if (r0 + RO_ITEMS <= t.n_rows) {
        uint32_t meta[RO_ITEMS], klen[RO_ITEMS], vlen[RO_ITEMS];
#pragma unroll
        for (int q = 0; q < RO_ITEMS / 4; q++) {
            const uint4 m = __ldg(reinterpret_cast<const uint4 *>(t.meta + r0) + q);
            const uint4 k = __ldg(reinterpret_cast<const uint4 *>(t.klen + r0) + q);
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(t.vlen + r0) + q);
            meta[q * 4] = m.x; meta[q * 4 + 1] = m.y; meta[q * 4 + 2] = m.z; meta[q * 4 + 3] = m.w;
            klen[q * 4] = k.x; klen[q * 4 + 1] = k.y; klen[q * 4 + 2] = k.z; klen[q * 4 + 3] = k.w;
            vlen[q * 4] = v.x; vlen[q * 4 + 1] = v.y; vlen[q * 4 + 2] = v.z; vlen[q * 4 + 3] = v.w;
        }
#pragma unroll
        for (int j = 0; j < RO_ITEMS; j++) {
            Row w = {meta[j], 0, klen[j], 0, vlen[j]};
            if (row_needs_bytes(w)) { w.koff = __ldg(t.koff + r0 + j); w.voff = __ldg(t.voff + r0 + j); }
            len[j] = row_out_len(w, t.arena);
            sum += len[j];
        }
    } else {
#pragma unroll
        for (int j = 0; j < RO_ITEMS; j++) {
            len[j] = 0;
            if (r0 + j < t.n_rows) len[j] = row_out_len(load_row(t, r0 + j), t.arena);
            sum += len[j];
        }
    }

    // inclusive scan of the per-thread sums inside the warp, then across the 8 warps
    unsigned long long incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long v = __shfl_up_sync(FULL, incl, d);
        if (lane >= (uint32_t)d) incl += v;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    unsigned long long warp_excl = 0, block_sum = 0;
#pragma unroll
    for (int i = 0; i < WARPS; i++) {
        const unsigned long long v = s_warp_sum[i];
        if ((uint32_t)i < warp) warp_excl += v;
        block_sum += v;
    }

    // decoupled look-back over the tile status words (warp 0)
    unsigned long long *status = scratch + 2;
    if (warp == 0) {
        unsigned long long prefix = 0;
        if (tile == 0) {
            if (lane == 0) st_status(status, ST_PREFIX | block_sum);
        } else {
            if (lane == 0) st_status(status + tile, ST_AGG | block_sum);
            int look = (int)tile - 1;
            for (;;) {
                const int idx = look - (int)lane;
                unsigned long long s = idx >= 0 ? ld_status(status + idx) : ST_PREFIX;
                while (__any_sync(FULL, (s >> 62) == 0)) {
                    if ((s >> 62) == 0) s = ld_status(status + idx);
                }
                const unsigned pm = __ballot_sync(FULL, (s >> 62) == 2);
                const int first = pm ? __ffs(pm) - 1 : 32;
                unsigned long long v = ((int)lane <= first) ? (s & ST_MASK) : 0ull;
#pragma unroll
                for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
                prefix += v;
                if (pm) break;
                look -= 32;
            }
            if (lane == 0) st_status(status + tile, ST_PREFIX | (prefix + block_sum));
        }
        if (lane == 0) {
            s_prefix = prefix;
            if ((unsigned long long)(tile + 1) * RO_TILE >= t.n_rows) scratch[1] = prefix + block_sum;  // total
        }
    }
    __syncthreads();
    unsigned long long o = s_prefix + warp_excl + (incl - sum);
#pragma unroll
    for (int j = 0; j < RO_ITEMS; j++) {
        if (r0 + j < t.n_rows) off[r0 + j] = (uint32_t)o;
        o += len[j];
        if (r0 + j + 1 == t.n_rows) off[r0 + j + 1] = o >= 0xffffffffull ? 0xffffffffu : (uint32_t)o;   // (4 GiB or more: the offsets have wrapped)
    }
}

// ---- K2: emit -------------------------------------------------------------------------------

constexpr uint32_t LANE_COPY_MAX = 16;    // generic path: plain runs up to this many bytes are copied by the row's own lane
constexpr uint32_t LANE_WORDS_MAX = 128;  // shared-memory path: up to this many bytes, 4 at a time

// A plain byte run of one row that is left for the whole warp to copy together (emit_kernel,
// phase 2), so that one long string does not hold 31 lanes idle.
struct Deferred { uint32_t src, dst, len; };

template <bool SMEM>
__device__ __forceinline__ uint32_t put_plain(uint8_t *dst, uint32_t pos, const uint8_t *src, uint32_t off, uint32_t len, Deferred *d) {
    if (SMEM) {
        // both buffers are shared memory: copy aligned 32-bit words, realigning the source with a
        // funnel shift (the staged slice is readable a word past the run)
        if (len > LANE_WORDS_MAX) { d->src = off; d->dst = pos; d->len = len; return pos + len; }
        uint32_t n = len, di = pos, si = off;
        while (n && (di & 3u)) { dst[di++] = src[si++]; n--; }
        if (n >= 4) {
            const uint32_t sh = (si & 3u) * 8u;
            const uint32_t *sw = reinterpret_cast<const uint32_t *>(src + (si & ~3u));
            uint32_t *dw = reinterpret_cast<uint32_t *>(dst + di);
            uint32_t lo = sw[0];
            const uint32_t words = n >> 2;
            for (uint32_t k = 0; k < words; k++) {
                const uint32_t hi = sw[k + 1];
                dw[k] = __funnelshift_r(lo, hi, sh);
                lo = hi;
            }
            di += words * 4; si += words * 4; n &= 3u;
        }
        while (n) { dst[di++] = src[si++]; n--; }
        return pos + len;
    }
    if (len <= LANE_COPY_MAX) {
        for (uint32_t i = 0; i < len; i++) dst[pos + i] = src[off + i];
    } else {
        d->src = off; d->dst = pos; d->len = len;
    }
    return pos + len;
}

// one line, written at dst[pos...]; src is indexed by arena offset
template <bool SMEM>
__device__ __forceinline__ void put_line(uint8_t *dst, uint32_t pos, const Row &w, const uint8_t *src, Deferred *dk, Deferred *dv) {
    const uint32_t depth = w.meta >> META_DEPTH_SHIFT;
    for (uint32_t i = 0; i < depth; i++) dst[pos++] = '\t';
    if (w.meta & META_HAS_KEY) {
        dst[pos++] = '"';
        if (w.meta & META_KEY_PLAIN) pos = put_plain<SMEM>(dst, pos, src, w.koff, w.klen, dk);
        else pos = put_escaped(dst, pos, src + w.koff, w.klen);
        dst[pos++] = '"'; dst[pos++] = ':'; dst[pos++] = ' ';
    }
    switch (w.meta & 15u) {
    case ROW_STR:
        dst[pos++] = '"';
        if (w.meta & META_VAL_PLAIN) pos = put_plain<SMEM>(dst, pos, src, w.voff, w.vlen, dv);
        else pos = put_escaped(dst, pos, src + w.voff, w.vlen);
        dst[pos++] = '"';
        break;
    case ROW_RAW: pos = put_plain<SMEM>(dst, pos, src, w.voff, w.vlen, dv); break;
    case ROW_OPEN_OBJ: dst[pos++] = '{'; break;
    case ROW_OPEN_ARR: dst[pos++] = '['; break;
    case ROW_CLOSE_OBJ: dst[pos++] = '}'; break;
    case ROW_CLOSE_ARR: dst[pos++] = ']'; break;
    case ROW_EMPTY_OBJ: dst[pos++] = '{'; dst[pos++] = '}'; break;
    default: dst[pos++] = '['; dst[pos++] = ']'; break;
    }
    if (w.meta & META_COMMA) dst[pos++] = ',';
    if (!(w.meta & META_LAST)) dst[pos++] = '\n';
}

__device__ __forceinline__ void warp_copy_deferred(uint8_t *dst, const uint8_t *src, const Deferred &d, uint32_t lane) {
    for (unsigned m = __ballot_sync(FULL, d.len != 0); m; m &= m - 1) {
        const int l = __ffs(m) - 1;
        const uint32_t so = __shfl_sync(FULL, d.src, l), dof = __shfl_sync(FULL, d.dst, l), n = __shfl_sync(FULL, d.len, l);
        for (uint32_t i = lane; i < n; i += 32) dst[dof + i] = src[so + i];
    }
}

// Input that is already in output form.  A state file that this library or the reference wrote
// is json.MarshalIndent text, so when it is loaded and serialised again, line r of the output is
// byte for byte a stretch of the input, and the 32 lines of a warp's tile one contiguous stretch.
// The warp copies text[ls0, ls0 + span) -- ls0 = where the first row's line would start -- into
// its output slice (realigned to the destination's 16-byte phase) and every lane then CHECKS its
// own line there: the tabs, the quotes, `": `, the comma and the newline by value, the key, the
// value and the brackets by position (they are referred to in place, so "at the right place"
// means "the right bytes").  One lane saying no sends the tile down the general path.
__device__ __forceinline__ bool emit_verbatim(const TapeView &t, const Row &w, uint32_t o, bool valid, uint32_t lane, uint4 *staging,
                                              uint8_t *slice, uint32_t out0, uint32_t span, uint32_t shift, uint64_t *bar, uint32_t *parity) {
    const uint32_t d = w.meta >> META_DEPTH_SHIFT, kind = w.meta & 15u;
    // text position of the first byte after the indentation, as the row's own fields give it
    uint32_t q;
    if (w.meta & META_HAS_KEY) q = w.koff - 1;
    else if (kind == ROW_STR) q = w.voff - 1;
    else if (kind == ROW_RAW) q = w.voff;
    else q = w.vlen;   // containers: the bracket (TapeView::text_len)
    const uint32_t ls0 = __shfl_sync(FULL, q - d, 0);
    if (ls0 > t.text_len || span > t.text_len - ls0 || ls0 < shift) return false;
    // text[ls0 - shift ...] -> slice[0 ...].  Where nothing before this tile has changed length, the
    // text sits at the output's own 16-byte phase and ONE bulk copy brings it straight into the
    // slice; otherwise the text goes in 16-byte chunks into the input staging and from there word
    // by word into the slice, the two phases reconciled by a funnel shift.
    if (((ls0 - shift) & 15u) == 0u) {
        const uint32_t bytes = ((shift + span + 15u) >> 4) << 4;   // <= WARP_OUT_BYTES (the tile fits), within the padded text
        if (lane == 0) bulk_load(slice, t.arena + (ls0 - shift), bytes, bar);
        mbar_wait(bar, *parity);
        *parity ^= 1u;
    } else {
        const uint32_t first = ls0 - shift, a0 = first & ~15u, delta = first - a0;
        const uint32_t chunks = (delta + shift + span + 15u) >> 4;
        if (chunks * 16u + 16u > (uint32_t)WARP_IN_BYTES) return false;
        const uint4 *src = reinterpret_cast<const uint4 *>(t.arena + a0);
        // (one chunk more than the span: the funnel reads a word ahead; all loads issued before any is used)
        constexpr int ROUNDS = WARP_IN_BYTES / 16 / 32;
        uint4 v[ROUNDS];
#pragma unroll
        for (int k = 0; k < ROUNDS; k++) {
            const uint32_t c = lane + 32u * k;
            if (c <= chunks) v[k] = __ldg(src + c);
        }
#pragma unroll
        for (int k = 0; k < ROUNDS; k++) {
            const uint32_t c = lane + 32u * k;
            if (c <= chunks) staging[c] = v[k];
        }
        __syncwarp();
        const uint32_t *sw = reinterpret_cast<const uint32_t *>(staging) + (delta >> 2);
        uint32_t *dw = reinterpret_cast<uint32_t *>(slice);
        const uint32_t words = (shift + span + 3u) >> 2, sh = (delta & 3u) * 8u;
        for (uint32_t j = lane; j < words; j += 32) dw[j] = __funnelshift_r(sw[j], sw[j + 1], sh);
    }
    __syncwarp();
    bool ok = true;
    if (valid) {
        uint32_t c = shift + (o - out0);       // slice offset of this line
        uint32_t tq = ls0 + (o - out0) + d;    // text position its first non-tab byte must have
        ok = q == tq && d <= 64;
        for (uint32_t i = 0; ok && i < d; i++) ok = slice[c + i] == '\t';
        c += d;
        if (ok && (w.meta & META_HAS_KEY)) {
            ok = (w.meta & META_KEY_PLAIN) && slice[c] == '"' && slice[c + 1 + w.klen] == '"' && slice[c + 2 + w.klen] == ':' &&
                 slice[c + 3 + w.klen] == ' ';
            c += w.klen + 4;
            tq += w.klen + 4;
        }
        if (ok) {
            switch (kind) {
            case ROW_STR:
                ok = (w.meta & META_VAL_PLAIN) && w.voff == tq + 1 && slice[c] == '"' && slice[c + 1 + w.vlen] == '"';
                c += w.vlen + 2;
                break;
            case ROW_RAW: ok = w.voff == tq; c += w.vlen; break;
            case ROW_EMPTY_OBJ: ok = w.vlen == tq && slice[c + 1] == '}'; c += 2; break;
            case ROW_EMPTY_ARR: ok = w.vlen == tq && slice[c + 1] == ']'; c += 2; break;
            default: ok = w.vlen == tq; c += 1; break;   // { [ } ]
            }
        }
        if (ok && (w.meta & META_COMMA)) { ok = slice[c] == ','; c++; }
        if (ok && !(w.meta & META_LAST)) ok = slice[c] == '\n';
    }
    return __all_sync(FULL, ok);
}

__global__ void __launch_bounds__(EMIT_THREADS) emit_kernel(TapeView t, const uint32_t *__restrict__ off,
                                                           uint8_t *__restrict__ out) {
    extern __shared__ uint4 emit_smem[];  // [WARPS][WARP_IN_BYTES/16] then [WARPS][WARP_OUT_BYTES/16]
    uint4(*s_in)[WARP_IN_BYTES / 16] = reinterpret_cast<uint4(*)[WARP_IN_BYTES / 16]>(emit_smem);
    uint4(*s_out)[WARP_OUT_BYTES / 16] = reinterpret_cast<uint4(*)[WARP_OUT_BYTES / 16]>(emit_smem + WARPS * (WARP_IN_BYTES / 16));

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t r0 = blockIdx.x * EMIT_ROWS_PER_BLOCK + warp * 32;  // first row of this warp's tile
    if (r0 >= t.n_rows) return;
    const uint32_t r = r0 + lane;
    const bool valid = r < t.n_rows;

    Row w = {0, 0, 0, 0, 0};
    uint32_t o = 0;
    if (valid) { w = load_row(t, r); o = __ldg(off + r); }

    const uint32_t r_end = min(r0 + 32u, t.n_rows);
    const uint32_t out0 = __ldg(off + r0), out1 = __ldg(off + r_end);  // off has n_rows + 1 entries
    const uint32_t shift = out0 & 15u;                                  // same 16-byte phase as global
    const uint32_t span = out1 - out0;
    const bool fits = shift + span <= (uint32_t)WARP_OUT_BYTES;        // else: straight to global memory

    uint8_t *slice = reinterpret_cast<uint8_t *>(s_out[warp]);
    __shared__ __align__(8) uint64_t s_bar[WARPS];   // one mbarrier per warp: its bulk loads complete on it, one after another
    if (lane == 0) mbar_init(&s_bar[warp], 1);
    __syncwarp();
    uint32_t parity = 0;
    Deferred dk = {0, 0, 0}, dv = {0, 0, 0};
    uint32_t base = 0;
    bool staged = false;
    const bool verbatim = t.text_len != 0 && fits && emit_verbatim(t, w, o, valid, lane, s_in[warp], slice, out0, span, shift, &s_bar[warp], &parity);
    if (!verbatim) staged = stage_warp_input(t, w, valid, s_in[warp], &base, &s_bar[warp], &parity);
    if (verbatim) {
        // the slice already holds the 32 lines
    } else if (staged && fits) {
        // the usual case, kept apart so that every access below is a 32-bit shared-memory access:
        // arena byte x is s_src[x - base]; output byte y is slice[y - out0 + shift]
        const uint8_t *s_src = reinterpret_cast<const uint8_t *>(s_in[warp]);
        Row v = w;
        v.koff -= base; v.voff -= base;
        // phase 1: every lane writes its own line, leaving long plain runs for phase 2
        if (valid) put_line<true>(slice, o - out0 + shift, v, s_src, &dk, &dv);
        // phase 2: the warp copies the long runs together, 32 bytes per step
        warp_copy_deferred(slice, s_src, dk, lane);
        warp_copy_deferred(slice, s_src, dv, lane);
    } else {
        // a tile that outgrew its staging: same steps on generic pointers
        const uint8_t *src = staged ? reinterpret_cast<const uint8_t *>(s_in[warp]) - base : t.arena;
        uint8_t *dst = fits ? slice : out;
        if (valid) put_line<false>(dst, fits ? o - out0 + shift : o, w, src, &dk, &dv);
        warp_copy_deferred(dst, src, dk, lane);
        warp_copy_deferred(dst, src, dv, lane);
    }
    if (!fits) return;
    // copy out: the whole 16-byte chunks of the slice in ONE bulk copy; the ragged first and last
    // chunk (shared with the neighbouring warps' tiles) a byte per lane -- lanes 0..15 the first,
    // 16..31 the last
    uint8_t *g = out + (out0 - shift);  // 16-byte aligned
    const uint32_t end = shift + span;
    const uint32_t c0 = (shift + 15u) >> 4, c1 = end >> 4;                 // whole chunks: [c0, c1)
    const uint32_t head_end = min(c0 << 4, end), tail_start = max(c1 << 4, head_end);
    if (!verbatim) fence_smem_for_bulk();   // (the lines were written with ordinary stores)
    __syncwarp();
    if (lane == 0 && c1 > c0) bulk_store(g + (c0 << 4), slice + (c0 << 4), (c1 - c0) << 4);
    {
        const uint32_t b = lane < 16 ? shift + lane : tail_start + (lane - 16);
        if (lane < 16 ? b < head_end : b < end) g[b] = slice[b];
    }
    if (lane == 0 && c1 > c0) bulk_store_wait_read();   // the slice must outlive the copy unit's reads of it
}

// ---- K3: key prefix classification -------------------------------------------------------------

constexpr int PREFIX_THREADS = 256;
constexpr int PREFIX_SMEM = 256;

__global__ void __launch_bounds__(PREFIX_THREADS) prefix_match_kernel(KeyIndexView k, const uint8_t *__restrict__ prefix,
                                                                     uint32_t plen, uint32_t *hits) {
    __shared__ uint8_t s_prefix[PREFIX_SMEM];
    for (uint32_t i = threadIdx.x; i < min(plen, (uint32_t)PREFIX_SMEM); i += blockDim.x) s_prefix[i] = prefix[i];
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t n_round = (k.n_keys + 31u) & ~31u;  // keep warps converged for the ballot
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x) {
        bool match = false;
        if (i < k.n_keys) {
            const uint32_t b = __ldg(k.koffs + i), e = __ldg(k.koffs + i + 1);
            if (e - b >= plen) {
                const uint8_t *key = k.keys + b;
                uint32_t j = 0;
                const uint32_t ps = min(plen, (uint32_t)PREFIX_SMEM);
                while (j < ps && __ldg(key + j) == s_prefix[j]) j++;
                if (j == ps) while (j < plen && __ldg(key + j) == __ldg(prefix + j)) j++;
                match = (j == plen);
            }
        }
        const unsigned m = __ballot_sync(FULL, match);
        if (m) {
            uint32_t slot = 0;
            if (lane == (uint32_t)(__ffs(m) - 1)) slot = atomicAdd(hits, (uint32_t)__popc(m));
            slot = __shfl_sync(FULL, slot, __ffs(m) - 1);
            if (match) hits[1 + slot + __popc(m & ((1u << lane) - 1))] = i;
        }
    }
}

}  // namespace

cudaError_t launch_row_offsets(const TapeView &t, uint32_t *off, unsigned long long *scratch, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(scratch, 0, scan_scratch_words(t.n_rows) * sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
    row_offsets_kernel<<<row_offsets_tiles(t.n_rows), EMIT_THREADS, 0, s>>>(t, off, scratch);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_emit(const TapeView &t, const uint32_t *off, uint8_t *out, cudaStream_t s) {
    constexpr int smem = WARPS * (WARP_IN_BYTES + WARP_OUT_BYTES);
    // the opt-in above 48 KiB is a per-device (per-context) attribute of the function: asked for
    // once on every device a handle of this process launches on
    static std::atomic<bool> configured[64];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const bool slot = dev >= 0 && dev < 64;
    if (!slot || !configured[dev].load(std::memory_order_acquire)) {
        if ((e = cudaFuncSetAttribute(emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != cudaSuccess) return e;
        if (slot) configured[dev].store(true, std::memory_order_release);
    }
    emit_kernel<<<emit_tiles(t.n_rows), EMIT_THREADS, smem, s>>>(t, off, out);
    g_launches++;
    return cudaGetLastError();
}

cudaError_t launch_prefix_match(const KeyIndexView &k, const uint8_t *prefix, uint32_t prefix_len, uint32_t *hits,
                                cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(hits, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    if (k.n_keys == 0) return cudaSuccess;
    const uint32_t blocks = min((k.n_keys + PREFIX_THREADS - 1) / PREFIX_THREADS, 148u * 8u);
    prefix_match_kernel<<<blocks, PREFIX_THREADS, 0, s>>>(k, prefix, prefix_len, hits);
    g_launches++;
    return cudaGetLastError();
}

}  // namespace tk

// ---------------------------------------------------------------------------------------------
// tk_row_offsets_noalloc: the line-offset pass for callers that must not allocate.
//
// Who this is for: a host program that has handed all of its GPU memory to its own pool or
// arena up front (so a cudaMalloc from a library would either fail or eat memory the program
// has budgeted for something else), and code on a latency-critical path where cudaMalloc /
// cudaFree are unwelcome because they can synchronise the whole device and stall unrelated
// streams.  Such a caller owns every input and output buffer; the only thing the offset pass
// needs beyond those is its small scan scratch (one tile counter, one total, one look-back status
// word per tile of 2,048 rows), and here that lives in the fixed array below instead of a
// cudaMalloc'd buffer.  The price is a size limit and one shared scratch per GPU.
// ---------------------------------------------------------------------------------------------

#include <cstdio>
#include <mutex>

#include "tk_state.h"

namespace {

constexpr uint32_t NOALLOC_MAX_TILES = TK_NOALLOC_MAX_ROWS / tk::RO_TILE;
static_assert(TK_NOALLOC_MAX_ROWS % tk::RO_TILE == 0, "the row limit is a whole number of tiles");
// 2 + 1,024 words = 8,208 bytes of device memory per GPU that loads this library, held until
// the process exits.
static __device__ __align__(256) unsigned long long g_noalloc_scratch[2 + NOALLOC_MAX_TILES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
unsigned long long *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

void noalloc_err(char *err, size_t cap, const char *msg, cudaError_t e = cudaSuccess) {
    if (!err || !cap) return;
    if (e == cudaSuccess) snprintf(err, cap, "%s", msg);
    else snprintf(err, cap, "%s: %s", msg, cudaGetErrorString(e));
}

}  // namespace

extern "C" int tk_row_offsets_noalloc(const uint32_t *meta, const uint32_t *koff, const uint32_t *klen,
                                      const uint32_t *voff, const uint32_t *vlen, const uint8_t *arena,
                                      uint32_t arena_bytes, uint32_t n_rows, uint32_t *off_out, void *stream,
                                      char *err, size_t err_cap) {
    if (!meta || !koff || !klen || !voff || !vlen || !off_out || (arena_bytes && !arena)) {
        noalloc_err(err, err_cap, "tk_row_offsets_noalloc: NULL argument");
        return TK_EARG;
    }
    if ((reinterpret_cast<uintptr_t>(meta) | reinterpret_cast<uintptr_t>(klen) | reinterpret_cast<uintptr_t>(vlen)) & 15u) {
        noalloc_err(err, err_cap, "tk_row_offsets_noalloc: meta, klen and vlen must be 16-byte aligned");
        return TK_EARG;
    }
    if (n_rows == 0 || n_rows > TK_NOALLOC_MAX_ROWS) {
        if (err && err_cap)
            snprintf(err, err_cap, "tk_row_offsets_noalloc: n_rows = %u is outside 1..%u (the fixed scratch holds one status word per %d rows)",
                     n_rows, (unsigned)TK_NOALLOC_MAX_ROWS, tk::RO_TILE);
        return TK_ELIMIT;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        noalloc_err(err, err_cap, "tk_row_offsets_noalloc: cudaGetDevice", e);
        (void)cudaGetLastError();
        return TK_ECUDA;
    }
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) {
        if (err && err_cap)
            snprintf(err, err_cap, "tk_row_offsets_noalloc: device index %d is outside 0..%d, the devices this entry point keeps an address for",
                     dev, NOALLOC_MAX_DEVICES - 1);
        return TK_ELIMIT;
    }
    unsigned long long *scratch;
    {
        // Set-up, once per device and with that device current: in host code the array's name is
        // only a stand-in, so ask the runtime for the address the kernel must be given.
        std::lock_guard<std::mutex> lock(g_noalloc_mu);
        if (!g_noalloc_ptr[dev]) {
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) {
                noalloc_err(err, err_cap, "tk_row_offsets_noalloc: cudaGetSymbolAddress", e);
                (void)cudaGetLastError();
                return TK_ECUDA;
            }
            g_noalloc_ptr[dev] = static_cast<unsigned long long *>(p);
        }
        scratch = g_noalloc_ptr[dev];
    }
    tk::TapeView t;
    t.n_rows = n_rows; t.arena_bytes = arena_bytes;
    t.meta = meta; t.koff = koff; t.klen = klen; t.voff = voff; t.vlen = vlen; t.arena = arena;
    t.text_len = 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // the array still holds the previous call's counter and status words: they must start at zero
    e = cudaMemsetAsync(scratch, 0, tk::scan_scratch_words(n_rows) * sizeof(unsigned long long), s);
    if (e == cudaSuccess) {
        tk::row_offsets_kernel<<<tk::row_offsets_tiles(n_rows), tk::EMIT_THREADS, 0, s>>>(t, off_out, scratch);
        tk::g_launches++;
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) {
        noalloc_err(err, err_cap, "tk_row_offsets_noalloc: launch", e);
        return TK_ECUDA;
    }
    return TK_OK;
}
