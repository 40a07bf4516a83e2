// lines.cu -- state.New for text that is already in encoding/json's indented form: see lines.cuh.
//
// Reference semantics restated: encoding/json Indent("", "\t") and encodeState (what the indented
// form looks like, line by line), the scanner's grammar (what makes a line sequence valid JSON),
// and map keys in sorted order (state/state.go:88-90 via gabs.go:395-402).  The general pipeline
// (parse.cu) is the authority: this pass only ever accepts a subset of what that one keeps on the
// device, and builds the same rows for it.
#include "lines.cuh"

#include <algorithm>
#include <atomic>

#include "bulkcopy.cuh"
#include "index.cuh"
#include "pdl.cuh"
#include "unquote.cuh"

namespace tk {
namespace {

constexpr unsigned FULL = 0xffffffffu;
typedef unsigned long long u64;
constexpr int LN_THREADS = 128, LN_CHUNK = 64;
static_assert(LN_THREADS * LN_CHUNK == LN_TILE, "a tile is one 64-byte chunk per thread");

__device__ __forceinline__ u64 low_bits(uint32_t k) { return k >= 64 ? ~0ull : ((1ull << k) - 1); }

// bit i: byte i of the 64 bytes in x[0..3] is '\n'
__device__ __forceinline__ u64 newline_mask(const uint4 (&x)[4]) {
    u64 m = 0;
#pragma unroll
    for (int v = 0; v < 4; v++) {
        const uint32_t w[4] = {x[v].x, x[v].y, x[v].z, x[v].w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t t = w[k] ^ 0x0a0a0a0au;
            const uint32_t z = ~(((t & 0x7f7f7f7fu) + 0x7f7f7f7fu) | t) & 0x80808080u;   // 0x80 in every byte that was '\n'
            m |= (u64)(((z >> 7) * 0x01020408u) >> 24) << ((v * 4 + k) * 4);
        }
    }
    return m;
}

// line ends of the 64-byte chunk of the arena that starts at `chunk` (16-byte aligned): newlines, and position n
// itself (the end of the text)
__device__ __forceinline__ u64 chunk_line_ends(const uint8_t *raw, u64 chunk, u64 n) {
    if (chunk > n) return 0;
    uint4 x[4];
#pragma unroll
    for (int v = 0; v < 4; v++) x[v] = __ldg(reinterpret_cast<const uint4 *>(raw + chunk) + v);
    u64 m = newline_mask(x);
    if (n - chunk < 64) m = (m & low_bits((uint32_t)(n - chunk))) | (1ull << (n - chunk));
    return m;
}

// ---- stage 0: count -------------------------------------------------------------------------
// COPY: the text is still where the caller has it on the device (src, 16-byte aligned) and is brought into the arena
// on the way -- one pass over it instead of a copy and then this one; only bytes below n are read from src and
// written to raw (the padding behind the text is the caller's memset).
template <bool COPY>
__global__ void __launch_bounds__(LN_THREADS) lines_count_kernel(uint8_t *raw, const uint8_t *__restrict__ src, u64 n, uint32_t *__restrict__ counts,
                                                                uint32_t *__restrict__ last_end, uint32_t *__restrict__ open_state,
                                                                uint32_t open_words) {
    __shared__ uint32_t s_cnt[LN_THREADS / 32], s_last[LN_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    {   // the row pass notes its openers in open_state with atomicMax: cleared here, a slice per block
        const uint32_t per = (open_words + gridDim.x - 1) / gridDim.x, w0 = blockIdx.x * per;
        for (uint32_t i = threadIdx.x; i < per && w0 + i < open_words; i += LN_THREADS) open_state[w0 + i] = 0;
    }
    const u64 chunk = (u64)blockIdx.x * LN_TILE + (u64)threadIdx.x * LN_CHUNK;
    u64 m;
    if (COPY) {
        // The tile goes through shared memory: one bulk copy in (the copy unit, no registers in between), the line ends
        // read off it there, one bulk copy out to the arena.  Bulk copies move multiples of 16 bytes: what is left of
        // the text's last tile behind the last multiple is moved by hand.
        __shared__ __align__(16) uint8_t s_tile[LN_TILE];
        __shared__ __align__(8) uint64_t s_bar;
        const u64 tile0 = (u64)blockIdx.x * LN_TILE;
        const uint32_t avail = n > tile0 ? (uint32_t)min((u64)LN_TILE, n - tile0) : 0u, bulk = avail & ~15u;
        if (threadIdx.x == 0) mbar_init(&s_bar, 1);
        __syncthreads();
        if (threadIdx.x == 0 && bulk) bulk_load(s_tile, src + tile0, bulk, &s_bar);
        if (threadIdx.x < avail - bulk) {
            const uint8_t b = src[tile0 + bulk + threadIdx.x];
            s_tile[bulk + threadIdx.x] = b;
            raw[tile0 + bulk + threadIdx.x] = b;
        }
        if (bulk) mbar_wait(&s_bar, 0);
        __syncthreads();   // (the bytes moved by hand)
        if (threadIdx.x == 0 && bulk) bulk_store(raw + tile0, s_tile, bulk);
        if (chunk > n) m = 0;
        else {
            uint4 x[4];
#pragma unroll
            for (int v = 0; v < 4; v++) x[v] = reinterpret_cast<const uint4 *>(s_tile + threadIdx.x * LN_CHUNK)[v];
            m = newline_mask(x);   // (bytes of the tile behind the end of the text are whatever was there: masked off)
            if (n - chunk < 64) m = (m & low_bits((uint32_t)(n - chunk))) | (1ull << (n - chunk));
        }
        if (threadIdx.x == 0 && bulk) bulk_store_wait_read();
    } else {
        m = chunk_line_ends(raw, chunk, n);
    }
    uint32_t c = (uint32_t)__popcll(m);
    uint32_t last = m ? (uint32_t)(chunk + 63 - __clzll((long long)m)) + 1u : 0u;   // position + 1; 0: none
#pragma unroll
    for (int d = 16; d; d >>= 1) { c += __shfl_xor_sync(FULL, c, d); last = max(last, __shfl_xor_sync(FULL, last, d)); }
    if (lane == 0) { s_cnt[warp] = c; s_last[warp] = last; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tc = 0, tl = 0;
        for (int w = 0; w < LN_THREADS / 32; w++) { tc += s_cnt[w]; tl = max(tl, s_last[w]); }
        counts[blockIdx.x] = tc;
        last_end[blockIdx.x] = tl;
    }
}

// in place: counts[t] -> line ends before tile t (the total is the number of lines), last_end[t] ->
// position + 1 of the last line end before tile t (0: none).  Text with fewer than one line end
// per LN_MIN_DENSITY bytes is not what this path is for (compact JSON is one line): left to the
// general parser before the row pass looks at it.
constexpr u64 LN_MIN_DENSITY = 512;
__global__ void __launch_bounds__(1024) lines_prefix_kernel(uint32_t *counts, uint32_t *last_end, uint32_t n_tiles, u64 n, ParseResult *res) {
    __shared__ uint32_t s_w[32], s_m[32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t run = 0, run_end = 0;
    griddep_wait();   // (launched while the count pass is still running)
    for (uint32_t base = 0; base < n_tiles; base += 1024 * 4) {
        const uint32_t first = base + threadIdx.x * 4;
        uint32_t v[4], e[4], sum = 0, top = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            v[j] = first + j < n_tiles ? counts[first + j] : 0u;
            e[j] = first + j < n_tiles ? last_end[first + j] : 0u;
            sum += v[j];
            top = max(top, e[j]);
        }
        uint32_t incl = sum, mincl = top;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, incl, o), z = __shfl_up_sync(FULL, mincl, o);
            if (lane >= (uint32_t)o) { incl += y; mincl = max(mincl, z); }
        }
        __syncthreads();
        if (lane == 31) { s_w[warp] = incl; s_m[warp] = mincl; }
        __syncthreads();
        uint32_t w = s_w[lane], wm = s_m[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, w, o), z = __shfl_up_sync(FULL, wm, o);
            if (lane >= (uint32_t)o) { w += y; wm = max(wm, z); }
        }
        const uint32_t total = __shfl_sync(FULL, w, 31), total_end = __shfl_sync(FULL, wm, 31);
        uint32_t before = __shfl_sync(FULL, w, warp ? warp - 1 : 0), before_end = __shfl_sync(FULL, wm, warp ? warp - 1 : 0);
        if (warp == 0) { before = 0; before_end = 0; }
        uint32_t lane_end = __shfl_up_sync(FULL, mincl, 1);
        if (lane == 0) lane_end = 0;
        uint32_t x = run + before + incl - sum, m = max(run_end, max(before_end, lane_end));
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (first + j < n_tiles) { counts[first + j] = x; last_end[first + j] = m; }
            x += v[j];
            m = max(m, e[j]);
        }
        run += total;
        run_end = max(run_end, total_end);
    }
    if (threadIdx.x == 0) {
        res->n_rows = run;
        if ((u64)run * LN_MIN_DENSITY < n) res->fast_reject = 1;
    }
}

// ---- stage 1: one row per line --------------------------------------------------------------

// ---- reading the text eight bytes at a time ----------------------------------------------------
constexpr u64 ONES = 0x0101010101010101ull, HIGHS = 0x8080808080808080ull;
// The text as the line parser reads it: a window of it sits in shared memory (positions lo .. hi,
// brought in by one bulk copy), the rest is fetched from global memory.  Thirty-two lanes reading
// thirty-two different lines are thirty-two different sectors to the L1 cache -- one tag look-up
// each -- but only a few bank conflicts to shared memory.
struct Text {
    const uint8_t *win;     // shared memory; win[0] is position lo (lo a multiple of 16)
    u64 lo, hi;
    const uint8_t *raw;     // global memory, readable 64 bytes past the end of the text
    // the 8 bytes at position p (any alignment), first byte lowest
    __device__ __forceinline__ u64 load8(u64 p) const {
        const u64 a = p & ~7ull;
        u64 x0, x1;
        if (a >= lo && a + 16 <= hi) {
            const u64 *w = reinterpret_cast<const u64 *>(win + (a - lo));
            x0 = w[0]; x1 = w[1];
        } else {
            const u64 *w = reinterpret_cast<const u64 *>(raw + a);
            x0 = __ldg(w); x1 = __ldg(w + 1);
        }
        const uint32_t sh = (uint32_t)(p & 7u) * 8u;
        return sh ? (x0 >> sh) | (x1 << (64u - sh)) : x0;
    }
    __device__ __forceinline__ uint32_t byte(u64 p) const { return p >= lo && p < hi ? (uint32_t)win[p - lo] : (uint32_t)__ldg(raw + p); }
    __device__ __forceinline__ uint32_t tabs(u64 p) const;                 // tabs at p: 0 .. 8, or 9 for "more than 8"
    __device__ __forceinline__ u64 string_end(u64 p, u64 end) const;      // see below
};

// 0x80 in bytes of x that equal c -- and possibly in bytes ABOVE such a byte (a borrow), never below
// the lowest one: the first flagged byte is always a true one, which is all the callers look at
__device__ __forceinline__ u64 eq_bytes(u64 x, uint32_t c) {
    const u64 v = x ^ (ONES * c);
    return (v - ONES) & ~v & HIGHS;
}
// likewise for bytes that may not stand for themselves in a string of the indented form: below 0x20,
// 0x80 and above, and \ < > &   (the quote is looked for separately)
__device__ __forceinline__ u64 bad_bytes(u64 x) {
    return (x & HIGHS) | ((x - ONES * 0x20u) & ~x & HIGHS) | eq_bytes(x, '\\') | eq_bytes(x, '<') | eq_bytes(x, '>') | eq_bytes(x, '&');
}
// from p (the byte after an opening quote) to the closing quote, which must come before `end`, every
// byte on the way one that stands for itself: its position, or 0
__device__ __forceinline__ u64 Text::string_end(u64 p, u64 end) const {
    const Text &raw = *this;
    for (; p < end; p += 8) {
        const u64 x = raw.load8(p);
        u64 q = eq_bytes(x, '"'), bad = bad_bytes(x);
        if (end - p < 8) { const u64 keep = (1ull << ((end - p) * 8)) - 1; q &= keep; bad &= keep; }
        if (q | bad) {
            const uint32_t iq = q ? (uint32_t)__ffsll((long long)q) : 65u, ib = bad ? (uint32_t)__ffsll((long long)bad) : 65u;
            return iq < ib ? p + ((iq - 1) >> 3) : 0;
        }
    }
    return 0;
}
// true / false / null, or an integer floatEncoder prints back unchanged (at most 15 digits, no leading zero)
template <class Reader>
__device__ __forceinline__ bool canonical_scalar(const Reader &raw, u64 a, u64 b) {
    const uint32_t len = (uint32_t)(b - a);
    if (len == 0 || len > 16) return false;
    const uint32_t c0 = raw.byte(a);
    if (c0 == 't') return len == 4 && raw.byte(a + 1) == 'r' && raw.byte(a + 2) == 'u' && raw.byte(a + 3) == 'e';
    if (c0 == 'n') return len == 4 && raw.byte(a + 1) == 'u' && raw.byte(a + 2) == 'l' && raw.byte(a + 3) == 'l';
    if (c0 == 'f') return len == 5 && raw.byte(a + 1) == 'a' && raw.byte(a + 2) == 'l' && raw.byte(a + 3) == 's' && raw.byte(a + 4) == 'e';
    u64 p = a;
    if (c0 == '-') p++;
    const uint32_t digits = (uint32_t)(b - p);
    if (digits == 0 || digits > 15) return false;
    if (raw.byte(p) == '0' && digits > 1) return false;
    for (; p < b; p++) { const uint32_t c = raw.byte(p); if (c < '0' || c > '9') return false; }
    return true;
}
// the same reads straight from global memory (the structure pass compares a few keys far apart)
struct GlobalText {
    const uint8_t *raw;
    __device__ __forceinline__ u64 load8(u64 p) const {
        const u64 *w = reinterpret_cast<const u64 *>(raw + (p & ~7ull));
        const u64 x0 = __ldg(w), x1 = __ldg(w + 1);
        const uint32_t sh = (uint32_t)(p & 7u) * 8u;
        return sh ? (x0 >> sh) | (x1 << (64u - sh)) : x0;
    }
};
// bytes of x reversed, so that comparing numbers compares byte strings
__device__ __forceinline__ u64 big_endian(u64 x) {
    return ((u64)__byte_perm((uint32_t)x, 0, 0x0123) << 32) | __byte_perm((uint32_t)(x >> 32), 0, 0x0123);
}
// The string whose first byte is at p (right after its opening quote): where its closing quote is, which must be
// before `end`, and how the row can refer to it:
//   STR_PLAIN    every byte stands for itself;
//   STR_IN_PLACE bytes from 0x80 up, all of them valid UTF-8 spelling neither U+2028 nor U+2029: that is how
//                json.Marshal writes such a string, so the decoded string is the bytes themselves, where they are;
//   STR_DECODE   a valid JSON string that reads differently once decoded (escape sequences, < > & as they are,
//                invalid UTF-8, the two separators): the caller decodes it into the side arena (unquote.cuh).
// 0: not a JSON string at all (a control byte, a bad escape, no closing quote) -- the general parser has the words for it.
enum { STR_PLAIN = 0, STR_IN_PLACE = 1, STR_DECODE = 2 };
template <class Reader>
__device__ __forceinline__ u64 string_close(const Reader &raw, u64 p, u64 end, int *mode) {
    *mode = STR_PLAIN;
    const u64 q = raw.string_end(p, end);
    if (q) return q;
    int m = STR_IN_PLACE;
    for (u64 i = p; i < end;) {
        const uint32_t c = raw.byte(i);
        if (c == '"') { *mode = m; return i; }
        if (c < 0x20) return 0;
        if (c == '\\') {
            if (i + 1 >= end) return 0;
            const uint32_t e = raw.byte(i + 1);
            if (e == 'u') {
                if (i + 6 > end) return 0;
                for (uint32_t k = 2; k < 6; k++)
                    if (!is_hex(raw.byte(i + k))) return 0;
                i += 6;
            } else if (e == '"' || e == '\\' || e == '/' || e == 'b' || e == 'f' || e == 'n' || e == 'r' || e == 't') {
                i += 2;
            } else return 0;
            m = STR_DECODE;
            continue;
        }
        if (c < 0x80) { if (c == '<' || c == '>' || c == '&') m = STR_DECODE; i++; continue; }
        uint32_t lo = 0x80, hi = 0xBF, len = 0;
        if (c >= 0xC2 && c <= 0xDF) len = 2;
        else if (c >= 0xE0 && c <= 0xEF) { len = 3; if (c == 0xE0) lo = 0xA0; if (c == 0xED) hi = 0x9F; }
        else if (c >= 0xF0 && c <= 0xF4) { len = 4; if (c == 0xF0) lo = 0x90; if (c == 0xF4) hi = 0x8F; }
        bool ok = len != 0 && i + len <= end;
        if (ok) {
            const uint32_t c1 = raw.byte(i + 1);
            ok = c1 >= lo && c1 <= hi;
            for (uint32_t k = 2; ok && k < len; k++) ok = (raw.byte(i + k) & 0xC0u) == 0x80u;
            if (ok && c == 0xE2 && c1 == 0x80 && (raw.byte(i + 2) | 1u) == 0xA9u) ok = false;   // U+2028, U+2029: written escaped
        }
        if (ok) i += len;
        else { m = STR_DECODE; i++; }   // (decodes to U+FFFD, or to the separator itself)
    }
    return 0;
}
// A string value that has to be decoded: room for it taken from the side arena (behind the text, rb.side_base
// onwards), the string written there.  false: the arena is full (the document goes to the general parser).
template <class Reader>
__device__ __forceinline__ bool decode_value(const Reader &raw, u64 a, u64 b, const RowBuffers &rb, ParseResult *res, uint32_t *voff, uint32_t *vlen) {
    const uint8_t *src = raw.raw + a;
    const uint32_t n = (uint32_t)(b - a), len = unquote(src, n, nullptr);
    const unsigned long long at = atomicAdd(&res->side_bytes, (unsigned long long)len);
    if (at + len > rb.side_cap) return false;
    unquote(src, n, rb.side + rb.side_base + at);
    *voff = rb.side_base + (uint32_t)at;
    *vlen = len;
    return true;
}
// strictly ascending, bytewise (the shorter first on a tie)?
template <class T>
__device__ __forceinline__ bool key_less(const T &raw, u64 a, uint32_t al, u64 b, uint32_t bl) {
    const uint32_t m = al < bl ? al : bl;
    for (uint32_t i = 0; i < m; i += 8) {
        u64 x = big_endian(raw.load8(a + i)), y = big_endian(raw.load8(b + i));
        if (m - i < 8) { const u64 keep = ~0ull << ((8 - (m - i)) * 8); x &= keep; y &= keep; }
        if (x != y) return x < y;
    }
    return al < bl;
}
// tabs at raw + p: 0 .. 8, or 9 for "more than 8"
__device__ __forceinline__ uint32_t Text::tabs(u64 p) const {
    const Text &raw = *this;
    const u64 t = raw.load8(p) ^ (ONES * 0x09u);
    const u64 nz = (((t & ~HIGHS) + ~HIGHS) | t) & HIGHS;     // 0x80 in every byte that is not a tab (exact)
    if (nz) return ((uint32_t)__ffsll((long long)nz) - 1) >> 3;
    return raw.byte(p + 8) == '\t' ? 9u : 8u;
}

// the row written, with what the module index and the verdict want to know about it
// Stage 2 wants, per block of LN_ROWS_PER_BLOCK rows and per depth, the last row that opens a container there,
// as (row + 1) << 1 | is-an-object: every opening row says so as it is written (about one row in seven).
__device__ __forceinline__ void note_opener(const RowBuffers &rb, uint32_t r, uint32_t kind, uint32_t d) {
    if ((kind == ROW_OPEN_OBJ || kind == ROW_OPEN_ARR) && d <= LINES_MAX_DEPTH)
        atomicMax(rb.open + (r / LN_ROWS_PER_BLOCK) * (LINES_MAX_DEPTH + 1) + d, ((r + 1) << 1) | (kind == ROW_OPEN_OBJ ? 1u : 0u));
}
template <class Reader>
__device__ __forceinline__ void finish_row(const Reader &raw, uint32_t r, uint32_t d, uint32_t meta, uint32_t koff, uint32_t klen, uint32_t voff,
                                           uint32_t vlen, RowBuffers rb, ParseResult *res) {
    const uint32_t kind = meta & 15u;
    if (d == 3 && kind == ROW_STR && (meta & META_HAS_KEY) && (klen == 4 || klen == 8)) {   // for the module index (index.cu)
        const u64 k8 = raw.load8(koff);
        if (klen == 4 && (uint32_t)k8 == 0x656d616eu) meta |= META_KEY_IS_NAME;                           // "name"
        else if (klen == 8 && k8 == 0x656d616e74736f68ull) meta |= META_KEY_IS_HOSTNAME;                  // "hostname"
    }
    rb.meta[r] = meta; rb.koff[r] = koff; rb.klen[r] = klen; rb.voff[r] = voff; rb.vlen[r] = vlen;
    note_opener(rb, r, kind, d);
    if (r == 0) res->root_kind = kind;
    if (d == 1 && (meta & META_HAS_KEY) && klen == 6 && (raw.load8(koff) & 0xffffffffffffull) == 0x656c75646f6dull) {   // "module"
        res->module_row = r + 1;
        res->module_kind = kind;
    }
}

// The line raw[s, e) becomes row r.  Returns false for anything that is not the indented form.
template <class Reader>
__device__ __forceinline__ bool line_to_row(const Reader &raw, u64 s, u64 e, u64 n, uint32_t r, RowBuffers rb, ParseResult *res) {
    const uint32_t d = raw.tabs(s);
    u64 p = s + d;
    if (d > LINES_MAX_DEPTH + 1 || p >= e) return false;
    if (r == 0 && d != 0) return false;
    uint32_t meta = 0, koff = 0, klen = 0, voff = 0, vlen = 0;
    uint32_t c = raw.byte(p);
    u64 after;   // the byte after the value
    bool value_done = false;
    if (c == '"') {
        int mode;
        const u64 q = string_close(raw, p + 1, e, &mode);
        if (!q) return false;
        if (q + 1 < e && raw.byte(q + 1) == ':') {
            // "key": value   (a key that would have to be decoded: left to the general parser -- the order of keys is
            // checked on the text as it stands)
            if (q + 2 >= e || raw.byte(q + 2) != ' ' || mode == STR_DECODE) return false;
            meta |= META_HAS_KEY | (mode == STR_PLAIN ? META_KEY_PLAIN : 0u);
            koff = (uint32_t)(p + 1);
            klen = (uint32_t)(q - p - 1);
            p = q + 3;
            if (p >= e) return false;
            c = raw.byte(p);
        } else {
            meta |= ROW_STR | (mode == STR_PLAIN ? META_VAL_PLAIN : 0u);
            voff = (uint32_t)(p + 1);
            vlen = (uint32_t)(q - p - 1);
            if (mode == STR_DECODE && !decode_value(raw, p + 1, q, rb, res, &voff, &vlen)) return false;
            after = q + 1;
            value_done = true;
        }
    }
    if (!value_done) {
        if (c == '"') {
            int mode;
            const u64 q = string_close(raw, p + 1, e, &mode);
            if (!q) return false;
            meta |= ROW_STR | (mode == STR_PLAIN ? META_VAL_PLAIN : 0u);
            voff = (uint32_t)(p + 1);
            vlen = (uint32_t)(q - p - 1);
            if (mode == STR_DECODE && !decode_value(raw, p + 1, q, rb, res, &voff, &vlen)) return false;
            after = q + 1;
        } else if (c == '{' || c == '[') {
            const uint32_t closer = c == '{' ? '}' : ']';
            vlen = (uint32_t)p;   // container rows keep the text position of their bracket (TapeView::text_len)
            if (p + 1 < e && raw.byte(p + 1) == closer) { meta |= c == '{' ? ROW_EMPTY_OBJ : ROW_EMPTY_ARR; after = p + 2; }
            else { meta |= c == '{' ? ROW_OPEN_OBJ : ROW_OPEN_ARR; after = p + 1; }
        } else if (c == '}' || c == ']') {
            if (meta & META_HAS_KEY) return false;
            meta |= c == '}' ? ROW_CLOSE_OBJ : ROW_CLOSE_ARR;
            vlen = (uint32_t)p;
            after = p + 1;
        } else {
            u64 q = p;
            while (q < e && raw.byte(q) != ',') q++;
            if (!canonical_scalar(raw, p, q)) return false;
            meta |= ROW_RAW;
            voff = (uint32_t)p;
            vlen = (uint32_t)(q - p);
            after = q;
        }
    }
    const uint32_t kind = meta & 15u;
    const bool is_open = kind == ROW_OPEN_OBJ || kind == ROW_OPEN_ARR;
    bool comma = false;
    if (after < e && raw.byte(after) == ',') { comma = true; after++; }
    if (after != e || (comma && is_open)) return false;
    if (comma) meta |= META_COMMA;
    meta |= d << META_DEPTH_SHIFT;

    // ---- this line against the next one: indentation, comma, key order
    if (e >= n) {
        if (d != 0 || comma || is_open) return false;
        meta |= META_LAST;
    } else {
        const uint32_t dn = raw.tabs(e + 1);
        const u64 q = e + 1 + dn;
        if (dn > LINES_MAX_DEPTH + 1 || q >= n) return false;   // (too deep, or an empty last line)
        const uint32_t cn = raw.byte(q);
        const bool next_closes = cn == '}' || cn == ']';
        if (is_open) {
            if (dn != d + 1 || next_closes) return false;   // ("{" then "}" is written "{}")
        } else if (next_closes) {
            if (d == 0 || dn != d - 1 || comma) return false;
        } else {
            if (dn != d || !comma) return false;
            if (meta & META_HAS_KEY) {
                // the next member of the same object: its key must sort strictly after this one
                if (cn != '"') return false;
                int next_mode;
                const u64 qe = string_close(raw, q + 1, n, &next_mode);    // (a newline on the way is not a byte that stands for itself)
                if (!qe || next_mode == STR_DECODE || !key_less(raw, koff, klen, q + 1, (uint32_t)(qe - q - 1))) return false;
            }
        }
    }
    finish_row(raw, r, d, meta, koff, klen, voff, vlen, rb, res);
    return true;
}

// ---- the same reads answered from byte-class masks ------------------------------------------------
// A tile's bytes are classified once, 64 bytes per thread and every thread doing the same thing
// (positioned table look-ups, as in the general scan): one bit per byte and class -- line feed,
// tab, quote, "stop" (a byte that may not stand for itself in a string: control bytes, 0x80 and
// above, \ < > &).  A line then finds its tabs, its quotes and whether anything in it stops a
// string by bit arithmetic on those masks instead of walking its bytes, which is what made the
// byte-wise parser thirty times slower: lanes walking different lines run different code.
constexpr bool ln_plain(uint32_t c) { return c >= 0x20 && c < 0x80 && c != '"' && c != '\\' && c != '<' && c != '>' && c != '&'; }
struct LineLut { uint32_t v[4 * 256]; };
constexpr LineLut make_line_lut() {
    LineLut t{};
    for (uint32_t j = 0; j < 4; j++)
        for (uint32_t c = 0; c < 256; c++)
            t.v[j * 256 + c] = ((c == '\n' ? 1u : 0u) << j) | ((c == '\t' ? 1u : 0u) << (8 + j)) | ((c == '"' ? 1u : 0u) << (16 + j)) |
                               ((!(ln_plain(c) || c == '"') ? 1u : 0u) << (24 + j));   // stop: includes line feed and tab
    return t;
}
__device__ const LineLut g_line_lut = make_line_lut();

__device__ __forceinline__ uint32_t lut_word(const uint32_t *lut, uint32_t x) {
    const char *b = reinterpret_cast<const char *>(lut);
    const uint32_t i0 = (x * 4u) & 0x3fcu, i1 = __umulhi(x, 1u << 26) & 0x3fcu, i2 = __umulhi(x, 1u << 18) & 0x3fcu, i3 = __umulhi(x, 1u << 10) & 0x3fcu;
    return *reinterpret_cast<const uint32_t *>(b + i0) | *reinterpret_cast<const uint32_t *>(b + 1024 + i1) |
           *reinterpret_cast<const uint32_t *>(b + 2048 + i2) | *reinterpret_cast<const uint32_t *>(b + 3072 + i3);
}
struct LineMasks { u64 NL, TAB, Q, STOP; };
__device__ __forceinline__ LineMasks classify_chunk(const uint32_t *s_lut, const uint4 *chunk) {
    uint32_t m[4][2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
        uint32_t y[4];
#pragma unroll
        for (int v = 0; v < 2; v++) {
            const uint4 q = chunk[h * 2 + v];
            y[v * 2 + 0] = lut_word(s_lut, q.x) | (lut_word(s_lut, q.y) << 4);
            y[v * 2 + 1] = lut_word(s_lut, q.z) | (lut_word(s_lut, q.w) << 4);
        }
        // 4 x 4 byte transpose: byte c of y[r] (class c of text bytes 8r .. 8r+7) -> byte r of class word c
        const uint32_t t01l = __byte_perm(y[0], y[1], 0x5140), t23l = __byte_perm(y[2], y[3], 0x5140);
        const uint32_t t01h = __byte_perm(y[0], y[1], 0x7362), t23h = __byte_perm(y[2], y[3], 0x7362);
        m[0][h] = __byte_perm(t01l, t23l, 0x5410);
        m[1][h] = __byte_perm(t01l, t23l, 0x7632);
        m[2][h] = __byte_perm(t01h, t23h, 0x5410);
        m[3][h] = __byte_perm(t01h, t23h, 0x7632);
    }
    LineMasks r;
    r.NL = (u64)m[0][0] | ((u64)m[0][1] << 32);
    r.TAB = (u64)m[1][0] | ((u64)m[1][1] << 32);
    r.Q = (u64)m[2][0] | ((u64)m[2][1] << 32);
    r.STOP = (u64)m[3][0] | ((u64)m[3][1] << 32);
    return r;
}

constexpr int WIN_BEFORE = 512, WIN_AFTER = 256, WIN_BYTES = WIN_BEFORE + LN_TILE + WIN_AFTER, WIN_CHUNKS = WIN_BYTES / 64;
constexpr int TILE_LINES = 2048;   // line ends of a tile listed in shared memory; a tile with more (lines under 4 bytes) is declined

struct MaskText {
    const uint8_t *win;               // the text window in shared memory; win[0] is position lo
    u64 lo, hi;
    const uint8_t *raw;
    const u64 *tabm, *qm, *stopm;     // one word per 64 bytes of the window [w0, w1); w0 may lie before the text (tile 0)
    long long w0, w1;
    mutable bool overrun;             // a question reached past the window: the caller asks the byte-wise reader instead
    __device__ __forceinline__ u64 load8(u64 p) const {
        const u64 a = p & ~7ull;
        u64 x0, x1;
        if (a >= lo && a + 16 <= hi) { const u64 *w = reinterpret_cast<const u64 *>(win + (a - lo)); x0 = w[0]; x1 = w[1]; }
        else { const u64 *w = reinterpret_cast<const u64 *>(raw + a); x0 = __ldg(w); x1 = __ldg(w + 1); }
        const uint32_t sh = (uint32_t)(p & 7u) * 8u;
        return sh ? (x0 >> sh) | (x1 << (64u - sh)) : x0;
    }
    __device__ __forceinline__ uint32_t byte(u64 p) const { return p >= lo && p < hi ? (uint32_t)win[p - lo] : (uint32_t)__ldg(raw + p); }
    // the same for an offset into the window the caller knows to be inside it (16 bytes to spare)
    // (three aligned 32-bit words and two funnel shifts: a 64-bit shift by a variable amount costs several times that)
    __device__ __forceinline__ u64 load8_in(uint32_t off) const {
        const uint32_t *w = reinterpret_cast<const uint32_t *>(win + (off & ~3u));
        const uint32_t x0 = w[0], x1 = w[1], x2 = w[2], sh = (off & 3u) * 8u;
        return (u64)__funnelshift_r(x0, x1, sh) | ((u64)__funnelshift_r(x1, x2, sh) << 32);
    }
    // 64 mask bits from position p on (p inside the window; the arrays have a spare word behind them).  Read as
    // 32-bit words and put together with funnel shifts: a 64-bit shift by a variable amount costs several times that.
    __device__ __forceinline__ u64 bits(const u64 *m, u64 p) const {
        const uint32_t b = (uint32_t)((long long)p - w0), i = b >> 5, sh = b & 31u;
        const uint32_t *m32 = reinterpret_cast<const uint32_t *>(m);
        const uint32_t x0 = m32[i], x1 = m32[i + 1], x2 = m32[i + 2];
        return (u64)__funnelshift_r(x0, x1, sh) | ((u64)__funnelshift_r(x1, x2, sh) << 32);
    }
    // 32 mask bits from window bit b on
    __device__ __forceinline__ uint32_t bits32(const u64 *m, uint32_t b) const {
        const uint32_t *m32 = reinterpret_cast<const uint32_t *>(m);
        return __funnelshift_r(m32[b >> 5], m32[(b >> 5) + 1], b & 31u);
    }
    __device__ __forceinline__ uint32_t tabs(u64 p) const {
        if ((long long)p + 64 > w1) { overrun = true; return 9; }
        const u64 t = bits(tabm, p);
        return min(t == ~0ull ? 64u : (uint32_t)__ffsll((long long)~t) - 1u, 9u);
    }
    __device__ __forceinline__ u64 string_end(u64 p, u64 end) const {
        const u64 lim = (long long)end < w1 ? end : (u64)w1;
        for (u64 a = p; a < lim; a += 64) {
            u64 q = bits(qm, a), st = bits(stopm, a);
            if (lim - a < 64) { const u64 keep = (1ull << (lim - a)) - 1; q &= keep; st &= keep; }
            if (q | st) {
                const uint32_t iq = q ? (uint32_t)__ffsll((long long)q) : 65u, is = st ? (uint32_t)__ffsll((long long)st) : 65u;
                return iq < is ? a + iq - 1 : 0;
            }
        }
        if ((long long)end > w1) overrun = true;
        return 0;
    }
};

// strictly ascending, bytewise, two keys inside the window (offsets into it); a8: the first eight bytes at a
__device__ __forceinline__ bool key_less_in(const MaskText &t, u64 a8, uint32_t a, uint32_t al, uint32_t b, uint32_t bl) {
    const uint32_t m = al < bl ? al : bl;
    u64 x = a8, y = t.load8_in(b);
    for (uint32_t i = 0; i < m; i += 8) {
        if (i) { x = t.load8_in(a + i); y = t.load8_in(b + i); }
        // the first byte in which they differ (32-bit halves: no byte reversal, no 64-bit shifts)
        const uint32_t xl = (uint32_t)x, yl = (uint32_t)y, xh = (uint32_t)(x >> 32), yh = (uint32_t)(y >> 32), dl = xl ^ yl, dh = xh ^ yh;
        if (dl | dh) {
            const uint32_t dw = dl ? dl : dh, xw = dl ? xl : xh, yw = dl ? yl : yh;
            const uint32_t bit = (uint32_t)__ffs((int)dw) - 1u, sh = bit & 24u, at = i + (dl ? 0u : 4u) + (bit >> 3);
            if (at < m) return ((xw >> sh) & 0xffu) < ((yw >> sh) & 0xffu);
            break;   // the same as far as the shorter one goes
        }
    }
    return al < bl;
}
// A line of at most 120 bytes that is not the last one, decided by arithmetic on the class masks: how many
// quotes and where, nothing between the indentation and the end that stops a string, the few bytes
// that must be exactly ':' ' ' ',' or a bracket read by position.  1: accepted, row written; 0: not
// the indented form; 2: not for this function (too long, at the edge of the window, the last line) --
// line_to_row decides.  Same rules as line_to_row, same rows.
__device__ __forceinline__ int line_fast(const MaskText &t, u64 s, u64 e, u64 n, uint32_t r, RowBuffers rb, ParseResult *res) {
    const uint32_t L = (uint32_t)(e - s);
    // (288: every read below -- this line, the start of the next, its key -- stays inside the window)
    if (L == 0 || L > 120 || (long long)s + 288 > t.w1 || s + 288 > t.hi || e >= n) return 2;
    // Straight-line code from here on: what is wrong is collected in `bad`, not returned at once, so that the lanes
    // of a warp -- thirty-two different lines -- stay together and the rows are written by all of them at once.
    const uint8_t *w = t.win + (uint32_t)(s - t.lo);      // w[x]: the byte x places into the line
    const uint32_t so = (uint32_t)(s - t.lo), s32 = (uint32_t)s;   // (the text is under 4 GiB)
    const uint32_t sb = (uint32_t)((long long)s - t.w0);          // the line's first bit in the masks
    const uint32_t d = (uint32_t)__ffs((int)~t.bits32(t.tabm, sb)) - 1u;   // (32 tabs: 0xffffffff -> far too deep)
    bool bad = d > LINES_MAX_DEPTH + 1 || d >= L || (r == 0 && d != 0);
    const u64 Q0 = t.bits(t.qm, s), Q1 = t.bits(t.qm, s + 64);
    const u64 S0 = t.bits(t.stopm, s), S1 = t.bits(t.stopm, s + 64);
    const u64 in0 = low_bits(L), in1 = L > 64 ? low_bits(L - 64) : 0ull;       // the line's own bits
    const bool stops = ((S0 & in0 & ~low_bits(min(d, 63u))) | (S1 & in1)) != 0;   // something that cannot stand for itself: line_to_row looks closer
    u64 q0 = Q0 & in0, q1 = Q1 & in1;
    const uint32_t nq = (uint32_t)__popcll(q0) + (uint32_t)__popcll(q1);
    bad = bad || (nq != 0 && nq != 2 && nq != 4);
    // Of the quotes only two places matter: whether one stands right after the indentation (then it is the first:
    // tabs are no quotes) and where the next one is (the end of the key).  A string value then starts with a quote
    // by the look of its first byte and must end with one right before the end of the line or its comma; that
    // there are no others is a matter of counting.  127: no such quote (inside the window all the same).
    const bool quote_at_d = ((q0 >> min(d, 63u)) & 1ull) != 0;
    const u64 r0 = q0 & ~low_bits(min(d, 63u) + 1u);
    const uint32_t p1 = r0 ? (uint32_t)__ffsll((long long)r0) - 1u : (q1 ? 64u + (uint32_t)__ffsll((long long)q1) - 1u : 127u);
    // "key": -- a quote right after the indentation, the next quote followed by a colon and a space
    const bool key_shape = nq >= 2 && quote_at_d && p1 + 1 < L && w[p1 + 1] == ':';
    bad = bad || (key_shape && (p1 + 2 >= L || w[p1 + 2] != ' ' || p1 + 3 >= L));
    const uint32_t used = key_shape ? 2u : 0u;                  // quotes accounted for
    const uint32_t v = key_shape ? min(p1 + 3, 127u) : min(d, 127u);   // where the value starts
    const uint32_t koff = key_shape ? s32 + d + 1 : 0u, klen = key_shape ? p1 - d - 1 : 0u;
    const uint32_t c = w[v], c1 = w[v + 1];
    const bool is_str = c == '"', is_brace = c == '{', is_brack = c == '[', is_cl = c == '}' || c == ']';
    const bool is_empty = (is_brace && c1 == '}' && v + 1 < L) || (is_brack && c1 == ']' && v + 1 < L);
    const bool is_open = (is_brace || is_brack) && !is_empty, is_scalar = !(is_str || is_brace || is_brack || is_cl);
    const bool trail = w[L - 1] == ',';
    const uint32_t qb = L - 1u - (trail ? 1u : 0u);                             // the value's closing quote, if it is a string
    bad = bad || (is_str ? (nq != used + 2 || qb <= v || qb >= L || w[min(qb, 127u)] != '"') : nq != used);   // a quote where no string can be, or one too few
    bad = bad || (is_cl && key_shape);
    uint32_t kind, after, voff = 0, vlen = 0;
    if (is_str) { kind = ROW_STR | META_VAL_PLAIN; voff = s32 + v + 1; vlen = qb - v - 1; after = qb + 1; }
    else if (is_empty) { kind = is_brace ? ROW_EMPTY_OBJ : ROW_EMPTY_ARR; vlen = s32 + v; after = v + 2; }
    else if (is_open) { kind = is_brace ? ROW_OPEN_OBJ : ROW_OPEN_ARR; vlen = s32 + v; after = v + 1; }
    else if (is_cl) { kind = c == '}' ? ROW_CLOSE_OBJ : ROW_CLOSE_ARR; vlen = s32 + v; after = v + 1; }
    else { kind = ROW_RAW; after = trail ? L - 1 : L; voff = s32 + v; vlen = after - v; }
    if (is_scalar && !bad) bad = after <= v || !canonical_scalar(t, s + v, s + after);   // (container rows keep the position of their bracket in vlen)
    const bool comma = after < L && w[min(after, 127u)] == ',';
    after += comma ? 1u : 0u;
    bad = bad || after != L || (comma && is_open);
    uint32_t meta = kind | (key_shape ? META_HAS_KEY | META_KEY_PLAIN : 0u) | (comma ? META_COMMA : 0u) | (d << META_DEPTH_SHIFT);
    // ---- this line against the next one: indentation, comma, key order
    const uint32_t nb = L + 1;                                             // where the next line starts, from s
    const uint32_t dn = (uint32_t)__ffs((int)~t.bits32(t.tabm, sb + nb)) - 1u;
    bad = bad || dn > LINES_MAX_DEPTH + 1 || e + 1 + dn >= n;
    const uint32_t cn = w[nb + min(dn, 16u)];
    const bool next_closes = cn == '}' || cn == ']';
    bad = bad || (is_open ? (dn != d + 1 || next_closes) : next_closes ? (d == 0 || dn != d - 1 || comma) : (dn != d || !comma));
    const u64 k8 = t.load8_in(so + d + 1);   // the key's first bytes (or whatever stands there: used only if key_shape)
    int verdict = 1;
    if (!bad && key_shape && !is_open && !next_closes) {
        // the next member of the same object: its key -- up to the next quote; whether that line is well formed is
        // its own thread's business -- must sort strictly after this one
        const uint32_t from = nb + dn + 1;                                // first byte of the next key, from s
        const u64 qn = t.bits(t.qm, s + from);                             // (from <= 131: inside the 288 bytes the window has behind s)
        if (cn != '"') bad = true;
        else if (!qn) verdict = 2;
        else bad = !key_less_in(t, k8, so + d + 1, klen, so + from, (uint32_t)__ffsll((long long)qn) - 1u);
    }
    if (bad) return 0;
    if (verdict == 2 || stops) return 2;
    // the row, with what the module index and the verdict want to know about it (finish_row, from the window)
    if (d == 3 && (kind & 15u) == ROW_STR && key_shape) {
        if (klen == 4 && (uint32_t)k8 == 0x656d616eu) meta |= META_KEY_IS_NAME;                    // "name"
        else if (klen == 8 && k8 == 0x656d616e74736f68ull) meta |= META_KEY_IS_HOSTNAME;           // "hostname"
    }
    rb.meta[r] = meta; rb.koff[r] = koff; rb.klen[r] = klen; rb.voff[r] = voff; rb.vlen[r] = vlen;
    note_opener(rb, r, kind & 15u, d);
    if (r == 0) res->root_kind = kind & 15u;
    if (d == 1 && key_shape && klen == 6 && (k8 & 0xffffffffffffull) == 0x656c75646f6dull) {       // "module"
        res->module_row = r + 1;
        res->module_kind = kind & 15u;
    }
    return 1;
}

constexpr int LR_THREADS = 192;   // 128 classify the tile (the first 12 also its margins); about 186 lines of config #4 per tile: one line each
__global__ void __launch_bounds__(LR_THREADS, 8) lines_rows_kernel(const uint8_t *__restrict__ raw, u64 n, u64 readable_end, uint32_t n_tiles,
                                                               const uint32_t *__restrict__ first_line, const uint32_t *__restrict__ last_end,
                                                               RowBuffers rb, uint32_t row_cap, ParseResult *res) {
    __shared__ uint32_t s_lut[4 * 256];
    __shared__ __align__(16) uint8_t s_text[WIN_BYTES];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ u64 s_tab[WIN_CHUNKS + 1], s_q[WIN_CHUNKS + 1], s_stop[WIN_CHUNKS + 1];   // (a spare word: MaskText::bits reads one ahead)
    __shared__ uint16_t s_end[TILE_LINES];
    __shared__ uint32_t s_warp[LR_THREADS / 32];
    __shared__ u64 s_prev_end;      // position + 1 of the last line end before the tile
    for (int i = threadIdx.x; i < 4 * 256; i += LR_THREADS) s_lut[i] = g_line_lut.v[i];   // (before the wait: nothing of the kernels before)
    griddep_wait();
    if (res->fast_reject) return;
    if (res->n_rows > row_cap) { if (threadIdx.x == 0 && blockIdx.x == 0) res->fast_reject = 1; return; }
    if (threadIdx.x == 0) { s_tab[WIN_CHUNKS] = 0; s_q[WIN_CHUNKS] = 0; s_stop[WIN_CHUNKS] = 0; }
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t parity = 0;
    bool ok = true;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const u64 tile0 = (u64)tile * LN_TILE;
        // the tile, the end of the line that runs into it and the start of the one that runs out of it: one bulk copy
        const u64 lo = tile0 >= WIN_BEFORE ? tile0 - WIN_BEFORE : 0, hi = min(tile0 + LN_TILE + WIN_AFTER, readable_end);
        const long long w0 = (long long)tile0 - WIN_BEFORE;
        if (threadIdx.x == 0) {
            bulk_load(s_text, raw + lo, (uint32_t)(hi - lo), &s_bar);
            s_prev_end = (u64)last_end[tile];
        }
        mbar_wait(&s_bar, parity);
        parity ^= 1u;
        // classes of every 64 bytes of the window: a chunk per thread, the few chunks of the margins by the first threads
        u64 my_ends = 0;
        for (int j = threadIdx.x; j < WIN_CHUNKS; j += LR_THREADS) {
            const int wc = j < LN_THREADS ? j + WIN_BEFORE / 64 : (j - LN_THREADS < WIN_BEFORE / 64 ? j - LN_THREADS : j);   // the tile's chunks first
            const long long at = w0 + (long long)wc * 64;
            LineMasks m = {0, 0, 0, 0};
            if (at >= 0 && (u64)at + 64 <= hi) {
                m = classify_chunk(s_lut, reinterpret_cast<const uint4 *>(s_text + ((u64)at - lo)));
                if ((u64)at + 64 > n) {   // past the end of the text nothing counts; the end itself ends the last line
                    const u64 keep = (u64)at >= n ? 0ull : low_bits((uint32_t)(n - (u64)at));
                    m.NL &= keep; m.TAB &= keep; m.Q &= keep; m.STOP &= keep;
                    if ((u64)at <= n) { m.NL |= 1ull << (n - (u64)at); m.STOP |= 1ull << (n - (u64)at); }
                }
            }
            s_tab[wc] = m.TAB; s_q[wc] = m.Q; s_stop[wc] = m.STOP;
            if (j < LN_THREADS) my_ends = m.NL;
        }
        // the tile's line ends listed in order
        const uint32_t c = (uint32_t)__popcll(my_ends);
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, incl, o);
            if (lane >= (uint32_t)o) incl += y;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t before = incl - c, total = 0;
        for (uint32_t w = 0; w < LR_THREADS / 32; w++) { if (w < warp) before += s_warp[w]; total += s_warp[w]; }
        {
            uint32_t k = before;
            u64 left = my_ends;
#pragma unroll
            for (int it = 0; it < 2; it++) {   // (a 64-byte chunk of config #4 holds one or two line ends: no loop for those)
                if (left && k < TILE_LINES) s_end[k] = (uint16_t)(threadIdx.x * LN_CHUNK + (uint32_t)__ffsll((long long)left) - 1);
                k += left ? 1u : 0u;
                left &= left - 1;
            }
            for (; left; left &= left - 1, k++)
                if (k < TILE_LINES) s_end[k] = (uint16_t)(threadIdx.x * LN_CHUNK + (uint32_t)__ffsll((long long)left) - 1);
        }
        __syncthreads();
        if (total > TILE_LINES) { ok = false; total = 0; }
        // a line per thread, in order
        const uint32_t row0 = first_line[tile];
        for (uint32_t k = threadIdx.x; k < total; k += LR_THREADS) {
            const u64 e = tile0 + s_end[k], s = k ? tile0 + s_end[k - 1] + 1 : s_prev_end;
            bool done = false;
            if ((long long)s >= w0 && s >= lo) {
                const MaskText mt{s_text, lo, hi, raw, s_tab, s_q, s_stop, w0, w0 + WIN_BYTES, false};
                const int fast = line_fast(mt, s, e, n, row0 + k, rb, res);
                if (fast != 2) { ok = ok && fast == 1; done = true; }
                else {
                    const bool good = line_to_row(mt, s, e, n, row0 + k, rb, res);
                    if (!mt.overrun) { ok = ok && good; done = true; }
                }
            }
            if (!done) {   // a line that starts before the window, or looks past it: byte by byte
                const Text text{s_text, lo, hi, raw};
                ok = ok && line_to_row(text, s, e, n, row0 + k, rb, res);
            }
        }
        __syncthreads();   // (the window and the lists are rewritten by the next tile)
    }
    if (!ok) res->fast_reject = 1;
}

// ---- stage 2: closers and openers, containers -----------------------------------------------
// The state that travels along the rows: for every depth k <= LINES_MAX_DEPTH the last opening row
// seen at that depth, as (row + 1) << 1 | is-an-object (0: none yet).  Later rows have larger
// numbers, so "the last one" is a maximum: the state of a stretch of rows is combined by max.
// Stage 1 has left that maximum per block of LN_ROWS_PER_BLOCK rows (note_opener).
constexpr int ST_DEPTHS = LINES_MAX_DEPTH + 1;
constexpr int SR_THREADS = 256, SR_ITEMS = LN_ROWS_PER_BLOCK / SR_THREADS;

// in place: agg[b][k] -> the last opening row at depth k in the blocks before b.  One block: thread
// (c, k) folds depth k over the c-th of 128 runs of blocks -- its loads are independent of each
// other, so one memory round trip for all of them --, a warp per depth scans the 128 run totals.
constexpr int OP_RUNS = 128;
__global__ void __launch_bounds__(OP_RUNS * ST_DEPTHS) lines_open_prefix_kernel(uint32_t *agg, ParseResult *res) {
    __shared__ uint32_t s_top[ST_DEPTHS][OP_RUNS];
    griddep_wait();
    const uint32_t n_blocks = (res->n_rows + LN_ROWS_PER_BLOCK - 1) / LN_ROWS_PER_BLOCK;
    if (res->fast_reject) {   // what the blocks of the next launch go by (they set fast_reject themselves)
        if (threadIdx.x == 0) res->reject_early = 1;
        return;
    }
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, k = threadIdx.x % ST_DEPTHS, c = threadIdx.x / ST_DEPTHS;
    const uint32_t per = (n_blocks + OP_RUNS - 1) / OP_RUNS, b0 = min(c * per, n_blocks), b1 = min(b0 + per, n_blocks);
    uint32_t top = 0;
#pragma unroll 4
    for (uint32_t b = b0; b < b1; b++) top = max(top, agg[b * ST_DEPTHS + k]);
    s_top[k][c] = top;
    __syncthreads();
    if (warp < ST_DEPTHS) {
        constexpr int PER_LANE = OP_RUNS / 32;
        uint32_t v[PER_LANE], mine = 0;
#pragma unroll
        for (int j = 0; j < PER_LANE; j++) { v[j] = s_top[warp][lane * PER_LANE + j]; mine = max(mine, v[j]); }
        uint32_t incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, incl, o);
            if (lane >= (uint32_t)o) incl = max(incl, y);
        }
        uint32_t x = __shfl_up_sync(FULL, incl, 1);
        if (lane == 0) x = 0;
#pragma unroll
        for (int j = 0; j < PER_LANE; j++) { s_top[warp][lane * PER_LANE + j] = x; x = max(x, v[j]); }
    }
    __syncthreads();
    uint32_t x = s_top[k][c];
#pragma unroll 4
    for (uint32_t b = b0; b < b1; b++) {
        const uint32_t v = agg[b * ST_DEPTHS + k];
        agg[b * ST_DEPTHS + k] = x;
        x = max(x, v);
    }
}

// The block's 1,024 rows as bit masks: for every depth the rows that open a container there (a
// word per 32 rows, made by four ballots), and which of those words are non-zero.  "The last
// opener at depth k before row r" is then two count-leading-zeros away, or the block's prefix.
//
// The module index's field lists (index.cuh) are made here as well, since every row has just been told
// its container: a "name" / "hostname" string three levels down (marked by stage 1) whose container is
// an object whose container is the root's "module".  Per block the hits are listed in row order (ballots
// again), sized, and placed in the lists by a decoupled look-back over the blocks -- which therefore take
// their number from a ticket counter, so that a block only ever waits for blocks that have started.
__device__ __forceinline__ u64 ld_relaxed_u64(const u64 *p) {
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(u64 *p, u64 v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
constexpr int FL_CAP = LN_ROWS_PER_BLOCK / 2;   // a hit needs the row of its own child as well: at most 512 per field and block
__global__ void __launch_bounds__(SR_THREADS, 6) lines_resolve_kernel(const uint8_t *__restrict__ raw, RowBuffers rb, const uint32_t *__restrict__ prefix,
                                                                  FieldIndexBuffers fx, ParseResult *res) {
    constexpr int WORDS = LN_ROWS_PER_BLOCK / 32;
    static_assert(WORDS == 32 && SR_THREADS / 32 == ST_DEPTHS, "a warp per depth folds the 32 words of that depth");
    __shared__ uint32_t s_open[ST_DEPTHS][WORDS], s_obj[WORDS], s_nz[ST_DEPTHS], s_before[ST_DEPTHS], s_meta[LN_ROWS_PER_BLOCK], s_jobs;
    __shared__ uint2 s_job[LN_ROWS_PER_BLOCK / 2];   // (a job needs a closing row before it)
    __shared__ uint32_t s_block, s_fmask[2][WORDS], s_fpre[2][WORDS], s_fcnt[2];
    __shared__ uint16_t s_frow[2][FL_CAP];            // the block's hits per field, in row order: the row (local), ...
    __shared__ uint32_t s_fchild[2][FL_CAP], s_fpos[2][FL_CAP];   // ... its child's row, the bytes of the records before it in the block
    __shared__ u64 s_fbase[2];                        // records before the block | record bytes before it << 32
    griddep_wait();
    const uint32_t n_rows = res->n_rows, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t n_blocks = (n_rows + LN_ROWS_PER_BLOCK - 1) / LN_ROWS_PER_BLOCK;
    // (reject_early, not fast_reject: that one may be set by a block of this very launch, and a block that left
    //  because of it would be waited for by its successors for ever)
    if (res->reject_early) return;
    bool ok = true;
    // Persistent blocks: as many as fit on the device at once, each taking the next 1,024 rows from the ticket
    // counter until none are left (measured the same as one short-lived block per 1,024 rows; kept because a
    // block then only ever waits for tickets that running blocks hold).
    for (;;) {
        __syncthreads();   // (the lists in shared memory are written again)
        if (threadIdx.x == 0) s_block = (uint32_t)atomicAdd(fx.status, 1ull);
        __syncthreads();
        const uint32_t block = s_block;
        if (block >= n_blocks) break;
        const uint32_t block0 = block * LN_ROWS_PER_BLOCK;
        uint32_t meta[SR_ITEMS];
#pragma unroll
        for (int j = 0; j < SR_ITEMS; j++) {   // rows j * 256 + thread: a warp's 32 rows are one word of the masks
            const uint32_t rl = j * SR_THREADS + threadIdx.x, r = block0 + rl;
            meta[j] = r < n_rows ? __ldg(rb.meta + r) : 0xffffffffu;
            s_meta[rl] = meta[j];
            const uint32_t kind = meta[j] & 15u, d = meta[j] >> META_DEPTH_SHIFT;
            const bool opens = r < n_rows && (kind == ROW_OPEN_OBJ || kind == ROW_OPEN_ARR) && d <= LINES_MAX_DEPTH;
            const uint32_t op = __ballot_sync(FULL, opens), b0 = __ballot_sync(FULL, d & 1u), b1 = __ballot_sync(FULL, d & 2u),
                           b2 = __ballot_sync(FULL, d & 4u), ob = __ballot_sync(FULL, kind == ROW_OPEN_OBJ);
            if (lane < ST_DEPTHS)
                s_open[lane][j * (SR_THREADS / 32) + warp] = op & (lane & 1u ? b0 : ~b0) & (lane & 2u ? b1 : ~b1) & (lane & 4u ? b2 : ~b2);
            if (lane == ST_DEPTHS) s_obj[j * (SR_THREADS / 32) + warp] = ob;
        }
        if (threadIdx.x < ST_DEPTHS) s_before[threadIdx.x] = __ldg(prefix + block * ST_DEPTHS + threadIdx.x);
        if (threadIdx.x == 0) s_jobs = 0;
        __syncthreads();
        {
            const uint32_t nz = __ballot_sync(FULL, s_open[warp][lane] != 0);
            if (lane == 0) s_nz[warp] = nz;
        }
        __syncthreads();
        // (row + 1) << 1 | is-an-object of the last opener at depth k before local row rl; 0: none
        auto state_at = [&](uint32_t k, uint32_t rl) -> uint32_t {
            uint32_t w = rl >> 5, m = s_open[k][w] & ((1u << (rl & 31u)) - 1u);
            if (!m) {
                const uint32_t nz = s_nz[k] & ((1u << w) - 1u);
                if (!nz) return s_before[k];
                w = 31u - (uint32_t)__clz((int)nz);
                m = s_open[k][w];
            }
            const uint32_t at = 31u - (uint32_t)__clz((int)m);
            return ((block0 + w * 32u + at + 1u) << 1) | ((s_obj[w] >> at) & 1u);
        };
        auto s_meta_of = [&](uint32_t row) -> uint32_t { return row >= block0 ? s_meta[row - block0] : __ldg(rb.meta + row); };
        const uint32_t module_row = res->module_row;   // 1 + row, 0: none (stage 1 has finished)
        // "module" as the index wants it: an object member of a root object; as a container: (row + 1) << 1 | 1
        const uint32_t module_enc = res->root_kind == ROW_OPEN_OBJ && module_row != 0 && res->module_kind == ROW_OPEN_OBJ ? (module_row << 1) | 1u : 0u;
        // ---- the module index's hits first: their number and sizes are what the blocks after this one wait for
        uint32_t hit[SR_ITEMS];   // child row << 2 | field (0 "name", 1 "hostname", 2 none)
#pragma unroll
        for (int j = 0; j < SR_ITEMS; j++) {
            const uint32_t rl = j * SR_THREADS + threadIdx.x, r = block0 + rl, m = meta[j];
            hit[j] = 2u;
            if (r < n_rows && (m & (META_KEY_IS_NAME | META_KEY_IS_HOSTNAME)) && (m >> META_DEPTH_SHIFT) == 3 && module_enc) {
                const uint32_t o = state_at(2, rl);   // its container: an object, whose container is "module"
                if (o != 0 && (o & 1u) && state_at(1, rl) == module_enc) hit[j] = (((o >> 1) - 1) << 2) | ((m & META_KEY_IS_HOSTNAME) ? 1u : 0u);
            }
            const uint32_t h0 = __ballot_sync(FULL, (hit[j] & 3u) == 0u), h1 = __ballot_sync(FULL, (hit[j] & 3u) == 1u);
            if (lane == 0) { s_fmask[0][j * (SR_THREADS / 32) + warp] = h0; s_fmask[1][j * (SR_THREADS / 32) + warp] = h1; }
        }
        __syncthreads();

        // ---- the module index's lists
        if (warp < 2) {   // hits before every 32-row word, per field
            const uint32_t c = (uint32_t)__popc(s_fmask[warp][lane]);
            uint32_t incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(FULL, incl, o);
                if (lane >= (uint32_t)o) incl += y;
            }
            s_fpre[warp][lane] = incl - c;
            if (lane == 31) s_fcnt[warp] = incl;
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < SR_ITEMS; j++) {
            const uint32_t f = hit[j] & 3u;
            if (f == 2u) continue;
            const uint32_t w = j * (SR_THREADS / 32) + warp, at = s_fpre[f][w] + (uint32_t)__popc(s_fmask[f][w] & ((1u << lane) - 1u));
            if (at < (uint32_t)FL_CAP) { s_frow[f][at] = (uint16_t)(j * SR_THREADS + threadIdx.x); s_fchild[f][at] = hit[j] >> 2; }
        }
        __syncthreads();
        const uint32_t cnt0 = min(s_fcnt[0], (uint32_t)FL_CAP), cnt1 = min(s_fcnt[1], (uint32_t)FL_CAP);   // (the cap cannot be met: see FL_CAP)
        for (uint32_t i = threadIdx.x; i < cnt0 + cnt1; i += SR_THREADS) {
            const uint32_t f = i >= cnt0 ? 1u : 0u, k = i - (f ? cnt0 : 0u);
            s_fpos[f][k] = 8u + __ldg(rb.vlen + block0 + s_frow[f][k]) + __ldg(rb.klen + s_fchild[f][k]);   // { u32 len, field, u32 len, key }
        }
        __syncthreads();
        if (warp < 2) {
            // warp 0 carries field 0, warp 1 field 1: record bytes before each hit of the block, then the block's
            // place among the blocks.  Payload of the look-back: records (30 bits) | record bytes (32 bits), as in index.cu
            const uint32_t f = warp, cnt = f ? cnt1 : cnt0;
            u64 run = 0;
            for (uint32_t base = 0; base < cnt; base += 32) {
                const uint32_t k = base + lane, v = k < cnt ? s_fpos[f][k] : 0u;
                u64 incl = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const u64 y = __shfl_up_sync(FULL, incl, o);
                    if (lane >= (uint32_t)o) incl += y;
                }
                if (k < cnt) s_fpos[f][k] = (uint32_t)(run + incl - v);
                run += __shfl_sync(FULL, incl, 31);
            }
            constexpr u64 ST_AGG = 1ull << 62, ST_PREFIX = 2ull << 62, ST_MASK = (1ull << 62) - 1;
            const u64 agg = ((u64)cnt & 0x3fffffffull) | ((run & 0xffffffffull) << 30);
            u64 *status = fx.status + 2 + (size_t)f * n_blocks;
            u64 before = 0;
            if (block == 0) {
                if (lane == 0) st_relaxed_u64(status, ST_PREFIX | agg);
            } else {
                if (lane == 0) st_relaxed_u64(status + block, ST_AGG | agg);
                int look = (int)block - 1;   // (128 status words per step instead of 32 was measured: slower, 50 us against 44)
                for (;;) {
                    const int idx = look - (int)lane;
                    u64 st = idx >= 0 ? ld_relaxed_u64(status + idx) : ST_PREFIX;
                    while (__any_sync(FULL, (st >> 62) == 0)) { if ((st >> 62) == 0) st = ld_relaxed_u64(status + idx); }
                    const unsigned pm = __ballot_sync(FULL, (st >> 62) == 2);
                    const int first = pm ? __ffs(pm) - 1 : 32;
                    u64 x = ((int)lane <= first) ? (st & ST_MASK) : 0ull;
#pragma unroll
                    for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(FULL, x, o);
                    before += x;
                    if (pm) break;
                    look -= 32;
                }
                if (lane == 0) st_relaxed_u64(status + block, ST_PREFIX | ((before + agg) & ST_MASK));
            }
            if (lane == 0) {
                s_fbase[f] = (before & 0x3fffffffull) | (((before >> 30) & 0xffffffffull) << 32);
                if (run) atomicAdd(&res->field_bytes[f], run);      // (the 64-bit total: a blob of 4 GiB or more is refused)
                if (block + 1 == n_blocks) {
                    const uint32_t m = (uint32_t)((before + agg) & 0x3fffffffull);
                    res->field_count[f] = m;
                    if (m <= (f ? fx.cap[1] : fx.cap[0])) (f ? fx.off[1] : fx.off[0])[m] = (uint32_t)(((before + agg) >> 30) & 0xffffffffull);
                }
            }
        } else {
            // ---- meanwhile the other six warps: every row of the block against its container
            constexpr uint32_t WORK_WARPS = SR_THREADS / 32 - 2;
            for (uint32_t wd = warp - 2; wd < (uint32_t)WORDS; wd += WORK_WARPS) {
                const uint32_t rl = wd * 32u + lane, r = block0 + rl;
                if (r >= n_rows) continue;
                const uint32_t m = s_meta[rl], kind = m & 15u, d = m >> META_DEPTH_SHIFT;
                const bool closes = kind == ROW_CLOSE_OBJ || kind == ROW_CLOSE_ARR;
                // one look-up for every row: a closer wants its opener (same depth), a value its container (one up)
                const uint32_t dq = closes ? d : d - 1u;
                const uint32_t o = dq <= LINES_MAX_DEPTH ? state_at(dq, rl) : 0u;
                if (closes) {
                    // the container it closes opened at the same depth, and was of the same kind
                    if (o == 0 || (o & 1u) != (kind == ROW_CLOSE_OBJ ? 1u : 0u)) ok = false;
                    else {
                        const uint32_t orow = (o >> 1) - 1;
                        rb.koff[r] = orow;        // (TapeView::rows_direct: rows, not bracket numbers)
                        rb.voff[orow] = r;
                        if (module_row && orow == module_row - 1) res->module_end = r;
                        // key order across the container that has just closed: if the next row is a member of the same
                        // object, the member before it is the one that opened here.  About one row in fourteen at
                        // config #4: listed, and compared below by neighbouring threads
                        const uint32_t nm = rl + 1 < (uint32_t)LN_ROWS_PER_BLOCK ? s_meta[rl + 1] : (r + 1 < n_rows ? __ldg(rb.meta + r + 1) : 0xffffffffu);
                        const uint32_t nk = nm & 15u;
                        if (nm != 0xffffffffu && nk != ROW_CLOSE_OBJ && nk != ROW_CLOSE_ARR && (nm & META_HAS_KEY) && (nm >> META_DEPTH_SHIFT) == d) {
                            const uint32_t at = atomicAdd(&s_jobs, 1u);
                            s_job[at] = make_uint2(r + 1, orow);
                        }
                    }
                } else if (d == 0) {
                    if (r != 0) ok = false;       // at the top there is exactly one value
                } else if (o == 0 || ((o & 1u) != 0) != ((m & META_HAS_KEY) != 0)) {
                    ok = false;                   // inside an object a value has a key, inside an array none
                }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(WORK_WARPS * 32) : "memory");   // (the six warps only: the job list is complete)
            for (uint32_t i = threadIdx.x - 64u; i < s_jobs; i += WORK_WARPS * 32) {
                const uint32_t r = s_job[i].x, orow = s_job[i].y;
                const uint32_t ao = __ldg(rb.koff + orow), al = __ldg(rb.klen + orow), bo = __ldg(rb.koff + r), bl = __ldg(rb.klen + r);
                if (!(s_meta_of(orow) & META_HAS_KEY) || !key_less(GlobalText{raw}, ao, al, bo, bl)) ok = false;
            }
            if (!ok) res->fast_reject = 1;
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < cnt0 + cnt1; i += SR_THREADS) {
            const uint32_t f = i >= cnt0 ? 1u : 0u, k = i - (f ? cnt0 : 0u), r = block0 + s_frow[f][k], c = s_fchild[f][k];
            const uint32_t at = (uint32_t)s_fbase[f] + k;
            if (at < (f ? fx.cap[1] : fx.cap[0])) {   // (selects, not fx.cap[f]: indexing a kernel parameter by a run-time value copies it to local memory)
                (f ? fx.refs[1] : fx.refs[0])[at] = FieldRef{__ldg(rb.voff + r), __ldg(rb.vlen + r), __ldg(rb.koff + c), __ldg(rb.klen + c)};
                (f ? fx.off[1] : fx.off[0])[at] = (uint32_t)(s_fbase[f] >> 32) + s_fpos[f][k];
            }
        }
    }   // next 1,024 rows
}

}  // namespace

cudaError_t launch_lines_count(uint8_t *raw, const uint8_t *src, size_t n, uint32_t *scratch, uint32_t *open_state, uint32_t row_cap,
                               ParseResult *res, cudaStream_t s) {
    const uint32_t tiles = lines_tiles(n);
    cudaError_t e = cudaMemsetAsync(res, 0, sizeof(ParseResult), s);
    if (e != cudaSuccess) return e;
    if (src) lines_count_kernel<true><<<tiles, LN_THREADS, 0, s>>>(raw, src, (u64)n, scratch, scratch + tiles, open_state, (uint32_t)lines_open_state_words(row_cap));
    else lines_count_kernel<false><<<tiles, LN_THREADS, 0, s>>>(raw, nullptr, (u64)n, scratch, scratch + tiles, open_state, (uint32_t)lines_open_state_words(row_cap));
    if ((e = launch_dependent(lines_prefix_kernel, dim3(1), dim3(1024), s, scratch, scratch + tiles, tiles, (u64)n, res)) != cudaSuccess) return e;
    count_parse_launch(); count_parse_launch();
    return cudaGetLastError();
}

cudaError_t launch_lines_rows(const uint8_t *raw, size_t n, const uint32_t *scratch, RowBuffers rb, uint32_t row_cap, ParseResult *res,
                              cudaStream_t s) {
    const uint32_t tiles = lines_tiles(n);
    // (the text is padded to a whole tile of the general scan and 64 bytes more: engine.cu, upload_text)
    const size_t readable_end = (size_t)p1_tiles(n) * P1_TILE + 64;
    static std::atomic<uint32_t> cached[64];   // resident blocks of the persistent kernel, per device
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    uint32_t grid = dev >= 0 && dev < 64 ? cached[dev].load() : 0;
    if (!grid) {
        int per_sm = 0, sms = 0;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lines_rows_kernel, LR_THREADS, 0)) != cudaSuccess) return e;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        grid = (uint32_t)(per_sm > 0 ? per_sm : 1) * (uint32_t)(sms > 0 ? sms : 1);
        if (dev >= 0 && dev < 64) cached[dev].store(grid);
    }
    if ((e = launch_dependent(lines_rows_kernel, dim3(std::min(tiles, grid)), dim3(LR_THREADS), s, raw, (u64)n, (u64)readable_end, tiles, scratch, scratch + tiles, rb, row_cap,
                              res)) != cudaSuccess) return e;
    count_parse_launch();
    return cudaGetLastError();
}

cudaError_t launch_lines_structure(const uint8_t *raw, RowBuffers rb, uint32_t row_cap, uint32_t *open_state, FieldIndexBuffers fx, ParseResult *res,
                                   cudaStream_t s) {
    const uint32_t blocks = lines_row_blocks(row_cap);
    if (!blocks) return cudaSuccess;
    // (the ticket counter and the look-back words of the field lists sit behind the open-container words: cleared by stage 0)
    fx.status = reinterpret_cast<unsigned long long *>(open_state + (size_t)blocks * ST_DEPTHS);
    cudaError_t e0 = launch_dependent(lines_open_prefix_kernel, dim3(1), dim3(OP_RUNS * ST_DEPTHS), s, open_state, res);
    if (e0 != cudaSuccess) return e0;
    static std::atomic<uint32_t> cached[64];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    uint32_t grid = dev >= 0 && dev < 64 ? cached[dev].load() : 0;
    if (!grid) {
        int per_sm = 0, sms = 0;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lines_resolve_kernel, SR_THREADS, 0)) != cudaSuccess) return e;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        grid = (uint32_t)(per_sm > 0 ? per_sm : 1) * (uint32_t)(sms > 0 ? sms : 1);
        if (dev >= 0 && dev < 64) cached[dev].store(grid);
    }
    if ((e = launch_dependent(lines_resolve_kernel, dim3(std::min(blocks, grid)), dim3(SR_THREADS), s, raw, rb, open_state, fx, res)) != cudaSuccess) return e;
    count_parse_launch(); count_parse_launch();
    return cudaGetLastError();
}

}  // namespace tk
