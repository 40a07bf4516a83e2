// parse.cu -- sm_100a kernels of state.New (see parse.cuh for the pipeline).
//
// Reference semantics restated: Go encoding/json's scanner (validity), unquote (string
// decoding: escapes, surrogate pairs, invalid UTF-8 -> U+FFFD) and objectInterface (member
// order irrelevant, last duplicate wins -- the latter is resolved on the host when the GPU
// reports that keys are not strictly ascending).
#include "parse.cuh"
#include "unquote.cuh"

#include <algorithm>
#include <atomic>

namespace tk {

static std::atomic<unsigned long long> g_parse_launches{0};
unsigned long long parse_launch_count() { return g_parse_launches.load(); }

namespace {

constexpr unsigned FULL = 0xffffffffu;
typedef unsigned long long u64;

// ---- status words for decoupled look-back -----------------------------------------------------
constexpr u64 ST_AGG = 1ull << 62, ST_PREFIX = 2ull << 62, ST_MASK = (1ull << 62) - 1;

__device__ __forceinline__ u64 ld_relaxed(const u64 *p) {
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(u64 *p, u64 v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ u64 low_mask(int i) { return i >= 64 ? ~0ull : ((1ull << i) - 1); }

// =================================================================================================
// P1: bytes -> tokens
// =================================================================================================

// ---- byte classes of the carry-protocol kernel (scan_tokens_kernel): one flag byte per text byte
constexpr uint32_t C_QUOTE = 1, C_BS = 2, C_WS = 4, C_ST = 8, C_CTRL = 16, C_HTML = 32;

__device__ __forceinline__ uint32_t byte_class(uint32_t c) {
    uint32_t k = 0;
    if (c == '"') k |= C_QUOTE;
    if (c == '\\') k |= C_BS;
    if (c == ' ' || c == '\t' || c == '\n' || c == '\r') k |= C_WS;
    if (c == '{' || c == '}' || c == '[' || c == ']' || c == ':' || c == ',') k |= C_ST;
    if (c < 0x20 && c != '\t' && c != '\n' && c != '\r') k |= C_CTRL;  // tab, newline, return: found through C_WS
    if (c == '<' || c == '>' || c == '&') k |= C_HTML;  // escaped on output: such strings are not plain
    return k;
}

// 4 per-byte flag bits (bit j of every byte of x) gathered into a nibble
__device__ __forceinline__ uint32_t gather4(uint32_t x, int j) {
    return (((x >> j) & 0x01010101u) * 0x01020408u) >> 24;
}

// ---- byte classes of the fast kernel (scan_tokens_fast_kernel).
// The kernel needs, per 64-byte chunk, one 64-bit mask per class.  Each class flag
// of a byte is looked up ALREADY AT ITS PLACE: table j (for byte j of a 32-bit word) holds the
// quote flag at bit j, white space at bit 8 + j, structural at bit 16 + j and "rare" at bit 24 + j,
// so OR-ing the four look-ups of a word gives one nibble per class, an odd word is shifted left by
// 4, and a byte permute of four such pairs yields 32 mask bits per class without any bit gathering.
// "Rare" bytes (backslash, control bytes other than tab/newline/return, < > &, non-ASCII) are told
// apart afterwards, one by one, only in chunks that have any.
constexpr bool is_ws(uint32_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
constexpr bool is_struct(uint32_t c) { return c == '{' || c == '}' || c == '[' || c == ']' || c == ':' || c == ','; }
constexpr bool is_rare(uint32_t c) { return c == '\\' || (c < 0x20 && !is_ws(c)) || c == '<' || c == '>' || c == '&' || c >= 0x80; }

struct ClassLut { uint32_t v[4 * 256]; };
constexpr ClassLut make_class_lut() {
    ClassLut t{};
    for (uint32_t j = 0; j < 4; j++)
        for (uint32_t c = 0; c < 256; c++)
            t.v[j * 256 + c] = ((c == '"' ? 1u : 0u) << j) | ((is_ws(c) ? 1u : 0u) << (8 + j)) | ((is_struct(c) ? 1u : 0u) << (16 + j)) |
                               ((is_rare(c) ? 1u : 0u) << (24 + j));
    return t;
}
__device__ const ClassLut g_class_lut = make_class_lut();

// token kind of the byte a token starts at (a closing quote is T_STR_OPEN + 1)
struct KindLut { uint8_t v[256]; };
constexpr KindLut make_kind_lut() {
    KindLut t{};
    for (uint32_t c = 0; c < 256; c++)
        t.v[c] = c == '{' ? T_LBRACE : c == '}' ? T_RBRACE : c == '[' ? T_LBRACKET : c == ']' ? T_RBRACKET : c == ':' ? T_COLON
               : c == ',' ? T_COMMA : c == '"' ? T_STR_OPEN : is_ws(c) ? T_SCALAR_END : T_SCALAR;
    return t;
}
__device__ const KindLut g_kind_lut = make_kind_lut();
static_assert(T_STR_CLOSE == T_STR_OPEN + 1, "a closing quote is the opening kind plus the in-string bit");

// simdjson's branch-free escape finder: bits that are escaped by a preceding backslash run.
// carry: in = first byte of this word is escaped; out = first byte of the next word is escaped.
__device__ __forceinline__ u64 find_escaped(u64 backslash, u64 &carry) {
    backslash &= ~carry;
    const u64 follows_escape = (backslash << 1) | carry;
    const u64 even_bits = 0x5555555555555555ull;
    const u64 odd_sequence_starts = backslash & ~even_bits & ~follows_escape;
    const u64 sum = odd_sequence_starts + backslash;
    carry = sum < odd_sequence_starts ? 1ull : 0ull;
    const u64 invert_mask = sum << 1;
    return (even_bits ^ invert_mask) & follows_escape;
}

__device__ __forceinline__ u64 prefix_xor(u64 x) {
    x ^= x << 1; x ^= x << 2; x ^= x << 4; x ^= x << 8; x ^= x << 16; x ^= x << 32;
    return x;
}

// token mask of one 64-byte chunk, given the state before its first byte
__device__ __forceinline__ u64 token_mask(u64 RQ, u64 SBx, u64 WS, u64 STm, bool s_in, bool prev_sc_if_out) {
    const u64 SB = s_in ? ~SBx : SBx;    // state before each byte: 1 = inside a string
    const u64 O = ~SB;
    const u64 SC = O & ~(WS | STm | RQ);  // scalar characters
    const u64 prev = (SC << 1) | ((!s_in && prev_sc_if_out) ? 1ull : 0ull);
    return (STm & O) | RQ | (SC & ~prev) | (WS & O & prev);
}

__global__ void __launch_bounds__(P1_THREADS) scan_tokens_kernel(const uint8_t *__restrict__ raw, u64 n, TokenBuffers tb,
                                                                u64 *scratch, ParseResult *res) {
    __shared__ uint8_t s_lut[256];
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp_esc[P1_THREADS / 32];   // bit0 all_bs, bit1 trail
    __shared__ uint32_t s_cin;                          // escape carry into the tile
    __shared__ uint32_t s_warp_par[P1_THREADS / 32];
    __shared__ u64 s_warp_cnt[P1_THREADS / 32];
    __shared__ uint32_t s_v;                            // in-string state entering the tile
    __shared__ u64 s_base;                              // tokens before the tile

    for (int c = threadIdx.x; c < 256; c += P1_THREADS) s_lut[c] = (uint8_t)byte_class(c);
    if (threadIdx.x == 0) s_tile = (uint32_t)atomicAdd(&scratch[0], 1ull);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t n_tiles = (uint32_t)((n + P1_TILE - 1) / P1_TILE);
    u64 *statusA = scratch + 4;             // escape carry
    u64 *statusB = scratch + 4 + n_tiles;   // quote parity + token counts
    const u64 chunk = (u64)tile * P1_TILE + (u64)threadIdx.x * P1_BYTES_PER_THREAD;

    // ---- load 64 bytes, build the class masks
    u64 Q = 0, B = 0, WS = 0, STm = 0, CT = 0, NA = 0, HT = 0;
    {
        const uint4 *p = reinterpret_cast<const uint4 *>(raw + chunk);
#pragma unroll
        for (int v = 0; v < 4; v++) {
            const uint4 q = __ldg(p + v);
            const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t x = w[k];
                const uint32_t c4 = (uint32_t)s_lut[x & 0xff] | ((uint32_t)s_lut[(x >> 8) & 0xff] << 8) |
                                    ((uint32_t)s_lut[(x >> 16) & 0xff] << 16) | ((uint32_t)s_lut[x >> 24] << 24);
                const int sh = (v * 4 + k) * 4;
                Q |= (u64)gather4(c4, 0) << sh;
                WS |= (u64)gather4(c4, 2) << sh;
                STm |= (u64)gather4(c4, 3) << sh;
                // backslashes, control bytes, < > & and non-ASCII bytes are rare in these documents:
                // their masks are only gathered for a word that holds one
                if ((c4 & 0x32323232u) | (x & 0x80808080u)) {
                    B |= (u64)gather4(c4, 1) << sh;
                    CT |= (u64)gather4(c4, 4) << sh;
                    HT |= (u64)gather4(c4, 5) << sh;
                    NA |= (u64)gather4(x, 7) << sh;
                }
            }
        }
    }

    // ---- stage 1: escape carry.  A chunk's carry-out depends on its carry-in only if the chunk
    //      is nothing but backslashes.
    const bool all_bs = (B == ~0ull);
    const uint32_t trail = all_bs ? 0u : ((uint32_t)__clzll((long long)~B) & 1u);  // odd trailing backslash run
    const unsigned m_nb = __ballot_sync(FULL, !all_bs), m_tr = __ballot_sync(FULL, trail);
    if (lane == 0) {
        uint32_t e = 0;
        if (m_nb == 0) e = 1;
        else e = ((m_tr >> (31 - __clz(m_nb))) & 1u) << 1;
        s_warp_esc[warp] = e;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // tile aggregate: published at once, so no tile ever waits on a chain
        uint32_t agg = 1;  // all_bs so far
        for (int i = 0; i < P1_THREADS / 32; i++)
            if (!(s_warp_esc[i] & 1)) agg = s_warp_esc[i];
        st_relaxed(statusA + tile, ST_AGG | agg);
        uint32_t cin = 0;
        for (int t = (int)tile - 1; t >= 0; t--) {
            u64 s;
            do { s = ld_relaxed(statusA + t); } while ((s >> 62) == 0);
            if (!(s & 1)) { cin = (uint32_t)(s >> 1) & 1u; break; }
        }
        s_cin = cin;
    }
    __syncthreads();
    u64 carry;
    {
        // carry into this thread's chunk: trail of the nearest earlier chunk that is not all backslashes
        uint32_t c = s_cin;
        for (uint32_t i = 0; i < warp; i++)
            if (!(s_warp_esc[i] & 1)) c = (s_warp_esc[i] >> 1) & 1u;
        const unsigned lower = m_nb & ((1u << lane) - 1);
        if (lower) c = (m_tr >> (31 - __clz(lower))) & 1u;
        carry = c;
    }
    const u64 E = find_escaped(B, carry);
    const u64 RQ = Q & ~E;                       // real (unescaped) quotes
    const u64 SBx = prefix_xor(RQ) ^ RQ;         // state before each byte, if the chunk started outside

    // ---- stage 2: quote parity + token counts for both possible entry states
    uint32_t pb = ' ';
    if (chunk > 0) pb = __ldg(raw + chunk - 1);
    const bool prev_sc_if_out = !(s_lut[pb] & (C_WS | C_ST | C_QUOTE));
    const u64 T0 = token_mask(RQ, SBx, WS, STm, false, prev_sc_if_out);
    const u64 T1 = token_mask(RQ, SBx, WS, STm, true, prev_sc_if_out);
    const uint32_t par = (uint32_t)__popcll(RQ) & 1u;

    // exclusive xor-scan of parity over the tile
    uint32_t lp = par;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(FULL, lp, d);
        if (lane >= (uint32_t)d) lp ^= v;
    }
    if (lane == 31) s_warp_par[warp] = lp;
    __syncthreads();
    uint32_t before = 0, tile_par = 0;
    for (uint32_t i = 0; i < P1_THREADS / 32; i++) {
        if (i < warp) before ^= s_warp_par[i];
        tile_par ^= s_warp_par[i];
    }
    lp = lp ^ par ^ before;                      // parity of the quotes in earlier threads of the tile
    // a: tokens if the tile is entered outside a string, b: if entered inside one
    const uint32_t a = (uint32_t)__popcll(lp ? T1 : T0), b = (uint32_t)__popcll(lp ? T0 : T1);
    u64 ab = (u64)a | ((u64)b << 32);
    u64 incl = ab;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u64 v = __shfl_up_sync(FULL, incl, d);
        if (lane >= (uint32_t)d) incl += v;
    }
    if (lane == 31) s_warp_cnt[warp] = incl;
    __syncthreads();
    u64 wbefore = 0, tile_cnt = 0;
    for (uint32_t i = 0; i < P1_THREADS / 32; i++) {
        if (i < warp) wbefore += s_warp_cnt[i];
        tile_cnt += s_warp_cnt[i];
    }
    const u64 excl = wbefore + incl - ab;
    const uint32_t cnt0 = (uint32_t)tile_cnt, cnt1 = (uint32_t)(tile_cnt >> 32);  // <= 16384 each

    if (warp == 0) {
        // look-back over (parity, cnt0, cnt1) aggregates; composition is ordered
        uint32_t V = 0;
        u64 base = 0;
        if (tile > 0) {
            if (lane == 0) st_relaxed(statusB + tile, ST_AGG | tile_par | ((u64)cnt0 << 1) | ((u64)cnt1 << 17));
            uint32_t rp = 0, r0 = 0, r1 = 0;  // running aggregate of the tiles nearer than the window
            int look = (int)tile - 1;
            for (;;) {
                const int idx = look - (int)lane;
                u64 s = idx >= 0 ? ld_relaxed(statusB + idx) : ST_PREFIX;  // before tile 0: outside, 0 tokens
                while (__any_sync(FULL, (s >> 62) == 0)) {
                    if ((s >> 62) == 0) s = ld_relaxed(statusB + idx);
                }
                const unsigned pm = __ballot_sync(FULL, (s >> 62) == 2);
                const int first = pm ? __ffs(pm) - 1 : 32;
                // window aggregate of lanes [0, first): lane 0 is nearest.  Ordered reduction:
                // (far then near): parity = pf ^ pn; cnt_s = cf_s + cn_(s ^ pf)
                uint32_t p = 0, c0 = 0, c1 = 0;
                if ((int)lane < first) { p = (uint32_t)s & 1u; c0 = (uint32_t)(s >> 1) & 0xffffu; c1 = (uint32_t)(s >> 17) & 0xffffu; }
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t fp = __shfl_down_sync(FULL, p, d), f0 = __shfl_down_sync(FULL, c0, d), f1 = __shfl_down_sync(FULL, c1, d);
                    if (lane + d < 32) {
                        const uint32_t n0 = f0 + (fp ? c1 : c0), n1 = f1 + (fp ? c0 : c1);
                        p ^= fp; c0 = n0; c1 = n1;
                    }
                }
                p = __shfl_sync(FULL, p, 0); c0 = __shfl_sync(FULL, c0, 0); c1 = __shfl_sync(FULL, c1, 0);
                // fold: (window [farther]) then (running [nearer])
                {
                    const uint32_t n0 = c0 + (p ? r1 : r0), n1 = c1 + (p ? r0 : r1);
                    rp ^= p; r0 = n0; r1 = n1;
                }
                if (pm) {
                    const u64 pre = __shfl_sync(FULL, s, first);
                    const uint32_t ps = (uint32_t)pre & 1u;
                    const u64 pc = (pre & ST_MASK) >> 1;
                    V = ps ^ rp;
                    base = pc + (ps ? r1 : r0);
                    break;
                }
                look -= 32;
            }
        }
        const uint32_t s_out = V ^ tile_par;
        const u64 total = base + (V ? cnt1 : cnt0);
        if (lane == 0) {
            st_relaxed(statusB + tile, ST_PREFIX | s_out | (total << 1));
            s_v = V;
            s_base = base;
            if (tile + 1 == n_tiles) { res->n_tokens = total; res->end_in_string = s_out; }
        }
    }
    __syncthreads();
    const bool s_in = (s_v ^ lp) != 0;
    const u64 T = s_in ? T1 : T0;
    const u64 SB = s_in ? ~SBx : SBx;
    u64 idx = s_base + (s_v ? (excl >> 32) : (excl & 0xffffffffull));

    // ---- string body checks (scanner.go stateInString / stateInStringEsc*)
    const u64 content = SB & ~RQ;
    bool bad = (CT & content) != 0;
    for (u64 m = WS & content; m; m &= m - 1)  // white space inside a string: only the space itself is allowed raw
        if (__ldg(raw + chunk + (__ffsll((long long)m) - 1)) != ' ') bad = true;
    for (u64 e = E & content; e; e &= e - 1) {
        const int i = __ffsll((long long)e) - 1;
        const uint32_t c = __ldg(raw + chunk + i);
        if (c == 'u') {
            for (int k = 1; k <= 4; k++) {
                const u64 at = chunk + i + k;
                const uint32_t h = at < n ? __ldg(raw + at) : 0;
                if (!((h >= '0' && h <= '9') || ((h | 0x20) >= 'a' && (h | 0x20) <= 'f'))) bad = true;
            }
        } else if (!(c == '"' || c == '\\' || c == '/' || c == 'b' || c == 'f' || c == 'n' || c == 'r' || c == 't')) {
            bad = true;
        }
    }
    if (bad) res->invalid = 1;

    // ---- strings that are not plain: a backslash, a non-ASCII byte or one of < > & inside.  They
    //      are decoded into the side arena and take the escaping path when serialised.
    for (u64 dmask = (B | NA | HT) & content; dmask;) {
        const int i = __ffsll((long long)dmask) - 1;
        const u64 t = idx + (u64)__popcll(T & low_mask(i)) - 1;  // the string's opening-quote token
        if (t < tb.cap) atomicOr(tb.decode_bits + (t >> 5), 1u << (t & 31));
        res->n_decode = 1;
        const u64 nq = RQ & ~low_mask(i + 1);
        dmask = nq ? (dmask & ~low_mask(__ffsll((long long)nq) - 1)) : 0;
    }

    // ---- emit tokens in order
    for (u64 t = T; t; t &= t - 1, idx++) {
        const int i = __ffsll((long long)t) - 1;
        if (idx >= tb.cap) continue;
        const uint32_t c = __ldg(raw + chunk + i);
        uint32_t k;
        if ((RQ >> i) & 1) k = ((SB >> i) & 1) ? T_STR_CLOSE : T_STR_OPEN;
        else if (c == '{') k = T_LBRACE;
        else if (c == '}') k = T_RBRACE;
        else if (c == '[') k = T_LBRACKET;
        else if (c == ']') k = T_RBRACKET;
        else if (c == ':') k = T_COLON;
        else if (c == ',') k = T_COMMA;
        else if ((WS >> i) & 1) k = T_SCALAR_END;
        else k = T_SCALAR;
        tb.pos[idx] = (uint32_t)(chunk + i);
        tb.kind[idx] = (uint8_t)k;
    }
}

// =================================================================================================
// P1, fast path (count / prefix / emit): see scan_fast.cuh
// =================================================================================================
#include "scan_fast.cuh"

// =================================================================================================
// generic exclusive scan of u32 (single pass, decoupled look-back); total -> scratch[1]
// =================================================================================================

constexpr int SCAN_THREADS = 256, SCAN_ITEMS = 4, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// T = uint32_t or u64 (two 32-bit running sums packed into one word); sums are carried as u64
// In: anything with `T operator()(uint32_t i) const` (a plain array, or values computed on the fly)
template <class T> struct ArrayIn {
    const T *p;
    __device__ __forceinline__ T operator()(uint32_t i) const { return p[i]; }
};
template <class T, class In>
__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(const In in, T *__restrict__ out, uint32_t n, u64 *scratch) {
    __shared__ uint32_t s_tile;
    __shared__ u64 s_warp[SCAN_THREADS / 32];
    __shared__ u64 s_prefix;
    if (threadIdx.x == 0) s_tile = (uint32_t)atomicAdd(&scratch[0], 1ull);
    __syncthreads();
    const uint32_t tile = s_tile, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    T v[SCAN_ITEMS];
    u64 sum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) { v[i] = base + i < n ? in(base + i) : T(0); sum += v[i]; }
    u64 incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u64 x = __shfl_up_sync(FULL, incl, d);
        if (lane >= (uint32_t)d) incl += x;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    u64 wbefore = 0, total = 0;
    for (uint32_t i = 0; i < SCAN_THREADS / 32; i++) { if (i < warp) wbefore += s_warp[i]; total += s_warp[i]; }
    // status words carry 62 payload bits: a packed pair must keep its high half below 2^30
    u64 *status = scratch + 2;
    if (warp == 0) {
        u64 prefix = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed(status, ST_PREFIX | total);
        } else {
            if (lane == 0) st_relaxed(status + tile, ST_AGG | total);
            int look = (int)tile - 1;
            for (;;) {
                const int idx = look - (int)lane;
                u64 s = idx >= 0 ? ld_relaxed(status + idx) : ST_PREFIX;
                while (__any_sync(FULL, (s >> 62) == 0)) { if ((s >> 62) == 0) s = ld_relaxed(status + idx); }
                const unsigned pm = __ballot_sync(FULL, (s >> 62) == 2);
                const int first = pm ? __ffs(pm) - 1 : 32;
                u64 x = ((int)lane <= first) ? (s & ST_MASK) : 0ull;
#pragma unroll
                for (int d = 16; d; d >>= 1) x += __shfl_xor_sync(FULL, x, d);
                prefix += x;
                if (pm) break;
                look -= 32;
            }
            if (lane == 0) st_relaxed(status + tile, ST_PREFIX | (prefix + total));
        }
        if (lane == 0) {
            s_prefix = prefix;
            if ((u64)(tile + 1) * SCAN_TILE >= n) scratch[1] = prefix + total;
        }
    }
    __syncthreads();
    u64 run = s_prefix + wbefore + incl - sum;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) { if (base + i < n) out[base + i] = (T)run; run += v[i]; }
}

inline size_t scan_scratch_words_u32(uint32_t n) { return 2 + (size_t)((n + SCAN_TILE - 1) / SCAN_TILE); }

}  // namespace

cudaError_t scan_u32(const uint32_t *in, uint32_t *out, uint32_t n, unsigned long long *scratch, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(scratch, 0, scan_scratch_words_u32(n) * 8, s);
    if (e != cudaSuccess) return e;
    if (n == 0) return cudaSuccess;
    scan_kernel<uint32_t><<<(n + SCAN_TILE - 1) / SCAN_TILE, SCAN_THREADS, 0, s>>>(ArrayIn<uint32_t>{in}, out, n, scratch);
    g_parse_launches++;
    return cudaGetLastError();
}
cudaError_t scan_u64(const unsigned long long *in, unsigned long long *out, uint32_t n, unsigned long long *scratch, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(scratch, 0, scan_scratch_words_u32(n) * 8, s);
    if (e != cudaSuccess) return e;
    if (n == 0) return cudaSuccess;
    scan_kernel<u64><<<(n + SCAN_TILE - 1) / SCAN_TILE, SCAN_THREADS, 0, s>>>(ArrayIn<u64>{in}, out, n, scratch);
    g_parse_launches++;
    return cudaGetLastError();
}
void count_parse_launch() { g_parse_launches++; }

namespace {

// =================================================================================================
// P2a: per-tile prefixes over the token stream
// =================================================================================================

// T_LBRACE 0, T_RBRACE 1, T_LBRACKET 2, T_RBRACKET 3: bit 1 tells braces from brackets, bit 0 openers from closers
__device__ __forceinline__ bool is_opener(uint32_t k) { return (k & ~2u) == 0u; }
__device__ __forceinline__ bool is_closer(uint32_t k) { return (k & ~2u) == 1u; }
static_assert(T_LBRACE == 0 && T_RBRACE == 1 && T_LBRACKET == 2 && T_RBRACKET == 3, "is_opener / is_closer");

// Which tokens start a tape row: scalars, openers, the opening quote of a string VALUE (no colon
// two tokens on), and closers other than the one of an empty container (which rides on the
// opener's row); keys ride on their value's row.  See emits_row_k and pass 1 of rows_kernel.

// ---- token passes with TP_ITEMS consecutive tokens per thread: one block-wide scan per 4,096
//      tokens; the per-256-token TilePrefix array is still what the later passes index.

constexpr int TP_ITEMS = 16;
constexpr int TP_THREADS = 256;
constexpr int TP_BLOCK = TP_THREADS * TP_ITEMS;        // 4,096 tokens = 2 blocks of TOK_TILE
constexpr int TP_TILES = TP_BLOCK / TOK_TILE;

// kinds of tokens [base - 4, base + TP_ITEMS + 4) as bytes k[0 .. 23]; 0xff outside the stream
struct KindWindow {
    uint32_t w[6];
    __device__ __forceinline__ uint32_t at(int j) const { return (w[(j + 4) >> 2] >> (((j + 4) & 3) * 8)) & 0xffu; }  // j in [-4, 20)
};
__device__ __forceinline__ KindWindow load_kinds(const uint8_t *kind, uint32_t base, uint32_t n_tok) {
    KindWindow kw;
    const uint32_t *p = reinterpret_cast<const uint32_t *>(kind);  // base is a multiple of 16
#pragma unroll
    for (int q = 0; q < 6; q++) {
        const long long first = (long long)base - 4 + q * 4;
        uint32_t v = 0xffffffffu;
        if (first >= 0 && first < (long long)n_tok) {
            v = __ldg(p + (first >> 2));
            if (first + 4 > (long long)n_tok) v |= 0xffffffffu << ((n_tok - (uint32_t)first) * 8);  // tokens past the end read as 0xff
        }
        kw.w[q] = v;
    }
    return kw;
}
__device__ __forceinline__ bool emits_row_k(uint32_t k, uint32_t prev, uint32_t next2) {
    // (bitwise on purpose: no branches)
    return ((uint32_t)(k == T_SCALAR) | (uint32_t)is_opener(k) | ((uint32_t)(k == T_STR_OPEN) & (uint32_t)(next2 != T_COLON)) |
            ((uint32_t)is_closer(k) & (uint32_t)!is_opener(prev))) != 0u;
}

// exclusive block scan of (depth delta, rows, brackets) over TP_THREADS threads, plus block totals
__device__ __forceinline__ void block_scan3(int d, uint32_t r, uint32_t b, int *ed, uint32_t *er, uint32_t *eb, int *td, uint32_t *tr,
                                            uint32_t *tbk) {
    __shared__ int s_d[TP_THREADS / 32];
    __shared__ uint32_t s_r[TP_THREADS / 32], s_b[TP_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int xd = d;
    uint32_t xr = r, xb = b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int yd = __shfl_up_sync(FULL, xd, o);
        const uint32_t yr = __shfl_up_sync(FULL, xr, o), yb = __shfl_up_sync(FULL, xb, o);
        if (lane >= (uint32_t)o) { xd += yd; xr += yr; xb += yb; }
    }
    if (lane == 31) { s_d[warp] = xd; s_r[warp] = xr; s_b[warp] = xb; }
    __syncthreads();
    int wd = 0, sd = 0;
    uint32_t wr = 0, wb = 0, sr = 0, sb = 0;
#pragma unroll
    for (uint32_t i = 0; i < TP_THREADS / 32; i++) {
        if (i < warp) { wd += s_d[i]; wr += s_r[i]; wb += s_b[i]; }
        sd += s_d[i]; sr += s_r[i]; sb += s_b[i];
    }
    *ed = wd + xd - d; *er = wr + xr - r; *eb = wb + xb - b;
    *td = sd; *tr = sr; *tbk = sb;
    __syncthreads();
}

// ---- P2a in two launches, neither waiting on another block:
//   token_agg     per block of TOK_TILE (2,048) tokens: depth change, rows, brackets, and the lowest /
//                 highest depth reached inside it relative to its start;
//   token_blocks  one block folds those in order: the TilePrefix of every token block, the totals
//                 and the depth checks (a closer with nothing open, the 10,000 limit, the end at 0).
struct TokAgg { int d, mn, mx; uint32_t rows, brackets; };

// kinds of tokens [base - 4, base + 12) as bytes; 0xff outside the stream (base is a multiple of 8)
struct KindWindow8 {
    uint32_t w[4];
    __device__ __forceinline__ uint32_t at(int j) const { return (w[(j + 4) >> 2] >> (((j + 4) & 3) * 8)) & 0xffu; }  // j in [-4, 12)
};
__device__ __forceinline__ KindWindow8 load_kinds8(const uint8_t *kind, uint32_t base, uint32_t n_tok) {
    KindWindow8 kw;
    const uint32_t *p = reinterpret_cast<const uint32_t *>(kind);
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const long long first = (long long)base - 4 + q * 4;
        uint32_t v = 0xffffffffu;
        if (first >= 0 && first < (long long)n_tok) {
            v = __ldg(p + (first >> 2));
            if (first + 4 > (long long)n_tok) v |= 0xffffffffu << ((n_tok - (uint32_t)first) * 8);  // tokens past the end read as 0xff
        }
        kw.w[q] = v;
    }
    return kw;
}

constexpr int TA_ITEMS = TOK_TILE / TP_THREADS;   // 8 tokens per thread
static_assert(TA_ITEMS == 8, "token_agg_kernel reads its tokens through KindWindow8");

__global__ void __launch_bounds__(TP_THREADS) token_agg_kernel(TokenBuffers tb, uint32_t n_tok, TokAgg *__restrict__ agg) {
    __shared__ int s_d[TP_THREADS / 32], s_mn[TP_THREADS / 32], s_mx[TP_THREADS / 32];
    __shared__ uint32_t s_r[TP_THREADS / 32], s_b[TP_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t base = blockIdx.x * TOK_TILE + threadIdx.x * TA_ITEMS;
    const KindWindow8 kw = load_kinds8(tb.kind, base, n_tok);
    int d = 0, mn = 0x7fffffff, mx = -0x7fffffff;
    uint32_t r = 0, b = 0;
#pragma unroll
    for (int j = 0; j < TA_ITEMS; j++) {
        const uint32_t k = kw.at(j);
        const int dd = is_opener(k) ? 1 : is_closer(k) ? -1 : 0;
        d += dd;
        b += dd != 0;
        r += emits_row_k(k, kw.at(j - 1), kw.at(j + 2)) ? 1u : 0u;
        if (k != 0xff) { mn = min(mn, d); mx = max(mx, d); }
    }
    // depth before this thread (inclusive warp scan, then the warps before), to place mn / mx
    int xd = d;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(FULL, xd, o);
        if (lane >= (uint32_t)o) xd += y;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) { r += __shfl_xor_sync(FULL, r, o); b += __shfl_xor_sync(FULL, b, o); }
    if (lane == 31) s_d[warp] = xd;
    if (lane == 0) { s_r[warp] = r; s_b[warp] = b; }
    __syncthreads();
    int before = xd - d;
    for (uint32_t i = 0; i < warp; i++) before += s_d[i];
    if (mn != 0x7fffffff) { mn += before; mx += before; }
#pragma unroll
    for (int o = 16; o; o >>= 1) { mn = min(mn, __shfl_xor_sync(FULL, mn, o)); mx = max(mx, __shfl_xor_sync(FULL, mx, o)); }
    if (lane == 0) { s_mn[warp] = mn; s_mx[warp] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        TokAgg a;
        a.d = 0; a.rows = 0; a.brackets = 0; a.mn = 0x7fffffff; a.mx = -0x7fffffff;
        for (uint32_t i = 0; i < TP_THREADS / 32; i++) {
            a.d += s_d[i]; a.rows += s_r[i]; a.brackets += s_b[i];
            a.mn = min(a.mn, s_mn[i]); a.mx = max(a.mx, s_mx[i]);
        }
        agg[blockIdx.x] = a;
    }
}

constexpr int TBP_THREADS = 1024, TBP_ITEMS = 4;   // one round covers 4,096 token blocks = 8.4 M tokens
__global__ void __launch_bounds__(TBP_THREADS) token_blocks_kernel(const TokAgg *__restrict__ agg, uint32_t n_blocks, TilePrefix *__restrict__ tiles,
                                                                  ParseResult *res) {
    __shared__ int s_d[TBP_THREADS / 32];
    __shared__ uint32_t s_r[TBP_THREADS / 32], s_b[TBP_THREADS / 32];
    __shared__ int s_lo, s_hi;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { s_lo = 0; s_hi = 0; }
    int run_d = 0;               // totals of the rounds before (the same in every thread)
    uint32_t run_r = 0, run_b = 0;
    for (uint32_t base = 0; base < n_blocks; base += TBP_THREADS * TBP_ITEMS) {
        const uint32_t first = base + threadIdx.x * TBP_ITEMS;
        TokAgg a[TBP_ITEMS];
        int d = 0;
        uint32_t r = 0, b = 0;
#pragma unroll
        for (int j = 0; j < TBP_ITEMS; j++) {
            a[j].d = 0; a[j].rows = 0; a[j].brackets = 0; a[j].mn = 0x7fffffff; a[j].mx = -0x7fffffff;
            if (first + j < n_blocks) a[j] = agg[first + j];
            d += a[j].d; r += a[j].rows; b += a[j].brackets;
        }
        int xd = d;
        uint32_t xr = r, xb = b;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int yd = __shfl_up_sync(FULL, xd, o);
            const uint32_t yr = __shfl_up_sync(FULL, xr, o), yb = __shfl_up_sync(FULL, xb, o);
            if (lane >= (uint32_t)o) { xd += yd; xr += yr; xb += yb; }
        }
        __syncthreads();   // the previous round has finished with the shared totals (and s_lo / s_hi are set)
        if (lane == 31) { s_d[warp] = xd; s_r[warp] = xr; s_b[warp] = xb; }
        __syncthreads();
        // the warps before this one: every warp folds the 32 warp totals with one shuffle scan
        int wd = s_d[lane];
        uint32_t wr = s_r[lane], wb = s_b[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int yd = __shfl_up_sync(FULL, wd, o);
            const uint32_t yr = __shfl_up_sync(FULL, wr, o), yb = __shfl_up_sync(FULL, wb, o);
            if (lane >= (uint32_t)o) { wd += yd; wr += yr; wb += yb; }
        }
        const int td = __shfl_sync(FULL, wd, 31);
        const uint32_t tr = __shfl_sync(FULL, wr, 31), tbk = __shfl_sync(FULL, wb, 31);
        const uint32_t src = warp ? warp - 1 : 0;
        int bd = __shfl_sync(FULL, wd, src);
        uint32_t br = __shfl_sync(FULL, wr, src), bb2 = __shfl_sync(FULL, wb, src);
        if (warp == 0) { bd = 0; br = 0; bb2 = 0; }
        TilePrefix p;
        p.depth = run_d + bd + xd - d; p.rows = run_r + br + xr - r; p.brackets = run_b + bb2 + xb - b; p.pad = 0;
        int lo = 0, hi = 0;   // lowest / highest depth reached inside this thread's blocks (0: nothing to report)
#pragma unroll
        for (int j = 0; j < TBP_ITEMS; j++) {
            if (first + j < n_blocks) {
                tiles[first + j] = p;
                if (a[j].mn != 0x7fffffff) { lo = min(lo, p.depth + a[j].mn); hi = max(hi, p.depth + a[j].mx); }
                p.depth += a[j].d; p.rows += a[j].rows; p.brackets += a[j].brackets;
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) { lo = min(lo, __shfl_xor_sync(FULL, lo, o)); hi = max(hi, __shfl_xor_sync(FULL, hi, o)); }
        if (lane == 0) {   // one shared-memory atomic per warp, not per block of tokens
            if (lo < 0) atomicMin(&s_lo, lo);
            if (hi > 0) atomicMax(&s_hi, hi);
        }
        run_d += td; run_r += tr; run_b += tbk;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (s_lo < 0) res->invalid = 1;                    // a closer with nothing open
        if (s_hi > 10000) res->invalid = 1;                // exceeded max depth
        res->max_level = (unsigned)min(s_hi, 16383);
        res->n_rows = run_r;
        res->n_brackets = run_b;
        res->final_depth = run_d;
        if (run_d != 0) res->invalid = 1;
    }
}

// =================================================================================================
// P2b: bracket tokens with their level; identity permutation for the sort
// =================================================================================================

__global__ void __launch_bounds__(TP_THREADS) brackets_kernel(TokenBuffers tb, uint32_t n_tok, const TilePrefix *__restrict__ tiles,
                                                             BracketBuffers bb) {
    const uint32_t base = blockIdx.x * TP_BLOCK + threadIdx.x * TP_ITEMS;
    const KindWindow kw = load_kinds(tb.kind, base, n_tok);
    int d = 0;
    uint32_t b = 0;
#pragma unroll
    for (int j = 0; j < TP_ITEMS; j++) {
        const uint32_t k = kw.at(j);
        const int dd = is_opener(k) ? 1 : is_closer(k) ? -1 : 0;
        d += dd;
        b += dd != 0;
    }
    int ed, td;
    uint32_t er, eb, tr, tbk;
    block_scan3(d, 0, b, &ed, &er, &eb, &td, &tr, &tbk);
    if (b != 0) {
        const TilePrefix p = tiles[blockIdx.x * TP_TILES];
        int depth = p.depth + ed;
        uint32_t bi = p.brackets + eb;
#pragma unroll
        for (int j = 0; j < TP_ITEMS; j++) {
            const uint32_t k = kw.at(j);
            const int dd = is_opener(k) ? 1 : is_closer(k) ? -1 : 0;
            if (dd != 0) {
                int level = depth + (dd > 0 ? 1 : 0);
                level = max(0, min(level, 16383));
                bb.tok[bi] = base + j;
                bb.lvkind[bi] = ((uint32_t)level << 8) | k;
                bb.sort_a[bi] = bi;
                bb.key_a[bi] = (uint32_t)level;   // (parent / match / row / ct are cleared in bulk by the launch wrapper)
                bi++;
            }
            depth += dd;
        }
    }
}

// Stable sort of (key, value) pairs by 3 key bits at a time, in three launches per digit and with
// no block waiting on another: per-tile histograms, one block that turns them into the start of
// every (bucket, tile) run, and a scatter that ranks each pair inside its tile.  Nesting levels up
// to 7 -- every document the reference writes -- take a single digit.
constexpr int RS_BITS = 3, RS_BUCKETS = 1 << RS_BITS;

// eight 16-bit counters (a tile has 1,024 pairs) in two words, so that a scan adds them all at once
struct Count8 {
    u64 a, b;
    __device__ __forceinline__ void inc(uint32_t bucket) { if (bucket < 4) a += 1ull << (16 * bucket); else b += 1ull << (16 * (bucket - 4)); }
    __device__ __forceinline__ uint32_t get(uint32_t bucket) const { return (uint32_t)((bucket < 4 ? a >> (16 * bucket) : b >> (16 * (bucket - 4))) & 0xffffu); }
};

__global__ void __launch_bounds__(SCAN_THREADS) radix_hist_kernel(const uint32_t *__restrict__ key, uint32_t n, int shift, uint32_t *__restrict__ hist,
                                                                 uint32_t tiles) {
    __shared__ u64 s_a[SCAN_THREADS / 32], s_b[SCAN_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    Count8 c{0, 0};
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++)
        if (base + i < n) c.inc((key[base + i] >> shift) & (RS_BUCKETS - 1));
#pragma unroll
    for (int o = 16; o; o >>= 1) { c.a += __shfl_xor_sync(FULL, c.a, o); c.b += __shfl_xor_sync(FULL, c.b, o); }
    if (lane == 0) { s_a[warp] = c.a; s_b[warp] = c.b; }
    __syncthreads();
    if (threadIdx.x < RS_BUCKETS) {
        Count8 t{0, 0};
        for (uint32_t w = 0; w < SCAN_THREADS / 32; w++) { t.a += s_a[w]; t.b += s_b[w]; }
        hist[threadIdx.x * tiles + blockIdx.x] = t.get(threadIdx.x);   // bucket-major
    }
}

// in place: hist[bucket][tile] -> number of pairs that sort before the first pair of that run
constexpr int RB_THREADS = 1024, RB_ITEMS = 4;
__global__ void __launch_bounds__(RB_THREADS) radix_base_kernel(uint32_t *hist, uint32_t entries) {
    __shared__ uint32_t s_w[RB_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t run = 0;
    for (uint32_t base = 0; base < entries; base += RB_THREADS * RB_ITEMS) {
        const uint32_t first = base + threadIdx.x * RB_ITEMS;
        uint32_t v[RB_ITEMS], sum = 0;
#pragma unroll
        for (int j = 0; j < RB_ITEMS; j++) { v[j] = first + j < entries ? hist[first + j] : 0u; sum += v[j]; }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, incl, o);
            if (lane >= (uint32_t)o) incl += y;
        }
        __syncthreads();   // the previous round has finished with s_w
        if (lane == 31) s_w[warp] = incl;
        __syncthreads();
        uint32_t w = s_w[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL, w, o);
            if (lane >= (uint32_t)o) w += y;
        }
        const uint32_t total = __shfl_sync(FULL, w, 31);
        uint32_t before = __shfl_sync(FULL, w, warp ? warp - 1 : 0);
        if (warp == 0) before = 0;
        uint32_t x = run + before + incl - sum;
#pragma unroll
        for (int j = 0; j < RB_ITEMS; j++) {
            if (first + j < entries) hist[first + j] = x;
            x += v[j];
        }
        run += total;
    }
}

__global__ void __launch_bounds__(SCAN_THREADS) radix_scatter_kernel(const uint32_t *__restrict__ key, const uint32_t *__restrict__ val, uint32_t n,
                                                                    int shift, const uint32_t *__restrict__ run_base, uint32_t tiles,
                                                                    uint32_t *__restrict__ key_out, uint32_t *__restrict__ val_out) {
    __shared__ u64 s_a[SCAN_THREADS / 32], s_b[SCAN_THREADS / 32];
    __shared__ uint32_t s_base[RS_BUCKETS];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    if (threadIdx.x < RS_BUCKETS) s_base[threadIdx.x] = run_base[threadIdx.x * tiles + blockIdx.x];
    uint32_t k[SCAN_ITEMS], v[SCAN_ITEMS];
    Count8 mine{0, 0};
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        k[i] = base + i < n ? key[base + i] : 0u;
        v[i] = base + i < n ? val[base + i] : 0u;
        if (base + i < n) mine.inc((k[i] >> shift) & (RS_BUCKETS - 1));
    }
    Count8 incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u64 ya = __shfl_up_sync(FULL, incl.a, o), yb = __shfl_up_sync(FULL, incl.b, o);
        if (lane >= (uint32_t)o) { incl.a += ya; incl.b += yb; }
    }
    if (lane == 31) { s_a[warp] = incl.a; s_b[warp] = incl.b; }
    __syncthreads();
    Count8 before{incl.a - mine.a, incl.b - mine.b};   // pairs of each bucket earlier in the tile
    for (uint32_t w = 0; w < warp; w++) { before.a += s_a[w]; before.b += s_b[w]; }
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        if (base + i >= n) break;
        const uint32_t bucket = (k[i] >> shift) & (RS_BUCKETS - 1);
        const uint32_t dst = s_base[bucket] + before.get(bucket);
        before.inc(bucket);
        key_out[dst] = k[i];
        val_out[dst] = v[i];
    }
}

// =================================================================================================
// P2d: pairs and enclosing openers from the level-sorted order
// =================================================================================================

enum { CT_TOP = 0, CT_OBJ = 1, CT_ARR = 2 };

__global__ void match_kernel(BracketBuffers bb, const uint32_t *__restrict__ sorted, const uint32_t *__restrict__ skey, uint32_t n_br,
                             ParseResult *res) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_br) return;
    // With depth never negative and ending at 0, every level's brackets alternate opener, closer:
    // even positions of the sorted order are openers and j+1 is the partner.
    if (j & 1) return;
    const uint32_t o = sorted[j];
    if (j + 1 >= n_br) { res->invalid = 1; return; }
    const uint32_t c = sorted[j + 1];
    const uint32_t ko = bb.lvkind[o] & 0xff, kc = bb.lvkind[c] & 0xff;
    if (!((ko == T_LBRACE && kc == T_RBRACE) || (ko == T_LBRACKET && kc == T_RBRACKET)) || skey[j] != skey[j + 1] || c < o) {
        res->invalid = 1;
        return;
    }
    bb.match[o] = c;
    bb.match[c] = o;
    const uint32_t level = skey[j];
    const uint32_t inside = ko == T_LBRACE ? CT_OBJ : CT_ARR;   // the container this pair delimits
    uint32_t outside = CT_TOP;                                    // the container the pair sits in
    if (level > 1) {
        // enclosing opener: the last bracket of level-1 that lies before o
        uint32_t lo = 0, hi = n_br;  // first position whose (key, bracket) >= (level-1, o)
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            const uint32_t mk = skey[mid];
            if (mk < level - 1 || (mk == level - 1 && sorted[mid] < o)) lo = mid + 1; else hi = mid;
        }
        if (lo == 0 || skey[lo - 1] != level - 1) { res->invalid = 1; return; }
        const uint32_t par = sorted[lo - 1];
        bb.parent[o] = par;
        outside = (bb.lvkind[par] & 0xff) == T_LBRACE ? CT_OBJ : CT_ARR;
    }
    // per bracket: bits 0-1 the container tokens AFTER it sit in, bits 2-3 the container the
    // bracket token itself is judged in (an opener: where it sits; a closer: what it closes)
    bb.ct[o] = inside | (outside << 2);
    bb.ct[c] = outside | (inside << 2);
}

// =================================================================================================
// P3: strings that need decoding
// =================================================================================================

__global__ void decode_count_kernel(const uint32_t *__restrict__ bits, uint32_t n_words, uint32_t *__restrict__ counts) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w < n_words) counts[w] = __popc(bits[w]);
}
__global__ void decode_list_kernel(const uint32_t *__restrict__ bits, const uint32_t *__restrict__ offs, uint32_t n_words,
                                   uint32_t *__restrict__ list) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    uint32_t o = offs[w];
    for (uint32_t b = bits[w]; b; b &= b - 1) list[o++] = w * 32 + (__ffs(b) - 1);
}

__global__ void decode_len_kernel(const uint8_t *__restrict__ raw, TokenBuffers tb, DecodeBuffers db, uint32_t n_dec) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_dec) return;
    const uint32_t t = db.tok[j];
    const uint32_t a = tb.pos[t] + 1, b = tb.pos[t + 1];  // the closing quote is the next token
    db.len[j] = unquote(raw + a, b > a ? b - a : 0, nullptr);
}
__global__ void decode_write_kernel(uint8_t *arena, uint32_t side_base, TokenBuffers tb, DecodeBuffers db, uint32_t n_dec) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_dec) return;
    const uint32_t t = db.tok[j];
    const uint32_t a = tb.pos[t] + 1, b = tb.pos[t + 1];
    const uint32_t o = db.off[j] + side_base;
    db.off[j] = o;
    unquote(arena + a, b > a ? b - a : 0, arena + o);
}

// =================================================================================================
// P2e: grammar + rows
// =================================================================================================

// where string token t (an opening quote) keeps its bytes: in place, or decoded in the side arena
__device__ __forceinline__ void string_ref(const TokenBuffers &tb, const DecodeBuffers &db, uint32_t n_dec, uint32_t t, uint32_t *off,
                                           uint32_t *len) {
    if ((tb.decode_bits[t >> 5] >> (t & 31)) & 1u) {
        uint32_t lo = 0, hi = n_dec;
        while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (db.tok[mid] < t) lo = mid + 1; else hi = mid; }
        *off = db.off[lo];
        *len = db.len[lo];
        return;
    }
    *off = tb.pos[t] + 1;
    *len = tb.pos[t + 1] - tb.pos[t] - 1;
}

// Go's number grammar (scanner.go state1/state0/stateDot/stateE...) and literals.  Returns 0 bad,
// 1 valid, 2 valid and already in the text form floatEncoder would print.
__device__ uint32_t check_scalar(const uint8_t *s, uint32_t n) {
    if (n == 0) return 0;
    if (s[0] == 't') return (n == 4 && s[1] == 'r' && s[2] == 'u' && s[3] == 'e') ? 2 : 0;
    if (s[0] == 'f') return (n == 5 && s[1] == 'a' && s[2] == 'l' && s[3] == 's' && s[4] == 'e') ? 2 : 0;
    if (s[0] == 'n') return (n == 4 && s[1] == 'u' && s[2] == 'l' && s[3] == 'l') ? 2 : 0;
    uint32_t i = 0;
    if (s[0] == '-') { i = 1; if (n == 1) return 0; }
    if (s[i] < '0' || s[i] > '9') return 0;
    const uint32_t int_start = i;
    if (s[i] == '0') i++;
    else while (i < n && s[i] >= '0' && s[i] <= '9') i++;
    const uint32_t int_digits = i - int_start;
    bool plain = true;
    if (i < n && s[i] == '.') {
        plain = false;
        i++;
        if (i >= n || s[i] < '0' || s[i] > '9') return 0;
        while (i < n && s[i] >= '0' && s[i] <= '9') i++;
    }
    if (i < n && (s[i] == 'e' || s[i] == 'E')) {
        plain = false;
        i++;
        if (i < n && (s[i] == '+' || s[i] == '-')) i++;
        if (i >= n || s[i] < '0' || s[i] > '9') return 0;
        while (i < n && s[i] >= '0' && s[i] <= '9') i++;
    }
    if (i != n) return 0;
    // integers of up to 15 digits are exact in float64 and print back unchanged ("-0" included)
    return (plain && int_digits <= 15) ? 2 : 1;
}



constexpr int RW_ITEMS = 8;                        // consecutive tokens per thread: about one object member
constexpr int RW_BLOCK = TP_THREADS * RW_ITEMS;   // 2,048 tokens per block = TOK_TILE

// Go's scanner grammar as a table: for the container a token sits in and the token's kind, the
// set of kinds the previous significant token may have.  A string is a KEY exactly when a colon
// follows its closing quote, so "key in value position", "value in key position", "key in an
// array" and the like all show up as a (previous, current) pair the table does not allow.
enum { G_START = 0, G_LBRACE, G_RBRACE, G_LBRACKET, G_RBRACKET, G_COLON, G_COMMA, G_KEY_OPEN, G_KEY_CLOSE, G_VAL_OPEN, G_VAL_CLOSE, G_SCALAR, G_N };
constexpr uint32_t gbit(int k) { return 1u << k; }
constexpr uint32_t G_VALUE_END = gbit(G_VAL_CLOSE) | gbit(G_SCALAR) | gbit(G_RBRACE) | gbit(G_RBRACKET);
// allowed[container][current kind] = mask of allowed previous kinds
__constant__ uint32_t g_allowed[3][G_N] = {
    // CT_TOP: one value, nothing after it
    {0, gbit(G_START), 0, gbit(G_START), 0, 0, 0, 0, 0, gbit(G_START), gbit(G_VAL_OPEN), gbit(G_START)},
    // CT_OBJ
    {0, gbit(G_COLON), gbit(G_LBRACE) | G_VALUE_END, gbit(G_COLON), 0, gbit(G_KEY_CLOSE), G_VALUE_END,
     gbit(G_LBRACE) | gbit(G_COMMA), gbit(G_KEY_OPEN), gbit(G_COLON), gbit(G_VAL_OPEN), gbit(G_COLON)},
    // CT_ARR
    {0, gbit(G_LBRACKET) | gbit(G_COMMA), 0, gbit(G_LBRACKET) | gbit(G_COMMA), gbit(G_LBRACKET) | G_VALUE_END, 0, G_VALUE_END,
     0, 0, gbit(G_LBRACKET) | gbit(G_COMMA), gbit(G_VAL_OPEN), gbit(G_LBRACKET) | gbit(G_COMMA)},
};

// kind of a token in grammar terms, without branches.  colon_after: for an opening quote, is token
// i + 2 a colon (open, close, colon); for a closing quote, is token i + 1 one.  k is 0 .. 8.
__device__ __forceinline__ uint32_t gkind(uint32_t k, bool colon_after) {
    // T:  {  }  [  ]  :  ,  "open  "close  scalar   ->   G:  1  2  3  4  5  6  7|9  8|10  11
    const uint32_t base = k >= 8 ? (uint32_t)G_SCALAR : ((0x87654321u >> (k * 4)) & 15u);
    return base + (((k - T_STR_OPEN) < 2u && !colon_after) ? 2u : 0u);
}
static_assert(G_KEY_OPEN == 7 && G_VAL_OPEN == 9 && G_KEY_CLOSE == 8 && G_VAL_CLOSE == 10 && G_SCALAR == 11 && G_COMMA == 6, "gkind's table");

// Three passes over a block of 2,048 tokens (256 threads x 8 consecutive tokens):
//   1  per token: depth change, bracket, "owns a row" -> block scan;
//   2  per token, walking the thread's 8 in order with the previous token's grammar kind carried
//      along: the grammar check, and every row-owning token is listed in shared memory at its
//      row's place (token, brackets before it, depth before it);
//   3  per ROW: thread t takes the block's t-th row, so that neighbouring threads do the same kind
//      of work and write neighbouring rows.
__global__ void __launch_bounds__(TP_THREADS) rows_kernel(const uint8_t *__restrict__ raw, u64 n, TokenBuffers tb, uint32_t n_tok,
                                                         const TilePrefix *__restrict__ tiles, BracketBuffers bb, DecodeBuffers db,
                                                         uint32_t n_dec, RowBuffers rb, uint32_t n_rows, ParseResult *res) {
    constexpr int HALO = 8;
    __shared__ uint8_t s_kind[RW_BLOCK + 2 * HALO];
    __shared__ uint32_t s_pos[RW_BLOCK + 2 * HALO];
    __shared__ uint8_t s_ct[RW_BLOCK + 1];
    __shared__ uint32_t s_dbits[RW_BLOCK / 32 + 2];
    __shared__ uint32_t s_allowed[3 * G_N];
    __shared__ uint32_t s_row[RW_BLOCK];    // token (11 bits) | brackets before it, from the block's first (12 bits)
    __shared__ uint16_t s_rowd[RW_BLOCK];   // depth before the token
    const long long blk0 = (long long)blockIdx.x * RW_BLOCK - HALO;
    for (int x = threadIdx.x; x < RW_BLOCK + 2 * HALO; x += TP_THREADS) {
        const long long i = blk0 + x;
        const bool in = i >= 0 && i < (long long)n_tok;
        s_kind[x] = in ? tb.kind[i] : (uint8_t)0xff;
        s_pos[x] = in ? tb.pos[i] : (uint32_t)n;   // positions past the last token read as the text length
    }
    if (threadIdx.x < 3 * G_N) s_allowed[threadIdx.x] = g_allowed[threadIdx.x / G_N][threadIdx.x % G_N];
    __syncthreads();
    const uint8_t *kind = s_kind - blk0;   // indexable by global token number inside the staged window
    const uint32_t *pos = s_pos - blk0;
    const uint32_t tok0 = blockIdx.x * RW_BLOCK;
    const uint32_t base = tok0 + threadIdx.x * RW_ITEMS;

    // ---- pass 1
    int d = 0;
    uint32_t r = 0, b = 0, rowbits = 0;
#pragma unroll
    for (int j = 0; j < RW_ITEMS; j++) {
        const uint32_t i = base + j;
        const uint32_t k = kind[i];   // 0xff past the end
        const bool op = is_opener(k), cl = is_closer(k);
        // (bitwise on purpose: no branches, both neighbours are read unconditionally)
        const uint32_t row = (uint32_t)(k == T_SCALAR) | (uint32_t)op | ((uint32_t)(k == T_STR_OPEN) & (uint32_t)(kind[i + 2] != T_COLON)) |
                             ((uint32_t)cl & (uint32_t)!is_opener(kind[i - 1]));
        d += (op ? 1 : 0) - (cl ? 1 : 0);
        b += (op || cl) ? 1u : 0u;
        r += row;
        rowbits |= row << j;
    }
    int ed, td;
    uint32_t er, eb, tr, tbk;
    block_scan3(d, r, b, &ed, &er, &eb, &td, &tr, &tbk);
    const TilePrefix p = tiles[blockIdx.x];
    // the block's slice of the per-bracket container table and of the decode bitmap
    const uint32_t b_first = p.brackets > 0 ? p.brackets - 1 : 0;           // first bracket any token here can ask about
    const uint32_t b_count = p.brackets + tbk - b_first;                     // <= RW_BLOCK + 1
    for (uint32_t x = threadIdx.x; x < b_count; x += TP_THREADS) s_ct[x] = (uint8_t)bb.ct[b_first + x];
    const long long w_first = (blk0 < 0 ? 0 : blk0) >> 5;                    // decode-bit words from the halo's first token on
    for (uint32_t x = threadIdx.x; x < RW_BLOCK / 32 + 2; x += TP_THREADS) s_dbits[x] = tb.decode_bits[w_first + x];
    __syncthreads();
    const uint8_t *ct_s = s_ct - b_first;
    const uint32_t *dbits = s_dbits - w_first;

    // ---- pass 2
    {
        uint32_t g_prev = G_START;   // grammar kind of the significant token before the thread's first one
        if (base > 0) {
            uint32_t pi = base - 1;
            if (kind[pi] == T_SCALAR_END) pi--;
            const uint32_t kp = kind[pi];
            if (kp != 0xffu) g_prev = gkind(kp, kind[pi + (kp == T_STR_OPEN ? 2 : 1)] == T_COLON);
        }
        int depth = p.depth + ed;
        uint32_t brb = p.brackets + eb, slot = er;
        bool bad = false;
#pragma unroll 2
        for (int j = 0; j < RW_ITEMS; j++) {
            const uint32_t i = base + j, k = kind[i];
            const bool op = is_opener(k), cl = is_closer(k), br = op || cl;
            if (k != 0xffu && k != T_SCALAR_END) {
                const uint32_t g = gkind(k, kind[i + (k == T_STR_OPEN ? 2 : 1)] == T_COLON);
                // container: the bracket's own entry for a bracket, else what follows the bracket before
                uint32_t ct = CT_TOP;
                if (br || brb > 0) {
                    const uint32_t e = ct_s[br ? brb : brb - 1];
                    ct = br ? (e >> 2) & 3u : e & 3u;
                    if (ct > CT_ARR) ct = CT_TOP;   // unmatched bracket: match_kernel has already said invalid
                }
                if (k != T_STR_CLOSE && !((s_allowed[ct * G_N + g] >> g_prev) & 1u)) bad = true;
                if (g == G_VAL_OPEN && kind[i + 1] != T_STR_CLOSE) bad = true;   // unterminated (cannot pass P1, kept as a guard)
                g_prev = g;
                if ((rowbits >> j) & 1u) {
                    s_row[slot] = (i - tok0) | ((brb - p.brackets) << 11);
                    s_rowd[slot] = (uint16_t)min(max(depth, 0), 65535);
                    slot++;
                }
            }
            depth += (op ? 1 : 0) - (cl ? 1 : 0);
            brb += br ? 1u : 0u;
        }
        if (bad) res->invalid = 1;
    }
    __syncthreads();

    // ---- pass 3
    for (uint32_t slot = threadIdx.x; slot < tr; slot += TP_THREADS) {
        const uint32_t e = s_row[slot];
        const uint32_t i = tok0 + (e & 2047u), my_b = p.brackets + (e >> 11);
        const uint32_t depth = s_rowd[slot];
        const uint32_t rr = p.rows + slot;
        if (rr >= n_rows) { res->invalid = 1; continue; }
        const uint32_t k = kind[i], k1 = kind[i + 1], k2 = kind[i + 2];
        const bool op = is_opener(k), cl = is_closer(k);
        uint32_t ct = CT_TOP;
        if (op || cl || my_b > 0) {
            const uint32_t c = ct_s[(op || cl) ? my_b : my_b - 1];
            ct = (op || cl) ? (c >> 2) & 3u : c & 3u;
            if (ct > CT_ARR) ct = CT_TOP;
        }
        uint32_t meta = 0, koff = 0, klen = 0, voff = 0, vlen = 0;
        uint32_t after = k1 == T_SCALAR_END ? k2 : k1;   // the significant token after the value, to look for the comma
        if (cl) {
            if (depth < 1) { res->invalid = 1; continue; }
            meta = (k == T_RBRACE ? ROW_CLOSE_OBJ : ROW_CLOSE_ARR) | ((depth - 1) << META_DEPTH_SHIFT);
            koff = bb.match[my_b];  // the opener's bracket: key_order and the module index find its row through bb.row
            vlen = pos[i];          // (container rows: where the bracket is in the text, see TapeView::text_len)
            bb.row[my_b] = rr;
        } else {
            if (op) {
                vlen = pos[i];
                if (is_closer(k1)) { meta = (k == T_LBRACE ? ROW_EMPTY_OBJ : ROW_EMPTY_ARR); after = k2 == T_SCALAR_END ? kind[i + 3] : k2; }
                else { meta = (k == T_LBRACE ? ROW_OPEN_OBJ : ROW_OPEN_ARR); bb.row[my_b] = rr; voff = my_b; }  // voff of an opening row: its bracket
            } else if (k == T_STR_OPEN) {
                meta = ROW_STR;
                if (!((dbits[i >> 5] >> (i & 31)) & 1u)) { meta |= META_VAL_PLAIN; voff = pos[i] + 1; vlen = pos[i + 1] - voff; }
                else { tb.pos = const_cast<uint32_t *>(pos); string_ref(tb, db, n_dec, i, &voff, &vlen); }
                after = k2 == T_SCALAR_END ? kind[i + 3] : k2;
            } else {  // scalar
                meta = ROW_RAW;
                voff = pos[i];
                vlen = pos[i + 1] - voff;
                const uint32_t c = check_scalar(raw + voff, vlen);
                if (c == 0) { res->invalid = 1; continue; }
                if (c == 1) { res->need_host_numbers = 1; res->not_canonical = 1; }
            }
            meta |= depth << META_DEPTH_SHIFT;
            if (ct == CT_OBJ && i >= 3) {   // (i < 3 only in text the grammar pass has already rejected)
                // "key" : value -> i-3 opening quote, i-2 closing quote, i-1 colon (the table has vouched for them)
                const uint32_t ko = i - 3;
                meta |= META_HAS_KEY;
                if (!((dbits[ko >> 5] >> (ko & 31)) & 1u)) { meta |= META_KEY_PLAIN; koff = pos[ko] + 1; klen = pos[ko + 1] - koff; }
                else { tb.pos = const_cast<uint32_t *>(pos); string_ref(tb, db, n_dec, ko, &koff, &klen); }
                if (ko > 0) { uint32_t q = ko - 1; if (kind[q] == T_SCALAR_END) q--; if (kind[q] == T_COMMA) meta |= (1u << 7); }  // has an earlier sibling
            }
        }
        // an opening row never carries the comma: it follows the closer and rides on the closing row
        const uint32_t rk = meta & 15u;
        if (rk != ROW_OPEN_OBJ && rk != ROW_OPEN_ARR && after == T_COMMA) meta |= META_COMMA;
        if (rr + 1 == n_rows) meta |= META_LAST;
        rb.meta[rr] = meta; rb.koff[rr] = koff; rb.klen[rr] = klen; rb.voff[rr] = voff; rb.vlen[rr] = vlen;
        // what the module index (index.cu) starts from: the root's kind and its "module" member
        if (rr == 0) res->root_kind = meta & 15u;
        if (depth == 1 && (meta & META_HAS_KEY) && klen == 6) {
            const uint8_t *kp = raw + koff;
            if (kp[0] == 'm' && kp[1] == 'o' && kp[2] == 'd' && kp[3] == 'u' && kp[4] == 'l' && kp[5] == 'e') {
                res->module_row = rr + 1;
                res->module_kind = meta & 15u;
            }
        }
    }
}


// =================================================================================================
// P4: are the keys of every object strictly ascending (bytewise, decoded)?
// =================================================================================================

__global__ void key_order_kernel(const uint8_t *__restrict__ arena, RowBuffers rb, uint32_t n_rows, BracketBuffers bb, ParseResult *res) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    // Text that rows_kernel has just found invalid leaves rows that refer to nothing -- with whatever an earlier, larger
    // document left in the recycled buffers as bracket numbers and offsets (tools/stress_lines.py, seed 1147): nothing
    // is followed from a row before the verdict so far has been looked at.  (Asked for here, looked at after the row's
    // own word has been asked for too: the two loads travel together.)
    const uint32_t invalid = __ldg(&res->invalid);
    uint32_t meta = rb.meta[r];
    if (invalid) return;
    if (r == 0 && res->module_row != 0 && res->module_kind == ROW_OPEN_OBJ) {
        // for the module index, which runs next: the row that closes "module" (voff of an opening row is its bracket)
        const uint32_t cb = bb.match[rb.voff[res->module_row - 1]];
        res->module_end = cb != 0xffffffffu ? bb.row[cb] : 0u;
    }
    const uint32_t kind = meta & 15u;
    if (kind == ROW_CLOSE_OBJ || kind == ROW_CLOSE_ARR) {
        // nothing to compare; leave the opener's row in koff for the sibling lookup below
        return;
    }
    // for the module index: a string member three levels down called "name" / "hostname" (the key's bytes are about
    // to be read for the comparison below anyway)
    uint32_t marks = 0;
    if ((meta >> META_DEPTH_SHIFT) == 3 && (meta & (15u | META_HAS_KEY)) == (ROW_STR | META_HAS_KEY)) {
        const uint32_t kl = rb.klen[r];
        if (kl == 4 || kl == 8) {
            const uint32_t ko = rb.koff[r], w0 = load4_be(arena, ko);
            if (kl == 4) { if (w0 == 0x6e616d65u) marks = META_KEY_IS_NAME; }
            else if (w0 == 0x686f7374u && load4_be(arena, ko + 4) == 0x6e616d65u) marks = META_KEY_IS_HOSTNAME;
        }
    }
    const bool has_sibling = (meta & (1u << 7)) != 0;
    if (has_sibling || marks) rb.meta[r] = (meta & ~(1u << 7)) | marks;
    if (!has_sibling || r == 0) return;
    uint32_t s = r - 1;  // previous sibling: the row before, or the opener of the container that just closed
    const uint32_t sk = rb.meta[s] & 15u;
    if (sk == ROW_CLOSE_OBJ || sk == ROW_CLOSE_ARR) {
        const uint32_t ob = rb.koff[s];
        if (ob == 0xffffffffu) { res->not_canonical = 1; return; }
        s = bb.row[ob];
        if (s == 0xffffffffu || s >= n_rows) { res->not_canonical = 1; return; }
    }
    const uint32_t ao = rb.koff[s], al = rb.klen[s], bo = rb.koff[r], bl = rb.klen[r];
    int cmp = compare_bytes(arena, ao, al, bo, bl);
    if (cmp >= 0) res->not_canonical = 1;  // out of order or duplicated: the host re-orders
}

}  // namespace

// =================================================================================================
// launch wrappers
// =================================================================================================

// resident blocks of the persistent fast kernel on the current device (asked once per device)
static cudaError_t fast_scan_grid(uint32_t *count_blocks, uint32_t *emit_blocks) {
    static std::atomic<uint32_t> cached[64][2];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const bool slot = dev >= 0 && dev < 64;
    uint32_t g0 = slot ? cached[dev][0].load() : 0, g1 = slot ? cached[dev][1].load() : 0;
    if (!g0 || !g1) {
        int a = 0, b = 0, sms = 0;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, scan_count_kernel, P1_THREADS, 0)) != cudaSuccess) return e;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, scan_emit_kernel, P1_THREADS, 0)) != cudaSuccess) return e;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        g0 = (uint32_t)(a > 0 ? a : 1) * (uint32_t)(sms > 0 ? sms : 1);
        g1 = (uint32_t)(b > 0 ? b : 1) * (uint32_t)(sms > 0 ? sms : 1);
        if (slot) { cached[dev][0].store(g0); cached[dev][1].store(g1); }
    }
    *count_blocks = g0;
    *emit_blocks = g1;
    return cudaSuccess;
}

cudaError_t launch_scan_tokens(const uint8_t *raw, size_t n, TokenBuffers tb, unsigned long long *scratch, ParseResult *res,
                               bool carry_protocol, int stage, cudaStream_t s) {
    cudaError_t e;
    const size_t decode_words = ((size_t)tb.cap + 31) / 32 + 1;
    if (stage == 0 && (e = cudaMemsetAsync(res, 0, sizeof(ParseResult), s)) != cudaSuccess) return e;
    if (n == 0) return stage == 0 ? cudaMemsetAsync(tb.decode_bits, 0, decode_words * 4, s) : cudaSuccess;
    const uint32_t n_tiles = p1_tiles(n);
    if (carry_protocol) {
        if (stage != 0) return cudaSuccess;
        if ((e = cudaMemsetAsync(tb.decode_bits, 0, decode_words * 4, s)) != cudaSuccess) return e;
        if ((e = cudaMemsetAsync(scratch, 0, (4 + 2 * (size_t)n_tiles) * 8, s)) != cudaSuccess) return e;   // ticket and the two status tables
        scan_tokens_kernel<<<n_tiles, P1_THREADS, 0, s>>>(raw, (u64)n, tb, scratch, res);
        g_parse_launches++;
        return cudaGetLastError();
    }
    // fast path: every word of both tables is written before it is read, nothing to clear
    u64 *tile_pre = scratch + 4;
    uint32_t *tile_agg = reinterpret_cast<uint32_t *>(scratch + 4 + n_tiles);
    uint32_t g_count = 0, g_emit = 0;
    if ((e = fast_scan_grid(&g_count, &g_emit)) != cudaSuccess) return e;
    // per-chunk masks handed from count to emit, behind the two tile tables (p1_scratch_words covers them)
    const size_t chunks = (size_t)n_tiles * P1_THREADS;
    u64 *masks = scratch + 4 + n_tiles + (n_tiles + 1) / 2;
    const ChunkStore cs{masks, masks + chunks, masks + 2 * chunks, masks + 3 * chunks, reinterpret_cast<uint32_t *>(masks + 4 * chunks)};
    if (stage == 0)   // (also clears the "needs decoding" bitmap the emit kernel sets bits in)
        scan_count_kernel<<<std::min(n_tiles, g_count), P1_THREADS, 0, s>>>(raw, n_tiles, tile_agg, tb.decode_bits, (uint32_t)decode_words, cs, res);
    else if (stage == 1) scan_prefix_kernel<<<1, PFX_THREADS, 0, s>>>(tile_agg, n_tiles, tile_pre, res);
    else scan_emit_kernel<<<std::min(n_tiles, g_emit), P1_THREADS, 0, s>>>(raw, (u64)n, n_tiles, tb, tile_pre, cs, res);
    g_parse_launches++;
    return cudaGetLastError();
}

cudaError_t launch_token_prefix(TokenBuffers tb, uint32_t n_tok, TilePrefix *tiles, unsigned long long *scratch, ParseResult *res,
                                cudaStream_t s) {
    const uint32_t blocks = tok_tiles(n_tok);
    TokAgg *agg = reinterpret_cast<TokAgg *>(scratch);   // tok_scratch_words(n_tok) words hold it
    token_agg_kernel<<<blocks, TP_THREADS, 0, s>>>(tb, n_tok, agg);
    token_blocks_kernel<<<1, TBP_THREADS, 0, s>>>(agg, blocks, tiles, res);
    g_parse_launches += 2;
    return cudaGetLastError();
}

cudaError_t launch_brackets(TokenBuffers tb, uint32_t n_tok, const TilePrefix *tiles, BracketBuffers bb, uint32_t n_br, cudaStream_t s) {
    cudaError_t e;
    // match / parent / row / ct are four equal slices of one block (engine.cu): "none" everywhere.
    // For ct that reads as container kind 3, which rows_kernel treats as the top level.
    if ((e = cudaMemsetAsync(bb.match, 0xff, (size_t)(bb.ct - bb.match) * 4 + (size_t)n_br * 4, s)) != cudaSuccess) return e;
    brackets_kernel<<<(n_tok + TP_BLOCK - 1) / TP_BLOCK, TP_THREADS, 0, s>>>(tb, n_tok, tiles, bb);
    g_parse_launches++;
    return cudaGetLastError();
}

cudaError_t launch_sort_brackets(BracketBuffers bb, uint32_t n_br, uint32_t max_level, unsigned long long *scratch, bool *in_a,
                                 cudaStream_t s) {
    *in_a = true;
    if (n_br == 0) return cudaSuccess;
    int bits = 0;
    while ((1u << bits) <= max_level) bits++;
    const uint32_t tiles = (n_br + SCAN_TILE - 1) / SCAN_TILE;
    uint32_t *hist = reinterpret_cast<uint32_t *>(scratch);   // RS_BUCKETS x tiles words
    for (int shift = 0; shift < bits; shift += RS_BITS) {
        const uint32_t *key = *in_a ? bb.key_a : bb.key_b, *val = *in_a ? bb.sort_a : bb.sort_b;
        uint32_t *key_o = *in_a ? bb.key_b : bb.key_a, *val_o = *in_a ? bb.sort_b : bb.sort_a;
        radix_hist_kernel<<<tiles, SCAN_THREADS, 0, s>>>(key, n_br, shift, hist, tiles);
        radix_base_kernel<<<1, RB_THREADS, 0, s>>>(hist, RS_BUCKETS * tiles);
        radix_scatter_kernel<<<tiles, SCAN_THREADS, 0, s>>>(key, val, n_br, shift, hist, tiles, key_o, val_o);
        g_parse_launches += 3;
        *in_a = !*in_a;
    }
    return cudaGetLastError();
}

cudaError_t launch_match(BracketBuffers bb, const uint32_t *sorted, const uint32_t *sorted_key, uint32_t n_br, ParseResult *res,
                         cudaStream_t s) {
    if (n_br == 0) return cudaSuccess;
    match_kernel<<<(n_br + 255) / 256, 256, 0, s>>>(bb, sorted, sorted_key, n_br, res);
    g_parse_launches++;
    return cudaGetLastError();
}

cudaError_t launch_decode_list(TokenBuffers tb, uint32_t n_tok, DecodeBuffers db, uint32_t *counts, uint32_t *offs,
                               unsigned long long *scratch, cudaStream_t s) {
    const uint32_t words = (n_tok + 31) / 32;
    decode_count_kernel<<<(words + 255) / 256, 256, 0, s>>>(tb.decode_bits, words, counts);
    g_parse_launches++;
    cudaError_t e = scan_u32(counts, offs, words, scratch, s);  // total -> scratch[1]
    if (e != cudaSuccess) return e;
    decode_list_kernel<<<(words + 255) / 256, 256, 0, s>>>(tb.decode_bits, offs, words, db.tok);
    g_parse_launches++;
    return cudaGetLastError();
}

cudaError_t launch_decode_lengths(const uint8_t *raw, TokenBuffers tb, DecodeBuffers db, uint32_t n_dec, unsigned long long *scratch,
                                  cudaStream_t s) {
    decode_len_kernel<<<(n_dec + 127) / 128, 128, 0, s>>>(raw, tb, db, n_dec);
    g_parse_launches++;
    return scan_u32(db.len, db.off, n_dec, scratch, s);  // total decoded bytes -> scratch[1]
}

cudaError_t launch_decode_write(uint8_t *arena, uint32_t side_base, TokenBuffers tb, DecodeBuffers db, uint32_t n_dec, cudaStream_t s) {
    decode_write_kernel<<<(n_dec + 127) / 128, 128, 0, s>>>(arena, side_base, tb, db, n_dec);
    g_parse_launches++;
    return cudaGetLastError();
}

cudaError_t launch_rows(const uint8_t *raw, size_t n, TokenBuffers tb, uint32_t n_tok, const TilePrefix *tiles, BracketBuffers bb,
                        DecodeBuffers db, uint32_t n_dec, RowBuffers rb, uint32_t n_rows, ParseResult *res, cudaStream_t s) {
    rows_kernel<<<(n_tok + RW_BLOCK - 1) / RW_BLOCK, TP_THREADS, 0, s>>>(raw, (u64)n, tb, n_tok, tiles, bb, db, n_dec, rb, n_rows, res);
    g_parse_launches++;
    return cudaGetLastError();
}

cudaError_t launch_key_order(const uint8_t *arena, RowBuffers rb, uint32_t n_rows, BracketBuffers bb, ParseResult *res, cudaStream_t s) {
    key_order_kernel<<<(n_rows + 255) / 256, 256, 0, s>>>(arena, rb, n_rows, bb, res);
    g_parse_launches++;
    return cudaGetLastError();
}

size_t parse_scan_scratch_words(uint32_t n) { return 4 * scan_scratch_words_u32(n); }   // (the radix sort keeps 8 x u32 per tile)

}  // namespace tk
