// scan_fast.cuh -- P1, fast path: bytes -> tokens in three launches with no kernel waiting on another.
// Included by parse.cu inside its anonymous namespace (it uses find_escaped, prefix_xor,
// token_mask and the tables defined there).
//
// A single-pass scan has to hand the string state and the token count from tile to tile IN ORDER;
// measured on config #4 the blocks of such a kernel spent most of their time waiting for their
// predecessors (60 % of the executed instructions were the polling loop).  Here the chain is cut:
//   count   every tile, independently: (quote parity, tokens if entered outside a string, tokens if
//           entered inside) -> one 32-bit word per tile;
//   prefix  one block folds those words in order: state entering each tile, tokens before it;
//   emit    every tile, independently, recomputes its masks (the text is still in L2) and writes
//           its tokens at their final places.
// Both tile kernels are persistent (tables loaded once per block), build the class masks from
// positioned look-ups (g_class_lut: no bit gathering), take a chunk's backslash carry from its left
// neighbour or from the bytes before the chunk, and emit stages tokens in shared memory so that the
// block writes them out coalesced with the kind looked up from the byte.

// Parity of the backslash run that ends right before byte `at` (a multiple of 64): whether the
// first byte of the chunk starting there is escaped.  Looks at most 512 bytes behind; a longer run
// sets res->slow_scan and the caller falls back to scan_tokens_kernel.
__device__ __noinline__ uint32_t carry_before(const uint8_t *raw, u64 at, ParseResult *res) {
    const u64 BS8 = 0x5c5c5c5c5c5c5c5cull;
    uint32_t run = 0;
#pragma unroll 1
    for (int step = 0; step < 64; step++) {
        if (at == 0) return run & 1u;
        const u64 w = __ldg(reinterpret_cast<const u64 *>(raw + at - 8)) ^ BS8;
        if (w) return (run + ((uint32_t)__clzll((long long)w) >> 3)) & 1u;  // little endian: the byte nearest to `at` is the top one
        run += 8;
        at -= 8;
    }
    res->slow_scan = 1;
    return 0;
}

// (quote parity, tokens if entered outside a string, tokens if entered inside) of a run of bytes,
// packed as bit 0 | bits 1..15 | bits 16..30 (a tile has at most 16,384 tokens); `far` precedes
// `near` in the text.  0 is the identity.
__device__ __forceinline__ uint32_t compose(uint32_t far, uint32_t near) {
    const uint32_t swapped = (near & 1u) | ((near >> 15) & 0xfffeu) | ((near & 0xfffeu) << 15);
    const uint32_t x = (far & 1u) ? swapped : near;
    return ((far & ~1u) + (x & ~1u)) | ((far ^ x) & 1u);
}
static_assert(P1_TILE <= 32767, "token counts of a tile must fit the 15-bit fields of compose()");

// The four look-ups of one text word.  Shifts, logic ops and compares all issue on the integer ALU
// pipe (one warp instruction per two cycles per scheduler), the busiest pipe of these kernels, so
// the byte offsets are formed with multiplies -- x >> s as the high half of x * 2^(32-s) -- which
// issue on the FMA pipe, leaving one AND per byte on the ALU pipe.
__device__ __forceinline__ uint32_t class_word(const uint32_t *lut, uint32_t x) {
    const char *b = reinterpret_cast<const char *>(lut);
    const uint32_t i0 = (x * 4u) & 0x3fcu, i1 = __umulhi(x, 1u << 26) & 0x3fcu, i2 = __umulhi(x, 1u << 18) & 0x3fcu,
                   i3 = __umulhi(x, 1u << 10) & 0x3fcu;
    return *reinterpret_cast<const uint32_t *>(b + i0) | *reinterpret_cast<const uint32_t *>(b + 1024 + i1) |
           *reinterpret_cast<const uint32_t *>(b + 2048 + i2) | *reinterpret_cast<const uint32_t *>(b + 3072 + i3);
}

struct ChunkMasks { u64 Q, WS, ST, R; };   // quotes, white space, structural, rare -- one bit per byte of a 64-byte chunk

struct Chunk64 { uint4 q[4]; };   // the 64 text bytes of one thread
__device__ __forceinline__ Chunk64 load_chunk(const uint8_t *chunk_ptr) {
    Chunk64 c;
    const uint4 *p = reinterpret_cast<const uint4 *>(chunk_ptr);
#pragma unroll
    for (int v = 0; v < 4; v++) c.q[v] = __ldg(p + v);
    return c;
}

__device__ __forceinline__ ChunkMasks classify64(const uint32_t *s_lut, const Chunk64 &c) {
    uint32_t Qw[2], WSw[2], STw[2], Rw[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
        uint32_t y[4];
#pragma unroll
        for (int v = 0; v < 2; v++) {
            const uint4 q = c.q[h * 2 + v];
            y[v * 2 + 0] = class_word(s_lut, q.x) | (class_word(s_lut, q.y) << 4);
            y[v * 2 + 1] = class_word(s_lut, q.z) | (class_word(s_lut, q.w) << 4);
        }
        // 4 x 4 byte transpose: byte c of y[r] (class c of text bytes 8r .. 8r+7) -> byte r of class word c
        const uint32_t t01l = __byte_perm(y[0], y[1], 0x5140), t23l = __byte_perm(y[2], y[3], 0x5140);
        const uint32_t t01h = __byte_perm(y[0], y[1], 0x7362), t23h = __byte_perm(y[2], y[3], 0x7362);
        Qw[h] = __byte_perm(t01l, t23l, 0x5410);
        WSw[h] = __byte_perm(t01l, t23l, 0x7632);
        STw[h] = __byte_perm(t01h, t23h, 0x5410);
        Rw[h] = __byte_perm(t01h, t23h, 0x7632);
    }
    ChunkMasks m;
    m.Q = (u64)Qw[0] | ((u64)Qw[1] << 32);
    m.WS = (u64)WSw[0] | ((u64)WSw[1] << 32);
    m.ST = (u64)STw[0] | ((u64)STw[1] << 32);
    m.R = (u64)Rw[0] | ((u64)Rw[1] << 32);
    return m;
}

// What both tile kernels derive from a chunk: real quotes, string state (if the chunk is entered
// outside a string), escaped bytes and the token masks for both entry states.
struct ChunkStrings { u64 RQ, SBx, E, T0, T1; };

__device__ __forceinline__ ChunkStrings chunk_strings(const uint8_t *__restrict__ raw, u64 chunk, uint32_t lane, const uint8_t *s_klut,
                                                      const ChunkMasks &m, u64 B, ParseResult *res, u64 *carry_in = nullptr) {
    const bool all_bs = (B == ~0ull);
    const uint32_t trail = all_bs ? 0u : ((uint32_t)__clzll((long long)~B) & 1u);  // odd trailing backslash run
    const u64 SCc = ~(m.WS | m.ST | m.Q);   // bytes that would be scalar characters outside a string
    const uint32_t mine_nb = (all_bs ? 1u : 0u) | (trail << 1) | ((uint32_t)(SCc >> 63) << 2);
    const uint32_t nb = __shfl_up_sync(FULL, mine_nb, 1);
    u64 carry;
    bool prev_sc_if_out;
    if (lane > 0) {
        prev_sc_if_out = (nb >> 2) & 1u;
        carry = (nb & 1u) ? carry_before(raw, chunk, res) : ((nb >> 1) & 1u);
    } else {
        prev_sc_if_out = chunk ? s_klut[__ldg(raw + chunk - 1)] == T_SCALAR : false;
        carry = carry_before(raw, chunk, res);
    }
    if (carry_in) *carry_in = carry;
    ChunkStrings s;
    s.E = find_escaped(B, carry);
    s.RQ = m.Q & ~s.E;                      // real (unescaped) quotes
    s.SBx = prefix_xor(s.RQ) ^ s.RQ;        // state before each byte, if the chunk started outside
    s.T0 = token_mask(s.RQ, s.SBx, m.WS, m.ST, false, prev_sc_if_out);
    s.T1 = token_mask(s.RQ, s.SBx, m.WS, m.ST, true, prev_sc_if_out);
    return s;
}

__device__ __forceinline__ uint32_t chunk_triple(const ChunkStrings &s) {
    return ((uint32_t)__popcll(s.RQ) & 1u) | ((uint32_t)__popcll(s.T0) << 1) | ((uint32_t)__popcll(s.T1) << 16);
}

__device__ __forceinline__ void load_tables(uint32_t *s_lut, uint8_t *s_klut) {
    for (int i = threadIdx.x; i < 4 * 256; i += P1_THREADS) s_lut[i] = g_class_lut.v[i];
    for (int i = threadIdx.x; i < 256; i += P1_THREADS) s_klut[i] = g_kind_lut.v[i];
}

// ---- count: one packed triple per tile
// What count hands to emit per 64-byte chunk, so that emit does not classify the text again: the
// token masks for both entry states, the string state and the white-space mask (T0 & T1 are the
// real quotes).  A chunk with a rare byte, or one entered right after an odd run of backslashes,
// is marked in rare_bits (one bit per chunk, a word per warp) and emit classifies it itself.
struct ChunkStore { u64 *T0, *T1, *SBx, *WS; uint32_t *rare_bits; };

__global__ void __launch_bounds__(P1_THREADS) scan_count_kernel(const uint8_t *__restrict__ raw, uint32_t n_tiles, uint32_t *__restrict__ tile_agg,
                                                               uint32_t *__restrict__ decode_bits, uint32_t decode_words, ChunkStore cs,
                                                               ParseResult *res) {
    __shared__ uint32_t s_lut[4 * 256];
    __shared__ uint8_t s_klut[256];
    __shared__ uint32_t s_warp_agg[2][P1_THREADS / 32];
    load_tables(s_lut, s_klut);
    // the "needs decoding" bitmap, for the emit kernel that follows on the stream
    for (uint32_t w = blockIdx.x * P1_THREADS + threadIdx.x; w < decode_words; w += gridDim.x * P1_THREADS) decode_bits[w] = 0u;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t it = 0;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it ^= 1) {
        const u64 chunk = (u64)tile * P1_TILE + (u64)threadIdx.x * P1_BYTES_PER_THREAD;
        const ChunkMasks m = classify64(s_lut, load_chunk(raw + chunk));
        u64 B = 0;
        for (u64 r = m.R; r; r &= r - 1) {   // only backslashes matter here
            const int i = __ffsll((long long)r) - 1;
            if (__ldg(raw + chunk + i) == '\\') B |= 1ull << i;
        }
        u64 carry_in;
        const ChunkStrings s = chunk_strings(raw, chunk, lane, s_klut, m, B, res, &carry_in);
        {
            const size_t ci = (size_t)tile * P1_THREADS + threadIdx.x;
            cs.T0[ci] = s.T0; cs.T1[ci] = s.T1; cs.SBx[ci] = s.SBx; cs.WS[ci] = m.WS;
            const unsigned rare = __ballot_sync(FULL, m.R != 0 || carry_in != 0);
            if (lane == 0) cs.rare_bits[ci >> 5] = rare;
        }
        uint32_t agg = chunk_triple(s);
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {   // ordered fold: lane L ends up with lanes L .. L + 2d - 1
            const uint32_t v = __shfl_down_sync(FULL, agg, d);
            if (lane + d < 32) agg = compose(agg, v);
        }
        if (lane == 0) s_warp_agg[it][warp] = agg;
        __syncthreads();   // (the two buffers alternate, so one barrier per tile is enough)
        if (threadIdx.x == 0) {
            uint32_t t = 0;
#pragma unroll
            for (int w = 0; w < P1_THREADS / 32; w++) t = compose(t, s_warp_agg[it][w]);
            tile_agg[tile] = t;
        }
    }
}

// ---- prefix: state entering every tile and tokens before it (one block; the fold is ordered)
constexpr int PFX_THREADS = 1024, PFX_ITEMS = 8;
struct Tri { uint32_t p, c0, c1; };   // the triple unpacked: the counts of a whole text do not fit 15 bits
__device__ __forceinline__ Tri tri_compose(const Tri &far, const Tri &near) {
    Tri r;
    r.p = far.p ^ near.p;
    r.c0 = far.c0 + (far.p ? near.c1 : near.c0);
    r.c1 = far.c1 + (far.p ? near.c0 : near.c1);
    return r;
}
__device__ __forceinline__ Tri tri_shfl_up(const Tri &t, int d) {
    Tri r;
    r.p = __shfl_up_sync(FULL, t.p, d); r.c0 = __shfl_up_sync(FULL, t.c0, d); r.c1 = __shfl_up_sync(FULL, t.c1, d);
    return r;
}
__device__ __forceinline__ Tri tri_warp_scan(Tri t, uint32_t lane) {   // inclusive
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const Tri f = tri_shfl_up(t, d);
        if (lane >= (uint32_t)d) t = tri_compose(f, t);
    }
    return t;
}

__global__ void __launch_bounds__(PFX_THREADS) scan_prefix_kernel(const uint32_t *__restrict__ tile_agg, uint32_t n_tiles, u64 *__restrict__ tile_pre,
                                                                 ParseResult *res) {
    __shared__ Tri s_warp[PFX_THREADS / 32];   // inclusive per warp, then exclusive
    __shared__ Tri s_total;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Tri run = {0, 0, 0};   // everything before this round (the same in every thread)
    for (uint32_t base = 0; base < n_tiles; base += PFX_THREADS * PFX_ITEMS) {
        const uint32_t first = base + threadIdx.x * PFX_ITEMS;
        uint32_t a[PFX_ITEMS];
        Tri mine = {0, 0, 0};
#pragma unroll
        for (int j = 0; j < PFX_ITEMS; j++) {
            a[j] = first + j < n_tiles ? tile_agg[first + j] : 0u;
            const Tri t = {a[j] & 1u, (a[j] >> 1) & 0x7fffu, a[j] >> 16};
            mine = tri_compose(mine, t);
        }
        const Tri incl = tri_warp_scan(mine, lane);
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const Tri w = s_warp[lane];
            const Tri wi = tri_warp_scan(w, lane);
            Tri we = tri_shfl_up(wi, 1);
            if (lane == 0) we = Tri{0, 0, 0};
            s_warp[lane] = we;
            if (lane == 31) s_total = wi;
        }
        __syncthreads();
        Tri ex = tri_shfl_up(incl, 1);   // the earlier lanes of this warp
        if (lane == 0) ex = Tri{0, 0, 0};
        ex = tri_compose(run, tri_compose(s_warp[warp], ex));
#pragma unroll
        for (int j = 0; j < PFX_ITEMS; j++) {
            if (first + j < n_tiles) {
                // the text starts outside a string: the "entered outside" count of the prefix is the real one
                tile_pre[first + j] = (u64)ex.p | ((u64)ex.c0 << 1);
                const Tri t = {a[j] & 1u, (a[j] >> 1) & 0x7fffu, a[j] >> 16};
                ex = tri_compose(ex, t);
                if (first + j + 1 == n_tiles) { res->n_tokens = ex.c0; res->end_in_string = ex.p; }
            }
        }
        run = tri_compose(run, s_total);
        __syncthreads();
    }
}

// ---- emit: tokens of every tile at their final places, string checks, decode marks
__global__ void __launch_bounds__(P1_THREADS) scan_emit_kernel(const uint8_t *__restrict__ raw, u64 n, uint32_t n_tiles, TokenBuffers tb,
                                                              const u64 *__restrict__ tile_pre, ChunkStore cs, ParseResult *res) {
    __shared__ uint32_t s_lut[4 * 256];
    __shared__ uint8_t s_klut[256];
    constexpr int TOK_STAGE = 2048;                     // (JSON text averages one token per 8 to 9 bytes: 8 KiB -> about 1,000)
    __shared__ uint16_t s_tok[TOK_STAGE];               // tokens in order: byte offset in the tile | closing-quote bit
    __shared__ uint32_t s_warp_agg[2][P1_THREADS / 32];
    constexpr uint32_t TOK_IN_STRING = 0x8000u;
    static_assert(P1_TILE <= 0x8000, "a staged token keeps its offset in 15 bits");
    load_tables(s_lut, s_klut);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t it = 0;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it ^= 1) {
        const u64 tile_base = (u64)tile * P1_TILE;
        const u64 chunk = tile_base + (u64)threadIdx.x * P1_BYTES_PER_THREAD;
        const u64 pre = tile_pre[tile];
        const uint32_t V = (uint32_t)pre & 1u;          // in-string state entering the tile
        const u64 base = pre >> 1;                      // tokens before the tile
        // What count left for this chunk; a warp with a rare chunk classifies its text again (all its
        // lanes: the neighbours' backslash runs travel by shuffle).
        const size_t ci = (size_t)tile * P1_THREADS + threadIdx.x;
        const unsigned rare_warp = __ldg(cs.rare_bits + (ci >> 5));
        ChunkMasks m;
        ChunkStrings s;
        u64 B = 0, CT = 0, NA = 0, HT = 0;
        if (rare_warp == 0) {
            s.T0 = __ldg(cs.T0 + ci); s.T1 = __ldg(cs.T1 + ci); s.SBx = __ldg(cs.SBx + ci); m.WS = __ldg(cs.WS + ci);
            s.RQ = s.T0 & s.T1;   // a real quote is a token whatever the state; anything else only outside a string
            s.E = 0;
            m.Q = s.RQ; m.ST = 0; m.R = 0;
        } else {
            m = classify64(s_lut, load_chunk(raw + chunk));
            // rare bytes, told apart one by one
            for (u64 r = m.R; r; r &= r - 1) {
                const int i = __ffsll((long long)r) - 1;
                const uint32_t c = __ldg(raw + chunk + i);
                const u64 bit = 1ull << i;
                if (c == '\\') B |= bit;
                else if (c >= 0x80) NA |= bit;
                else if (c < 0x20) CT |= bit;
                else HT |= bit;
            }
            s = chunk_strings(raw, chunk, lane, s_klut, m, B, res);
        }

        // exclusive ordered scan of the triples over the tile
        uint32_t incl = chunk_triple(s);
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(FULL, incl, d);
            if (lane >= (uint32_t)d) incl = compose(v, incl);
        }
        uint32_t excl = __shfl_up_sync(FULL, incl, 1);
        if (lane == 0) excl = 0;
        if (lane == 31) s_warp_agg[it][warp] = incl;
        // (two alternating buffers: this barrier also tells that the previous tile's write-out has
        //  finished with s_tok)
        __syncthreads();
        uint32_t wbefore = 0, tile_agg = 0;
#pragma unroll
        for (uint32_t i = 0; i < P1_THREADS / 32; i++) {
            if (i == warp) wbefore = tile_agg;
            tile_agg = compose(tile_agg, s_warp_agg[it][i]);
        }
        excl = compose(wbefore, excl);
        const uint32_t cnt = V ? (tile_agg >> 16) : ((tile_agg >> 1) & 0x7fffu);   // tokens of the tile
        const bool s_in = ((V ^ excl) & 1u) != 0;
        const u64 T = s_in ? s.T1 : s.T0;
        const u64 SB = s_in ? ~s.SBx : s.SBx;
        uint32_t lr = V ? (excl >> 16) : ((excl >> 1) & 0x7fffu);   // rank of this thread's first token in the tile
        const u64 idx = base + lr;

        // ---- string body checks (scanner.go stateInString / stateInStringEsc*)
        const u64 content = SB & ~s.RQ;
        bool bad = (CT & content) != 0;
        for (u64 w = m.WS & content; w; w &= w - 1)  // white space inside a string: only the space itself is allowed raw
            if (__ldg(raw + chunk + (__ffsll((long long)w) - 1)) != ' ') bad = true;
        for (u64 e = s.E & content; e; e &= e - 1) {
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

        // ---- strings that are not plain: a backslash, a non-ASCII byte or one of < > & inside
        for (u64 dmask = (B | NA | HT) & content; dmask;) {
            const int i = __ffsll((long long)dmask) - 1;
            const u64 t = idx + (u64)__popcll(T & low_mask(i)) - 1;  // the string's opening-quote token
            if (t < tb.cap) atomicOr(tb.decode_bits + (t >> 5), 1u << (t & 31));
            res->n_decode = 1;
            const u64 nq = s.RQ & ~low_mask(i + 1);
            dmask = nq ? (dmask & ~low_mask(__ffsll((long long)nq) - 1)) : 0;
        }

        // ---- tokens: staged in order in shared memory, then written out by the whole block.  The
        //      staging area holds TOK_STAGE tokens; a denser tile takes more than one round.
        for (uint32_t b0 = 0; b0 < cnt; b0 += TOK_STAGE) {
            if (b0) __syncthreads();   // the previous round's write-out has finished with s_tok
            {
                const uint32_t off = threadIdx.x * P1_BYTES_PER_THREAD;
                const uint32_t sl = (uint32_t)SB, sh = (uint32_t)(SB >> 32);
                uint32_t r = lr - b0;   // slot of this thread's next token (wraps below the round's window)
                for (uint32_t t = (uint32_t)T; t; t &= t - 1, r++) {
                    const uint32_t i = __ffs(t) - 1;
                    if (r < TOK_STAGE) s_tok[r] = (uint16_t)((off + i) | (((sl >> i) & 1u) << 15));
                }
                for (uint32_t t = (uint32_t)(T >> 32); t; t &= t - 1, r++) {
                    const uint32_t i = __ffs(t) - 1;
                    if (r < TOK_STAGE) s_tok[r] = (uint16_t)((off + 32 + i) | (((sh >> i) & 1u) << 15));
                }
            }
            __syncthreads();
            // position, and the kind of the byte there (only a closing quote is a token whose
            // state-before is "inside")
            const uint32_t m_round = min((uint32_t)TOK_STAGE, cnt - b0);
            const u64 g0 = base + b0;
#pragma unroll 4
            for (uint32_t j = threadIdx.x; j < m_round; j += P1_THREADS) {
                const uint32_t e = s_tok[j], off = e & (TOK_IN_STRING - 1);
                const uint32_t k = s_klut[__ldg(raw + tile_base + off)] + (e >> 15);
                if (g0 + j < tb.cap) {
                    tb.pos[g0 + j] = (uint32_t)(tile_base + off);
                    tb.kind[g0 + j] = (uint8_t)k;
                }
            }
        }
    }
}
