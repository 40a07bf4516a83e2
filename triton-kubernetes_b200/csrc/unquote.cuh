// unquote.cuh -- encoding/json's unquote on the device: the bytes between the quotes of a JSON string turned into
// the string they spell (escapes, surrogate pairs, U+FFFD for invalid UTF-8).  Used by the general parser's decode
// kernels (parse.cu) and by the line pass for the few strings it cannot keep in place (lines.cu).
#pragma once
#include <cstdint>

namespace tk {

__device__ __forceinline__ uint32_t hex_val(uint32_t c) { return c <= '9' ? c - '0' : (c | 0x20) - 'a' + 10; }
__device__ __forceinline__ bool is_hex(uint32_t c) { return (c >= '0' && c <= '9') || ((c | 0x20) >= 'a' && (c | 0x20) <= 'f'); }

__device__ __forceinline__ uint32_t utf8_valid_len(const uint8_t *s, uint32_t i, uint32_t n) {
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

__device__ __forceinline__ uint32_t put_rune(uint8_t *dst, uint32_t w, uint32_t r) {
    if (r > 0x10FFFF || (r >= 0xD800 && r <= 0xDFFF)) r = 0xFFFD;
    if (r < 0x80) { if (dst) dst[w] = (uint8_t)r; return w + 1; }
    if (r < 0x800) { if (dst) { dst[w] = 0xC0 | (r >> 6); dst[w + 1] = 0x80 | (r & 0x3F); } return w + 2; }
    if (r < 0x10000) { if (dst) { dst[w] = 0xE0 | (r >> 12); dst[w + 1] = 0x80 | ((r >> 6) & 0x3F); dst[w + 2] = 0x80 | (r & 0x3F); } return w + 3; }
    if (dst) { dst[w] = 0xF0 | (r >> 18); dst[w + 1] = 0x80 | ((r >> 12) & 0x3F); dst[w + 2] = 0x80 | ((r >> 6) & 0x3F); dst[w + 3] = 0x80 | (r & 0x3F); }
    return w + 4;
}

// encoding/json unquote: s[0, n) is the text between the quotes.  dst == nullptr: length only.
__device__ inline uint32_t unquote(const uint8_t *s, uint32_t n, uint8_t *dst) {
    uint32_t w = 0;
    for (uint32_t i = 0; i < n;) {
        const uint32_t c = s[i];
        if (c == '\\') {
            if (i + 1 >= n) break;  // cannot happen in text that passed P1
            const uint32_t e = s[i + 1];
            i += 2;
            uint32_t r;
            switch (e) {
            case 'b': r = '\b'; break;
            case 'f': r = '\f'; break;
            case 'n': r = '\n'; break;
            case 'r': r = '\r'; break;
            case 't': r = '\t'; break;
            case 'u': {
                if (i + 4 > n) { i = n; continue; }
                r = (hex_val(s[i]) << 12) | (hex_val(s[i + 1]) << 8) | (hex_val(s[i + 2]) << 4) | hex_val(s[i + 3]);
                i += 4;
                if (r >= 0xD800 && r <= 0xDFFF) {
                    bool pair = r < 0xDC00 && i + 6 <= n && s[i] == '\\' && s[i + 1] == 'u' && is_hex(s[i + 2]) && is_hex(s[i + 3]) &&
                                is_hex(s[i + 4]) && is_hex(s[i + 5]);
                    uint32_t r2 = 0;
                    if (pair) {
                        r2 = (hex_val(s[i + 2]) << 12) | (hex_val(s[i + 3]) << 8) | (hex_val(s[i + 4]) << 4) | hex_val(s[i + 5]);
                        pair = r2 >= 0xDC00 && r2 <= 0xDFFF;
                    }
                    if (pair) { r = 0x10000 + ((r - 0xD800) << 10) + (r2 - 0xDC00); i += 6; }
                    else r = 0xFFFD;
                }
                break;
            }
            default: r = e;  // " \ /
            }
            w = put_rune(dst, w, r);
            continue;
        }
        if (c < 0x80) { if (dst) dst[w] = (uint8_t)c; w++; i++; continue; }
        const uint32_t l = utf8_valid_len(s, i, n);
        if (l == 0) { w = put_rune(dst, w, 0xFFFD); i++; continue; }
        for (uint32_t k = 0; k < l; k++) { if (dst) dst[w + k] = s[i + k]; }
        w += l;
        i += l;
    }
    return w;
}


}  // namespace tk
