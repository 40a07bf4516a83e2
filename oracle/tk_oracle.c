/*
 * tk_oracle.c -- CPU restatement of the reference's state.State path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may build, load or call it, and only as the checker or the
 * timed CPU baseline.  The product (triton-kubernetes_b200/csrc) never links it.
 *
 * What it restates (paths relative to /root/reference):
 *   state/state.go:15-160                          State: New/Get/Set* /Add* /Delete/Bytes/Clusters/Nodes
 *   vendor/github.com/Jeffail/gabs/gabs.go:81-114  Path/Search incl. array fan-out
 *   vendor/github.com/Jeffail/gabs/gabs.go:164-173 ChildrenMap
 *   vendor/github.com/Jeffail/gabs/gabs.go:180-208 Set/SetP (nil root, nil intermediates, collisions)
 *   vendor/github.com/Jeffail/gabs/gabs.go:276-301 Delete/DeleteP
 *   vendor/github.com/Jeffail/gabs/gabs.go:385-402 Bytes/BytesIndent ("{}" for a nil root)
 *   vendor/github.com/Jeffail/gabs/gabs.go:425-433 ParseJSON
 *   create/node.go:350-380                         getNewHostnames
 *
 * Third-party arithmetic that is NOT under /root/reference: Go's standard library
 * encoding/json (toolchain pinned only as "go: 1.9.x", .travis.yml:2-3), reached at
 * gabs.go:387,397,428.  Its published behaviour is restated here: the scanner state
 * machine and its error texts, Unmarshal into interface{} (objects -> maps with the
 * last duplicate key winning, numbers -> float64, invalid UTF-8 -> U+FFFD, \u escapes
 * with surrogate pairing), Marshal (map keys sorted bytewise, HTML-safe string
 * escaping, ES6-style float formatting) and Indent ("" prefix, "\t" indent).
 *
 * PARITY PINNING: the method semantics are pinned by the reference's own tests
 * (state/state_test.go, create/node_test.go, destroy/ and get/ *_test.go files and the
 * compact-output vectors in the vendored gabs_test.go); tests/test_oracle_reference_vectors.py
 * replays them.  The BYTES of State.Bytes() (tab indented) are asserted by no reference
 * test and the Go toolchain is absent here, so for that output parity is UNPINNED by the
 * reference; it rests on the rules above plus a cross-check against CPython's json module.
 *
 * Deliberate, documented deviations:
 *   - nesting deeper than ORA_MAX_DEPTH (10000) is rejected with "exceeded max depth"
 *     (Go 1.9 has no limit and would exhaust its stack; Go >= 1.15 uses this same limit);
 *   - Clusters()/Nodes() return pairs sorted by name; on duplicate names the bytewise
 *     greatest module key wins (Go: random map order decides; SURVEY.md 8a note 3).
 *
 * "Records" model Go struct values stored with SetP (SURVEY.md 8a note 1): members keep
 * their given (declaration) order when serialised and are opaque to Search/Get/Clusters/
 * Nodes, Set-through and Delete-through until the document is re-parsed.
 */
#define _GNU_SOURCE
#include <errno.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ORA_MAX_DEPTH 10000

typedef enum { V_NULL = 0, V_BOOL, V_NUM, V_STR, V_ARR, V_MAP, V_REC } vkind;

typedef struct val val;
typedef struct {
    char *key;
    size_t klen;
    uint64_t hash;
    val *v;
} entry;

struct val {
    vkind k;
    int b;          /* V_BOOL */
    double num;     /* V_NUM */
    char *s;        /* V_STR */
    size_t slen;
    val **items;    /* V_ARR */
    size_t n, cap;  /* V_ARR: items; V_MAP / V_REC: entries */
    entry *ent;     /* V_MAP (unordered, unique keys) / V_REC (ordered, as given) */
    uint32_t *idx;  /* V_MAP hash index: 0 empty, 0xffffffff tombstone, else entry+1 */
    size_t idxcap;
    size_t tombs;
};

typedef struct {
    val *root; /* NULL == Go nil */
} doc;

/* ------------------------------------------------------------------ util */

static void *xmalloc(size_t n) {
    void *p = malloc(n ? n : 1);
    if (!p) { fprintf(stderr, "oracle: out of memory\n"); abort(); }
    return p;
}
static void *xrealloc(void *p, size_t n) {
    p = realloc(p, n ? n : 1);
    if (!p) { fprintf(stderr, "oracle: out of memory\n"); abort(); }
    return p;
}
static char *xmemdup(const char *s, size_t n) {
    char *p = xmalloc(n + 1);
    if (n) memcpy(p, s, n);
    p[n] = 0;
    return p;
}

typedef struct { char *p; size_t n, cap; } buf;
static void buf_put(buf *b, const void *s, size_t n) {
    if (b->n + n + 1 > b->cap) {
        size_t c = b->cap ? b->cap : 256;
        while (c < b->n + n + 1) c *= 2;
        b->p = xrealloc(b->p, c);
        b->cap = c;
    }
    if (n) memcpy(b->p + b->n, s, n);
    b->n += n;
    b->p[b->n] = 0;
}
static void buf_putc(buf *b, char c) { buf_put(b, &c, 1); }
static void buf_puts(buf *b, const char *s) { buf_put(b, s, strlen(s)); }

static uint64_t fnv1a(const char *s, size_t n) {
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; i++) { h ^= (unsigned char)s[i]; h *= 1099511628211ull; }
    return h ? h : 1;
}

/* ---------------------------------------------------------------- values */

static val *v_new(vkind k) {
    val *v = xmalloc(sizeof *v);
    memset(v, 0, sizeof *v);
    v->k = k;
    return v;
}
static val *v_str(const char *s, size_t n) {
    val *v = v_new(V_STR);
    v->s = xmemdup(s, n);
    v->slen = n;
    return v;
}
static void v_free(val *v) {
    if (!v) return;
    if (v->k == V_STR) free(v->s);
    if (v->k == V_ARR) { for (size_t i = 0; i < v->n; i++) v_free(v->items[i]); free(v->items); }
    if (v->k == V_MAP || v->k == V_REC) {
        for (size_t i = 0; i < v->n; i++) { free(v->ent[i].key); v_free(v->ent[i].v); }
        free(v->ent);
        free(v->idx);
    }
    free(v);
}
static void arr_push(val *a, val *x) {
    if (a->n == a->cap) { a->cap = a->cap ? a->cap * 2 : 4; a->items = xrealloc(a->items, a->cap * sizeof *a->items); }
    a->items[a->n++] = x;
}

/* map: Go map[string]interface{} -- unordered, unique keys */
static void map_rehash(val *m, size_t want) {
    size_t cap = 8;
    while (cap < want * 2) cap *= 2;
    free(m->idx);
    m->idx = xmalloc(cap * sizeof *m->idx);
    memset(m->idx, 0, cap * sizeof *m->idx);
    m->idxcap = cap;
    m->tombs = 0;
    for (size_t i = 0; i < m->n; i++) {
        size_t s = m->ent[i].hash & (cap - 1);
        while (m->idx[s]) s = (s + 1) & (cap - 1);
        m->idx[s] = (uint32_t)(i + 1);
    }
}
/* returns index slot of key or (size_t)-1 */
static size_t map_slot(const val *m, const char *k, size_t klen, uint64_t h) {
    if (!m->idxcap) return (size_t)-1;
    size_t s = h & (m->idxcap - 1);
    for (;;) {
        uint32_t e = m->idx[s];
        if (e == 0) return (size_t)-1;
        if (e != 0xffffffffu) {
            const entry *en = &m->ent[e - 1];
            if (en->hash == h && en->klen == klen && memcmp(en->key, k, klen) == 0) return s;
        }
        s = (s + 1) & (m->idxcap - 1);
    }
}
/* Go: v, present := m[k]; absent keys read as nil (NULL here) */
static val *map_get(const val *m, const char *k, size_t klen, int *present) {
    size_t s = map_slot(m, k, klen, fnv1a(k, klen));
    if (present) *present = (s != (size_t)-1);
    if (s == (size_t)-1) return NULL;
    return m->ent[m->idx[s] - 1].v;
}
/* Go: m[k] = v (v may be NULL == nil interface, i.e. JSON null) */
static void map_put(val *m, const char *k, size_t klen, val *v) {
    uint64_t h = fnv1a(k, klen);
    size_t s = map_slot(m, k, klen, h);
    if (s != (size_t)-1) {
        entry *en = &m->ent[m->idx[s] - 1];
        v_free(en->v);
        en->v = v;
        return;
    }
    if (m->n == m->cap) { m->cap = m->cap ? m->cap * 2 : 4; m->ent = xrealloc(m->ent, m->cap * sizeof *m->ent); }
    if ((m->n + m->tombs + 1) * 2 > m->idxcap) map_rehash(m, m->n + 1);
    m->ent[m->n].key = xmemdup(k, klen);
    m->ent[m->n].klen = klen;
    m->ent[m->n].hash = h;
    m->ent[m->n].v = v;
    m->n++;
    size_t p = h & (m->idxcap - 1);
    while (m->idx[p] && m->idx[p] != 0xffffffffu) p = (p + 1) & (m->idxcap - 1);
    if (m->idx[p] == 0xffffffffu) m->tombs--;
    m->idx[p] = (uint32_t)m->n;
}
/* Go: delete(m, k) -- no-op when absent */
static void map_del(val *m, const char *k, size_t klen) {
    size_t s = map_slot(m, k, klen, fnv1a(k, klen));
    if (s == (size_t)-1) return;
    size_t e = m->idx[s] - 1;
    m->idx[s] = 0xffffffffu;
    m->tombs++;
    free(m->ent[e].key);
    v_free(m->ent[e].v);
    size_t last = m->n - 1;
    if (e != last) {
        /* move the last entry into the hole and repoint its index slot */
        size_t ls = map_slot(m, m->ent[last].key, m->ent[last].klen, m->ent[last].hash);
        m->ent[e] = m->ent[last];
        m->idx[ls] = (uint32_t)(e + 1);
    }
    m->n--;
}
static void rec_push(val *r, const char *k, size_t klen, val *v) {
    if (r->n == r->cap) { r->cap = r->cap ? r->cap * 2 : 4; r->ent = xrealloc(r->ent, r->cap * sizeof *r->ent); }
    r->ent[r->n].key = xmemdup(k, klen);
    r->ent[r->n].klen = klen;
    r->ent[r->n].hash = 0;
    r->ent[r->n].v = v;
    r->n++;
}

/* --------------------------------------------------- UTF-8 (Go unicode/utf8) */

/* utf8.DecodeRune: returns size; *r = 0xFFFD with size 1 for any invalid byte */
static int utf8_decode(const unsigned char *p, size_t n, uint32_t *r) {
    if (n == 0) { *r = 0xFFFD; return 0; }
    unsigned c = p[0];
    if (c < 0x80) { *r = c; return 1; }
    unsigned lo = 0x80, hi = 0xBF;
    int size;
    if (c >= 0xC2 && c <= 0xDF) size = 2;
    else if (c == 0xE0) { size = 3; lo = 0xA0; }
    else if (c >= 0xE1 && c <= 0xEC) size = 3;
    else if (c == 0xED) { size = 3; hi = 0x9F; }
    else if (c >= 0xEE && c <= 0xEF) size = 3;
    else if (c == 0xF0) { size = 4; lo = 0x90; }
    else if (c >= 0xF1 && c <= 0xF3) size = 4;
    else if (c == 0xF4) { size = 4; hi = 0x8F; }
    else { *r = 0xFFFD; return 1; }
    if ((size_t)size > n) { *r = 0xFFFD; return 1; }
    if (p[1] < lo || p[1] > hi) { *r = 0xFFFD; return 1; }
    for (int i = 2; i < size; i++)
        if (p[i] < 0x80 || p[i] > 0xBF) { *r = 0xFFFD; return 1; }
    if (size == 2) *r = ((c & 0x1F) << 6) | (p[1] & 0x3F);
    else if (size == 3) *r = ((c & 0x0F) << 12) | ((p[1] & 0x3F) << 6) | (p[2] & 0x3F);
    else *r = ((c & 0x07) << 18) | ((p[1] & 0x3F) << 12) | ((p[2] & 0x3F) << 6) | (p[3] & 0x3F);
    return size;
}
/* utf8.EncodeRune (surrogates and out-of-range become U+FFFD) */
static void utf8_encode(buf *b, uint32_t r) {
    if (r > 0x10FFFF || (r >= 0xD800 && r <= 0xDFFF)) r = 0xFFFD;
    char t[4];
    if (r < 0x80) { t[0] = (char)r; buf_put(b, t, 1); }
    else if (r < 0x800) { t[0] = (char)(0xC0 | (r >> 6)); t[1] = (char)(0x80 | (r & 0x3F)); buf_put(b, t, 2); }
    else if (r < 0x10000) { t[0] = (char)(0xE0 | (r >> 12)); t[1] = (char)(0x80 | ((r >> 6) & 0x3F)); t[2] = (char)(0x80 | (r & 0x3F)); buf_put(b, t, 3); }
    else { t[0] = (char)(0xF0 | (r >> 18)); t[1] = (char)(0x80 | ((r >> 12) & 0x3F)); t[2] = (char)(0x80 | ((r >> 6) & 0x3F)); t[3] = (char)(0x80 | (r & 0x3F)); buf_put(b, t, 4); }
}

/* ------------------------------------------- Go encoding/json scanner (checkValid) */

/* quoteChar from encoding/json/scanner.go */
static void quote_char(unsigned char c, char *out /* >= 16 */) {
    if (c == '\'') { strcpy(out, "'\\''"); return; }
    if (c == '"') { strcpy(out, "'\"'"); return; }
    /* strconv.Quote(string(rune(c))) with the outer quotes swapped for ' */
    if (c == '\\') { strcpy(out, "'\\\\'"); return; }
    if (c == '\a') { strcpy(out, "'\\a'"); return; }
    if (c == '\b') { strcpy(out, "'\\b'"); return; }
    if (c == '\f') { strcpy(out, "'\\f'"); return; }
    if (c == '\n') { strcpy(out, "'\\n'"); return; }
    if (c == '\r') { strcpy(out, "'\\r'"); return; }
    if (c == '\t') { strcpy(out, "'\\t'"); return; }
    if (c == '\v') { strcpy(out, "'\\v'"); return; }
    if (c < 0x20 || c == 0x7f) { sprintf(out, "'\\x%02x'", c); return; }
    if (c < 0x80) { sprintf(out, "'%c'", c); return; }
    /* byte -> rune U+0080..U+00FF; strconv.IsPrint decides */
    if (c < 0xA0) { sprintf(out, "'\\u00%02x'", c); return; }      /* C1 controls: not printable */
    if (c == 0xAD) { strcpy(out, "'\\u00ad'"); return; }             /* soft hyphen: not printable */
    if (c == 0xA0) { strcpy(out, "'\\u00a0'"); return; }             /* nbsp: not printable for Quote */
    out[0] = '\''; out[1] = (char)(0xC0 | (c >> 6)); out[2] = (char)(0x80 | (c & 0x3F)); out[3] = '\''; out[4] = 0;
}

enum { P_OBJKEY, P_OBJVAL, P_ARRVAL };
enum {
    S_BEGIN_VALUE, S_BEGIN_STRING_OR_EMPTY, S_BEGIN_STRING, S_BEGIN_VALUE_OR_EMPTY, S_END_VALUE, S_END_TOP,
    S_IN_STRING, S_IN_STRING_ESC, S_ESC_U, S_ESC_U1, S_ESC_U12, S_ESC_U123,
    S_NEG, S_1, S_0, S_DOT, S_DOT0, S_E, S_ESIGN, S_E0,
    S_T, S_TR, S_TRU, S_F, S_FA, S_FAL, S_FALS, S_N, S_NU, S_NUL, S_ERROR
};
typedef struct {
    int step;
    unsigned char *stack;
    size_t depth, cap;
    int end_top;
    char err[128];
} scanner;

static int is_space(unsigned char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }
static int sc_error(scanner *s, unsigned char c, const char *ctx) {
    char q[16];
    quote_char(c, q);
    snprintf(s->err, sizeof s->err, "invalid character %s %s", q, ctx);
    s->step = S_ERROR;
    return 0;
}
static int sc_push(scanner *s, int p) {
    if (s->depth >= ORA_MAX_DEPTH) { snprintf(s->err, sizeof s->err, "exceeded max depth"); s->step = S_ERROR; return 0; }
    if (s->depth == s->cap) { s->cap = s->cap ? s->cap * 2 : 64; s->stack = xrealloc(s->stack, s->cap); }
    s->stack[s->depth++] = (unsigned char)p;
    return 1;
}
static int sc_step(scanner *s, unsigned char c);
static int sc_end_value(scanner *s, unsigned char c) {
    if (s->depth == 0) { s->step = S_END_TOP; s->end_top = 1; return sc_step(s, c); }
    if (is_space(c)) { s->step = S_END_VALUE; return 1; }
    switch (s->stack[s->depth - 1]) {
    case P_OBJKEY:
        if (c == ':') { s->stack[s->depth - 1] = P_OBJVAL; s->step = S_BEGIN_VALUE; return 1; }
        return sc_error(s, c, "after object key");
    case P_OBJVAL:
        if (c == ',') { s->stack[s->depth - 1] = P_OBJKEY; s->step = S_BEGIN_STRING; return 1; }
        if (c == '}') { s->depth--; s->step = S_END_VALUE; if (s->depth == 0) { s->step = S_END_TOP; s->end_top = 1; } return 1; }
        return sc_error(s, c, "after object key:value pair");
    default:
        if (c == ',') { s->step = S_BEGIN_VALUE; return 1; }
        if (c == ']') { s->depth--; s->step = S_END_VALUE; if (s->depth == 0) { s->step = S_END_TOP; s->end_top = 1; } return 1; }
        return sc_error(s, c, "after array element");
    }
}
static int is_hex(unsigned char c) { return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F'); }
static int sc_begin_value(scanner *s, unsigned char c) {
    if (is_space(c)) return 1;
    switch (c) {
    case '{': s->step = S_BEGIN_STRING_OR_EMPTY; return sc_push(s, P_OBJKEY);
    case '[': s->step = S_BEGIN_VALUE_OR_EMPTY; return sc_push(s, P_ARRVAL);
    case '"': s->step = S_IN_STRING; return 1;
    case '-': s->step = S_NEG; return 1;
    case '0': s->step = S_0; return 1;
    case 't': s->step = S_T; return 1;
    case 'f': s->step = S_F; return 1;
    case 'n': s->step = S_N; return 1;
    }
    if (c >= '1' && c <= '9') { s->step = S_1; return 1; }
    return sc_error(s, c, "looking for beginning of value");
}
static int sc_step(scanner *s, unsigned char c) {
    switch (s->step) {
    case S_BEGIN_VALUE: return sc_begin_value(s, c);
    case S_BEGIN_STRING_OR_EMPTY:
        if (is_space(c)) return 1;
        if (c == '}') { s->stack[s->depth - 1] = P_OBJVAL; return sc_end_value(s, c); }
        /* fallthrough */
    case S_BEGIN_STRING:
        if (is_space(c)) return 1;
        if (c == '"') { s->step = S_IN_STRING; return 1; }
        return sc_error(s, c, "looking for beginning of object key string");
    case S_BEGIN_VALUE_OR_EMPTY:
        if (is_space(c)) return 1;
        if (c == ']') return sc_end_value(s, c);
        return sc_begin_value(s, c);
    case S_END_VALUE: return sc_end_value(s, c);
    case S_END_TOP:
        if (!is_space(c)) return sc_error(s, c, "after top-level value");
        return 1;
    case S_IN_STRING:
        if (c == '"') { s->step = S_END_VALUE; return 1; }
        if (c == '\\') { s->step = S_IN_STRING_ESC; return 1; }
        if (c < 0x20) return sc_error(s, c, "in string literal");
        return 1;
    case S_IN_STRING_ESC:
        switch (c) {
        case 'b': case 'f': case 'n': case 'r': case 't': case '\\': case '/': case '"': s->step = S_IN_STRING; return 1;
        case 'u': s->step = S_ESC_U; return 1;
        }
        return sc_error(s, c, "in string escape code");
    case S_ESC_U: case S_ESC_U1: case S_ESC_U12: case S_ESC_U123:
        if (is_hex(c)) { s->step = (s->step == S_ESC_U123) ? S_IN_STRING : s->step + 1; return 1; }
        return sc_error(s, c, "in \\u hexadecimal character escape");
    case S_NEG:
        if (c == '0') { s->step = S_0; return 1; }
        if (c >= '1' && c <= '9') { s->step = S_1; return 1; }
        return sc_error(s, c, "in numeric literal");
    case S_1:
        if (c >= '0' && c <= '9') return 1;
        /* fallthrough */
    case S_0:
        if (c == '.') { s->step = S_DOT; return 1; }
        if (c == 'e' || c == 'E') { s->step = S_E; return 1; }
        return sc_end_value(s, c);
    case S_DOT:
        if (c >= '0' && c <= '9') { s->step = S_DOT0; return 1; }
        return sc_error(s, c, "after decimal point in numeric literal");
    case S_DOT0:
        if (c >= '0' && c <= '9') return 1;
        if (c == 'e' || c == 'E') { s->step = S_E; return 1; }
        return sc_end_value(s, c);
    case S_E:
        if (c == '+' || c == '-') { s->step = S_ESIGN; return 1; }
        /* fallthrough */
    case S_ESIGN:
        if (c >= '0' && c <= '9') { s->step = S_E0; return 1; }
        return sc_error(s, c, "in exponent of numeric literal");
    case S_E0:
        if (c >= '0' && c <= '9') return 1;
        return sc_end_value(s, c);
    case S_T: if (c == 'r') { s->step = S_TR; return 1; } return sc_error(s, c, "in literal true (expecting 'r')");
    case S_TR: if (c == 'u') { s->step = S_TRU; return 1; } return sc_error(s, c, "in literal true (expecting 'u')");
    case S_TRU: if (c == 'e') { s->step = S_END_VALUE; return 1; } return sc_error(s, c, "in literal true (expecting 'e')");
    case S_F: if (c == 'a') { s->step = S_FA; return 1; } return sc_error(s, c, "in literal false (expecting 'a')");
    case S_FA: if (c == 'l') { s->step = S_FAL; return 1; } return sc_error(s, c, "in literal false (expecting 'l')");
    case S_FAL: if (c == 's') { s->step = S_FALS; return 1; } return sc_error(s, c, "in literal false (expecting 's')");
    case S_FALS: if (c == 'e') { s->step = S_END_VALUE; return 1; } return sc_error(s, c, "in literal false (expecting 'e')");
    case S_N: if (c == 'u') { s->step = S_NU; return 1; } return sc_error(s, c, "in literal null (expecting 'u')");
    case S_NU: if (c == 'l') { s->step = S_NUL; return 1; } return sc_error(s, c, "in literal null (expecting 'l')");
    case S_NUL: if (c == 'l') { s->step = S_END_VALUE; return 1; } return sc_error(s, c, "in literal null (expecting 'l')");
    default: return 0;
    }
}
/* checkValid: returns 1 if data is one valid JSON value; else fills err */
static int check_valid(const unsigned char *p, size_t n, char *err, size_t errcap) {
    scanner s;
    memset(&s, 0, sizeof s);
    s.step = S_BEGIN_VALUE;
    int ok = 1;
    for (size_t i = 0; i < n && ok; i++) ok = sc_step(&s, p[i]);
    if (ok) {
        /* scanner.eof() */
        if (!s.end_top) {
            sc_step(&s, ' ');
            if (s.step == S_ERROR) ok = 0;
            else if (!s.end_top) { snprintf(s.err, sizeof s.err, "unexpected end of JSON input"); ok = 0; }
        }
    }
    if (!ok && err) snprintf(err, errcap, "%s", s.err);
    free(s.stack);
    return ok;
}

/* ---------------------------------------------------- decode (Unmarshal) */

typedef struct {
    const unsigned char *p;
    size_t n, i;
    char *saved_err; /* first UnmarshalTypeError, reported after the decode completes */
    int record_mode; /* objects become ordered records */
} decoder;

static void skip_ws(decoder *d) { while (d->i < d->n && is_space(d->p[d->i])) d->i++; }

static unsigned hexval(unsigned char c) { return c <= '9' ? c - '0' : (c | 0x20) - 'a' + 10; }

/* unquote: d->i is just past the opening quote; input already validated */
static void dec_string(decoder *d, buf *out) {
    out->n = 0;
    buf_put(out, "", 0);
    const unsigned char *p = d->p;
    size_t i = d->i;
    for (;;) {
        unsigned char c = p[i];
        if (c == '"') { i++; break; }
        if (c == '\\') {
            unsigned char e = p[i + 1];
            i += 2;
            switch (e) {
            case '"': buf_putc(out, '"'); break;
            case '\\': buf_putc(out, '\\'); break;
            case '/': buf_putc(out, '/'); break;
            case 'b': buf_putc(out, '\b'); break;
            case 'f': buf_putc(out, '\f'); break;
            case 'n': buf_putc(out, '\n'); break;
            case 'r': buf_putc(out, '\r'); break;
            case 't': buf_putc(out, '\t'); break;
            case 'u': {
                uint32_t rr = (hexval(p[i]) << 12) | (hexval(p[i + 1]) << 8) | (hexval(p[i + 2]) << 4) | hexval(p[i + 3]);
                i += 4;
                if (rr >= 0xD800 && rr <= 0xDFFF) {
                    /* utf16.IsSurrogate: try to pair with a following \uXXXX */
                    int paired = 0;
                    if (rr < 0xDC00 && p[i] == '\\' && p[i + 1] == 'u' && is_hex(p[i + 2]) && is_hex(p[i + 3]) &&
                        is_hex(p[i + 4]) && is_hex(p[i + 5])) {
                        uint32_t r2 = (hexval(p[i + 2]) << 12) | (hexval(p[i + 3]) << 8) | (hexval(p[i + 4]) << 4) | hexval(p[i + 5]);
                        if (r2 >= 0xDC00 && r2 <= 0xDFFF) {
                            utf8_encode(out, 0x10000 + ((rr - 0xD800) << 10) + (r2 - 0xDC00));
                            i += 6;
                            paired = 1;
                        }
                    }
                    if (!paired) utf8_encode(out, 0xFFFD);
                } else {
                    utf8_encode(out, rr);
                }
                break;
            }
            }
            continue;
        }
        if (c < 0x80) { buf_putc(out, (char)c); i++; continue; }
        uint32_t r;
        /* the closing quote bounds the string, so decoding may look ahead safely up to n */
        int size = utf8_decode(p + i, d->n - i, &r);
        if (r == 0xFFFD && size == 1) { utf8_encode(out, 0xFFFD); i += 1; }
        else { buf_put(out, p + i, (size_t)size); i += (size_t)size; }
    }
    d->i = i;
}

static val *dec_value(decoder *d);

static val *dec_object(decoder *d) {
    val *m = v_new(d->record_mode ? V_REC : V_MAP);
    buf key = {0};
    d->i++; /* { */
    skip_ws(d);
    if (d->p[d->i] == '}') { d->i++; free(key.p); return m; }
    for (;;) {
        skip_ws(d);
        d->i++; /* opening quote */
        dec_string(d, &key);
        size_t klen = key.n;
        char *k = xmemdup(key.p, klen);
        skip_ws(d);
        d->i++; /* : */
        val *v = dec_value(d);
        if (m->k == V_MAP) map_put(m, k, klen, v); /* last duplicate wins */
        else rec_push(m, k, klen, v);
        free(k);
        skip_ws(d);
        unsigned char c = d->p[d->i++];
        if (c == '}') break;
    }
    free(key.p);
    return m;
}
static val *dec_array(decoder *d) {
    val *a = v_new(V_ARR);
    d->i++;
    skip_ws(d);
    if (d->p[d->i] == ']') { d->i++; return a; }
    for (;;) {
        arr_push(a, dec_value(d));
        skip_ws(d);
        unsigned char c = d->p[d->i++];
        if (c == ']') break;
    }
    return a;
}
/* returns NULL for JSON null (Go nil interface) */
static val *dec_value(decoder *d) {
    skip_ws(d);
    unsigned char c = d->p[d->i];
    if (c == '{') return dec_object(d);
    if (c == '[') return dec_array(d);
    if (c == '"') {
        buf s = {0};
        d->i++;
        dec_string(d, &s);
        val *v = v_str(s.p, s.n);
        free(s.p);
        return v;
    }
    if (c == 't') { d->i += 4; val *v = v_new(V_BOOL); v->b = 1; return v; }
    if (c == 'f') { d->i += 5; val *v = v_new(V_BOOL); v->b = 0; return v; }
    if (c == 'n') { d->i += 4; return NULL; }
    /* number: convertNumber -> strconv.ParseFloat(s, 64) */
    size_t st = d->i;
    while (d->i < d->n) {
        unsigned char x = d->p[d->i];
        if ((x >= '0' && x <= '9') || x == '-' || x == '+' || x == '.' || x == 'e' || x == 'E') d->i++;
        else break;
    }
    char *txt = xmemdup((const char *)d->p + st, d->i - st);
    double f = strtod(txt, NULL);
    val *v = v_new(V_NUM);
    if (isinf(f)) {
        if (!d->saved_err) {
            size_t cap = strlen(txt) + 80;
            d->saved_err = xmalloc(cap);
            snprintf(d->saved_err, cap, "json: cannot unmarshal number %s into Go value of type float64", txt);
        }
        f = 0;
    }
    v->num = f;
    free(txt);
    return v;
}

/* json.Unmarshal(data, &interface{}).  Returns 0 and fills err on failure. */
static int unmarshal(const char *raw, size_t len, int record_mode, val **out, char *err, size_t errcap) {
    if (!check_valid((const unsigned char *)raw, len, err, errcap)) return 0;
    decoder d = { (const unsigned char *)raw, len, 0, NULL, record_mode };
    val *v = dec_value(&d);
    if (d.saved_err) {
        snprintf(err, errcap, "%s", d.saved_err);
        free(d.saved_err);
        v_free(v);
        return 0;
    }
    *out = v;
    return 1;
}

/* ------------------------------------------------------- encode (Marshal) */

static const char HEX[] = "0123456789abcdef";

/* encodeState.string with escapeHTML = true */
static void enc_string(buf *b, const char *s, size_t n) {
    buf_putc(b, '"');
    size_t start = 0, i = 0;
    const unsigned char *p = (const unsigned char *)s;
    while (i < n) {
        unsigned char c = p[i];
        if (c < 0x80) {
            int safe = c >= 0x20 && c != '"' && c != '\\' && c != '<' && c != '>' && c != '&';
            if (safe) { i++; continue; }
            if (start < i) buf_put(b, s + start, i - start);
            switch (c) {
            case '\\': case '"': buf_putc(b, '\\'); buf_putc(b, (char)c); break;
            case '\n': buf_put(b, "\\n", 2); break;
            case '\r': buf_put(b, "\\r", 2); break;
            case '\t': buf_put(b, "\\t", 2); break;
            default:
                buf_put(b, "\\u00", 4);
                buf_putc(b, HEX[c >> 4]);
                buf_putc(b, HEX[c & 0xF]);
            }
            i++;
            start = i;
            continue;
        }
        uint32_t r;
        int size = utf8_decode(p + i, n - i, &r);
        if (r == 0xFFFD && size == 1) {
            if (start < i) buf_put(b, s + start, i - start);
            buf_put(b, "\\ufffd", 6);
            i += 1;
            start = i;
            continue;
        }
        if (r == 0x2028 || r == 0x2029) {
            if (start < i) buf_put(b, s + start, i - start);
            buf_put(b, "\\u202", 5);
            buf_putc(b, HEX[r & 0xF]);
            i += (size_t)size;
            start = i;
            continue;
        }
        i += (size_t)size;
    }
    if (start < n) buf_put(b, s + start, n - start);
    buf_putc(b, '"');
}

/* floatEncoder (64 bit): ES6-style; shortest digits that round-trip */
static void enc_float(buf *b, double f) {
    if (f == 0) { buf_puts(b, signbit(f) ? "-0" : "0"); return; }
    char t[64];
    int prec;
    for (prec = 0; prec < 17; prec++) {
        snprintf(t, sizeof t, "%.*e", prec, f);
        if (strtod(t, NULL) == f) break;
    }
    /* t = [-]d[.ddd]e[+-]XX */
    char digits[32];
    int nd = 0, neg = 0;
    const char *q = t;
    if (*q == '-') { neg = 1; q++; }
    for (; *q && *q != 'e'; q++) if (*q != '.') digits[nd++] = *q;
    int e10 = atoi(q + 1);
    while (nd > 1 && digits[nd - 1] == '0') nd--;
    double a = fabs(f);
    if (neg) buf_putc(b, '-');
    if (a < 1e-6 || a >= 1e21) {
        buf_putc(b, digits[0]);
        if (nd > 1) { buf_putc(b, '.'); buf_put(b, digits + 1, (size_t)nd - 1); }
        buf_putc(b, 'e');
        buf_putc(b, e10 < 0 ? '-' : '+');
        int ae = e10 < 0 ? -e10 : e10;
        char et[16];
        /* strconv pads to two digits; encoding/json then trims "e-0X" to "e-X" */
        if (ae < 10 && e10 >= 0) snprintf(et, sizeof et, "0%d", ae);
        else snprintf(et, sizeof et, "%d", ae);
        buf_puts(b, et);
        return;
    }
    if (e10 >= 0) {
        for (int i = 0; i <= e10; i++) buf_putc(b, i < nd ? digits[i] : '0');
        if (nd > e10 + 1) { buf_putc(b, '.'); buf_put(b, digits + e10 + 1, (size_t)(nd - e10 - 1)); }
    } else {
        buf_puts(b, "0.");
        for (int i = 0; i < -e10 - 1; i++) buf_putc(b, '0');
        buf_put(b, digits, (size_t)nd);
    }
}

static int cmp_entry(const void *a, const void *b) {
    const entry *x = *(const entry *const *)a, *y = *(const entry *const *)b;
    size_t m = x->klen < y->klen ? x->klen : y->klen;
    int c = memcmp(x->key, y->key, m);
    if (c) return c;
    return (x->klen > y->klen) - (x->klen < y->klen);
}
static const entry **sorted_entries(const val *m) {
    const entry **v = xmalloc((m->n + 1) * sizeof *v);
    for (size_t i = 0; i < m->n; i++) v[i] = &m->ent[i];
    qsort(v, m->n, sizeof *v, cmp_entry);
    return v;
}

/* json.Indent's newline(): "\n" + prefix + depth * indent; fmt == NULL means compact Marshal */
typedef struct { const char *prefix, *indent; } indent_fmt;
static void newline(buf *b, const indent_fmt *indent, int depth) {
    if (!indent) return;
    buf_putc(b, '\n');
    buf_puts(b, indent->prefix);
    for (int i = 0; i < depth; i++) buf_puts(b, indent->indent);
}
/* Marshal followed by Indent(prefix "", indent "\t") when indent != 0, else compact Marshal */
static void enc_value(buf *b, const val *v, const indent_fmt *indent, int depth) {
    if (!v) { buf_puts(b, "null"); return; }
    switch (v->k) {
    case V_NULL: buf_puts(b, "null"); break;
    case V_BOOL: buf_puts(b, v->b ? "true" : "false"); break;
    case V_NUM: enc_float(b, v->num); break;
    case V_STR: enc_string(b, v->s, v->slen); break;
    case V_ARR:
        if (v->n == 0) { buf_puts(b, "[]"); break; }
        buf_putc(b, '[');
        for (size_t i = 0; i < v->n; i++) {
            if (i) buf_putc(b, ',');
            newline(b, indent, depth + 1);
            enc_value(b, v->items[i], indent, depth + 1);
        }
        newline(b, indent, depth);
        buf_putc(b, ']');
        break;
    case V_MAP:
    case V_REC: {
        if (v->n == 0) { buf_puts(b, "{}"); break; }
        const entry **ord = NULL;
        if (v->k == V_MAP) ord = sorted_entries(v);
        buf_putc(b, '{');
        for (size_t i = 0; i < v->n; i++) {
            const entry *e = ord ? ord[i] : &v->ent[i];
            if (i) buf_putc(b, ',');
            newline(b, indent, depth + 1);
            enc_string(b, e->key, e->klen);
            buf_putc(b, ':');
            if (indent) buf_putc(b, ' ');
            enc_value(b, e->v, indent, depth + 1);
        }
        newline(b, indent, depth);
        buf_putc(b, '}');
        free(ord);
        break;
    }
    }
}

/* ----------------------------------------------------------- gabs paths */

typedef struct { const char *p; size_t n; } part;
/* strings.Split(path, ".") */
static part *split(const char *s, size_t n, char sep, size_t *count) {
    size_t c = 1;
    for (size_t i = 0; i < n; i++) if (s[i] == sep) c++;
    part *v = xmalloc(c * sizeof *v);
    size_t k = 0, st = 0;
    for (size_t i = 0; i <= n; i++)
        if (i == n || s[i] == sep) { v[k].p = s + st; v[k].n = i - st; k++; st = i + 1; }
    *count = c;
    return v;
}

/* temp arrays created by Search's array fan-out; items are borrowed */
typedef struct tmp { val *v; struct tmp *next; } tmp;
static val *search(val *obj, const part *h, size_t nh, tmp **tmps) {
    for (size_t t = 0; t < nh; t++) {
        if (obj && obj->k == V_MAP) {
            obj = map_get(obj, h[t].p, h[t].n, NULL);
        } else if (obj && obj->k == V_ARR) {
            val *a = v_new(V_ARR);
            tmp *x = xmalloc(sizeof *x);
            x->v = a; x->next = *tmps; *tmps = x;
            for (size_t i = 0; i < obj->n; i++) {
                val *r = search(obj->items[i], h + t, nh - t, tmps);
                if (r) arr_push(a, r);
            }
            if (a->n == 0) return NULL;
            return a;
        } else {
            return NULL; /* strings, numbers, bools, nil and records (Go structs) */
        }
    }
    return obj;
}
static void free_tmps(tmp *t) {
    while (t) { tmp *n = t->next; free(t->v->items); free(t->v); free(t); t = n; }
}

#define ERR_COLLISION "encountered value collision whilst building path"
#define ERR_NOT_OBJ "not an object"

/* gabs Set: value ownership passes to the tree on success; on failure the caller frees it */
static int gabs_set(doc *d, val *value, const part *path, size_t np, char *err, size_t errcap) {
    if (np == 0) { v_free(d->root); d->root = value; return 1; }
    if (!d->root) d->root = v_new(V_MAP);
    val *object = d->root;
    for (size_t t = 0; t < np; t++) {
        if (object && object->k == V_MAP) {
            if (t == np - 1) {
                map_put(object, path[t].p, path[t].n, value);
            } else if (map_get(object, path[t].p, path[t].n, NULL) == NULL) {
                map_put(object, path[t].p, path[t].n, v_new(V_MAP));
            }
            object = map_get(object, path[t].p, path[t].n, NULL);
        } else {
            snprintf(err, errcap, "%s", ERR_COLLISION);
            return 0;
        }
    }
    return 1;
}
static int gabs_delete(doc *d, const part *path, size_t np, char *err, size_t errcap) {
    if (!d->root) { snprintf(err, errcap, "%s", ERR_NOT_OBJ); return 0; }
    val *object = d->root;
    for (size_t t = 0; t < np; t++) {
        if (object && object->k == V_MAP) {
            if (t == np - 1) {
                map_del(object, path[t].p, path[t].n);
            } else if (map_get(object, path[t].p, path[t].n, NULL) == NULL) {
                snprintf(err, errcap, "%s", ERR_NOT_OBJ);
                return 0;
            }
            object = map_get(object, path[t].p, path[t].n, NULL);
        } else {
            snprintf(err, errcap, "%s", ERR_NOT_OBJ);
            return 0;
        }
    }
    return 1;
}

/* ------------------------------------------------------------ public API */

#define API __attribute__((visibility("default")))

API void *ora_new(const char *raw, size_t len, char *err, size_t errcap) {
    val *root = NULL;
    if (!unmarshal(raw, len, 0, &root, err, errcap)) return NULL;
    doc *d = xmalloc(sizeof *d);
    d->root = root;
    return d;
}
API void ora_free(void *h) {
    doc *d = h;
    if (!d) return;
    v_free(d->root);
    free(d);
}
API void ora_free_buf(char *p) { free(p); }

/* state.Get: the string at path or "" (state/state.go:27-34).  *out borrows from the tree. */
API int ora_get(void *h, const char *path, size_t plen, const char **out, size_t *outlen) {
    doc *d = h;
    size_t np;
    part *ps = split(path, plen, '.', &np);
    tmp *tmps = NULL;
    val *r = search(d->root, ps, np, &tmps);
    if (r && r->k == V_STR) { *out = r->s; *outlen = r->slen; }
    else { *out = ""; *outlen = 0; }
    free_tmps(tmps);
    free(ps);
    return 0;
}

/* SetP(value, path): value is JSON text; is_record keeps member order and makes it opaque */
API int ora_set(void *h, const char *path, size_t plen, const char *json, size_t jlen, int is_record, char *err, size_t errcap) {
    doc *d = h;
    val *v = NULL;
    if (!unmarshal(json, jlen, is_record, &v, err, errcap)) return 1;
    size_t np;
    part *ps = split(path, plen, '.', &np);
    int ok = gabs_set(d, v, ps, np, err, errcap);
    free(ps);
    if (!ok) { v_free(v); return 1; }
    return 0;
}

/* getClusterKeyParts (state/state.go:149-160) */
static int cluster_key_parts(const char *ck, size_t cklen, part *provider, part *name, char *err, size_t errcap) {
    size_t np;
    part *ps = split(ck, cklen, '_', &np);
    if (np < 3) {
        snprintf(err, errcap, "Could not get cluster key parts, cluster does not follow format `cluster_{provider}_{clusterName}` '%.*s'", (int)cklen, ck);
        free(ps);
        return 0;
    }
    *provider = ps[1];
    *name = ps[2];
    free(ps);
    return 1;
}

API int ora_add_cluster(void *h, const char *provider, size_t prlen, const char *name, size_t nlen,
                        const char *json, size_t jlen, int is_record, char *err, size_t errcap) {
    buf p = {0};
    buf_puts(&p, "module.cluster_");
    buf_put(&p, provider, prlen);
    buf_putc(&p, '_');
    buf_put(&p, name, nlen);
    int rc = ora_set(h, p.p, p.n, json, jlen, is_record, err, errcap);
    free(p.p);
    return rc;
}
API int ora_add_node(void *h, const char *ck, size_t cklen, const char *name, size_t nlen,
                     const char *json, size_t jlen, int is_record, char *err, size_t errcap) {
    part provider, cname;
    if (!cluster_key_parts(ck, cklen, &provider, &cname, err, errcap)) return 1;
    buf p = {0};
    buf_puts(&p, "module.node_");
    buf_put(&p, provider.p, provider.n);
    buf_putc(&p, '_');
    buf_put(&p, cname.p, cname.n);
    buf_putc(&p, '_');
    buf_put(&p, name, nlen);
    int rc = ora_set(h, p.p, p.n, json, jlen, is_record, err, errcap);
    free(p.p);
    return rc;
}
API int ora_delete(void *h, const char *path, size_t plen, char *err, size_t errcap) {
    doc *d = h;
    size_t np;
    part *ps = split(path, plen, '.', &np);
    int ok = gabs_delete(d, ps, np, err, errcap);
    free(ps);
    return ok ? 0 : 1;
}

/* state.Bytes(): BytesIndent("", "\t"); a nil root gives "{}" (gabs.go:395-402) */
API int ora_bytes_indent(void *h, const char *prefix, const char *indent, char **out, size_t *outlen) {
    doc *d = h;
    buf b = {0};
    indent_fmt f = { prefix ? prefix : "", indent };
    if (!d->root) buf_puts(&b, "{}");
    else enc_value(&b, d->root, indent ? &f : NULL, 0);
    *out = b.p;
    *outlen = b.n;
    return 0;
}
API int ora_bytes(void *h, int indent, char **out, size_t *outlen) {
    return ora_bytes_indent(h, "", indent ? "\t" : NULL, out, outlen);
}

/* Container.Path(path).String(): compact Marshal of the Search result, "{}" for nil.
 * Exists only to replay the vendored gabs_test.go vectors. */
API int ora_path_string(void *h, const char *path, size_t plen, char **out, size_t *outlen) {
    doc *d = h;
    size_t np = 0;
    part *ps = NULL;
    tmp *tmps = NULL;
    val *r = d->root;
    if (path) { ps = split(path, plen, '.', &np); r = search(d->root, ps, np, &tmps); }
    buf b = {0};
    if (!r) buf_puts(&b, "{}");
    else enc_value(&b, r, NULL, 0);
    free_tmps(tmps);
    free(ps);
    *out = b.p;
    *outlen = b.n;
    return 0;
}

typedef struct { const char *name; size_t nlen; const char *key; size_t klen; } pair;
static int cmp_pair(const void *a, const void *b) {
    const pair *x = a, *y = b;
    size_t m = x->nlen < y->nlen ? x->nlen : y->nlen;
    int c = memcmp(x->name, y->name, m);
    if (c) return c;
    if (x->nlen != y->nlen) return (x->nlen > y->nlen) - (x->nlen < y->nlen);
    m = x->klen < y->klen ? x->klen : y->klen;
    c = memcmp(x->key, y->key, m);
    if (c) return c;
    return (x->klen > y->klen) - (x->klen < y->klen);
}
/* shared body of Clusters()/Nodes(): children of "module" whose key has the prefix, mapped
 * field -> key.  Output: repeated { u32 nlen, name, u32 klen, key }, sorted by name. */
static int children_by_prefix(doc *d, const char *prefix, size_t plen, const char *field,
                              char **out, size_t *outlen, size_t *count, char *err, size_t errcap) {
    val *module = NULL;
    if (d->root && d->root->k == V_MAP) module = map_get(d->root, "module", 6, NULL);
    else if (d->root && d->root->k == V_ARR) {
        /* S("module") on an array root fans out; the result is an array or nil: never a map */
        module = NULL;
    }
    if (!module || module->k != V_MAP) { snprintf(err, errcap, "%s", ERR_NOT_OBJ); return 1; }
    pair *ps = xmalloc((module->n + 1) * sizeof *ps);
    size_t np = 0;
    size_t flen = strlen(field);
    for (size_t i = 0; i < module->n; i++) {
        const entry *e = &module->ent[i];
        if (e->klen < plen || memcmp(e->key, prefix, plen) != 0) continue;
        /* child.Path(field).Data().(string): child must be a map whose field is a string */
        val *c = e->v;
        if (!c || c->k != V_MAP) continue;
        val *f = map_get(c, field, flen, NULL);
        if (!f || f->k != V_STR) continue;
        ps[np].name = f->s; ps[np].nlen = f->slen; ps[np].key = e->key; ps[np].klen = e->klen;
        np++;
    }
    qsort(ps, np, sizeof *ps, cmp_pair);
    buf b = {0};
    size_t kept = 0;
    for (size_t i = 0; i < np; i++) {
        /* duplicate names: the greatest key (last in this order) wins */
        if (i + 1 < np && ps[i + 1].nlen == ps[i].nlen && memcmp(ps[i + 1].name, ps[i].name, ps[i].nlen) == 0) continue;
        uint32_t l = (uint32_t)ps[i].nlen;
        buf_put(&b, &l, 4);
        buf_put(&b, ps[i].name, ps[i].nlen);
        l = (uint32_t)ps[i].klen;
        buf_put(&b, &l, 4);
        buf_put(&b, ps[i].key, ps[i].klen);
        kept++;
    }
    free(ps);
    if (!b.p) buf_put(&b, "", 0);
    *out = b.p;
    *outlen = b.n;
    *count = kept;
    return 0;
}
API int ora_clusters(void *h, char **out, size_t *outlen, size_t *count, char *err, size_t errcap) {
    return children_by_prefix(h, "cluster_", 8, "name", out, outlen, count, err, errcap);
}
API int ora_nodes(void *h, const char *ck, size_t cklen, char **out, size_t *outlen, size_t *count, char *err, size_t errcap) {
    part provider, cname;
    if (!cluster_key_parts(ck, cklen, &provider, &cname, err, errcap)) return 1;
    buf p = {0};
    buf_puts(&p, "node_");
    buf_put(&p, provider.p, provider.n);
    buf_putc(&p, '_');
    buf_put(&p, cname.p, cname.n);
    buf_putc(&p, '_');
    int rc = children_by_prefix(h, p.p, p.n, "hostname", out, outlen, count, err, errcap);
    free(p.p);
    return rc;
}

/* strconv.Atoi: optional sign, decimal digits only, int64 range */
static int go_atoi(const char *s, size_t n, long long *out) {
    if (n == 0) return 0;
    size_t i = 0;
    int neg = 0;
    if (s[0] == '+' || s[0] == '-') { neg = s[0] == '-'; i = 1; if (n == 1) return 0; }
    unsigned long long v = 0;
    for (; i < n; i++) {
        if (s[i] < '0' || s[i] > '9') return 0;
        unsigned d = (unsigned)(s[i] - '0');
        if (v > (0xffffffffffffffffull - d) / 10) return 0;
        v = v * 10 + d;
    }
    if (!neg && v > 9223372036854775807ull) return 0;
    if (neg && v > 9223372036854775808ull) return 0;
    *out = neg ? (long long)(0 - v) : (long long)v;
    return 1;
}
/* getNewHostnames (create/node.go:350-380).  existing: n names given as (ptr,len) pairs.
 * Output: the new names joined by '\n' (names never contain one here). */
API int ora_new_hostnames(const char *const *existing, const size_t *elens, size_t n, const char *node_name,
                          size_t nnlen, long long to_add, char **out, size_t *outlen) {
    buf b = {0};
    buf_put(&b, "", 0);
    if (to_add >= 1) {
        long long start = 1;
        for (size_t i = 0; i < n; i++) {
            if (elens[i] < nnlen + 1 || memcmp(existing[i], node_name, nnlen) != 0 || existing[i][nnlen] != '-') continue;
            long long num;
            if (!go_atoi(existing[i] + nnlen + 1, elens[i] - nnlen - 1, &num)) continue;
            if (num >= start) start = num + 1;
        }
        for (long long i = 0; i < to_add; i++) {
            char t[32];
            if (i) buf_putc(&b, '\n');
            buf_put(&b, node_name, nnlen);
            snprintf(t, sizeof t, "-%lld", start + i);
            buf_puts(&b, t);
        }
    }
    *out = b.p;
    *outlen = b.n;
    return 0;
}

/* json.Valid-style helper for tests: 1 if raw parses, else 0 with the Go error text */
API int ora_check(const char *raw, size_t len, char *err, size_t errcap) {
    val *v = NULL;
    if (!unmarshal(raw, len, 0, &v, err, errcap)) return 0;
    v_free(v);
    return 1;
}
