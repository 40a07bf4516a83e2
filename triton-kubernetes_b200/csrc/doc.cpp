// doc.cpp -- host tree, gabs path operations, host value parser (see doc.hpp).
#include "doc.hpp"

#include <charconv>
#include <clocale>
#include <locale.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace tk {

// strtod in the "C" locale whatever the host program's LC_NUMERIC is (a decimal-comma locale would
// stop at the '.'): the library is loaded into other people's processes.
double strtod_c(const char *s) {
    static const locale_t c_locale = newlocale(LC_ALL_MASK, "C", (locale_t)0);
    return c_locale ? strtod_l(s, nullptr, c_locale) : strtod(s, nullptr);
}


const char *const kErrCollision = "encountered value collision whilst building path";
const char *const kErrNotObj = "not an object";

void free_node(Node *n) {
    if (!n) return;
    switch (n->kind) {
    case Kind::Bool: delete static_cast<BoolNode *>(n); break;
    case Kind::Num: delete static_cast<NumNode *>(n); break;
    case Kind::Str: delete static_cast<StrNode *>(n); break;
    case Kind::Arr: {
        auto *a = static_cast<ArrNode *>(n);
        for (Node *c : a->items) free_node(c);
        delete a;
        break;
    }
    case Kind::Map: {
        auto *m = static_cast<MapNode *>(n);
        for (auto &kv : m->m) free_node(kv.second);
        delete m;
        break;
    }
    case Kind::Rec: {
        auto *r = static_cast<RecNode *>(n);
        for (auto &kv : r->m) free_node(kv.second);
        delete r;
        break;
    }
    }
}

std::vector<std::string_view> split(std::string_view s, char sep) {
    std::vector<std::string_view> out;
    size_t st = 0;
    for (size_t i = 0; i <= s.size(); i++) {
        if (i == s.size() || s[i] == sep) {
            out.emplace_back(s.data() + st, i - st);
            st = i + 1;
        }
    }
    return out;
}

// ------------------------------------------------------------------ UTF-8 helpers

// unicode/utf8.DecodeRune: length of the valid sequence at p, or 0 if p[0] starts none
static int utf8_valid_len(const unsigned char *p, size_t n) {
    unsigned c = p[0];
    unsigned lo = 0x80, hi = 0xBF;
    int len;
    if (c < 0x80) return 1;
    if (c >= 0xC2 && c <= 0xDF) len = 2;
    else if (c == 0xE0) { len = 3; lo = 0xA0; }
    else if (c >= 0xE1 && c <= 0xEF) { len = 3; if (c == 0xED) hi = 0x9F; }
    else if (c == 0xF0) { len = 4; lo = 0x90; }
    else if (c >= 0xF1 && c <= 0xF3) len = 4;
    else if (c == 0xF4) { len = 4; hi = 0x8F; }
    else return 0;
    if ((size_t)len > n) return 0;
    if (p[1] < lo || p[1] > hi) return 0;
    for (int i = 2; i < len; i++)
        if ((p[i] & 0xC0) != 0x80) return 0;
    return len;
}

static void append_rune(std::string *s, uint32_t r) {
    if (r > 0x10FFFF || (r >= 0xD800 && r <= 0xDFFF)) r = 0xFFFD;
    if (r < 0x80) {
        s->push_back((char)r);
    } else if (r < 0x800) {
        s->push_back((char)(0xC0 | (r >> 6)));
        s->push_back((char)(0x80 | (r & 0x3F)));
    } else if (r < 0x10000) {
        s->push_back((char)(0xE0 | (r >> 12)));
        s->push_back((char)(0x80 | ((r >> 6) & 0x3F)));
        s->push_back((char)(0x80 | (r & 0x3F)));
    } else {
        s->push_back((char)(0xF0 | (r >> 18)));
        s->push_back((char)(0x80 | ((r >> 12) & 0x3F)));
        s->push_back((char)(0x80 | ((r >> 6) & 0x3F)));
        s->push_back((char)(0x80 | (r & 0x3F)));
    }
}

// ------------------------------------------------------------------ value parser

namespace {

// scanner.go quoteChar
std::string quote_char(unsigned char c) {
    char t[16];
    switch (c) {
    case '\'': return "'\\''";
    case '"': return "'\"'";
    case '\\': return "'\\\\'";
    case '\a': return "'\\a'";
    case '\b': return "'\\b'";
    case '\f': return "'\\f'";
    case '\n': return "'\\n'";
    case '\r': return "'\\r'";
    case '\t': return "'\\t'";
    case '\v': return "'\\v'";
    }
    if (c < 0x20 || c == 0x7f) { snprintf(t, sizeof t, "'\\x%02x'", c); return t; }
    if (c < 0x80) { snprintf(t, sizeof t, "'%c'", c); return t; }
    if (c <= 0xA0 || c == 0xAD) { snprintf(t, sizeof t, "'\\u00%02x'", c); return t; }  // not printable
    std::string s = "'";
    s.push_back((char)(0xC0 | (c >> 6)));
    s.push_back((char)(0x80 | (c & 0x3F)));
    s.push_back('\'');
    return s;
}

inline bool is_ws(unsigned char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }
inline bool is_digit(unsigned char c) { return c >= '0' && c <= '9'; }
inline bool is_hex(unsigned char c) { return is_digit(c) || ((c | 0x20) >= 'a' && (c | 0x20) <= 'f'); }
inline unsigned hexv(unsigned char c) { return c <= '9' ? c - '0' : (c | 0x20) - 'a' + 10; }

// Recursive descent over the same grammar as Go's scanner, phrased so that the FIRST error it
// meets carries Go's message for it.  End of input behaves as Go's scanner.eof(): one virtual
// ' ' is fed; states that would skip it report "unexpected end of JSON input".
class Parser {
public:
    Parser(std::string_view t, bool rec, bool build) : p_((const unsigned char *)t.data()), n_(t.size()), rec_(rec), build_(build) {}

    bool run(Node **out, std::string *err) {
        Node *v = nullptr;
        bool ok = value(&v);
        if (ok) {
            ws();
            if (i_ < n_) ok = fail(p_[i_], "after top-level value");
        }
        if (ok && !range_err_.empty()) { err_ = range_err_; ok = false; }
        if (!ok) {
            free_node(v);
            if (err) *err = err_;
            return false;
        }
        if (out) *out = v; else free_node(v);
        return true;
    }

private:
    const unsigned char *p_;
    size_t n_, i_ = 0;
    bool rec_, build_;
    int depth_ = 0;
    std::string err_, range_err_;

    void ws() { while (i_ < n_ && is_ws(p_[i_])) i_++; }
    bool fail(unsigned char c, const char *ctx) {
        err_ = "invalid character " + quote_char(c) + " " + ctx;
        return false;
    }
    bool eof_fail() { err_ = "unexpected end of JSON input"; return false; }
    // error at position i_: a real byte, or the virtual ' ' at end of input
    bool fail_here(const char *ctx) { return fail(i_ < n_ ? p_[i_] : ' ', ctx); }

    bool literal(const char *word, const char *name) {
        // first byte already matched
        static const char *const fmt = "in literal %s (expecting '%c')";
        for (size_t k = 1; word[k]; k++) {
            i_++;
            if (i_ >= n_ || p_[i_] != (unsigned char)word[k]) {
                char ctx[64];
                snprintf(ctx, sizeof ctx, fmt, name, word[k]);
                return fail_here(ctx);
            }
        }
        i_++;
        return true;
    }

    bool number(Node **out) {
        size_t st = i_;
        if (p_[i_] == '-') {
            i_++;
            if (i_ >= n_ || !is_digit(p_[i_])) return fail_here("in numeric literal");
        }
        if (p_[i_] == '0') i_++;
        else while (i_ < n_ && is_digit(p_[i_])) i_++;
        if (i_ < n_ && p_[i_] == '.') {
            i_++;
            if (i_ >= n_ || !is_digit(p_[i_])) return fail_here("after decimal point in numeric literal");
            while (i_ < n_ && is_digit(p_[i_])) i_++;
        }
        if (i_ < n_ && (p_[i_] == 'e' || p_[i_] == 'E')) {
            i_++;
            if (i_ < n_ && (p_[i_] == '+' || p_[i_] == '-')) i_++;
            if (i_ >= n_ || !is_digit(p_[i_])) return fail_here("in exponent of numeric literal");
            while (i_ < n_ && is_digit(p_[i_])) i_++;
        }
        std::string txt((const char *)p_ + st, i_ - st);
        double v = strtod_c(txt.c_str());
        if (std::isinf(v)) {
            // strconv.ParseFloat range error: reported after the whole input proved valid
            if (range_err_.empty()) range_err_ = "json: cannot unmarshal number " + txt + " into Go value of type float64";
            v = 0;
        }
        if (build_) *out = new NumNode(v);
        return true;
    }

    // i_ is just past the opening quote
    bool string(std::string *s) {
        for (;;) {
            if (i_ >= n_) return eof_fail();
            unsigned char c = p_[i_];
            if (c == '"') { i_++; return true; }
            if (c < 0x20) return fail(c, "in string literal");
            if (c == '\\') {
                i_++;
                if (i_ >= n_) return fail(' ', "in string escape code");
                unsigned char e = p_[i_];
                switch (e) {
                case '"': case '\\': case '/': s->push_back((char)e); i_++; break;
                case 'b': s->push_back('\b'); i_++; break;
                case 'f': s->push_back('\f'); i_++; break;
                case 'n': s->push_back('\n'); i_++; break;
                case 'r': s->push_back('\r'); i_++; break;
                case 't': s->push_back('\t'); i_++; break;
                case 'u': {
                    i_++;
                    uint32_t r = 0;
                    for (int k = 0; k < 4; k++, i_++) {
                        if (i_ >= n_ || !is_hex(p_[i_])) return fail_here("in \\u hexadecimal character escape");
                        r = (r << 4) | hexv(p_[i_]);
                    }
                    if (r >= 0xD800 && r <= 0xDFFF) {
                        uint32_t r2 = 0;
                        bool pair = r < 0xDC00 && i_ + 6 <= n_ && p_[i_] == '\\' && p_[i_ + 1] == 'u' && is_hex(p_[i_ + 2]) &&
                                    is_hex(p_[i_ + 3]) && is_hex(p_[i_ + 4]) && is_hex(p_[i_ + 5]);
                        if (pair) {
                            r2 = (hexv(p_[i_ + 2]) << 12) | (hexv(p_[i_ + 3]) << 8) | (hexv(p_[i_ + 4]) << 4) | hexv(p_[i_ + 5]);
                            pair = r2 >= 0xDC00 && r2 <= 0xDFFF;
                        }
                        if (pair) { append_rune(s, 0x10000 + ((r - 0xD800) << 10) + (r2 - 0xDC00)); i_ += 6; }
                        else append_rune(s, 0xFFFD);
                    } else {
                        append_rune(s, r);
                    }
                    break;
                }
                default: return fail(e, "in string escape code");
                }
                continue;
            }
            if (c < 0x80) { s->push_back((char)c); i_++; continue; }
            int len = utf8_valid_len(p_ + i_, n_ - i_);
            if (len == 0) { s->append("\xEF\xBF\xBD"); i_++; }
            else { s->append((const char *)p_ + i_, (size_t)len); i_ += (size_t)len; }
        }
    }

    bool object(Node **out) {
        if (depth_ >= 10000) { err_ = "exceeded max depth"; return false; }
        depth_++;
        MapNode *m = nullptr;
        RecNode *r = nullptr;
        if (build_) { if (rec_) { r = new RecNode(); *out = r; } else { m = new MapNode(); *out = m; } }
        i_++;  // {
        ws();
        if (i_ >= n_) return eof_fail();
        if (p_[i_] == '}') { i_++; depth_--; return true; }
        for (;;) {
            ws();
            if (i_ >= n_) return eof_fail();
            if (p_[i_] != '"') return fail(p_[i_], "looking for beginning of object key string");
            i_++;
            std::string key;
            if (!string(&key)) return false;
            ws();
            if (i_ >= n_) return eof_fail();
            if (p_[i_] != ':') return fail(p_[i_], "after object key");
            i_++;
            Node *v = nullptr;
            bool ok = value(&v);
            if (build_) {
                if (m) {  // last duplicate wins
                    auto it = m->m.find(key);
                    if (it != m->m.end()) { free_node(it->second); it->second = v; }
                    else m->m.emplace(std::move(key), v);
                } else {
                    r->m.emplace_back(std::move(key), v);
                }
            }
            if (!ok) return false;
            ws();
            if (i_ >= n_) return eof_fail();
            unsigned char c = p_[i_];
            if (c == ',') { i_++; continue; }
            if (c == '}') { i_++; depth_--; return true; }
            return fail(c, "after object key:value pair");
        }
    }

    bool array(Node **out) {
        if (depth_ >= 10000) { err_ = "exceeded max depth"; return false; }
        depth_++;
        ArrNode *a = nullptr;
        if (build_) { a = new ArrNode(); *out = a; }
        i_++;  // [
        ws();
        if (i_ >= n_) return eof_fail();
        if (p_[i_] == ']') { i_++; depth_--; return true; }
        for (;;) {
            Node *v = nullptr;
            bool ok = value(&v);
            if (build_) a->items.push_back(v);
            if (!ok) return false;
            ws();
            if (i_ >= n_) return eof_fail();
            unsigned char c = p_[i_];
            if (c == ',') { i_++; continue; }
            if (c == ']') { i_++; depth_--; return true; }
            return fail(c, "after array element");
        }
    }

    bool value(Node **out) {
        ws();
        if (i_ >= n_) return eof_fail();
        unsigned char c = p_[i_];
        switch (c) {
        case '{': return object(out);
        case '[': return array(out);
        case '"': {
            i_++;
            std::string s;
            if (!string(&s)) return false;
            if (build_) *out = new StrNode(std::move(s));
            return true;
        }
        case 't':
            if (!literal("true", "true")) return false;
            if (build_) *out = new BoolNode(true);
            return true;
        case 'f':
            if (!literal("false", "false")) return false;
            if (build_) *out = new BoolNode(false);
            return true;
        case 'n':
            return literal("null", "null");  // nil
        default:
            if (c == '-' || is_digit(c)) return number(out);
            return fail(c, "looking for beginning of value");
        }
    }
};

}  // namespace

bool parse_value(std::string_view text, bool record_mode, Node **out, std::string *err) {
    Parser p(text, record_mode, true);
    return p.run(out, err);
}

bool syntax_error_text(std::string_view text, std::string *err) {
    Parser p(text, false, false);
    return !p.run(nullptr, err);
}

// ------------------------------------------------------------------ float formatting

void format_go_float(double v, std::string *out) {
    if (v == 0) { out->append(std::signbit(v) ? "-0" : "0"); return; }
    char t[64];
    auto res = std::to_chars(t, t + sizeof t, v, std::chars_format::scientific);  // shortest round trip
    std::string_view sv(t, (size_t)(res.ptr - t));
    bool neg = sv[0] == '-';
    if (neg) sv.remove_prefix(1);
    size_t epos = sv.find('e');
    std::string digits;
    for (char c : sv.substr(0, epos)) if (c != '.') digits.push_back(c);
    int e10 = atoi(std::string(sv.substr(epos + 1)).c_str());
    while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    int nd = (int)digits.size();
    double a = std::fabs(v);
    if (neg) out->push_back('-');
    if (a < 1e-6 || a >= 1e21) {
        out->push_back(digits[0]);
        if (nd > 1) { out->push_back('.'); out->append(digits, 1, std::string::npos); }
        out->push_back('e');
        out->push_back(e10 < 0 ? '-' : '+');
        out->append(std::to_string(e10 < 0 ? -e10 : e10));  // "e-7", "e+21": two digits only arise for >= 10
        return;
    }
    if (e10 >= 0) {
        for (int i = 0; i <= e10; i++) out->push_back(i < nd ? digits[(size_t)i] : '0');
        if (nd > e10 + 1) { out->push_back('.'); out->append(digits, (size_t)e10 + 1, std::string::npos); }
    } else {
        out->append("0.");
        out->append((size_t)(-e10 - 1), '0');
        out->append(digits);
    }
}

// ------------------------------------------------------------------ gabs path operations

// Search restricted to what Get can observe: Get wants a string, and any array met before the
// end of the path turns the result into an array or nil (gabs.go:96-108), never a string.
std::string_view Document::get(std::string_view path) const {
    const Node *obj = root;
    for (std::string_view part : split(path, '.')) {
        if (!obj || obj->kind != Kind::Map) return {};
        auto &m = static_cast<const MapNode *>(obj)->m;
        auto it = m.find(part);
        obj = it == m.end() ? nullptr : it->second;
    }
    if (obj && obj->kind == Kind::Str) return static_cast<const StrNode *>(obj)->s;
    return {};
}

int Document::set(std::string_view path, Node *value, std::string *err) {
    auto parts = split(path, '.');
    if (!root) { root = new MapNode(); version++; }
    Node *obj = root;
    for (size_t t = 0; t < parts.size(); t++) {
        if (!obj || obj->kind != Kind::Map) {
            *err = kErrCollision;
            free_node(value);
            return 2;
        }
        auto &m = static_cast<MapNode *>(obj)->m;
        auto it = m.find(parts[t]);
        if (t + 1 == parts.size()) {
            if (it != m.end()) { free_node(it->second); it->second = value; }
            else it = m.emplace(std::string(parts[t]), value).first;
            version++;
        } else if (it == m.end()) {
            it = m.emplace(std::string(parts[t]), new MapNode()).first;
            version++;
        } else if (it->second == nullptr) {
            it->second = new MapNode();  // a JSON null on the way is replaced (mmap[k] == nil)
            version++;
        }
        obj = it->second;
    }
    return 0;
}

int Document::del(std::string_view path, std::string *err) {
    if (!root) { *err = kErrNotObj; return 3; }
    auto parts = split(path, '.');
    Node *obj = root;
    for (size_t t = 0; t < parts.size(); t++) {
        if (!obj || obj->kind != Kind::Map) { *err = kErrNotObj; return 3; }
        auto &m = static_cast<MapNode *>(obj)->m;
        auto it = m.find(parts[t]);
        if (t + 1 == parts.size()) {
            if (it != m.end()) { free_node(it->second); m.erase(it); version++; }
            return 0;
        }
        if (it == m.end() || it->second == nullptr) { *err = kErrNotObj; return 3; }
        obj = it->second;
    }
    return 0;
}

const MapNode *Document::module_children() const {
    if (!root || root->kind != Kind::Map) return nullptr;
    auto &m = static_cast<const MapNode *>(root)->m;
    auto it = m.find(std::string_view("module"));
    if (it == m.end() || !it->second || it->second->kind != Kind::Map) return nullptr;
    return static_cast<const MapNode *>(it->second);
}

bool cluster_key_parts(std::string_view cluster_key, std::string_view *provider, std::string_view *name, std::string *err) {
    // strings.Split(clusterKey, "_"): parts[1] and parts[2]; at least three parts = two underscores
    const size_t a = cluster_key.find('_');
    const size_t b = a == std::string_view::npos ? a : cluster_key.find('_', a + 1);
    if (b == std::string_view::npos) {
        *err = "Could not get cluster key parts, cluster does not follow format `cluster_{provider}_{clusterName}` '";
        err->append(cluster_key);
        err->push_back('\'');
        return false;
    }
    const size_t c = cluster_key.find('_', b + 1);
    *provider = cluster_key.substr(a + 1, b - a - 1);
    *name = cluster_key.substr(b + 1, c == std::string_view::npos ? std::string_view::npos : c - b - 1);
    return true;
}

// strconv.Atoi
static bool go_atoi(std::string_view s, long long *out) {
    if (s.empty()) return false;
    size_t i = 0;
    bool neg = false;
    if (s[0] == '+' || s[0] == '-') { neg = s[0] == '-'; i = 1; if (s.size() == 1) return false; }
    unsigned long long v = 0;
    for (; i < s.size(); i++) {
        if (s[i] < '0' || s[i] > '9') return false;
        unsigned d = (unsigned)(s[i] - '0');
        if (v > (~0ull - d) / 10) return false;
        v = v * 10 + d;
    }
    if (!neg && v > 9223372036854775807ull) return false;
    if (neg && v > 9223372036854775808ull) return false;
    *out = neg ? (long long)(0 - v) : (long long)v;
    return true;
}

std::vector<std::string> new_hostnames(const std::vector<std::string_view> &existing, std::string_view node_name,
                                       long long nodes_to_add) {
    std::vector<std::string> out;
    if (nodes_to_add < 1) return out;
    long long start = 1;
    std::string prefix(node_name);
    prefix.push_back('-');
    for (std::string_view e : existing) {
        if (e.size() < prefix.size() || e.compare(0, prefix.size(), prefix) != 0) continue;
        long long num;
        if (!go_atoi(e.substr(prefix.size()), &num)) continue;
        if (num >= start) start = num + 1;
    }
    for (long long i = 0; i < nodes_to_add; i++) out.push_back(std::string(node_name) + "-" + std::to_string(start + i));
    return out;
}

}  // namespace tk
