// engine.cu -- per-handle GPU context (see engine.hpp).
#include "engine.hpp"

#include <algorithm>
#include <cstring>

#include "kernels.cuh"
#include "tk_state.h"

namespace tk {

template <class T>
bool PinVec<T>::reserve(size_t want) {
    if (want <= cap) return true;
    size_t c = cap ? cap : 1024;
    while (c < want) c *= 2;
    T *q = nullptr;
    if (cudaHostAlloc(reinterpret_cast<void **>(&q), c * sizeof(T), cudaHostAllocDefault) != cudaSuccess) {
        (void)cudaGetLastError();
        failed = true;
        return false;
    }
    if (n) memcpy(q, p, n * sizeof(T));
    if (p) cudaFreeHost(p);
    p = q;
    cap = c;
    return true;
}
template struct PinVec<uint32_t>;
template struct PinVec<uint8_t>;
template struct PinVec<unsigned long long>;

cudaError_t DevBuf::reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    size_t c = cap ? cap : 4096;
    while (c < bytes) c *= 2;
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, c);
    if (e != cudaSuccess) return e;
    if (p) cudaFree(p);  // contents are never carried over: every buffer is rewritten after a grow
    p = q;
    cap = c;
    return cudaSuccess;
}

const char *kernel_name(int kind) {
    static const char *const names[K_COUNT] = {"row_offsets_kernel", "emit_kernel", "prefix_match_kernel"};
    return kind >= 0 && kind < K_COUNT ? names[kind] : "";
}

cudaError_t Engine::copy_h2d(void *dst, const void *src, size_t bytes) {
    h2d_bytes_ += bytes;
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream_);
}
cudaError_t Engine::copy_d2h(void *dst, const void *src, size_t bytes) {
    d2h_bytes_ += bytes;
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream_);
}

// With timing on, every kernel launch is bracketed by two events on the handle's stream; the
// elapsed times are summed per kernel when collect_times() is called.
template <class F>
cudaError_t Engine::timed(int kind, F &&launch) {
    launches_[kind]++;
    if (!timing_) return launch();
    cudaEvent_t a, b;
    if (free_events_.size() >= 2) {
        a = free_events_.back(); free_events_.pop_back();
        b = free_events_.back(); free_events_.pop_back();
    } else {
        cudaError_t e;
        if ((e = cudaEventCreate(&a)) != cudaSuccess) return e;
        if ((e = cudaEventCreate(&b)) != cudaSuccess) return e;
    }
    cudaEventRecord(a, stream_);
    cudaError_t e = launch();
    cudaEventRecord(b, stream_);
    pending_.push_back({kind, a, b});
    return e;
}

void Engine::set_timing(bool on) { timing_ = on; }

int Engine::collect_times(std::string *err) {
    if (!stream_) return TK_OK;
    cudaError_t e = cudaStreamSynchronize(stream_);
    if (e != cudaSuccess) return fail(e, "synchronize", err);
    for (auto &p : pending_) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) { kernel_ms_[p.kind] += ms; timed_launches_[p.kind]++; }
        free_events_.push_back(p.a);
        free_events_.push_back(p.b);
    }
    pending_.clear();
    return TK_OK;
}

Engine::~Engine() {
    for (auto &p : pending_) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto ev : free_events_) cudaEventDestroy(ev);
    if (stream_) {
        cudaStreamSynchronize(stream_);
        cudaStreamDestroy(stream_);
    }
}

int Engine::fail(cudaError_t e, const char *what, std::string *err) {
    *err = std::string("cuda: ") + what + ": " + cudaGetErrorString(e);
    (void)cudaGetLastError();
    return TK_ECUDA;
}

int Engine::init(std::string *err) {
    if (ready_) {
        cudaError_t e = cudaSetDevice(device_);
        if (e != cudaSuccess) return fail(e, "cudaSetDevice", err);
        return TK_OK;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        *err = std::string("cuda: no usable CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
               "); libtkstate has no CPU fallback";
        (void)cudaGetLastError();
        return TK_ECUDA;
    }
    e = cudaGetDevice(&device_);
    if (e != cudaSuccess) return fail(e, "cudaGetDevice", err);
    e = cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking);
    if (e != cudaSuccess) return fail(e, "cudaStreamCreate", err);
    ready_ = true;
    return TK_OK;
}

// ---------------------------------------------------------------------------- flatten

void Engine::push_row(uint32_t kind, uint32_t depth, const std::string *key, const char *val, size_t vlen, bool comma, bool last) {
    uint32_t meta = kind | (depth << META_DEPTH_SHIFT);
    uint32_t koff = 0, klen = 0, voff = 0;
    if (key) {
        meta |= META_HAS_KEY;
        koff = (uint32_t)h_arena_.append(reinterpret_cast<const uint8_t *>(key->data()), key->size());
        klen = (uint32_t)key->size();
    }
    if (vlen) voff = (uint32_t)h_arena_.append(reinterpret_cast<const uint8_t *>(val), vlen);
    if (comma) meta |= META_COMMA;
    if (last) meta |= META_LAST;
    h_meta_.push(meta);
    h_koff_.push(koff);
    h_klen_.push(klen);
    h_voff_.push(voff);
    h_vlen_.push((uint32_t)vlen);
}

// One row per output line, in the order encoding/json writes them: map keys ascending, record
// members as given, arrays in index order.
void Engine::flat_value(const Node *n, uint32_t depth, const std::string *key, bool comma, bool last) {
    if (!n) { push_row(ROW_RAW, depth, key, "null", 4, comma, last); return; }
    switch (n->kind) {
    case Kind::Bool: {
        bool v = static_cast<const BoolNode *>(n)->v;
        push_row(ROW_RAW, depth, key, v ? "true" : "false", v ? 4 : 5, comma, last);
        break;
    }
    case Kind::Num: {
        std::string t;
        format_go_float(static_cast<const NumNode *>(n)->v, &t);
        push_row(ROW_RAW, depth, key, t.data(), t.size(), comma, last);
        break;
    }
    case Kind::Str: {
        const std::string &s = static_cast<const StrNode *>(n)->s;
        push_row(ROW_STR, depth, key, s.data(), s.size(), comma, last);
        break;
    }
    case Kind::Arr: {
        auto &items = static_cast<const ArrNode *>(n)->items;
        if (items.empty()) { push_row(ROW_EMPTY_ARR, depth, key, nullptr, 0, comma, last); break; }
        push_row(ROW_OPEN_ARR, depth, key, nullptr, 0, false, false);
        for (size_t i = 0; i < items.size(); i++) flat_value(items[i], depth + 1, nullptr, i + 1 < items.size(), false);
        push_row(ROW_CLOSE_ARR, depth, nullptr, nullptr, 0, comma, last);
        break;
    }
    case Kind::Map: {
        auto &m = static_cast<const MapNode *>(n)->m;
        if (m.empty()) { push_row(ROW_EMPTY_OBJ, depth, key, nullptr, 0, comma, last); break; }
        push_row(ROW_OPEN_OBJ, depth, key, nullptr, 0, false, false);
        size_t i = 0;
        for (auto &kv : m) flat_value(kv.second, depth + 1, &kv.first, ++i < m.size(), false);
        push_row(ROW_CLOSE_OBJ, depth, nullptr, nullptr, 0, comma, last);
        break;
    }
    case Kind::Rec: {
        auto &m = static_cast<const RecNode *>(n)->m;
        if (m.empty()) { push_row(ROW_EMPTY_OBJ, depth, key, nullptr, 0, comma, last); break; }
        push_row(ROW_OPEN_OBJ, depth, key, nullptr, 0, false, false);
        for (size_t i = 0; i < m.size(); i++) flat_value(m[i].second, depth + 1, &m[i].first, i + 1 < m.size(), false);
        push_row(ROW_CLOSE_OBJ, depth, nullptr, nullptr, 0, comma, last);
        break;
    }
    }
}

void Engine::flatten(const Document &doc) {
    h_meta_.clear(); h_koff_.clear(); h_klen_.clear(); h_voff_.clear(); h_vlen_.clear(); h_arena_.clear();
    // the LAST flag (no trailing newline) lands on the root's own row, or on its closing row
    if (!doc.root) push_row(ROW_EMPTY_OBJ, 0, nullptr, nullptr, 0, false, true);  // gabs: nil -> "{}"
    else flat_value(doc.root, 0, nullptr, false, true);
    pin_failed_ = pin_failed_ || h_meta_.failed || h_koff_.failed || h_klen_.failed || h_voff_.failed || h_vlen_.failed ||
                  h_arena_.failed;
}

// packed keys of module's children, in map (= bytewise) order
void Engine::flatten_keys(const Document &doc) {
    h_koffs_.clear(); h_keys_.clear();
    entries_.clear();
    const MapNode *mod = doc.module_children();
    module_ok_ = mod != nullptr;
    h_koffs_.push(0);
    if (mod) {
        entries_.reserve(mod->m.size());
        for (auto &kv : mod->m) {
            h_keys_.append(reinterpret_cast<const uint8_t *>(kv.first.data()), kv.first.size());
            h_koffs_.push((uint32_t)h_keys_.n);
            entries_.push_back(&kv);
        }
    }
    pin_failed_ = pin_failed_ || h_koffs_.failed || h_keys_.failed;
}

int Engine::stage(const Document &doc, std::string *err) {
    int rc = stage_tape(doc, err);
    return rc ? rc : stage_keys(doc, err);
}

int Engine::stage_tape(const Document &doc, std::string *err) {
    int rc = init(err);
    if (rc) return rc;
    if (staged_version_ == doc.version) return TK_OK;
    flatten(doc);
    if (pin_failed_) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
    if (h_arena_.n >= 0xfffffff0ull || h_meta_.n >= 0xfffffff0ull) {
        *err = "document too large: string arena and row count must stay below 4 GiB";
        return TK_ELIMIT;
    }
    const size_t n = h_meta_.n;
    cudaError_t e;
    DevBuf *rows[5] = {&d_meta_, &d_koff_, &d_klen_, &d_voff_, &d_vlen_};
    PinVec<uint32_t> *hrows[5] = {&h_meta_, &h_koff_, &h_klen_, &h_voff_, &h_vlen_};
    for (int i = 0; i < 5; i++) {
        if ((e = rows[i]->reserve(n * 4)) != cudaSuccess) return fail(e, "cudaMalloc(rows)", err);
        if ((e = copy_h2d(rows[i]->p, hrows[i]->p, n * 4)) != cudaSuccess)
            return fail(e, "copy rows", err);
    }
    if ((e = d_arena_.reserve(h_arena_.n + 32)) != cudaSuccess) return fail(e, "cudaMalloc(arena)", err);
    if (h_arena_.n && (e = copy_h2d(d_arena_.p, h_arena_.p, h_arena_.n)) != cudaSuccess)
        return fail(e, "copy arena", err);
    if ((e = d_off_.reserve((n + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(off)", err);
    if ((e = d_scratch_.reserve(scan_scratch_words((uint32_t)n) * 8)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
    tape_.n_rows = (uint32_t)n;
    tape_.arena_bytes = (uint32_t)h_arena_.n;
    tape_.meta = d_meta_.as<uint32_t>(); tape_.koff = d_koff_.as<uint32_t>(); tape_.klen = d_klen_.as<uint32_t>();
    tape_.voff = d_voff_.as<uint32_t>(); tape_.vlen = d_vlen_.as<uint32_t>(); tape_.arena = d_arena_.as<uint8_t>();
    staged_version_ = doc.version;
    return TK_OK;
}

int Engine::stage_keys(const Document &doc, std::string *err) {
    int rc = init(err);
    if (rc) return rc;
    if (keys_version_ == doc.version) return TK_OK;
    flatten_keys(doc);
    if (pin_failed_) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
    if (h_keys_.n >= 0xfffffff0ull) { *err = "document too large: module keys must stay below 4 GiB"; return TK_ELIMIT; }
    cudaError_t e;
    const size_t nk = entries_.size();
    if ((e = d_koffs_.reserve((nk + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(koffs)", err);
    if ((e = d_keys_.reserve(h_keys_.n + 32)) != cudaSuccess) return fail(e, "cudaMalloc(keys)", err);
    if ((e = d_hits_.reserve((nk + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(hits)", err);
    if ((e = copy_h2d(d_koffs_.p, h_koffs_.p, (nk + 1) * 4)) != cudaSuccess)
        return fail(e, "copy key offsets", err);
    if (h_keys_.n && (e = copy_h2d(d_keys_.p, h_keys_.p, h_keys_.n)) != cudaSuccess)
        return fail(e, "copy keys", err);
    keys_.n_keys = (uint32_t)nk;
    keys_.key_bytes = (uint32_t)h_keys_.n;
    keys_.koffs = d_koffs_.as<uint32_t>();
    keys_.keys = d_keys_.as<uint8_t>();
    keys_version_ = doc.version;
    return TK_OK;
}

// ---------------------------------------------------------------------------- Bytes

int Engine::bytes(const Document &doc, bool to_host, const void **out, size_t *len, std::string *err) {
    int rc = stage_tape(doc, err);
    if (rc) return rc;
    cudaError_t e;
    if (!h_total_.reserve(1)) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
    unsigned long long *scratch = d_scratch_.as<unsigned long long>();
    if ((e = timed(K_ROW_OFFSETS, [&] { return launch_row_offsets(tape_, d_off_.as<uint32_t>(), scratch, stream_); })) != cudaSuccess) return fail(e, "row_offsets", err);
    if (total_version_ != doc.version) {
        // first serialisation of this version: the output size must be known before the text
        // buffer can be sized, which costs one synchronisation
        if ((e = copy_d2h(h_total_.p, scratch + 1, 8)) != cudaSuccess) return fail(e, "copy total", err);
        if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "row_offsets", err);
        if (h_total_.p[0] >= 0xffffffffull) { *err = "document too large: serialised text must stay below 4 GiB"; return TK_ELIMIT; }
        out_total_ = (size_t)h_total_.p[0];
        total_version_ = doc.version;
        if ((e = d_out_.reserve(out_total_ + 32)) != cudaSuccess) return fail(e, "cudaMalloc(out)", err);
    }
    if ((e = timed(K_EMIT, [&] { return launch_emit(tape_, d_off_.as<uint32_t>(), d_out_.as<uint8_t>(), stream_); })) != cudaSuccess) return fail(e, "emit", err);
    *len = out_total_;
    if (!to_host) { *out = d_out_.p; return TK_OK; }
    if (!h_out_.reserve(out_total_ + 1)) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
    if ((e = copy_d2h(h_out_.p, d_out_.p, out_total_)) != cudaSuccess) return fail(e, "copy text", err);
    if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "emit", err);
    *out = h_out_.p;
    return TK_OK;
}

// ---------------------------------------------------------------------------- Clusters / Nodes

int Engine::prefix_hits(const Document &doc, std::string_view prefix, bool to_host, std::vector<uint32_t> *hits,
                        size_t *count, std::string *err) {
    int rc = stage_keys(doc, err);
    if (rc) return rc;
    cudaError_t e;
    const size_t nk = keys_.n_keys;
    if ((e = d_prefix_.reserve(prefix.size() + 16)) != cudaSuccess) return fail(e, "cudaMalloc(prefix)", err);
    if (prefix.size() && (e = copy_h2d(d_prefix_.p, prefix.data(), prefix.size())) != cudaSuccess)
        return fail(e, "copy prefix", err);
    uint32_t *d_hits = d_hits_.as<uint32_t>();
    if ((e = timed(K_PREFIX_MATCH, [&] { return launch_prefix_match(keys_, d_prefix_.as<uint8_t>(), (uint32_t)prefix.size(), d_hits, stream_); })) != cudaSuccess)
        return fail(e, "prefix_match", err);
    // first chunk: the count plus as many indices as usually suffice, in one copy
    const size_t first = std::min<size_t>(nk + 1, 4096);
    if (!h_hits_.reserve(nk + 1)) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
    if ((e = copy_d2h(h_hits_.p, d_hits, first * 4)) != cudaSuccess) return fail(e, "copy hits", err);
    if (!to_host && count == nullptr) return TK_OK;
    if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "prefix_match", err);
    const size_t c = h_hits_.p[0];
    if (count) *count = c;
    if (!to_host) return TK_OK;
    if (c + 1 > first) {
        if ((e = copy_d2h(h_hits_.p + first, d_hits + first, (c + 1 - first) * 4)) != cudaSuccess)
            return fail(e, "copy hits", err);
        if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "copy hits", err);
    }
    hits->assign(h_hits_.p + 1, h_hits_.p + 1 + c);
    return TK_OK;
}

}  // namespace tk
