// engine.cu -- per-handle GPU context (see engine.hpp).
#include "engine.hpp"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <thread>

#include "kernels.cuh"
#include "lines.cuh"
#include "tk_state.h"

namespace tk {

// ---------------------------------------------------------------------------- buffer pools

namespace {

size_t size_class(size_t bytes) {
    if (bytes <= 256) return 256;
    if (bytes > (size_t(1) << 30)) return (bytes + (size_t(1) << 28) - 1) & ~((size_t(1) << 28) - 1);
    size_t c = 256;
    while (c < bytes) c <<= 1;
    return c;
}

struct Pool {
    std::mutex mu;
    std::map<std::pair<int, size_t>, std::vector<void *>> free_blocks;  // (device, size class) -> blocks
};
Pool g_dev_pool, g_pin_pool;

}  // namespace

void *pool_alloc_device(size_t bytes, size_t *got) {
    const size_t c = size_class(bytes);
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
    {
        std::lock_guard<std::mutex> lock(g_dev_pool.mu);
        auto &v = g_dev_pool.free_blocks[{dev, c}];
        if (!v.empty()) { void *p = v.back(); v.pop_back(); *got = c; return p; }
    }
    void *p = nullptr;
    if (cudaMalloc(&p, c) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
    *got = c;
    return p;
}
void pool_free_device(void *p, size_t bytes) {
    if (!p) return;
    int dev = 0;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) == cudaSuccess) dev = a.device; else (void)cudaGetLastError();
    std::lock_guard<std::mutex> lock(g_dev_pool.mu);
    g_dev_pool.free_blocks[{dev, bytes}].push_back(p);
}
void *pool_alloc_pinned(size_t bytes, size_t *got) {
    const size_t c = size_class(bytes);
    {
        std::lock_guard<std::mutex> lock(g_pin_pool.mu);
        auto &v = g_pin_pool.free_blocks[{0, c}];
        if (!v.empty()) { void *p = v.back(); v.pop_back(); *got = c; return p; }
    }
    void *p = nullptr;
    if (cudaHostAlloc(&p, c, cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
    *got = c;
    return p;
}
void pool_free_pinned(void *p, size_t bytes) {
    if (!p) return;
    std::lock_guard<std::mutex> lock(g_pin_pool.mu);
    g_pin_pool.free_blocks[{0, bytes}].push_back(p);
}

static thread_local cudaStream_t tl_scope_stream = nullptr;
StreamScope::StreamScope(cudaStream_t s) : prev_(tl_scope_stream) { tl_scope_stream = s; }
StreamScope::~StreamScope() { tl_scope_stream = prev_; }
void pool_release_sync() {
    if (tl_scope_stream && cudaStreamSynchronize(tl_scope_stream) != cudaSuccess) (void)cudaGetLastError();
}

cudaError_t DevBuf::reserve(size_t bytes, size_t keep_bytes, cudaStream_t s) {
    if (bytes <= cap) return cudaSuccess;
    size_t got = 0;
    void *q = pool_alloc_device(bytes, &got);
    if (!q) return cudaErrorMemoryAllocation;
    if (p && keep_bytes) {
        cudaError_t e = cudaMemcpyAsync(q, p, keep_bytes, cudaMemcpyDeviceToDevice, s);
        if (e != cudaSuccess) { pool_free_device(q, got); return e; }
        // the old block goes back to the pool; work queued on other streams never uses it
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) { pool_free_device(q, got); return e; }
    }
    if (p) { pool_release_sync(); pool_free_device(p, cap); }
    p = q;
    cap = got;
    return cudaSuccess;
}

// Development aid: TK_TRACE=1 prints, per handle, the host time at which each phase of New / Bytes /
// the index build was reached (microseconds since the phase list was started), to stderr.
namespace {
struct Trace {
    bool on;
    std::chrono::steady_clock::time_point t0;
    std::string line;
    Trace() : on(getenv("TK_TRACE") != nullptr) { if (on) t0 = std::chrono::steady_clock::now(); }
    void mark(const char *what) {
        if (!on) return;
        const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
        char buf[64];
        snprintf(buf, sizeof buf, " %s=%.0f", what, us);
        line += buf;
    }
    ~Trace() { if (on && !line.empty()) fprintf(stderr, "[tk trace]%s\n", line.c_str()); }
};
}  // namespace

static std::atomic<bool> g_default_timing{false};
void set_default_timing(bool on) { g_default_timing = on; }
static std::atomic<int> g_line_pass{-1};   // -1: not asked yet (TK_NO_FAST_PATH in the environment switches it off)
void set_line_pass(bool on) { g_line_pass.store(on ? 1 : 0); }
// bytes of side arena the line pass may fill with decoded strings (more than that: the general pipeline takes the text)
static size_t lines_side_cap(size_t n) { return std::min<size_t>(n / 8 + 4096, 0xffffff00ull - std::min<size_t>(0xffffff00ull, (size_t)p1_tiles(n) * P1_TILE + 128)); }
static bool line_pass_on() {
    int v = g_line_pass.load();
    if (v < 0) { v = getenv("TK_NO_FAST_PATH") ? 0 : 1; g_line_pass.store(v); }
    return v != 0;
}

const char *kernel_name(int kind) {
    static const char *const names[K_COUNT] = {"row_offsets_kernel", "emit_kernel", "prefix_match_kernel", "scan_tokens_kernel (carry protocol)",
                                               "token_agg + token_blocks kernels", "brackets_kernel", "sort_brackets (radix passes)", "match_kernel",
                                               "decode kernels", "rows_kernel", "key_order_kernel", "field_pack_kernel (module index)", "field_find_kernel",
                                               "lines_count + lines_prefix kernels", "lines_rows_kernel", "lines_open_prefix + lines_resolve kernels (containers, field lists)",
                                               "scan_count_kernel", "scan_prefix_kernel", "scan_emit_kernel"};
    return kind >= 0 && kind < K_COUNT ? names[kind] : "";
}

cudaError_t Engine::copy_h2d(void *dst, const void *src, size_t bytes) {
    h2d_bytes_ += bytes;
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream_);
}
// Small results (counts, the parse verdict: at most a few words) are written into pinned host
// memory by a kernel.  A DMA copy of them would wait in the copy engine's queue behind the bulk
// transfers of every other document in flight on this GPU, and the host waits for each of them.
__global__ void publish_words_kernel(uint32_t *host_dst, const uint32_t *dev_src, uint32_t words) {
    for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) host_dst[i] = dev_src[i];
}

cudaError_t Engine::copy_d2h(void *dst, const void *src, size_t bytes) {
    d2h_bytes_ += bytes;
    if (bytes <= 256 && bytes % 4 == 0 && (reinterpret_cast<uintptr_t>(dst) & 3) == 0 && (reinterpret_cast<uintptr_t>(src) & 3) == 0) {
        // every small destination in this file is a pinned block from pool_alloc_pinned (mapped)
        publish_words_kernel<<<1, 32, 0, stream_>>>(static_cast<uint32_t *>(dst), static_cast<const uint32_t *>(src), (uint32_t)(bytes / 4));
        count_parse_launch();
        return cudaGetLastError();
    }
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

// streams are recycled like the buffers: a handle takes an idle one and gives it back drained
namespace {
struct StreamPool {
    std::mutex mu;
    std::map<int, std::vector<cudaStream_t>> idle;   // per device
} g_stream_pool;
}  // namespace

Engine::~Engine() {
    drop_edit_state();
    for (auto &ev : stage_ev_) if (ev) cudaEventDestroy(ev);
    for (auto &p : pending_) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto ev : free_events_) cudaEventDestroy(ev);
    if (stream_) {
        if (cudaStreamSynchronize(stream_) == cudaSuccess) {
            std::lock_guard<std::mutex> lock(g_stream_pool.mu);
            g_stream_pool.idle[device_].push_back(stream_);
        } else {
            (void)cudaGetLastError();
            cudaStreamDestroy(stream_);
        }
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
    {
        std::lock_guard<std::mutex> lock(g_stream_pool.mu);
        auto &idle = g_stream_pool.idle[device_];
        if (!idle.empty()) { stream_ = idle.back(); idle.pop_back(); }
    }
    if (!stream_) {
        e = cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking);
        if (e != cudaSuccess) return fail(e, "cudaStreamCreate", err);
    }
    timing_ = g_default_timing.load();
    ready_ = true;
    return TK_OK;
}

// ---------------------------------------------------------------------------- flatten

// every byte printable ASCII other than " \\ < > & : the escaped form is the byte string itself
static bool is_plain(const char *s, size_t n) {
    static const struct Table {
        bool plain[256];
        Table() {
            for (int c = 0; c < 256; c++)
                plain[c] = c >= 0x20 && c < 0x80 && c != '"' && c != '\\' && c != '<' && c != '>' && c != '&';
        }
    } t;
    for (size_t i = 0; i < n; i++)
        if (!t.plain[(unsigned char)s[i]]) return false;
    return true;
}

// Where the rows of a flatten go: the handle's pinned arrays (FlatPinned) or, for a worker thread
// of a wide object, plain vectors that are appended to them afterwards (FlatLocal).
struct FlatPinned {
    PinVec<uint32_t> &meta, &koff, &klen, &voff, &vlen;
    PinVec<uint8_t> &arena;
    static constexpr bool may_fork = true;
    size_t put(const char *p, size_t n) { return arena.append(reinterpret_cast<const uint8_t *>(p), n); }
    void row(uint32_t m, uint32_t ko, uint32_t kl, uint32_t vo, uint32_t vl) { meta.push(m); koff.push(ko); klen.push(kl); voff.push(vo); vlen.push(vl); }
};
struct FlatLocal {
    std::vector<uint32_t> meta, koff, klen, voff, vlen;
    std::vector<uint8_t> arena;
    static constexpr bool may_fork = false;
    size_t put(const char *p, size_t n) { const size_t at = arena.size(); arena.insert(arena.end(), p, p + n); return at; }
    void row(uint32_t m, uint32_t ko, uint32_t kl, uint32_t vo, uint32_t vl) {
        meta.push_back(m); koff.push_back(ko); klen.push_back(kl); voff.push_back(vo); vlen.push_back(vl);
    }
};

template <class Sink>
static void push_row(Sink &out, uint32_t kind, uint32_t depth, const std::string *key, const char *val, size_t vlen, bool comma, bool last) {
    uint32_t meta = kind | (depth << META_DEPTH_SHIFT);
    uint32_t koff = 0, klen = 0, voff = 0;
    if (kind == ROW_STR && is_plain(val, vlen)) meta |= META_VAL_PLAIN;
    if (key) {
        meta |= META_HAS_KEY;
        if (is_plain(key->data(), key->size())) meta |= META_KEY_PLAIN;
        koff = (uint32_t)out.put(key->data(), key->size());
        klen = (uint32_t)key->size();
    }
    if (vlen) voff = (uint32_t)out.put(val, vlen);
    if (comma) meta |= META_COMMA;
    if (last) meta |= META_LAST;
    out.row(meta, koff, klen, voff, (uint32_t)vlen);
}

template <class Sink>
static void flat_value(Sink &out, const Node *n, uint32_t depth, const std::string *key, bool comma, bool last);

// An object with very many members (the "module" of a large state document): the members are
// flattened by several threads, each into its own vectors, and appended in order.  The tree is
// only read.  Offsets of a worker's rows are relative to its own arena and are moved on append.
constexpr size_t FORK_MIN_MEMBERS = 8192;
typedef std::pair<const std::string *, const Node *> MemberRef;
static void flat_pairs_forked(FlatPinned &out, const std::vector<MemberRef> &members, uint32_t depth, bool comma_after_last);
static void flat_members_forked(FlatPinned &out, const MapNode *node, uint32_t depth) {
    std::vector<MemberRef> members;
    members.reserve(node->m.size());
    for (auto &kv : node->m) members.emplace_back(&kv.first, kv.second);
    flat_pairs_forked(out, members, depth, false);
}
// members (key, value) flattened one after another at `depth`, a comma after each but -- unless
// comma_after_last -- the last
static void flat_pairs_forked(FlatPinned &out, const std::vector<MemberRef> &members, uint32_t depth, bool comma_after_last) {
    const size_t n = members.size();
    if (n < 1024) {
        for (size_t i = 0; i < n; i++) flat_value(out, members[i].second, depth, members[i].first, comma_after_last || i + 1 < n, false);
        return;
    }
    const size_t hw = std::max<size_t>(1, std::thread::hardware_concurrency());
    const size_t workers = std::max<size_t>(2, std::min<size_t>({hw, 16, n / 1024}));
    std::vector<FlatLocal> parts(workers);
    std::vector<std::thread> threads;
    std::atomic<bool> failed{false};
    for (size_t w = 0; w < workers; w++)
        threads.emplace_back([&, w] {
            try {
                FlatLocal &p = parts[w];
                for (size_t i = w * n / workers; i < (w + 1) * n / workers; i++)
                    flat_value(p, members[i].second, depth, members[i].first, comma_after_last || i + 1 < n, false);
            } catch (...) {   // out of memory: reported like a failed pinned allocation
                failed = true;
            }
        });
    for (auto &t : threads) t.join();
    if (failed) { out.arena.failed = true; return; }
    // one growth step for everything the workers produced (pinned allocations are not cheap)
    size_t rows = 0, bytes = 0;
    for (const FlatLocal &p : parts) { rows += p.meta.size(); bytes += p.arena.size(); }
    out.arena.reserve(out.arena.n + bytes + 4096);
    for (PinVec<uint32_t> *v : {&out.meta, &out.koff, &out.klen, &out.voff, &out.vlen}) v->reserve(v->n + rows + 64);
    for (const FlatLocal &p : parts) {
        const uint32_t base = (uint32_t)out.arena.append(p.arena.data(), p.arena.size());
        for (size_t r = 0; r < p.meta.size(); r++)
            out.row(p.meta[r], p.klen[r] ? p.koff[r] + base : 0u, p.klen[r], p.vlen[r] ? p.voff[r] + base : 0u, p.vlen[r]);
    }
}

// One row per output line, in the order encoding/json writes them: map keys ascending, record
// members as given, arrays in index order.
template <class Sink>
static void flat_value(Sink &out, const Node *n, uint32_t depth, const std::string *key, bool comma, bool last) {
    if (!n) { push_row(out, ROW_RAW, depth, key, "null", 4, comma, last); return; }
    switch (n->kind) {
    case Kind::Bool: {
        bool v = static_cast<const BoolNode *>(n)->v;
        push_row(out, ROW_RAW, depth, key, v ? "true" : "false", v ? 4 : 5, comma, last);
        break;
    }
    case Kind::Num: {
        std::string t;
        format_go_float(static_cast<const NumNode *>(n)->v, &t);
        push_row(out, ROW_RAW, depth, key, t.data(), t.size(), comma, last);
        break;
    }
    case Kind::Str: {
        const std::string &s = static_cast<const StrNode *>(n)->s;
        push_row(out, ROW_STR, depth, key, s.data(), s.size(), comma, last);
        break;
    }
    case Kind::Arr: {
        auto &items = static_cast<const ArrNode *>(n)->items;
        if (items.empty()) { push_row(out, ROW_EMPTY_ARR, depth, key, nullptr, 0, comma, last); break; }
        push_row(out, ROW_OPEN_ARR, depth, key, nullptr, 0, false, false);
        for (size_t i = 0; i < items.size(); i++) flat_value(out, items[i], depth + 1, nullptr, i + 1 < items.size(), false);
        push_row(out, ROW_CLOSE_ARR, depth, nullptr, nullptr, 0, comma, last);
        break;
    }
    case Kind::Map: {
        auto &m = static_cast<const MapNode *>(n)->m;
        if (m.empty()) { push_row(out, ROW_EMPTY_OBJ, depth, key, nullptr, 0, comma, last); break; }
        push_row(out, ROW_OPEN_OBJ, depth, key, nullptr, 0, false, false);
        if constexpr (Sink::may_fork) {
            if (m.size() >= FORK_MIN_MEMBERS) {
                flat_members_forked(out, static_cast<const MapNode *>(n), depth + 1);
                push_row(out, ROW_CLOSE_OBJ, depth, nullptr, nullptr, 0, comma, last);
                break;
            }
        }
        size_t i = 0;
        for (auto &kv : m) flat_value(out, kv.second, depth + 1, &kv.first, ++i < m.size(), false);
        push_row(out, ROW_CLOSE_OBJ, depth, nullptr, nullptr, 0, comma, last);
        break;
    }
    case Kind::Rec: {
        auto &m = static_cast<const RecNode *>(n)->m;
        if (m.empty()) { push_row(out, ROW_EMPTY_OBJ, depth, key, nullptr, 0, comma, last); break; }
        push_row(out, ROW_OPEN_OBJ, depth, key, nullptr, 0, false, false);
        for (size_t i = 0; i < m.size(); i++) flat_value(out, m[i].second, depth + 1, &m[i].first, i + 1 < m.size(), false);
        push_row(out, ROW_CLOSE_OBJ, depth, nullptr, nullptr, 0, comma, last);
        break;
    }
    }
}

void Engine::flatten(const Document &doc) {
    h_meta_.clear(); h_koff_.clear(); h_klen_.clear(); h_voff_.clear(); h_vlen_.clear(); h_arena_.clear();
    FlatPinned out{h_meta_, h_koff_, h_klen_, h_voff_, h_vlen_, h_arena_};
    // the LAST flag (no trailing newline) lands on the root's own row, or on its closing row
    if (!doc.root) push_row(out, ROW_EMPTY_OBJ, 0, nullptr, nullptr, 0, false, true);  // gabs: nil -> "{}"
    else flat_value(out, doc.root, 0, nullptr, false, true);
    pin_failed_ = pin_failed_ || h_meta_.failed || h_koff_.failed || h_klen_.failed || h_voff_.failed || h_vlen_.failed ||
                  h_arena_.failed;
}

// packed keys of module's children, in map (= bytewise) order
void Engine::flatten_keys(const Document &doc) {
    h_koffs_.clear(); h_keys_.clear();
    entries_.clear();
    const MapNode *mod = doc.module_children();
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
    StreamScope scope(stream_);
    if (staged_version_ == doc.version) return TK_OK;
    if (!doc.host_valid && edit_) return apply_edits(doc, err);   // edited on the device: the tape is rebuilt there
    if (!doc.host_valid) { *err = "internal: no host tree to flatten"; return TK_EARG; }
    tape_from_parse_ = false;
    // copies of the previous staging may still be reading the pinned arrays flatten() rewrites
    { cudaError_t e0 = cudaStreamSynchronize(stream_); if (e0 != cudaSuccess) return fail(e0, "synchronize", err); }
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
    tape_.text_len = 0;
    tape_.rows_direct = false;
    tape_.meta = d_meta_.as<uint32_t>(); tape_.koff = d_koff_.as<uint32_t>(); tape_.klen = d_klen_.as<uint32_t>();
    tape_.voff = d_voff_.as<uint32_t>(); tape_.vlen = d_vlen_.as<uint32_t>(); tape_.arena = d_arena_.as<uint8_t>();
    staged_version_ = doc.version;
    return TK_OK;
}

int Engine::stage_keys(const Document &doc, std::string *err) {
    int rc = init(err);
    if (rc) return rc;
    StreamScope scope(stream_);
    if (keys_version_ == doc.version) return TK_OK;
    { cudaError_t e0 = cudaStreamSynchronize(stream_); if (e0 != cudaSuccess) return fail(e0, "synchronize", err); }
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
    StreamScope scope(stream_);
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
    StreamScope scope(stream_);
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

// ============================================================================ state.New on the GPU

namespace tk {

int Engine::fetch_result(std::string *err) {
    cudaError_t e;
    if (!h_res_.reserve(1)) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
    if ((e = copy_d2h(h_res_.p, d_res_.p, sizeof(ParseResult))) != cudaSuccess) return fail(e, "copy parse result", err);
    if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "parse", err);
    return TK_OK;
}

// text -> start of the arena, space padded to a whole number of 16 KiB tiles (+32)
int Engine::upload_text(const char *raw, size_t n, std::string *err) {
    int rc = init(err);
    if (rc) return rc;
    StreamScope scope(stream_);
    pending_text_ = nullptr;
    if (n >= 0xfff00000ull) { *err = "document too large: the text must stay below 4 GiB"; return TK_ELIMIT; }
    const size_t padded = (size_t)p1_tiles(n) * P1_TILE + 64;
    cudaError_t e;
    // (with room behind the text for the strings the line pass has to decode: lines_side_cap)
    if ((e = d_arena_.reserve(line_pass_on() ? ((padded + 15) & ~size_t(15)) + lines_side_cap(n) + 32 : padded)) != cudaSuccess) return fail(e, "cudaMalloc(arena)", err);
    // Text in ordinary (pageable) host memory -- what a Go caller holds: ioutil.ReadFile's slice,
    // backend/local/backend.go:60-65 -- is not copied by the driver's own staging (a few GB/s) but
    // through two pinned blocks of this handle: the calling thread fills one while the copy engine
    // drains the other.  Pinned or registered memory is copied from where it is.
    bool pageable = false;
    if (n >= STAGE_MIN) {
        cudaPointerAttributes attr;
        if (cudaPointerGetAttributes(&attr, raw) != cudaSuccess) { (void)cudaGetLastError(); pageable = true; }
        else pageable = attr.type == cudaMemoryTypeUnregistered;
    }
    if (pageable) {
        if (!h_stage_[0].reserve(STAGE_CHUNK) || !h_stage_[1].reserve(STAGE_CHUNK)) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
        for (int k = 0; k < 2; k++)
            if (!stage_ev_[k] && (e = cudaEventCreateWithFlags(&stage_ev_[k], cudaEventDisableTiming)) != cudaSuccess) return fail(e, "cudaEventCreate", err);
        bool busy[2] = {false, false};
        int k = 0;
        for (size_t off = 0; off < n; off += STAGE_CHUNK, k ^= 1) {
            const size_t len = std::min(STAGE_CHUNK, n - off);
            if (busy[k] && (e = cudaEventSynchronize(stage_ev_[k])) != cudaSuccess) return fail(e, "copy text", err);
            memcpy(h_stage_[k].p, raw + off, len);
            if ((e = copy_h2d(d_arena_.as<uint8_t>() + off, h_stage_[k].p, len)) != cudaSuccess) return fail(e, "copy text", err);
            if ((e = cudaEventRecord(stage_ev_[k], stream_)) != cudaSuccess) return fail(e, "copy text", err);
            busy[k] = true;
        }
    } else if (n && (e = copy_h2d(d_arena_.p, raw, n)) != cudaSuccess) {
        return fail(e, "copy text", err);
    }
    if ((e = cudaMemsetAsync(d_arena_.as<uint8_t>() + n, ' ', padded - n, stream_)) != cudaSuccess) return fail(e, "pad text", err);
    raw_len_ = n;
    arena_used_ = padded;
    staged_version_ = keys_version_ = total_version_ = 0;
    return TK_OK;
}

// the same for text that is already in device memory
int Engine::upload_text_device(const void *dev_raw, size_t n, std::string *err) {
    int rc = init(err);
    if (rc) return rc;
    StreamScope scope(stream_);
    if (n >= 0xfff00000ull) { *err = "document too large: the text must stay below 4 GiB"; return TK_ELIMIT; }
    const size_t padded = (size_t)p1_tiles(n) * P1_TILE + 64;
    cudaError_t e;
    // (with room behind the text for the strings the line pass has to decode: lines_side_cap)
    if ((e = d_arena_.reserve(line_pass_on() ? ((padded + 15) & ~size_t(15)) + lines_side_cap(n) + 32 : padded)) != cudaSuccess) return fail(e, "cudaMalloc(arena)", err);
    // (the line pass brings the text over itself, in its first pass over it: parse_lines)
    pending_text_ = n && line_pass_on() && (reinterpret_cast<uintptr_t>(dev_raw) & 15u) == 0 ? static_cast<const uint8_t *>(dev_raw) : nullptr;
    if (n && !pending_text_ && (e = cudaMemcpyAsync(d_arena_.p, dev_raw, n, cudaMemcpyDeviceToDevice, stream_)) != cudaSuccess) return fail(e, "copy text", err);
    if ((e = cudaMemsetAsync(d_arena_.as<uint8_t>() + n, ' ', padded - n, stream_)) != cudaSuccess) return fail(e, "pad text", err);
    raw_len_ = n;
    arena_used_ = padded;
    staged_version_ = keys_version_ = total_version_ = 0;
    return TK_OK;
}

// state.New for text in the indented form (lines.cu): every launch queued at once -- sizes come from
// bounds, the real counts stay on the device -- and one wait at the end for the verdict.
int Engine::parse_lines(size_t n, ParseResult *out, bool *accepted, std::string *err) {
    *accepted = false;
    cudaError_t e;
    const uint8_t *raw = d_arena_.as<uint8_t>();
    ParseResult *d_res = d_res_.as<ParseResult>();
    // (a line of the indented form is rarely under 16 bytes; a text with more lines than that is declined)
    const size_t row_cap = std::min<size_t>(n / 16 + 4096, 0x7fffffffu);
    DevBuf *rows[5] = {&d_meta_, &d_koff_, &d_klen_, &d_voff_, &d_vlen_};
    for (int i = 0; i < 5; i++)
        if ((e = rows[i]->reserve((row_cap + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(rows)", err);
    if ((e = d_scratch_.reserve(lines_scratch_words(n) * 4 + 64)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
    if ((e = d_tiles_.reserve(lines_open_state_words((uint32_t)row_cap) * 4 + 64)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
    RowBuffers rb{d_meta_.as<uint32_t>(), d_koff_.as<uint32_t>(), d_klen_.as<uint32_t>(), d_voff_.as<uint32_t>(), d_vlen_.as<uint32_t>(),
                  d_tiles_.as<uint32_t>()};
    // strings that have to be decoded go behind the text, as in the general parse (what upload_text left room for)
    const size_t side_base = ((size_t)p1_tiles(n) * P1_TILE + 64 + 15) & ~size_t(15);
    rb.side = d_arena_.as<uint8_t>();
    rb.side_base = (uint32_t)side_base;
    rb.side_cap = d_arena_.cap > side_base + 32 ? (uint32_t)std::min<size_t>(d_arena_.cap - side_base - 32, lines_side_cap(n)) : 0u;
    uint32_t *scratch = d_scratch_.as<uint32_t>();
    // the module index's lists are made by the last stage (the row count is still on the device: sized by bounds)
    for (int f = 0; f < 2; f++) {
        const size_t cap = std::min<size_t>(row_cap, n / 256 + 1024);
        if ((e = d_frefs_[f].reserve(cap * sizeof(FieldRef))) != cudaSuccess || (e = d_foff_[f].reserve((cap + 1) * 4)) != cudaSuccess)
            return fail(e, "cudaMalloc(index)", err);
        fx_.refs[f] = d_frefs_[f].as<FieldRef>();
        fx_.off[f] = d_foff_[f].as<uint32_t>();
        fx_.cap[f] = (uint32_t)cap;
    }
    if ((e = d_find_status_.reserve(field_status_words((uint32_t)row_cap) * 8)) != cudaSuccess) return fail(e, "cudaMalloc(index)", err);
    fx_.status = d_find_status_.as<unsigned long long>();   // (for a second run of field_find with larger lists; the first is part of stage 2)
    const uint8_t *src = pending_text_;
    pending_text_ = nullptr;
    if ((e = timed(K_LINES_COUNT, [&] { return launch_lines_count(d_arena_.as<uint8_t>(), src, n, scratch, rb.open, (uint32_t)row_cap, d_res, stream_); })) != cudaSuccess)
        return fail(e, "lines_count", err);
    if ((e = timed(K_LINES_ROWS, [&] { return launch_lines_rows(raw, n, scratch, rb, (uint32_t)row_cap, d_res, stream_); })) != cudaSuccess)
        return fail(e, "lines_rows", err);
    if ((e = timed(K_LINES_STRUCTURE, [&] { return launch_lines_structure(raw, rb, (uint32_t)row_cap, rb.open, fx_, d_res, stream_); })) != cudaSuccess)
        return fail(e, "lines_structure", err);
    tape_.n_rows = (uint32_t)row_cap;
    tape_.arena_bytes = (uint32_t)arena_used_;
    tape_.text_len = (uint32_t)n;
    tape_.rows_direct = true;
    tape_.meta = rb.meta; tape_.koff = rb.koff; tape_.klen = rb.klen; tape_.voff = rb.voff; tape_.vlen = rb.vlen;
    tape_.arena = raw;
    int rc = fetch_result(err);
    if (rc) return rc;
    if (h_res_.p->fast_reject || h_res_.p->n_rows == 0 || h_res_.p->n_rows > row_cap) return TK_OK;
    *out = *h_res_.p;
    *accepted = true;
    if (out->side_bytes) {   // (more may have been asked for than rows refer to -- a line looked at twice --: all of it counts as used)
        arena_used_ = side_base + (size_t)std::min<unsigned long long>(out->side_bytes, rb.side_cap);
        tape_.arena_bytes = (uint32_t)arena_used_;
    }
    const uint32_t n_rows = out->n_rows;
    if ((e = d_off_.reserve(((size_t)n_rows + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(off)", err);
    if ((e = d_scratch_.reserve(scan_scratch_words(n_rows) * 8)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
    tape_.n_rows = n_rows;
    tape_from_parse_ = true;
    bb_ = BracketBuffers{};
    parsed_ = *out;
    index_version_ = 0;
    return TK_OK;
}

// Every kernel of the pipeline, on the text already resident at the start of the arena.  Leaves
// the tape (rows in input order) on the device and the verdict in *out.  Host synchronisations:
// after P1 (token count sizes the later buffers), after P2a (row / bracket counts), after the
// last kernel; two more only if some string needs decoding.
int Engine::parse_resident(size_t n, ParseResult *out, std::string *err) {
    cudaError_t e;
    StreamScope scope(stream_);
    const uint8_t *raw = d_arena_.as<uint8_t>();
    memset(out, 0, sizeof *out);
    Trace tr;
    tr.mark("new:start");
    if ((e = d_res_.reserve(sizeof(ParseResult))) != cudaSuccess) return fail(e, "cudaMalloc(result)", err);
    ParseResult *d_res = d_res_.as<ParseResult>();
    if (n == 0) { out->invalid = 1; return TK_OK; }

    // ---- text that is already in the indented form (every file this library or the reference wrote):
    //      the line-by-line pass builds the tape; anything else it declines, and the general pipeline runs
    if (pending_text_ && !line_pass_on()) {   // (switched off between the upload and here)
        if ((e = cudaMemcpyAsync(d_arena_.p, pending_text_, n, cudaMemcpyDeviceToDevice, stream_)) != cudaSuccess) return fail(e, "copy text", err);
        pending_text_ = nullptr;
    }
    if (line_pass_on()) {   // (tk_set_line_pass / TK_NO_FAST_PATH: the tests run every vector through both)
        bool accepted = false;
        int rc = parse_lines(n, out, &accepted, err);
        tr.mark("lines:done");
        if (rc || accepted) return rc;
        memset(out, 0, sizeof *out);
    }

    // ---- P1 (re-run with a larger token buffer in the rare case the estimate was too small)
    size_t cap = n / 4 + 4096;
    TokenBuffers tb{};
    bool carry_protocol = false;  // set when the fast kernel met a backslash run longer than its look-behind
    for (;;) {
        if ((e = d_tok_pos_.reserve((cap + 2) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(tokens)", err);
        if ((e = d_tok_kind_.reserve(cap + 4096 + 64)) != cudaSuccess) return fail(e, "cudaMalloc(tokens)", err);
        if ((e = d_dec_bits_.reserve((cap + 31) / 32 * 4 + 1024)) != cudaSuccess) return fail(e, "cudaMalloc(tokens)", err);
        if ((e = d_scratch_.reserve(p1_scratch_words(n) * 8)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
        tb.pos = d_tok_pos_.as<uint32_t>(); tb.kind = d_tok_kind_.as<uint8_t>(); tb.decode_bits = d_dec_bits_.as<uint32_t>();
        tb.cap = (uint32_t)cap;
        static const int kinds[2][3] = {{K_SCAN_COUNT, K_SCAN_PREFIX, K_SCAN_EMIT}, {K_SCAN_TOKENS, K_SCAN_TOKENS, K_SCAN_TOKENS}};
        for (int stage = 0; stage < (carry_protocol ? 1 : 3); stage++)
            if ((e = timed(kinds[carry_protocol ? 1 : 0][stage], [&] {
                     return launch_scan_tokens(raw, n, tb, d_scratch_.as<unsigned long long>(), d_res, carry_protocol, stage, stream_);
                 })) != cudaSuccess)
                return fail(e, "scan_tokens", err);
        tr.mark("scan:queued");
        int rc = fetch_result(err);
        tr.mark("scan:done");
        if (rc) return rc;
        if (h_res_.p->slow_scan && !carry_protocol) { carry_protocol = true; continue; }
        if (h_res_.p->n_tokens <= cap) break;
        if (h_res_.p->n_tokens >= 0x7ffffff0ull) { *err = "document too large: more than 2^31 tokens"; return TK_ELIMIT; }
        cap = (size_t)h_res_.p->n_tokens;
    }
    *out = *h_res_.p;
    if (out->n_tokens == 0 || out->end_in_string) out->invalid = 1;
    if (out->invalid) return TK_OK;
    const uint32_t n_tok = (uint32_t)out->n_tokens;

    // ---- P2a
    if ((e = d_tiles_.reserve((size_t)tok_tiles(n_tok) * sizeof(TilePrefix))) != cudaSuccess) return fail(e, "cudaMalloc(tiles)", err);
    if ((e = d_scratch_.reserve(tok_scratch_words(n_tok) * 8)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
    TilePrefix *tiles = d_tiles_.as<TilePrefix>();
    if ((e = timed(K_TOKEN_PREFIX, [&] { return launch_token_prefix(tb, n_tok, tiles, d_scratch_.as<unsigned long long>(), d_res, stream_); })) != cudaSuccess)
        return fail(e, "token_prefix", err);
    tr.mark("tokprefix:queued");
    int rc = fetch_result(err);
    tr.mark("tokprefix:done");
    if (rc) return rc;
    *out = *h_res_.p;
    if (out->invalid) return TK_OK;
    const uint32_t n_rows = out->n_rows, n_br = out->n_brackets;

    // ---- brackets: compact, sort by level, pair
    BracketBuffers bb{};
    {
        uint32_t **slots[6] = {&bb.tok, &bb.lvkind, &bb.sort_a, &bb.sort_b, &bb.key_a, &bb.key_b};
        for (int i = 0; i < 6; i++) {
            if ((e = d_br_[i].reserve(((size_t)n_br + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(brackets)", err);
            *slots[i] = d_br_[i].as<uint32_t>();
        }
        // match / parent / row / ct share one block: launch_brackets fills all four with 0xff at once
        const size_t stride = ((size_t)n_br + 4) & ~size_t(3);
        if ((e = d_br_[6].reserve(4 * stride * 4)) != cudaSuccess) return fail(e, "cudaMalloc(brackets)", err);
        bb.match = d_br_[6].as<uint32_t>();
        bb.parent = bb.match + stride;
        bb.row = bb.parent + stride;
        bb.ct = bb.row + stride;
    }
    if ((e = d_scratch_.reserve(parse_scan_scratch_words(std::max(n_br, (n_tok + 31) / 32)) * 8 + 64)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
    if ((e = timed(K_BRACKETS, [&] { return launch_brackets(tb, n_tok, tiles, bb, n_br, stream_); })) != cudaSuccess)
        return fail(e, "brackets", err);
    bool in_a = true;
    if ((e = timed(K_SORT_BRACKETS, [&] {
             return launch_sort_brackets(bb, n_br, out->max_level, d_scratch_.as<unsigned long long>(), &in_a, stream_);
         })) != cudaSuccess)
        return fail(e, "sort_brackets", err);
    if ((e = timed(K_MATCH, [&] { return launch_match(bb, in_a ? bb.sort_a : bb.sort_b, in_a ? bb.key_a : bb.key_b, n_br, d_res, stream_); })) != cudaSuccess)
        return fail(e, "match", err);

    // ---- strings with escapes or non-ASCII bytes are decoded into a side arena behind the text
    DecodeBuffers db{};
    uint32_t n_dec = 0;
    if (out->n_decode) {
        const uint32_t words = (n_tok + 31) / 32;
        if ((e = d_dec_counts_.reserve(((size_t)words + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(decode)", err);
        if ((e = d_dec_offs_.reserve(((size_t)words + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(decode)", err);
        if ((e = d_dec_tok_.reserve(((size_t)n_tok / 2 + 2) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(decode)", err);
        db.tok = d_dec_tok_.as<uint32_t>();
        if (!h_words_.reserve(4)) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
        unsigned long long *scratch = d_scratch_.as<unsigned long long>();
        if ((e = timed(K_DECODE, [&] { return launch_decode_list(tb, n_tok, db, d_dec_counts_.as<uint32_t>(), d_dec_offs_.as<uint32_t>(), scratch, stream_); })) != cudaSuccess)
            return fail(e, "decode_list", err);
        if ((e = copy_d2h(h_words_.p, scratch + 1, 8)) != cudaSuccess) return fail(e, "copy decode count", err);
        if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "decode_list", err);
        n_dec = (uint32_t)h_words_.p[0];
        if ((e = d_dec_off_.reserve(((size_t)n_dec + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(decode)", err);
        if ((e = d_dec_len_.reserve(((size_t)n_dec + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(decode)", err);
        db.off = d_dec_off_.as<uint32_t>(); db.len = d_dec_len_.as<uint32_t>();
        if ((e = d_scratch_.reserve(parse_scan_scratch_words(n_dec) * 8 + 64)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
        scratch = d_scratch_.as<unsigned long long>();
        if ((e = timed(K_DECODE, [&] { return launch_decode_lengths(raw, tb, db, n_dec, scratch, stream_); })) != cudaSuccess)
            return fail(e, "decode_lengths", err);
        if ((e = copy_d2h(h_words_.p, scratch + 1, 8)) != cudaSuccess) return fail(e, "copy decode size", err);
        if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "decode_lengths", err);
        const size_t side = (size_t)h_words_.p[0];
        const size_t side_base = ((size_t)p1_tiles(n) * P1_TILE + 64 + 15) & ~size_t(15);
        if (side_base + side >= 0xfffffff0ull) { *err = "document too large: text plus decoded strings must stay below 4 GiB"; return TK_ELIMIT; }
        if ((e = d_arena_.reserve(side_base + side + 32, side_base, stream_)) != cudaSuccess) return fail(e, "cudaMalloc(arena)", err);
        raw = d_arena_.as<uint8_t>();
        if ((e = timed(K_DECODE, [&] { return launch_decode_write(d_arena_.as<uint8_t>(), (uint32_t)side_base, tb, db, n_dec, stream_); })) != cudaSuccess)
            return fail(e, "decode_write", err);
        arena_used_ = side_base + side;
        out->side_bytes = side;
    }

    // ---- rows + grammar, then key order
    DevBuf *rows[5] = {&d_meta_, &d_koff_, &d_klen_, &d_voff_, &d_vlen_};
    for (int i = 0; i < 5; i++)
        if ((e = rows[i]->reserve(((size_t)n_rows + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(rows)", err);
    RowBuffers rb{d_meta_.as<uint32_t>(), d_koff_.as<uint32_t>(), d_klen_.as<uint32_t>(), d_voff_.as<uint32_t>(), d_vlen_.as<uint32_t>()};
    if ((e = timed(K_ROWS, [&] { return launch_rows(raw, n, tb, n_tok, tiles, bb, db, n_dec, rb, n_rows, d_res, stream_); })) != cudaSuccess)
        return fail(e, "rows", err);
    if (n_rows && (e = timed(K_KEY_ORDER, [&] { return launch_key_order(raw, rb, n_rows, bb, d_res, stream_); })) != cudaSuccess)
        return fail(e, "key_order", err);
    // module index lists (index.cu), while the verdict is on its way: sized by guess (a "name" /
    // "hostname" member per 256 bytes of text is four times what config #4 has); a list that outgrows
    // its buffer is counted, and build_index runs the kernel again with room for it
    tape_.n_rows = n_rows;
    tape_.arena_bytes = (uint32_t)arena_used_;
    tape_.text_len = (uint32_t)n;
    tape_.rows_direct = false;
    tape_.meta = rb.meta; tape_.koff = rb.koff; tape_.klen = rb.klen; tape_.voff = rb.voff; tape_.vlen = rb.vlen;
    tape_.arena = d_arena_.as<uint8_t>();
    for (int f = 0; f < 2; f++) {
        const size_t cap = std::min<size_t>(n_rows, n / 256 + 1024);
        if ((e = d_frefs_[f].reserve(cap * sizeof(FieldRef))) != cudaSuccess || (e = d_foff_[f].reserve((cap + 1) * 4)) != cudaSuccess)
            return fail(e, "cudaMalloc(index)", err);
        fx_.refs[f] = d_frefs_[f].as<FieldRef>();
        fx_.off[f] = d_foff_[f].as<uint32_t>();
        fx_.cap[f] = (uint32_t)cap;
    }
    if ((e = d_find_status_.reserve(field_status_words(n_rows) * 8)) != cudaSuccess) return fail(e, "cudaMalloc(index)", err);
    fx_.status = d_find_status_.as<unsigned long long>();
    if (n_rows && (e = timed(K_FIELD_FIND, [&] { return launch_field_find(tape_, fx_, d_res, stream_); })) != cudaSuccess)
        return fail(e, "field_find", err);
    tr.mark("rest:queued");
    rc = fetch_result(err);
    tr.mark("rest:done");
    if (rc) return rc;
    const unsigned long long side = out->side_bytes;
    *out = *h_res_.p;
    out->side_bytes = side;
    if (out->invalid) return TK_OK;

    if ((e = d_off_.reserve(((size_t)n_rows + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(off)", err);
    if ((e = d_scratch_.reserve(scan_scratch_words(n_rows) * 8)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
    tape_from_parse_ = true;
    bb_ = bb;
    parsed_ = *out;
    index_version_ = 0;
    return TK_OK;
}

int Engine::parse(const char *raw, size_t n, bool raw_on_device, Document *doc, std::string *err) {
    int rc = raw_on_device ? upload_text_device(raw, n, err) : upload_text(raw, n, err);
    if (rc) return rc;
    ParseResult r;
    rc = parse_resident(n, &r, err);
    if (rc) return rc;
    if (r.invalid) {
        // The GPU has established that json.Unmarshal fails; Go's wording of the FIRST error in
        // byte order comes from re-reading the text with the scanner restatement.
        std::string copy;
        if (raw_on_device) {
            copy.resize(n);
            cudaError_t e = cudaMemcpy(copy.data(), d_arena_.p, n, cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) return fail(e, "copy text back", err);
            raw = copy.data();
        }
        if (!syntax_error_text(std::string_view(raw, n), err))
            *err = "internal: device and host validators disagree (device says invalid)";
        return TK_ESYNTAX;
    }
    free_node(doc->root);
    doc->root = nullptr;
    doc->version = 1;
    doc->host_valid = false;
    staged_version_ = doc->version;  // the tape on the device IS this version of the document
    if (r.not_canonical || r.n_rows == 1) {
        // (a lone scalar root takes this path too: a nil root serialises as "{}", gabs.go:395-402)
        // keys out of order or duplicated, or numbers whose text must be re-formatted: the ordered
        // map insertions of the tree build put that right; Bytes() then flattens the tree
        rc = materialise(doc, err);
        if (rc) return rc;
        staged_version_ = 0;
    }
    return TK_OK;
}

// Builds tree nodes from tape rows (rows in input order).  Map insertion keeps the last duplicate
// (encoding/json objectInterface) and orders keys bytewise.  One builder per thread: the only state
// it writes is its own first range error.
struct TreeBuilder {
    const uint32_t *meta, *koff, *klen, *voff, *vlen, *close_of;   // close_of[r]: the closing row of opener r
    const char *arena;
    std::string range_err;

    void attach(Node *parent, Node *v, uint32_t r) const {
        if (parent->kind == Kind::Map) {
            auto &m = static_cast<MapNode *>(parent)->m;
            std::string_view key(arena + koff[r], klen[r]);
            auto hint = m.end();
            if (!m.empty() && std::prev(hint)->first < key) {
                m.emplace_hint(hint, std::string(key), v);  // ascending input: O(1) append
            } else {
                auto it = m.find(key);
                if (it != m.end()) { free_node(it->second); it->second = v; }
                else m.emplace(std::string(key), v);
            }
        } else {
            static_cast<ArrNode *>(parent)->items.push_back(v);
        }
    }

    Node *scalar(uint32_t r) {
        if ((meta[r] & 15u) == ROW_STR) return new StrNode(std::string(arena + voff[r], vlen[r]));
        std::string_view t(arena + voff[r], vlen[r]);
        if (t == "true") return new BoolNode(true);
        if (t == "false") return new BoolNode(false);
        if (t == "null") return nullptr;
        const std::string txt(t);
        double d = strtod_c(txt.c_str());
        if (std::isinf(d)) {
            if (range_err.empty()) range_err = "json: cannot unmarshal number " + txt + " into Go value of type float64";
            d = 0;
        }
        return new NumNode(d);
    }

    // the value that starts at row r; *next = the row after it.  Containers are walked with an
    // explicit stack (nesting may be 10,000 deep).
    Node *value(uint32_t r, uint32_t *next) {
        const uint32_t kind = meta[r] & 15u;
        if (kind == ROW_EMPTY_OBJ) { *next = r + 1; return new MapNode(); }
        if (kind == ROW_EMPTY_ARR) { *next = r + 1; return new ArrNode(); }
        if (kind != ROW_OPEN_OBJ && kind != ROW_OPEN_ARR) { *next = r + 1; return scalar(r); }
        const uint32_t end = close_of[r];
        Node *top = kind == ROW_OPEN_OBJ ? static_cast<Node *>(new MapNode()) : static_cast<Node *>(new ArrNode());
        std::vector<Node *> stack{top};
        for (uint32_t i = r + 1; i < end; i++) {
            const uint32_t k = meta[i] & 15u;
            if (k == ROW_CLOSE_OBJ || k == ROW_CLOSE_ARR) { stack.pop_back(); continue; }
            Node *v;
            if (k == ROW_OPEN_OBJ) v = new MapNode();
            else if (k == ROW_OPEN_ARR) v = new ArrNode();
            else if (k == ROW_EMPTY_OBJ) v = new MapNode();
            else if (k == ROW_EMPTY_ARR) v = new ArrNode();
            else v = scalar(i);
            attach(stack.back(), v, i);
            if (k == ROW_OPEN_OBJ || k == ROW_OPEN_ARR) stack.push_back(v);
        }
        *next = end + 1;
        return top;
    }
};

// A container that spans very many rows.  With many members (the "module" of a large state
// document) they are built by several threads, each with its own builder, and attached in order;
// with few (the root above it) each member is looked at the same way in turn.
constexpr uint32_t FORK_MIN_ROWS = 65536, FORK_MIN_MEMBERS_BUILD = 1024;

static Node *build_value(TreeBuilder &tb, uint32_t r, uint32_t *next, int levels) {
    const uint32_t kind = tb.meta[r] & 15u;
    if ((kind != ROW_OPEN_OBJ && kind != ROW_OPEN_ARR) || levels == 0 || tb.close_of[r] - r < FORK_MIN_ROWS) return tb.value(r, next);
    const uint32_t end = tb.close_of[r];
    std::vector<uint32_t> starts;   // first row of every member
    for (uint32_t i = r + 1; i < end;) {
        starts.push_back(i);
        const uint32_t k = tb.meta[i] & 15u;
        i = (k == ROW_OPEN_OBJ || k == ROW_OPEN_ARR) ? tb.close_of[i] + 1 : i + 1;
    }
    Node *top = kind == ROW_OPEN_OBJ ? static_cast<Node *>(new MapNode()) : static_cast<Node *>(new ArrNode());
    const size_t n = starts.size();
    std::vector<Node *> nodes(n, nullptr);
    if (n < FORK_MIN_MEMBERS_BUILD) {
        for (size_t i = 0; i < n; i++) { uint32_t nx; nodes[i] = build_value(tb, starts[i], &nx, levels - 1); }
    } else {
        const size_t hw = std::max<size_t>(1, std::thread::hardware_concurrency());
        const size_t workers = std::max<size_t>(2, std::min<size_t>({hw, 16, n / 256}));
        std::vector<TreeBuilder> builders(workers, tb);
        std::vector<std::thread> threads;
        std::atomic<bool> failed{false};
        for (size_t w = 0; w < workers; w++) {
            builders[w].range_err.clear();
            threads.emplace_back([&, w] {
                try {
                    for (size_t i = w * n / workers; i < (w + 1) * n / workers; i++) { uint32_t nx; nodes[i] = builders[w].value(starts[i], &nx); }
                } catch (...) {
                    failed = true;
                }
            });
        }
        for (auto &t : threads) t.join();
        if (failed) throw std::bad_alloc();
        for (size_t w = 0; w < workers; w++)   // the first error in row order, as the single-threaded walk reports it
            if (tb.range_err.empty() && !builders[w].range_err.empty()) tb.range_err = builders[w].range_err;
    }
    for (size_t i = 0; i < n; i++) tb.attach(top, nodes[i], starts[i]);
    *next = end + 1;
    return top;
}

// Rows arrive in input order; a stack of open containers rebuilds the tree.  Map insertion keeps
// the last duplicate (encoding/json objectInterface) and orders keys bytewise.
int Engine::materialise(Document *doc, std::string *err) {
    if (doc->host_valid) return TK_OK;
    StreamScope scope(stream_);
    if (edit_) { int rc = apply_edits(*doc, err); if (rc) return rc; }   // the tree is built from the tape as edited
    if (!tape_from_parse_ || !tape_.n_rows) { *err = "internal: no device tape to build the tree from"; return TK_EARG; }
    const size_t n = tape_.n_rows;
    cudaError_t e;
    PinVec<uint32_t> *hrows[5] = {&h_meta_, &h_koff_, &h_klen_, &h_voff_, &h_vlen_};
    const uint32_t *drows[5] = {tape_.meta, tape_.koff, tape_.klen, tape_.voff, tape_.vlen};
    for (int i = 0; i < 5; i++) {
        if (!hrows[i]->reserve(n)) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
        if ((e = copy_d2h(hrows[i]->p, drows[i], n * 4)) != cudaSuccess) return fail(e, "copy rows", err);
    }
    if (!h_arena_.reserve(arena_used_ + 1)) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
    if ((e = copy_d2h(h_arena_.p, d_arena_.p, arena_used_)) != cudaSuccess) return fail(e, "copy arena", err);
    if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "copy tape", err);
    for (int i = 0; i < 5; i++) hrows[i]->n = 0;  // scratch use only: the flattener starts from empty
    h_arena_.n = 0;

    // the closing row of every opener, then the tree (wide containers by several threads)
    const char *arena = reinterpret_cast<const char *>(h_arena_.p);
    std::vector<uint32_t> close_of(n, 0);
    {
        std::vector<uint32_t> open;
        for (uint32_t r = 0; r < n; r++) {
            const uint32_t k = h_meta_.p[r] & 15u;
            if (k == ROW_OPEN_OBJ || k == ROW_OPEN_ARR) open.push_back(r);
            else if ((k == ROW_CLOSE_OBJ || k == ROW_CLOSE_ARR) && !open.empty()) { close_of[open.back()] = r; open.pop_back(); }
        }
    }
    TreeBuilder tb{h_meta_.p, h_koff_.p, h_klen_.p, h_voff_.p, h_vlen_.p, close_of.data(), arena, {}};
    uint32_t next = 0;
    Node *root = build_value(tb, 0, &next, 3);
    const std::string &range_err = tb.range_err;
    if (!range_err.empty()) {
        free_node(root);
        *err = range_err;
        return TK_ESYNTAX;
    }
    free_node(doc->root);
    doc->root = root;
    doc->host_valid = true;
    adopt_added_values(doc);
    return TK_OK;
}

}  // namespace tk

// ============================================================================ module index

namespace tk {

// Once per document version: the field lists the parse left on the device (rows_kernel) are packed
// into result records (csrc/index.cu) and copied to the host with their offsets and order bitmaps.
// The sizes of everything are known from the parse verdict, so the host waits once, at the end.
int Engine::build_index(const Document &doc, std::string *err) {
    if (index_version_ == doc.version || edit_) return TK_OK;   // (once edits begin, the index is what was built then plus the edit list)
    StreamScope scope(stream_);
    if (!device_resident(doc)) { *err = "internal: module index needs the parser's tape"; return TK_EARG; }
    cudaError_t e;
    index_module_ok_ = parsed_.root_kind == ROW_OPEN_OBJ && parsed_.module_row != 0 &&
                       (parsed_.module_kind == ROW_OPEN_OBJ || parsed_.module_kind == ROW_EMPTY_OBJ);
    index_m_[0] = index_m_[1] = 0;
    index_hint_[0] = index_hint_[1] = 0;
    if (parsed_.index_slow) { *err = "internal: module index not available for this document"; return TK_EARG; }
    if (index_module_ok_ && (parsed_.field_count[0] > fx_.cap[0] || parsed_.field_count[1] > fx_.cap[1])) {
        // a list outgrew its buffer: the kernel has counted it, run it again with room (same tape, same verdict)
        for (int f = 0; f < 2; f++) {
            const size_t cap = std::max<size_t>(parsed_.field_count[f], fx_.cap[f]);
            if ((e = d_frefs_[f].reserve(cap * sizeof(FieldRef))) != cudaSuccess || (e = d_foff_[f].reserve((cap + 1) * 4)) != cudaSuccess)
                return fail(e, "cudaMalloc(index)", err);
            fx_.refs[f] = d_frefs_[f].as<FieldRef>();
            fx_.off[f] = d_foff_[f].as<uint32_t>();
            fx_.cap[f] = (uint32_t)cap;
        }
        e = timed(K_FIELD_FIND, [&] { return launch_field_find(tape_, fx_, d_res_.as<ParseResult>(), stream_); });
        if (e != cudaSuccess) return fail(e, "field_find", err);
    }
    uint32_t m[2] = {0, 0};
    size_t bytes[2] = {0, 0}, bit_words[2] = {0, 0};
    uint32_t *viol[2] = {nullptr, nullptr};
    uint8_t *blob[2] = {nullptr, nullptr};
    for (int f = 0; index_module_ok_ && f < 2; f++) {
        m[f] = parsed_.field_count[f];
        if (!m[f]) continue;
        if (parsed_.field_bytes[f] >= 0xfffffff0ull) { *err = "document too large: the module index must stay below 4 GiB"; return TK_ELIMIT; }
        bytes[f] = (size_t)parsed_.field_bytes[f];
        bit_words[f] = ((size_t)m[f] + 31) / 32 + 1;
        if ((e = d_fviolbits_[f].reserve(bit_words[f] * 4)) != cudaSuccess || (e = d_fblob_[f].reserve(bytes[f] + 64)) != cudaSuccess)
            return fail(e, "cudaMalloc(index)", err);
        if (!h_foff_[f].reserve((size_t)m[f] + 1) || !h_fviolbits_[f].reserve(bit_words[f]) || !h_fblob_[f].reserve(bytes[f] + 1)) {
            *err = "cuda: pinned host allocation failed";
            return TK_ECUDA;
        }
        viol[f] = d_fviolbits_[f].as<uint32_t>();
        blob[f] = d_fblob_[f].as<uint8_t>();
    }
    if (m[0] || m[1]) {
        e = timed(K_MODULE_INDEX, [&] {
            cudaError_t x;
            for (int f = 0; f < 2; f++)
                if (m[f] && (x = cudaMemsetAsync(viol[f], 0, bit_words[f] * 4, stream_)) != cudaSuccess) return x;
            return launch_field_pack(tape_.arena, fx_, m, viol, blob, stream_);
        });
        if (e != cudaSuccess) return fail(e, "module index", err);
        for (int f = 0; f < 2; f++) {
            if (!m[f]) continue;
            if ((e = copy_d2h(h_foff_[f].p, fx_.off[f], ((size_t)m[f] + 1) * 4)) != cudaSuccess) return fail(e, "copy index", err);
            if ((e = copy_d2h(h_fviolbits_[f].p, viol[f], (((size_t)m[f] + 31) / 32) * 4)) != cudaSuccess) return fail(e, "copy index", err);
            if ((e = copy_d2h(h_fblob_[f].p, blob[f], bytes[f])) != cudaSuccess) return fail(e, "copy index", err);
            index_m_[f] = m[f];
        }
        if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "module index", err);
    }
    for (int f = 0; f < 2; f++)
        if (index_m_[f] && h_foff_[f].p[index_m_[f]] != (uint32_t)bytes[f]) {
            *err = "internal: module index sizes disagree";
            return TK_ECUDA;
        }
    index_version_ = doc.version;
    return TK_OK;
}

// Record j of field f (the j-th child, in key order, that has the field): { u32 field_len, field,
// u32 key_len, key } at blob + off[j].
struct Engine::IndexView {
    const uint32_t *off;
    const char *blob;
    uint32_t m;
    std::string_view key(uint32_t j) const {
        const char *rec = blob + off[j];
        uint32_t fl, kl;
        memcpy(&fl, rec, 4);
        memcpy(&kl, rec + 4 + fl, 4);
        return std::string_view(rec + 8 + fl, kl);
    }
    std::string_view field(uint32_t j) const {
        const char *rec = blob + off[j];
        uint32_t fl;
        memcpy(&fl, rec, 4);
        return std::string_view(rec + 4, fl);
    }
    bool has_prefix(uint32_t j, std::string_view prefix) const {
        const std::string_view k = key(j);
        return k.size() >= prefix.size() && memcmp(k.data(), prefix.data(), prefix.size()) == 0;
    }
};

// first and one-past-last record of field f whose key starts with prefix (records are in bytewise key order)
void Engine::index_range(int f, std::string_view prefix, uint32_t *lo_out, uint32_t *hi_out) {
    const IndexView v{h_foff_[f].p, reinterpret_cast<const char *>(h_fblob_[f].p), index_m_[f]};
    uint32_t lo = 0;
    const uint32_t hint = index_hint_[f];
    if (hint < v.m && v.has_prefix(hint, prefix) && (hint == 0 || v.key(hint - 1) < prefix)) {
        lo = hint;  // a sweep over clusters in order: this run starts where the previous one ended
    } else {
        uint32_t hi = v.m;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (v.key(mid) < prefix) lo = mid + 1; else hi = mid;
        }
    }
    // end of the run: gallop, then bisect (runs are short next to the index); keys with the prefix
    // are contiguous, so has_prefix is true ... true false ... from lo on
    uint32_t a = lo;
    if (lo < v.m && v.has_prefix(lo, prefix)) {
        uint32_t last_good = lo, step = 1, bad = v.m;
        for (;;) {
            const uint64_t probe = (uint64_t)last_good + step;
            if (probe >= v.m) break;
            if (v.has_prefix((uint32_t)probe, prefix)) { last_good = (uint32_t)probe; step <<= 1; }
            else { bad = (uint32_t)probe; break; }
        }
        a = last_good + 1;
        uint32_t b = bad;
        while (a < b) {
            const uint32_t mid = (a + b) >> 1;
            if (v.has_prefix(mid, prefix)) a = mid + 1; else b = mid;
        }
    }
    *lo_out = lo;
    *hi_out = a;
    index_hint_[f] = a;
}

// the run's (field value, key) pairs, for the caller to sort: used when the run is not in field order
int Engine::index_lookup(const Document &doc, std::string_view prefix, bool hostname, bool *module_ok,
                         std::vector<std::pair<std::string_view, std::string_view>> *pairs, std::string *err) {
    int rc = build_index(doc, err);
    if (rc) return rc;
    *module_ok = index_module_ok_;
    pairs->clear();
    const int f = hostname ? 1 : 0;
    if (!index_module_ok_) return TK_OK;
    if (index_m_[f] == 0) { edit_added_pairs(f, prefix, pairs); return TK_OK; }
    uint32_t lo, hi;
    index_range(f, prefix, &lo, &hi);
    const IndexView v{h_foff_[f].p, reinterpret_cast<const char *>(h_fblob_[f].p), index_m_[f]};
    for (uint32_t j = lo; j < hi; j++)
        if (!edit_record_dead(f, j)) pairs->emplace_back(v.field(j), v.key(j));
    edit_added_pairs(f, prefix, pairs);
    return TK_OK;
}

// The same query answered without copying when possible: the finished C-ABI result is a byte range
// of the field's record blob.
int Engine::index_lookup_direct(const Document &doc, std::string_view prefix, bool hostname, bool *module_ok, bool *direct,
                                const uint8_t **blob, size_t *blob_len, size_t *count, std::string *err) {
    int rc = build_index(doc, err);
    if (rc) return rc;
    *module_ok = index_module_ok_;
    *direct = true;
    *blob = reinterpret_cast<const uint8_t *>("");
    *blob_len = 0;
    *count = 0;
    const int f = hostname ? 1 : 0;
    if (!index_module_ok_) return TK_OK;
    if (edit_added_any(f, prefix)) { *direct = false; return TK_OK; }   // children added since: the pairs are put together one by one
    if (index_m_[f] == 0) return TK_OK;
    uint32_t lo, hi;
    index_range(f, prefix, &lo, &hi);
    if (lo == hi) return TK_OK;
    for (uint32_t j = lo; edit_ && j < hi; j++)
        if (edit_record_dead(f, j)) { *direct = false; return TK_OK; }     // children deleted since
    // order violations strictly inside the run (record lo may compare with a child outside it):
    // any bit set among records lo + 1 .. hi - 1
    const uint32_t *bits = h_fviolbits_[f].p;
    for (uint32_t j = lo + 1; j < hi;) {
        const uint32_t w = j >> 5, first = j & 31, last = std::min<uint32_t>(32, hi - (w << 5));   // bits [first, last) of word w
        uint32_t mask = last >= 32 ? 0xffffffffu : ((1u << last) - 1);
        mask &= ~((1u << first) - 1);
        if (bits[w] & mask) { *direct = false; return TK_OK; }
        j = (w + 1) << 5;
    }
    const uint32_t *off = h_foff_[f].p;
    *blob = h_fblob_[f].p + off[lo];
    *blob_len = off[hi] - off[lo];
    *count = hi - lo;
    return TK_OK;
}

}  // namespace tk

// ============================================================================ edits on the device

namespace tk {

// What the host knows about module's children while a document is edited on the device.
struct Engine::EditState {
    struct Base { uint32_t key_off, key_len, rows, tape_row; uint32_t rec[2]; bool live; };   // a child the parser found; rec: its index records (or none)
    struct Added {                                                               // a child set since
        Node *value = nullptr;      // as given (a record stays a record: opaque, members in the given order)
        bool on_device = false;
        uint32_t rows = 0, tape_row = 0;
        bool has[2] = {false, false};   // a map with a string "name" / "hostname": visible to Clusters() / Nodes()
        std::string field[2];
    };
    std::vector<Base> base;         // in key order
    const char *keys = nullptr;     // their keys, one after another (h_ckeys_)
    std::map<std::string, Added, std::less<>> added;
    std::vector<uint8_t> dead[2];   // per field: index records whose child has been deleted or replaced
    uint32_t n_rows = 0;            // rows of the device tape as last rebuilt
    uint32_t head_rows = 0;         // [0, head_rows): everything before module's first child
    uint32_t tail_row = 0;          // [tail_row, n_rows): module's closing row and everything after it
    size_t live = 0;                // children alive, found and added
    // the child whose last row carries no comma on the device tape: the last in key order when it was last rebuilt
    bool last_is_added = false;
    size_t last_base = 0;
    std::string last_added;
    bool dirty = false;             // edited since the tape was last rebuilt
    ~EditState() { for (auto &kv : added) free_node(kv.second.value); }
    std::string_view key(const Base &b) const { return std::string_view(keys + b.key_off, b.key_len); }
    Base *find_base(std::string_view k) {
        size_t lo = 0, hi = base.size();
        while (lo < hi) { const size_t mid = (lo + hi) >> 1; if (key(base[mid]) < k) lo = mid + 1; else hi = mid; }
        return lo < base.size() && key(base[lo]) == k ? &base[lo] : nullptr;
    }
};

void Engine::drop_edit_state() { delete edit_; edit_ = nullptr; }

bool Engine::edit_record_dead(int f, uint32_t j) const { return edit_ && j < edit_->dead[f].size() && edit_->dead[f][j]; }

bool Engine::edit_added_any(int f, std::string_view prefix) const {
    if (!edit_) return false;
    for (auto it = edit_->added.lower_bound(prefix); it != edit_->added.end(); ++it) {
        if (it->first.size() < prefix.size() || it->first.compare(0, prefix.size(), prefix) != 0) break;
        if (it->second.has[f]) return true;
    }
    return false;
}

void Engine::edit_added_pairs(int f, std::string_view prefix, std::vector<std::pair<std::string_view, std::string_view>> *pairs) const {
    if (!edit_) return;
    for (auto it = edit_->added.lower_bound(prefix); it != edit_->added.end(); ++it) {
        if (it->first.size() < prefix.size() || it->first.compare(0, prefix.size(), prefix) != 0) break;
        if (it->second.has[f]) pairs->emplace_back(it->second.field[f], it->first);
    }
}

// After the tree has been built from the edited tape: the children set since take their original
// values back (the tape knows objects only; a record must stay opaque and keep its member order
// until the document is parsed again -- SURVEY.md 8a note 1), and the edit list is done with.
void Engine::adopt_added_values(Document *doc) {
    if (!edit_) return;
    MapNode *mod = nullptr;
    if (doc->root && doc->root->kind == Kind::Map) {
        auto &m = static_cast<MapNode *>(doc->root)->m;
        auto it = m.find(std::string_view("module"));
        if (it != m.end() && it->second && it->second->kind == Kind::Map) mod = static_cast<MapNode *>(it->second);
    }
    for (auto &kv : edit_->added) {
        if (!mod) break;
        auto it = mod->m.find(kv.first);
        if (it == mod->m.end()) continue;
        free_node(it->second);
        it->second = kv.second.value;
        kv.second.value = nullptr;
    }
    drop_edit_state();
}

// "module.<key>" on a document the device path can edit: *key set, true.  Anything else: false
// (*rc != 0: a CUDA failure on the way into edit mode).
bool Engine::edit_path_key(const Document &doc, std::string_view path, std::string_view *key, std::string *err, int *rc) {
    *rc = TK_OK;
    if (doc.host_valid) return false;
    const size_t dot = path.find('.');
    if (dot == std::string_view::npos || path.substr(0, dot) != "module" || path.find('.', dot + 1) != std::string_view::npos) return false;
    if (!edit_) {
        if (!index_usable(doc)) return false;
        if (!(parsed_.root_kind == ROW_OPEN_OBJ && parsed_.module_row != 0 && parsed_.module_kind == ROW_OPEN_OBJ && parsed_.module_end != 0)) return false;
        *rc = enter_edit_mode(doc, err);
        if (*rc || !edit_) return false;
    }
    *key = path.substr(dot + 1);
    return true;
}

// The list of module's children, once, from the tape as parsed: flags -> ranks -> (first row, last
// row, key) per child -> keys packed -> host.
int Engine::enter_edit_mode(const Document &doc, std::string *err) {
    StreamScope scope(stream_);
    int rc = build_index(doc, err);
    if (rc) return rc;
    cudaError_t e;
    const uint32_t n = tape_.n_rows, module_row = parsed_.module_row - 1, module_end = parsed_.module_end;
    if (!h_words_.reserve(8)) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
    if ((e = d_cflags_.reserve((size_t)n * 4)) != cudaSuccess || (e = d_crank_.reserve((size_t)n * 4)) != cudaSuccess ||
        (e = d_scratch_.reserve(parse_scan_scratch_words(n) * 8 + 64)) != cudaSuccess)
        return fail(e, "cudaMalloc(children)", err);
    unsigned long long *scratch = d_scratch_.as<unsigned long long>();
    if ((e = launch_child_flags(tape_, module_row, module_end, d_cflags_.as<uint32_t>(), stream_)) != cudaSuccess ||
        (e = scan_u32(d_cflags_.as<uint32_t>(), d_crank_.as<uint32_t>(), n, scratch, stream_)) != cudaSuccess ||
        (e = copy_d2h(h_words_.p, scratch + 1, 8)) != cudaSuccess || (e = cudaStreamSynchronize(stream_)) != cudaSuccess)
        return fail(e, "child list", err);
    const uint32_t nc = (uint32_t)h_words_.p[0];
    if (nc == 0) return TK_OK;   // (cannot be for a non-empty object; the caller then builds the host tree)
    for (int i = 0; i < 4; i++)
        if ((e = d_clist_[i].reserve((size_t)nc * 4)) != cudaSuccess) return fail(e, "cudaMalloc(children)", err);
    if ((e = d_ckeyoff_.reserve(((size_t)nc + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(children)", err);
    ChildList cl{d_clist_[0].as<uint32_t>(), d_clist_[1].as<uint32_t>(), d_clist_[2].as<uint32_t>(), d_clist_[3].as<uint32_t>()};
    if ((e = launch_child_entries(tape_, d_cflags_.as<uint32_t>(), d_crank_.as<uint32_t>(), bb_.match, bb_.row, cl, stream_)) != cudaSuccess ||
        (e = scan_u32(cl.klen, d_ckeyoff_.as<uint32_t>(), nc, scratch, stream_)) != cudaSuccess ||
        (e = copy_d2h(h_words_.p, scratch + 1, 8)) != cudaSuccess || (e = cudaStreamSynchronize(stream_)) != cudaSuccess)
        return fail(e, "child list", err);
    const size_t key_bytes = (size_t)h_words_.p[0];
    if ((e = d_ckeys_.reserve(key_bytes + 32)) != cudaSuccess) return fail(e, "cudaMalloc(children)", err);
    if (!h_clist_[0].reserve(nc) || !h_clist_[1].reserve(nc) || !h_clist_[2].reserve((size_t)nc + 1) || !h_ckeys_.reserve(key_bytes + 1)) {
        *err = "cuda: pinned host allocation failed";
        return TK_ECUDA;
    }
    if ((e = launch_child_keys(tape_.arena, cl, d_ckeyoff_.as<uint32_t>(), nc, d_ckeys_.as<uint8_t>(), stream_)) != cudaSuccess ||
        (e = copy_d2h(h_clist_[0].p, cl.row, (size_t)nc * 4)) != cudaSuccess || (e = copy_d2h(h_clist_[1].p, cl.end, (size_t)nc * 4)) != cudaSuccess ||
        (e = copy_d2h(h_clist_[2].p, d_ckeyoff_.p, (size_t)nc * 4)) != cudaSuccess ||
        (key_bytes && (e = copy_d2h(h_ckeys_.p, d_ckeys_.p, key_bytes)) != cudaSuccess) || (e = cudaStreamSynchronize(stream_)) != cudaSuccess)
        return fail(e, "child list", err);
    h_clist_[2].p[nc] = (uint32_t)key_bytes;
    auto *E = new EditState();
    E->keys = reinterpret_cast<const char *>(h_ckeys_.p);
    E->base.resize(nc);
    for (uint32_t j = 0; j < nc; j++) {
        const uint32_t row = h_clist_[0].p[j], end = h_clist_[1].p[j];
        E->base[j] = EditState::Base{h_clist_[2].p[j], h_clist_[2].p[j + 1] - h_clist_[2].p[j], end - row + 1, row, {0xffffffffu, 0xffffffffu}, true};
    }
    // which index records belong to which child: both lists are in key order, one walk each
    for (int f = 0; f < 2; f++) {
        const IndexView v{h_foff_[f].p, reinterpret_cast<const char *>(h_fblob_[f].p), index_m_[f]};
        uint32_t c = 0;
        for (uint32_t j = 0; j < v.m; j++) {
            const std::string_view k = v.key(j);
            while (c < nc && E->key(E->base[c]) < k) c++;
            if (c < nc && E->key(E->base[c]) == k) E->base[c].rec[f] = j;
        }
    }
    E->n_rows = n;
    E->head_rows = module_row + 1;
    E->tail_row = module_end;
    E->live = nc;
    E->last_base = nc - 1;
    for (int f = 0; f < 2; f++) E->dead[f].assign(index_m_[f], 0);
    edit_ = E;
    return TK_OK;
}

int Engine::edit_delete(Document *doc, std::string_view path, bool *handled, std::string *err) {
    *handled = false;
    std::string_view key;
    int rc;
    if (!edit_path_key(*doc, path, &key, err, &rc)) return rc;
    EditState &E = *edit_;
    auto it = E.added.find(key);
    EditState::Base *b = it == E.added.end() ? E.find_base(key) : nullptr;
    const bool present = it != E.added.end() || (b && b->live);
    if (present && E.live == 1) return TK_OK;   // module would become empty ("{}" is one row, not two): left to the host tree
    *handled = true;
    if (!present) return TK_OK;                  // gabs: deleting an absent key succeeds and changes nothing
    if (it != E.added.end()) {
        free_node(it->second.value);
        it->second.value = nullptr;
        E.added.erase(it);
    } else {
        b->live = false;
        for (int f = 0; f < 2; f++)
            if (b->rec[f] != 0xffffffffu) E.dead[f][b->rec[f]] = 1;
    }
    E.live--;
    E.dirty = true;
    doc->version++;
    return TK_OK;
}

int Engine::edit_set(Document *doc, std::string_view path, Node *value, bool *handled, std::string *err) {
    *handled = false;
    std::string_view key;
    int rc;
    if (!edit_path_key(*doc, path, &key, err, &rc)) return rc;
    EditState &E = *edit_;
    EditState::Added a;   // (its rows are made when the tape is next rebuilt)
    if (value && value->kind == Kind::Map) {
        auto &m = static_cast<const MapNode *>(value)->m;
        static const char *const names[2] = {"name", "hostname"};
        for (int f = 0; f < 2; f++) {
            auto fi = m.find(std::string_view(names[f]));
            if (fi != m.end() && fi->second && fi->second->kind == Kind::Str) { a.has[f] = true; a.field[f] = static_cast<const StrNode *>(fi->second)->s; }
        }
    }
    a.value = value;
    *handled = true;
    auto it = E.added.find(key);
    if (it != E.added.end()) {
        free_node(it->second.value);
        it->second = std::move(a);
    } else {
        if (EditState::Base *b = E.find_base(key); b && b->live) {
            b->live = false;
            for (int f = 0; f < 2; f++)
                if (b->rec[f] != 0xffffffffu) E.dead[f][b->rec[f]] = 1;
        } else {
            E.live++;
        }
        E.added.emplace(std::string(key), std::move(a));
    }
    E.dirty = true;
    doc->version++;
    return TK_OK;
}

// The device tape rebuilt to match the child list: runs of surviving rows and the rows of the
// children set since (uploaded behind the old rows first), in key order, gathered into the second
// set of row arrays; then the comma after the last child taken off and the one after the child
// that used to be last put back.
int Engine::apply_edits(const Document &doc, std::string *err) {
    if (!edit_) return TK_OK;
    EditState &E = *edit_;
    if (!E.dirty) { staged_version_ = doc.version; return TK_OK; }
    StreamScope scope(stream_);
    cudaError_t e;
    // (the pinned staging below may still be in flight from the previous rebuild)
    if ((e = cudaStreamSynchronize(stream_)) != cudaSuccess) return fail(e, "synchronize", err);
    h_segs_.clear(); h_addarena_.clear();
    for (auto &v : h_addrows_) v.clear();
    // the rows of the children set since the last rebuild: flattened (many of them: by several threads)
    // straight into the pinned upload buffers, one child after another in key order
    const size_t side_base = (arena_used_ + 15) & ~size_t(15);
    {
        std::vector<MemberRef> fresh;
        std::vector<EditState::Added *> fresh_added;
        for (auto &kv : E.added)
            if (!kv.second.on_device) { fresh.emplace_back(&kv.first, kv.second.value); fresh_added.push_back(&kv.second); }
        if (!fresh.empty()) {
            FlatPinned sink{h_addrows_[0], h_addrows_[1], h_addrows_[2], h_addrows_[3], h_addrows_[4], h_addarena_};
            try {
                flat_pairs_forked(sink, fresh, 2, true);   // children of module: depth 2, a comma after each (taken off the last one below)
            } catch (...) {
                *err = "out of memory";
                return TK_ELIMIT;
            }
            // how many rows each child took: a child ends where depth 2 is reached again
            const uint32_t *meta = h_addrows_[0].p;
            const size_t total = h_addrows_[0].n;
            size_t r = 0;
            for (EditState::Added *a : fresh_added) {
                const size_t start = r;
                const uint32_t kind = meta[r] & 15u;
                r++;
                if (kind == ROW_OPEN_OBJ || kind == ROW_OPEN_ARR) {
                    while (r < total && !((meta[r] >> META_DEPTH_SHIFT) == 2 && ((meta[r] & 15u) == ROW_CLOSE_OBJ || (meta[r] & 15u) == ROW_CLOSE_ARR))) r++;
                    r++;   // its closing row
                }
                a->rows = (uint32_t)(r - start);
            }
            // (FlatPinned keeps offsets relative to its own arena, the second and fourth array: moved behind the device arena)
            uint32_t *ko = h_addrows_[1].p, *kl = h_addrows_[2].p, *vo = h_addrows_[3].p, *vl = h_addrows_[4].p;
            for (size_t i = 0; i < total; i++) {
                ko[i] = kl[i] ? ko[i] + (uint32_t)side_base : 0u;
                vo[i] = vl[i] ? vo[i] + (uint32_t)side_base : 0u;
            }
        }
    }
    uint32_t cur = 0, upload_rows = 0;
    auto add_seg = [&](uint32_t src, uint32_t n) {
        if (!n) return;
        if (h_segs_.n && h_segs_.p[h_segs_.n - 1].src + h_segs_.p[h_segs_.n - 1].n == src) h_segs_.p[h_segs_.n - 1].n += n;
        else h_segs_.push(RowSegment{cur, src, n});
        cur += n;
    };
    add_seg(0, E.head_rows);
    // what used to be last, by new row number (0xffffffff: gone), and what is last now
    uint32_t old_last_row = 0xffffffffu, new_last_row = 0xffffffffu;
    bool new_last_is_added = false;
    size_t new_last_base = 0;
    const std::string *new_last_added = nullptr;
    size_t bi = 0;
    auto ai = E.added.begin();
    while (bi < E.base.size() || ai != E.added.end()) {
        while (bi < E.base.size() && !E.base[bi].live) bi++;
        const bool take_base = bi < E.base.size() && (ai == E.added.end() || E.key(E.base[bi]) < std::string_view(ai->first));
        if (take_base) {
            EditState::Base &b = E.base[bi];
            const uint32_t at = cur;
            add_seg(b.tape_row, b.rows);
            b.tape_row = at;
            if (!E.last_is_added && E.last_base == bi) old_last_row = cur - 1;
            new_last_row = cur - 1; new_last_is_added = false; new_last_base = bi;
            bi++;
        } else if (ai != E.added.end()) {
            EditState::Added &a = ai->second;
            const uint32_t at = cur;
            if (a.on_device) {
                add_seg(a.tape_row, a.rows);
            } else {
                // its rows were flattened above, in this order: they go behind the old rows, its strings behind the arena
                add_seg(E.n_rows + upload_rows, a.rows);
                upload_rows += a.rows;
                a.on_device = true;
            }
            a.tape_row = at;
            if (E.last_is_added && E.last_added == ai->first) old_last_row = cur - 1;
            new_last_row = cur - 1; new_last_is_added = true; new_last_added = &ai->first;
            ++ai;
        }
    }
    const uint32_t new_tail = cur;
    add_seg(E.tail_row, E.n_rows - E.tail_row);
    const uint32_t new_n = cur;
    bool pin_ok = !h_segs_.failed && !h_addarena_.failed;
    for (auto &v : h_addrows_) pin_ok = pin_ok && !v.failed;
    if (!pin_ok) { *err = "cuda: pinned host allocation failed"; return TK_ECUDA; }
    if (side_base + h_addarena_.n >= 0xfffffff0ull || (size_t)E.n_rows + upload_rows >= 0xfffffff0ull) {
        *err = "document too large: string arena and row count must stay below 4 GiB";
        return TK_ELIMIT;
    }
    // new rows and their strings
    DevBuf *rows[5] = {&d_meta_, &d_koff_, &d_klen_, &d_voff_, &d_vlen_};
    if (upload_rows) {
        for (int i = 0; i < 5; i++) {
            if ((e = rows[i]->reserve(((size_t)E.n_rows + upload_rows) * 4, (size_t)E.n_rows * 4, stream_)) != cudaSuccess) return fail(e, "cudaMalloc(rows)", err);
            if ((e = copy_h2d(rows[i]->as<uint32_t>() + E.n_rows, h_addrows_[i].p, (size_t)upload_rows * 4)) != cudaSuccess) return fail(e, "copy rows", err);
        }
        if ((e = d_arena_.reserve(side_base + h_addarena_.n + 32, arena_used_, stream_)) != cudaSuccess) return fail(e, "cudaMalloc(arena)", err);
        if (h_addarena_.n && (e = copy_h2d(d_arena_.as<uint8_t>() + side_base, h_addarena_.p, h_addarena_.n)) != cudaSuccess) return fail(e, "copy arena", err);
        arena_used_ = side_base + h_addarena_.n;
    }
    // gather
    if ((e = d_segs_.reserve(h_segs_.n * sizeof(RowSegment))) != cudaSuccess) return fail(e, "cudaMalloc(segments)", err);
    if ((e = copy_h2d(d_segs_.p, h_segs_.p, h_segs_.n * sizeof(RowSegment))) != cudaSuccess) return fail(e, "copy segments", err);
    for (int i = 0; i < 5; i++)
        if ((e = d_rows2_[i].reserve(((size_t)new_n + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(rows)", err);
    const RowArrays from{d_meta_.as<uint32_t>(), d_koff_.as<uint32_t>(), d_klen_.as<uint32_t>(), d_voff_.as<uint32_t>(), d_vlen_.as<uint32_t>()};
    const RowArrays to{d_rows2_[0].as<uint32_t>(), d_rows2_[1].as<uint32_t>(), d_rows2_[2].as<uint32_t>(), d_rows2_[3].as<uint32_t>(), d_rows2_[4].as<uint32_t>()};
    if ((e = launch_gather_rows(from, to, d_segs_.as<RowSegment>(), (uint32_t)h_segs_.n, new_n, stream_)) != cudaSuccess) return fail(e, "gather_rows", err);
    if (old_last_row == new_last_row) old_last_row = 0xffffffffu;
    if ((e = launch_patch_commas(to.meta, old_last_row, new_last_row, stream_)) != cudaSuccess) return fail(e, "patch_commas", err);
    for (int i = 0; i < 5; i++) { std::swap(rows[i]->p, d_rows2_[i].p); std::swap(rows[i]->cap, d_rows2_[i].cap); }
    tape_.n_rows = new_n;
    tape_.meta = d_meta_.as<uint32_t>(); tape_.koff = d_koff_.as<uint32_t>(); tape_.klen = d_klen_.as<uint32_t>();
    tape_.voff = d_voff_.as<uint32_t>(); tape_.vlen = d_vlen_.as<uint32_t>();
    tape_.arena = d_arena_.as<uint8_t>();
    tape_.arena_bytes = (uint32_t)arena_used_;
    if ((e = d_off_.reserve(((size_t)new_n + 1) * 4)) != cudaSuccess) return fail(e, "cudaMalloc(off)", err);
    if ((e = d_scratch_.reserve(scan_scratch_words(new_n) * 8)) != cudaSuccess) return fail(e, "cudaMalloc(scratch)", err);
    E.n_rows = new_n;
    E.tail_row = new_tail;
    E.last_is_added = new_last_is_added;
    E.last_base = new_last_base;
    E.last_added = new_last_added ? *new_last_added : std::string();
    E.dirty = false;
    staged_version_ = doc.version;
    return TK_OK;
}

}  // namespace tk

