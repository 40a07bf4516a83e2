// engine.hpp -- per-handle GPU context: the stream, device and pinned buffers (drawn from
// process-wide pools so that creating a handle costs no cudaMalloc after warm-up), the GPU parse
// driver (state.New), the host flattener (tree -> line tape) and the bulk operations.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstring>
#include <string>
#include <string_view>
#include <vector>

#include "doc.hpp"
#include "edit.cuh"
#include "index.cuh"
#include "parse.cuh"
#include "tape.hpp"

namespace tk {

// Process-wide caches of device and pinned-host blocks, keyed by power-of-two size.  Blocks go
// back to the cache when a handle releases them and to the driver only at process exit.
void *pool_alloc_device(size_t bytes, size_t *got);
void pool_free_device(void *p, size_t bytes);
void *pool_alloc_pinned(size_t bytes, size_t *got);
void pool_free_pinned(void *p, size_t bytes);
// A handle's buffers grow by taking a larger block and handing the old one back to the pool, where
// another handle (another thread, another stream) may pick it up at once.  Work this handle has
// queued may still be using the old block, so the handle's stream is drained first: every Engine
// entry point names its stream for the calling thread (StreamScope), and the two reserve()s below
// call pool_release_sync() before they give a block back.  Growth of a block that is in use is
// rare (a handle re-staging a document that got larger), so the wait costs nothing that matters.
void pool_release_sync();
struct StreamScope {
    explicit StreamScope(cudaStream_t s);
    ~StreamScope();
    StreamScope(const StreamScope &) = delete;
    StreamScope &operator=(const StreamScope &) = delete;
private:
    cudaStream_t prev_;
};

// grow-only pinned host array: host<->device copies from it are true DMA
template <class T>
struct PinVec {
    T *p = nullptr;
    size_t n = 0, cap = 0, cap_bytes = 0;
    bool failed = false;
    ~PinVec() { if (p) pool_free_pinned(p, cap_bytes); }
    PinVec() = default;
    PinVec(const PinVec &) = delete;
    PinVec &operator=(const PinVec &) = delete;
    bool reserve(size_t want) {
        if (want <= cap) return true;
        size_t got = 0;
        T *q = static_cast<T *>(pool_alloc_pinned(want * sizeof(T), &got));
        if (!q) { failed = true; return false; }
        if (n) memcpy(q, p, n * sizeof(T));
        if (p) { pool_release_sync(); pool_free_pinned(p, cap_bytes); }
        p = q;
        cap_bytes = got;
        cap = got / sizeof(T);
        return true;
    }
    void clear() { n = 0; }
    inline void push(const T &v) {
        if (n == cap && !reserve(n ? n * 2 : 1024)) return;
        p[n++] = v;
    }
    inline size_t append(const T *src, size_t len) {
        if (n + len > cap && !reserve((n + len) * 2)) return 0;
        if (len) memcpy(p + n, src, len * sizeof(T));
        const size_t at = n;
        n += len;
        return at;
    }
};

// grow-only device array; contents survive a grow only if keep_bytes says how many to carry over
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    ~DevBuf() { if (p) pool_free_device(p, cap); }
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    cudaError_t reserve(size_t bytes, size_t keep_bytes = 0, cudaStream_t s = nullptr);
    template <class T> T *as() const { return static_cast<T *>(p); }
};

enum KernelKind {
    K_ROW_OFFSETS = 0, K_EMIT, K_PREFIX_MATCH, K_SCAN_TOKENS, K_TOKEN_PREFIX, K_BRACKETS, K_SORT_BRACKETS, K_MATCH,
    K_DECODE, K_ROWS, K_KEY_ORDER, K_MODULE_INDEX, K_FIELD_FIND, K_LINES_COUNT, K_LINES_ROWS, K_LINES_STRUCTURE, K_SCAN_COUNT, K_SCAN_PREFIX, K_SCAN_EMIT, K_COUNT
};
const char *kernel_name(int kind);
void set_line_pass(bool on);       // state.New tries the line-by-line pass first (default) or goes straight to the general pipeline
void set_default_timing(bool on);  // engines created from now on start with per-kernel timing on/off

class Engine {
public:
    Engine() = default;
    ~Engine();
    Engine(const Engine &) = delete;
    Engine &operator=(const Engine &) = delete;

    int init(std::string *err);  // binds to the current device; TK_ECUDA + text if there is none
    cudaStream_t stream() const { return stream_; }

    // state.New: raw text -> line tape on the device.  On success either the tape is in output
    // order (doc->host_valid stays false: the tree is built only if a point operation needs it)
    // or the tree has been built from the tape on the host (keys out of order, duplicates, numbers
    // that need re-formatting).  Invalid text: TK_ESYNTAX with Go's message.
    int parse(const char *raw, size_t n, bool raw_on_device, Document *doc, std::string *err);
    // builds doc->root from the device tape (rows in input order); used by parse() and lazily by
    // the point operations
    int materialise(Document *doc, std::string *err);
    // the device-only part of parse(), for bench.py: text already at the start of the arena
    int parse_resident(size_t n, ParseResult *out, std::string *err);
    int upload_text(const char *raw, size_t n, std::string *err);
    int upload_text_device(const void *dev_raw, size_t n, std::string *err);
    const uint8_t *pending_text_ = nullptr;   // text still in the caller's device buffer: the line pass copies it as it counts (lines.cu)

    int stage(const Document &doc, std::string *err);
    int stage_tape(const Document &doc, std::string *err);
    int stage_keys(const Document &doc, std::string *err);

    int bytes(const Document &doc, bool to_host, const void **out, size_t *len, std::string *err);
    int prefix_hits(const Document &doc, std::string_view prefix, bool to_host, std::vector<uint32_t> *hits, size_t *count,
                    std::string *err);
    const std::pair<const std::string, Node *> *module_entry(uint32_t i) const { return entries_[i]; }

    // Clusters()/Nodes() for a document that lives on the device as the parser's tape: children of
    // "module" whose key starts with prefix and that have the string field ("name" / "hostname").
    // The packed index is built on the GPU once per document version; the lookup is a binary search.
    // *module_ok false: "module" is missing or not an object (gabs.ErrNotObj).
    bool device_resident(const Document &doc) const { return tape_from_parse_ && staged_version_ == doc.version; }
    // false for the rare document whose children have more members than the index pass walks through
    bool index_usable(const Document &doc) const { return device_resident(doc) && !parsed_.index_slow; }
    int index_lookup(const Document &doc, std::string_view prefix, bool hostname, bool *module_ok,
                     std::vector<std::pair<std::string_view, std::string_view>> *pairs, std::string *err);
    // The same query answered without copying when possible: *blob / *blob_len / *count describe the
    // finished C-ABI result inside the pre-packed blob of that field.  *direct false: the run is not
    // in strictly ascending field order (or has gaps) -- use index_lookup and sort.
    int index_lookup_direct(const Document &doc, std::string_view prefix, bool hostname, bool *module_ok, bool *direct,
                            const uint8_t **blob, size_t *blob_len, size_t *count, std::string *err);

    // ---- edits of module's children on a document that lives on the device (State.Delete,
    // AddNode, AddCluster, SetManager: a two-part path "module.<key>").  The host keeps the list of
    // children and what was done to it; Clusters() / Nodes() answer from the index plus that list;
    // the device tape is rebuilt from row segments when Bytes() or the host tree next need it
    // (apply_edits).  *handled false: not an edit this path takes -- the caller builds the host tree.
    bool edits_active() const { return edit_ != nullptr; }
    int edit_delete(Document *doc, std::string_view path, bool *handled, std::string *err);
    int edit_set(Document *doc, std::string_view path, Node *value, bool *handled, std::string *err);   // owns value if handled
    int apply_edits(const Document &doc, std::string *err);

    // measurement hooks
    void set_timing(bool on) { timing_ = on; }
    int collect_times(std::string *err);
    double kernel_ms(int kind) const { return kernel_ms_[kind]; }
    unsigned long long timed_launches(int kind) const { return timed_launches_[kind]; }
    unsigned long long launches(int kind) const { return launches_[kind]; }
    unsigned long long h2d_bytes() const { return h2d_bytes_; }
    unsigned long long d2h_bytes() const { return d2h_bytes_; }

private:
    int fail(cudaError_t e, const char *what, std::string *err);
    cudaError_t copy_h2d(void *dst, const void *src, size_t bytes);
    cudaError_t copy_d2h(void *dst, const void *src, size_t bytes);
    template <class F> cudaError_t timed(int kind, F &&launch);
    int fetch_result(std::string *err);
    int parse_lines(size_t n, ParseResult *out, bool *accepted, std::string *err);
    int build_index(const Document &doc, std::string *err);
    struct IndexView;
    void index_range(int field, std::string_view prefix, uint32_t *lo, uint32_t *hi);

    void flatten(const Document &doc);
    void flatten_keys(const Document &doc);

    bool ready_ = false;
    int device_ = -1;
    cudaStream_t stream_ = nullptr;
    uint64_t staged_version_ = 0, keys_version_ = 0;
    bool pin_failed_ = false;

    struct Pending { int kind; cudaEvent_t a, b; };
    std::vector<Pending> pending_;
    std::vector<cudaEvent_t> free_events_;
    bool timing_ = false;
    double kernel_ms_[K_COUNT] = {};
    unsigned long long timed_launches_[K_COUNT] = {}, launches_[K_COUNT] = {};
    unsigned long long h2d_bytes_ = 0, d2h_bytes_ = 0;

    // staging of text that arrives in pageable host memory (upload_text)
    static constexpr size_t STAGE_CHUNK = size_t(4) << 20, STAGE_MIN = size_t(1) << 20;
    PinVec<uint8_t> h_stage_[2];
    cudaEvent_t stage_ev_[2] = {nullptr, nullptr};

    // host (pinned) tape + key index
    PinVec<uint32_t> h_meta_, h_koff_, h_klen_, h_voff_, h_vlen_, h_koffs_;
    PinVec<uint8_t> h_arena_, h_keys_;
    std::vector<const std::pair<const std::string, Node *> *> entries_;

    // device tape, key index and work buffers
    DevBuf d_meta_, d_koff_, d_klen_, d_voff_, d_vlen_, d_arena_, d_koffs_, d_keys_;
    DevBuf d_off_, d_out_, d_scratch_, d_hits_, d_prefix_;
    TapeView tape_{};
    KeyIndexView keys_{};
    size_t raw_len_ = 0;      // bytes of the parsed text at the start of d_arena_
    size_t arena_used_ = 0;   // raw text (padded) + side arena
    bool tape_from_parse_ = false;

    // parse work buffers
    DevBuf d_tok_pos_, d_tok_kind_, d_dec_bits_, d_tiles_, d_res_;
    DevBuf d_br_[7], d_dec_tok_, d_dec_off_, d_dec_len_, d_dec_counts_, d_dec_offs_;
    PinVec<ParseResult> h_res_;
    PinVec<unsigned long long> h_words_;

    // module index: the field lists field_find_kernel leaves at the end of the parse, packed on demand (csrc/index.cu)
    DevBuf d_frefs_[2], d_find_status_;
    DevBuf d_fviolbits_[2], d_fblob_[2], d_foff_[2];
    PinVec<uint8_t> h_fblob_[2];
    PinVec<uint32_t> h_foff_[2], h_fviolbits_[2];
    uint64_t index_version_ = 0;
    uint32_t index_m_[2] = {0, 0}, index_hint_[2] = {0, 0};   // per field: records in the index, where the last run ended
    bool index_module_ok_ = false;
    ParseResult parsed_{};  // the verdict of the last parse: where "module" is, the sizes of the field lists
    FieldIndexBuffers fx_{};   // the lists field_find_kernel filled (kept for a second run with more room)
    BracketBuffers bb_{};      // bracket tables of the last parse (the child list of a first edit reads match / row)

    // edits (see above): host state, the child list's device and host buffers, the second set of row arrays
    struct EditState;
    EditState *edit_ = nullptr;
    void drop_edit_state();
    bool edit_record_dead(int f, uint32_t j) const;
    bool edit_added_any(int f, std::string_view prefix) const;
    void edit_added_pairs(int f, std::string_view prefix, std::vector<std::pair<std::string_view, std::string_view>> *pairs) const;
    void adopt_added_values(Document *doc);
    int enter_edit_mode(const Document &doc, std::string *err);
    bool edit_path_key(const Document &doc, std::string_view path, std::string_view *key, std::string *err, int *rc);
    DevBuf d_cflags_, d_crank_, d_clist_[4], d_ckeyoff_, d_ckeys_, d_rows2_[5], d_segs_;
    PinVec<uint32_t> h_clist_[3], h_addrows_[5];
    PinVec<uint8_t> h_ckeys_, h_addarena_;
    PinVec<RowSegment> h_segs_;

    // results
    PinVec<uint8_t> h_out_;
    PinVec<uint32_t> h_hits_;
    PinVec<unsigned long long> h_total_;
    uint64_t total_version_ = 0;
    size_t out_total_ = 0;
};

}  // namespace tk
