// engine.hpp -- per-handle GPU context: pinned staging buffers, device buffers, the stream, and
// the host flattener that turns the tree into the line tape (tape.hpp).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <string_view>
#include <vector>

#include "doc.hpp"
#include "tape.hpp"

namespace tk {

// grow-only pinned host array (cudaHostAlloc): host->device copies from it are true DMA
template <class T>
struct PinVec {
    T *p = nullptr;
    size_t n = 0, cap = 0;
    bool failed = false;
    ~PinVec() { if (p) cudaFreeHost(p); }
    PinVec() = default;
    PinVec(const PinVec &) = delete;
    PinVec &operator=(const PinVec &) = delete;
    bool reserve(size_t want);
    void clear() { n = 0; }
    inline void push(const T &v) {
        if (n == cap && !reserve(n + 1)) return;
        p[n++] = v;
    }
    // appends len items and returns the index of the first
    inline size_t append(const T *src, size_t len) {
        if (n + len > cap && !reserve(n + len)) return 0;
        if (len) memcpy(p + n, src, len * sizeof(T));
        size_t at = n;
        n += len;
        return at;
    }
};

// grow-only device array
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    ~DevBuf() { if (p) cudaFree(p); }
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    cudaError_t reserve(size_t bytes);
    template <class T> T *as() const { return static_cast<T *>(p); }
};

enum KernelKind { K_ROW_OFFSETS = 0, K_EMIT = 1, K_PREFIX_MATCH = 2, K_COUNT = 3 };
const char *kernel_name(int kind);

class Engine {
public:
    // per-kernel device time, measured with events on the handle's stream while timing is on
    void set_timing(bool on);
    int collect_times(std::string *err);  // synchronises the stream
    double kernel_ms(int kind) const { return kernel_ms_[kind]; }
    unsigned long long timed_launches(int kind) const { return timed_launches_[kind]; }
    unsigned long long launches(int kind) const { return launches_[kind]; }
    unsigned long long h2d_bytes() const { return h2d_bytes_; }
    unsigned long long d2h_bytes() const { return d2h_bytes_; }

    Engine() = default;
    ~Engine();
    Engine(const Engine &) = delete;
    Engine &operator=(const Engine &) = delete;

    // binds to the current device and creates the stream; TK_ECUDA + text if there is no device
    int init(std::string *err);
    cudaStream_t stream() const { return stream_; }

    // flatten + upload whatever differs from doc.version: the line tape (for Bytes) and the
    // packed module keys (for Clusters/Nodes) are staged independently
    int stage(const Document &doc, std::string *err);
    int stage_tape(const Document &doc, std::string *err);
    int stage_keys(const Document &doc, std::string *err);

    // State.Bytes(): text into pinned host memory (to_host) or left on the device
    int bytes(const Document &doc, bool to_host, const void **out, size_t *len, std::string *err);

    // prefix classification over the staged module keys; hit indices (unordered) into *hits
    int prefix_hits(const Document &doc, std::string_view prefix, bool to_host, std::vector<uint32_t> *hits,
                    size_t *count, std::string *err);

    // module child i of the staged key index (valid while doc is unchanged)
    const std::pair<const std::string, Node *> *module_entry(uint32_t i) const { return entries_[i]; }
    bool module_is_object() const { return module_ok_; }

private:
    int fail(cudaError_t e, const char *what, std::string *err);
    cudaError_t copy_h2d(void *dst, const void *src, size_t bytes);
    cudaError_t copy_d2h(void *dst, const void *src, size_t bytes);
    template <class F> cudaError_t timed(int kind, F &&launch);
    struct Pending { int kind; cudaEvent_t a, b; };
    std::vector<Pending> pending_;
    std::vector<cudaEvent_t> free_events_;
    bool timing_ = false;
    double kernel_ms_[K_COUNT] = {0, 0, 0};
    unsigned long long timed_launches_[K_COUNT] = {0, 0, 0}, launches_[K_COUNT] = {0, 0, 0};
    unsigned long long h2d_bytes_ = 0, d2h_bytes_ = 0;
    void flatten(const Document &doc);
    void flatten_keys(const Document &doc);
    void flat_value(const Node *n, uint32_t depth, const std::string *key, bool comma, bool last);
    void push_row(uint32_t kind, uint32_t depth, const std::string *key, const char *val, size_t vlen, bool comma, bool last);

    bool ready_ = false;
    int device_ = -1;
    cudaStream_t stream_ = nullptr;
    uint64_t staged_version_ = 0, keys_version_ = 0;
    bool pin_failed_ = false;

    // host (pinned) tape + key index
    PinVec<uint32_t> h_meta_, h_koff_, h_klen_, h_voff_, h_vlen_, h_koffs_;
    PinVec<uint8_t> h_arena_, h_keys_;
    std::vector<const std::pair<const std::string, Node *> *> entries_;
    bool module_ok_ = false;

    // device mirrors and work buffers
    DevBuf d_meta_, d_koff_, d_klen_, d_voff_, d_vlen_, d_arena_, d_koffs_, d_keys_;
    DevBuf d_off_, d_out_, d_scratch_, d_hits_, d_prefix_;
    TapeView tape_{};
    KeyIndexView keys_{};

    // results
    PinVec<uint8_t> h_out_;
    PinVec<uint32_t> h_hits_;
    PinVec<unsigned long long> h_total_;
    uint64_t total_version_ = 0;  // version for which out_total_ is known
    size_t out_total_ = 0;
};

}  // namespace tk
