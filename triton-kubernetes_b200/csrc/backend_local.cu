// backend_local.cu -- the local backend (backend/local/backend.go:14-132): one directory per
// cluster manager under <root> (default ~/.triton-kubernetes), holding main.tf.json, the
// document exactly as State.Bytes() gives it, and terraform's own terraform.tfstate next to it.
//
// Byte-compatible with the reference's on-disk layout.  One deliberate improvement
// (SURVEY.md 8f-1): PersistState writes to a temporary file in the same directory and renames it
// over main.tf.json, so a crash never leaves a half-written document; the reference writes in place.
#include <dirent.h>
#include <errno.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "engine.hpp"
#include "handle.hpp"
#include "records.hpp"
#include "tk_state.h"

using namespace tk;

struct tk_backend {
    std::string root;  // expanded root directory
    std::string err;
    std::string out1, out2;
};

namespace {

// "open /x/y: no such file or directory" -- the shape of Go's *os.PathError text
std::string path_error(const char *op, const std::string &path, int e) {
    std::string m = strerror(e);
    if (!m.empty() && m[0] >= 'A' && m[0] <= 'Z') m[0] = (char)(m[0] - 'A' + 'a');
    return std::string(op) + " " + path + ": " + m;
}

int mkdir_all(const std::string &path, std::string *err) {  // os.MkdirAll(path, os.ModePerm)
    struct stat st;
    if (stat(path.c_str(), &st) == 0) {
        if (S_ISDIR(st.st_mode)) return 0;
        *err = path_error("mkdir", path, ENOTDIR);
        return -1;
    }
    size_t slash = path.find_last_of('/');
    if (slash != std::string::npos && slash > 0 && mkdir_all(path.substr(0, slash), err)) return -1;
    if (mkdir(path.c_str(), 0777) != 0 && errno != EEXIST) { *err = path_error("mkdir", path, errno); return -1; }
    return 0;
}

int remove_all(const std::string &path, std::string *err) {  // os.RemoveAll: a missing path is not an error
    struct stat st;
    if (lstat(path.c_str(), &st) != 0) {
        if (errno == ENOENT) return 0;
        *err = path_error("lstat", path, errno);
        return -1;
    }
    if (S_ISDIR(st.st_mode)) {
        DIR *d = opendir(path.c_str());
        if (!d) { *err = path_error("open", path, errno); return -1; }
        std::vector<std::string> names;
        while (dirent *e = readdir(d))
            if (strcmp(e->d_name, ".") && strcmp(e->d_name, "..")) names.push_back(e->d_name);
        closedir(d);
        for (auto &n : names)
            if (remove_all(path + "/" + n, err)) return -1;
        if (rmdir(path.c_str()) != 0 && errno != ENOENT) { *err = path_error("remove", path, errno); return -1; }
        return 0;
    }
    if (unlink(path.c_str()) != 0 && errno != ENOENT) { *err = path_error("remove", path, errno); return -1; }
    return 0;
}

int fail(tk_backend *b, int code, std::string msg) { b->err = std::move(msg); return code; }
std::string_view sv(const char *p, size_t n) { return p ? std::string_view(p, n) : std::string_view(); }

}  // namespace

extern "C" {

static thread_local std::string g_backend_err;  // for a failed tk_local_backend_new
const char *tk_backend_last_error(tk_backend *b) { return b ? b->err.c_str() : g_backend_err.c_str(); }

// local.New() -- backend/local/backend.go:28-41
int tk_local_backend_new(const char *root_dir, size_t root_dir_len, tk_backend **out) {
    if (!out) return TK_EARG;
    *out = nullptr;
    auto *b = new tk_backend();
    if (root_dir && root_dir_len) {
        b->root.assign(root_dir, root_dir_len);
    } else {
        const char *home = getenv("HOME");  // go-homedir: $HOME first
        if (!home || !*home) { delete b; g_backend_err = "cannot expand ~: $HOME is not set"; return TK_EIO; }
        b->root = std::string(home) + "/.triton-kubernetes";
    }
    while (b->root.size() > 1 && b->root.back() == '/') b->root.pop_back();
    std::string err;
    if (mkdir_all(b->root, &err)) { delete b; g_backend_err = err; return TK_EIO; }
    *out = b;
    return TK_OK;
}

void tk_backend_free(tk_backend *b) { delete b; }

// Backend.State(name) -- backend/local/backend.go:43-66
int tk_backend_state(tk_backend *b, const char *name, size_t name_len, tk_state **out) {
    if (!b || !out) return TK_EARG;
    const std::string path = b->root + "/" + std::string(sv(name, name_len)) + "/main.tf.json";
    struct stat st;
    if (stat(path.c_str(), &st) != 0) {
        // os.IsNotExist is true for ENOENT only: any other stat error (ENOTDIR, EACCES ...) is returned
        if (errno == ENOENT) return tk_state_new_empty(name, name_len, out);  // state.New(name, []byte("{}"))
        return fail(b, TK_EIO, path_error("stat", path, errno));
    }
    int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) return fail(b, TK_EIO, path_error("open", path, errno));
    // The file is read straight into a pinned block from the library's pool, so that the
    // host->device copy of state.New is a DMA from where read(2) put the bytes (a pageable buffer
    // would be staged through the driver's bounce buffers first).  Without a device there is no
    // pinned memory; plain memory then carries the text to tk_state_new, which reports TK_ECUDA.
    size_t cap = (size_t)st.st_size + 65536, pin_bytes = 0, got = 0;
    char *pin = static_cast<char *>(pool_alloc_pinned(cap, &pin_bytes));
    std::string plain;
    char *buf = pin;
    if (pin) cap = pin_bytes; else { plain.resize(cap); buf = plain.data(); }
    int io_errno = 0;
    for (;;) {
        if (got == cap) {   // the file grew under us: carry on in a larger block
            size_t more_bytes = 0;
            char *more = pin ? static_cast<char *>(pool_alloc_pinned(cap * 2, &more_bytes)) : nullptr;
            if (more) {
                memcpy(more, buf, got);
                pool_free_pinned(pin, pin_bytes);
                pin = buf = more; pin_bytes = cap = more_bytes;
            } else {
                if (pin) { plain.assign(buf, got); pool_free_pinned(pin, pin_bytes); pin = nullptr; }
                plain.resize(cap * 2);
                buf = plain.data(); cap = plain.size();
            }
        }
        ssize_t r = read(fd, buf + got, cap - got);
        if (r < 0) { if (errno == EINTR) continue; io_errno = errno; break; }
        if (r == 0) break;
        got += (size_t)r;
    }
    close(fd);
    int rc = io_errno ? TK_EIO : tk_state_new(name, name_len, buf, got, out);
    if (pin) pool_free_pinned(pin, pin_bytes);   // tk_state_new has waited for its copy of the text
    if (io_errno) return fail(b, TK_EIO, path_error("read", path, io_errno));
    if (rc) return fail(b, rc, tk_last_error(nullptr));
    return TK_OK;
}

// Backend.DeleteState(name) -- backend/local/backend.go:68-77
int tk_backend_delete_state(tk_backend *b, const char *name, size_t name_len) {
    if (!b) return TK_EARG;
    std::string err;
    if (remove_all(b->root + "/" + std::string(sv(name, name_len)), &err)) return fail(b, TK_EIO, err);
    return TK_OK;
}

// Backend.PersistState(state) -- backend/local/backend.go:79-100
int tk_backend_persist_state(tk_backend *b, tk_state *s) {
    if (!b || !s) return TK_EARG;
    const std::string dir = b->root + "/" + s->name;
    std::string err;
    (void)mkdir_all(dir, &err);  // the reference ignores this error too; the write below reports the problem
    const char *text;
    size_t n;
    int rc = tk_state_bytes(s, &text, &n);
    if (rc) return fail(b, rc, tk_last_error(s));
    // a temporary file of its own name (two processes may persist the same manager at once)
    const std::string path = dir + "/main.tf.json";
    // (O_EXCL with mode 0644, not mkstemp: the kernel then applies the umask as it does for the
    //  reference's ioutil.WriteFile(path, bytes, 0644))
    static std::atomic<unsigned> serial{0};
    std::string tmp;
    int fd = -1;
    for (int attempt = 0; attempt < 100 && fd < 0; attempt++) {
        tmp = dir + "/.main.tf.json." + std::to_string((long)getpid()) + "." + std::to_string(serial++) + ".tmp";
        fd = open(tmp.c_str(), O_WRONLY | O_CREAT | O_EXCL, 0644);
        if (fd < 0 && errno != EEXIST) break;
    }
    if (fd < 0) return fail(b, TK_EIO, path_error("open", tmp, errno));
    size_t put = 0;
    while (put < n) {
        ssize_t w = write(fd, text + put, n - put);
        if (w < 0) { int e = errno; close(fd); unlink(tmp.c_str()); return fail(b, TK_EIO, path_error("write", path, e)); }
        put += (size_t)w;
    }
    if (close(fd) != 0) { int e = errno; unlink(tmp.c_str()); return fail(b, TK_EIO, path_error("close", path, e)); }
    if (rename(tmp.c_str(), path.c_str()) != 0) { int e = errno; unlink(tmp.c_str()); return fail(b, TK_EIO, path_error("rename", path, e)); }
    return TK_OK;
}

// Backend.States() -- backend/local/backend.go:102-121: names of the directories under root, sorted
int tk_backend_states(tk_backend *b, const char **names, size_t *names_len, size_t *count) {
    if (!b || !names || !names_len || !count) return TK_EARG;
    DIR *d = opendir(b->root.c_str());
    if (!d) return fail(b, TK_EIO, path_error("open", b->root, errno));
    std::vector<std::string> v;
    while (dirent *e = readdir(d)) {
        if (!strcmp(e->d_name, ".") || !strcmp(e->d_name, "..")) continue;
        struct stat st;
        if (lstat((b->root + "/" + e->d_name).c_str(), &st) == 0 && S_ISDIR(st.st_mode)) v.push_back(e->d_name);
    }
    closedir(d);
    std::sort(v.begin(), v.end());
    b->out1.clear();
    for (size_t i = 0; i < v.size(); i++) { if (i) b->out1.push_back('\n'); b->out1 += v[i]; }
    *names = b->out1.data();
    *names_len = b->out1.size();
    *count = v.size();
    return TK_OK;
}

// Backend.StateTerraformConfig(name) -- backend/local/backend.go:123-132
int tk_backend_state_terraform_config(tk_backend *b, const char *name, size_t name_len, const char **path, size_t *path_len,
                                      const char **record_json, size_t *record_len) {
    if (!b || !path || !path_len || !record_json || !record_len) return TK_EARG;
    b->out1 = "terraform.backend.local";
    StrNode p(b->root + "/" + std::string(sv(name, name_len)) + "/terraform.tfstate");
    MapNode settings;
    settings.m.emplace("path", &p);
    b->out2.clear();
    std::string err;
    const bool ok = render_record(*find_record("localTerraformBackendConfig"), &settings, &b->out2, &err);
    settings.m.clear();  // p lives on the stack
    if (!ok) return fail(b, TK_EARG, err);
    *path = b->out1.data(); *path_len = b->out1.size();
    *record_json = b->out2.data(); *record_len = b->out2.size();
    return TK_OK;
}

}  // extern "C"
