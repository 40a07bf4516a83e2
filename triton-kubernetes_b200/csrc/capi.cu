// capi.cu -- the extern "C" surface declared in include/tk_state.h.
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "doc.hpp"
#include "engine.hpp"
#include "handle.hpp"
#include "kernels.cuh"
#include "records.hpp"
#include "tk_state.h"
#include "tk_state_bench.h"

using namespace tk;


static thread_local std::string g_err;        // for calls that have no handle
static thread_local std::string g_hostnames;  // tk_new_hostnames result

static int set_err(tk_state *s, int code, std::string msg) {
    if (s) s->err = std::move(msg); else g_err = std::move(msg);
    return code;
}
static std::string_view sv(const char *p, size_t n) { return p ? std::string_view(p, n) : std::string_view(); }

// The point operations work on the host tree; a document that so far exists only as the device
// tape state.New produced gets its tree built here, once.
static int ensure_host(tk_state *s) {
    if (s->doc.host_valid) return TK_OK;
    std::string err;
    int rc = s->eng.materialise(&s->doc, &err);
    if (rc) return set_err(s, rc, err);
    return TK_OK;
}

extern "C" {

const char *tk_version(void) { return "libtkstate 0.1 (sm_100a)"; }
unsigned long long tk_kernel_launches(void) { return kernel_launch_count() + parse_launch_count(); }

int tk_set_device(int device) {
    const cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) return TK_OK;
    g_err = std::string("cuda: cudaSetDevice: ") + cudaGetErrorString(e);
    (void)cudaGetLastError();
    return TK_ECUDA;
}

const char *tk_last_error(tk_state *s) { return s ? s->err.c_str() : g_err.c_str(); }

static int new_from(const char *name, size_t name_len, const char *raw, size_t raw_len, bool on_device, tk_state **out);

int tk_state_new(const char *name, size_t name_len, const char *raw, size_t raw_len, tk_state **out) {
    return new_from(name, name_len, raw, raw_len, false, out);
}
int tk_state_new_device(const char *name, size_t name_len, const void *dev_raw, size_t raw_len, tk_state **out) {
    return new_from(name, name_len, static_cast<const char *>(dev_raw), raw_len, true, out);
}
int tk_set_default_timing(int enable) { set_default_timing(enable != 0); return TK_OK; }
int tk_set_line_pass(int enable) { set_line_pass(enable != 0); return TK_OK; }

static int new_from(const char *name, size_t name_len, const char *raw, size_t raw_len, bool on_device, tk_state **out) {
    if (!out) return set_err(nullptr, TK_EARG, "tk_state_new: out is NULL");
    *out = nullptr;
    auto *s = new tk_state();
    s->name.assign(sv(name, name_len));
    std::string err;
    // whole documents are parsed on the GPU; without a device this fails, loudly
    int rc = s->eng.parse(raw, raw_len, on_device, &s->doc, &err);
    if (rc) { delete s; return set_err(nullptr, rc, err); }
    *out = s;
    return TK_OK;
}

int tk_state_new_empty(const char *name, size_t name_len, tk_state **out) {
    if (!out) return set_err(nullptr, TK_EARG, "tk_state_new_empty: out is NULL");
    auto *s = new tk_state();
    s->name.assign(sv(name, name_len));
    s->doc.root = new MapNode();
    *out = s;
    return TK_OK;
}

void tk_state_free(tk_state *s) { delete s; }

int tk_state_name(tk_state *s, const char **name, size_t *name_len) {
    if (!s || !name || !name_len) return set_err(s, TK_EARG, "tk_state_name: NULL argument");
    *name = s->name.data();
    *name_len = s->name.size();
    return TK_OK;
}

int tk_state_get(tk_state *s, const char *path, size_t path_len, const char **value, size_t *value_len) {
    if (!s || !value || !value_len) return set_err(s, TK_EARG, "tk_state_get: NULL argument");
    if (!s->doc.host_valid && (s->eng.edits_active() || s->eng.index_usable(s->doc))) {
        // "module.<key>.name" / "module.<key>.hostname" of a document on the device -- what the flows ask a loaded state
        // (create/node.go, destroy/*.go) -- is in the module index: no host tree for it.  Anything the index does
        // not hold (no such child, the member missing or not a string) is "" for Get as well.
        const std::string_view p = sv(path, path_len);
        const size_t d1 = p.find('.'), d2 = d1 == std::string_view::npos ? d1 : p.find('.', d1 + 1);
        if (d2 != std::string_view::npos && p.find('.', d2 + 1) == std::string_view::npos && p.substr(0, d1) == "module") {
            const std::string_view key = p.substr(d1 + 1, d2 - d1 - 1), member = p.substr(d2 + 1);
            if (member == "name" || member == "hostname") {
                std::string err;
                bool module_ok = false;
                int rc = s->eng.index_lookup(s->doc, key, member == "hostname", &module_ok, &s->pairs, &err);
                if (rc) return set_err(s, rc, err);
                if (module_ok) {
                    s->blob.clear();
                    for (auto &pr : s->pairs)
                        if (pr.second == key) { s->blob.assign(pr.first.data(), pr.first.size()); break; }
                    *value = s->blob.data();
                    *value_len = s->blob.size();
                    return TK_OK;
                }
            }
        }
    }
    if (int rc = ensure_host(s)) return rc;
    std::string_view v = s->doc.get(sv(path, path_len));
    *value = v.data() ? v.data() : "";
    *value_len = v.size();
    return TK_OK;
}

int tk_state_set(tk_state *s, const char *path, size_t path_len, const char *json, size_t json_len, int value_kind) {
    if (!s) return set_err(s, TK_EARG, "tk_state_set: NULL handle");
    Node *v = nullptr;
    std::string err;
    if (!s->doc.host_valid) {
        // a document on the device: a child of "module" is set there (no host tree, no second upload);
        // for any other path the tree is built first, as for every other point operation
        if (!parse_value(sv(json, json_len), value_kind == TK_VALUE_RECORD, &v, &err)) {
            if (int rc = ensure_host(s)) return rc;      // (gabs walks the path before json.Marshal's result is looked at: same order of errors)
            return set_err(s, TK_ESYNTAX, err);
        }
        bool handled = false;
        int rc = s->eng.edit_set(&s->doc, sv(path, path_len), v, &handled, &err);
        if (rc) { free_node(v); return set_err(s, rc, err); }
        if (handled) return TK_OK;
        if ((rc = ensure_host(s)) != 0) { free_node(v); return rc; }
        rc = s->doc.set(sv(path, path_len), v, &err);
        if (rc) return set_err(s, rc, err);
        return TK_OK;
    }
    if (!parse_value(sv(json, json_len), value_kind == TK_VALUE_RECORD, &v, &err)) return set_err(s, TK_ESYNTAX, err);
    int rc = s->doc.set(sv(path, path_len), v, &err);
    if (rc) return set_err(s, rc, err);
    return TK_OK;
}

int tk_state_set_manager(tk_state *s, const char *json, size_t json_len, int value_kind) {
    static const char path[] = "module.cluster-manager";
    return tk_state_set(s, path, sizeof path - 1, json, json_len, value_kind);
}

int tk_state_set_terraform_backend_config(tk_state *s, const char *path, size_t path_len, const char *json, size_t json_len,
                                          int value_kind) {
    return tk_state_set(s, path, path_len, json, json_len, value_kind);
}

int tk_state_add_cluster(tk_state *s, const char *provider, size_t provider_len, const char *name, size_t name_len,
                         const char *json, size_t json_len, int value_kind) {
    std::string path = "module.cluster_";
    path.append(sv(provider, provider_len));
    path.push_back('_');
    path.append(sv(name, name_len));
    return tk_state_set(s, path.data(), path.size(), json, json_len, value_kind);
}

int tk_state_add_node(tk_state *s, const char *cluster_key, size_t cluster_key_len, const char *name, size_t name_len,
                      const char *json, size_t json_len, int value_kind) {
    if (!s) return set_err(s, TK_EARG, "tk_state_add_node: NULL handle");
    std::string_view provider, cname;
    std::string err;
    if (!cluster_key_parts(sv(cluster_key, cluster_key_len), &provider, &cname, &err)) return set_err(s, TK_EKEYPARTS, err);
    std::string path = "module.node_";
    path.append(provider);
    path.push_back('_');
    path.append(cname);
    path.push_back('_');
    path.append(sv(name, name_len));
    return tk_state_set(s, path.data(), path.size(), json, json_len, value_kind);
}

int tk_state_delete(tk_state *s, const char *path, size_t path_len) {
    if (!s) return set_err(s, TK_EARG, "tk_state_delete: NULL handle");
    std::string err;
    if (!s->doc.host_valid) {
        bool handled = false;
        int rc0 = s->eng.edit_delete(&s->doc, sv(path, path_len), &handled, &err);   // a child of "module", on the device
        if (rc0) return set_err(s, rc0, err);
        if (handled) return TK_OK;
    }
    if (int rc = ensure_host(s)) return rc;
    int rc = s->doc.del(sv(path, path_len), &err);
    if (rc) return set_err(s, rc, err);
    return TK_OK;
}

int tk_state_stage(tk_state *s) {
    if (!s) return set_err(s, TK_EARG, "tk_state_stage: NULL handle");
    std::string err;
    if (int rc = ensure_host(s)) return rc;
    int rc = s->eng.stage(s->doc, &err);
    if (rc) return set_err(s, rc, err);
    return TK_OK;
}

int tk_state_bytes(tk_state *s, const char **out, size_t *out_len) {
    if (!s || !out || !out_len) return set_err(s, TK_EARG, "tk_state_bytes: NULL argument");
    std::string err;
    const void *p = nullptr;
    int rc = s->eng.bytes(s->doc, true, &p, out_len, &err);
    if (rc) return set_err(s, rc, err);
    *out = static_cast<const char *>(p);
    return TK_OK;
}

int tk_state_bytes_device(tk_state *s, const void **dev_out, size_t *out_len) {
    if (!s || !dev_out || !out_len) return set_err(s, TK_EARG, "tk_state_bytes_device: NULL argument");
    std::string err;
    int rc = s->eng.bytes(s->doc, false, dev_out, out_len, &err);
    if (rc) return set_err(s, rc, err);
    return TK_OK;
}

int tk_state_prefix_count_device(tk_state *s, const char *prefix, size_t prefix_len, size_t *count) {
    if (!s) return set_err(s, TK_EARG, "tk_state_prefix_count_device: NULL handle");
    std::string err;
    if (int rc = ensure_host(s)) return rc;
    int rc = s->eng.prefix_hits(s->doc, sv(prefix, prefix_len), false, nullptr, count, &err);
    if (rc) return set_err(s, rc, err);
    return TK_OK;
}

int tk_state_timing(tk_state *s, int enable) {
    if (!s) return set_err(s, TK_EARG, "tk_state_timing: NULL handle");
    s->eng.set_timing(enable != 0);
    return TK_OK;
}

int tk_state_kernel_stats(tk_state *s, int kind, const char **name, double *ms, unsigned long long *timed_launches,
                          unsigned long long *launches) {
    if (!s) return set_err(s, TK_EARG, "tk_state_kernel_stats: NULL handle");
    if (kind < 0 || kind >= K_COUNT) return set_err(s, TK_EARG, "tk_state_kernel_stats: no such kernel");
    std::string err;
    int rc = s->eng.collect_times(&err);
    if (rc) return set_err(s, rc, err);
    if (name) *name = kernel_name(kind);
    if (ms) *ms = s->eng.kernel_ms(kind);
    if (timed_launches) *timed_launches = s->eng.timed_launches(kind);
    if (launches) *launches = s->eng.launches(kind);
    return TK_OK;
}

int tk_state_copy_stats(tk_state *s, unsigned long long *h2d_bytes, unsigned long long *d2h_bytes) {
    if (!s) return set_err(s, TK_EARG, "tk_state_copy_stats: NULL handle");
    if (h2d_bytes) *h2d_bytes = s->eng.h2d_bytes();
    if (d2h_bytes) *d2h_bytes = s->eng.d2h_bytes();
    return TK_OK;
}

int tk_state_host_tree_built(tk_state *s) { return s && s->doc.host_valid ? 1 : 0; }

void *tk_state_stream(tk_state *s) { return s ? (void *)s->eng.stream() : nullptr; }

// Shared body of Clusters()/Nodes() (state/state.go:94-147): children of "module" whose key has
// the prefix (classified on the GPU), mapped <field> -> key for children that are maps with a
// string <field>.
static int children_by_prefix(tk_state *s, std::string_view prefix, std::string_view field, const uint8_t **blob,
                              size_t *blob_len, size_t *count) {
    std::string err;
    auto &pairs = s->pairs;  // (field value, module key); kept on the handle so a sweep of calls re-uses its storage
    pairs.clear();
    if (!s->doc.host_valid && (s->eng.edits_active() || s->eng.index_usable(s->doc))) {
        // the document lives on the device: one GPU pass per version builds the packed index, and
        // the finished answer is usually a byte range of the result records packed alongside it
        bool module_ok = false, direct = false;
        int rc0 = s->eng.index_lookup_direct(s->doc, prefix, field == "hostname", &module_ok, &direct, blob, blob_len, count, &err);
        if (rc0) return set_err(s, rc0, err);
        if (!module_ok) return set_err(s, TK_ENOTOBJ, kErrNotObj);
        if (direct) return TK_OK;
        int rc = s->eng.index_lookup(s->doc, prefix, field == "hostname", &module_ok, &pairs, &err);
        if (rc) return set_err(s, rc, err);
        if (!module_ok) return set_err(s, TK_ENOTOBJ, kErrNotObj);
    } else {
        if (int rc = ensure_host(s)) return rc;
        const MapNode *mod = s->doc.module_children();
        if (!mod) return set_err(s, TK_ENOTOBJ, kErrNotObj);
        // the tree keeps keys in bytewise order: keys with the prefix form one contiguous run
        for (auto it = mod->m.lower_bound(prefix); it != mod->m.end(); ++it) {
            const std::string &key = it->first;
            if (key.size() < prefix.size() || key.compare(0, prefix.size(), prefix) != 0) break;
            const Node *child = it->second;
            if (!child || child->kind != Kind::Map) continue;  // records (Go structs) are opaque
            auto &m = static_cast<const MapNode *>(child)->m;
            auto f = m.find(field);
            if (f == m.end() || !f->second || f->second->kind != Kind::Str) continue;
            pairs.emplace_back(static_cast<const StrNode *>(f->second)->s, key);
        }
    }
    // pairs arrive in key order, which for the usual naming is name order too: sort only if needed
    if (!std::is_sorted(pairs.begin(), pairs.end())) std::sort(pairs.begin(), pairs.end());
    size_t bytes = 0;
    for (auto &p : pairs) bytes += 8 + p.first.size() + p.second.size();
    s->blob.resize(bytes);
    char *w = s->blob.data();
    size_t kept = 0;
    for (size_t i = 0; i < pairs.size(); i++) {
        if (i + 1 < pairs.size() && pairs[i + 1].first == pairs[i].first) continue;  // greatest key wins
        uint32_t l = (uint32_t)pairs[i].first.size();
        memcpy(w, &l, 4); w += 4;
        memcpy(w, pairs[i].first.data(), l); w += l;
        l = (uint32_t)pairs[i].second.size();
        memcpy(w, &l, 4); w += 4;
        memcpy(w, pairs[i].second.data(), l); w += l;
        kept++;
    }
    s->blob.resize((size_t)(w - s->blob.data()));
    *blob = reinterpret_cast<const uint8_t *>(s->blob.data());
    *blob_len = s->blob.size();
    *count = kept;
    return TK_OK;
}

int tk_state_clusters(tk_state *s, const uint8_t **blob, size_t *blob_len, size_t *count) {
    if (!s || !blob || !blob_len || !count) return set_err(s, TK_EARG, "tk_state_clusters: NULL argument");
    return children_by_prefix(s, "cluster_", "name", blob, blob_len, count);
}

int tk_state_nodes(tk_state *s, const char *cluster_key, size_t cluster_key_len, const uint8_t **blob, size_t *blob_len,
                   size_t *count) {
    if (!s || !blob || !blob_len || !count) return set_err(s, TK_EARG, "tk_state_nodes: NULL argument");
    std::string_view provider, cname;
    std::string err;
    if (!cluster_key_parts(sv(cluster_key, cluster_key_len), &provider, &cname, &err)) return set_err(s, TK_EKEYPARTS, err);
    std::string &prefix = s->prefix;  // kept on the handle: no allocation per call in a sweep
    prefix.assign("node_");
    prefix.append(provider);
    prefix.push_back('_');
    prefix.append(cname);
    prefix.push_back('_');
    return children_by_prefix(s, prefix, "hostname", blob, blob_len, count);
}

int tk_new_hostnames(const char *const *existing, const size_t *existing_len, size_t n, const char *node_name,
                     size_t node_name_len, long long nodes_to_add, const char **out, size_t *out_len) {
    if (!out || !out_len || (n && (!existing || !existing_len))) return set_err(nullptr, TK_EARG, "tk_new_hostnames: NULL argument");
    std::vector<std::string_view> ex;
    ex.reserve(n);
    for (size_t i = 0; i < n; i++) ex.push_back(sv(existing[i], existing_len[i]));
    g_hostnames.clear();
    bool firstname = true;
    for (auto &h : new_hostnames(ex, sv(node_name, node_name_len), nodes_to_add)) {
        if (!firstname) g_hostnames.push_back('\n');
        g_hostnames.append(h);
        firstname = false;
    }
    *out = g_hostnames.data();
    *out_len = g_hostnames.size();
    return TK_OK;
}

// ---- renderer records -------------------------------------------------------

static thread_local std::string g_record;  // result of the record calls below

int tk_render_record(const char *struct_name, size_t struct_name_len, const char *settings_json, size_t settings_len,
                     const char **out, size_t *out_len) {
    if (!out || !out_len) return set_err(nullptr, TK_EARG, "tk_render_record: NULL argument");
    const RecordSpec *spec = find_record(sv(struct_name, struct_name_len));
    if (!spec) return set_err(nullptr, TK_EARG, "tk_render_record: no such record struct: " + std::string(sv(struct_name, struct_name_len)));
    Node *settings = nullptr;
    std::string err;
    if (settings_len && !parse_value(sv(settings_json, settings_len), false, &settings, &err)) return set_err(nullptr, TK_ESYNTAX, err);
    g_record.clear();
    const bool ok = render_record(*spec, settings, &g_record, &err);
    free_node(settings);
    if (!ok) return set_err(nullptr, TK_EARG, err);
    *out = g_record.data();
    *out_len = g_record.size();
    return TK_OK;
}

int tk_record_struct_name(const char *kind, const char *provider, const char **out, size_t *out_len) {
    if (!kind || !provider || !out || !out_len) return set_err(nullptr, TK_EARG, "tk_record_struct_name: NULL argument");
    g_record = record_struct_name(kind, provider);
    if (g_record.empty()) return set_err(nullptr, TK_EARG, std::string("no record struct for ") + kind + " / " + provider);
    *out = g_record.data();
    *out_len = g_record.size();
    return TK_OK;
}

int tk_module_source(const char *kind, const char *provider, const char *source_url, const char *source_ref, const char **out,
                     size_t *out_len) {
    if (!kind || !provider || !out || !out_len) return set_err(nullptr, TK_EARG, "tk_module_source: NULL argument");
    if (!module_source(kind, provider, source_url ? source_url : "", source_ref ? source_ref : "", &g_record))
        return set_err(nullptr, TK_EARG, std::string("no terraform module for ") + kind + " / " + provider);
    *out = g_record.data();
    *out_len = g_record.size();
    return TK_OK;
}

int tk_record_table(size_t index, const char **struct_name, const char **fields_json, size_t *fields_len) {
    const RecordSpec *spec = record_at(index);
    if (!spec) return TK_EARG;
    g_record.clear();
    record_table_json(*spec, &g_record);
    if (struct_name) *struct_name = spec->go_name;
    if (fields_json) *fields_json = g_record.data();
    if (fields_len) *fields_len = g_record.size();
    return TK_OK;
}

}  // extern "C"
