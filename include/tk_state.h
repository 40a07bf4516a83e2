/*
 * tk_state.h -- C ABI of libtkstate: the state.State path of triton-kubernetes,
 * B200-native (CUDA, sm_100a).
 *
 * The reference has no FFI boundary of its own: state.State is a concrete Go
 * type called directly from create/, destroy/, get/, shell/ and backend/.  This header is the
 * boundary a maintainer would bind with cgo to swap that type's body for this library; each
 * entry point cites the reference method it replaces (paths relative to the reference root).
 * INTEGRATION.md shows the cgo stub.  The measurement hooks bench.py uses (per-kernel timing,
 * copy and launch counters) are not part of the drop-in surface: they live in tk_state_bench.h.
 *
 * Conventions
 *   - Every function returns TK_OK (0) or a TK_E* code; the Go error text (err.Error()) of the
 *     failing call is available verbatim from tk_last_error(handle) -- or tk_last_error(NULL)
 *     for a failed tk_state_new -- until the next call on that handle / thread.
 *   - Strings are (pointer, length) pairs and may contain any byte, NUL included.
 *   - Inputs are borrowed for the duration of the call only.
 *   - Returned buffers belong to the handle and stay valid until the next call that returns the
 *     same kind of buffer on that handle, or tk_state_free.
 *   - A handle is not thread-safe (neither is state.State: plain Go maps); distinct handles are
 *     independent.  A handle binds to the CUDA device that was current when it was created.
 *   - Values (the `obj interface{}` arguments of the Go methods) cross the boundary as JSON
 *     text, i.e. what json.Marshal(obj) gives on the Go side, plus a flag that says whether obj
 *     was a map (TK_VALUE_MAP: keys sorted on output, visible to Get/Clusters/Nodes) or a
 *     struct (TK_VALUE_RECORD: member order kept as given, opaque until the document is
 *     re-parsed -- SURVEY.md 8a note 1, create/cluster.go:146-152).
 *   - The bulk passes run on the GPU and have no CPU fallback: parsing a document (New),
 *     serialising it (Bytes) and, for a document that New left on the device, classifying the
 *     children of "module" (Clusters, Nodes).  Without a usable CUDA device they fail with
 *     TK_ECUDA and say so.
 */
#ifndef TK_STATE_H
#define TK_STATE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define TK_API __attribute__((visibility("default")))
#else
#define TK_API
#endif

typedef struct tk_state tk_state;

enum {
    TK_OK = 0,
    TK_ESYNTAX = 1,   /* json.Unmarshal failed (state.New, or a value argument) */
    TK_ECOLLISION = 2,/* gabs.ErrPathCollision */
    TK_ENOTOBJ = 3,   /* gabs.ErrNotObj */
    TK_EKEYPARTS = 4, /* getClusterKeyParts: cluster key has fewer than three '_' parts */
    TK_ECUDA = 5,     /* no device, or a CUDA call failed */
    TK_ELIMIT = 6,    /* document exceeds a documented size limit */
    TK_EARG = 7,      /* NULL where a pointer is required */
    TK_EIO = 8        /* local backend: file system error */
};

enum { TK_VALUE_MAP = 0, TK_VALUE_RECORD = 1 };

/* Size limits (documented in DESIGN.md): serialised document and string arena < 4 GiB each. */
#define TK_MAX_DEPTH 10000

/* state.New(name, raw) -- state/state.go:15-25 (gabs.ParseJSON, gabs.go:425-433). */
TK_API int tk_state_new(const char *name, size_t name_len, const char *raw, size_t raw_len, tk_state **out);

/* state.New(name, []byte("{}")) -- what backend/local/backend.go:54 and backend/manta/backend.go:131
 * do for a manager that has no file yet.  Pure host work: needs no device, so the path
 * operations (Get/Set/Add/Delete) can be exercised on a machine without a GPU. */
TK_API int tk_state_new_empty(const char *name, size_t name_len, tk_state **out);

/* Releases the document, its device buffers and its stream. */
TK_API void tk_state_free(tk_state *s);

/* State.Name -- state/state.go:11. */
TK_API int tk_state_name(tk_state *s, const char **name, size_t *name_len);

/* err.Error() of the last failing call; "" if none.  s may be NULL (thread-local slot). */
TK_API const char *tk_last_error(tk_state *s);

/* State.Get(path) -- state/state.go:27-34: the string at the dotted path, or "" for every
 * kind of miss (never an error). */
TK_API int tk_state_get(tk_state *s, const char *path, size_t path_len, const char **value, size_t *value_len);

/* configJSON.SetP(obj, path) -- the body shared by SetManager / SetTerraformBackendConfig /
 * AddCluster / AddNode (gabs.go:180-208). */
TK_API int tk_state_set(tk_state *s, const char *path, size_t path_len, const char *json, size_t json_len, int value_kind);

/* State.SetManager(obj) -- state/state.go:36-43 (path "module.cluster-manager"). */
TK_API int tk_state_set_manager(tk_state *s, const char *json, size_t json_len, int value_kind);

/* State.SetTerraformBackendConfig(path, obj) -- state/state.go:45-52. */
TK_API int tk_state_set_terraform_backend_config(tk_state *s, const char *path, size_t path_len, const char *json,
                                                 size_t json_len, int value_kind);

/* State.AddCluster(provider, name, obj) -- state/state.go:55-62. */
TK_API int tk_state_add_cluster(tk_state *s, const char *provider, size_t provider_len, const char *name,
                                size_t name_len, const char *json, size_t json_len, int value_kind);

/* State.AddNode(clusterKey, name, obj) -- state/state.go:65-77. */
TK_API int tk_state_add_node(tk_state *s, const char *cluster_key, size_t cluster_key_len, const char *name,
                             size_t name_len, const char *json, size_t json_len, int value_kind);

/* State.Delete(path) -- state/state.go:79-86 (gabs.go:276-301). */
TK_API int tk_state_delete(tk_state *s, const char *path, size_t path_len);

/* State.Bytes() -- state/state.go:88-90: json.MarshalIndent(doc, "", "\t"); "{}" for a nil
 * root.  Serialised on the GPU; *out is host memory owned by the handle. */
TK_API int tk_state_bytes(tk_state *s, const char **out, size_t *out_len);

/* State.Clusters() -- state/state.go:94-113 and State.Nodes(clusterKey) -- :117-147.
 * The Go map[string]string result comes back as *count pairs packed into *blob as
 *   { uint32 name_len; name bytes; uint32 key_len; key bytes }   (little endian)
 * sorted by name (Go's map order is random; on duplicate names the bytewise greatest module
 * key is kept, where Go keeps a random one).  For a document that lives on the device the
 * children of "module" are classified and packed by one GPU pass per document version and each
 * call is a binary search over that index; once the host tree exists (after a point operation)
 * the same range lookup runs on the tree's ordered map. */
TK_API int tk_state_clusters(tk_state *s, const uint8_t **blob, size_t *blob_len, size_t *count);
TK_API int tk_state_nodes(tk_state *s, const char *cluster_key, size_t cluster_key_len, const uint8_t **blob,
                          size_t *blob_len, size_t *count);

/* getNewHostnames(existingNames, nodeName, nodesToAdd) -- create/node.go:350-380.
 * existing: n (pointer, length) pairs.  Result: the new names joined by '\n', handle-less and
 * owned by the calling thread until its next call of this function. */
TK_API int tk_new_hostnames(const char *const *existing, const size_t *existing_len, size_t n, const char *node_name,
                            size_t node_name_len, long long nodes_to_add, const char **out, size_t *out_len);

/* ---- renderer records -- SURVEY.md 8a row 14 ------------------------------------------------
 * The reference builds one Go struct per provider x {manager, cluster, node} and hands it to
 * SetManager / AddCluster / AddNode (base structs create/manager.go:16-27, create/cluster.go:24-43,
 * create/node.go:19-41; provider structs in create/{manager,cluster,node}_<provider>.go).
 * tk_render_record turns a flat JSON object of settings {json field name: value} into that
 * struct's JSON text as encoding/json would marshal it -- declaration order, embedded base struct
 * first, omitempty applied -- ready to pass on with TK_VALUE_RECORD.  Results are owned by the
 * calling thread until its next record call. */
TK_API int tk_render_record(const char *struct_name, size_t struct_name_len, const char *settings_json, size_t settings_len,
                            const char **out, size_t *out_len);
/* ("manager"|"cluster"|"node", "triton"|"aws"|"gcp"|"azure"|"baremetal"|"vsphere") -> Go struct name */
TK_API int tk_record_struct_name(const char *kind, const char *provider, const char **out, size_t *out_len);
/* `<url>//terraform/modules/<dir>?ref=<ref>` -- create/manager.go:160-171, create/cluster.go:299-311,
 * create/node.go:209-221.  NULL url / ref: the defaults github.com/joyent/triton-kubernetes, master. */
TK_API int tk_module_source(const char *kind, const char *provider, const char *source_url, const char *source_ref,
                            const char **out, size_t *out_len);
/* table number `index` (0 .. until TK_EARG): struct name and its fields as JSON, for the tests */
TK_API int tk_record_table(size_t index, const char **struct_name, const char **fields_json, size_t *fields_len);

/* ---- local backend -- backend/local/backend.go:14-132 ------------------
 * backend.Backend (backend/backend.go:7-27) over <root>/<name>/main.tf.json.  root_dir NULL means
 * ~/.triton-kubernetes as in the reference; tests point it elsewhere. */
typedef struct tk_backend tk_backend;
TK_API int tk_local_backend_new(const char *root_dir, size_t root_dir_len, tk_backend **out);   /* local.New, :28-41 */
TK_API void tk_backend_free(tk_backend *b);
TK_API const char *tk_backend_last_error(tk_backend *b);                                       /* b may be NULL */
TK_API int tk_backend_state(tk_backend *b, const char *name, size_t name_len, tk_state **out); /* State, :43-66 */
TK_API int tk_backend_delete_state(tk_backend *b, const char *name, size_t name_len);          /* DeleteState, :68-77 */
TK_API int tk_backend_persist_state(tk_backend *b, tk_state *s);                               /* PersistState, :79-100 */
/* States, :102-121: directory names joined by '\n', sorted */
TK_API int tk_backend_states(tk_backend *b, const char **names, size_t *names_len, size_t *count);
/* StateTerraformConfig, :123-132: the path "terraform.backend.local" and the record {"path": ...} */
TK_API int tk_backend_state_terraform_config(tk_backend *b, const char *name, size_t name_len, const char **path,
                                             size_t *path_len, const char **record_json, size_t *record_len);

/* ---- device-resident variants: what bench.py times with inputs already in HBM ------------ */

/* state.New on text that is already in device memory (dev_raw: device pointer, borrowed for the
 * call).  Same result as tk_state_new; no host->device copy of the text. */
TK_API int tk_state_new_device(const char *name, size_t name_len, const void *dev_raw, size_t raw_len, tk_state **out);

/* Makes the device copy of the document current (flatten + host->device copy if the document
 * changed since the last staging).  Bytes/Clusters/Nodes call this implicitly. */
TK_API int tk_state_stage(tk_state *s);

/* State.Bytes() leaving the text in device memory: *dev_out is a device pointer owned by the
 * handle.  Runs only the kernels (the document must have been staged). */
TK_API int tk_state_bytes_device(tk_state *s, const void **dev_out, size_t *out_len);

/* Allocation-free line-offset pass.  (No counterpart in the reference: this entry point exists
 * only on the GPU side of the boundary.)
 *
 * For callers that cannot let a library allocate GPU memory: programs that gave all of it to
 * their own pool at start-up, and latency-critical paths where cudaMalloc/cudaFree are unwelcome
 * because they may synchronise the device and stall other streams.  It runs the first pass of
 * State.Bytes() -- the escaped length of every output line and the running sum of those lengths
 * -- on a line tape the caller already holds in device memory (layout: csrc/tape.hpp), writing
 * n_rows + 1 offsets to off_out (off_out[r] = byte offset of line r, off_out[n_rows] = length of
 * the whole text).  All pointers except err are device pointers owned by the caller; the row
 * arrays must be 16-byte aligned and arena readable for 16 bytes past arena_bytes.  The work is enqueued on `stream` (a cudaStream_t,
 * NULL = default stream) and the call returns without waiting for it.
 *
 * It calls no cudaMalloc, cudaFree or other allocating routine.  The scan's scratch (a tile
 * counter, a total and one look-back status word per tile of 2,048 rows) is a fixed array inside
 * the library, 8,208 bytes on every GPU that loads it, which limits n_rows to TK_NOALLOC_MAX_ROWS
 * (1,024 tiles); larger tapes are rejected with TK_ELIMIT and a message in err.  Things to know:
 *   - the first call on each device looks up the array's device address; make that first call
 *     with the device current and before any stream capture begins;
 *   - every call on a GPU uses the same array, which keeps the previous call's contents: the
 *     part a call needs is cleared on `stream` first, and two calls whose GPU work overlaps
 *     (different streams, or different threads) race on it -- serialise them;
 *   - offsets are 32-bit.  The length of the whole text is summed in 64 bits inside the library's
 *     scratch, where the caller cannot see it, so a text of 4 GiB or more is reported in band:
 *     off_out[n_rows] is then 0xffffffff and the other offsets have wrapped and must not be used
 *     (a real length never reaches 0xffffffff: tk_state_bytes refuses such documents too).
 * err (may be NULL) receives the error text; nothing is allocated for it either. */
#define TK_NOALLOC_MAX_ROWS 2097152u
TK_API int tk_row_offsets_noalloc(const uint32_t *meta, const uint32_t *koff, const uint32_t *klen,
                                  const uint32_t *voff, const uint32_t *vlen, const uint8_t *arena,
                                  uint32_t arena_bytes, uint32_t n_rows, uint32_t *off_out, void *stream,
                                  char *err, size_t err_cap);

/* 1 if the document's host tree exists.  After state.New on text whose object keys are already
 * in encoding/json's output order (strictly ascending, hence unique) the document lives only as
 * the line tape on the device and Bytes() runs straight from it; the tree is built on the first
 * point operation (Get/Set/Add/Delete/Clusters/Nodes), or at once when the keys need re-ordering. */
TK_API int tk_state_host_tree_built(tk_state *s);

/* The handle's CUDA stream (a cudaStream_t), so a caller can time or order work on it. */
TK_API void *tk_state_stream(tk_state *s);

/* Library/ABI version string. */
TK_API const char *tk_version(void);

/* Selects the CUDA device for the calling thread: handles the thread creates from now on bind to
 * it (a handle binds to the device current at creation).  For callers that have no CUDA binding
 * of their own -- a Go program driving several GPUs, one goroutine locked to a thread per device.
 * No counterpart in the reference.  TK_ECUDA (text from tk_last_error(NULL)) if there is no such
 * device. */
TK_API int tk_set_device(int device);

#ifdef __cplusplus
}
#endif
#endif /* TK_STATE_H */
