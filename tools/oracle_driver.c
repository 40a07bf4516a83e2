/* oracle_driver.c -- the benchmark step and the per-operation loops of tools/bench_driver.c over
 * the CPU oracle (oracle/tk_oracle.c), through function pointers bench.py / tools/op_mix.py hand
 * in.  Links neither the oracle nor the product: the process that runs `bench.py --impl reference`
 * loads this library and the oracle, and nothing else of this repository.  Bench infrastructure.
 */
#define _GNU_SOURCE
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <time.h>

static double now_ms(void) {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return t.tv_sec * 1e3 + t.tv_nsec * 1e-6;
}

typedef int (*ora_get_fn)(void *, const char *, size_t, const char **, size_t *);
typedef int (*ora_add_node_fn)(void *, const char *, size_t, const char *, size_t, const char *, size_t, int, char *, size_t);
typedef int (*ora_delete_fn)(void *, const char *, size_t, char *, size_t);
typedef int (*ora_nodes_fn2)(void *, const char *, size_t, char **, size_t *, size_t *, char *, size_t);
typedef void (*ora_free_buf_fn2)(char *);

double tkb_ora_get_loop(void *fn, void *h, const char *const *paths, const size_t *lens, size_t n, size_t total, size_t *value_bytes) {
    ora_get_fn f = (ora_get_fn)fn;
    const char *v;
    size_t vl, sum = 0;
    double t0 = now_ms();
    for (size_t i = 0, k = 0; i < total; i++, k = (k + 1 == n ? 0 : k + 1)) {
        if (f(h, paths[k], lens[k], &v, &vl)) return -1.0;
        sum += vl;
    }
    *value_bytes = sum;
    return (now_ms() - t0) * 1e-3;
}

double tkb_ora_add_node_loop(void *fn, void *h, const char *const *cks, const size_t *ck_lens, const char *const *names,
                             const size_t *name_lens, const char *const *jsons, const size_t *json_lens, size_t n, int is_record) {
    ora_add_node_fn f = (ora_add_node_fn)fn;
    char err[256];
    double t0 = now_ms();
    for (size_t i = 0; i < n; i++)
        if (f(h, cks[i], ck_lens[i], names[i], name_lens[i], jsons[i], json_lens[i], is_record, err, sizeof err)) return -1.0;
    return (now_ms() - t0) * 1e-3;
}

double tkb_ora_delete_loop(void *fn, void *h, const char *const *paths, const size_t *lens, size_t n) {
    ora_delete_fn f = (ora_delete_fn)fn;
    char err[256];
    double t0 = now_ms();
    for (size_t i = 0; i < n; i++)
        if (f(h, paths[i], lens[i], err, sizeof err)) return -1.0;
    return (now_ms() - t0) * 1e-3;
}

double tkb_ora_nodes_loop(void *fn_nodes, void *fn_free, void *h, const char *const *cks, const size_t *ck_lens, size_t n, size_t *pairs) {
    ora_nodes_fn2 f = (ora_nodes_fn2)fn_nodes;
    ora_free_buf_fn2 fr = (ora_free_buf_fn2)fn_free;
    char err[256], *out;
    size_t out_len, count, sum = 0;
    double t0 = now_ms();
    for (size_t i = 0; i < n; i++) {
        if (f(h, cks[i], ck_lens[i], &out, &out_len, &count, err, sizeof err)) return -1.0;
        fr(out);
        sum += count;
    }
    *pairs = sum;
    return (now_ms() - t0) * 1e-3;
}

double tkb_ora_destroy_sweep(void *fn_nodes, void *fn_free, void *fn_delete, void *h, const char *const *cks, const size_t *ck_lens,
                             const char *const *cluster_paths, const size_t *cluster_path_lens, size_t n, size_t *deleted) {
    ora_nodes_fn2 f = (ora_nodes_fn2)fn_nodes;
    ora_free_buf_fn2 fr = (ora_free_buf_fn2)fn_free;
    ora_delete_fn del = (ora_delete_fn)fn_delete;
    char err[256], *out, path[512];
    size_t out_len, count, done = 0;
    double t0 = now_ms();
    memcpy(path, "module.", 7);
    for (size_t i = 0; i < n; i++) {
        if (f(h, cks[i], ck_lens[i], &out, &out_len, &count, err, sizeof err)) return -1.0;
        for (size_t at = 0, k = 0; k < count; k++) {
            uint32_t fl, kl;
            memcpy(&fl, out + at, 4); at += 4 + fl;
            memcpy(&kl, out + at, 4); at += 4;
            if (kl + 7 >= sizeof path) { fr(out); return -2.0; }
            memcpy(path + 7, out + at, kl); at += kl;
            if (del(h, path, 7 + kl, err, sizeof err)) { fr(out); return -1.0; }
            done++;
        }
        fr(out);
        if (del(h, cluster_paths[i], cluster_path_lens[i], err, sizeof err)) return -1.0;
        done++;
    }
    *deleted = done;
    return (now_ms() - t0) * 1e-3;
}

/* The same step through the CPU oracle (oracle/tk_oracle.c), via its exported functions. */
typedef void *(*ora_new_fn)(const char *, size_t, char *, size_t);
typedef void (*ora_free_fn)(void *);
typedef void (*ora_free_buf_fn)(char *);
typedef int (*ora_clusters_fn)(void *, char **, size_t *, size_t *, char *, size_t);
typedef int (*ora_nodes_fn)(void *, const char *, size_t, char **, size_t *, size_t *, char *, size_t);
typedef int (*ora_bytes_fn)(void *, int, char **, size_t *);

int tkb_oracle_step(void **fns, const char *text, size_t n, const char *const *cks, const size_t *ck_lens, size_t n_cks,
                    size_t want_clusters, size_t want_nodes, size_t want_len) {
    ora_new_fn f_new = (ora_new_fn)fns[0];
    ora_free_fn f_free = (ora_free_fn)fns[1];
    ora_free_buf_fn f_free_buf = (ora_free_buf_fn)fns[2];
    ora_clusters_fn f_clusters = (ora_clusters_fn)fns[3];
    ora_nodes_fn f_nodes = (ora_nodes_fn)fns[4];
    ora_bytes_fn f_bytes = (ora_bytes_fn)fns[5];
    char err[512];
    char *out;
    size_t out_len, count, nodes = 0;
    void *h = f_new(text, n, err, sizeof err);
    if (!h) return 1;
    if (f_clusters(h, &out, &out_len, &count, err, sizeof err)) { f_free(h); return 1; }
    f_free_buf(out);
    if (count != want_clusters) { f_free(h); return -1; }
    for (size_t i = 0; i < n_cks; i++) {
        if (f_nodes(h, cks[i], ck_lens[i], &out, &out_len, &count, err, sizeof err)) { f_free(h); return 1; }
        f_free_buf(out);
        nodes += count;
    }
    if (nodes != want_nodes) { f_free(h); return -1; }
    f_bytes(h, 1, &out, &out_len);
    int ok = out_len == want_len;
    f_free_buf(out);
    f_free(h);
    return ok ? 0 : -1;
}
