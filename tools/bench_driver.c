/* bench_driver.c -- the benchmark step as a compiled loop over the C ABI.
 *
 * bench.py drives the library from Python; a Python-level loop of 1,000 Nodes() calls costs
 * ~3 us of ctypes overhead per call, which would be measured as if it were the library.  This
 * driver runs one whole step (New -> Clusters -> Nodes(ck) for every cluster key -> Bytes) in C,
 * for the product, through include/tk_state.h.  The same loops over the CPU oracle live in
 * tools/oracle_driver.c, a library of its own, so that the process that times the reference arm
 * never maps libtkstate.so.  Bench infrastructure, not product.
 */
#define _GNU_SOURCE
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <time.h>

#include "tk_state.h"
#include "tk_state_bench.h"

static double now_ms(void) {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return t.tv_sec * 1e3 + t.tv_nsec * 1e-6;
}

/* One step through the product.  text: host pointer (device_text == 0) or device pointer.
 * to_host: Bytes() into host memory (e2e) or left on the device (value).
 * phase_ms[5]: new, clusters, nodes, bytes, (unused).  Returns 0 on success, else a TK_E* code
 * (or -1 for a wrong result).  *keep receives the handle (the caller frees it after its events). */
int tkb_product_step(const void *text, size_t n, int device_text, int to_host, const char *const *cks, const size_t *ck_lens,
                     size_t n_cks, size_t want_clusters, size_t want_nodes, double *phase_ms, tk_state **keep,
                     const char **out_text, size_t *out_len) {
    tk_state *s = NULL;
    const uint8_t *blob;
    size_t blob_len, count, nodes = 0;
    double t0 = now_ms();
    int rc = device_text ? tk_state_new_device("bench", 5, text, n, &s) : tk_state_new("bench", 5, (const char *)text, n, &s);
    if (rc) return rc;
    *keep = s;
    double t1 = now_ms();
    rc = tk_state_clusters(s, &blob, &blob_len, &count);
    if (rc) return rc;
    if (count != want_clusters) return -1;
    double t2 = now_ms();
    for (size_t i = 0; i < n_cks; i++) {
        rc = tk_state_nodes(s, cks[i], ck_lens[i], &blob, &blob_len, &count);
        if (rc) return rc;
        nodes += count;
    }
    if (nodes != want_nodes) return -1;
    double t3 = now_ms();
    if (to_host) rc = tk_state_bytes(s, out_text, out_len);
    else rc = tk_state_bytes_device(s, (const void **)out_text, out_len);
    if (rc) return rc;
    double t4 = now_ms();
    phase_ms[0] += t1 - t0; phase_ms[1] += t2 - t1; phase_ms[2] += t3 - t2; phase_ms[3] += t4 - t3;
    return 0;
}

/* ---- `blocks` timed blocks of `steps` whole steps each, with `inflight` documents in flight: that
 * many threads, each running whole steps on its own handle (own stream, own buffers).  The threads
 * first run the warm-up steps; then, for every block, they meet at a barrier, the caller's
 * `at_start(block)` runs on the calling thread (bench.py synchronises the device and records its
 * CUDA start event there, on an idle but still warm device), the steps of the block are handed out
 * from a shared counter, and when every thread has finished its last one `at_end(block)` runs.
 * texts[i % n_texts] is the input of step i.
 * pageable != 0 (host text only): the text is what a Go caller holds -- plain malloc'd memory, not
 * pinned -- and the result of Bytes() is copied out of the handle's buffer into the worker's own
 * malloc'd buffer, as C.GoBytes does.
 * Returns 0, or the first failing step's code. */
#include <pthread.h>
#include <sched.h>
#include <stdlib.h>

typedef struct {
    const void *const *texts; size_t n_texts, n; int device, device_text, to_host, pageable;
    const char *const *cks; const size_t *ck_lens; size_t n_cks, want_clusters, want_nodes;
    size_t steps, blocks, warmup_each; int worker;
    pthread_barrier_t *go, *done;
    size_t *next; pthread_mutex_t *mu; int *rc;
    double phase_ms[5]; size_t out_len; unsigned long long h2d, d2h;
    const char *expect_head, *expect_tail; size_t expect_n;
    char *copy_out; size_t copy_cap;
    const int *cpus; size_t n_cpus;   /* the cores this rank's workers stay on (n_cpus 0: wherever the scheduler puts them) */
} tkb_worker;

static int tkb_one(tkb_worker *w, size_t i, double *phase) {
    tk_state *s = NULL;
    const char *out = NULL;
    size_t out_len = 0;
    int rc = tkb_product_step(w->texts[i % w->n_texts], w->n, w->device_text, w->to_host, w->cks, w->ck_lens, w->n_cks, w->want_clusters,
                              w->want_nodes, phase, &s, &out, &out_len);
    if (!rc && out_len != w->expect_n) rc = -1;
    if (!rc && w->to_host && w->pageable) {
        if (w->copy_cap < out_len) { free(w->copy_out); w->copy_out = (char *)malloc(out_len ? out_len : 1); w->copy_cap = out_len; }
        if (!w->copy_out) rc = -3;
        else { memcpy(w->copy_out, out, out_len); out = w->copy_out; }
    }
    if (!rc && w->to_host && w->expect_n >= 64 &&
        (memcmp(out, w->expect_head, 64) != 0 || memcmp(out + out_len - 64, w->expect_tail, 64) != 0)) rc = -1;
    if (s) {
        tk_state_copy_stats(s, &w->h2d, &w->d2h);
        tk_state_free(s);   /* waits for the handle's stream: the emit kernel of a resident step is still queued */
    }
    return rc;
}

static void *tkb_worker_main(void *arg) {
    tkb_worker *w = (tkb_worker *)arg;
    double scratch_phase[5] = {0, 0, 0, 0, 0};
    if (w->n_cpus) {
        /* one core per worker, out of this rank's share: spinning waits of one rank do not push another rank's
         * workers off their cores, and the pinned buffers a worker allocates are first touched near where it runs */
        cpu_set_t set;
        CPU_ZERO(&set);
        CPU_SET(w->cpus[(size_t)w->worker % w->n_cpus], &set);
        pthread_setaffinity_np(pthread_self(), sizeof set, &set);
    }
    int rc = tk_set_device(w->device);
    for (size_t k = 0; !rc && k < w->warmup_each; k++) rc = tkb_one(w, (size_t)w->worker + k * 64, scratch_phase);
    for (size_t b = 0; b < w->blocks; b++) {
        pthread_barrier_wait(w->done);   /* warm-up / previous block finished everywhere */
        pthread_barrier_wait(w->go);     /* the caller has started its clock */
        while (!rc) {
            pthread_mutex_lock(w->mu);
            const size_t i = (*w->next)++;
            const int stop = *w->rc != 0;
            pthread_mutex_unlock(w->mu);
            if (i >= (b + 1) * w->steps || stop) {
                if (i >= (b + 1) * w->steps) { pthread_mutex_lock(w->mu); (*w->next)--; pthread_mutex_unlock(w->mu); }
                break;
            }
            rc = tkb_one(w, i, w->phase_ms);
        }
        if (rc) { pthread_mutex_lock(w->mu); if (!*w->rc) *w->rc = rc; pthread_mutex_unlock(w->mu); }
    }
    pthread_barrier_wait(w->done);
    free(w->copy_out);
    return NULL;
}

int tkb_product_run(const void *const *texts, size_t n_texts, size_t n, int device, int device_text, int to_host, int pageable,
                    const char *const *cks, const size_t *ck_lens, size_t n_cks, size_t want_clusters, size_t want_nodes, size_t steps,
                    size_t blocks, size_t warmup, size_t inflight, const char *expect_head, const char *expect_tail, size_t expect_n,
                    void (*at_start)(int), void (*at_end)(int), const int *cpus, size_t n_cpus, double *phase_ms, unsigned long long *h2d,
                    unsigned long long *d2h) {
    if (inflight < 1) inflight = 1;
    if (inflight > 64) inflight = 64;
    if (blocks < 1) blocks = 1;
    pthread_barrier_t go, done;
    pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    pthread_t th[64];
    tkb_worker w[64];
    size_t next = 0;
    int rc = 0;
    pthread_barrier_init(&go, NULL, (unsigned)inflight + 1);
    pthread_barrier_init(&done, NULL, (unsigned)inflight + 1);
    for (size_t t = 0; t < inflight; t++) {
        memset(&w[t], 0, sizeof w[t]);
        w[t].texts = texts; w[t].n_texts = n_texts; w[t].n = n; w[t].device = device; w[t].device_text = device_text; w[t].to_host = to_host;
        w[t].pageable = pageable;
        w[t].cks = cks; w[t].ck_lens = ck_lens; w[t].n_cks = n_cks; w[t].want_clusters = want_clusters; w[t].want_nodes = want_nodes;
        w[t].steps = steps; w[t].blocks = blocks;
        w[t].warmup_each = (warmup + inflight - 1) / inflight; if (!w[t].warmup_each) w[t].warmup_each = 1;
        w[t].worker = (int)t; w[t].go = &go; w[t].done = &done; w[t].next = &next; w[t].mu = &mu; w[t].rc = &rc;
        w[t].expect_head = expect_head; w[t].expect_tail = expect_tail; w[t].expect_n = expect_n;
        w[t].cpus = cpus; w[t].n_cpus = n_cpus;
        pthread_create(&th[t], NULL, tkb_worker_main, &w[t]);
    }
    for (size_t b = 0; b < blocks; b++) {
        pthread_barrier_wait(&done);
        if (b && at_end) at_end((int)b - 1);
        if (at_start) at_start((int)b);
        pthread_barrier_wait(&go);
    }
    pthread_barrier_wait(&done);
    if (at_end) at_end((int)blocks - 1);
    for (size_t t = 0; t < inflight; t++) pthread_join(th[t], NULL);
    for (int k = 0; k < 5; k++) phase_ms[k] = 0;
    for (size_t t = 0; t < inflight; t++) {
        for (int k = 0; k < 5; k++) phase_ms[k] += w[t].phase_ms[k];
        if (w[t].h2d) { *h2d = w[t].h2d; *d2h = w[t].d2h; }
    }
    pthread_barrier_destroy(&go);
    pthread_barrier_destroy(&done);
    return rc;
}

/* ---- per-operation loops (SURVEY.md 8d, config #4 op mix; tools/op_mix.py).  Each returns the
 * seconds the loop took, or a negative number on the first failing call.  strs/lens are parallel
 * arrays; `total` calls cycle through the n entries in order. */
double tkb_get_loop(tk_state *s, const char *const *paths, const size_t *lens, size_t n, size_t total, size_t *value_bytes) {
    const char *v;
    size_t vl, sum = 0;
    double t0 = now_ms();
    for (size_t i = 0, k = 0; i < total; i++, k = (k + 1 == n ? 0 : k + 1)) {
        if (tk_state_get(s, paths[k], lens[k], &v, &vl)) return -1.0;
        sum += vl;
    }
    *value_bytes = sum;
    return (now_ms() - t0) * 1e-3;
}

double tkb_add_node_loop(tk_state *s, const char *const *cks, const size_t *ck_lens, const char *const *names, const size_t *name_lens,
                         const char *const *jsons, const size_t *json_lens, size_t n, int value_kind) {
    double t0 = now_ms();
    for (size_t i = 0; i < n; i++)
        if (tk_state_add_node(s, cks[i], ck_lens[i], names[i], name_lens[i], jsons[i], json_lens[i], value_kind)) return -1.0;
    return (now_ms() - t0) * 1e-3;
}

double tkb_delete_loop(tk_state *s, const char *const *paths, const size_t *lens, size_t n) {
    double t0 = now_ms();
    for (size_t i = 0; i < n; i++)
        if (tk_state_delete(s, paths[i], lens[i])) return -1.0;
    return (now_ms() - t0) * 1e-3;
}

double tkb_nodes_loop(tk_state *s, const char *const *cks, const size_t *ck_lens, size_t n, size_t *pairs) {
    const uint8_t *blob;
    size_t blob_len, count, sum = 0;
    double t0 = now_ms();
    for (size_t i = 0; i < n; i++) {
        if (tk_state_nodes(s, cks[i], ck_lens[i], &blob, &blob_len, &count)) return -1.0;
        sum += count;
    }
    *pairs = sum;
    return (now_ms() - t0) * 1e-3;
}

/* config #5 at library granularity: per cluster Nodes(ck), Delete("module." + key) for each pair, then
 * Delete of the cluster key.  cluster_paths[i] = "module." + cks[i]. */
double tkb_destroy_sweep(tk_state *s, const char *const *cks, const size_t *ck_lens, const char *const *cluster_paths,
                         const size_t *cluster_path_lens, size_t n, size_t *deleted) {
    const uint8_t *blob;
    size_t blob_len, count, done = 0;
    char path[512];
    double t0 = now_ms();
    memcpy(path, "module.", 7);
    for (size_t i = 0; i < n; i++) {
        if (tk_state_nodes(s, cks[i], ck_lens[i], &blob, &blob_len, &count)) return -1.0;
        /* the blob belongs to the handle and a Delete may recycle it: take the keys out first */
        size_t bytes = blob_len;
        uint8_t *copy = (uint8_t *)__builtin_alloca(bytes ? bytes : 1);
        memcpy(copy, blob, bytes);
        for (size_t at = 0, k = 0; k < count; k++) {
            uint32_t fl, kl;
            memcpy(&fl, copy + at, 4); at += 4 + fl;
            memcpy(&kl, copy + at, 4); at += 4;
            if (kl + 7 >= sizeof path) return -2.0;
            memcpy(path + 7, copy + at, kl); at += kl;
            if (tk_state_delete(s, path, 7 + kl)) return -1.0;
            done++;
        }
        if (tk_state_delete(s, cluster_paths[i], cluster_path_lens[i])) return -1.0;
        done++;
    }
    *deleted = done;
    return (now_ms() - t0) * 1e-3;
}
