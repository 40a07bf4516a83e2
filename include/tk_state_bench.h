/*
 * tk_state_bench.h -- measurement hooks of libtkstate.
 *
 * None of these has a counterpart in the reference and none belongs to the drop-in surface
 * (include/tk_state.h): bench.py, tools/op_mix.py and the tests use them to time kernels, count
 * launches and copies, and to run one classification step in isolation.
 */
#ifndef TK_STATE_BENCH_H
#define TK_STATE_BENCH_H

#include "tk_state.h"

#ifdef __cplusplus
extern "C" {
#endif

/* The classification step of Clusters()/Nodes() alone: matching module keys are counted and
 * their indices left in device memory; no host copy of the pairs is built. */
TK_API int tk_state_prefix_count_device(tk_state *s, const char *prefix, size_t prefix_len, size_t *count);

/* Measurement hooks for bench.py.  With timing enabled every kernel launch on this handle is
 * bracketed by CUDA events on the handle's stream; tk_state_kernel_stats synchronises the stream
 * and reports, for kernel number `kind` (0 .. until it returns TK_EARG), its name, the summed
 * device time of the launches made while timing was on, how many those were, and all launches. */
TK_API int tk_state_timing(tk_state *s, int enable);
/* Handles created after this call start with timing on (or off): lets state.New itself be timed. */
TK_API int tk_set_default_timing(int enable);
/* state.New for handles parsed after this call: 1 (default) tries the line-by-line pass for text in
 * json.MarshalIndent form first, 0 sends everything through the general pipeline.  Both build the same
 * tape and give the same verdict; the switch exists so that tests and measurements can run both. */
TK_API int tk_set_line_pass(int enable);
TK_API int tk_state_kernel_stats(tk_state *s, int kind, const char **name, double *ms,
                                 unsigned long long *timed_launches, unsigned long long *launches);
/* Bytes this handle has copied host->device and device->host so far. */
TK_API int tk_state_copy_stats(tk_state *s, unsigned long long *h2d_bytes, unsigned long long *d2h_bytes);

/* Number of kernels this library has launched in this process (for bench.py's gpu_launches). */
TK_API unsigned long long tk_kernel_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* TK_STATE_BENCH_H */
