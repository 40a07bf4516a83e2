#!/usr/bin/env python
"""bench.py -- state.State ops/sec on BASELINE.json config #4 (1,000 clusters x 100 nodes).

One STEP is one pass of the hot path over the 62,177,405-byte synthetic document:
    state.New(text) -> Clusters() -> Nodes(ck) for each of the 1,000 cluster keys -> Bytes()
= 1,003 state.State operations (destroy/cluster.go:64-161 and get/cluster.go:65-118 walk a
document this way).  The step itself is a compiled loop over the C ABI (tools/bench_driver.c), for
both arms, and so is the loop over steps with its worker threads (tkb_product_run), so that
Python call overhead and the GIL are not measured as if they were the library.

  value  the step with the text already resident in HBM and the result left in HBM
  e2e    the same step through the public C ABI with HOST buffers: pinned host text in, the
         host->device copy, kernels, and the device->host copy of the result text, all inside the
         timed region

  e2e_pageable  e2e as the reference's own callers would drive it (backend/local/backend.go:60-65,94):
         the text in plain malloc'd memory, the result copied out of the library's buffer

All are whole-job throughput: K steps with `--inflight` documents in flight per GPU (default 4) --
that many host threads each run whole steps on their own handle, i.e. their own stream and
buffers, so one document's copies and host-side waits overlap another's kernels.  The path shards
by document and the CPU arm runs one document per host thread the same way.  After
the warm-up the block of K steps is run BLOCKS (7) times back to back; every block is bracketed by
CUDA events recorded on an idle device before its first step and after every thread has waited for
its last one, and the reported number is the MEDIAN block (min and max beside it in
config.blocks).  The latency of a single document (one in flight; for `value` with L2 flushed
before every step) is reported in config.latency_ms_per_step, and the per-kernel times for the
roofline come from that serial, L2-flushed run.
  --impl reference   the CPU oracle (oracle/tk_oracle.c, a restatement of the reference's Go
         code; the Go toolchain is absent) on every host core, same step, same metric.  That
         process loads the oracle and tools/oracle_driver.c only, never the product library.

Under torchrun (N > 1) every rank runs the step on its own document on its own GPU (the path
shards by document: one manager per worker, no exchange step, so no collective in the data
path); time is the max over ranks, value the aggregate.
"""
import argparse
import ctypes
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from ctypes import POINTER, byref, c_char_p, c_double, c_int, c_size_t, c_void_p

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

OPS_PER_STEP = 1003
BLOCKS = 7   # timed blocks of --steps steps each; the median block is the reported one
WORKLOAD = ("config #4: synthetic state document, 1000 clusters x 100 nodes (62,177,405 B tab-indented); "
            "step = New + Clusters + 1000 x Nodes(ck) + Bytes = 1003 state.State ops")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--clusters", type=int, default=1000, help=argparse.SUPPRESS)
    ap.add_argument("--nodes", type=int, default=100, help=argparse.SUPPRESS)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help=argparse.SUPPRESS)
    ap.add_argument("--inflight", type=int, default=0,
                    help="documents in flight per GPU: that many host threads each run whole steps on their own handle and "
                         "stream (default: 4, fewer if the host has under 4 cores per rank)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------ driver

def _build(src, lib, extra=()):
    if not os.path.exists(lib) or os.path.getmtime(lib) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-pthread", "-I" + os.path.join(ROOT, "include"), "-o", lib, src] + list(extra))
    return lib


def driver():
    """tools/libtkbench.so, the step loop over the product's C ABI: built on first use (gcc is part
    of the image, here and on the GPU box)."""
    pkg = os.path.join(ROOT, "triton-kubernetes_b200")
    lib = _build(os.path.join(ROOT, "tools", "bench_driver.c"), os.path.join(ROOT, "tools", "libtkbench.so"),
                 ["-L" + pkg, "-ltkstate", "-Wl,-rpath," + pkg])
    importlib.import_module("triton-kubernetes_b200").state.lib()   # libtkstate first, by full path
    D = ctypes.CDLL(lib)
    D.tkb_product_step.argtypes = [c_void_p, c_size_t, c_int, c_int, POINTER(c_char_p), POINTER(c_size_t), c_size_t, c_size_t,
                                   c_size_t, POINTER(c_double), POINTER(c_void_p), POINTER(c_void_p), POINTER(c_size_t)]
    return D


def oracle_driver():
    """tools/libtkbench_oracle.so, the same loops over the CPU oracle through function pointers.
    Links nothing of this repository: the reference arm's process never maps libtkstate.so."""
    D = ctypes.CDLL(_build(os.path.join(ROOT, "tools", "oracle_driver.c"), os.path.join(ROOT, "tools", "libtkbench_oracle.so")))
    D.tkb_oracle_step.argtypes = [POINTER(c_void_p), c_char_p, c_size_t, POINTER(c_char_p), POINTER(c_size_t), c_size_t, c_size_t,
                                  c_size_t, c_size_t]
    return D


def load_synth():
    """triton-kubernetes_b200/synth.py by file path: the document generator is plain Python, and the
    reference arm must not import the product package."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("tk_synth", os.path.join(ROOT, "triton-kubernetes_b200", "synth.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def key_arrays(cks):
    arr = (c_char_p * len(cks))(*cks)
    lens = (c_size_t * len(cks))(*[len(k) for k in cks])
    return arr, lens


# ------------------------------------------------------------------------------------ clocks

class ClockSampler:
    """nvidia-smi clocks and throttle reasons sampled every 100 ms from before the first timed block to after the last."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------ CPU arm

class OracleArm:
    def __init__(self, text, cks, clusters, nodes):
        from oracle import oracle as O
        O.build()
        L = O.lib()
        self.D = oracle_driver()
        names = ["ora_new", "ora_free", "ora_free_buf", "ora_clusters", "ora_nodes", "ora_bytes"]
        self.fns = (c_void_p * len(names))(*[ctypes.cast(getattr(L, n), c_void_p).value for n in names])
        self.text, self.cks, self.clusters, self.nodes = text, cks, clusters, nodes
        self.arr, self.lens = key_arrays(cks)

    def step(self):
        rc = self.D.tkb_oracle_step(self.fns, self.text, len(self.text), self.arr, self.lens, len(self.cks), self.clusters,
                                    self.clusters * self.nodes, len(self.text))
        if rc:
            raise SystemExit("bench.py: oracle step failed (%d)" % rc)

    def run(self, threads, steps_per_thread=None, seconds=None):
        """Whole steps on `threads` host threads (ctypes releases the GIL); returns (steps, seconds)."""
        done = [0] * threads
        t0 = time.perf_counter()

        def work(i):
            while True:
                self.step()
                done[i] += 1
                if steps_per_thread is not None and done[i] >= steps_per_thread:
                    return
                if seconds is not None and time.perf_counter() - t0 >= seconds:
                    return

        ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
        for t in ts: t.start()
        for t in ts: t.join()
        return sum(done), time.perf_counter() - t0


def oracle_text(args):
    synth = load_synth()
    from oracle import oracle as O
    O.build()
    return O.New("bench", synth.document(args.clusters, args.nodes)).Bytes()


def run_reference(args, rank, world):
    if rank != 0:
        return
    synth = load_synth()
    text = oracle_text(args)
    cks = [synth.cluster_key(c).encode() for c in range(args.clusters)]
    arm = OracleArm(text, cks, args.clusters, args.nodes)
    threads = os.cpu_count() or 1
    arm.run(threads, steps_per_thread=min(args.warmup, 1))
    steps, dt = arm.run(threads, steps_per_thread=max(1, args.steps))
    value = steps * OPS_PER_STEP / dt
    line = {
        "impl": "reference", "metric": "state.State ops/sec", "value": value, "unit": "ops/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / max(1, args.steps) * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU oracle (C restatement of the reference's Go path; no Go toolchain here), "
                   "%d host threads each running %d whole steps on its own parse of the document" % (threads, max(1, args.steps))},
        "cpu_baseline": {"value": value, "unit": "ops/s", "cores": threads, "kind": "port",
                         "sample": "%d full passes in %.1f s" % (steps, dt)},
        "e2e": {"value": value, "unit": "ops/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------ GPU arm

def run_ours(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    tk = importlib.import_module("triton-kubernetes_b200")
    synth = importlib.import_module("triton-kubernetes_b200.synth")
    D = driver()
    L = tk.state.lib()
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        peaks = {}
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")

    # ---- synthetic input: the persisted (tab-indented) form of the document, in pinned memory
    seed = tk.New("bench", synth.document(args.clusters, args.nodes))
    text = seed.Bytes()
    seed.close()
    cks = [synth.cluster_key(c).encode() for c in range(args.clusters)]
    arr, lens = key_arrays(cks)
    pinned = torch.empty(len(text), dtype=torch.uint8, pin_memory=True)
    ctypes.memmove(pinned.data_ptr(), text, len(text))
    dev_text = torch.empty(len(text), dtype=torch.uint8, device=dev)
    dev_text.copy_(pinned)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    R_COPIES = 4   # the resident input rotates over 4 device copies (249 MB > L2) when steps overlap
    dev_copies = [dev_text] + [dev_text.clone() for _ in range(R_COPIES - 1)]
    torch.cuda.synchronize()

    def step(src_ptr, device_text, to_host, phase):
        """One step through the C driver; returns the handle (caller frees) and the result view."""
        h, out, n = c_void_p(), c_void_p(), c_size_t()
        rc = D.tkb_product_step(src_ptr, len(text), 1 if device_text else 0, 1 if to_host else 0, arr, lens, len(cks), args.clusters,
                                args.clusters * args.nodes, phase, byref(h), byref(out), byref(n))
        if rc or n.value != len(text):
            raise SystemExit("bench.py: step failed (%d): %s" % (rc, L.tk_last_error(h if h.value else None)))
        return h, out.value

    def kernel_stats(h, acc):
        kind = 0
        while True:
            name, ms, tl, ln = c_char_p(), c_double(), ctypes.c_ulonglong(), ctypes.c_ulonglong()
            if L.tk_state_kernel_stats(h, kind, byref(name), byref(ms), byref(tl), byref(ln)) != 0:
                return
            a = acc.setdefault(name.value.decode(), {"ms": 0.0, "timed": 0})
            a["ms"] += ms.value; a["timed"] += tl.value
            kind += 1

    def timed_resident(steps, per_kernel):
        """Serial steps (one document at a time), L2 flushed before each: latency and per-kernel times."""
        L.tk_set_default_timing(1 if per_kernel else 0)
        total, stats, phase = 0.0, {}, (c_double * 5)()
        for _ in range(steps):
            flush.zero_()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            h, _ = step(dev_text.data_ptr(), True, False, phase)
            b.record(torch.cuda.ExternalStream(L.tk_state_stream(h), device=dev))
            b.synchronize()
            total += a.elapsed_time(b)
            if per_kernel:
                kernel_stats(h, stats)
            L.tk_state_free(h)
        L.tk_set_default_timing(0)
        return total, stats

    AT_BLOCK = ctypes.CFUNCTYPE(None, c_int)
    D.tkb_product_run.argtypes = [POINTER(c_void_p), c_size_t, c_size_t, c_int, c_int, c_int, c_int, POINTER(c_char_p), POINTER(c_size_t),
                                  c_size_t, c_size_t, c_size_t, c_size_t, c_size_t, c_size_t, c_size_t, c_char_p, c_char_p, c_size_t,
                                  AT_BLOCK, AT_BLOCK, POINTER(c_int), c_size_t, POINTER(c_double), POINTER(ctypes.c_ulonglong),
                                  POINTER(ctypes.c_ulonglong)]
    # this rank's share of the host cores: its worker threads are pinned inside it (one core each)
    try:
        allowed = sorted(os.sched_getaffinity(0))
    except AttributeError:
        allowed = list(range(os.cpu_count() or 1))
    per_rank = max(1, len(allowed) // max(1, world))
    my_cpus = allowed[local_rank * per_rank:(local_rank + 1) * per_rank] if world > 1 else []
    cpu_arr = (c_int * max(1, len(my_cpus)))(*(my_cpus or [0]))
    dev_ptrs = (c_void_p * R_COPIES)(*[t.data_ptr() for t in dev_copies])
    host_ptrs = (c_void_p * 1)(pinned.data_ptr())
    # what a Go caller holds: the text in ordinary heap memory (ioutil.ReadFile's slice), not pinned
    libc = ctypes.CDLL(None)
    libc.malloc.restype, libc.malloc.argtypes = c_void_p, [c_size_t]
    heap_text = libc.malloc(len(text))
    ctypes.memmove(heap_text, text, len(text))
    heap_ptrs = (c_void_p * 1)(heap_text)

    def run_overlapped(steps, inflight, mode, warmup, blocks=BLOCKS):
        """`blocks` timed blocks of `steps` whole steps, `inflight` at a time: tools/bench_driver.c starts
        that many threads, each running whole steps on its own handle (own stream, own buffers).  The
        threads first run the warm-up steps; before every block they meet at a barrier, so that the
        clock starts on a device that is idle but has not had time to drop its clocks; every block
        is bracketed by CUDA events (start after that barrier and a device synchronize, end after
        every thread has waited for its last step and another synchronize).
        mode: "resident" (device text in, device text out), "pinned" or "pageable" (host text in and out).
        Returns (per-block ms, per-step phases, copy stats, kernel launches of one block)."""
        marks = [dict() for _ in range(blocks)]

        def at_start(b):
            barrier()
            m = marks[b]
            m["launches"] = tk.state.kernel_launches()
            m["a"], m["b"] = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            m["a"].record()
            m["t0"] = time.perf_counter()

        def at_end(b):
            m = marks[b]
            torch.cuda.synchronize()
            m["b"].record()
            m["b"].synchronize()
            m["wall"] = (time.perf_counter() - m["t0"]) * 1e3
            m["n_launches"] = tk.state.kernel_launches() - m["launches"]

        to_host = mode != "resident"
        ptrs, n_ptrs = {"resident": (dev_ptrs, R_COPIES), "pinned": (host_ptrs, 1), "pageable": (heap_ptrs, 1)}[mode]
        phase, h2d, d2h = (c_double * 5)(), ctypes.c_ulonglong(), ctypes.c_ulonglong()
        cb_a, cb_b = AT_BLOCK(at_start), AT_BLOCK(at_end)
        rc = D.tkb_product_run(ptrs, n_ptrs, len(text), local_rank, 0 if to_host else 1, 1 if to_host else 0,
                               1 if mode == "pageable" else 0, arr, lens, len(cks), args.clusters, args.clusters * args.nodes, steps,
                               blocks, warmup, max(1, inflight), text[:64], text[-64:], len(text), cb_a, cb_b, cpu_arr, len(my_cpus), phase,
                               byref(h2d), byref(d2h))
        if rc:
            raise SystemExit("bench.py: a step failed (%d): %s" % (rc, L.tk_last_error(None)))
        ms = [max(m["a"].elapsed_time(m["b"]), m["wall"]) for m in marks]
        return ms, [phase[i] / (steps * blocks) for i in range(4)], (h2d.value, d2h.value), marks[len(marks) // 2]["n_launches"]

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor(ms, device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.tolist()
        return list(ms)

    def spread(ms, steps):
        per = sorted(x / steps for x in ms)
        return {"median": statistics.median(per), "min": per[0], "max": per[-1], "blocks": len(per)}

    # documents in flight: 6 (3, 4, 6 and 8 measured on a 16-core box: value 0.305, 0.271, 0.256, 0.265 ms per step; e2e the same
    # for all), but never more spinning worker threads than this rank's share of the cores less one
    # (left for the main thread, NCCL's proxy thread and the driver's own threads)
    inflight = args.inflight if args.inflight > 0 else max(1, min(6, per_rank - 1 if world > 1 else (os.cpu_count() or 1) - 1))
    sampler = ClockSampler(local_rank)
    sampler.start()
    # ---- value: text resident in HBM, result left in HBM
    ms_value, _, _, launches = run_overlapped(args.steps, inflight, "resident", args.warmup)
    barrier()
    n_serial = min(args.steps, 10)
    ms_serial, _ = timed_resident(n_serial, per_kernel=False)
    ms_serial /= n_serial

    # ---- e2e: pinned host text in, host text out; and the same with pageable buffers on both sides
    ms_e2e, ph, copy_stats, _ = run_overlapped(args.steps, inflight, "pinned", args.warmup)
    phase_ms = {k: ph[i] for i, k in enumerate(("new", "clusters", "nodes_x%d" % len(cks), "bytes"))}
    barrier()
    ms_page, ph_page, _, _ = run_overlapped(args.steps, inflight, "pageable", args.warmup)
    barrier()
    ms_e2e_serial, _, _, _ = run_overlapped(n_serial, 1, "pinned", 1, blocks=1)
    ms_e2e_serial = ms_e2e_serial[0] / n_serial
    clocks = sampler.stop()

    # ---- per-kernel device times for the roofline (the same steps again, events around every launch)
    _, stats = timed_resident(args.steps, per_kernel=True)
    kern = {}
    for name, st in stats.items():
        if st["timed"]:
            kern[name] = {"launches_per_step": st["timed"] / args.steps, "avg_us": st["ms"] / st["timed"] * 1e3,
                          "ms_per_step": st["ms"] / args.steps}
    # algorithmic bytes per launch: New = input length, Bytes = output length
    algo = {"scan_count_kernel": len(text), "scan_emit_kernel": len(text), "emit_kernel": len(text), "row_offsets_kernel": len(text),
            "lines_count + lines_prefix kernels": len(text), "lines_rows_kernel": len(text)}
    for k, v in kern.items():
        if k in algo:
            v["algorithmic_bytes"] = algo[k]
            v["gbs"] = algo[k] / (v["avg_us"] * 1e-6) / 1e9
    dominant = max((k for k in kern if k in algo), key=lambda k: kern[k]["ms_per_step"])
    achieved = kern[dominant]["gbs"]
    # DRAM bytes of that kernel from the committed ncu capture (tools/ncu_traffic.py), if there is one
    traffic, traffic_src = None, None
    try:
        import glob
        newest = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_traffic.json")))[-1]
        k = json.load(open(newest))["kernels"].get(dominant)
        if k:
            traffic, traffic_src = k["dram_read_bytes"] + k["dram_write_bytes"], os.path.relpath(newest, ROOT)
    except (IndexError, OSError, ValueError, KeyError):
        pass
    # the whole step against the same peak: New reads the text once, Bytes writes it once, the module
    # index is the key bytes + 4 B per key -- over the sum of all kernel times of a step
    key_bytes = sum(len(synth.cluster_key(c)) for c in range(args.clusters)) + len("cluster-manager") + \
        sum(len(synth.node_key(c, i)) for c in range(args.clusters) for i in range(1, args.nodes + 1))
    n_keys = 1 + args.clusters * (1 + args.nodes)
    step_bytes = 2 * len(text) + key_bytes + 4 * n_keys
    kernel_ms = sum(v["ms_per_step"] for v in kern.values())
    roofline = {"kernel": dominant, "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": algo[dominant], "avg_launch_us": kern[dominant]["avg_us"],
                "kernel_ms_per_step_total": kernel_ms, "step_algorithmic_bytes": step_bytes,
                "step_frac": step_bytes / (kernel_ms * 1e-3) / 1e9 / hbm_peak if kernel_ms else None,
                "step_frac_serial_latency": step_bytes / (ms_serial * 1e-3) / 1e9 / hbm_peak,
                "value_serial_l2_flushed_ms": ms_serial, "kernels": kern}

    # ---- max over ranks, block by block; the median block is the reported one
    ms_value, ms_e2e, ms_page = max_over_ranks(ms_value), max_over_ranks(ms_e2e), max_over_ranks(ms_page)
    if rank == 0:
        sv, se, sp = spread(ms_value, args.steps), spread(ms_e2e, args.steps), spread(ms_page, args.steps)
        ops = lambda ms_step: world * OPS_PER_STEP / (ms_step * 1e-3)
        line = {
            "metric": "state.State ops/sec", "value": ops(sv["median"]), "unit": "ops/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sv["median"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "documents_in_flight": inflight,
                       "blocks": {"note": "the block of --steps steps is timed %d times back to back after the warm-up; value / e2e are the "
                                          "median block, ms per step" % BLOCKS,
                                  "value": sv, "e2e": se, "e2e_pageable": sp},
                       "l2": ("every step reads one of %d device copies of the text in turn and streams about 400 MB of its own "
                              "buffers: far more than the 126 MB L2 between two uses of anything" % R_COPIES) if inflight > 1 else
                             "inputs rotate over %d device copies (249 MB > L2); the serial latency run flushes L2 (256 MiB write) "
                             "before every step" % R_COPIES,
                       "latency_ms_per_step": {"value_serial_l2_flushed": ms_serial, "e2e_serial": ms_e2e_serial},
                       "value_covers": "the whole step with the text already in HBM and the result left in HBM "
                                       "(the module index is copied to the host once per step: Clusters/Nodes look it up there)",
                       "e2e_variants": {"pinned": {"value": ops(se["median"]), "ms_per_step": se["median"],
                                                   "what": "caller's text in pinned host memory; result read in place from the handle's pinned buffer"},
                                        "pageable": {"value": ops(sp["median"]), "ms_per_step": sp["median"],
                                                     "what": "caller's text in malloc'd memory (ioutil.ReadFile's slice, backend/local/backend.go:60-65); "
                                                             "result memcpy'd out of the handle's buffer (C.GoBytes, :94)",
                                                     "phase_ms": {k: ph_page[i] for i, k in enumerate(("new", "clusters", "nodes", "bytes"))}}},
                       "driver": "tools/bench_driver.c (the step is a compiled loop over the C ABI)",
                       "host_threads": {"workers_per_rank": inflight, "cores_per_rank": per_rank if world > 1 else len(allowed),
                                        "pinned_to": my_cpus if world > 1 else "not pinned (one rank)"},
                       "pcie_per_rank_GBps": {"e2e_h2d": copy_stats[0] / (se["median"] * 1e-3) / 1e9, "e2e_d2h": copy_stats[1] / (se["median"] * 1e-3) / 1e9,
                                              "note": "bytes copied per step / e2e step time, per GPU; with both directions busy one B200 on "
                                                      "this pool's hosts sustains about 49 GB/s each way alone"},
                       "clusters": args.clusters, "nodes_per_cluster": args.nodes, "document_bytes": len(text)},
            "e2e": {"value": ops(se["median"]), "unit": "ops/s",
                    "ms_per_step": se["median"], "h2d_bytes_per_step": copy_stats[0], "d2h_bytes_per_step": copy_stats[1],
                    "phase_ms": phase_ms, "pageable": {"value": ops(sp["median"]), "ms_per_step": sp["median"]}},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
        }
        if world == 1:
            arm = OracleArm(text, cks, args.clusters, args.nodes)
            steps, dt = arm.run(1, seconds=args.cpu_seconds)
            line["cpu_baseline"] = {"value": steps * OPS_PER_STEP / dt, "unit": "ops/s", "cores": 1, "kind": "port",
                                    "ms_per_step": dt / steps * 1e3,
                                    "sample": "%d full passes of the %d-byte document in %.1f s (oracle/tk_oracle.c, gcc -O2, one thread)"
                                              % (steps, len(text), dt)}
        print(json.dumps(line), flush=True)
    libc.free.argtypes = [c_void_p]
    libc.free(heap_text)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1:
        # not launched by torchrun: relaunch under it (one process per GPU)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
