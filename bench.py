#!/usr/bin/env python
"""bench.py -- state.State ops/sec on BASELINE.json config #4 (1,000 clusters x 100 nodes).

One STEP is one pass of the hot path over the 62,177,405-byte synthetic document:
    state.New(text) -> Clusters() -> Nodes(ck) for each of the 1,000 cluster keys -> Bytes()
= 1,003 state.State operations (destroy/cluster.go:64-161 and get/cluster.go:65-118 walk a
document this way).

  value  the step with its input already resident in HBM (kernels only, output left in HBM),
         timed with CUDA events on the stream the library launches on
  e2e    the same step through the public C ABI with HOST buffers: pinned host text in,
         host->device copies, kernels, device->host copy of the result text, all inside the
         timed region
  --impl reference   the CPU oracle (oracle/tk_oracle.c, a restatement of the reference's Go
         code; the Go toolchain is absent) on the host cores, same step, same metric

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

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

OPS_PER_STEP = 1003
WORKLOAD = ("config #4: synthetic state document, 1000 clusters x 100 nodes (62,177,405 B tab-indented); "
            "step = New + Clusters + 1000 x Nodes(ck) + Bytes = 1003 state.State ops")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--clusters", type=int, default=1000, help=argparse.SUPPRESS)
    ap.add_argument("--nodes", type=int, default=100, help=argparse.SUPPRESS)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help=argparse.SUPPRESS)
    return ap.parse_args()


# ------------------------------------------------------------------------------------ clocks

class ClockSampler:
    """nvidia-smi clocks and throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
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

def oracle_step(O, text, cluster_keys):
    s = O.New("bench", text)
    cl = s.Clusters()
    n = 0
    for ck in cluster_keys:
        n += s.NodesCount(ck)
    out = s.Bytes()
    return len(cl), n, out


def cpu_baseline(text, cluster_keys, seconds, threads=1):
    """The oracle timed on `threads` host threads (ctypes releases the GIL), each running whole steps."""
    from oracle import oracle as O
    O.build()
    O.lib()
    done = [0] * threads
    stop_at = [None]

    def work(i):
        while True:
            oracle_step(O, text, cluster_keys)
            done[i] += 1
            if time.perf_counter() >= stop_at[0]:
                return

    t0 = time.perf_counter()
    stop_at[0] = t0 + seconds
    ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    for t in ts: t.start()
    for t in ts: t.join()
    dt = time.perf_counter() - t0
    steps = sum(done)
    return {"value": steps * OPS_PER_STEP / dt, "unit": "ops/s", "cores": threads, "kind": "port",
            "sample": "%d full passes of the 62,177,405-byte document in %.1f s (oracle/tk_oracle.c, gcc -O2)" % (steps, dt),
            "ms_per_step": dt / steps * 1e3 * threads}


def run_reference(args, rank, world):
    if rank != 0:
        return
    synth = importlib.import_module("triton-kubernetes_b200.synth")
    from oracle import oracle as O
    O.build()
    text = O.New("bench", synth.document(args.clusters, args.nodes)).Bytes()
    cks = [synth.cluster_key(c).encode() for c in range(args.clusters)]
    threads = os.cpu_count() or 1
    for _ in range(min(args.warmup, 1)):
        oracle_step(O, text, cks)
    per_thread = max(1, args.steps)

    def work():
        for _ in range(per_thread):
            oracle_step(O, text, cks)

    t0 = time.perf_counter()
    ts = [threading.Thread(target=work) for _ in range(threads)]
    for t in ts: t.start()
    for t in ts: t.join()
    dt = time.perf_counter() - t0
    steps = per_thread * threads
    value = steps * OPS_PER_STEP / dt
    line = {
        "impl": "reference", "metric": "state.State ops/sec", "value": value, "unit": "ops/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / per_thread * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU oracle (C restatement of the reference's Go path; no Go toolchain here), "
                   "%d host threads each running %d whole steps on its own copy of the document" % (threads, per_thread)},
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
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")

    # ---- synthetic input: the persisted (tab-indented) form of the document, in pinned memory
    seed = tk.New("bench", synth.document(args.clusters, args.nodes))
    text = seed.Bytes()
    seed.close()
    cks = [synth.cluster_key(c).encode() for c in range(args.clusters)]
    node_prefixes = [("node_baremetal_c%04d_" % c).encode() for c in range(args.clusters)]
    pinned = torch.empty(len(text), dtype=torch.uint8, pin_memory=True)
    ctypes.memmove(pinned.data_ptr(), text, len(text))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- value: text resident in HBM; New + Clusters + 1000 x Nodes + Bytes, result left in HBM
    dev_text = torch.empty(len(text), dtype=torch.uint8, device=dev)
    dev_text.copy_(pinned)
    torch.cuda.synchronize()

    def resident_step():
        s = tk.state.NewDevice("bench", dev_text.data_ptr(), len(text))
        nc = s.ClustersCount()
        n = 0
        for ck in cks:
            n += s.NodesCount(ck)
        _, ln = s.BytesDevice()
        return s, (nc == args.clusters and n == args.clusters * args.nodes and ln == len(text))

    def timed_resident(steps, per_kernel):
        tk.state.set_default_timing(per_kernel)
        total, stats = 0.0, {}
        for _ in range(steps):
            flush.zero_()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            s, ok = resident_step()
            b.record(torch.cuda.ExternalStream(s.stream, device=dev))
            b.synchronize()
            total += a.elapsed_time(b)
            if not ok:
                raise SystemExit("bench.py: resident step returned a wrong result")
            if per_kernel:
                for name, st in s.KernelStats().items():
                    acc = stats.setdefault(name, {"ms": 0.0, "timed": 0})
                    acc["ms"] += st["ms"]; acc["timed"] += st["timed"]
            s.close()
        tk.state.set_default_timing(False)
        return total, stats

    for _ in range(args.warmup):
        s, ok = resident_step()
        s.close()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = tk.state.kernel_launches()
    ms_value, _ = timed_resident(args.steps, per_kernel=False)
    launches = tk.state.kernel_launches() - launches0
    barrier()

    # ---- e2e: host text in, host text out, through the public API
    phase = {"new": 0.0, "clusters": 0.0, "nodes_x1000": 0.0, "bytes": 0.0, "close": 0.0}

    def e2e_step():
        t0 = time.perf_counter()
        s = tk.New("bench", (pinned.data_ptr(), len(text)))
        t1 = time.perf_counter()
        nc = s.ClustersCount()
        t2 = time.perf_counter()
        n = 0
        for ck in cks:
            n += s.NodesCount(ck)
        t3 = time.perf_counter()
        addr, ln = s.BytesView()
        t4 = time.perf_counter()
        stats = s.CopyStats()
        phase["new"] += t1 - t0; phase["clusters"] += t2 - t1; phase["nodes_x1000"] += t3 - t2; phase["bytes"] += t4 - t3
        ok = (nc == args.clusters and n == args.clusters * args.nodes and ln == len(text)
              and ctypes.string_at(addr + ln - 64, 64) == text[-64:] and ctypes.string_at(addr, 64) == text[:64])
        t5 = time.perf_counter()
        s.close()
        phase["close"] += time.perf_counter() - t5
        return ok, stats

    for _ in range(args.warmup):
        ok, copy_stats = e2e_step()
        if not ok:
            raise SystemExit("bench.py: e2e step returned a wrong result")
    barrier()
    for k in phase:
        phase[k] = 0.0
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ok, copy_stats = e2e_step()
    torch.cuda.synchronize()
    b.record()
    b.synchronize()
    ms_e2e = max(a.elapsed_time(b), (time.perf_counter() - t0) * 1e3)
    barrier()
    clocks = sampler.stop()

    # ---- per-kernel device times for the roofline (same steps again, events around every launch)
    _, stats = timed_resident(args.steps, per_kernel=True)
    kern = {}
    for name, st in stats.items():
        if st["timed"]:
            kern[name] = {"launches_per_step": st["timed"] / args.steps, "avg_us": st["ms"] / st["timed"] * 1e3,
                          "ms_per_step": st["ms"] / args.steps}
    # algorithmic bytes per launch: New = input length, Bytes = output length
    algo = {"scan_tokens_kernel": len(text), "emit_kernel": len(text), "row_offsets_kernel": len(text)}
    dominant = max((k for k in kern if k in algo), key=lambda k: kern[k]["ms_per_step"])
    achieved = algo[dominant] / (kern[dominant]["avg_us"] * 1e-6) / 1e9
    for k, v in kern.items():
        if k in algo:
            v["algorithmic_bytes"] = algo[k]
            v["gbs"] = algo[k] / (v["avg_us"] * 1e-6) / 1e9
    roofline = {"kernel": dominant, "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": algo[dominant], "avg_launch_us": kern[dominant]["avg_us"], "kernels": kern}

    # ---- max over ranks
    if world > 1:
        t = torch.tensor([ms_value, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_value, ms_e2e = t.tolist()
    if rank == 0:
        line = {
            "metric": "state.State ops/sec", "value": world * OPS_PER_STEP * args.steps / (ms_value * 1e-3), "unit": "ops/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_value / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "l2": "flushed (256 MiB write) between timed steps",
                       "value_covers": "the whole step with the text already in HBM and the result left in HBM "
                                       "(the module index is copied to the host once per step: Clusters/Nodes look it up there)",
                       "clusters": args.clusters, "nodes_per_cluster": args.nodes, "document_bytes": len(text)},
            "e2e": {"value": world * OPS_PER_STEP * args.steps / (ms_e2e * 1e-3), "unit": "ops/s",
                    "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": copy_stats[0], "d2h_bytes_per_step": copy_stats[1],
                    "phase_ms": {k: v / args.steps * 1e3 for k, v in phase.items()}},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
        }
        if world == 1:
            line["cpu_baseline"] = cpu_baseline(text, cks, args.cpu_seconds, threads=1)
        print(json.dumps(line), flush=True)
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
