"""Prints the numbers of a bench.py JSON line that matter when comparing two builds."""
import json
import sys

for path in sys.argv[1:]:
    d = json.loads(open(path).read().strip().splitlines()[-1])
    b = d["config"].get("blocks", {})
    print("== %s" % path)
    print("value %.0f ops/s  %.4f ms/step  (blocks %s)  launches/block %s" % (d["value"], d["ms_per_step"],
          {k: "%.3f-%.3f" % (v["min"], v["max"]) for k, v in b.items() if isinstance(v, dict)}, d.get("gpu_launches")))
    print("e2e   %.0f ops/s  %.4f ms/step   pageable %.4f ms/step   serial: value %.3f e2e %.3f" % (
        d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"].get("pageable", {}).get("ms_per_step", float("nan")),
        d["config"]["latency_ms_per_step"]["value_serial_l2_flushed"], d["config"]["latency_ms_per_step"]["e2e_serial"]))
    r = d["roofline"]
    print("roofline %s frac %.3f (%.1f us)  kernel sum %.4f ms  step_frac %.4f  clocks %s" % (
        r["kernel"], r["frac"], r["avg_launch_us"], r["kernel_ms_per_step_total"], r.get("step_frac") or 0, d.get("clocks")))
    for k, v in sorted(r["kernels"].items(), key=lambda kv: -kv[1]["ms_per_step"]):
        print("  %-42s %6.1f us x%.1f  %.4f ms/step" % (k, v["avg_us"], v["launches_per_step"], v["ms_per_step"]))
    if "cpu_baseline" in d:
        print("cpu_baseline %.1f ops/s" % d["cpu_baseline"]["value"])
