"""DRAM traffic per kernel launch from an `ncu --set full` capture:

    python tools/ncu_traffic.py gpurun_out/prof.ncu-rep > profiles/<round>_ncu_traffic.json

bench.py reads the newest profiles/*_ncu_traffic.json for the `roofline.traffic` field
(dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, averaged over the captured
launches).  ncu replays every kernel with caches flushed, so these are cold-cache figures."""
import csv, io, json, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = rows[0]
units = rows[1]


def col(name):
    return h.index(name)


def to_bytes(v, unit):
    x = float(v.replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


per = {}
for r in rows[2:]:
    name = r[col("Kernel Name")].split("(")[0].split("::")[-1]
    rd = to_bytes(r[col("dram__bytes_read.sum")], units[col("dram__bytes_read.sum")])
    wr = to_bytes(r[col("dram__bytes_write.sum")], units[col("dram__bytes_write.sum")])
    us = float(r[col("gpu__time_duration.sum")].replace(",", "")) / {"ns": 1e3, "us": 1, "ms": 1e-3}.get(units[col("gpu__time_duration.sum")], 1e3)
    per.setdefault(name, []).append((rd, wr, us))
out = {}
for name, ls in per.items():
    # launches that had nothing to do (the line pass on text it declines exits at once; the seed document of
    # tools/one_step.py is such a text) would halve the averages: only launches within 5x of the longest count
    longest = max(us for _, _, us in ls)
    kept = [l for l in ls if l[2] * 5 >= longest]
    n = len(kept)
    out[name] = {"launches": n, "skipped_idle_launches": len(ls) - n, "dram_read_bytes": round(sum(l[0] for l in kept) / n),
                 "dram_write_bytes": round(sum(l[1] for l in kept) / n), "us": round(sum(l[2] for l in kept) / n, 2)}
print(json.dumps({"source": rep.split("/")[-1], "note": "per launch, ncu --set full (caches flushed before every replay)", "kernels": out}, indent=1))
