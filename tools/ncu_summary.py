"""Per-kernel one-line summary of an .ncu-rep (needs ncu on PATH; no GPU): python tools/ncu_summary.py file.ncu-rep"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
cols = [("gpu__time_duration.sum", "us", 1e-3 if units[idx["gpu__time_duration.sum"]] == "ns" else 1.0),
        ("dram__bytes_read.sum", "rdMB", None), ("dram__bytes_write.sum", "wrMB", None),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 1.0),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%", 1.0),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "l1%", 1.0),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%", 1.0),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 1.0),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 1.0),
        ("launch__registers_per_thread", "regs", 1.0),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst", 1.0)]


def to_mb(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(unit, 1.0)


print("%-24s %10s" % ("kernel", "grid") + "".join("%9s" % c[1] for c in cols))
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0].split("::")[-1]
    out = "%-24s %10s" % (name[:24], r[idx["Grid Size"]].split(",")[0].strip("("))
    for key, label, scale in cols:
        if key not in idx:
            out += "%9s" % "-"
            continue
        v = r[idx[key]]
        if scale is None:
            out += "%9.2f" % to_mb(v, units[idx[key]])
        else:
            out += "%9.1f" % (float(v.replace(",", "")) * scale)
    print(out)
