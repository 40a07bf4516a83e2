"""Launch list (ncu --metrics gpu__time_duration.sum --csv --log-file X) -> per-kernel totals and shares:

    python tools/launch_summary.py gpurun_out/launches.csv "the command that was profiled" > profiles/<round>_launches_summary.txt

ncu serialises the launches and flushes caches before each, so compare the SHARES with bench.py's
live event timings, not the absolute times."""
import csv
import sys

path = sys.argv[1]
cmd = sys.argv[2] if len(sys.argv) > 2 else ""
rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10 and r[0].isdigit()]
agg = {}
for r in rows:
    name = r[4].split("(")[0].split("::")[-1]
    if "<" in r[4].split("(")[0]:
        name = r[4].split("(")[0].split("::")[-1]
    key = (name, r[8], r[7])
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += float(r[-1]) / 1e3
total = sum(v[1] for v in agg.values()) or 1.0
print("# ncu --metrics gpu__time_duration.sum --clock-control none  %s" % cmd)
print("# cold-cache, serialised launches: compare SHARES with bench.py's live event timings, not absolutes")
print("# %d launches, %.1f us in all" % (len(rows), total))
for (name, grid, block), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-44s grid x block %-16s x %-14s n=%5d total=%10.1f us avg=%8.2f us share=%5.1f%%" % (name[:44], grid, block, n, us, us / n, 100 * us / total))
