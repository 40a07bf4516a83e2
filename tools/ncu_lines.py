"""Per-source-line cost of one kernel in an .ncu-rep: python tools/ncu_lines.py rep kernel [N]

Needs a capture made with --import-source on from a build with -lineinfo.  Prints the N source
lines with the most executed warp-instructions, next to their share of the stall samples."""
import csv, io, subprocess, sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kern],
                     capture_output=True, text=True).stdout
lines = {}
hdr = None
for r in csv.reader(io.StringIO(raw)):
    if len(r) > 8 and r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r)}
        # "Source" appears twice (CUDA text, then SASS): the first index is the CUDA text
        src_i = r.index("Source")
        continue
    if hdr is None or len(r) < 9 or not r[0].isdigit():
        continue
    try:
        inst = int(r[hdr["Instructions Executed"]]); smp = int(r[hdr["# Samples"]])
        thr = int(r[hdr["Thread Instructions Executed"]])
    except ValueError:
        continue
    k = int(r[0])
    a = lines.setdefault(k, [0, 0, 0, r[src_i].strip()])
    a[0] += inst; a[1] += smp; a[2] += thr
ti = sum(v[0] for v in lines.values()) or 1
ts = sum(v[1] for v in lines.values()) or 1
print("%s: %d warp-instructions, %d samples over %d source lines" % (kern, ti, ts, len(lines)))
for k, v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5d  %5.1f%% inst %5.1f%% smp  thr/inst %4.1f  %s" % (k, 100.0 * v[0] / ti, 100.0 * v[1] / ts, v[2] / max(v[0], 1), v[3][:100]))
