"""Hottest SASS instructions of one kernel in an .ncu-rep: python tools/ncu_hot.py rep kernel [N]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[2:] if len(r) == len(hdr) and r[ix["Instructions Executed"]].isdigit()]
tot_i = sum(int(r[ix["Instructions Executed"]]) for r in body)
tot_s = sum(int(r[ix["# Samples"]]) for r in body)
print("%s: %d SASS lines, %d warp-instructions, %d samples" % (kern, len(body), tot_i, tot_s))
# cumulative by opcode
ops = {}
for r in body:
    op = r[ix["Source"]].split()[0] if not r[ix["Source"]].strip().startswith("@") else r[ix["Source"]].split()[1]
    op = op.split(".")[0]
    ops[op] = ops.get(op, 0) + int(r[ix["Instructions Executed"]])
print("by opcode:", ", ".join("%s %.1f%%" % (k, 100.0 * v / tot_i) for k, v in sorted(ops.items(), key=lambda x: -x[1])[:14]))
for r in sorted(body, key=lambda r: -int(r[ix["# Samples"]]))[:top]:
    print("%6.2f%% smp %6.2f%% inst  thr %5s  %s" % (100.0 * int(r[ix["# Samples"]]) / max(tot_s, 1), 100.0 * int(r[ix["Instructions Executed"]]) / tot_i,
                                                  r[ix["Avg. Threads Executed"]], r[ix["Source"]].strip()[:90]))
