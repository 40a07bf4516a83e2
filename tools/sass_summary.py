"""Per-kernel counts of the SASS instructions that move data, from the built library:

    python tools/sass_summary.py [kernel-name-substring ...] > profiles/<round>_sass_<what>.txt

UBLKCP / UTMA* are the bulk-copy (TMA) unit, SYNCS.* the mbarrier operations that go with it,
LDG/STG global loads and stores through registers, LDS/STS shared-memory accesses, LDGSTS the
Ampere-style cp.async.  Needs cuobjdump (no GPU)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.environ.get("TK_STATE_LIB") or os.path.join(ROOT, "triton-kubernetes_b200", "libtkstate.so")
want = sys.argv[1:]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
kern, counts, lines = None, collections.OrderedDict(), {}
for ln in out.split("\n"):
    m = re.search(r"Function : (\S+)", ln)
    if m:
        kern = m.group(1)
        counts[kern] = collections.Counter()
        lines[kern] = []
        continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", ln)
    if kern and m:
        op = m.group(1)
        if re.match(r"(UBLKCP|UTMA|SYNCS|LDGSTS|LDG|STG|LDS|STS|RED|ATOM)", op):
            counts[kern][op] += 1
            if re.match(r"(UBLKCP|UTMA|SYNCS)", op):
                lines[kern].append(ln.strip().split("/*")[1].split("*/")[1].strip() if "*/" in ln else ln.strip())
for k, c in counts.items():
    dem = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()
    names = re.findall(r"(?:::|^)(\w+)[(<]", dem)
    short = names[0] if names else dem
    if want and not any(w in short for w in want):
        continue
    print("%s" % short)
    print("    " + ", ".join("%s x%d" % (op, n) for op, n in sorted(c.items())))
    for l in lines[k]:
        print("        " + l)
