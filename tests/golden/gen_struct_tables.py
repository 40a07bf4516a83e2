"""Generates tests/golden/struct_tables.json from the reference's Go sources.

Run in the build container (needs /root/reference): python tests/golden/gen_struct_tables.py
It reads every `type X struct` under create/ and backend/local/ whose fields carry `json:` tags
(the records the renderers hand to state.SetManager / AddCluster / AddNode /
SetTerraformBackendConfig -- SURVEY.md 8a row 14) and records, per struct, the field order as
encoding/json will emit it: declaration order, an embedded struct's fields promoted in place,
`json:"-"` dropped, `omitempty` noted, Go field type kept.  The JSON file is the pinned golden
vector for the product's record tables (tests/test_records.py); the tests never read
/root/reference themselves.
"""
import glob
import json
import os
import re

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "struct_tables.json")

structs = {}
for path in sorted(glob.glob(REF + "/create/*.go") + glob.glob(REF + "/backend/local/*.go")):
    if path.endswith("_test.go"):
        continue
    lines = open(path).read().split("\n")
    i = 0
    while i < len(lines):
        m = re.match(r"^type (\w+) struct \{", lines[i])
        if not m:
            i += 1
            continue
        name, start = m.group(1), i + 1
        fields = []
        i += 1
        while not lines[i].startswith("}"):
            ln = lines[i].strip()
            i += 1
            if not ln or ln.startswith("//"):
                continue
            fm = re.match(r'^(\w+)\s+([\w\[\]\.\*]+)\s+`json:"([^"]*)"`', ln)
            if fm:
                tag = fm.group(3).split(",")
                fields.append({"go": fm.group(1), "type": fm.group(2), "json": tag[0], "omitempty": "omitempty" in tag[1:],
                               "line": i})
            elif re.match(r"^\w+$", ln):
                fields.append({"embed": ln, "line": i})
        structs[name] = {"file": os.path.relpath(path, REF), "line": start, "fields": fields}

def flatten(name):
    out = []
    for f in structs[name]["fields"]:
        if "embed" in f:
            out.extend(flatten(f["embed"]))
        elif f["json"] != "-":
            out.append({"json": f["json"], "omitempty": f["omitempty"], "type": f["type"]})
    return out

tables = {}
for name, s in structs.items():
    if not any("json" in f or "embed" in f for f in s["fields"]):
        continue
    tables[name] = {"source": "%s:%d" % (s["file"], s["line"]), "fields": flatten(name)}
json.dump(tables, open(OUT, "w"), indent=1, sort_keys=True)
print("%d structs -> %s" % (len(tables), OUT))
for k in sorted(tables):
    print(k, [f["json"] + ("?" if f["omitempty"] else "") for f in tables[k]["fields"]][:6], "...")
