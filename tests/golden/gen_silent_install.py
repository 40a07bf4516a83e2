"""Copies the reference's silent-install example configurations into tests/golden/silent-install/.

Run in the build container (needs /root/reference): python tests/golden/gen_silent_install.py
The two YAML files (examples/silent-install/create-manager.yaml, cluster-triton-ha.yaml) are input
fixtures for the settings loader (flows/config.py, cmd/root.go:48-67); docs_keys.json lists the
parameter names the reference documents for them (docs/guide/silent-install-yaml.md:11-49), which
tests/test_flows.py compares with what the loader sees.  The tests never read /root/reference.
"""
import json
import os
import re
import shutil

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "silent-install")
os.makedirs(OUT, exist_ok=True)
for name in ("create-manager.yaml", "cluster-triton-ha.yaml"):
    shutil.copyfile(os.path.join(REF, "examples", "silent-install", name), os.path.join(OUT, name))

doc = open(os.path.join(REF, "docs", "guide", "silent-install-yaml.md")).read()
sections = {}
current = None
for line in doc.split("\n"):
    m = re.match(r"^## (.+?)\s*$", line)
    if m:
        current = m.group(1)
        sections[current] = []
        continue
    if current and line.startswith("| `"):
        sections[current] += re.findall(r"`([a-z0-9_]+)`", line.split("|")[1])
json.dump({k: sorted(set(v)) for k, v in sections.items() if v}, open(os.path.join(OUT, "docs_keys.json"), "w"), indent=1, sort_keys=True)
print({k: len(v) for k, v in sections.items() if v})
