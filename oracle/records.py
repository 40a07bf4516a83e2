"""Oracle restatement of how encoding/json marshals the reference's renderer structs
(TEST INFRASTRUCTURE ONLY).

The field tables are read from tests/golden/struct_tables.json, which
tests/golden/gen_struct_tables.py extracted from the reference's Go sources (create/*.go,
backend/local/backend.go).  Rules restated: fields in declaration order with an embedded struct's
fields in its place; `json:"-"` skipped; `omitempty` drops "" and nil/empty slices but never a
struct value; a nil slice without omitempty is `null`.
"""
import json
import os

from .oracle import Record

_T = None


def tables():
    global _T
    if _T is None:
        here = os.path.dirname(os.path.abspath(__file__))
        _T = json.load(open(os.path.join(here, "..", "tests", "golden", "struct_tables.json")))
    return _T


def render(go_struct: str, settings: dict) -> Record:
    fields = tables()[go_struct]["fields"]
    unknown = set(settings) - {f["json"] for f in fields}
    if unknown:
        raise KeyError("%s has no field %s" % (go_struct, sorted(unknown)))
    members = []
    for f in fields:
        v = settings.get(f["json"])
        if f["type"] == "string":
            v = "" if v is None else v
            if v == "" and f["omitempty"]:
                continue
        elif f["type"] == "[]string":
            if not v and f["omitempty"]:
                continue
        else:
            v = render(f["type"], v or {})
        members.append((f["json"], v))
    return Record(members)


PROVIDER_DIRS = {"triton": "triton", "aws": "aws", "gcp": "gcp", "azure": "azure", "baremetal": "bare-metal", "vsphere": "vsphere"}
KIND_SUFFIX = {"manager": "-rancher", "cluster": "-rancher-k8s", "node": "-rancher-k8s-host"}


def module_source(kind, provider, source_url=None, source_ref=None):
    """create/manager.go:160-171, create/cluster.go:299-311, create/node.go:209-221."""
    return "%s//terraform/modules/%s%s?ref=%s" % (source_url or "github.com/joyent/triton-kubernetes",
                                                    PROVIDER_DIRS[provider], KIND_SUFFIX[kind], source_ref or "master")
