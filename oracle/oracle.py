"""ctypes front end of the CPU oracle (oracle/tk_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package.

The class mirrors the reference's state.State method set (state/state.go:15-147) so the
parity tests read like state/state_test.go.  Values handed to SetManager/AddCluster/... are
either a dict (Go map[string]interface{}: sorted on output, visible to Get) or a Record
(a Go struct value: declaration order on output, opaque until re-parsed; SURVEY.md 8a note 1).
"""
from __future__ import annotations

import ctypes
import json
import os
import subprocess
from ctypes import POINTER, byref, c_char_p, c_int, c_longlong, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "tk_oracle.c")
_LIB = os.path.join(_HERE, "libtk_oracle.so")


def build(force: bool = False) -> str:
    """Compile the C restatement with gcc (idempotent)."""
    if force or not os.path.exists(_LIB) or (
        os.path.exists(_SRC) and os.path.getmtime(_LIB) < os.path.getmtime(_SRC)
    ):
        subprocess.check_call(
            ["gcc", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-o", _LIB, _SRC, "-lm"]
        )
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB)
        L.ora_new.restype = c_void_p
        L.ora_new.argtypes = [c_char_p, c_size_t, c_char_p, c_size_t]
        L.ora_free.argtypes = [c_void_p]
        L.ora_free_buf.argtypes = [c_void_p]
        L.ora_get.argtypes = [c_void_p, c_char_p, c_size_t, POINTER(c_void_p), POINTER(c_size_t)]
        L.ora_set.argtypes = [c_void_p, c_char_p, c_size_t, c_char_p, c_size_t, c_int, c_char_p, c_size_t]
        L.ora_add_cluster.argtypes = [c_void_p, c_char_p, c_size_t, c_char_p, c_size_t, c_char_p, c_size_t, c_int, c_char_p, c_size_t]
        L.ora_add_node.argtypes = [c_void_p, c_char_p, c_size_t, c_char_p, c_size_t, c_char_p, c_size_t, c_int, c_char_p, c_size_t]
        L.ora_delete.argtypes = [c_void_p, c_char_p, c_size_t, c_char_p, c_size_t]
        L.ora_bytes.argtypes = [c_void_p, c_int, POINTER(c_void_p), POINTER(c_size_t)]
        L.ora_bytes_indent.argtypes = [c_void_p, c_char_p, c_char_p, POINTER(c_void_p), POINTER(c_size_t)]
        L.ora_path_string.argtypes = [c_void_p, c_char_p, c_size_t, POINTER(c_void_p), POINTER(c_size_t)]
        L.ora_clusters.argtypes = [c_void_p, POINTER(c_void_p), POINTER(c_size_t), POINTER(c_size_t), c_char_p, c_size_t]
        L.ora_nodes.argtypes = [c_void_p, c_char_p, c_size_t, POINTER(c_void_p), POINTER(c_size_t), POINTER(c_size_t), c_char_p, c_size_t]
        L.ora_new_hostnames.argtypes = [POINTER(c_char_p), POINTER(c_size_t), c_size_t, c_char_p, c_size_t, c_longlong, POINTER(c_void_p), POINTER(c_size_t)]
        L.ora_check.argtypes = [c_char_p, c_size_t, c_char_p, c_size_t]
        _lib = L
    return _lib


class OracleError(Exception):
    """Carries the Go error text (err.Error()) verbatim."""


class Record:
    """A Go struct value: ordered (name, value) members, omitempty already applied."""

    def __init__(self, members):
        self.members = list(members.items()) if isinstance(members, dict) else list(members)

    def to_json(self) -> bytes:
        return _to_json(self)


def _b(s) -> bytes:
    return s if isinstance(s, (bytes, bytearray)) else str(s).encode("utf-8", "surrogateescape")


def _enc(o) -> str:
    if isinstance(o, Record):
        return "{" + ",".join(json.dumps(k, ensure_ascii=False) + ":" + _enc(v) for k, v in o.members) + "}"
    if isinstance(o, dict):
        return "{" + ",".join(json.dumps(k, ensure_ascii=False) + ":" + _enc(v) for k, v in o.items()) + "}"
    if isinstance(o, (list, tuple)):
        return "[" + ",".join(_enc(v) for v in o) + "]"
    return json.dumps(o, ensure_ascii=False)


def _to_json(o) -> bytes:
    """Transport encoding only: the C side decodes it again, so escaping choices here do not matter."""
    return _enc(o).encode("utf-8", "surrogateescape")


def _is_record(o) -> int:
    return 1 if isinstance(o, Record) else 0


def _take(ptr: c_void_p, n: c_size_t) -> bytes:
    data = ctypes.string_at(ptr.value, n.value) if n.value else b""
    lib().ora_free_buf(ptr)
    return data


def _pairs(blob: bytes) -> dict:
    out, i = {}, 0
    while i < len(blob):
        nl = int.from_bytes(blob[i:i + 4], "little"); i += 4
        name = blob[i:i + nl]; i += nl
        kl = int.from_bytes(blob[i:i + 4], "little"); i += 4
        key = blob[i:i + kl]; i += kl
        out[name] = key
    return out


class State:
    """state.State (state/state.go:10-13)."""

    def __init__(self, name, raw):
        raw = _b(raw)
        err = ctypes.create_string_buffer(512)
        h = lib().ora_new(raw, len(raw), err, 512)
        if not h:
            raise OracleError(err.value.decode("utf-8", "replace"))
        self.Name = name
        self._h = h

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and _lib is not None:
            _lib.ora_free(h)

    # -- reads ---------------------------------------------------------------
    def Get(self, path) -> bytes:
        p = _b(path)
        out, n = c_void_p(), c_size_t()
        lib().ora_get(self._h, p, len(p), byref(out), byref(n))
        return ctypes.string_at(out.value, n.value) if n.value else b""

    def Bytes(self) -> bytes:
        out, n = c_void_p(), c_size_t()
        lib().ora_bytes(self._h, 1, byref(out), byref(n))
        return _take(out, n)

    def BytesCompact(self) -> bytes:
        """gabs Container.Bytes() (compact json.Marshal) -- used only to replay gabs_test.go."""
        out, n = c_void_p(), c_size_t()
        lib().ora_bytes(self._h, 0, byref(out), byref(n))
        return _take(out, n)

    def BytesIndent(self, prefix, indent) -> bytes:
        """gabs Container.BytesIndent(prefix, indent) (gabs.go:395-402)."""
        out, n = c_void_p(), c_size_t()
        lib().ora_bytes_indent(self._h, _b(prefix), _b(indent), byref(out), byref(n))
        return _take(out, n)

    def PathString(self, path=None) -> bytes:
        """gabs Container.Path(path).String()."""
        out, n = c_void_p(), c_size_t()
        p = None if path is None else _b(path)
        lib().ora_path_string(self._h, p, 0 if p is None else len(p), byref(out), byref(n))
        return _take(out, n)

    def Clusters(self) -> dict:
        out, n, cnt = c_void_p(), c_size_t(), c_size_t()
        err = ctypes.create_string_buffer(512)
        if lib().ora_clusters(self._h, byref(out), byref(n), byref(cnt), err, 512):
            raise OracleError(err.value.decode("utf-8", "replace"))
        return _pairs(_take(out, n))

    def NodesCount(self, cluster_key) -> int:
        """len(Nodes(ck)) without building the Python dict (bench.py times the C code, not Python)."""
        ck = _b(cluster_key)
        out, n, cnt = c_void_p(), c_size_t(), c_size_t()
        err = ctypes.create_string_buffer(1024 + len(ck))
        if lib().ora_nodes(self._h, ck, len(ck), byref(out), byref(n), byref(cnt), err, len(err)):
            raise OracleError(err.value.decode("utf-8", "replace"))
        lib().ora_free_buf(out)
        return cnt.value

    def Nodes(self, cluster_key) -> dict:
        ck = _b(cluster_key)
        out, n, cnt = c_void_p(), c_size_t(), c_size_t()
        err = ctypes.create_string_buffer(1024 + len(ck))
        if lib().ora_nodes(self._h, ck, len(ck), byref(out), byref(n), byref(cnt), err, len(err)):
            raise OracleError(err.value.decode("utf-8", "replace"))
        return _pairs(_take(out, n))

    # -- writes --------------------------------------------------------------
    def _set(self, path, obj):
        p, j = _b(path), (obj if isinstance(obj, (bytes, bytearray)) else _to_json(obj))
        err = ctypes.create_string_buffer(512)
        if lib().ora_set(self._h, p, len(p), j, len(j), _is_record(obj), err, 512):
            raise OracleError(err.value.decode("utf-8", "replace"))

    def SetManager(self, obj):
        self._set("module.cluster-manager", obj)

    def SetTerraformBackendConfig(self, path, obj):
        self._set(path, obj)

    def AddCluster(self, provider, name, obj):
        pr, nm, j = _b(provider), _b(name), _to_json(obj)
        err = ctypes.create_string_buffer(512)
        if lib().ora_add_cluster(self._h, pr, len(pr), nm, len(nm), j, len(j), _is_record(obj), err, 512):
            raise OracleError(err.value.decode("utf-8", "replace"))

    def AddNode(self, cluster_key, name, obj):
        ck, nm, j = _b(cluster_key), _b(name), _to_json(obj)
        err = ctypes.create_string_buffer(1024 + len(ck))
        if lib().ora_add_node(self._h, ck, len(ck), nm, len(nm), j, len(j), _is_record(obj), err, len(err)):
            raise OracleError(err.value.decode("utf-8", "replace"))

    def Delete(self, path):
        p = _b(path)
        err = ctypes.create_string_buffer(512)
        if lib().ora_delete(self._h, p, len(p), err, 512):
            raise OracleError(err.value.decode("utf-8", "replace"))


def New(name, raw) -> State:
    """state.New (state/state.go:15-25)."""
    return State(name, raw)


def getNewHostnames(existing, node_name, nodes_to_add):
    """create/node.go:350-380."""
    ex = [_b(e) for e in existing]
    arr = (c_char_p * max(len(ex), 1))(*ex)
    lens = (c_size_t * max(len(ex), 1))(*[len(e) for e in ex])
    nn = _b(node_name)
    out, n = c_void_p(), c_size_t()
    lib().ora_new_hostnames(arr, lens, len(ex), nn, len(nn), int(nodes_to_add), byref(out), byref(n))
    blob = _take(out, n)
    return [x.decode() for x in blob.split(b"\n")] if blob else []


def check(raw):
    """(ok, go_error_text) for json.Unmarshal(raw, &interface{})."""
    raw = _b(raw)
    err = ctypes.create_string_buffer(512)
    ok = lib().ora_check(raw, len(raw), err, 512)
    return bool(ok), err.value.decode("utf-8", "replace")
