"""Host-side mirror of the reference's `state` package over the C ABI (include/tk_state.h).

The reference is Go; this image has no Go toolchain, so the test/bench harness binds the
library with ctypes.  Method names, argument meaning and error text follow state/state.go so
the parity tests read like state/state_test.go.  INTEGRATION.md shows the equivalent cgo stub.

There is no CPU fallback in here: every method is one call into libtkstate.so, and loading
fails loudly if the library has not been built (python __graft_entry__.py build).
"""
from __future__ import annotations

import ctypes
import json
import os
from ctypes import POINTER, byref, c_char_p, c_int, c_longlong, c_size_t, c_ulonglong, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TK_STATE_LIB") or os.path.join(_HERE, "libtkstate.so")   # the override is for A/B builds in development

TK_OK, TK_ESYNTAX, TK_ECOLLISION, TK_ENOTOBJ, TK_EKEYPARTS, TK_ECUDA, TK_ELIMIT, TK_EARG, TK_EIO = range(9)
TK_VALUE_MAP, TK_VALUE_RECORD = 0, 1

# every symbol include/tk_state.h declares (tests check the library exports each one)
SYMBOLS = [
    "tk_state_new", "tk_state_new_empty", "tk_state_free", "tk_state_name", "tk_last_error", "tk_state_get",
    "tk_state_set", "tk_state_set_manager", "tk_state_set_terraform_backend_config", "tk_state_add_cluster",
    "tk_state_add_node", "tk_state_delete", "tk_state_bytes", "tk_state_clusters", "tk_state_nodes",
    "tk_new_hostnames", "tk_state_stage", "tk_state_bytes_device", "tk_state_prefix_count_device",
    "tk_state_stream", "tk_kernel_launches", "tk_version", "tk_state_timing", "tk_state_kernel_stats",
    "tk_state_copy_stats", "tk_row_offsets_noalloc", "tk_state_host_tree_built", "tk_state_new_device",
    "tk_set_default_timing", "tk_set_line_pass", "tk_render_record", "tk_record_struct_name", "tk_module_source", "tk_record_table",
    "tk_local_backend_new", "tk_backend_free", "tk_backend_last_error", "tk_backend_state", "tk_backend_delete_state",
    "tk_backend_persist_state", "tk_backend_states", "tk_backend_state_terraform_config", "tk_set_device",
]

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libtkstate.so is not built (%s missing): run `python __graft_entry__.py build`; "
            "there is no Python or CPU fallback" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    P = c_void_p
    L.tk_state_new.argtypes = [c_char_p, c_size_t, c_char_p, c_size_t, POINTER(P)]
    L.tk_state_new_device.argtypes = [c_char_p, c_size_t, P, c_size_t, POINTER(P)]
    L.tk_set_default_timing.argtypes = [c_int]
    L.tk_set_line_pass.argtypes = [c_int]
    L.tk_render_record.argtypes = [c_char_p, c_size_t, c_char_p, c_size_t, POINTER(P), POINTER(c_size_t)]
    L.tk_record_struct_name.argtypes = [c_char_p, c_char_p, POINTER(P), POINTER(c_size_t)]
    L.tk_module_source.argtypes = [c_char_p, c_char_p, c_char_p, c_char_p, POINTER(P), POINTER(c_size_t)]
    L.tk_record_table.argtypes = [c_size_t, POINTER(c_char_p), POINTER(P), POINTER(c_size_t)]
    L.tk_local_backend_new.argtypes = [c_char_p, c_size_t, POINTER(P)]
    L.tk_backend_free.argtypes = [P]
    L.tk_backend_free.restype = None
    L.tk_backend_last_error.argtypes = [P]
    L.tk_backend_last_error.restype = c_char_p
    L.tk_backend_state.argtypes = [P, c_char_p, c_size_t, POINTER(P)]
    L.tk_backend_delete_state.argtypes = [P, c_char_p, c_size_t]
    L.tk_backend_persist_state.argtypes = [P, P]
    L.tk_backend_states.argtypes = [P, POINTER(P), POINTER(c_size_t), POINTER(c_size_t)]
    L.tk_backend_state_terraform_config.argtypes = [P, c_char_p, c_size_t, POINTER(P), POINTER(c_size_t), POINTER(P), POINTER(c_size_t)]
    L.tk_state_new_empty.argtypes = [c_char_p, c_size_t, POINTER(P)]
    L.tk_state_free.argtypes = [P]
    L.tk_state_free.restype = None
    L.tk_state_name.argtypes = [P, POINTER(P), POINTER(c_size_t)]
    L.tk_last_error.argtypes = [P]
    L.tk_last_error.restype = c_char_p
    L.tk_state_get.argtypes = [P, c_char_p, c_size_t, POINTER(P), POINTER(c_size_t)]
    L.tk_state_set.argtypes = [P, c_char_p, c_size_t, c_char_p, c_size_t, c_int]
    L.tk_state_set_manager.argtypes = [P, c_char_p, c_size_t, c_int]
    L.tk_state_set_terraform_backend_config.argtypes = [P, c_char_p, c_size_t, c_char_p, c_size_t, c_int]
    L.tk_state_add_cluster.argtypes = [P, c_char_p, c_size_t, c_char_p, c_size_t, c_char_p, c_size_t, c_int]
    L.tk_state_add_node.argtypes = [P, c_char_p, c_size_t, c_char_p, c_size_t, c_char_p, c_size_t, c_int]
    L.tk_state_delete.argtypes = [P, c_char_p, c_size_t]
    L.tk_state_bytes.argtypes = [P, POINTER(P), POINTER(c_size_t)]
    L.tk_state_clusters.argtypes = [P, POINTER(P), POINTER(c_size_t), POINTER(c_size_t)]
    L.tk_state_nodes.argtypes = [P, c_char_p, c_size_t, POINTER(P), POINTER(c_size_t), POINTER(c_size_t)]
    L.tk_new_hostnames.argtypes = [POINTER(c_char_p), POINTER(c_size_t), c_size_t, c_char_p, c_size_t, c_longlong,
                                   POINTER(P), POINTER(c_size_t)]
    L.tk_state_stage.argtypes = [P]
    L.tk_state_bytes_device.argtypes = [P, POINTER(P), POINTER(c_size_t)]
    L.tk_state_prefix_count_device.argtypes = [P, c_char_p, c_size_t, POINTER(c_size_t)]
    L.tk_state_stream.argtypes = [P]
    L.tk_state_stream.restype = P
    L.tk_state_timing.argtypes = [P, c_int]
    L.tk_state_kernel_stats.argtypes = [P, c_int, POINTER(c_char_p), POINTER(ctypes.c_double), POINTER(c_ulonglong),
                                        POINTER(c_ulonglong)]
    L.tk_state_copy_stats.argtypes = [P, POINTER(c_ulonglong), POINTER(c_ulonglong)]
    L.tk_row_offsets_noalloc.argtypes = [P, P, P, P, P, P, ctypes.c_uint32, ctypes.c_uint32, P, P, c_char_p, c_size_t]
    L.tk_state_host_tree_built.argtypes = [P]
    L.tk_kernel_launches.restype = c_ulonglong
    L.tk_version.restype = c_char_p
    _lib = L
    return L


class StateError(Exception):
    """err.Error() of the failing call, verbatim; .code is the TK_E* value."""

    def __init__(self, code, text):
        super().__init__(text)
        self.code = code


class Record:
    """A Go struct value: ordered (name, value) members with omitempty already applied."""

    def __init__(self, members):
        self.members = list(members.items()) if isinstance(members, dict) else list(members)


def _b(s) -> bytes:
    return bytes(s) if isinstance(s, (bytes, bytearray)) else str(s).encode("utf-8", "surrogateescape")


def _enc(o) -> str:
    if isinstance(o, Record):
        return "{" + ",".join(json.dumps(k, ensure_ascii=False) + ":" + _enc(v) for k, v in o.members) + "}"
    if isinstance(o, dict):
        return "{" + ",".join(json.dumps(k, ensure_ascii=False) + ":" + _enc(v) for k, v in o.items()) + "}"
    if isinstance(o, (list, tuple)):
        return "[" + ",".join(_enc(v) for v in o) + "]"
    return json.dumps(o, ensure_ascii=False)


def _value(obj):
    """(json text, TK_VALUE_*) -- the transport form of an `obj interface{}` argument."""
    kind = TK_VALUE_RECORD if isinstance(obj, Record) else TK_VALUE_MAP
    if isinstance(obj, (bytes, bytearray)):
        return bytes(obj), kind
    return _enc(obj).encode("utf-8", "surrogateescape"), kind


def _pairs(ptr, n) -> dict:
    blob = ctypes.string_at(ptr.value, n.value) if n.value else b""
    out, i = {}, 0
    while i < len(blob):
        nl = int.from_bytes(blob[i:i + 4], "little"); i += 4
        name = blob[i:i + nl]; i += nl
        kl = int.from_bytes(blob[i:i + 4], "little"); i += 4
        out[name] = blob[i:i + kl]; i += kl
    return out


class State:
    """state.State (state/state.go:10-13) backed by a tk_state handle."""

    def __init__(self, name, raw=None, device_ptr=None, handle=None):
        L = lib()
        nm = _b(name)
        h = c_void_p()
        if handle is not None:  # adopt a handle another call produced (the local backend)
            self._h = handle
            return
        if device_ptr is not None:  # (device address, length): text already in HBM
            rc = L.tk_state_new_device(nm, len(nm), device_ptr[0], device_ptr[1], byref(h))
        elif raw is None:
            rc = L.tk_state_new_empty(nm, len(nm), byref(h))
        elif isinstance(raw, tuple):  # (address, length) of a caller-owned (e.g. pinned) buffer
            rc = L.tk_state_new(nm, len(nm), ctypes.cast(raw[0], c_char_p), raw[1], byref(h))
        else:
            raw = _b(raw)
            rc = L.tk_state_new(nm, len(nm), raw, len(raw), byref(h))
        if rc != TK_OK:
            raise StateError(rc, L.tk_last_error(None).decode("utf-8", "replace"))
        self._h = h

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h and _lib is not None:
            _lib.tk_state_free(h)

    __del__ = close

    def _check(self, rc):
        if rc != TK_OK:
            raise StateError(rc, lib().tk_last_error(self._h).decode("utf-8", "replace"))

    @property
    def Name(self) -> bytes:
        p, n = c_void_p(), c_size_t()
        self._check(lib().tk_state_name(self._h, byref(p), byref(n)))
        return ctypes.string_at(p.value, n.value) if n.value else b""

    # -- reads -----------------------------------------------------------------------------
    def Get(self, path) -> bytes:
        p = _b(path)
        v, n = c_void_p(), c_size_t()
        self._check(lib().tk_state_get(self._h, p, len(p), byref(v), byref(n)))
        return ctypes.string_at(v.value, n.value) if n.value else b""

    def Bytes(self) -> bytes:
        v, n = c_void_p(), c_size_t()
        self._check(lib().tk_state_bytes(self._h, byref(v), byref(n)))
        return ctypes.string_at(v.value, n.value) if n.value else b""

    def BytesView(self):
        """Bytes() without the copy into a Python object: (address, length) of the handle's buffer."""
        v, n = c_void_p(), c_size_t()
        self._check(lib().tk_state_bytes(self._h, byref(v), byref(n)))
        return v.value, n.value

    def Clusters(self) -> dict:
        b, n, c = c_void_p(), c_size_t(), c_size_t()
        self._check(lib().tk_state_clusters(self._h, byref(b), byref(n), byref(c)))
        return _pairs(b, n)

    def ClustersCount(self) -> int:
        """len(Clusters()) without building the Python dict."""
        b, n, c = c_void_p(), c_size_t(), c_size_t()
        self._check(lib().tk_state_clusters(self._h, byref(b), byref(n), byref(c)))
        return c.value

    def NodesCount(self, cluster_key) -> int:
        """len(Nodes(ck)) without building the Python dict; out-parameters are reused across calls
        so that a sweep over many clusters measures the library, not ctypes object churn."""
        out = self.__dict__.get("_nc")
        if out is None:
            b, n, c = c_void_p(), c_size_t(), c_size_t()
            out = self._nc = (lib().tk_state_nodes, byref(b), byref(n), byref(c), c)
        ck = cluster_key if isinstance(cluster_key, bytes) else _b(cluster_key)
        rc = out[0](self._h, ck, len(ck), out[1], out[2], out[3])
        if rc != TK_OK:
            self._check(rc)
        return out[4].value

    def Nodes(self, cluster_key) -> dict:
        ck = _b(cluster_key)
        b, n, c = c_void_p(), c_size_t(), c_size_t()
        self._check(lib().tk_state_nodes(self._h, ck, len(ck), byref(b), byref(n), byref(c)))
        return _pairs(b, n)

    # -- writes ----------------------------------------------------------------------------
    def _set(self, path, obj):
        p = _b(path)
        j, kind = _value(obj)
        self._check(lib().tk_state_set(self._h, p, len(p), j, len(j), kind))

    def SetManager(self, obj):
        j, kind = _value(obj)
        self._check(lib().tk_state_set_manager(self._h, j, len(j), kind))

    def SetTerraformBackendConfig(self, path, obj):
        p = _b(path)
        j, kind = _value(obj)
        self._check(lib().tk_state_set_terraform_backend_config(self._h, p, len(p), j, len(j), kind))

    def AddCluster(self, provider, name, obj):
        pr, nm = _b(provider), _b(name)
        j, kind = _value(obj)
        self._check(lib().tk_state_add_cluster(self._h, pr, len(pr), nm, len(nm), j, len(j), kind))

    def AddNode(self, cluster_key, name, obj):
        ck, nm = _b(cluster_key), _b(name)
        j, kind = _value(obj)
        self._check(lib().tk_state_add_node(self._h, ck, len(ck), nm, len(nm), j, len(j), kind))

    def Delete(self, path):
        p = _b(path)
        self._check(lib().tk_state_delete(self._h, p, len(p)))

    # -- device-resident variants (bench.py) ----------------------------------------------------
    def Stage(self):
        self._check(lib().tk_state_stage(self._h))

    def BytesDevice(self):
        v, n = c_void_p(), c_size_t()
        self._check(lib().tk_state_bytes_device(self._h, byref(v), byref(n)))
        return v.value, n.value

    def PrefixCountDevice(self, prefix, want_count=True):
        p = _b(prefix)
        c = c_size_t()
        self._check(lib().tk_state_prefix_count_device(self._h, p, len(p), byref(c) if want_count else None))
        return c.value

    @property
    def host_tree_built(self) -> bool:
        return bool(lib().tk_state_host_tree_built(self._h))

    def Timing(self, on: bool):
        self._check(lib().tk_state_timing(self._h, 1 if on else 0))

    def KernelStats(self) -> dict:
        """{kernel name: {"ms": summed device time while timing was on, "timed": n, "launches": n}}"""
        out, kind = {}, 0
        while True:
            name, ms, tl, ln = c_char_p(), ctypes.c_double(), c_ulonglong(), c_ulonglong()
            if lib().tk_state_kernel_stats(self._h, kind, byref(name), byref(ms), byref(tl), byref(ln)) != TK_OK:
                break
            out[name.value.decode()] = {"ms": ms.value, "timed": tl.value, "launches": ln.value}
            kind += 1
        return out

    def CopyStats(self):
        a, b = c_ulonglong(), c_ulonglong()
        self._check(lib().tk_state_copy_stats(self._h, byref(a), byref(b)))
        return a.value, b.value

    @property
    def stream(self) -> int:
        return lib().tk_state_stream(self._h) or 0


def New(name, raw) -> State:
    """state.New(name, raw) -- state/state.go:15-25."""
    return State(name, raw)


def getNewHostnames(existing, node_name, nodes_to_add):
    """create/node.go:350-380 through tk_new_hostnames."""
    ex = [_b(e) for e in existing]
    arr = (c_char_p * max(len(ex), 1))(*ex)
    lens = (c_size_t * max(len(ex), 1))(*[len(e) for e in ex])
    nn = _b(node_name)
    out, n = c_void_p(), c_size_t()
    rc = lib().tk_new_hostnames(arr, lens, len(ex), nn, len(nn), int(nodes_to_add), byref(out), byref(n))
    if rc != TK_OK:
        raise StateError(rc, lib().tk_last_error(None).decode("utf-8", "replace"))
    blob = ctypes.string_at(out.value, n.value) if n.value else b""
    return [x.decode() for x in blob.split(b"\n")] if blob else []


def NewDevice(name, device_address, length) -> State:
    """state.New on text already in device memory."""
    return State(name, device_ptr=(device_address, length))


def set_default_timing(on: bool):
    lib().tk_set_default_timing(1 if on else 0)


def set_line_pass(on: bool):
    """tk_set_line_pass: state.New with (default) or without the line-by-line pass for indented text."""
    lib().tk_set_line_pass(1 if on else 0)


def set_device(device: int):
    """tk_set_device: the CUDA device for handles the calling thread creates from now on."""
    if lib().tk_set_device(int(device)) != TK_OK:
        raise StateError(TK_ECUDA, lib().tk_last_error(None).decode("utf-8", "replace"))


def kernel_launches() -> int:
    return int(lib().tk_kernel_launches())
