"""Renderer records over the C ABI: the Go structs the reference hands to
SetManager / AddCluster / AddNode, rendered in encoding/json's struct order."""
from __future__ import annotations

import ctypes
import json
from ctypes import byref, c_char_p, c_size_t, c_void_p

from .state import Record, StateError, TK_OK, _b, _enc, lib


def _take(p, n):
    return ctypes.string_at(p.value, n.value) if n.value else b""


def struct_name(kind: str, provider: str) -> str:
    p, n = c_void_p(), c_size_t()
    rc = lib().tk_record_struct_name(_b(kind), _b(provider), byref(p), byref(n))
    if rc != TK_OK:
        raise StateError(rc, lib().tk_last_error(None).decode())
    return _take(p, n).decode()


def module_source(kind: str, provider: str, source_url=None, source_ref=None) -> str:
    p, n = c_void_p(), c_size_t()
    rc = lib().tk_module_source(_b(kind), _b(provider), None if source_url is None else _b(source_url),
                                None if source_ref is None else _b(source_ref), byref(p), byref(n))
    if rc != TK_OK:
        raise StateError(rc, lib().tk_last_error(None).decode())
    return _take(p, n).decode()


def render_json(go_struct: str, settings: dict) -> bytes:
    """The struct's JSON text exactly as the library renders it."""
    s, j = _b(go_struct), _enc(settings).encode("utf-8", "surrogateescape")
    p, n = c_void_p(), c_size_t()
    rc = lib().tk_render_record(s, len(s), j, len(j), byref(p), byref(n))
    if rc != TK_OK:
        raise StateError(rc, lib().tk_last_error(None).decode())
    return _take(p, n)


def render(go_struct: str, settings: dict) -> Record:
    """A Record (ordered members) ready for State.SetManager / AddCluster / AddNode."""
    return Record(json.loads(render_json(go_struct, settings), object_pairs_hook=lambda kv: Record(kv)).members)


def tables() -> dict:
    """{Go struct name: [{"json", "omitempty", "type"}, ...]} as compiled into the library."""
    out, i = {}, 0
    while True:
        name, p, n = c_char_p(), c_void_p(), c_size_t()
        if lib().tk_record_table(i, byref(name), byref(p), byref(n)) != TK_OK:
            return out
        out[name.value.decode()] = json.loads(_take(p, n))
        i += 1
