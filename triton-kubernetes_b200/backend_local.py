"""backend.Backend (backend/backend.go:7-27) implemented by the local backend
(backend/local/backend.go:14-132) over the C ABI."""
from __future__ import annotations

import ctypes
from ctypes import byref, c_size_t, c_void_p

from .state import Record, State, StateError, TK_OK, _b, lib
import json


class LocalBackend:
    def __init__(self, root_dir=None):
        """local.New(); root_dir None means ~/.triton-kubernetes as in the reference."""
        h = c_void_p()
        r = None if root_dir is None else _b(root_dir)
        rc = lib().tk_local_backend_new(r, 0 if r is None else len(r), byref(h))
        if rc != TK_OK:
            raise StateError(rc, lib().tk_backend_last_error(None).decode("utf-8", "replace"))
        self._b = h

    def close(self):
        b, self._b = getattr(self, "_b", None), None
        if b:
            lib().tk_backend_free(b)

    __del__ = close

    def _check(self, rc):
        if rc != TK_OK:
            raise StateError(rc, lib().tk_backend_last_error(self._b).decode("utf-8", "replace"))

    def State(self, name) -> State:
        nm, h = _b(name), c_void_p()
        self._check(lib().tk_backend_state(self._b, nm, len(nm), byref(h)))
        return State(name, handle=h)

    def DeleteState(self, name):
        nm = _b(name)
        self._check(lib().tk_backend_delete_state(self._b, nm, len(nm)))

    def PersistState(self, state: State):
        self._check(lib().tk_backend_persist_state(self._b, state._h))

    def States(self) -> list:
        p, n, c = c_void_p(), c_size_t(), c_size_t()
        self._check(lib().tk_backend_states(self._b, byref(p), byref(n), byref(c)))
        blob = ctypes.string_at(p.value, n.value) if n.value else b""
        return [x.decode("utf-8", "surrogateescape") for x in blob.split(b"\n")] if c.value else []

    def StateTerraformConfig(self, name):
        nm = _b(name)
        p, pl, r, rl = c_void_p(), c_size_t(), c_void_p(), c_size_t()
        self._check(lib().tk_backend_state_terraform_config(self._b, nm, len(nm), byref(p), byref(pl), byref(r), byref(rl)))
        rec = json.loads(ctypes.string_at(r.value, rl.value), object_pairs_hook=lambda kv: Record(kv))
        return ctypes.string_at(p.value, pl.value).decode(), rec
