"""triton-kubernetes_b200: the state.State path of triton-kubernetes on B200 (sm_100a).

The package directory name carries a hyphen (it follows the reference's name), so import it
with importlib:  tk = importlib.import_module("triton-kubernetes_b200").
"""
from . import backend_local, records, state  # noqa: F401
from .backend_local import LocalBackend  # noqa: F401
from .state import New, Record, State, StateError, getNewHostnames  # noqa: F401
