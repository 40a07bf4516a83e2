"""Non-interactive create / destroy / get flows: the reference's
create/, destroy/, get/, shell/ and cmd/ packages restated for `--non-interactive` use, on top of
the State / Backend / record interfaces.  Prompts, cloud SDK calls and the Manta backend are out
of scope, so providers whose renderers call a cloud API are refused by name.

Every flow takes an `Env`: the state module (New, Record, getNewHostnames), a backend and the
record renderer -- the product's (GPU) or the oracle's (CPU) -- so the same flow code produces the
files the parity tests compare.
"""
from .config import Config  # noqa: F401
from .env import Env, FlowError  # noqa: F401
from . import backend, create, destroy, get, shell  # noqa: F401
