from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable

from .config import Config


class FlowError(Exception):
    """err.Error() of the reference flow, verbatim."""


@dataclass
class Env:
    state: Any                      # module with New(name, raw), Record, getNewHostnames
    backend: Any                    # backend.Backend: State, DeleteState, PersistState, States, StateTerraformConfig
    render: Callable                # render(go_struct_name, settings) -> Record
    module_source: Callable         # module_source(kind, provider, url, ref) -> str
    config: Config = field(default_factory=Config)
    terraform: str = "terraform"    # looked up on PATH (tests put a stub there)
    out: list = field(default_factory=list)   # what the reference prints with fmt.Println
