"""util/backend_prompt.go:18-168 in non-interactive mode: which backend persists the documents."""
from __future__ import annotations

from .env import FlowError


def prompt_for_backend(config, local_backend_factory):
    """local_backend_factory() -> the local backend (product or oracle).  The Manta backend is a
    remote network service and out of scope: its validation ladder is kept, the
    backend itself is refused."""
    if not config.IsSet("backend_provider"):
        raise FlowError("backend_provider must be provided")
    provider = config.GetString("backend_provider")
    if provider == "local":
        return local_backend_factory()
    if provider == "manta":
        if not config.IsSet("triton_account"):
            raise FlowError("triton_account must be specified")
        if not config.IsSet("triton_key_path"):
            raise FlowError("triton_key_path must be specified")
        key_path = config.GetString("triton_key_path")
        if not config.IsSet("triton_key_id"):
            try:
                open(key_path).close()                                  # util/ssh_utils.go:13-42 reads the key to fingerprint it
            except OSError:
                raise FlowError("Unable to read private key: open %s: no such file or directory" % key_path)
        if not config.IsSet("triton_url"):
            raise FlowError("triton_url must be specified")
        if not config.IsSet("manta_url"):
            raise FlowError("manta_url must be specified")
        raise FlowError("the Manta backend is a remote service and is not part of this build")
    raise FlowError("Unsupported backend provider '%s'" % provider)
