"""get/manager.go:16-93 and get/cluster.go:17-141 (non-interactive): terraform output of one module."""
from __future__ import annotations

from . import shell
from .env import FlowError


def _manager(env):
    managers = env.backend.States()
    if not managers:
        raise FlowError("No cluster managers.")
    if not env.config.IsSet("cluster_manager"):
        raise FlowError("cluster_manager must be specified")
    selected = env.config.GetString("cluster_manager")
    if selected not in managers:
        raise FlowError("Selected cluster manager '%s' does not exist." % selected)
    return selected


def get_manager(env):
    state = env.backend.State(_manager(env))
    shell.run_terraform_output_with_state(env, state, "cluster-manager")        # get/manager.go:63-93


def get_cluster(env):
    state = env.backend.State(_manager(env))
    clusters = state.Clusters()                                                 # get/cluster.go:65
    if not clusters:
        raise FlowError("No clusters.")
    if not env.config.IsSet("cluster_name"):
        raise FlowError("cluster_name must be specified")
    name = env.config.GetString("cluster_name")
    key = clusters.get(name.encode())
    if key is None:
        raise FlowError("A cluster named '%s', does not exist." % name)
    shell.run_terraform_output_with_state(env, state, key.decode())             # :109-138
