"""destroy/manager.go:16-84, destroy/cluster.go:16-167, destroy/node.go:16-186 (non-interactive)."""
from __future__ import annotations

from . import shell
from .env import FlowError


def _manager(env, none_message):
    managers = env.backend.States()
    if not managers:
        raise FlowError(none_message)
    if not env.config.IsSet("cluster_manager"):
        raise FlowError("cluster_manager must be specified")
    selected = env.config.GetString("cluster_manager")
    if selected not in managers:
        raise FlowError("Selected cluster manager '%s' does not exist." % selected)
    return selected


def _cluster(env, state):
    clusters = state.Clusters()
    if not env.config.IsSet("cluster_name"):
        raise FlowError("cluster_name must be specified")
    name = env.config.GetString("cluster_name")
    key = clusters.get(name.encode())
    if key is None:
        raise FlowError("A cluster named '%s', does not exist." % name)
    return key.decode()


def delete_manager(env):
    manager = _manager(env, "No cluster managers, please create a cluster manager before creating a kubernetes cluster.")
    state = env.backend.State(manager)
    shell.run_terraform_destroy_with_state(env, state, [])                      # destroy/manager.go:70
    env.backend.DeleteState(manager)                                            # :76


def delete_cluster(env):
    manager = _manager(env, "No cluster managers.")
    state = env.backend.State(manager)
    cluster_key = _cluster(env, state)
    nodes = state.Nodes(cluster_key)                                            # destroy/cluster.go:126
    # Go ranges over the map in random order; any order is the reference's behaviour, so use the sorted one
    node_keys = [k.decode() for _, k in sorted(nodes.items())]
    args = ["-target=module.%s" % cluster_key] + ["-target=module.%s" % k for k in node_keys]
    shell.run_terraform_destroy_with_state(env, state, args)                    # :141
    state.Delete("module.%s" % cluster_key)                                     # :147
    for k in node_keys:
        state.Delete("module.%s" % k)                                           # :153-158
    env.backend.PersistState(state)


def delete_node(env):
    manager = _manager(env, "No cluster managers, please create a cluster manager before creating a kubernetes node.")
    state = env.backend.State(manager)
    cluster_key = _cluster(env, state)
    nodes = state.Nodes(cluster_key)                                            # destroy/node.go:110
    if not env.config.IsSet("hostname"):
        raise FlowError("hostname must be specified")
    hostname = env.config.GetString("hostname")
    node_key = nodes.get(hostname.encode())
    if node_key is None:
        raise FlowError("A node named '%s', does not exist." % hostname)
    # the node is still in the document while terraform runs; it is removed only after it succeeds
    shell.run_terraform_destroy_with_state(env, state, ["-target=module.%s" % node_key.decode()])   # :167-168
    state.Delete("module.%s" % node_key.decode())                               # :174
    env.backend.PersistState(state)                                             # :180
