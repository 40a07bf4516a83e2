"""create/manager.go:29-154, create/cluster.go:45-290, create/node.go:43-347, the bare-metal renderers
(create/{manager,cluster,node}_bare_metal.go) and the vSphere ones (create/{cluster,node}_vsphere.go),
non-interactive mode only."""
from __future__ import annotations

import os
import re

from . import shell
from .env import FlowError

CLOUD_API = {"triton", "aws", "gcp", "azure"}          # their renderers call a cloud API unconditionally
CLUSTER_NAME = re.compile(r"^[a-z0-9]([-a-z0-9]*[a-z0-9])?(\.[a-z0-9]([-a-z0-9]*[a-z0-9])?)*$")   # create/cluster.go:314


def _need(cfg, key, message=None):
    """the ladder: viper.IsSet -> value; else (non-interactive) the reference's error text"""
    if cfg.IsSet(key):
        return cfg.GetString(key)
    raise FlowError(message or "%s must be specified" % key)


def _source(env, kind, provider):
    c = env.config
    return env.module_source(kind, provider, c.GetString("source_url") if c.IsSet("source_url") else None,
                             c.GetString("source_ref") if c.IsSet("source_ref") else None)


def _registry(cfg, settings, prefix_key, user_key, pass_key, out_keys):
    """create/manager.go:176-218 / create/cluster.go:391-516: username and password only if a registry is given"""
    if cfg.IsSet(prefix_key):
        settings[out_keys[0]] = cfg.GetString(prefix_key)
    if settings.get(out_keys[0]):
        settings[out_keys[1]] = _need(cfg, user_key)
        settings[out_keys[2]] = _need(cfg, pass_key)


# ---------------------------------------------------------------------------------- manager

def new_manager(env):
    cfg = env.config
    provider = _need(cfg, "manager_cloud_provider")
    name = _need(cfg, "name")
    if name == "":
        raise FlowError("Invalid Cluster Manager Name")
    if name in env.backend.States():
        raise FlowError("A Cluster Manager with the name '%s' already exists." % name)
    state = env.backend.State(name)
    if provider == "baremetal":
        _new_bare_metal_manager(env, state, name)
    elif provider in CLOUD_API:
        raise FlowError("manager_cloud_provider '%s' needs cloud API calls, which this build leaves out" % provider)
    else:
        raise FlowError("Unsupported cloud provider '%s', cannot create manager" % provider)
    path, obj = env.backend.StateTerraformConfig(name)
    state.SetTerraformBackendConfig(path, obj)                                  # create/manager.go:140
    shell.run_terraform_apply_with_state(env, state)
    env.backend.PersistState(state)


def _base_manager(env, provider, name):
    cfg = env.config
    s = {"source": _source(env, "manager", provider), "name": name}
    _registry(cfg, s, "private_registry", "private_registry_username", "private_registry_password",
              ("rancher_registry", "rancher_registry_username", "rancher_registry_password"))
    for key in ("rancher_server_image", "rancher_agent_image"):
        if cfg.IsSet(key):
            s[key] = cfg.GetString(key)
    s["rancher_admin_password"] = _need(cfg, "rancher_admin_password", "UI Admin Password must be specified")
    return s


def _new_bare_metal_manager(env, state, name):                                  # create/manager_bare_metal.go:29-150
    cfg = env.config
    s = _base_manager(env, "baremetal", name)
    s["host"] = _need(cfg, "host")
    s["ssh_user"] = _need(cfg, "ssh_user")
    s["bastion_host"] = _need(cfg, "bastion_host")
    s["key_path"] = _need(cfg, "key_path")
    state.SetManager(env.render("bareMetalManagerTerraformConfig", s))


# ---------------------------------------------------------------------------------- cluster

def _select_manager(env, none_message):
    cfg = env.config
    managers = env.backend.States()
    if not managers:
        raise FlowError(none_message)
    selected = _need(cfg, "cluster_manager")
    if selected not in managers:
        raise FlowError("Selected cluster manager '%s' does not exist." % selected)
    return selected


def new_cluster(env):
    cfg = env.config
    manager = _select_manager(env, "No cluster managers, please create a cluster manager before creating a kubernetes cluster.")
    state = env.backend.State(manager)
    provider = _need(cfg, "cluster_cloud_provider")
    if provider == "baremetal":
        cluster_name = _new_bare_metal_cluster(env, state)
    elif provider == "vsphere":
        cluster_name = _new_vsphere_cluster(env, state)
    elif provider in CLOUD_API:
        raise FlowError("cluster_cloud_provider '%s' needs cloud API calls, which this build leaves out" % provider)
    else:
        raise FlowError("Unsupported cloud provider '%s', cannot create cluster" % provider)
    # struct values are invisible to Get/Clusters until the document is re-parsed (create/cluster.go:146-152)
    state = env.state.New(manager, state.Bytes())
    clusters = state.Clusters()
    cluster_key = clusters.get(cluster_name.encode())
    if cluster_key is None:
        raise FlowError("Couldn't find cluster key for cluster '%s'.\n" % cluster_name)
    if cfg.IsSet("nodes"):
        nodes = cfg.Get("nodes")
        if not isinstance(nodes, list):
            raise FlowError("Could not read 'nodes' configuration")
        for node in nodes:
            if not isinstance(node, dict):
                raise FlowError("Could not read node configuration")
            for key in ("rancher_host_label", "node_count", "hostname"):          # create/cluster.go:177-179
                cfg.Set(key, node.get(key))
            if provider == "baremetal":                                             # :203-208
                for key in ("ssh_user", "key_path", "bastion_host", "hosts"):
                    cfg.Set(key, node.get(key))
            hostnames = new_node_in(env, manager, cluster_key.decode(), state)
            _print_nodes_added(env, hostnames)
    shell.run_terraform_apply_with_state(env, state)
    env.backend.PersistState(state)


def _base_cluster(env, provider):                                                   # create/cluster.go:292-516
    cfg = env.config
    s = {"source": _source(env, "cluster", provider),
         "rancher_api_url": "${module.cluster-manager.rancher_url}",
         "rancher_access_key": "${module.cluster-manager.rancher_access_key}",
         "rancher_secret_key": "${module.cluster-manager.rancher_secret_key}"}
    s["name"] = _need(cfg, "name")
    if s["name"] == "" or not CLUSTER_NAME.match(s["name"]):
        raise FlowError("Invalid Cluster Name")
    s["k8s_version"] = _need(cfg, "k8s_version")
    s["k8s_network_provider"] = _need(cfg, "k8s_network_provider")
    _registry(cfg, s, "private_registry", "private_registry_username", "private_registry_password",
              ("rancher_registry", "rancher_registry_username", "rancher_registry_password"))
    _registry(cfg, s, "k8s_registry", "k8s_registry_username", "k8s_registry_password",
              ("k8s_registry", "k8s_registry_username", "k8s_registry_password"))
    return s


def _new_bare_metal_cluster(env, state):                                            # create/cluster_bare_metal.go:19-36
    s = _base_cluster(env, "baremetal")
    state.AddCluster("baremetal", s["name"], env.render("bareMetalClusterTerraformConfig", s))
    return s["name"]


VSPHERE_CLUSTER_KEYS = ("vsphere_user", "vsphere_password", "vsphere_server", "vsphere_datacenter_name", "vsphere_datastore_name",
                        "vsphere_resource_pool_name", "vsphere_network_name")


def _new_vsphere_cluster(env, state):                                               # create/cluster_vsphere.go:29-195
    cfg = env.config
    s = _base_cluster(env, "vsphere")
    for key in VSPHERE_CLUSTER_KEYS:                                                # the ladder, in the reference's order
        s[key] = _need(cfg, key, "%s must be specified." % key)
    state.AddCluster("vsphere", s["name"], env.render("vSphereClusterTerraformConfig", s))
    return s["name"]


# ---------------------------------------------------------------------------------- node

def new_node(env):                                                                  # create/node.go:43-169
    cfg = env.config
    manager = _select_manager(env, "No cluster managers, please create a cluster manager before creating a kubernetes cluster.")
    state = env.backend.State(manager)
    clusters = state.Clusters()
    if not clusters:
        raise FlowError("No clusters, please create a cluster before creating a kubernetes node.")
    cluster_name = _need(cfg, "cluster_name")
    cluster_key = clusters.get(cluster_name.encode())
    if cluster_key is None:
        raise FlowError("A cluster named '%s', does not exist." % cluster_name)
    hostnames = new_node_in(env, manager, cluster_key.decode(), state)
    shell.run_terraform_apply_with_state(env, state)
    env.backend.PersistState(state)
    _print_nodes_added(env, hostnames)
    return hostnames


def new_node_in(env, manager, cluster_key, state):                                  # create/node.go:171-195
    parts = cluster_key.split("_")
    if len(parts) < 3:
        raise FlowError("Could not determine cloud provider for cluster '%s'" % cluster_key)
    if parts[1] == "baremetal":
        return _new_bare_metal_node(env, cluster_key, state)
    if parts[1] == "vsphere":
        return _new_vsphere_node(env, cluster_key, state)
    if parts[1] in CLOUD_API:
        raise FlowError("provider '%s' needs cloud API calls, which this build leaves out" % parts[1])
    raise FlowError("Unsupported cloud provider '%s', cannot create node" % parts[1])


def _base_node(env, provider, cluster_key, state):                                  # create/node.go:197-347
    cfg = env.config
    get = lambda p: state.Get(p).decode()
    s = {"source": _source(env, "node", provider),
         "rancher_api_url": "${module.cluster-manager.rancher_url}",
         "rancher_cluster_registration_token": "${module.%s.rancher_cluster_registration_token}" % cluster_key,
         "rancher_cluster_ca_checksum": "${module.%s.rancher_cluster_ca_checksum}" % cluster_key,
         "rancher_agent_image": get("module.cluster-manager.rancher_agent_image"),
         "rancher_registry": get("module.%s.rancher_registry" % cluster_key),
         "rancher_registry_username": get("module.%s.rancher_registry_username" % cluster_key),
         "rancher_registry_password": get("module.%s.rancher_registry_password" % cluster_key)}
    label = _need(cfg, "rancher_host_label")
    if label not in ("worker", "etcd", "control"):
        raise FlowError("Invalid rancher_host_label '%s', must be 'worker', 'etcd' or 'control'" % label)
    s["rancher_host_labels"] = {label: "true"}
    count_input = _need(cfg, "node_count")
    try:
        if not re.fullmatch(r"[+-]?[0-9]+", count_input):
            raise ValueError
        count = int(count_input)
    except ValueError:
        raise FlowError("node_count must be a valid number. Found '%s'." % count_input)
    if count <= 0:
        raise FlowError("node_count must be greater than 0. Found '%d'." % count)
    s["hostname"] = _need(cfg, "hostname")
    if s["hostname"] == "":
        raise FlowError("Invalid Hostname")
    return s, count


def _new_bare_metal_node(env, cluster_key, state):                                  # create/node_bare_metal.go:34-181
    cfg = env.config
    s, count = _base_node(env, "baremetal", cluster_key, state)
    s["ssh_user"] = _need(cfg, "ssh_user")
    s["bastion_host"] = _need(cfg, "bastion_host")
    s["key_path"] = _need(cfg, "key_path")
    existing = [h.decode() for h in state.Nodes(cluster_key)]
    hostnames = env.state.getNewHostnames(existing, s["hostname"], count)
    if not cfg.IsSet("hosts"):
        raise FlowError("hosts must be specified")
    hosts = cfg.GetStringSlice("hosts")
    if len(hosts) != count:
        raise FlowError("not enough hosts")
    for host, hostname in zip(hosts, hostnames):
        state.AddNode(cluster_key, hostname, env.render("bareMetalNodeTerraformConfig", dict(s, hostname=hostname, host=host)))
    return hostnames


def _expand_home(path):
    """homedir.Expand: a leading ~ (alone or before a slash) becomes $HOME; anything else is left alone"""
    if path == "~" or path.startswith("~/"):
        return os.path.expanduser("~") + path[1:]
    if path.startswith("~"):
        raise FlowError("cannot expand user-specific home dir")
    return path


def _new_vsphere_node(env, cluster_key, state):                                     # create/node_vsphere.go:37-158
    cfg = env.config
    s, count = _base_node(env, "vsphere", cluster_key, state)
    get = lambda p: state.Get(p).decode()
    for key in ("vsphere_user", "vsphere_password", "vsphere_server"):              # copied from the cluster's block
        s[key] = get("module.%s.%s" % (cluster_key, key))
    for key in ("vsphere_datacenter_name", "vsphere_datastore_name", "vsphere_resource_pool_name", "vsphere_network_name"):
        s[key] = "${module.%s.%s}" % (cluster_key, key)                             # terraform outputs of the cluster module
    s["vsphere_template_name"] = _need(cfg, "vsphere_template_name")
    s["ssh_user"] = _need(cfg, "ssh_user")
    s["key_path"] = _expand_home(_need(cfg, "key_path"))
    existing = [h.decode() for h in state.Nodes(cluster_key)]
    hostnames = env.state.getNewHostnames(existing, s["hostname"], count)
    for hostname in hostnames:
        state.AddNode(cluster_key, hostname, env.render("vSphereNodeTerraformConfig", dict(s, hostname=hostname)))
    return hostnames


def _print_nodes_added(env, hostnames):                                             # create/cluster.go:518-525
    if not hostnames:
        return
    if len(hostnames) == 1:
        env.out.append("1 node added: %s" % hostnames[0])
    else:
        env.out.append("%d nodes added: %s" % (len(hostnames), ", ".join(hostnames)))
