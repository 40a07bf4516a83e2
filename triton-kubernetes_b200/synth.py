"""Deterministic synthetic state documents: BASELINE.json config #4.

1,000 clusters x 100 nodes of bare-metal records in their post-reload (all-map) form:
101,001 module keys, 3,115,015 key bytes, 62,177,405 bytes tab-indented, 55,435,335 compact.
No RNG: the document is a pure function of (clusters, nodes_per_cluster).
"""
from __future__ import annotations

import json

SRC = "github.com/joyent/triton-kubernetes//terraform/modules/"
REF = "?ref=master"


def manager_record() -> dict:
    return {
        "source": SRC + "bare-metal-rancher" + REF, "name": "bench", "rancher_admin_password": "admin",
        "host": "10.0.0.1", "ssh_user": "ubuntu", "key_path": "~/.ssh/id_rsa",
    }


def cluster_record(c: int) -> dict:
    return {
        "source": SRC + "bare-metal-rancher-k8s" + REF, "name": "c%04d" % c,
        "rancher_api_url": "${module.cluster-manager.rancher_url}",
        "rancher_access_key": "${module.cluster-manager.rancher_access_key}",
        "rancher_secret_key": "${module.cluster-manager.rancher_secret_key}",
        "k8s_version": "v1.9.5-rancher1-1", "k8s_network_provider": "calico",
    }


def node_record(c: int, i: int) -> dict:
    ck = "cluster_baremetal_c%04d" % c
    return {
        "source": SRC + "bare-metal-rancher-k8s-host" + REF, "hostname": "c%04d-w-%d" % (c, i),
        "rancher_api_url": "${module.cluster-manager.rancher_url}",
        "rancher_cluster_registration_token": "${module.%s.rancher_cluster_registration_token}" % ck,
        "rancher_cluster_ca_checksum": "${module.%s.rancher_cluster_ca_checksum}" % ck,
        "rancher_host_labels": {"worker": "true"},
        "host": "10.%d.%d.%d" % (c // 256, c % 256, i), "bastion_host": "", "ssh_user": "ubuntu",
        "key_path": "~/.ssh/id_rsa",
    }


def cluster_key(c: int) -> str:
    return "cluster_baremetal_c%04d" % c


def node_key(c: int, i: int) -> str:
    return "node_baremetal_c%04d_c%04d-w-%d" % (c, c, i)


def document(clusters: int = 1000, nodes_per_cluster: int = 100, with_nodes: bool = True) -> bytes:
    """Compact JSON text of the document, keys in generation order (NOT sorted)."""
    d = json.dumps
    parts = ['{"module":{"cluster-manager":', d(manager_record(), separators=(",", ":"))]
    for c in range(clusters):
        parts.append(',"%s":%s' % (cluster_key(c), d(cluster_record(c), separators=(",", ":"))))
        if with_nodes:
            for i in range(1, nodes_per_cluster + 1):
                parts.append(',"%s":%s' % (node_key(c, i), d(node_record(c, i), separators=(",", ":"))))
    parts.append('},"terraform":{"backend":{"local":{"path":"/root/.triton-kubernetes/bench/terraform.tfstate"}}}}')
    return "".join(parts).encode()
