"""Non-interactive flows.

The error ladders restate the reference's own flow tests (destroy/*_test.go, get/*_test.go,
util/backend_prompt_test.go: exact err.Error() strings, every test stopping before terraform
would run).  The end-to-end cases are BASELINE.json configs #2, #3 and #5: create manager /
cluster + 3 node groups / destroy, bare metal, with a terraform stub on PATH that records its
arguments and a hash of the main.tf.json it was handed; they run once on the oracle (CPU) and,
marked gpu, on the product, and the files are compared byte for byte."""
import hashlib
import importlib
import json
import os
import stat

import pytest

flows = importlib.import_module("triton-kubernetes_b200.flows")
Config, Env, FlowError = flows.Config, flows.Env, flows.FlowError

MOCK_CLUSTERS = b'{"module":{"cluster_1":{"name":"dev_cluster"},"cluster_2":{"name":"beta_cluster"},"cluster_3":{"name":"prod_cluster"}}}'
MOCK_NODE_HOST = b'{"module":{"cluster_triton_dev-cluster":{"name":"dev_cluster"}}}'


class MockBackend:
    """backend/mocks/Backend.go: States() and State(name) return what the test set up."""

    def __init__(self, states, docs=None):
        self.states, self.docs = states, docs or {}

    def States(self):
        return list(self.states)

    def State(self, name):
        return self.docs[name]


def oracle_env(oracle, backend, **settings):
    from oracle import records as orec
    cfg = Config(environ={})
    for k, v in settings.items():
        cfg.Set(k.replace("_DASH_", "-"), v)
    cfg.Set("non-interactive", True)
    return Env(state=oracle, backend=backend, render=orec.render, module_source=orec.module_source, config=cfg)


def err(fn, env):
    with pytest.raises(Exception) as e:      # FlowError, or the state's own error passed through
        fn(env)
    return str(e.value)


# ---- destroy/node_test.go -------------------------------------------------------------------

def test_delete_node_ladder(oracle):
    d = flows.destroy.delete_node
    assert err(d, oracle_env(oracle, MockBackend([]))) == \
        "No cluster managers, please create a cluster manager before creating a kubernetes node."          # :17-29
    two = ["dev-manager", "beta-manager"]
    assert err(d, oracle_env(oracle, MockBackend(two))) == "cluster_manager must be specified"            # :31-44
    assert err(d, oracle_env(oracle, MockBackend(two), cluster_manager="prod-manager")) == \
        "Selected cluster manager 'prod-manager' does not exist."                                          # :46-60
    docs = {"dev-manager": oracle.New("NodeState", MOCK_CLUSTERS)}
    assert err(d, oracle_env(oracle, MockBackend(two, docs), cluster_manager="dev-manager")) == "cluster_name must be specified"   # :61-79
    assert err(d, oracle_env(oracle, MockBackend(two, docs), cluster_manager="dev-manager", cluster_name="cluster_alpha")) == \
        "A cluster named 'cluster_alpha', does not exist."                                                 # :81-99
    docs = {"dev-manager": oracle.New("NodeState", MOCK_NODE_HOST)}
    assert err(d, oracle_env(oracle, MockBackend(two, docs), cluster_manager="dev-manager", cluster_name="dev_cluster")) == \
        "hostname must be specified"                                                                       # :101-119
    assert err(d, oracle_env(oracle, MockBackend(two, docs), cluster_manager="dev-manager", cluster_name="dev_cluster",
                             hostname="dev_node_host")) == "A node named 'dev_node_host', does not exist."  # :121-138


# ---- destroy/cluster_test.go, destroy/manager_test.go ------------------------------------------

def test_delete_cluster_and_manager_ladders(oracle):
    d = flows.destroy.delete_cluster
    two = ["dev-manager", "beta-manager"]
    assert err(d, oracle_env(oracle, MockBackend([]))) == "No cluster managers."
    assert err(d, oracle_env(oracle, MockBackend(two))) == "cluster_manager must be specified"
    assert err(d, oracle_env(oracle, MockBackend(two), cluster_manager="prod-manager")) == "Selected cluster manager 'prod-manager' does not exist."
    docs = {"dev-manager": oracle.New("ClusterState", MOCK_CLUSTERS)}
    assert err(d, oracle_env(oracle, MockBackend(two, docs), cluster_manager="dev-manager")) == "cluster_name must be specified"
    assert err(d, oracle_env(oracle, MockBackend(two, docs), cluster_manager="dev-manager", cluster_name="cluster_alpha")) == \
        "A cluster named 'cluster_alpha', does not exist."
    m = flows.destroy.delete_manager
    assert err(m, oracle_env(oracle, MockBackend([]))) == \
        "No cluster managers, please create a cluster manager before creating a kubernetes cluster."
    assert err(m, oracle_env(oracle, MockBackend(two))) == "cluster_manager must be specified"
    assert err(m, oracle_env(oracle, MockBackend(two), cluster_manager="prod-manager")) == "Selected cluster manager 'prod-manager' does not exist."


# ---- get/cluster_test.go, get/manager_test.go ----------------------------------------------------

def test_get_ladders(oracle):
    g = flows.get.get_cluster
    two = ["dev-manager", "beta-manager"]
    assert err(g, oracle_env(oracle, MockBackend([]))) == "No cluster managers."
    assert err(g, oracle_env(oracle, MockBackend(two))) == "cluster_manager must be specified"
    assert err(g, oracle_env(oracle, MockBackend(two), cluster_manager="prod-manager")) == "Selected cluster manager 'prod-manager' does not exist."
    docs = {"dev-manager": oracle.New("s", b'{"module":{}}')}
    assert err(g, oracle_env(oracle, MockBackend(two, docs), cluster_manager="dev-manager")) == "No clusters."       # get/cluster_test.go:63-87
    docs = {"dev-manager": oracle.New("s", MOCK_CLUSTERS)}
    assert err(g, oracle_env(oracle, MockBackend(two, docs), cluster_manager="dev-manager")) == "cluster_name must be specified"
    assert err(g, oracle_env(oracle, MockBackend(two, docs), cluster_manager="dev-manager", cluster_name="cluster_alpha")) == \
        "A cluster named 'cluster_alpha', does not exist."
    m = flows.get.get_manager
    assert err(m, oracle_env(oracle, MockBackend([]))) == "No cluster managers."
    assert err(m, oracle_env(oracle, MockBackend(two))) == "cluster_manager must be specified"
    assert err(m, oracle_env(oracle, MockBackend(two), cluster_manager="prod-manager")) == "Selected cluster manager 'prod-manager' does not exist."


# ---- util/backend_prompt_test.go:9-107 -----------------------------------------------------------

def test_backend_prompt_ladder():
    pb = flows.backend.prompt_for_backend

    def e(**kw):
        c = Config(environ={})
        for k, v in kw.items():
            c.Set(k, v)
        with pytest.raises(FlowError) as x:
            pb(c, lambda: None)
        return str(x.value)

    assert e() == "backend_provider must be provided"
    assert e(backend_provider="S3") == "Unsupported backend provider 'S3'"
    assert e(backend_provider="manta") == "triton_account must be specified"
    assert e(backend_provider="manta", triton_account="xyz") == "triton_key_path must be specified"
    assert e(backend_provider="manta", triton_account="xyz", triton_key_path="") == \
        "Unable to read private key: open : no such file or directory"
    assert e(backend_provider="manta", triton_account="xyz", triton_key_path="", triton_key_id="") == "triton_url must be specified"
    assert e(backend_provider="manta", triton_account="xyz", triton_key_path="", triton_key_id="", triton_url="xyz.triton.com") == \
        "manta_url must be specified"
    c = Config(environ={})
    c.Set("backend_provider", "local")
    assert pb(c, lambda: "the-local-backend") == "the-local-backend"


# ---- the settings store (cmd/root.go:48-67, docs/guide/silent-install-yaml.md) -------------------

def test_config_precedence_and_yaml(tmp_path):
    p = tmp_path / "c.yaml"
    p.write_text("Backend_Provider: local\nname: from-file\nnode_count: 3\nprivate_registry: \"\"\nnodes:\n  - hostname: w\n    node_count: 2\n    hosts: [a, b]\n")
    c = Config.from_yaml(str(p), environ={"NAME": "from-env", "SOURCE_REF": "v9"})
    assert c.GetString("backend_provider") == "local"                 # keys are case-insensitive
    assert c.GetString("name") == "from-env" and c.GetString("source_ref") == "v9"     # AutomaticEnv beats the file
    c.Set("name", "from-set")
    assert c.GetString("name") == "from-set"                          # viper.Set beats both
    c.Set("name", None)
    assert c.GetString("name") == "from-env"                          # a nil override does not shadow
    assert c.IsSet("private_registry") and c.GetString("private_registry") == ""    # set to the empty string IS set
    assert not c.IsSet("missing") and c.GetString("node_count") == "3"
    assert c.Get("nodes")[0]["hosts"] == ["a", "b"] and not c.GetBool("non-interactive")


# ---- configs #2, #3, #5 end to end ---------------------------------------------------------------

STUB = """#!/bin/sh
# terraform stub: records argv and a hash of the document it was handed, succeeds
echo "$@ $(sha256sum main.tf.json | cut -d' ' -f1)" >> "$TK_STUB_LOG"
exit 0
"""

MANAGER = dict(manager_cloud_provider="baremetal", name="dev", rancher_admin_password="s3cr<e>t&", host="10.0.0.1",
               ssh_user="ubuntu", bastion_host="", key_path="~/.ssh/id_rsa", private_registry="")
CLUSTER = dict(cluster_manager="dev", cluster_cloud_provider="baremetal", name="c1", k8s_version="v1.9.5-rancher1-1",
               k8s_network_provider="calico", private_registry="", k8s_registry="",
               nodes=[dict(rancher_host_label="etcd", node_count=3, hostname="c1-e", ssh_user="ubuntu", key_path="~/.ssh/id_rsa",
                           bastion_host="", hosts=["10.0.1.1", "10.0.1.2", "10.0.1.3"]),
                      dict(rancher_host_label="control", node_count=1, hostname="c1-c", ssh_user="ubuntu", key_path="~/.ssh/id_rsa",
                           bastion_host="", hosts=["10.0.2.1"]),
                      dict(rancher_host_label="worker", node_count=2, hostname="c1-w", ssh_user="ubuntu", key_path="~/.ssh/id_rsa",
                           bastion_host="bastion.example", hosts=["10.0.3.1", "10.0.3.2"])])


def run_lifecycle(make_env, root, log, monkeypatch):
    """create manager -> create cluster + 3 node groups -> add a node -> destroy node -> destroy cluster -> destroy manager"""
    stub_dir = os.path.join(root, "bin")
    os.makedirs(stub_dir, exist_ok=True)
    stub = os.path.join(stub_dir, "terraform")
    open(stub, "w").write(STUB)
    os.chmod(stub, os.stat(stub).st_mode | stat.S_IEXEC)
    monkeypatch.setenv("PATH", stub_dir + os.pathsep + os.environ["PATH"])
    monkeypatch.setenv("TK_STUB_LOG", log)
    snapshots = []

    def snap(env):
        p = os.path.join(root, "store", "dev", "main.tf.json")
        snapshots.append(open(p, "rb").read() if os.path.exists(p) else None)

    env = make_env(os.path.join(root, "store"), MANAGER)
    flows.create.new_manager(env); snap(env)                                        # config #2
    env = make_env(os.path.join(root, "store"), CLUSTER)
    flows.create.new_cluster(env); snap(env)                                        # config #3
    out = list(env.out)
    env = make_env(os.path.join(root, "store"), dict(cluster_manager="dev", cluster_name="c1", rancher_host_label="worker", node_count="1",
                                                     hostname="c1-w", ssh_user="ubuntu", key_path="k", bastion_host="", hosts=["10.0.3.9"]))
    assert flows.create.new_node(env) == ["c1-w-3"]; snap(env)                      # continues the series (create/node.go:350-380)
    env = make_env(os.path.join(root, "store"), dict(cluster_manager="dev", cluster_name="c1"))
    flows.get.get_cluster(env)
    env = make_env(os.path.join(root, "store"), dict(cluster_manager="dev", cluster_name="c1", hostname="c1-e-2"))
    flows.destroy.delete_node(env); snap(env)
    env = make_env(os.path.join(root, "store"), dict(cluster_manager="dev", cluster_name="c1"))
    flows.destroy.delete_cluster(env); snap(env)                                    # config #5 at CLI granularity
    env = make_env(os.path.join(root, "store"), dict(cluster_manager="dev"))
    flows.destroy.delete_manager(env); snap(env)
    return snapshots, out, open(log).read().splitlines()


def oracle_factory(oracle):
    from oracle import records as orec
    from oracle.backend_local import LocalBackend as OB

    def make(store, settings):
        cfg = Config(dict(settings), environ={})
        cfg.Set("non-interactive", True)
        return Env(state=oracle, backend=OB(store), render=orec.render, module_source=orec.module_source, config=cfg)
    return make


def product_factory(tk):
    def make(store, settings):
        cfg = Config(dict(settings), environ={})
        cfg.Set("non-interactive", True)
        return Env(state=tk, backend=tk.LocalBackend(store), render=tk.records.render, module_source=tk.records.module_source, config=cfg)
    return make


def check_lifecycle(snapshots, out, log, oracle):
    m, c, n, dn, dc, dm = snapshots
    doc = json.loads(c)
    assert sorted(doc["module"]) == ["cluster-manager", "cluster_baremetal_c1", "node_baremetal_c1_c1-c-1", "node_baremetal_c1_c1-e-1",
                                      "node_baremetal_c1_c1-e-2", "node_baremetal_c1_c1-e-3", "node_baremetal_c1_c1-w-1", "node_baremetal_c1_c1-w-2"]
    assert doc["module"]["node_baremetal_c1_c1-w-2"]["bastion_host"] == "bastion.example"
    assert doc["module"]["node_baremetal_c1_c1-e-2"]["rancher_host_labels"] == {"etcd": "true"}
    assert doc["terraform"]["backend"]["local"]["path"].endswith("/store/dev/terraform.tfstate")
    assert out == ["3 nodes added: c1-e-1, c1-e-2, c1-e-3", "1 node added: c1-c-1", "2 nodes added: c1-w-1, c1-w-2"]
    assert "node_baremetal_c1_c1-w-3" in json.loads(n)["module"] and "node_baremetal_c1_c1-e-2" not in json.loads(dn)["module"]
    assert sorted(json.loads(dc)["module"]) == ["cluster-manager"] and dm is None
    # what terraform was asked to do, in order; the hash shows which document it saw
    verbs = [" ".join(l.split()[:-1]) for l in log]
    assert verbs[:4] == ["init -force-copy", "apply -auto-approve", "init -force-copy", "apply -auto-approve"]
    assert verbs[6:8] == ["init -force-copy", "output -module cluster_baremetal_c1"]
    assert verbs[9] == "destroy -force -target=module.node_baremetal_c1_c1-e-2"
    assert verbs[11].startswith("destroy -force -target=module.cluster_baremetal_c1 -target=module.node_baremetal_c1_c1-c-1")
    assert verbs[13] == "destroy -force"
    assert log[1].split()[-1] == hashlib.sha256(m).hexdigest()       # apply saw exactly what was then persisted
    # destroy ran with the node still in the document -- as re-serialised after the reload, i.e. all maps, all sorted
    assert log[9].split()[-1] == hashlib.sha256(oracle.New("dev", n).Bytes()).hexdigest()


def test_lifecycle_on_the_oracle(oracle, tmp_path, monkeypatch):
    snaps, out, log = run_lifecycle(oracle_factory(oracle), str(tmp_path), str(tmp_path / "stub.log"), monkeypatch)
    check_lifecycle(snaps, out, log, oracle)


@pytest.mark.gpu
def test_lifecycle_product_matches_oracle(tk, oracle, tmp_path, monkeypatch):
    a = run_lifecycle(product_factory(tk), str(tmp_path / "gpu"), str(tmp_path / "gpu.log"), monkeypatch)
    b = run_lifecycle(oracle_factory(oracle), str(tmp_path / "ora"), str(tmp_path / "ora.log"), monkeypatch)
    check_lifecycle(*a, oracle)
    fix = lambda s: None if s is None else s.replace(str(tmp_path / "gpu").encode(), b"<root>").replace(str(tmp_path / "ora").encode(), b"<root>")
    assert [fix(s) for s in a[0]] == [fix(s) for s in b[0]]          # every persisted document, byte for byte
    assert a[1] == b[1] and [" ".join(l.split()[:-1]) for l in a[2]] == [" ".join(l.split()[:-1]) for l in b[2]]


def test_create_ladders(oracle, tmp_path):
    make = oracle_factory(oracle)
    e = lambda fn, s: err(fn, make(str(tmp_path / "s"), s))
    nm = flows.create.new_manager
    assert e(nm, {}) == "manager_cloud_provider must be specified"
    assert e(nm, dict(manager_cloud_provider="baremetal")) == "name must be specified"
    assert e(nm, dict(manager_cloud_provider="cloudy", name="x")) == "Unsupported cloud provider 'cloudy', cannot create manager"
    assert e(nm, dict(manager_cloud_provider="baremetal", name="x")) == "UI Admin Password must be specified"
    assert e(nm, dict(manager_cloud_provider="baremetal", name="x", rancher_admin_password="p")) == "host must be specified"
    assert e(nm, dict(manager_cloud_provider="baremetal", name="x", rancher_admin_password="p", private_registry="r")) == \
        "private_registry_username must be specified"
    nc = flows.create.new_cluster
    assert e(nc, {}) == "No cluster managers, please create a cluster manager before creating a kubernetes cluster."
    os.makedirs(tmp_path / "s" / "dev")
    assert e(nc, {}) == "cluster_manager must be specified"
    assert e(nc, dict(cluster_manager="nope")) == "Selected cluster manager 'nope' does not exist."
    assert e(nc, dict(cluster_manager="dev")) == "cluster_cloud_provider must be specified"
    assert e(nc, dict(cluster_manager="dev", cluster_cloud_provider="baremetal")) == "name must be specified"
    assert e(nc, dict(cluster_manager="dev", cluster_cloud_provider="baremetal", name="Bad_Name")) == "Invalid Cluster Name"
    assert e(nc, dict(cluster_manager="dev", cluster_cloud_provider="baremetal", name="ok")) == "k8s_version must be specified"
    nn = flows.create.new_node
    assert e(nn, dict(cluster_manager="dev")) == "not an object"       # an empty manager has no "module": Clusters() fails (create/node.go:97-100)


# ------------------------------------------------------------------ vSphere cluster + nodes (config #3, the other provider with no cloud API calls)

VSPHERE_CLUSTER = dict(cluster_manager="dev", cluster_cloud_provider="vsphere", name="v1", k8s_version="v1.9.5-rancher1-1",
                       k8s_network_provider="flannel", private_registry="", k8s_registry="",
                       vsphere_user="admin@vsphere.local", vsphere_password="p<w>&d", vsphere_server="vc.example", vsphere_datacenter_name="dc1",
                       vsphere_datastore_name="ds1", vsphere_resource_pool_name="pool", vsphere_network_name="VM Network",
                       vsphere_template_name="ubuntu-tpl", ssh_user="ubuntu", key_path="~/.ssh/id_rsa",
                       nodes=[dict(rancher_host_label="etcd", node_count=1, hostname="v1-e"),
                              dict(rancher_host_label="worker", node_count=2, hostname="v1-w")])


def run_vsphere(make_env, root, log, monkeypatch):
    stub_dir = os.path.join(root, "bin")
    os.makedirs(stub_dir, exist_ok=True)
    stub = os.path.join(stub_dir, "terraform")
    open(stub, "w").write(STUB)
    os.chmod(stub, os.stat(stub).st_mode | stat.S_IEXEC)
    monkeypatch.setenv("PATH", stub_dir + os.pathsep + os.environ["PATH"])
    monkeypatch.setenv("TK_STUB_LOG", log)
    monkeypatch.setenv("HOME", "/home/tester")
    store = os.path.join(root, "store")
    flows.create.new_manager(make_env(store, MANAGER))
    env = make_env(store, VSPHERE_CLUSTER)
    flows.create.new_cluster(env)
    created = open(os.path.join(store, "dev", "main.tf.json"), "rb").read()
    env2 = make_env(store, dict(cluster_manager="dev", cluster_name="v1", rancher_host_label="worker", node_count="2", hostname="v1-w",
                                vsphere_template_name="tpl2", ssh_user="root", key_path="/keys/k"))
    added = flows.create.new_node(env2)
    grown = open(os.path.join(store, "dev", "main.tf.json"), "rb").read()
    return created, grown, list(env.out), added


def check_vsphere(created, grown, out, added):
    doc = json.loads(created)
    assert sorted(doc["module"]) == ["cluster-manager", "cluster_vsphere_v1", "node_vsphere_v1_v1-e-1", "node_vsphere_v1_v1-w-1", "node_vsphere_v1_v1-w-2"]
    c = doc["module"]["cluster_vsphere_v1"]
    assert c["source"].endswith("//terraform/modules/vsphere-rancher-k8s?ref=master") and c["vsphere_network_name"] == "VM Network"
    n = doc["module"]["node_vsphere_v1_v1-w-2"]
    assert n["source"].endswith("//terraform/modules/vsphere-rancher-k8s-host?ref=master")
    assert n["vsphere_user"] == "admin@vsphere.local" and n["vsphere_password"] == "p<w>&d"        # copied from the cluster block
    assert n["vsphere_datastore_name"] == "${module.cluster_vsphere_v1.vsphere_datastore_name}"    # referenced, not copied
    assert n["vsphere_template_name"] == "ubuntu-tpl" and n["key_path"] == "/home/tester/.ssh/id_rsa" and n["hostname"] == "v1-w-2"
    assert b"p\\u003cw\\u003e\\u0026d" in created                                                  # escapeHTML on the way to disk
    assert out == ["1 node added: v1-e-1", "2 nodes added: v1-w-1, v1-w-2"]
    assert added == ["v1-w-3", "v1-w-4"]
    g = json.loads(grown)["module"]
    assert g["node_vsphere_v1_v1-w-4"]["vsphere_template_name"] == "tpl2" and g["node_vsphere_v1_v1-w-4"]["key_path"] == "/keys/k"


def test_vsphere_cluster_and_nodes_on_the_oracle(oracle, tmp_path, monkeypatch):
    check_vsphere(*run_vsphere(oracle_factory(oracle), str(tmp_path), str(tmp_path / "stub.log"), monkeypatch))


@pytest.mark.gpu
def test_vsphere_product_matches_oracle(tk, oracle, tmp_path, monkeypatch):
    a = run_vsphere(product_factory(tk), str(tmp_path / "gpu"), str(tmp_path / "gpu.log"), monkeypatch)
    b = run_vsphere(oracle_factory(oracle), str(tmp_path / "ora"), str(tmp_path / "ora.log"), monkeypatch)
    check_vsphere(*a)
    fix = lambda s: s.replace(str(tmp_path / "gpu").encode(), b"<root>").replace(str(tmp_path / "ora").encode(), b"<root>")
    assert fix(a[0]) == fix(b[0]) and fix(a[1]) == fix(b[1]) and a[2:] == b[2:]


def test_vsphere_ladders(oracle, tmp_path):
    make = oracle_factory(oracle)
    store = str(tmp_path / "s")
    e = lambda fn, s: err(fn, make(store, s))
    os.makedirs(tmp_path / "s" / "dev")
    base = dict(cluster_manager="dev", cluster_cloud_provider="vsphere", name="v1", k8s_version="v", k8s_network_provider="calico",
                private_registry="", k8s_registry="")
    nc = flows.create.new_cluster
    assert e(nc, base) == "vsphere_user must be specified."
    keys = ["vsphere_user", "vsphere_password", "vsphere_server", "vsphere_datacenter_name", "vsphere_datastore_name",
            "vsphere_resource_pool_name", "vsphere_network_name"]
    for i in range(1, len(keys)):
        assert e(nc, dict(base, **{k: "x" for k in keys[:i]})) == "%s must be specified." % keys[i]
    assert flows.create._expand_home("~/a") == os.path.expanduser("~") + "/a" and flows.create._expand_home("/abs") == "/abs"
    with pytest.raises(flows.FlowError):
        flows.create._expand_home("~other/a")


# ------------------------------------------------------------------ the reference's silent-install examples as loader fixtures

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "silent-install")


def test_silent_install_examples_load():
    """examples/silent-install/*.yaml (copied by tests/golden/gen_silent_install.py) through the settings
    loader: the keys it sees are the ones docs/guide/silent-install-yaml.md:11-49 documents.  (They need
    the Triton API to RENDER -- create/manager_triton.go:179 -- not to LOAD.)"""
    docs = json.load(open(os.path.join(GOLDEN, "docs_keys.json")))
    m = Config.from_yaml(os.path.join(GOLDEN, "create-manager.yaml"), environ={})
    keys = set(m._file)
    assert keys <= set(docs["Cluster Manager YAML"])                                   # nothing undocumented in the example
    assert set(docs["Cluster Manager YAML"]) - keys == {"manta_url"}                   # (documented for the manta backend only)
    assert m.GetString("backend_provider") == "local" and m.GetString("name") == "test-manager"
    assert m.IsSet("private_registry") and m.GetString("private_registry") == ""       # set to "" IS set: no prompt, no registry
    assert m.GetStringSlice("triton_network_names") == ["Joyent-SDC-Public"]
    assert m.GetString("triton_image_version") == "20180109" and m.GetString("rancher_admin_password") == "admin"
    assert not m.IsSet("triton_key_id") and not m.IsSet("manager_cloud_provider")      # the example leaves the provider to the prompt
    c = Config.from_yaml(os.path.join(GOLDEN, "cluster-triton-ha.yaml"), environ={"K8S_VERSION": "v1.10.0-rancher1-1"})
    assert set(docs["Cluster YAML"]) <= set(c._file)
    assert set(c._file) - set(docs["Cluster YAML"]) == {"triton_account", "triton_key_path", "triton_key_id", "triton_url"}
    assert c.GetString("cluster_cloud_provider") == "triton" and c.GetString("k8s_version") == "v1.10.0-rancher1-1"   # AutomaticEnv
    nodes = c.Get("nodes")
    assert [n["node_count"] for n in nodes] == [3, 3, 4] and [n["rancher_host_label"] for n in nodes] == ["etcd", "control", "worker"]
    assert all(n["triton_network_names"] == ["Joyent-SDC-Public"] for n in nodes)


def test_silent_install_cluster_example_stops_at_the_cloud_api(oracle, tmp_path):
    """The flow takes the example as far as the reference does without a network: manager selection and
    the ladder run, then the Triton renderer is refused by name."""
    store = tmp_path / "s"
    os.makedirs(store / "test-manager")
    from oracle import records as orec
    from oracle.backend_local import LocalBackend as OB
    cfg = Config.from_yaml(os.path.join(GOLDEN, "cluster-triton-ha.yaml"), environ={})
    cfg.Set("non-interactive", True)
    env = Env(state=oracle, backend=OB(str(store)), render=orec.render, module_source=orec.module_source, config=cfg)
    assert "triton" in err(flows.create.new_cluster, env)


# ------------------------------------------------------------------ the command line (cmd/root.go, cmd/create.go, cmd/destroy.go, cmd/get.go)

def cli(argv):
    main = importlib.import_module("triton-kubernetes_b200.__main__").main
    lines = []
    rc = main(argv, out=lines.append)
    return rc, lines


def test_cli_argument_checks(tmp_path):
    assert cli(["create"]) == (1, ['Error: "triton-kubernetes create" requires one argument'])
    assert cli(["destroy", "a", "b"]) == (1, ['Error: "triton-kubernetes destory" requires one argument'])      # (sic, cmd/destroy.go:22)
    assert cli(["get", "node"]) == (1, ['Error: invalid argument "node" for "triton-kubernetes get"'])
    assert cli(["create", "volume"]) == (1, ['Error: invalid argument "volume" for "triton-kubernetes create"'])
    rc, lines = cli(["create", "manager"])
    assert rc == 1 and "--non-interactive" in lines[-1]                                      # prompts are not part of this build
    empty = tmp_path / "empty.yaml"
    empty.write_text("{}\n")
    assert cli(["--non-interactive", "--config", str(empty), "create", "manager"]) == \
        (1, ["Running in non interactive mode", "Using config file: %s" % empty, "backend_provider must be provided"])
    cfg = tmp_path / "c.yaml"
    cfg.write_text("backend_provider: local\n")
    assert cli(["--config=%s" % cfg, "--non-interactive", "--store", str(tmp_path / "s"), "get", "manager"]) == \
        (1, ["Running in non interactive mode", "Using config file: %s" % cfg, "get manager called", "No cluster managers."])


def _yaml(path, settings):
    import yaml
    with open(path, "w") as f:
        yaml.safe_dump(settings, f)
    return str(path)


@pytest.mark.gpu
def test_cli_configs_2_and_3_match_the_oracle_flows(tk, oracle, tmp_path, monkeypatch):
    """BASELINE.json config #2 (`create manager --non-interactive`, local backend, terraform stubbed) and
    config #3 (`create cluster` with three node groups) through the command line, product underneath;
    the persisted documents equal what the same settings give through the oracle's flows."""
    stub_dir = tmp_path / "bin"
    os.makedirs(stub_dir)
    stub = stub_dir / "terraform"
    stub.write_text(STUB)
    os.chmod(stub, os.stat(stub).st_mode | stat.S_IEXEC)
    monkeypatch.setenv("PATH", str(stub_dir) + os.pathsep + os.environ["PATH"])
    monkeypatch.setenv("TK_STUB_LOG", str(tmp_path / "stub.log"))
    gpu, ora = str(tmp_path / "gpu"), str(tmp_path / "ora")
    rc, lines = cli(["--non-interactive", "--config", _yaml(tmp_path / "m.yaml", dict(MANAGER, backend_provider="local")), "--store", gpu,
                     "create", "manager"])
    assert (rc, lines[0], lines[2]) == (0, "Running in non interactive mode", "create manager called")
    rc, lines = cli(["--non-interactive", "--config", _yaml(tmp_path / "c.yaml", dict(CLUSTER, backend_provider="local")), "--store", gpu,
                     "create", "cluster"])
    assert rc == 0 and lines[-3:] == ["3 nodes added: c1-e-1, c1-e-2, c1-e-3", "1 node added: c1-c-1", "2 nodes added: c1-w-1, c1-w-2"]
    make = oracle_factory(oracle)
    flows.create.new_manager(make(ora, MANAGER))
    flows.create.new_cluster(make(ora, CLUSTER))
    a = open(os.path.join(gpu, "dev", "main.tf.json"), "rb").read().replace(gpu.encode(), b"<root>")
    b = open(os.path.join(ora, "dev", "main.tf.json"), "rb").read().replace(ora.encode(), b"<root>")
    assert a == b and len(json.loads(a)["module"]) == 8
    rc, lines = cli(["--non-interactive", "--config", _yaml(tmp_path / "g.yaml", dict(backend_provider="local", cluster_manager="dev", cluster_name="c1")),
                     "--store", gpu, "get", "cluster"])
    assert rc == 0 and lines[2] == "get cluster called"
    assert open(tmp_path / "stub.log").read().splitlines()[-1].startswith("output -module cluster_baremetal_c1 ")
