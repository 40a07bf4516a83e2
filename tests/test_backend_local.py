"""The local backend (backend/local/backend.go:14-132) against its oracle
restatement: same directory layout, same file bytes."""
import os

import pytest


def tree(root):
    out = {}
    for d, _, files in os.walk(root):
        for f in files:
            p = os.path.join(d, f)
            out[os.path.relpath(p, root)] = open(p, "rb").read()
    return out


def test_registry_operations_need_no_device(tk, tmp_path):
    from oracle.backend_local import LocalBackend as OB
    a, b = tk.LocalBackend(str(tmp_path / "a")), OB(str(tmp_path / "b"))
    assert a.States() == b.States() == []
    for name in ("zeta", "alpha", "mid.dle"):
        os.makedirs(tmp_path / "a" / name); os.makedirs(tmp_path / "b" / name)
    (tmp_path / "a" / "stray-file").write_text("x"); (tmp_path / "b" / "stray-file").write_text("x")
    assert a.States() == b.States() == ["alpha", "mid.dle", "zeta"]          # directories only, sorted (ioutil.ReadDir)
    a.DeleteState("alpha"); b.DeleteState("alpha")
    a.DeleteState("never-existed"); b.DeleteState("never-existed")             # os.RemoveAll: not an error
    assert a.States() == b.States() == ["mid.dle", "zeta"]
    pa, ra = a.StateTerraformConfig("zeta")
    pb, rb = b.StateTerraformConfig("zeta")
    assert pa == pb == "terraform.backend.local"
    assert ra.members == [("path", str(tmp_path / "a" / "zeta" / "terraform.tfstate"))]
    assert rb.members == [("path", str(tmp_path / "b" / "zeta" / "terraform.tfstate"))]
    s = a.State("brand-new")                                                  # no file yet: state.New(name, "{}") (:51-55)
    assert s.Name == b"brand-new" and s.Get("anything") == b""


def test_default_root_is_under_home(tk, tmp_path, monkeypatch):
    monkeypatch.setenv("HOME", str(tmp_path))
    b = tk.LocalBackend()
    assert os.path.isdir(tmp_path / ".triton-kubernetes") and b.States() == []


@pytest.mark.gpu
def test_create_manager_cluster_nodes_persist_and_reload(tk, tmp_path):
    """create/manager.go:103-148 then create/cluster.go:93-284 on the bare-metal records, through both backends."""
    from oracle import oracle as O
    from oracle import records as orec
    from oracle.backend_local import LocalBackend as OB
    results = []
    for be, mod, render in ((tk.LocalBackend(str(tmp_path / "gpu")), tk, tk.records.render), (OB(str(tmp_path / "ora")), O, orec.render)):
        st = be.State("dev")
        st.SetManager(render("bareMetalManagerTerraformConfig", {
            "source": orec.module_source("manager", "baremetal"), "name": "dev", "rancher_admin_password": "s3cr<e>t&",
            "host": "10.0.0.1", "ssh_user": "ubuntu", "key_path": "~/.ssh/id_rsa"}))
        path, cfg = be.StateTerraformConfig("dev")
        # the two roots differ; use one path text for both so the files can be compared byte for byte
        st.SetTerraformBackendConfig(path, mod.Record([("path", "/home/u/.triton-kubernetes/dev/terraform.tfstate")]))
        be.PersistState(st)
        st = be.State("dev")                                                     # reload: everything is a map now
        st.AddCluster("baremetal", "c1", render("bareMetalClusterTerraformConfig", {
            "source": orec.module_source("cluster", "baremetal"), "name": "c1",
            "rancher_api_url": "${module.cluster-manager.rancher_url}", "rancher_access_key": "${module.cluster-manager.rancher_access_key}",
            "rancher_secret_key": "${module.cluster-manager.rancher_secret_key}", "k8s_version": "v1.9.5-rancher1-1"}))
        st = mod.New("dev", st.Bytes())                                          # create/cluster.go:146-152
        ck = st.Clusters()[b"c1"]
        for i, host in enumerate(mod.getNewHostnames(list(st.Nodes(ck)), "c1-w", 3)):
            st.AddNode(ck, host, render("bareMetalNodeTerraformConfig", {
                "source": orec.module_source("node", "baremetal"), "hostname": host, "rancher_api_url": "${module.cluster-manager.rancher_url}",
                "rancher_host_labels": {"worker": "true"}, "host": "10.0.1.%d" % i}))
        be.PersistState(st)
        results.append((st.Bytes(), be.States(), sorted(be.State("dev").Nodes(ck).items())))
    assert results[0] == results[1]
    ga, oa = tree(tmp_path / "gpu"), tree(tmp_path / "ora")
    assert ga == oa and list(ga) == ["dev/main.tf.json"]
    text = ga["dev/main.tf.json"]
    # SURVEY.md 8a note 1: the reloaded cluster block is sorted, the node blocks of this run are in struct order
    assert text.index(b'"k8s_version"') < text.index(b'"name": "c1"')
    assert text.index(b'"source": "github.com/joyent/triton-kubernetes//terraform/modules/bare-metal-rancher-k8s-host') < text.index(b'"hostname": "c1-w-1"')
    assert b"s3cr\\u003ce\\u003et\\u0026" in text
    assert oct(os.stat(tmp_path / "gpu" / "dev" / "main.tf.json").st_mode & 0o777) == oct(0o644 & ~_umask())


def _umask():
    m = os.umask(0)
    os.umask(m)
    return m
