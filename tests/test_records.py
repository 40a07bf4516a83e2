"""Renderer records: the tables compiled into the library are pinned to the
reference's Go structs through tests/golden/struct_tables.json (extracted from create/*.go and
backend/local/backend.go by tests/golden/gen_struct_tables.py), and rendering follows
encoding/json's struct rules as restated in oracle/records.py."""
import json
import os
import random

import pytest

from oracle import records as orec

GOLDEN = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "struct_tables.json")))


def test_tables_match_the_reference_structs(tk):
    got = tk.records.tables()
    assert sorted(got) == sorted(GOLDEN) and len(got) == 22
    for name, fields in got.items():
        assert fields == GOLDEN[name]["fields"], name


def test_spot_checks_against_the_go_source():
    # hand-read from the reference, independent of the extraction script
    f = [x["json"] for x in GOLDEN["bareMetalNodeTerraformConfig"]["fields"]]          # create/node.go:19-41 + node_bare_metal.go:20-27
    assert f == ["source", "hostname", "rancher_api_url", "rancher_cluster_registration_token", "rancher_cluster_ca_checksum",
                 "rancher_host_labels", "rancher_agent_image", "rancher_registry", "rancher_registry_username",
                 "rancher_registry_password", "host", "bastion_host", "ssh_user", "key_path"]
    omit = {x["json"] for x in GOLDEN["bareMetalManagerTerraformConfig"]["fields"] if x["omitempty"]}  # manager.go:16-27 + manager_bare_metal.go:20-27
    assert omit == {"rancher_admin_password", "rancher_server_image", "rancher_agent_image", "rancher_registry",
                    "rancher_registry_username", "rancher_registry_password", "bastion_host", "ssh_user", "key_path"}
    assert [x["json"] for x in GOLDEN["tritonClusterTerraformConfig"]["fields"]][-4:] == \
        ["triton_account", "triton_key_path", "triton_key_id", "triton_url"]                       # cluster_triton.go:21-28
    assert all("node_count" != x["json"] and "-" != x["json"] for t in GOLDEN.values() for x in t["fields"])   # json:"-" (node.go:23)


@pytest.mark.parametrize("kind,provider,go", [
    ("manager", "triton", "tritonManagerTerraformConfig"), ("manager", "baremetal", "bareMetalManagerTerraformConfig"),
    ("cluster", "vsphere", "vSphereClusterTerraformConfig"), ("node", "aws", "awsNodeTerraformConfig"),
    ("node", "gcp", "gcpNodeTerraformConfig"), ("cluster", "azure", "azureClusterTerraformConfig")])
def test_struct_names_and_module_sources(tk, kind, provider, go):
    assert tk.records.struct_name(kind, provider) == go
    assert tk.records.module_source(kind, provider) == orec.module_source(kind, provider)
    assert tk.records.module_source(kind, provider, "example.com/fork", "v1") == orec.module_source(kind, provider, "example.com/fork", "v1")


def test_no_vsphere_manager(tk):
    with pytest.raises(tk.StateError):
        tk.records.struct_name("manager", "vsphere")          # the reference has no vSphere manager (create/manager.go:108-122)
    assert tk.records.module_source("cluster", "baremetal") == \
        "github.com/joyent/triton-kubernetes//terraform/modules/bare-metal-rancher-k8s?ref=master"   # create/cluster.go:310


def random_settings(rng, fields):
    s = {}
    for f in fields:
        r = rng.random()
        if r < 0.25:
            continue                                            # absent: Go zero value
        if f["type"] == "string":
            s[f["json"]] = rng.choice(["", "x", "${module.cluster-manager.rancher_url}", 'q"uo\\te', "tab\there", "é <&>"])
        elif f["type"] == "[]string":
            s[f["json"]] = rng.choice([[], ["a"], ["Joyent-SDC-Public", "n<2>"]])
        else:
            s[f["json"]] = random_settings(rng, GOLDEN[f["type"]]["fields"])
    return s


@pytest.mark.parametrize("go", sorted(GOLDEN))
def test_render_matches_oracle(tk, go):
    rng = random.Random(hash(go) & 0xffff)
    for _ in range(25):
        settings = random_settings(rng, GOLDEN[go]["fields"])
        got = json.loads(tk.records.render_json(go, settings), object_pairs_hook=list)
        want = json.loads(orec.render(go, settings).to_json(), object_pairs_hook=list)
        assert got == want


def test_render_rejects_unknown_fields_and_wrong_types(tk):
    with pytest.raises(tk.StateError) as e:
        tk.records.render_json("bareMetalClusterTerraformConfig", {"nam": "typo"})
    assert 'has no field "nam"' in str(e.value)
    with pytest.raises(tk.StateError):
        tk.records.render_json("bareMetalClusterTerraformConfig", {"name": 5})
    with pytest.raises(tk.StateError):
        tk.records.render_json("noSuchStruct", {})


def test_hand_derived_golden_block():
    """create/node_bare_metal.go:170-178 for one worker, marshalled by hand from the struct rules."""
    rec = orec.render("bareMetalNodeTerraformConfig", {
        "source": orec.module_source("node", "baremetal"), "hostname": "c1-w-1",
        "rancher_api_url": "${module.cluster-manager.rancher_url}",
        "rancher_cluster_registration_token": "${module.cluster_baremetal_c1.rancher_cluster_registration_token}",
        "rancher_cluster_ca_checksum": "${module.cluster_baremetal_c1.rancher_cluster_ca_checksum}",
        "rancher_host_labels": {"worker": "true"}, "host": "10.0.0.5", "ssh_user": "ubuntu", "key_path": "~/.ssh/id_rsa"})
    assert rec.to_json() == (
        b'{"source":"github.com/joyent/triton-kubernetes//terraform/modules/bare-metal-rancher-k8s-host?ref=master",'
        b'"hostname":"c1-w-1","rancher_api_url":"${module.cluster-manager.rancher_url}",'
        b'"rancher_cluster_registration_token":"${module.cluster_baremetal_c1.rancher_cluster_registration_token}",'
        b'"rancher_cluster_ca_checksum":"${module.cluster_baremetal_c1.rancher_cluster_ca_checksum}",'
        b'"rancher_host_labels":{"worker":"true"},"host":"10.0.0.5","bastion_host":"","ssh_user":"ubuntu","key_path":"~/.ssh/id_rsa"}')
