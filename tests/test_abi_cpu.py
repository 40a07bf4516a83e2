"""CPU-side checks of the product library: it loads, exports every symbol include/*.h
declare, its host logic (path operations, key parts, hostnames) matches the oracle, and the
bulk operations refuse to run without a GPU instead of falling back."""
import os
import random
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = "".join(open(os.path.join(ROOT, "include", h)).read() for h in sorted(os.listdir(os.path.join(ROOT, "include"))) if h.endswith(".h"))
    return re.findall(r"TK_API\s+[\w\s\*]+?\b(tk_\w+)\s*\(", text)


def test_library_exports_every_declared_symbol(tk):
    lib = tk.state.lib()
    declared = header_symbols()
    assert len(declared) >= 20
    assert sorted(declared) == sorted(tk.state.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.tk_version()


def test_library_contains_sm100a_code(tk):
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", tk.state.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(has_gpu(), reason="checks the no-device behaviour")
def test_bulk_ops_fail_loudly_without_a_device(tk):
    with pytest.raises(tk.StateError) as e:
        tk.New("x", b"{}")
    assert e.value.code == tk.state.TK_ECUDA and "no CPU fallback" in str(e.value)
    s = tk.State("x")
    s.SetManager({"name": "m"})
    for call in (s.Bytes, s.Stage):
        with pytest.raises(tk.StateError) as e:
            call()
        assert e.value.code == tk.state.TK_ECUDA
    # on a document whose tree is on the host, Clusters/Nodes are ordered-map range lookups: host logic
    assert s.Clusters() == {} and s.Nodes("cluster_a_b") == {}


def test_key_parts_error_needs_no_device(tk):
    s = tk.State("x")
    with pytest.raises(tk.StateError) as e:
        s.Nodes("cluster_dev-cluster")
    assert str(e.value) == ("Could not get cluster key parts, cluster does not follow format "
                            "`cluster_{provider}_{clusterName}` 'cluster_dev-cluster'")
    with pytest.raises(tk.StateError) as e:
        s.AddNode("nounderscore", "n", {})
    assert e.value.code == tk.state.TK_EKEYPARTS


def test_module_not_an_object_needs_no_device(tk):
    s = tk.State("x")
    with pytest.raises(tk.StateError) as e:
        s.Clusters()
    assert str(e.value) == "not an object"
    s._set("module", 1)
    with pytest.raises(tk.StateError) as e:
        s.Nodes("cluster_a_b")
    assert str(e.value) == "not an object"


def test_state_test_go_path_vectors_on_host(tk):
    # state/state_test.go:20-101 restated with values set through the ABI (no device needed)
    s = tk.State("AddState")
    assert s.Name == b"AddState"
    s.SetManager({"field": "test"})
    assert s.Get("module.cluster-manager.field") == b"test"
    s.SetTerraformBackendConfig("module.path.to.backend", {"field": "test"})
    assert s.Get("module.path.to.backend.field") == b"test"
    s.AddCluster("aws", "name", {"field": "test"})
    assert s.Get("module.cluster_aws_name.field") == b"test"
    s.AddNode("cluster_aws_cluster-name", "node-name", {"field": "test"})
    assert s.Get("module.node_aws_cluster-name_node-name.field") == b"test"
    s.Delete("module.cluster_aws_name.field")
    assert s.Get("module.cluster_aws_name.field") == b""


def test_gabs_error_values_on_host(tk):
    s = tk.State("x")
    s._set("foo", "bar")
    with pytest.raises(tk.StateError) as e:
        s._set("foo.bar", "new")                       # gabs_test.go:1045-1056
    assert str(e.value) == "encountered value collision whilst building path" and e.value.code == tk.state.TK_ECOLLISION
    s.Delete("nope")                                    # absent last key: no error (gabs.go:276-296)
    with pytest.raises(tk.StateError) as e:
        s.Delete("nope.deeper")
    assert str(e.value) == "not an object"
    with pytest.raises(tk.StateError) as e:
        s._set("a", b"{bad")
    assert e.value.code == tk.state.TK_ESYNTAX
    assert str(e.value) == "invalid character 'b' looking for beginning of object key string"


def test_value_parser_error_texts_match_oracle(tk, oracle):
    from test_oracle_encoding import ERROR_VECTORS
    for src, msg in ERROR_VECTORS:
        s = tk.State("x")
        with pytest.raises(tk.StateError) as e:
            s._set("a", src)
        assert str(e.value) == msg, src[:40]


@pytest.mark.parametrize("existing,name,count", [
    (["test-1", "test-2"], "test", 0), (["test-1", "test-2"], "test", -10), (["test-1", "test-2"], "bar", 1),
    (["test"], "test", 1), (["foo", "bar"], "test", 3), (["test-1", "test-2", "bar-3", "bar-4"], "test", 3),
    (["a-+7", "a--3", "a-9x", "a-", "a-007", "a-99999999999999999999"], "a", 2), ([], "n", 2),
])
def test_new_hostnames_match_oracle(tk, oracle, existing, name, count):
    assert tk.getNewHostnames(existing, name, count) == oracle.getNewHostnames(existing, name, count)


def random_ops(rng, n):
    keys = ["module", "a", "b", "terraform", "x.y", "", "module.cluster_p_c", "module.node_p_c_n", "a.b.c", "a.b"]
    vals = [{"name": "c", "hostname": "h"}, "s", 1, None, [1, {"k": "v"}], {}, {"deep": {"er": {"x": "y"}}}, True]
    for _ in range(n):
        op = rng.random()
        path = rng.choice(keys) if rng.random() < 0.7 else rng.choice(keys) + "." + rng.choice(keys)
        if op < 0.55:
            v = rng.choice(vals)
            yield ("set", path, v, rng.random() < 0.25 and isinstance(v, dict))
        elif op < 0.8:
            yield ("del", path)
        else:
            yield ("get", path)


@pytest.mark.parametrize("seed", range(20))
def test_random_path_ops_match_oracle_on_host(tk, oracle, seed):
    """Set/Delete/Get sequences, errors included, behave exactly like the oracle's (host logic only)."""
    rng = random.Random(seed)
    a, b = tk.State("x"), oracle.New("x", b"{}")
    for op in random_ops(rng, 300):
        ra = rb = None
        if op[0] == "set":
            va = tk.Record(op[2]) if op[3] else op[2]
            vb = oracle.Record(op[2]) if op[3] else op[2]
            try: a._set(op[1], va)
            except Exception as e: ra = str(e)
            try: b._set(op[1], vb)
            except Exception as e: rb = str(e)
        elif op[0] == "del":
            try: a.Delete(op[1])
            except Exception as e: ra = str(e)
            try: b.Delete(op[1])
            except Exception as e: rb = str(e)
        else:
            ra, rb = a.Get(op[1]), b.Get(op[1])
        assert ra == rb, op
