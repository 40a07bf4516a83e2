"""Every vector the reference's own tests hold for this path.

Each test names the reference test it restates (paths relative to /root/reference).  Tests
that take `impl` run twice: against the oracle on CPU (`-m "not gpu"`) and, marked gpu,
against the product through its C ABI on the B200.  Tests that take `oracle` replay the
vendored gabs vectors, which need gabs-only calls (compact String(), Path fan-out) that the
state.State surface does not have.
"""
import pytest

from oracle.oracle import OracleError, Record

KEY_PARTS_ERR = ("Could not get cluster key parts, cluster does not follow format "
                 "`cluster_{provider}_{clusterName}` 'cluster_dev-cluster'")


# ---- state/state_test.go ------------------------------------------------------------------

def test_get(impl):  # state/state_test.go:8-18
    s = impl.New("GetState", b'{"config":{"triton":{"key":"55fd4s","url":"https://api.storage.com"}}}')
    assert s.Get("config.triton.key") == b"55fd4s"


def test_set_manager(impl):  # state/state_test.go:20-34
    s = impl.New("AddState", b"{}")
    s.SetManager({"field": "test"})
    assert s.Get("module.cluster-manager.field") == b"test"


def test_set_terraform_backend_config(impl):  # state/state_test.go:36-50
    s = impl.New("AddState", b"{}")
    s.SetTerraformBackendConfig("module.path.to.backend", {"field": "test"})
    assert s.Get("module.path.to.backend.field") == b"test"


def test_add_cluster(impl):  # state/state_test.go:53-67
    s = impl.New("AddState", b"{}")
    s.AddCluster("aws", "name", {"field": "test"})
    assert s.Get("module.cluster_aws_name.field") == b"test"


def test_add_node(impl):  # state/state_test.go:69-83
    s = impl.New("AddState", b"{}")
    s.AddNode("cluster_aws_cluster-name", "node-name", {"field": "test"})
    assert s.Get("module.node_aws_cluster-name_node-name.field") == b"test"


def test_delete(impl):  # state/state_test.go:86-101
    s = impl.New("DelState", b'{"config":{"triton":{"key":"55fd4s","url":"https://api.storage.com"}}}')
    s.Delete("config.triton.key")
    assert s.Get("config.triton.key") == b""


CLUSTERS_DOC = b"""{
		"module":{
						"cluster_1":{"name":"dev_cluster"},
						"cluster_2":{"name":"beta_cluster"},
						"cluster_3":{"name":"prod_cluster"}
					}
				}"""

NODES_DOC = b"""{
    "module":{
      "node_triton_dev-cluster_1":{"hostname":"dev-worker1"},
      "node_triton_dev-cluster_2":{"hostname":"dev-etcd1"},
      "node_triton_dev-cluster_3":{"hostname":"dev-control1"},
      "node_aws_dev-cluster_1":{"hostname":"dev-control2"},
	  "node_aws_dev-cluster_2":{"hostname":"dev-control3"}
    }
    }"""


def test_get_clusters(impl):  # state/state_test.go:104-124
    s = impl.New("ClusterState", CLUSTERS_DOC)
    c = s.Clusters()
    assert len(c) == 3
    assert c == {b"dev_cluster": b"cluster_1", b"beta_cluster": b"cluster_2", b"prod_cluster": b"cluster_3"}


def test_get_nodes(impl):  # state/state_test.go:126-156
    ck = impl.New("ClusterState", b'{"config":{"triton":{"key":"cluster_triton_dev-cluster","url":"https://api.storage.com"}}}').Get("config.triton.key")
    s = impl.New("NodeState", NODES_DOC)
    n = s.Nodes(ck)
    assert len(n) == 3
    assert n == {b"dev-worker1": b"node_triton_dev-cluster_1", b"dev-etcd1": b"node_triton_dev-cluster_2",
                 b"dev-control1": b"node_triton_dev-cluster_3"}


def test_get_nodes_without_cloud_provider(impl):  # state/state_test.go:158-190, string at :184
    ck = impl.New("ClusterState", b'{"config":{"triton":{"key":"cluster_dev-cluster","url":"https://api.storage.com"}}}').Get("config.triton.key")
    s = impl.New("NodeState", NODES_DOC)
    with pytest.raises(Exception) as e:
        s.Nodes(ck)
    assert str(e.value) == KEY_PARTS_ERR


# ---- create/node_test.go:8-24 -------------------------------------------------------------

@pytest.mark.parametrize("existing,name,count,expected", [
    (["test-1", "test-2"], "test", 0, []),
    (["test-1", "test-2"], "test", -10, []),
    (["test-1", "test-2"], "bar", 1, ["bar-1"]),
    (["test"], "test", 1, ["test-1"]),
    (["foo", "bar"], "test", 3, ["test-1", "test-2", "test-3"]),
    (["test"], "test", 3, ["test-1", "test-2", "test-3"]),
    (["test-1", "test-2", "bar-3", "bar-4"], "test", 3, ["test-3", "test-4", "test-5"]),
])
def test_get_new_hostnames(impl, existing, name, count, expected):
    assert impl.getNewHostnames(existing, name, count) == expected


# ---- destroy/*_test.go, get/*_test.go fixtures that reach Clusters()/Nodes() ----------------

MOCK_CLUSTERS = b"""{
								"module":{
												"cluster_1":{"name":"dev_cluster"},
												"cluster_2":{"name":"beta_cluster"},
												"cluster_3":{"name":"prod_cluster"}
								}
}"""  # destroy/cluster_test.go:11-17
MOCK_NODE_HOST = b"""{
								"module":{
												"cluster_triton_dev-cluster":{"name":"dev_cluster"}
								}
}"""  # destroy/node_test.go:11-15


def test_destroy_fixtures(impl):
    # destroy/node_test.go:81-97: "cluster_alpha" is not among the cluster names
    assert b"cluster_alpha" not in impl.New("NodeState", MOCK_CLUSTERS).Clusters()
    # destroy/node_test.go:99-138: the cluster resolves, it has no nodes
    s = impl.New("NodeState", MOCK_NODE_HOST)
    assert s.Clusters() == {b"dev_cluster": b"cluster_triton_dev-cluster"}
    assert s.Nodes(b"cluster_triton_dev-cluster") == {}


def test_get_cluster_no_clusters(impl):  # get/cluster_test.go:63-87: empty "module" -> no clusters
    assert impl.New("s", b'{"module":{}}').Clusters() == {}


# ---- vendored gabs_test.go: compact, sorted-key output and the gabs error values ------------

def test_gabs_basic(oracle):  # gabs_test.go:11-43
    sample = b'{"test":{"value":10},"test2":20}'
    s = oracle.New("g", sample)
    assert s.PathString("test.value") == b"10"
    assert s.Get("test2.value") == b""
    assert s.PathString("test2") == b"20"
    assert s.BytesCompact() == sample


@pytest.mark.parametrize("path,exists", [  # gabs_test.go:45-75
    ("one.two.three", False), ("test", True), ("test.value", True), ("test2", True),
    ("test2.value", False), ("test.value2", False), ("test.VALUE", False)])
def test_gabs_exists(oracle, path, exists):
    s = oracle.New("g", b'{"test":{"value":10},"test2":20}')
    assert (s.PathString(path) != b"{}") == exists


@pytest.mark.parametrize("doc", [  # gabs_test.go:77-132
    b'{"foo":{"bar":{"baz":[10, 2, 3]}}}',
    b'{"foo":{"bar":[{"baz":10},{"baz":2},{"baz":3}]}}',
    b'{"foo":[{"bar":{"baz":10}},{"bar":{"baz":2}},{"bar":{"baz":3}}]}',
    b'[{"foo":{"bar":{"baz":10}}},{"foo":{"bar":{"baz":2}}},{"foo":{"bar":{"baz":3}}}]'])
def test_gabs_exists_with_arrays(oracle, doc):
    s = oracle.New("g", doc)
    assert s.PathString("foo.bar.baz") == b"[10,2,3]"
    assert s.PathString("foo.bar.baz_NOPE") == b"{}"


def test_gabs_find_array(oracle):  # gabs_test.go:180-236
    s = oracle.New("g", b'{"test":{"array":[{"value":1}, {"value":2}, {"value":3}]}}')
    assert s.PathString("test.array.value") == b"[1,2,3]"
    s = oracle.New("g", b"""{
			"test":{
				"array":[
						{"values":[{"more":1},{"more":2},{"more":3}]},
						{"values":[{"more":4},{"more":5},{"more":6}]},
						{"values":[{"more":7},{"more":8},{"more":9}]}
					]
				}
			}""")
    assert s.PathString("test.array.values.more") == b"[[1,2,3],[4,5,6],[7,8,9]]"


def test_gabs_deletes(oracle):  # gabs_test.go:238-265
    s = oracle.New("g", b"""{
		"outter":{
			"inner":{"value1":10,"value2":22,"value3":32},
			"alsoInner":{"value1":20,"value2":42,"value3":92}
		}
	}""")
    s.Delete("outter.inner.value2")
    s.Delete("outter.alsoInner.value1")
    assert s.BytesCompact() == b'{"outter":{"alsoInner":{"value2":42,"value3":92},"inner":{"value1":10,"value3":32}}}'


def test_gabs_examples2(oracle):  # gabs_test.go:314-359
    s = oracle.New("g", b"{}")
    s._set("outter.inner.value", 10)
    s._set("outter.inner.value2", 20)
    s._set("outter.inner2.value3", 30)
    assert s.BytesCompact() == b'{"outter":{"inner":{"value":10,"value2":20},"inner2":{"value3":30}}}'
    s = oracle.New("g", b"{}")
    s._set("array", [10, 20, 30])
    expected = b"""{
      "array": [
        10,
        20,
        30
      ]
    }"""
    assert s.BytesIndent("    ", "  ") == expected


def test_gabs_examples3(oracle):  # gabs_test.go:361-376
    s = oracle.New("g", b"{}")
    s._set("foo.array", [10, 20, 30])
    assert s.BytesCompact() == b'{"foo":{"array":[10,20,30]}}'


def test_gabs_modify(oracle):  # gabs_test.go:392-420
    s = oracle.New("g", b'{"test":{"value":10},"test2":20}')
    s._set("test.value", 45.0)
    assert s.PathString("test.value") == b"45"
    assert s.BytesCompact() == b'{"test":{"value":45},"test2":20}'
    assert s.PathString("test") == b'{"value":45}'


def test_gabs_children(oracle):  # gabs_test.go:422-494 (first block)
    s = oracle.New("g", b'{"objectOne":{},"objectTwo":{},"objectThree":{}}')
    for k in ("objectOne", "objectTwo", "objectThree"):
        s._set(k + ".child", "hello world")
    assert s.BytesCompact() == (b'{"objectOne":{"child":"hello world"},"objectThree":{"child":"hello world"}'
                                b',"objectTwo":{"child":"hello world"}}')


def test_gabs_shorthand(oracle):  # gabs_test.go:812-858
    s = oracle.New("g", b"""{
		"outter":{"inner":{"value":5,"value2":10,"value3":11},"inner2":{}},
		"outter2":{"inner":0}
	}""")
    assert s.PathString("outter.doesntexist.alsodoesntexist.inner.value") == b"{}"
    assert s.PathString("outter.inner.value2") == b"10"
    s._set("outter2.inner", {"value": 5, "value2": 10, "value3": 11})
    assert s.BytesCompact() == (b'{"outter":{"inner":{"value":5,"value2":10,"value3":11},"inner2":{}}'
                                b',"outter2":{"inner":{"value":5,"value2":10,"value3":11}}}')


@pytest.mark.parametrize("bad", [b'{dfads"', b"", b"{}\n{}"])  # gabs_test.go:860-876
def test_gabs_invalid(oracle, bad):
    with pytest.raises(OracleError):
        oracle.New("g", bad)


def test_gabs_missing_path_string(oracle):  # gabs_test.go:878-887
    assert oracle.New("g", b"{}").PathString("Doesn't exist") == b"{}"


def test_gabs_creation(oracle):  # gabs_test.go:890-914
    s = oracle.New("g", b"{}")
    s._set("test.inner", {})
    s._set("test.inner.first", 10)
    s._set("test.inner.second", 20)
    s._set("test.inner.array", ["first element of the array", 2, "three"])
    assert s.BytesCompact() == (b'{"test":{"inner":{"array":["first element of the array",2,"three"],'
                                b'"first":10,"second":20}}}')


def test_gabs_no_type_children(oracle):  # gabs_test.go:1017-1030: ChildrenMap on a number -> ErrNotObj
    s = oracle.New("g", b'{"module":1}')
    with pytest.raises(OracleError) as e:
        s.Clusters()
    assert str(e.value) == "not an object"


def test_gabs_nil_set(oracle):  # gabs_test.go:1045-1056
    s = oracle.New("g", b"null")          # Container{nil}
    s._set("foo", "bar")                   # nil root becomes a map
    with pytest.raises(OracleError) as e:
        s._set("foo.bar", "new")           # "foo" holds a string
    assert str(e.value) == "encountered value collision whilst building path"
    assert s.BytesCompact() == b'{"foo":"bar"}'


# ---- SURVEY.md 8a notes: behaviour the reference exhibits but does not test -----------------

def test_record_is_opaque_until_reparsed(impl):  # create/cluster.go:146-152 works around this
    s = impl.New("s", b"{}")
    s.AddCluster("baremetal", "c1", impl.Record([("source", "x"), ("name", "c1")]))
    assert s.Get("module.cluster_baremetal_c1.name") == b""
    assert s.Clusters() == {}
    s2 = impl.New("s", s.Bytes())
    assert s2.Get("module.cluster_baremetal_c1.name") == b"c1"
    assert s2.Clusters() == {b"c1": b"cluster_baremetal_c1"}


def test_dotted_cluster_name_nests(impl):  # gabs.go:206-208 + create/cluster.go:314
    s = impl.New("s", b"{}")
    s.AddCluster("x", "a.b", {"name": "a.b"})
    assert s.Bytes() == b'{\n\t"module": {\n\t\t"cluster_x_a": {\n\t\t\t"b": {\n\t\t\t\t"name": "a.b"\n\t\t\t}\n\t\t}\n\t}\n}'


def test_delete_absent_last_key_is_not_an_error(impl):  # gabs.go:276-296
    s = impl.New("s", b'{"module":{"a":{}}}')
    s.Delete("module.nope")
    s.Delete("module.a.nope")
    with pytest.raises(Exception) as e:
        s.Delete("module.nope.deeper")
    assert str(e.value) == "not an object"
    with pytest.raises(Exception):
        impl.New("s", b"null").Delete("a")
