"""Parity proper: the CUDA path through the C ABI against the oracle, bit for bit.

Everything here needs a B200 (`-m gpu`).  Nothing reads /root/reference.
"""
import importlib
import json
import random

import pytest

from test_abi_cpu import random_ops
from test_oracle_encoding import ERROR_VECTORS, FLOAT_VECTORS, STRING_VECTORS, rand_tree

pytestmark = pytest.mark.gpu

synth = importlib.import_module("triton-kubernetes_b200.synth")


def same(tk, oracle, raw):
    a, b = tk.New("x", raw), oracle.New("x", raw)
    assert a.Bytes() == b.Bytes()
    return a, b


def test_kernels_really_launch(tk):
    before = tk.state.kernel_launches()
    s = tk.New("x", b'{"module":{"cluster_a_b":{"name":"b"}}}')
    s.Bytes(); s.Clusters(); s.Nodes("cluster_a_b")
    assert tk.state.kernel_launches() - before >= 4   # the parse, row_offsets + emit, the module index


def test_new_keeps_canonical_documents_on_the_device(tk, oracle):
    """Text already in encoding/json's order is serialised straight from the tape the GPU parser built;
    anything that needs re-ordering, de-duplication or number re-formatting builds the host tree."""
    canon = oracle.New("x", synth.document(5, 4)).Bytes()
    for raw, resident in [(canon, True), (rb'{"a":{"b":1,"c":[true,null,"x"]},"b":"\u00e9\n"}', True),
                          (b'{"b":1,"a":2}', False), (b'{"a":1,"a":2}', False), (b'{"a":1.50}', False),
                          (b'{"a":{"y":1,"x":2}}', False), (b'[{"k":1},{"k":1}]', True), (rb'{"a":"x","a\u0000":"y"}', True)]:
        s = tk.New("x", raw)
        assert s.host_tree_built == (not resident), raw[:40]
        assert s.Bytes() == oracle.New("x", raw).Bytes()
        assert s.host_tree_built == (not resident)          # Bytes() alone never builds the tree
    s = tk.New("x", canon)
    assert s.Get("module.cluster-manager.name") == b"bench" and not s.host_tree_built    # (from the module index)
    assert s.Get("module.cluster-manager.source") == oracle.New("x", canon).Get("module.cluster-manager.source") and s.host_tree_built
    assert s.Bytes() == canon


@pytest.mark.parametrize("raw,want", [
    (b"null", b"{}"), (b" {} ", b"{}"), (b"[]", b"[]"), (b'"s"', b'"s"'), (b"[[],{}]", b"[\n\t[],\n\t{}\n]"),
    (b'{"a":{"b":[1]}}', b'{\n\t"a": {\n\t\t"b": [\n\t\t\t1\n\t\t]\n\t}\n}'), (b"true", b"true"), (b"-0", b"-0"),
    (b'{"":""}', b'{\n\t"": ""\n}'),
])
def test_root_forms(tk, oracle, raw, want):
    assert tk.New("x", raw).Bytes() == want == oracle.New("x", raw).Bytes()


@pytest.mark.parametrize("src,want", STRING_VECTORS)
def test_string_vectors(tk, src, want):
    assert tk.New("x", src).Bytes() == want
    assert tk.New("x", b"{" + src + b":1}").Bytes() == b"{\n\t" + want + b": 1\n}"


def test_float_vectors(tk, oracle):
    raw = ("[" + ",".join(s for s, _ in FLOAT_VECTORS) + "]").encode()
    assert tk.New("x", raw).Bytes() == ("[\n\t" + ",\n\t".join(w for _, w in FLOAT_VECTORS) + "\n]").encode()


@pytest.mark.parametrize("src,msg", ERROR_VECTORS)
def test_error_vectors(tk, src, msg):
    with pytest.raises(tk.StateError) as e:
        tk.New("x", src)
    assert str(e.value) == msg and e.value.code == tk.state.TK_ESYNTAX


def test_raw_invalid_utf8_in_path_key(tk, oracle):
    a, b = tk.State("x"), oracle.New("x", b"{}")
    for s in (a, b):
        s._set(b"a\xffb.\xe2\x80\xa8", b'"v\xc3"')
    assert a.Bytes() == b.Bytes() == b'{\n\t"a\\ufffdb": {\n\t\t"\\u2028": "v\xef\xbf\xbd"\n\t}\n}'


@pytest.mark.parametrize("seed", range(30))
def test_random_documents(tk, oracle, seed):
    rng = random.Random(77000 + seed)
    tree = {"k%d" % i: rand_tree(rng) for i in range(rng.randrange(1, 40))}
    raw = json.dumps(tree, ensure_ascii=rng.random() < 0.5).encode()
    a, b = same(tk, oracle, raw)
    # idempotence through the GPU path
    assert tk.New("x", a.Bytes()).Bytes() == a.Bytes()


def hostile_bytes(rng, n):
    pool = [b"a", b"\\n", b"\\\"", b"\\u0000", b"\\u2028", b"\\ud83d\\ude00", b"\\ud83d", b"<", b"&", b"\xc3\xa9", b"\xe2\x80\xa8",
            b"\xe2\x80\xa9", b"\xff", b"\xe2\x82", b"\xf0\x9f\x98\x80", b"\xed\xa0\x80", b"\x7f", b" ", b"\\t", b"\xc0\xaf", b"\xf4\x90\x80\x80"]
    return b"".join(rng.choice(pool) for _ in range(n))


@pytest.mark.parametrize("seed", range(10))
def test_hostile_strings(tk, oracle, seed):
    """Escapes, invalid UTF-8, U+2028/9 and multi-byte sequences at every alignment and across staging boundaries."""
    rng = random.Random(4242 + seed)
    members = []
    for i in range(200):
        k = hostile_bytes(rng, rng.randrange(0, 8))
        v = hostile_bytes(rng, rng.randrange(0, 60))
        members.append(b'"%d' % i + k + b'":"' + v + b'"')
    same(tk, oracle, b"{" + b",".join(members) + b"}")


@pytest.mark.parametrize("n", [1, 31, 32, 33, 255, 256, 257, 3000, 3071, 3072, 3073, 4095, 4096, 4097, 70000, 1 << 20])
def test_long_strings(tk, oracle, n):
    """One string of n bytes among short ones: covers a warp tile that outgrows its shared-memory staging."""
    rng = random.Random(n)
    body = hostile_bytes(rng, n // 2) if n < 5000 else (b"x<\\n" * (n // 4))
    raw = b'{"a":"1","b":"' + body + b'","c":["' + b"y" * n + b'"],"d":"2"}'
    same(tk, oracle, raw)


@pytest.mark.parametrize("run", [1, 2, 3, 7, 8, 9, 62, 63, 64, 65, 66, 127, 128, 129, 511, 512, 513, 575, 576, 577, 2047, 2048, 8191, 8192, 8193,
                                 16385, 40000, 40001])
@pytest.mark.parametrize("lead", [0, 1, 5, 61, 62, 63, 2040, 8185])
def test_backslash_runs(tk, oracle, run, lead):
    """A run of backslashes before a quote, of every parity and at every alignment against the
    64-byte chunks, the warps and the 8 KiB tiles of the token scan.  Short runs are resolved by the
    look-behind of the fast kernel; runs past 512 bytes send the parse through the carry protocol.
    An odd run escapes the quote: the string then ends at the next one (or the text is invalid)."""
    raw = b'{"k":"' + b"x" * lead + b"\\" * run + b'","z":"\\\\"}'
    try:
        want = oracle.New("x", raw).Bytes()
    except oracle.OracleError as e:
        with pytest.raises(tk.StateError) as got:
            tk.New("x", raw)
        assert str(got.value) == str(e)
        return
    assert tk.New("x", raw).Bytes() == want


@pytest.mark.parametrize("depth", [1, 2, 31, 100, 1000, 10000])
def test_deep_nesting(tk, oracle, depth):
    same(tk, oracle, b"[" * depth + b"]" * depth)
    same(tk, oracle, b'{"a":' * depth + b'"x"' + b"}" * depth)


@pytest.mark.parametrize("n", [0, 1, 255, 256, 257, 513, 5000])
def test_row_counts_around_tile_boundaries(tk, oracle, n):
    same(tk, oracle, b"[" + b",".join(b'"%d"' % i for i in range(n)) + b"]")
    same(tk, oracle, b"{" + b",".join(b'"k%d":%d' % (i, i) for i in range(n)) + b"}")


def test_duplicate_and_unsorted_keys(tk, oracle):
    same(tk, oracle, b'{"b":1,"a":1,"B":1,"\xc3\xa9":1,"aa":1,"":1,"a\\u0000":1,"~":1,"a":{"z":1,"y":2},"\\u0061":3}')


@pytest.mark.parametrize("seed", range(12))
def test_random_op_sequences(tk, oracle, seed):
    """Interleaved Set/Delete/Get with maps and records; Bytes/Clusters/Nodes compared along the way."""
    rng = random.Random(900 + seed)
    a, b = tk.New("x", b'{"module":{}}'), oracle.New("x", b'{"module":{}}')
    for step, op in enumerate(random_ops(rng, 200)):
        ra = rb = None
        for s, mod in ((a, tk), (b, oracle)):
            r = None
            try:
                if op[0] == "set":
                    s._set(op[1], mod.Record(op[2]) if op[3] else op[2])
                elif op[0] == "del":
                    s.Delete(op[1])
                else:
                    r = s.Get(op[1])
            except Exception as e:
                r = str(e)
            if s is a: ra = r
            else: rb = r
        assert ra == rb, op
        if step % 20 == 19:
            assert a.Bytes() == b.Bytes()
            for call in ("Clusters", lambda s: s.Nodes("cluster_p_c")):
                fa = (lambda s: s.Clusters()) if call == "Clusters" else call
                try: xa = fa(a)
                except Exception as e: xa = str(e)
                try: xb = fa(b)
                except Exception as e: xb = str(e)
                assert xa == xb


def test_records_serialise_in_given_order_and_stay_opaque(tk, oracle):
    for mod in (tk, oracle):
        s = mod.New("m", b"{}")
        s.SetManager(mod.Record([("source", "s"), ("name", "m"), ("host", "h")]))
        s.AddCluster("baremetal", "c1", mod.Record([("source", "x"), ("name", "c1"), ("k8s_version", "v")]))
        s.AddNode("cluster_baremetal_c1", "n1", mod.Record([("source", "y"), ("hostname", "n1"),
                  ("rancher_host_labels", mod.Record([("worker", "true")])), ("host", "10.0.0.1")]))
        s.AddNode("cluster_baremetal_c1", "n2", {"hostname": "n2", "source": "y"})
        if mod is tk:
            got = (s.Bytes(), s.Clusters(), s.Nodes("cluster_baremetal_c1"), s.Get("module.cluster-manager.name"))
        else:
            want = (s.Bytes(), s.Clusters(), s.Nodes("cluster_baremetal_c1"), s.Get("module.cluster-manager.name"))
    assert got == want
    assert got[1] == {} and got[2] == {b"n2": b"node_baremetal_c1_n2"} and got[3] == b""
    assert b'"source": "x",\n\t\t\t"name": "c1"' in got[0]       # declaration order kept
    re = tk.New("m", got[0])                                         # create/cluster.go:146-152: reload
    assert re.Clusters() == {b"c1": b"cluster_baremetal_c1"}
    assert set(re.Nodes("cluster_baremetal_c1")) == {b"n1", b"n2"}


def test_prefix_semantics(tk, oracle):
    """strings.Index(key, prefix) == 0: overlapping names, children that are not maps, missing fields, duplicates."""
    doc = {"module": {
        "cluster_a_x": {"name": "x"}, "cluster_a_x_y": {"name": "x"}, "cluster_": {"name": "empty"}, "cluster": {"name": "no"},
        "xcluster_a": {"name": "no"}, "cluster_b_arr": [{"name": "no"}], "cluster_b_str": "s", "cluster_b_null": None,
        "cluster_b_noname": {"nam": "x"}, "cluster_b_numname": {"name": 1}, "cluster_c_z": {"name": "z"},
        "node_a_x_1": {"hostname": "h1"}, "node_a_x_y_1": {"hostname": "h2"}, "node_a_x": {"hostname": "h3"},
        "node_a_x_": {"hostname": "h4"}, "node_a_xy_1": {"hostname": "h5"}, "node_a_x_2": {"hostname": "h1"},
        "node_a_x_3": {"hostname": 5}, "node_a_x_4": ["hostname"], "node_b_x_1": {"hostname": "h6"}}}
    raw = json.dumps(doc).encode()
    a, b = tk.New("x", raw), oracle.New("x", raw)
    assert a.Clusters() == b.Clusters()
    for ck in ("cluster_a_x", "cluster_a_x_y", "cluster_a_xy", "cluster_b_x", "cluster_a_", "cluster__", "a_b_c", "cluster_a_x_extra_parts"):
        assert a.Nodes(ck) == b.Nodes(ck), ck
    assert a.Nodes("cluster_a_x") == {b"h1": b"node_a_x_2", b"h2": b"node_a_x_y_1", b"h4": b"node_a_x_"}


def test_long_prefix(tk, oracle):
    name = "n" * 700                                   # longer than the kernel's shared-memory prefix buffer
    doc = {"module": {"node_p_%s_1" % name: {"hostname": "a"}, "node_p_%sX_1" % name[:-1]: {"hostname": "b"},
                      "node_p_%s_2" % name: {"hostname": "c"}}}
    raw = json.dumps(doc).encode()
    ck = "cluster_p_" + name
    assert tk.New("x", raw).Nodes(ck) == oracle.New("x", raw).Nodes(ck) and len(oracle.New("x", raw).Nodes(ck)) == 2


@pytest.mark.parametrize("clusters,nodes", [(3, 5), (40, 25)])
def test_config4_small(tk, oracle, clusters, nodes):
    raw = synth.document(clusters, nodes)
    a, b = same(tk, oracle, raw)
    assert a.Clusters() == b.Clusters() and len(a.Clusters()) == clusters
    for c in range(clusters):
        assert a.Nodes(synth.cluster_key(c)) == b.Nodes(synth.cluster_key(c))


def test_config5_destroy_sweep_small(tk, oracle):
    """destroy/cluster.go:126-158 at library granularity: Nodes(ck) -> Delete each node -> Delete the cluster."""
    raw = synth.document(30, 12)
    a, b = tk.New("x", raw), oracle.New("x", raw)
    for s in (a, b):
        for name, ck in sorted(s.Clusters().items()):
            for host, nk in sorted(s.Nodes(ck).items()):
                s.Delete(b"module." + nk)
            s.Delete(b"module." + ck)
            s.Delete(b"module." + ck)          # deleting an absent key is not an error
    want = json.dumps({"module": {"cluster-manager": synth.manager_record()},
                       "terraform": {"backend": {"local": {"path": "/root/.triton-kubernetes/bench/terraform.tfstate"}}}},
                      indent="\t", sort_keys=True).encode()
    assert a.Bytes() == b.Bytes() == want


def test_config4_full_size(tk, oracle):
    """BASELINE.json config #4 at full size: 1,000 clusters x 100 nodes, 62,177,405 bytes."""
    raw = synth.document()
    a = tk.New("bench", raw)
    text = a.Bytes()
    assert len(text) == 62177405
    b = oracle.New("bench", raw)
    assert text == b.Bytes()
    ca = a.Clusters()
    assert ca == b.Clusters() and len(ca) == 1000
    for c in (0, 1, 499, 998, 999):
        na = a.Nodes(synth.cluster_key(c))
        assert na == b.Nodes(synth.cluster_key(c)) and len(na) == 100
    # size-independent properties
    again = tk.New("bench", text)
    assert again.Bytes() == text                                    # serialise(parse(x)) is a fixed point
    assert sum(len(a.Nodes(ck)) for ck in list(ca.values())[::50]) == 20 * 100
    # an edit moves the document to the host tree; Bytes() then flattens it (the 101,001 members of
    # "module" by several threads) and serialises the uploaded tape
    for s in (a, b):
        s.Delete("module." + synth.node_key(7, 7))
        s.AddNode(synth.cluster_key(3), "extra<&>", {"hostname": "extra<&>", "n": [1, 2.5, None, True, {}], "s": "\u00e9\n"})
    assert len(a.Nodes(synth.cluster_key(7))) == 99 and a.Nodes(synth.cluster_key(3)) == b.Nodes(synth.cluster_key(3))
    edited = a.Bytes()
    assert edited == b.Bytes() and len(edited) != len(text)


def _module_doc(rng, n):
    """Children of "module" with every oddity Clusters()/Nodes() must cope with."""
    mod = {}
    for i in range(n):
        kind = rng.choice(["cluster", "node", "node", "node", "other"])
        prov = rng.choice(["a", "b", "aa"])
        cl = rng.choice(["x", "y", "x_y", "xy"])
        key = {"cluster": "cluster_%s_%s" % (prov, cl), "node": "node_%s_%s_%d" % (prov, cl, rng.randrange(40)),
               "other": "zz%d" % i}[kind]
        r = rng.random()
        name = rng.choice(["n%d" % rng.randrange(12), "dup", "", "é", "a<b"])
        if r < 0.6:
            child = {"name": name, "hostname": rng.choice(["h%d" % rng.randrange(30), "dup", "h<1>"]), "x": {"name": "inner"}}
        elif r < 0.7:
            child = {"name": name}
        elif r < 0.8:
            child = {"hostname": 5, "name": None}
        elif r < 0.9:
            child = rng.choice([[{"name": "in-array"}], "string", 7, None, {}])
        else:
            child = {"hostname": "only-host"}
        mod[key] = child
    return {"module": mod, "other": {"cluster_a_x": {"name": "not-in-module"}}}


@pytest.mark.parametrize("seed", range(25))
def test_index_lookups_on_device_resident_documents(tk, oracle, seed):
    """Clusters()/Nodes() answered from the GPU-built index (document never leaves the device): runs of
    children in ascending field order come straight out of the packed records, anything else (duplicate
    or descending names, children without the field) takes the sort path -- both must match the oracle."""
    rng = random.Random(5150 + seed)
    raw = oracle.New("x", json.dumps(_module_doc(rng, rng.choice([0, 1, 5, 60, 400]))).encode()).Bytes()   # canonical key order
    a, b = tk.New("x", raw), oracle.New("x", raw)
    assert not a.host_tree_built
    assert a.Clusters() == b.Clusters()
    for prov in ("a", "b", "aa", "zz"):
        for cl in ("x", "y", "xy", "x_y", ""):
            ck = "cluster_%s_%s" % (prov, cl)
            assert a.Nodes(ck) == b.Nodes(ck), ck
    assert not a.host_tree_built                        # none of it needed the host tree
    assert a.Bytes() == raw


@pytest.mark.parametrize("raw,err", [(b'{"module":[1]}', "not an object"), (b'{"module":"s"}', "not an object"), (b'{"x":1}', "not an object"),
                                     (b'[{"module":{}}]', "not an object"), (b'{"module":null}', "not an object")])
def test_index_module_not_an_object(tk, raw, err):
    s = tk.New("x", raw)
    for call in (s.Clusters, lambda: s.Nodes("cluster_a_b")):
        with pytest.raises(tk.StateError) as e:
            call()
        assert str(e.value) == err
    assert tk.New("x", b'{"module":{}}').Clusters() == {}


def test_document_larger_than_one_fold_round(tk, oracle):
    """1,400 clusters x 100 nodes (87 MB, 9.8 M tokens, 563 k brackets): the one-block folds of the
    parser (tile triples: 8,192 per round; token blocks: 4,096; radix run starts: 4,096 entries)
    all need more than one round, which config #4 does not reach."""
    b = oracle.New("big", synth.document(1400, 100))
    raw = b.Bytes()                                     # the persisted form: keys in order, so the document stays on the device
    a = tk.New("big", raw)
    text = a.Bytes()
    assert len(text) > 64 * 8192 * 128 and text == raw
    assert not a.host_tree_built
    ca = a.Clusters()
    assert ca == b.Clusters() and len(ca) == 1400
    for c in (0, 1, 777, 1399):
        assert a.Nodes(synth.cluster_key(c)) == b.Nodes(synth.cluster_key(c))
    assert tk.New("big", text).Bytes() == text


@pytest.mark.parametrize("bad_at", [1, 31, 32, 33, 63, 64, 100, 198])
def test_index_order_violation_anywhere_in_a_run(tk, oracle, bad_at):
    """200 nodes of one cluster whose hostnames ascend with their keys except at one place: the run
    is not in field order, which the index records in a bitmap (one bit per record) -- also for
    records past the first words of it -- and the lookup must then sort like the reference."""
    mod = {"cluster_a_c": {"name": "c"}}
    for i in range(200):
        host = "h%03d" % i
        if i == bad_at:
            host = "h%03d" % (i - 1) if i % 2 else "a-first"        # a duplicate of the previous name / one that sorts before all
        mod["node_a_c_%03d" % i] = {"hostname": host}
    mod["node_a_c_150x"] = {"name": "no hostname here"}              # a child without the field inside the run
    raw = oracle.New("x", json.dumps({"module": mod}).encode()).Bytes()
    a, b = tk.New("x", raw), oracle.New("x", raw)
    got = a.Nodes("cluster_a_c")
    assert got == b.Nodes("cluster_a_c") and len(got) in (199, 200)
    assert not a.host_tree_built
    assert a.Clusters() == b.Clusters() == {b"c": b"cluster_a_c"}


def test_wide_unsorted_object_is_rebuilt_by_several_threads(tk, oracle):
    """30,000 members in shuffled order with duplicates (the last one wins), one of them an
    out-of-range number nowhere / everywhere: the host tree of such a wide object is built by several
    threads (engine.cu, build_value) and flattened by several threads, and must come out exactly
    as the single-threaded oracle's."""
    rng = random.Random(31337)
    ids = list(range(30000)) + [rng.randrange(30000) for _ in range(500)]
    rng.shuffle(ids)
    body = ",".join('"k%05d":{"v":%d,"s":["a<%d>",null]}' % (i, pos, i) for pos, i in enumerate(ids))
    raw = ('{"z":1,"wide":{' + body + '},"a":[]}').encode()
    a, b = tk.New("x", raw), oracle.New("x", raw)
    assert a.host_tree_built                      # keys out of order: the tree is built at once
    assert a.Bytes() == b.Bytes()
    assert a.Get("wide.k00007.v") == b.Get("wide.k00007.v")
    # the first out-of-range number in text order is the one reported, whichever thread met it
    bad = raw.replace(b'"k%05d":{"v":' % ids[20000], b'"k%05d":{"v":1e999,"w":' % ids[20000], 1).replace(
        b'"k%05d":{"v":' % ids[100], b'"k%05d":{"v":2e999,"w":' % ids[100], 1)
    with pytest.raises(oracle.OracleError) as want:
        oracle.New("x", bad)
    with pytest.raises(tk.StateError) as got:
        tk.New("x", bad)
    assert str(got.value) == str(want.value) and "2e999" in str(got.value)


def test_input_already_in_output_form_and_near_misses(tk, oracle):
    """Text that is already json.MarshalIndent output is copied line by line after every line has
    been checked against the tape (emit_kernel, emit_verbatim); text that only nearly is -- spaces
    for tabs, a missing or extra blank, CRLF, blanks inside empty containers, a blank before a
    comma -- must come out in the one true form all the same."""
    doc = {"module": {"m%03d" % i: {"name": "n%d" % i, "list": [1, "two", {}, [], {"k": None}], "e": {}, "t": True, "s": "x y"}
                      for i in range(300)}, "z": [], "a": -0.5e-7}
    # -0.5e-7 is not in floatEncoder's form: that document goes through the host tree, the rest stays on the device
    doc.pop("a")
    pretty = oracle.New("x", json.dumps(doc).encode()).Bytes()
    a = tk.New("x", pretty)
    assert not a.host_tree_built and a.Bytes() == pretty
    rng = random.Random(99)
    lines = pretty.split(b"\n")
    mutations = [
        lambda l: l.replace(b"\t", b"    ", 1),                    # spaces for a tab
        lambda l: l.replace(b'": ', b'":', 1),                     # no blank after the colon
        lambda l: l.replace(b'": ', b'":  ', 1),                   # two blanks
        lambda l: l.replace(b'": ', b'" : ', 1),                   # a blank before the colon
        lambda l: l + b"\r",                                       # CRLF
        lambda l: l.replace(b"{}", b"{ }").replace(b"[]", b"[\t]"),
        lambda l: l[:-1] + b" ," if l.endswith(b",") else l + b" ",
        lambda l: b"\t" + l,                                       # one tab too many
        lambda l: l[1:] if l.startswith(b"\t") else l,             # one too few
    ]
    for trial in range(60):
        ls = list(lines)
        for _ in range(rng.choice([1, 1, 2, 5])):
            i = rng.randrange(len(ls))
            ls[i] = rng.choice(mutations)(ls[i])
        raw = b"\n".join(ls)
        s = tk.New("x", raw)
        assert s.Bytes() == pretty, trial
    # several lines on one line, and a line split in two
    assert tk.New("x", pretty.replace(b",\n\t\t\t", b", ", 7)).Bytes() == pretty
    assert tk.New("x", pretty.replace(b'": ', b'":\n ', 3)).Bytes() == pretty


# ---- several devices / several threads in one process (tk_set_device, include/tk_state.h)

def _new_bytes_on(tk, oracle, device, raw, out, slot):
    """What a Go program does per goroutine locked to a thread: pick the device, then use handles."""
    try:
        assert tk.state.lib().tk_set_device(device) == 0
        want = oracle.New("x", raw).Bytes()
        for _ in range(3):
            s = tk.New("x", raw)
            assert s.Bytes() == want
            assert dict(s.Clusters()) == dict(oracle.New("x", raw).Clusters())
            s.close()
        out[slot] = True
    except BaseException as e:   # reported by the test thread
        out[slot] = e


def test_set_device_from_two_threads_on_one_device(tk, oracle):
    """Two host threads select device 0 and run New + Clusters + Bytes at the same time: the per-device
    set-up of the kernels (opt-in shared memory of emit_kernel, the resident grid of the scans) is
    entered concurrently."""
    import threading
    raws = [oracle.New("x", synth.document(7, 5)).Bytes(), oracle.New("x", synth.document(3, 9)).Bytes()]
    out = [None, None]
    ts = [threading.Thread(target=_new_bytes_on, args=(tk, oracle, 0, raws[i], out, i)) for i in range(2)]
    for t in ts: t.start()
    for t in ts: t.join()
    assert out == [True, True], out


def test_handles_on_two_devices_in_one_process(tk, oracle):
    """Device 1 after device 0 in the same process: function attributes are per device, so the second
    device needs its own opt-in before emit_kernel can launch there (one process per GPU, the way
    torchrun runs bench.py, never sees this)."""
    import threading
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    raw = oracle.New("x", synth.document(9, 6)).Bytes()
    out = [None, None]
    _new_bytes_on(tk, oracle, 0, raw, out, 0)
    _new_bytes_on(tk, oracle, 1, raw, out, 1)
    assert out == [True, True], out
    ts = [threading.Thread(target=_new_bytes_on, args=(tk, oracle, d, raw, out, d)) for d in range(2)]
    for t in ts: t.start()
    for t in ts: t.join()
    assert tk.state.lib().tk_set_device(0) == 0
    assert out == [True, True], out


def test_index_lists_that_outgrow_their_first_buffer(tk, oracle):
    """The field lists of the module index are sized by a guess from the text length (one "name" /
    "hostname" member per 256 bytes); 6,000 children of a dozen bytes each are far denser than that,
    so the lists are counted on the first pass and filled on a second one."""
    mod = {"cluster_a_%04d" % i: {"name": "n%04d" % i} for i in range(3000)}
    mod.update({"node_a_0007_%04d" % i: {"hostname": "h%04d" % i} for i in range(3000)})
    raw = oracle.New("x", json.dumps({"module": mod}).encode()).Bytes()
    a, b = tk.New("x", raw), oracle.New("x", raw)
    ca = a.Clusters()
    assert ca == b.Clusters() and len(ca) == 3000
    na = a.Nodes("cluster_a_0007")
    assert na == b.Nodes("cluster_a_0007") and len(na) == 3000
    assert not a.host_tree_built


def test_index_gives_way_to_the_host_tree_for_a_child_with_thousands_of_members(tk, oracle):
    """The index pass of the general pipeline finds the child a "name" member belongs to by walking up through
    at most 4,096 rows; a child with more members than that before its name sends Clusters()/Nodes() to the
    host tree, with the same answers.  (The line pass knows every row's container and has no such limit.)"""
    wide = {"k%05d" % i: "v" for i in range(6000)}
    wide["name"] = "wide"                                # sorts after every k....
    mod = {"cluster_a_b": {"name": "b"}, "cluster_a_wide": wide, "node_a_b_1": {"hostname": "h1", "x": {"name": "too deep"}, "y": [{"hostname": "in array"}]}}
    raw = oracle.New("x", json.dumps({"module": mod, "name": "at the root", "m2": {"cluster_a_q": {"name": "q"}}}).encode()).Bytes()
    b = oracle.New("x", raw)
    try:
        for line_pass in (True, False):
            tk.state.set_line_pass(line_pass)
            a = tk.New("x", raw)
            assert a.Clusters() == b.Clusters() == {b"b": b"cluster_a_b", b"wide": b"cluster_a_wide"}
            assert a.Nodes("cluster_a_b") == b.Nodes("cluster_a_b") == {b"h1": b"node_a_b_1"}
            assert a.host_tree_built == (not line_pass)
    finally:
        tk.state.set_line_pass(True)
    small = oracle.New("x", json.dumps({"module": {"cluster_a_b": {"name": "b"}, "node_a_b_1": {"hostname": "h1", "x": {"name": "too deep"},
                                                   "y": [{"hostname": "in array"}]}}, "name": "at the root"}).encode()).Bytes()
    c = tk.New("x", small)
    assert c.Clusters() == {b"b": b"cluster_a_b"} and c.Nodes("cluster_a_b") == {b"h1": b"node_a_b_1"} and not c.host_tree_built


# ---- edits of module's children on a document that stays on the device ---------------------------

def test_config5_destroy_sweep_on_the_device(tk, oracle):
    """The destroy sweep (destroy/cluster.go:126-158) over a document loaded from its persisted form:
    Nodes / Delete work on the index and the child list, Bytes() rebuilds the tape on the device, and
    the host tree is never built."""
    raw = oracle.New("x", synth.document(30, 12)).Bytes()
    a, b = tk.New("x", raw), oracle.New("x", raw)
    assert not a.host_tree_built
    snaps = {id(a): [], id(b): []}
    for s in (a, b):
        for i, (name, ck) in enumerate(sorted(s.Clusters().items())):
            for host, nk in sorted(s.Nodes(ck).items()):
                s.Delete(b"module." + nk)
            assert s.Nodes(ck) == {}
            s.Delete(b"module." + ck)
            s.Delete(b"module." + ck)          # deleting an absent key is not an error
            if i in (0, 7, 28):
                snaps[id(s)].append((s.Bytes(), s.Clusters()))
    assert snaps[id(a)] == snaps[id(b)]
    want = json.dumps({"module": {"cluster-manager": synth.manager_record()},
                       "terraform": {"backend": {"local": {"path": "/root/.triton-kubernetes/bench/terraform.tfstate"}}}},
                      indent="\t", sort_keys=True).encode()
    assert a.Bytes() == b.Bytes() == want
    assert not a.host_tree_built
    a.Delete("module.cluster-manager"); b.Delete("module.cluster-manager")     # the last child: "module": {} -- left to the host tree
    assert a.Bytes() == b.Bytes() and a.host_tree_built


def test_device_edits_first_last_replace_and_records(tk, oracle):
    """SetManager / AddCluster / AddNode / Delete on children of "module", document on the device: a new
    child that sorts first, one that sorts last (the comma moves), a replaced one, a record (opaque
    to Clusters / Nodes / Get, members in the given order), add-then-delete, and what Get sees once
    the host tree is finally built."""
    raw = oracle.New("x", synth.document(3, 2)).Bytes()
    a, b = tk.New("x", raw), oracle.New("x", raw)
    steps = [
        lambda s, m: s._set("module.aaa-first", {"name": "first", "x": [1, {"y": None}], "e": {}}),
        lambda s, m: s._set("module.zzz-last", {"hostname": "h<last>&", "name": "zl"}),
        lambda s, m: s.Delete("module.zzz-last"),
        lambda s, m: s._set("module.zzz-last2", "a plain string child"),
        lambda s, m: s.AddCluster("baremetal", "c0001", {"name": "c0001-replaced", "k8s_version": "v2"}),
        lambda s, m: s.AddNode("cluster_baremetal_c0001", "n-rec", m.Record([("source", "s"), ("hostname", "rec-host"), ("a", m.Record([("z", "1"), ("b", "2")]))])),
        lambda s, m: s.AddNode("cluster_baremetal_c0001", "n-map", {"hostname": "map-host", "source": "s"}),
        lambda s, m: s.SetManager(m.Record([("source", "new"), ("name", "mgr")])),
        lambda s, m: s.Delete("module." + synth.node_key(0, 1)),
        lambda s, m: s.Delete("module.not-there"),
        lambda s, m: s._set("module.tmp", {"name": "tmp"}),
        lambda s, m: s.Delete("module.tmp"),
        lambda s, m: s._set("module.zzzz-null", None),
    ]
    for i, step in enumerate(steps):
        step(a, tk); step(b, oracle)
        assert a.Clusters() == b.Clusters(), i
        for c in range(3):
            assert a.Nodes(synth.cluster_key(c)) == b.Nodes(synth.cluster_key(c)), i
        if i % 3 == 2 or i + 1 == len(steps):
            assert a.Bytes() == b.Bytes(), i
        assert not a.host_tree_built, i
    # the first look through Get builds the tree from the edited tape; records are still opaque in it
    for path in ("module.aaa-first.name", "module.node_baremetal_c0001_n-rec.hostname", "module.node_baremetal_c0001_n-map.hostname",
                 "module.cluster-manager.name", "module.zzz-last2", "module.cluster_baremetal_c0001.k8s_version"):
        assert a.Get(path) == b.Get(path), path
    assert a.host_tree_built and a.Bytes() == b.Bytes()
    a._set("module.after", {"name": "after"}); b._set("module.after", {"name": "after"})
    assert a.Bytes() == b.Bytes() and a.Clusters() == b.Clusters()
    # and the text written along the way loads like the oracle's (records become maps: keys sorted from then on)
    assert tk.New("y", a.Bytes()).Bytes() == oracle.New("y", b.Bytes()).Bytes()


@pytest.mark.parametrize("seed", range(12))
def test_random_module_child_edits_stay_on_the_device(tk, oracle, seed):
    """Random Set (maps and records) / Delete of module children interleaved with Clusters, Nodes and
    Bytes on a document that starts on the device: the results match the oracle and the host tree is
    not built until something other than a module child is touched."""
    rng = random.Random(4100 + seed)
    raw = oracle.New("x", json.dumps(_module_doc(rng, rng.choice([3, 20, 120]))).encode()).Bytes()
    a, b = tk.New("x", raw), oracle.New("x", raw)
    if len(json.loads(raw)["module"]) == 0:
        return
    keys = sorted(json.loads(raw)["module"])
    for step in range(150):
        r = rng.random()
        key = rng.choice(keys + ["cluster_a_x", "node_a_x_%d" % rng.randrange(8), "zz-new-%d" % rng.randrange(5), "0-first"])
        if r < 0.35:
            if len(json.loads(b.Bytes())["module"]) <= 1:
                continue                       # (the last child goes through the host tree: another test)
            a.Delete("module." + key); b.Delete("module." + key)
        elif r < 0.75:
            val = rng.choice([{"name": "n%d" % rng.randrange(6), "hostname": "h%d" % rng.randrange(9)}, {"hostname": "dup"}, {"name": "é<"}, "str", 5, None,
                              {"x": {"name": "inner"}}, [1, 2]])
            rec = rng.random() < 0.3 and isinstance(val, dict)
            va = tk.Record(list(val.items())) if rec else val
            vb = oracle.Record(list(val.items())) if rec else val
            a._set("module." + key, va); b._set("module." + key, vb)
            if key not in keys:
                keys.append(key)
        elif r < 0.9:
            assert a.Clusters() == b.Clusters()
            ck = rng.choice(["cluster_a_x", "cluster_b_y", "cluster_aa_x_y", "cluster_a_xy"])
            assert a.Nodes(ck) == b.Nodes(ck)
        else:
            assert a.Bytes() == b.Bytes()
        assert not a.host_tree_built, step
    assert a.Bytes() == b.Bytes() and a.Clusters() == b.Clusters()
    assert not a.host_tree_built
    a._set("other.deep", {"k": "v"}); b._set("other.deep", {"k": "v"})          # not a module child: the tree takes over
    assert a.host_tree_built and a.Bytes() == b.Bytes() and a.Clusters() == b.Clusters()


# ---- text in the indented form: the line-by-line pass (csrc/lines.cu) and what it must decline --------

def _outcome(mod, raw):
    try:
        s = mod.New("x", raw)
        return ("ok", s.Bytes(), s.Clusters() if b'"module"' in raw[:40] else None)
    except Exception as e:       # the Go error text
        return ("err", str(e))


def _pretty(oracle, doc):
    return oracle.New("x", json.dumps(doc).encode()).Bytes()


NEAR_MISSES = [
    (b'"name": "c0001"', b'"name":"c0001"'), (b'"name": "c0001"', b'"name":  "c0001"'), (b'"name": "c0001",', b'"name": "c0001" ,'),
    (b'"name": "c0001",', b'"name": "c0001"'), (b'"name": "c0001",', b'"name": "c0001",,'), (b'"name": "c0001",', b'"name": "c0001", '),
    (b'"name": "c0001",', b'"name": "c0001",\r'), (b'\t\t\t"name": "c0001"', b'\t\t"name": "c0001"'), (b'\t\t\t"name": "c0001"', b'\t\t\t\t"name": "c0001"'),
    (b'\t\t\t"name": "c0001"', b'\t\t \t"name": "c0001"'), (b'"name": "c0001"', b'"name": "c0\\u0030001"'), (b'"name": "c0001"', b'"name": "c<1"'),
    (b'"name": "c0001"', b'"name": "c\xc3\xa91"'), (b'"name": "c0001"', b'"name": "c\xff1"'), (b'"name": "c0001"', b'"name": "c\x011"'),
    (b'"name": "c0001"', b'"name": "c\t1"'), (b'"name": "c0001"', b'"name": 17'), (b'"name": "c0001"', b'"name": 1.50'), (b'"name": "c0001"', b'"name": 1e3'),
    (b'"name": "c0001"', b'"name": 01'), (b'"name": "c0001"', b'"name": -0'), (b'"name": "c0001"', b'"name": 1234567890123456'),
    (b'"name": "c0001"', b'"name": 123456789012345'), (b'"name": "c0001"', b'"name": tru'), (b'"name": "c0001"', b'"name": null'),
    (b'"name": "c0001"', b'"name": {}'), (b'"name": "c0001"', b'"name": []'), (b'"name": "c0001"', b'"name": [ ]'), (b'"name": "c0001"', b'"name": "c0001" "x"'),
    (b'"name": "c0001"', b'"name" "c0001"'), (b'"name": "c0001"', b'"name": "c0001'), (b'"name": "c0001"', b'name: "c0001"'), (b'"name": "c0001"', b'"c0001"'),
    (b'"name": "c0001"', b'"aame": "c0001"'), (b'"name": "c0001"', b'"zame": "c0001"'), (b'"name": "c0001"', b'"k8s_version": "c0001"'),
    (b'"cluster_baremetal_c0001": {', b'"cluster_baremetal_c0001": ['), (b'"cluster_baremetal_c0001": {', b'"cluster_baremetal_c0001": {}'),
    (b'"cluster_baremetal_c0001": {', b'"cluster_baremetal_c0001": {,'), (b'"cluster_baremetal_c0001": {', b'"cluster_baremetal_b": {'),
    (b'"cluster_baremetal_c0001": {', b'"cluster_baremetal_c0000": {'), (b'"cluster_baremetal_c0001": {', b'"a": {'), (b'"cluster_baremetal_c0001": {', b'{'),
    (b'"worker": "true"\n', b'"worker": "true",\n'), (b'"worker": "true"\n', b'"worker": "true"\n\n'), (b'"worker": "true"\n', b'"worker": "true"\n\t\t\t\t"zz": 1\n'),
]


def test_near_misses_of_the_indented_form_are_declined_or_agree(tk, oracle):
    """A document in the indented form with one small thing wrong -- a space, a comma, a tab, a key out of
    order, a number that would be re-formatted, a byte that would be escaped, the wrong bracket -- at the
    start, in the middle and at the end of a text that spans several tiles of the line pass: whatever the
    product makes of it (a document, or Go's error text) is what the oracle makes of it."""
    base = _pretty(oracle, json.loads(synth.document(25, 8)))
    assert len(base) > 3 * 8192 and _outcome(tk, base) == _outcome(oracle, base)
    assert not tk.New("x", base).host_tree_built
    for old, new in NEAR_MISSES:
        for which in (0, -1):
            at = base.find(old) if which == 0 else base.rfind(old)
            if at < 0:
                continue
            raw = base[:at] + new + base[at + len(old):]
            assert _outcome(tk, raw) == _outcome(oracle, raw), (old, new, which)
    for raw in (base + b"\n", base + b" ", b"\n" + base, base[:-1], base[:-2] + b"]", base + b"}", base.replace(b"\n", b"\r\n"), base.replace(b"\t", b"  "),
                base[:1] + b"\n" + base[1:], base.replace(b"{\n\t\"module\"", b"[\n\t\"module\"", 1)):
        assert _outcome(tk, raw) == _outcome(oracle, raw), raw[:30]


@pytest.mark.parametrize("seed", range(10))
def test_random_damage_to_the_indented_form(tk, oracle, seed):
    """Random single-byte and single-line damage anywhere in an indented document (several tiles long)."""
    rng = random.Random(7700 + seed)
    base = _pretty(oracle, rand_tree(rng, 5) if seed % 2 else json.loads(synth.document(12, 6)))
    if seed % 2:
        base = _pretty(oracle, {"module": {("k%04d" % i): rand_tree(rng, 3) for i in range(300)}, "z": [1, "two", None, {}]})
    lines = base.split(b"\n")
    for _ in range(60):
        r = rng.random()
        if r < 0.4:
            i = rng.randrange(len(base))
            raw = base[:i] + bytes([rng.choice(b' \t\n",:{}[]\\0a<\xc3\x00')]) + base[i + (rng.random() < 0.7):]
        elif r < 0.7:
            i = rng.randrange(len(lines))
            raw = b"\n".join(lines[:i] + lines[i + 1:])
        elif r < 0.85:
            i, j = rng.randrange(len(lines)), rng.randrange(len(lines))
            l2 = list(lines); l2[i], l2[j] = l2[j], l2[i]
            raw = b"\n".join(l2)
        else:
            i = rng.randrange(len(lines))
            raw = b"\n".join(lines[:i] + [lines[i]] + lines[i:])
        assert _outcome(tk, raw) == _outcome(oracle, raw), raw[:60]


def test_indented_form_long_lines_deep_nesting_and_dense_tiles(tk, oracle):
    """Shapes at the limits of the line pass: strings longer than its fast path (120 bytes), than its window
    margins (512) and than a tile (8 KiB), nesting at and past the depth it tracks, a tile with thousands of
    tiny lines, empty containers everywhere -- all must come out as from the oracle, on the device when the
    general parse would have left them there."""
    docs = [
        {"module": {"a": {"name": "x" * n, "hostname": "y" * (n + 3)} for n in (100, 119, 120, 121, 127, 128, 129, 300, 511, 512, 513, 700)}},
        {"module": {"k%02d" % i: {"name": "n" * (8100 + i * 17)} for i in range(6)}, "tail": "t" * 20000},
        {"a": [list(range(3000))], "b": {"c": [[], {}, [[]], [{}], {"d": {}}]}},
    ]
    for depth in (6, 7, 8, 9, 10, 12):
        d = "leaf"
        for i in range(depth):
            d = {"k%d" % i: d, "l": [d]} if i % 2 else [d, {"m": i}]
        docs.append({"module": {"deep": d, "zz": {"name": "after"}}})
    for doc in docs:
        raw = _pretty(oracle, doc)
        assert _outcome(tk, raw) == _outcome(oracle, raw)
        s = tk.New("x", raw)
        assert not s.host_tree_built and s.Bytes() == raw


def _full_outcome(tk, raw):
    try:
        s = tk.New("x", raw)
        pairs = s.Clusters() if b'"module"' in raw[:40] else None
        nodes = [s.Nodes(k) for _, k in sorted(pairs.items())[:3]] if pairs else None
        return ("ok", s.host_tree_built, s.Bytes(), pairs, nodes)
    except Exception as e:
        return ("err", str(e))


def test_line_pass_and_general_pipeline_give_the_same_result(tk, oracle):
    """state.New has two parsers for text in the indented form (lines.cu first, parse.cu when that one
    declines).  Every shape here goes through the library twice -- tk_set_line_pass(1), then (0) -- and
    the verdict, the text, the pairs and where the document lives must not differ (and agree with the oracle)."""
    rng = random.Random(4242)
    base = _pretty(oracle, json.loads(synth.document(25, 8)))
    texts = [base, _pretty(oracle, json.loads(synth.document(3, 2))), _pretty(oracle, {}), _pretty(oracle, []), _pretty(oracle, {"a": 1}),
             b'{"module":{"b":{"name":"x"}}}', _pretty(oracle, {"module": {"a": {"name": "n" * 300, "hostname": "h" * 9000}}, "t": "t" * 20000})]
    texts += [_pretty(oracle, rand_tree(rng, 5)) for _ in range(10)]
    for old, new in NEAR_MISSES:
        at = base.rfind(old)
        if at >= 0:
            texts.append(base[:at] + new + base[at + len(old):])
    try:
        for raw in texts:
            tk.state.set_line_pass(True)
            with_lines = _full_outcome(tk, raw)
            tk.state.set_line_pass(False)
            without = _full_outcome(tk, raw)
            assert with_lines == without, raw[:60]
            assert with_lines[0] == _outcome(oracle, raw)[0]
            if with_lines[0] == "ok":
                assert with_lines[2] == oracle.New("x", raw).Bytes()
    finally:
        tk.state.set_line_pass(True)


def test_text_already_on_the_device(tk, oracle):
    """tk_state_new_device: the text is in the caller's device buffer.  The line pass copies it into the handle's
    arena while it counts the lines (16-byte aligned source), anything else goes through a device-to-device
    copy first; lengths around the 64-byte chunks and 8 KiB tiles of that pass, aligned and not, indented and
    compact, valid and not -- the outcome is that of tk_state_new on the same bytes."""
    import torch
    dev = torch.device("cuda", torch.cuda.current_device())
    base = _pretty(oracle, json.loads(synth.document(30, 9)))
    texts = [_pretty(oracle, {}), _pretty(oracle, {"a": "b"}), b'{"module":{"cluster_a_b":{"name":"b"}}}', base, base[:-1], base + b"\n", b"{", b"[1, 2"]
    for n in (63, 64, 65, 127, 128, 8191, 8192, 8193, 16384):
        filler = {"k": "v" * 4}
        while len(_pretty(oracle, filler)) < n:
            filler["k"] += "v"
        t = _pretty(oracle, filler)
        texts.append(t)                                   # (ends within a byte or two of n)
        texts.append(base[:n])                            # cut off in mid-document: invalid
    for raw in texts:
        for shift in (0, 16, 1):
            buf = torch.zeros(len(raw) + 64, dtype=torch.uint8, device=dev)
            buf[shift:shift + len(raw)] = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
            torch.cuda.synchronize()
            want = _outcome(tk, raw)
            try:
                s = tk.state.NewDevice("x", buf.data_ptr() + shift, len(raw))
                got = ("ok", s.Bytes(), s.Clusters() if b'"module"' in raw[:40] else None)
            except Exception as e:
                got = ("err", str(e))
            assert got == want == _outcome(oracle, raw), (raw[:40], len(raw), shift)


def test_get_of_a_module_childs_name_or_hostname_needs_no_host_tree(tk, oracle):
    """Get("module.<key>.name") / (".hostname") on a document that lives on the device is answered from the module
    index -- before and after edits of module's children --; every other path builds the host tree, as before.
    Whatever the index does not hold (no such child, no such member, not a string, a record) is "" for Get too."""
    mod = {"cluster_a_b": {"name": "b", "x": "1"}, "cluster_a_c": {"name": 7}, "cluster_a_d": ["name"], "cluster_a_e": {"hostname": {"name": "deep"}},
           "node_a_b_1": {"hostname": "h1"}, "node_a_b_10": {"hostname": "h10", "name": "n10"}, "node_a_b_2": {}}
    raw = oracle.New("x", json.dumps({"module": mod, "name": "root"}).encode()).Bytes()
    a, b = tk.New("x", raw), oracle.New("x", raw)
    paths = ["module.%s.%s" % (k, f) for k in list(mod) + ["absent", "node_a_b_", ""] for f in ("name", "hostname")]
    for p in paths:
        assert a.Get(p) == b.Get(p), p
    assert not a.host_tree_built
    for s in (a, b):
        s.Delete("module.node_a_b_1")
        s.AddNode("cluster_a_b", "zz", {"hostname": "hz"})
        s.AddCluster("a", "q", tk.Record([("name", "opaque")]) if s is a else oracle.Record([("name", "opaque")]))
    for p in paths + ["module.node_a_b_zz.hostname", "module.cluster_a_q.name"]:
        assert a.Get(p) == b.Get(p), p
    assert not a.host_tree_built and a.Bytes() == b.Bytes()
    assert a.Get("module.cluster_a_b.x") == b.Get("module.cluster_a_b.x") == b"1" and a.host_tree_built
    assert a.Get("name") == b.Get("name") == b"root"


def test_invalid_text_after_a_larger_document_in_recycled_buffers(tk, oracle):
    """Buffers are recycled between handles, so the kernels of the general pipeline that run after the one that finds
    a text invalid see, where that text left rows unwritten, whatever a larger document before it left there.
    (tools/stress_lines.py, seed 1147: key_order_kernel followed such a row's stale bracket number out of bounds.)"""
    big = _pretty(oracle, json.loads(synth.document(60, 30)))
    small = _pretty(oracle, {"module": {("cluster_a_%02d" % i): {"name": "n%d" % i, "deep": {"a": {"b": [1, {"c": "x" * 700}]}}} for i in range(40)},
                             "z": {"k": ["v", {"w": None}]}})
    damaged = []
    for old, new in ((b'",\n', b'"-\n'), (b'": {', b'": {-'), (b'}\n', b'}}\n'), (b'[\n', b'[,\n'), (b'": "', b'" "')):
        for which in (small.find(old), small.find(old, len(small) // 2), small.rfind(old)):
            if which >= 0:
                damaged.append(small[:which] + new + small[which + len(old):])
    try:
        for line_pass in (True, False):
            tk.state.set_line_pass(line_pass)
            for raw in damaged:
                s = tk.New("x", big)
                s.Bytes(); s.Clusters()
                s.close()
                assert _outcome(tk, raw) == _outcome(oracle, raw), raw[:40]
    finally:
        tk.state.set_line_pass(True)


def test_indented_form_with_text_beyond_ascii_stays_on_the_device(tk, oracle):
    """Strings with bytes from 0x80 up, as json.Marshal writes them (valid UTF-8 as it stands), are kept in place by
    the line pass; what Marshal would have written differently -- invalid sequences, U+2028 / U+2029, overlong
    forms, surrogates -- and everything around such bytes goes the way the oracle says, whichever parser sees it."""
    doc = {"module": {"cluster_a_é": {"name": "é√ß", "k8s": "日本語 \U0001F600"}, "node_a_é_1": {"hostname": "hôte-1", "ключ": "значение", "z": "x" * 130 + "é"},
                      "node_a_é_2": {"hostname": "\u07ff\u0800\uffff\U00010000\U0010ffff"}}, "é": ["√", {"ß": None}]}
    raw = _pretty(oracle, doc)
    a, b = tk.New("x", raw), oracle.New("x", raw)
    assert not a.host_tree_built and a.Bytes() == b.Bytes() == raw
    assert a.Clusters() == b.Clusters() and a.Nodes("cluster_a_é") == b.Nodes("cluster_a_é") and len(b.Nodes("cluster_a_é")) == 2
    assert a.Get("module.node_a_é_1.hostname") == b.Get("module.node_a_é_1.hostname") == "hôte-1".encode()
    assert not a.host_tree_built
    e = "é".encode()
    bad = [b"\xc3", b"\xa9", b"\xc0\xaf", b"\xe0\x80\xaf", b"\xed\xa0\x80", b"\xf4\x90\x80\x80", b"\xf8\x88\x80\x80\x80", b"\xff", b"\xe2\x80\xa8", b"\xe2\x80\xa9",
           b"\xe2\x80\xaa", b"\xc3\xa9\xc3", "\u2027".encode(), b"\xf0\x9f\x98", b"\xc2\x80", b"\xdf\xbf", b"\xef\xbf\xbf", b"\xf4\x8f\xbf\xbf"]
    try:
        for line_pass in (True, False):
            tk.state.set_line_pass(line_pass)
            for seq in bad:
                for at in (raw.find(e), raw.rfind(e)):
                    t = raw[:at] + seq + raw[at + len(e):]
                    assert _outcome(tk, t) == _outcome(oracle, t), (seq, at, line_pass)
    finally:
        tk.state.set_line_pass(True)


def test_indented_form_with_escaped_values_stays_on_the_device(tk, oracle):
    """String values that read differently once decoded -- escape sequences, < > & (written \\u003c ...), control
    characters, invalid UTF-8 left as it is -- are decoded into the side arena by the line pass, as the general
    parser does; keys of that kind, and more such text than the side arena has room for, send the document to the
    general parser.  Either way the results are the oracle's."""
    vals = ['a"b', "back\\slash", "tab\there", "line\nfeed\r", "<tag attr=\"x\">&amp;</tag>", "\x01\x1f\x7f", "é<√>", "  ", "/", "", "\\u0041 literally",
            "x" * 130 + "<", "q" * 9000 + "&" + "r" * 300, "\U0001F600&"]
    doc = {"module": {"cluster_a_b": {"name": 'n"<1>', "v": vals}, "node_a_b_1": {"hostname": "h\\1", "labels": {"k": "<&>"}}, "node_a_b_2": {"hostname": "plain"}},
           "z": vals[::-1]}
    raw = _pretty(oracle, doc)
    assert b"\\u003c" in raw and b'\\"' in raw and b"\\\\" in raw
    a, b = tk.New("x", raw), oracle.New("x", raw)
    assert not a.host_tree_built and a.Bytes() == b.Bytes() == raw
    assert a.Clusters() == b.Clusters() == {b'n"<1>': b"cluster_a_b"} and a.Nodes("cluster_a_b") == b.Nodes("cluster_a_b") and len(b.Nodes("cluster_a_b")) == 2
    assert a.Get("module.node_a_b_1.hostname") == b.Get("module.node_a_b_1.hostname") == b"h\\1" and not a.host_tree_built
    a.Delete("module.node_a_b_2"); b.Delete("module.node_a_b_2")
    a.AddNode("cluster_a_b", "3", {"hostname": "<3>"}); b.AddNode("cluster_a_b", "3", {"hostname": "<3>"})
    assert a.Bytes() == b.Bytes() and not a.host_tree_built
    # escapes written another way than Marshal writes them, invalid UTF-8 as it stands: valid JSON all the same
    for old, new in ((b"\\u003c", b"\\u003C"), (b"\\u003c", b"<"), (b"\\n", b"\\u000a"), (b"/", b"\\/"), (b"\\u0001", b"\\u0001\\ud83d\\ude00\\ud83d"), ("é".encode(), b"\xc3"),
                     ("é".encode(), b"\\u00e9"), (b"plain", b"pl\\u0061in"), (b'"hostname": "h', b'"host\\u006eame": "h'), (b'"z": [', b'"\\u007a": [')):
        assert old in raw, old
        t = raw.replace(old, new)
        assert _outcome(tk, t) == _outcome(oracle, t), (old, new)
    for old, new in ((b"\\u003c", b"\\u003"), (b"\\u003c", b"\\x3c"), (b"\\n", b"\n"), (b"\\t", b"\t"), (b'\\"', b'"'), (b"\\\\", b"\\")):
        t = raw.replace(old, new, 1)
        assert _outcome(tk, t) == _outcome(oracle, t), (old, new)
    # more decoded text than the side arena takes (an eighth of the text): still right
    many = _pretty(oracle, {"k%04d" % i: "<&>" * 40 for i in range(400)})
    s = tk.New("x", many)
    assert s.Bytes() == many == oracle.New("x", many).Bytes()
