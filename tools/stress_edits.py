"""Differential stress of point operations on documents that live on the device, against the oracle (not part of
the suite: run it on a GPU box for as many seeds as time allows).  Per seed: a random module-shaped document in
the indented form, then a few hundred random operations -- Delete / Set / AddNode / AddCluster / SetManager of
module's children (maps and records, plain and not), Get of names and of other paths, Clusters, Nodes, Bytes,
now and then an operation that needs the host tree, now and then the text written and loaded again -- each
compared with what the oracle returns or raises.

    python tools/stress_edits.py [seeds] [first_seed]"""
import importlib
import json
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
tk = importlib.import_module("triton-kubernetes_b200")
from oracle import oracle as O

seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 100
first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
NAMES = ["a", "b", "c1", "dev", "prod-1", "x_y", "é", "n<1>", ""]


def rand_child(rng, kind):
    if kind == "cluster":
        v = {"name": rng.choice(NAMES), "k8s_version": "v1.%d" % rng.randrange(20), "source": "s" * rng.choice([1, 30, 200])}
    else:
        v = {"hostname": rng.choice(NAMES) + "-%d" % rng.randrange(40), "labels": {"worker": rng.choice(["true", "false"])}, "count": rng.randrange(5)}
    r = rng.random()
    if r < 0.08:
        v["name"] = rng.choice([7, None, {"name": "deep"}, ["x"]])
    if r > 0.92:
        v["hostname"] = rng.choice(NAMES)
    if 0.08 <= r < 0.12:
        return rng.choice(["str", 5, None, [1, 2], {}])
    return v


def rand_document(rng):
    mod = {}
    if rng.random() < 0.8:
        mod["cluster-manager"] = {"name": "m", "source": "src"}
    for _ in range(rng.choice([0, 1, 2, 5, 40])):
        prov, cn = rng.choice(["a", "aws", "b"]), rng.choice(["x", "y", "x_y", "c%d" % rng.randrange(50)])
        mod["cluster_%s_%s" % (prov, cn)] = rand_child(rng, "cluster")
        for i in range(rng.choice([0, 1, 3, 30])):
            mod["node_%s_%s_%s" % (prov, cn, rng.choice(["w", "e"]) + str(i))] = rand_child(rng, "node")
    doc = {"module": mod}
    if rng.random() < 0.7:
        doc["terraform"] = {"backend": {"local": {"path": "/p"}}}
    return doc


def both(a, b, what, fa, fb):
    ra = rb = None
    try:
        ra = ("ok", fa())
    except Exception as e:
        ra = ("err", str(e))
    try:
        rb = ("ok", fb())
    except Exception as e:
        rb = ("err", str(e))
    if ra != rb:
        raise AssertionError("%s: product %r, oracle %r" % (what, str(ra)[:300], str(rb)[:300]))
    return ra


ops = bad = on_device = 0
for seed in range(first, first + seeds):
    rng = random.Random(7300000 + seed)
    raw = O.New("x", json.dumps(rand_document(rng), ensure_ascii=False).encode()).Bytes()
    log = []
    try:
        a, b = tk.New("x", raw), O.New("x", raw)
        for step in range(rng.choice([20, 100, 400])):
            keys = sorted(json.loads(b.Bytes()).get("module", {}).keys()) if step % 25 == 0 else keys
            key = rng.choice((keys or ["k"]) + ["cluster_a_x", "node_a_x_w1", "zz-%d" % rng.randrange(4), "0-first", "node_a_x_"])
            r = rng.random()
            if r < 0.22:
                log.append(("Delete", key)); both(a, b, log[-1], lambda: a.Delete("module." + key), lambda: b.Delete("module." + key))
            elif r < 0.40:
                val = rand_child(rng, rng.choice(["cluster", "node"]))
                rec = isinstance(val, dict) and rng.random() < 0.3
                log.append(("Set", key, val, rec))
                both(a, b, log[-1], lambda: a._set("module." + key, tk.Record(list(val.items())) if rec else val),
                     lambda: b._set("module." + key, O.Record(list(val.items())) if rec else val))
            elif r < 0.50:
                ck = rng.choice([k for k in keys if k.startswith("cluster_")] + ["cluster_a_x", "bad"])
                nm, val = "w%d" % rng.randrange(60), rand_child(rng, "node")
                log.append(("AddNode", ck, nm, val)); both(a, b, log[-1], lambda: a.AddNode(ck, nm, val), lambda: b.AddNode(ck, nm, val))
            elif r < 0.55:
                pr, nm, val = rng.choice(["a", "aws"]), rng.choice(["x", "n%d" % rng.randrange(9)]), rand_child(rng, "cluster")
                log.append(("AddCluster", pr, nm, val)); both(a, b, log[-1], lambda: a.AddCluster(pr, nm, val), lambda: b.AddCluster(pr, nm, val))
            elif r < 0.58:
                val = {"name": rng.choice(NAMES), "source": "s"}
                log.append(("SetManager", val)); both(a, b, log[-1], lambda: a.SetManager(val), lambda: b.SetManager(val))
            elif r < 0.72:
                p = "module.%s.%s" % (key, rng.choice(["name", "hostname"]))
                log.append(("Get", p)); both(a, b, log[-1], lambda: a.Get(p), lambda: b.Get(p))
            elif r < 0.82:
                log.append(("Clusters",)); cl = both(a, b, log[-1], lambda: sorted(a.Clusters().items()), lambda: sorted(b.Clusters().items()))
                if cl[0] == "ok" and cl[1]:
                    ck = rng.choice(cl[1])[1]
                    log.append(("Nodes", ck)); both(a, b, log[-1], lambda: sorted(a.Nodes(ck).items()), lambda: sorted(b.Nodes(ck).items()))
            elif r < 0.88:
                ck = rng.choice(["cluster_a_x", "cluster_b_y", "cluster_aws_x_y", "cluster_a_xy", "nope", "cluster_a"])
                log.append(("Nodes", ck)); both(a, b, log[-1], lambda: sorted(a.Nodes(ck).items()), lambda: sorted(b.Nodes(ck).items()))
            elif r < 0.96:
                log.append(("Bytes",)); both(a, b, log[-1], a.Bytes, b.Bytes)
            elif r < 0.98:
                log.append(("reload",)); t = b.Bytes(); both(a, b, log[-1], a.Bytes, b.Bytes); a, b = tk.New("x", t), O.New("x", t)
            elif r < 0.99:
                p = rng.choice(["module.%s.source" % key, "terraform.backend.local.path", "module", "nope.x"])
                log.append(("Get", p)); both(a, b, log[-1], lambda: a.Get(p), lambda: b.Get(p))
            else:
                log.append(("Backend",)); both(a, b, log[-1], lambda: a.SetTerraformBackendConfig("terraform.backend.local", {"path": "/q"}),
                                               lambda: b.SetTerraformBackendConfig("terraform.backend.local", {"path": "/q"}))
            ops += 1
        both(a, b, "final Bytes", a.Bytes, b.Bytes)
        on_device += not a.host_tree_built
    except AssertionError as e:
        bad += 1
        print("MISMATCH seed %d after %d operations: %s\n  last operations: %r" % (seed, len(log), e, log[-6:]))
        os.makedirs("gpurun_out", exist_ok=True)
        open("gpurun_out/stress_edits_fail_%d.json" % seed, "wb").write(raw)
        if bad >= 5:
            break
print("%d operations over %d seeds, %d mismatching seeds; %d documents were still on the device at the end" % (ops, seeds, bad, on_device))
sys.exit(1 if bad else 0)
