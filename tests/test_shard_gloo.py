"""The N > 1 path on CPU: two gloo ranks share one local-backend root, each takes its share of the
cluster managers (shard.assign) and works on them alone; the job time is the max over ranks.
The documents are written and read with the oracle here (no GPU); on the B200 the same split runs
with one GPU per rank (bench.py --gpus N)."""
import importlib
import json
import os
import sys

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, root, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shard = importlib.import_module("triton-kubernetes_b200.shard")
    tk = importlib.import_module("triton-kubernetes_b200")
    from oracle.backend_local import LocalBackend as OB
    names = tk.LocalBackend(root).States()                      # the registry listing is host logic of the product
    mine = shard.assign(names, world)[rank]
    be, result = OB(root), {}
    for name in mine:
        st = be.State(name)
        clusters = st.Clusters()
        result[name] = sum(len(st.Nodes(ck)) for ck in clusters.values())
    ms = shard.job_time_ms(10.0 * (rank + 1), dist)             # rank 1 is "slower"
    gathered = [None] * world
    dist.all_gather_object(gathered, result)
    if rank == 0:
        json.dump({"ms": ms, "per_rank": gathered, "value": shard.job_value(7, world, ms)}, open(os.path.join(out_dir, "out.json"), "w"))
    dist.destroy_process_group()


def test_two_ranks_split_the_managers(tmp_path):
    sys.path.insert(0, ROOT)
    from oracle import oracle as O
    from oracle.backend_local import LocalBackend as OB
    synth = importlib.import_module("triton-kubernetes_b200.synth")
    root = str(tmp_path / "root")
    be = OB(root)
    want = {}
    for i, name in enumerate(["delta", "alpha", "echo", "charlie", "bravo"]):
        st = O.New(name, synth.document(i + 1, 3))
        be.PersistState(st)
        want[name] = (i + 1) * 3
    mp.spawn(_worker, args=(2, root, 29531, str(tmp_path)), nprocs=2, join=True)
    out = json.load(open(tmp_path / "out.json"))
    assert out["ms"] == 20.0                                       # max over ranks
    assert out["value"] == pytest.approx(2 * 7 / 0.020)
    assert [sorted(r) for r in out["per_rank"]] == [["alpha", "charlie", "echo"], ["bravo", "delta"]]
    merged = {k: v for r in out["per_rank"] for k, v in r.items()}
    assert merged == want                                          # nothing lost, nothing done twice


def test_assign_is_a_partition():
    shard = importlib.import_module("triton-kubernetes_b200.shard")
    names = ["m%02d" % i for i in range(17)]
    for world in (1, 2, 3, 8, 32):
        parts = shard.assign(reversed(names), world)
        assert len(parts) == world and sorted(sum(parts, [])) == names
        assert max(map(len, parts)) - min(map(len, parts)) <= 1
