"""Three device-resident steps of the config #4 pass (New + Clusters + Nodes x1000 + Bytes) with no
timing around them: the command profiles/ are captured from (ncu replays each kernel many times,
so numbers printed under ncu are never bench values)."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

tk = importlib.import_module("triton-kubernetes_b200")
synth = importlib.import_module("triton-kubernetes_b200.synth")
clusters = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
nodes = int(sys.argv[2]) if len(sys.argv) > 2 else 100
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
seed = tk.New("bench", synth.document(clusters, nodes))
text = seed.Bytes()
seed.close()
dev = torch.device("cuda", 0)
dev_text = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev)
torch.cuda.synchronize()
before = tk.state.kernel_launches()
for _ in range(steps):
    s = tk.state.NewDevice("bench", dev_text.data_ptr(), len(text))
    assert s.ClustersCount() == clusters
    assert s.NodesCount(synth.cluster_key(0)) == nodes
    _, n = s.BytesDevice()
    assert n == len(text)
    torch.cuda.synchronize()
    s.close()
print("ok: %d kernel launches per step" % ((tk.state.kernel_launches() - before) // steps))
