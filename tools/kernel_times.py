"""Per-kernel device times of the config #4 step, spread over steps: python tools/kernel_times.py [steps]

Every step is a fresh handle with per-kernel event timing on (New from text resident on the device
+ Clusters + Nodes + Bytes left on the device); prints min / median / max per kernel in us.  A
development aid next to bench.py (which reports the averages)."""
import importlib
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

tk = importlib.import_module("triton-kubernetes_b200")
synth = importlib.import_module("triton-kubernetes_b200.synth")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 15
if os.environ.get("TK_TEXT_FROM_ORACLE"):      # (for experiments with kernels that deliberately compute the wrong thing)
    from oracle import oracle as O
    O.build()
    text = O.New("bench", synth.document(1000, 100)).Bytes()
else:
    seed = tk.New("bench", synth.document(1000, 100))
    text = seed.Bytes()
    seed.close()
dev_text = torch.frombuffer(bytearray(text), dtype=torch.uint8).to("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
torch.cuda.synchronize()
tk.state.set_default_timing(True)
rows = {}
for i in range(steps + 3):
    flush.fill_(i & 255)   # evict the text and the tables from L2, as bench.py does between steps
    torch.cuda.synchronize()
    s = tk.state.NewDevice("bench", dev_text.data_ptr(), len(text))
    if not os.environ.get("TK_SKIP_INDEX"):
        s.ClustersCount(); s.NodesCount(synth.cluster_key(0))
    s.BytesDevice()
    st = s.KernelStats()
    s.close()
    if i >= 3:
        for k, v in st.items():
            if v["timed"]:
                rows.setdefault(k, []).append(1e3 * v["ms"])
tk.state.set_default_timing(False)
print("%-32s %8s %8s %8s   (us per step, %d steps, %d-byte document)" % ("kernel", "min", "median", "max", steps, len(text)))
for k, v in rows.items():
    print("%-32s %8.1f %8.1f %8.1f" % (k, min(v), statistics.median(v), max(v)))
print("%-32s %8.1f" % ("sum of medians", sum(statistics.median(v) for v in rows.values())))
