"""Latency of state.New alone on text resident in HBM (one document, nothing else in flight): median and
spread of the host's wall clock around tk_state_new_device over many repetitions, L2 flushed before each.
For changes that the per-kernel timings cannot show (launch gaps, overlap of one launch with the next)."""
import importlib
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

tk = importlib.import_module("triton-kubernetes_b200")
synth = importlib.import_module("triton-kubernetes_b200.synth")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
seed = tk.New("bench", synth.document(1000, 100))
text = seed.Bytes()
seed.close()
dev_text = torch.frombuffer(bytearray(text), dtype=torch.uint8).to("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
new_ms, bytes_ms = [], []
for i in range(reps + 10):
    flush.fill_(i & 255)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    s = tk.state.NewDevice("bench", dev_text.data_ptr(), len(text))
    t1 = time.perf_counter()
    s.BytesDevice()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    s.close()
    if i >= 10:
        new_ms.append((t1 - t0) * 1e3)
        bytes_ms.append((t2 - t1) * 1e3)
for name, v in (("New", new_ms), ("Bytes (left on the device)", bytes_ms)):
    v.sort()
    print("%-28s median %.4f ms   p10 %.4f   p90 %.4f   (%d repetitions, %d-byte document)" % (name, statistics.median(v), v[len(v) // 10], v[9 * len(v) // 10], len(v), len(text)))
