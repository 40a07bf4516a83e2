"""Differential stress of state.New's two parsers against the oracle (not part of the test suite: run it on a GPU
box for as many seeds as time allows).  Random documents in the indented form -- plain ASCII mostly, so that the
line pass keeps them, with numbers, empty containers, long strings, deep nesting and module-shaped children mixed
in --, then random damage to them; for every text: outcome (document or Go's error text), serialised text,
Clusters / Nodes pairs and a Get must be the oracle's, with the line pass on and off.

    python tools/stress_lines.py [seeds] [first_seed] [threads]          (STRESS_TEXT_ON_DEVICE=1: through tk_state_new_device)"""
import importlib
import json
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
tk = importlib.import_module("triton-kubernetes_b200")
from oracle import oracle as O

seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 200
first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
PLAIN = "abcdefghijklmnopqrstuvwxyzABCXYZ0123456789_-.$/:~ '()=?#@!*+,;[]{}|^%`"
SPICE = "é√ <>&\t\n\"\\\x01"


UTF8 = "é√ß日本語\U0001F600\u07ff\u0800\uffff"


def rand_string(rng, spice):
    n = rng.choice([0, 1, 3, 7, 8, 9, 15, 16, 17, 30, 60, 100, 119, 120, 121, 130, 300, 600, 9000]) if rng.random() < 0.15 else rng.randrange(0, 14)
    alphabet = PLAIN + (SPICE if spice == 1 else UTF8 if spice == 2 else "")   # 2: text the line pass keeps in place
    return "".join(rng.choice(alphabet) for _ in range(n))


def rand_value(rng, depth, spice):
    r = rng.random()
    if depth > rng.choice([2, 3, 5, 8, 10]) or r < 0.5:
        odd = [1234567890123456, 1.5, 1e21, -0.0] if rng.random() < 0.03 else []   # (numbers that are re-formatted send the document to the host tree)
        return rng.choice([rand_string(rng, spice), rand_string(rng, spice), None, True, False, 0, -1, 7, 123456789012345] + odd)
    if r < 0.65:
        return [rand_value(rng, depth + 1, spice) for _ in range(rng.choice([0, 0, 1, 2, 3, 40]))]
    return {rand_string(rng, spice): rand_value(rng, depth + 1, spice) for _ in range(rng.choice([0, 0, 1, 2, 3, 5, 30]))}


def rand_document(rng):
    spice = rng.choice([0, 0, 0, 1, 1, 2, 2])
    mod = {}
    for c in range(rng.randrange(0, 6)):
        ck = "cluster_%s_%s" % (rng.choice(["a", "aws", "b"]), rand_string(rng, False)[:5] or "c")
        mod[ck] = {"name": rand_string(rng, spice), "x": rand_value(rng, 2, spice)} if rng.random() < 0.8 else rand_value(rng, 1, spice)
        for i in range(rng.randrange(0, 5)):
            child = {"hostname": rand_string(rng, spice), "labels": rand_value(rng, 2, spice)}
            if rng.random() < 0.2:
                child["name"] = rand_string(rng, spice)
            mod["node_%s_%d" % (ck[8:], i)] = child if rng.random() < 0.9 else rand_value(rng, 1, spice)
    for _ in range(rng.randrange(0, 4)):
        mod[rand_string(rng, spice)] = rand_value(rng, 1, spice)
    doc = {"module": mod} if rng.random() < 0.85 else {}
    for _ in range(rng.randrange(0, 4)):
        doc[rand_string(rng, spice)] = rand_value(rng, 1, spice)
    if rng.random() < 0.05:
        return rand_value(rng, 0, spice)
    return doc


ON_DEVICE = bool(os.environ.get("STRESS_TEXT_ON_DEVICE"))   # the product gets its text in device memory (tk_state_new_device)


def outcome(mod, raw):
    try:
        if ON_DEVICE and mod is tk and len(raw):
            import torch
            shift = len(raw) % 3 * 8                               # 0, 8 or 16: the line pass copies from 16-byte aligned text only
            buf = torch.empty(len(raw) + 32, dtype=torch.uint8, device="cuda")
            buf[shift:shift + len(raw)] = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to("cuda")
            torch.cuda.synchronize()
            s = mod.state.NewDevice("x", buf.data_ptr() + shift, len(raw))
        else:
            s = mod.New("x", raw)
    except Exception as e:
        return ("err", str(e))
    out = ["ok", s.Bytes()]
    try:
        cl = s.Clusters()
        out.append(sorted(cl.items()))
        for _, ck in sorted(cl.items())[:3]:
            try:
                out.append(sorted(s.Nodes(ck).items()))
            except Exception as e:
                out.append(str(e))
            out.append(s.Get(b"module." + ck + b".name"))
    except Exception as e:
        out.append(str(e))
    return tuple(out)


def damage(rng, base):
    lines = base.split(b"\n")
    r = rng.random()
    if r < 0.35:
        i = rng.randrange(len(base))
        return base[:i] + bytes([rng.choice(b' \t\n",:{}[]\\0a1-.eE<\xc3\x00\xe2\x80\xa8\xed\xa0\xf4\x90\xc0\xff')]) + base[i + (rng.random() < 0.7):]
    if r < 0.55:
        i = rng.randrange(len(lines))
        return b"\n".join(lines[:i] + lines[i + 1:])
    if r < 0.7:
        i, j = rng.randrange(len(lines)), rng.randrange(len(lines))
        lines[i], lines[j] = lines[j], lines[i]
        return b"\n".join(lines)
    if r < 0.8:
        i = rng.randrange(len(lines))
        return b"\n".join(lines[:i] + [lines[i]] + lines[i:])
    if r < 0.9:
        return base[:rng.randrange(len(base) + 1)]
    return base + rng.choice([b"\n", b" ", b"}", b",", b"\t"])


def one_seed(seed, toggle=True):
    """-> (outcomes checked, mismatches, undamaged document stayed on the device)"""
    rng = random.Random(9100000 + seed)
    base = O.New("x", json.dumps(rand_document(rng), ensure_ascii=False).encode("utf-8", "surrogatepass")).Bytes()
    texts = [base] + [damage(rng, base) for _ in range(6)]
    checked = bad = 0
    for raw in texts:
        want = outcome(O, raw)
        for line_pass in (True, False):
            if toggle:
                tk.state.set_line_pass(line_pass)
            got = outcome(tk, raw)
            checked += 1
            if got != want:
                bad += 1
                print("MISMATCH seed %d line_pass %s text %r...\n  got  %r\n  want %r" % (seed, line_pass, raw[:80], str(got)[:300], str(want)[:300]))
                os.makedirs("gpurun_out", exist_ok=True)
                open("gpurun_out/stress_fail_%d_%d.json" % (seed, bad), "wb").write(raw)
    if toggle:
        tk.state.set_line_pass(True)
    s = tk.New("x", base)
    return checked, bad, int(not s.host_tree_built)


threads = int(sys.argv[3]) if len(sys.argv) > 3 else 1
if threads <= 1:
    results = [one_seed(seed) for seed in range(first, first + seeds)]
else:
    # several handles at work at once, buffers and streams recycled between threads; the switch between the two
    # parsers is process-wide, so one more thread flips it while the others run
    import concurrent.futures
    import threading
    import time
    stop = threading.Event()

    def flipper():
        on = True
        while not stop.is_set():
            on = not on
            tk.state.set_line_pass(on)
            time.sleep(0.003)

    f = threading.Thread(target=flipper)
    f.start()
    with concurrent.futures.ThreadPoolExecutor(threads) as pool:
        results = list(pool.map(lambda seed: one_seed(seed, toggle=False), range(first, first + seeds)))
    stop.set()
    f.join()
    tk.state.set_line_pass(True)
checked, bad, kept = (sum(r[i] for r in results) for i in range(3))
print("%d outcomes checked over %d seeds on %d thread(s), %d mismatches; %d of the undamaged documents stayed on the device" % (checked, seeds, threads, bad, kept))
sys.exit(1 if bad else 0)
