"""tk_row_offsets_noalloc: the line-offset pass with its scratch in a fixed array inside the library.

Some users of this library cannot let it allocate GPU memory -- their program gave all of it to
its own pool at start-up, or they sit on a latency-critical path where a cudaMalloc could stall
the whole device.  For them the entry point takes every buffer from the caller.  This test plays
such a caller: it builds the line tape itself, keeps it in its own device buffers, and checks
the offsets it gets back against the text of the same document -- as the CPU oracle serialises it
and as the ordinary entry point (tk_state_bytes) does: line r must start exactly where the
allocation-free pass says it does.  Two calls in a row, the second on a different and smaller
document and on a stream of the caller's own, show that what the first call left in the shared
scratch does not leak into the second; around that second call the free device memory is read
(cudaMemGetInfo) before and after, and must not have moved.

Lives in a file of its own, named so that pytest collects it last.
"""
import ctypes
import importlib
import json

import pytest

pytestmark = pytest.mark.gpu

ROW_STR, ROW_RAW, OPEN_OBJ, OPEN_ARR, CLOSE_OBJ, CLOSE_ARR, EMPTY_OBJ, EMPTY_ARR = range(8)
HAS_KEY, COMMA, LAST, DEPTH_SHIFT = 1 << 4, 1 << 5, 1 << 6, 16


def build_tape(doc):
    """The caller's own flattener: one row per output line, keys in bytewise order (csrc/tape.hpp)."""
    meta, koff, klen, voff, vlen, arena = [], [], [], [], [], bytearray()

    def put(b):
        at = len(arena)
        arena.extend(b)
        return at

    def row(kind, depth, key, val, comma, last):
        m = kind | (depth << DEPTH_SHIFT) | (COMMA if comma else 0) | (LAST if last else 0)
        ko = kl = vo = 0
        if key is not None:
            m |= HAS_KEY
            kb = key.encode("utf-8", "surrogateescape")
            ko, kl = put(kb), len(kb)
        if val:
            vo = put(val)
        meta.append(m); koff.append(ko); klen.append(kl); voff.append(vo); vlen.append(len(val) if val else 0)

    def walk(v, depth, key, comma, last):
        if isinstance(v, str):
            row(ROW_STR, depth, key, v.encode("utf-8", "surrogateescape"), comma, last)
        elif v is None or isinstance(v, (bool, int)):
            row(ROW_RAW, depth, key, json.dumps(v).encode(), comma, last)
        elif isinstance(v, list):
            if not v:
                return row(EMPTY_ARR, depth, key, b"", comma, last)
            row(OPEN_ARR, depth, key, b"", False, False)
            for i, x in enumerate(v):
                walk(x, depth + 1, None, i + 1 < len(v), False)
            row(CLOSE_ARR, depth, None, b"", comma, last)
        else:
            if not v:
                return row(EMPTY_OBJ, depth, key, b"", comma, last)
            row(OPEN_OBJ, depth, key, b"", False, False)
            items = sorted(v.items(), key=lambda kv: kv[0].encode("utf-8", "surrogateescape"))
            for i, (k, x) in enumerate(items):
                walk(x, depth + 1, k, i + 1 < len(items), False)
            row(CLOSE_OBJ, depth, None, b"", comma, last)

    walk(doc, 0, None, False, True)
    return meta, koff, klen, voff, vlen, bytes(arena)


def offsets_noalloc(tk, torch, doc, own_stream=False, watch_memory=False):
    meta, koff, klen, voff, vlen, arena = build_tape(doc)
    n = len(meta)
    dev = torch.device("cuda", torch.cuda.current_device())
    cols = [torch.tensor(c, dtype=torch.int64, device=dev).to(torch.int32) for c in (meta, koff, klen, voff, vlen)]
    a = torch.zeros(len(arena) + 32, dtype=torch.uint8, device=dev)     # 16 readable bytes past the end
    a[:len(arena)] = torch.frombuffer(bytearray(arena or b"\0"), dtype=torch.uint8)[:len(arena)].to(dev)
    off = torch.full((n + 1,), -1, dtype=torch.int32, device=dev)
    err = ctypes.create_string_buffer(256)
    side = torch.cuda.Stream() if own_stream else None                     # (created before the memory reading below)
    stream = side.cuda_stream if own_stream else torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()                                               # the caller's buffers are ready on every stream
    free_before = torch.cuda.mem_get_info()[0]
    rc = tk.state.lib().tk_row_offsets_noalloc(*[c.data_ptr() for c in cols], a.data_ptr(), len(arena), n,
                                               off.data_ptr(), stream, err, 256)
    assert rc == 0, err.value
    torch.cuda.synchronize()
    if watch_memory:
        assert torch.cuda.mem_get_info()[0] == free_before                 # nothing was allocated, by anyone, during the call
    return [x & 0xffffffff for x in off.tolist()]


def offsets_from_text(text):
    """Where each line of the reference serialisation starts, plus the total length."""
    starts, i = [0], text.find(b"\n")
    while i != -1:
        starts.append(i + 1)
        i = text.find(b"\n", i + 1)
    return starts + [len(text)]


def documents():
    synth = importlib.import_module("triton-kubernetes_b200.synth")
    big = json.loads(synth.document(60, 30))                      # ~2,000 module blocks, ~26,000 lines
    big["strings"] = {"esc<>&": "tab\there \"quoted\" back\\slash   é \U0001F600 \x01\x7f", "": "",
                      "nested": [[], {}, [1, True, None, {"k": ["v", -5]}]], "long": "x<" * 3000}
    small = {"module": {"cluster_a_b": {"name": "b", "n": [1, 2, 3]}}, "z": "\n" * 50, "e": {}}
    return big, small


def test_two_calls_in_a_row_match_the_oracle_and_the_ordinary_entry_point(tk, oracle):
    import torch
    for i, doc in enumerate(documents()):                         # big first, then small: stale scratch must not matter
        raw = json.dumps(doc, ensure_ascii=False).encode("utf-8", "surrogateescape")
        want = oracle.New("noalloc", raw).Bytes()
        assert tk.New("noalloc", raw).Bytes() == want
        # the second call runs on a stream of the caller's own, with the free device memory watched
        assert offsets_noalloc(tk, torch, doc, own_stream=i == 1, watch_memory=i == 1) == offsets_from_text(want)


def test_sizes_beyond_the_fixed_scratch_are_rejected(tk):
    import torch
    L = tk.state.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    dummy = torch.zeros(64, dtype=torch.int32, device=dev)
    err = ctypes.create_string_buffer(256)
    p = dummy.data_ptr()
    for n in (0, 2097152 + 1, 0xffffffff):
        rc = L.tk_row_offsets_noalloc(p, p, p, p, p, p, 16, n, p, None, err, 256)
        assert rc == tk.state.TK_ELIMIT and b"outside 1..2097152" in err.value
    assert L.tk_row_offsets_noalloc(None, p, p, p, p, p, 16, 8, p, None, err, 256) == tk.state.TK_EARG
