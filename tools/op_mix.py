#!/usr/bin/env python
"""tools/op_mix.py -- the per-operation table of SURVEY.md 8(d), config #4 and config #5.

Every operation of the state.State method set is timed on its own, single handle, single host
thread, through the C ABI (compiled loops in tools/bench_driver.c and tools/oracle_driver.c), for the product (libtkstate on
the GPU in this box) and for the CPU oracle (oracle/tk_oracle.c), on the config #4 document
(1,000 clusters x 100 nodes, 62,177,405 bytes).  The document is built once outside the timed
regions; any loop that changes it gets a fresh handle.  Prints one JSON object; nothing here reads
/root/reference.

    python tools/op_mix.py [--clusters 1000] [--nodes 100] [--reps 5] [--skip-oracle]
"""
import argparse
import ctypes
import importlib
import json
import os
import sys
import time
from ctypes import POINTER, byref, c_char_p, c_double, c_int, c_size_t, c_void_p

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (for the driver build)


def arrays(strs):
    return (c_char_p * len(strs))(*strs), (c_size_t * len(strs))(*[len(s) for s in strs])


def best_mean(ts):
    return {"best_s": min(ts), "mean_s": sum(ts) / len(ts), "reps": len(ts)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clusters", type=int, default=1000)
    ap.add_argument("--nodes", type=int, default=100)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--gets", type=int, default=1000000)
    ap.add_argument("--skip-oracle", action="store_true")
    args = ap.parse_args()

    tk = importlib.import_module("triton-kubernetes_b200")
    synth = importlib.import_module("triton-kubernetes_b200.synth")
    from oracle import oracle as O
    O.build()
    OL = O.lib()
    D = bench.driver()
    DO = bench.oracle_driver()
    L = tk.state.lib()
    dbl = c_double
    D.tkb_get_loop.restype = D.tkb_add_node_loop.restype = D.tkb_delete_loop.restype = D.tkb_nodes_loop.restype = dbl
    D.tkb_destroy_sweep.restype = dbl
    DO.tkb_ora_get_loop.restype = DO.tkb_ora_add_node_loop.restype = DO.tkb_ora_delete_loop.restype = DO.tkb_ora_nodes_loop.restype = dbl
    DO.tkb_ora_destroy_sweep.restype = dbl
    SA, ZA = POINTER(c_char_p), POINTER(c_size_t)
    D.tkb_get_loop.argtypes = [c_void_p, SA, ZA, c_size_t, c_size_t, POINTER(c_size_t)]
    D.tkb_add_node_loop.argtypes = [c_void_p, SA, ZA, SA, ZA, SA, ZA, c_size_t, c_int]
    D.tkb_delete_loop.argtypes = [c_void_p, SA, ZA, c_size_t]
    D.tkb_nodes_loop.argtypes = [c_void_p, SA, ZA, c_size_t, POINTER(c_size_t)]
    D.tkb_destroy_sweep.argtypes = [c_void_p, SA, ZA, SA, ZA, c_size_t, POINTER(c_size_t)]
    DO.tkb_ora_get_loop.argtypes = [c_void_p, c_void_p, SA, ZA, c_size_t, c_size_t, POINTER(c_size_t)]
    DO.tkb_ora_add_node_loop.argtypes = [c_void_p, c_void_p, SA, ZA, SA, ZA, SA, ZA, c_size_t, c_int]
    DO.tkb_ora_delete_loop.argtypes = [c_void_p, c_void_p, SA, ZA, c_size_t]
    DO.tkb_ora_nodes_loop.argtypes = [c_void_p, c_void_p, c_void_p, SA, ZA, c_size_t, POINTER(c_size_t)]
    DO.tkb_ora_destroy_sweep.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, SA, ZA, SA, ZA, c_size_t, POINTER(c_size_t)]

    def fn(name):
        return ctypes.cast(getattr(OL, name), c_void_p)

    C, N = args.clusters, args.nodes
    text = O.New("bench", synth.document(C, N)).Bytes()                       # the persisted (tab-indented) form
    base = O.New("bench", synth.document(C, N, with_nodes=False)).Bytes()     # manager + clusters only
    cks = [synth.cluster_key(c).encode() for c in range(C)]
    node_keys = sorted(synth.node_key(c, i).encode() for c in range(C) for i in range(1, N + 1))
    get_paths = [b"module." + k + b".hostname" for k in node_keys]
    del_paths = [b"module." + k for k in node_keys]
    cluster_paths = [b"module." + k for k in cks]
    add = [(synth.cluster_key(c).encode(), ("c%04d-w-%d" % (c, i)).encode(),
            json.dumps(synth.node_record(c, i), separators=(",", ":")).encode()) for c in range(C) for i in range(1, N + 1)]
    a_cks, a_ck_lens = arrays([a[0] for a in add])
    a_names, a_name_lens = arrays([a[1] for a in add])
    a_jsons, a_json_lens = arrays([a[2] for a in add])
    g_arr, g_lens = arrays(get_paths)
    d_arr, d_lens = arrays(del_paths)
    ck_arr, ck_lens = arrays(cks)
    cp_arr, cp_lens = arrays(cluster_paths)
    MB = len(text) / 1e6
    out = {"document_bytes": len(text), "module_keys": len(node_keys) + C + 1, "clusters": C, "nodes_per_cluster": N,
           "host_threads": 1, "gets": args.gets}

    # ------------------------------------------------------------------ product (GPU + host tree)
    import torch
    pinned = torch.empty(len(text), dtype=torch.uint8, pin_memory=True)
    ctypes.memmove(pinned.data_ptr(), text, len(text))
    L.tk_state_new.argtypes = [c_char_p, c_size_t, c_void_p, c_size_t, POINTER(c_void_p)]

    def p_new(raw_ptr, n):
        h = c_void_p()
        rc = L.tk_state_new(b"bench", 5, raw_ptr, n, byref(h))
        if rc:
            raise SystemExit("op_mix: tk_state_new failed: %s" % L.tk_last_error(None))
        return h

    def p_bytes(h):
        o, n = c_void_p(), c_size_t()
        if L.tk_state_bytes(h, byref(o), byref(n)):
            raise SystemExit("op_mix: tk_state_bytes failed")
        return o.value, n.value

    prod = {}
    h = p_new(pinned.data_ptr(), len(text)); p_bytes(h)                      # warm-up: buffer pools, module load
    wb, wl, wc = c_void_p(), c_size_t(), c_size_t()
    L.tk_state_clusters(h, byref(wb), byref(wl), byref(wc)); L.tk_state_free(h)
    ts, tb = [], []
    for _ in range(args.reps):
        t0 = time.perf_counter(); h = p_new(pinned.data_ptr(), len(text)); ts.append(time.perf_counter() - t0)
        t0 = time.perf_counter(); _, n = p_bytes(h); tb.append(time.perf_counter() - t0)
        assert n == len(text)
        L.tk_state_free(h)
    prod["New"] = dict(best_mean(ts), docs_per_s=1 / min(ts), MB_per_s=MB / min(ts))
    prod["Bytes"] = dict(best_mean(tb), docs_per_s=1 / min(tb), MB_per_s=MB / min(tb))

    h = p_new(pinned.data_ptr(), len(text))
    blob, bl, cnt = c_void_p(), c_size_t(), c_size_t()
    t0 = time.perf_counter(); L.tk_state_clusters(h, byref(blob), byref(bl), byref(cnt)); first = time.perf_counter() - t0
    assert cnt.value == C
    t0 = time.perf_counter()
    for _ in range(100):
        L.tk_state_clusters(h, byref(blob), byref(bl), byref(cnt))
    prod["Clusters"] = {"first_call_s (builds the module index on the GPU)": first, "calls_per_s": 100 / (time.perf_counter() - t0), "calls": 100}
    pairs = c_size_t()
    dt = D.tkb_nodes_loop(h, ck_arr, ck_lens, C, byref(pairs))
    assert dt > 0 and pairs.value == C * N
    prod["Nodes"] = {"calls_per_s": C / dt, "calls": C, "pairs": pairs.value}
    vb = c_size_t()
    L.tk_state_host_tree_built.argtypes = [c_void_p]
    # module.<node>.hostname: answered from the module index, the document stays on the device
    dt = D.tkb_get_loop(h, g_arr, g_lens, len(get_paths), args.gets, byref(vb))
    assert dt > 0 and vb.value > 0 and not L.tk_state_host_tree_built(h)
    prod["Get (module.<key>.hostname, document on the device)"] = {"ops_per_s": args.gets / dt, "ops": args.gets}
    # module.<node>.source: any other path builds the host tree first
    src_paths = [b"module." + k + b".source" for k in node_keys]
    s_arr, s_lens = arrays(src_paths)
    t0 = time.perf_counter(); dt1 = D.tkb_get_loop(h, s_arr, s_lens, len(src_paths), 1, byref(vb)); first = time.perf_counter() - t0
    dt = D.tkb_get_loop(h, s_arr, s_lens, len(src_paths), args.gets, byref(vb))
    assert dt > 0 and dt1 > 0 and L.tk_state_host_tree_built(h)
    prod["Get"] = {"first_call_s (builds the host tree from the device tape)": first, "ops_per_s": args.gets / dt, "ops": args.gets}
    dt = D.tkb_delete_loop(h, d_arr, d_lens, len(del_paths))
    assert dt > 0
    prod["Delete"] = {"ops_per_s": len(del_paths) / dt, "ops": len(del_paths)}
    _, n = p_bytes(h)
    assert n == len(base), (n, len(base))
    L.tk_state_free(h)

    hb = ctypes.create_string_buffer(base, len(base))
    h = p_new(ctypes.addressof(hb), len(base))
    vb2 = c_size_t()
    D.tkb_get_loop(h, cp_arr, cp_lens, 1, 1, byref(vb2))   # (the host tree is built before the timed inserts)
    dt = D.tkb_add_node_loop(h, a_cks, a_ck_lens, a_names, a_name_lens, a_jsons, a_json_lens, len(add), tk.state.TK_VALUE_MAP)
    assert dt > 0
    prod["AddNode"] = {"ops_per_s": len(add) / dt, "ops": len(add)}
    t0 = time.perf_counter(); o, n = p_bytes(h); tflat = time.perf_counter() - t0
    prod["Bytes after edits"] = {"seconds (host tree flattened to a tape, uploaded, serialised on the GPU, copied back)": tflat,
                                 "MB_per_s": n / 1e6 / tflat}
    assert ctypes.string_at(o, n) == text
    L.tk_state_free(h)

    # the same edits on a document that stays on the device (no Get first: the host tree is never built)
    L.tk_state_host_tree_built.argtypes = [c_void_p]
    h = p_new(ctypes.addressof(hb), len(base))
    dt = D.tkb_add_node_loop(h, a_cks, a_ck_lens, a_names, a_name_lens, a_jsons, a_json_lens, len(add), tk.state.TK_VALUE_MAP)
    assert dt > 0 and not L.tk_state_host_tree_built(h)
    prod["AddNode (document on the device)"] = {"ops_per_s": len(add) / dt, "ops": len(add)}
    t0 = time.perf_counter(); o, n = p_bytes(h); tdev = time.perf_counter() - t0
    prod["Bytes after edits (document on the device)"] = {"seconds (tape rebuilt on the GPU from row segments, serialised, copied back)": tdev,
                                                          "MB_per_s": n / 1e6 / tdev}
    assert ctypes.string_at(o, n) == text and not L.tk_state_host_tree_built(h)
    L.tk_state_free(h)
    h = p_new(pinned.data_ptr(), len(text))
    dt = D.tkb_delete_loop(h, d_arr, d_lens, len(del_paths))
    assert dt > 0 and not L.tk_state_host_tree_built(h)
    t0 = time.perf_counter(); _, n = p_bytes(h); tdev = time.perf_counter() - t0
    assert n == len(base), (n, len(base))
    prod["Delete (document on the device)"] = {"ops_per_s": len(del_paths) / dt, "ops": len(del_paths), "Bytes_afterwards_s": tdev}
    L.tk_state_free(h)

    h = p_new(pinned.data_ptr(), len(text))
    deleted = c_size_t()
    dt = D.tkb_destroy_sweep(h, ck_arr, ck_lens, cp_arr, cp_lens, C, byref(deleted))
    assert dt > 0, dt
    t0 = time.perf_counter(); o, n = p_bytes(h); tbytes = time.perf_counter() - t0
    final = ctypes.string_at(o, n)
    built_after_sweep = L.tk_state_host_tree_built(h)
    L.tk_state_free(h)
    prod["config5_destroy_sweep"] = {"host_tree_built": bool(built_after_sweep), "seconds": dt, "deletes": deleted.value, "ops_per_s (Nodes + Delete calls)": (deleted.value + C) / dt,
                                     "final_Bytes_s": tbytes, "final_document_bytes": len(final)}
    out["product"] = prod

    # ------------------------------------------------------------------ oracle (CPU, one thread)
    if not args.skip_oracle:
        ora = {}
        err = ctypes.create_string_buffer(512)
        OL.ora_new.restype = c_void_p
        OL.ora_new.argtypes = [c_char_p, c_size_t, c_char_p, c_size_t]
        OL.ora_free.argtypes = [c_void_p]
        OL.ora_bytes.argtypes = [c_void_p, c_int, POINTER(c_void_p), POINTER(c_size_t)]
        OL.ora_free_buf.argtypes = [c_void_p]
        OL.ora_clusters.argtypes = [c_void_p, POINTER(c_void_p), POINTER(c_size_t), POINTER(c_size_t), c_char_p, c_size_t]
        ts, tb = [], []
        for _ in range(max(1, min(args.reps, 3))):
            t0 = time.perf_counter(); oh = OL.ora_new(text, len(text), err, 512); ts.append(time.perf_counter() - t0)
            o, n = c_void_p(), c_size_t()
            t0 = time.perf_counter(); OL.ora_bytes(oh, 1, byref(o), byref(n)); tb.append(time.perf_counter() - t0)
            assert n.value == len(text)
            OL.ora_free_buf(o); OL.ora_free(oh)
        ora["New"] = dict(best_mean(ts), docs_per_s=1 / min(ts), MB_per_s=MB / min(ts))
        ora["Bytes"] = dict(best_mean(tb), docs_per_s=1 / min(tb), MB_per_s=MB / min(tb))
        oh = OL.ora_new(text, len(text), err, 512)
        o, n, cnt = c_void_p(), c_size_t(), c_size_t()
        t0 = time.perf_counter()
        for _ in range(100):
            OL.ora_clusters(oh, byref(o), byref(n), byref(cnt), err, 512); OL.ora_free_buf(o)
        ora["Clusters"] = {"calls_per_s": 100 / (time.perf_counter() - t0), "calls": 100}
        dt = DO.tkb_ora_nodes_loop(fn("ora_nodes"), fn("ora_free_buf"), oh, ck_arr, ck_lens, C, byref(pairs))
        assert dt > 0 and pairs.value == C * N
        ora["Nodes"] = {"calls_per_s": C / dt, "calls": C, "pairs": pairs.value}
        dt = DO.tkb_ora_get_loop(fn("ora_get"), oh, g_arr, g_lens, len(get_paths), args.gets, byref(vb))
        assert dt > 0
        ora["Get"] = {"ops_per_s": args.gets / dt, "ops": args.gets}
        dt = DO.tkb_ora_delete_loop(fn("ora_delete"), oh, d_arr, d_lens, len(del_paths))
        assert dt > 0
        ora["Delete"] = {"ops_per_s": len(del_paths) / dt, "ops": len(del_paths)}
        OL.ora_free(oh)
        oh = OL.ora_new(base, len(base), err, 512)
        dt = DO.tkb_ora_add_node_loop(fn("ora_add_node"), oh, a_cks, a_ck_lens, a_names, a_name_lens, a_jsons, a_json_lens, len(add), 0)
        assert dt > 0
        ora["AddNode"] = {"ops_per_s": len(add) / dt, "ops": len(add)}
        o, n = c_void_p(), c_size_t()
        OL.ora_bytes(oh, 1, byref(o), byref(n))
        assert ctypes.string_at(o.value, n.value) == text
        OL.ora_free_buf(o); OL.ora_free(oh)
        oh = OL.ora_new(text, len(text), err, 512)
        dt = DO.tkb_ora_destroy_sweep(fn("ora_nodes"), fn("ora_free_buf"), fn("ora_delete"), oh, ck_arr, ck_lens, cp_arr, cp_lens, C, byref(deleted))
        assert dt > 0
        t0 = time.perf_counter(); OL.ora_bytes(oh, 1, byref(o), byref(n)); tbytes = time.perf_counter() - t0
        ofinal = ctypes.string_at(o.value, n.value)
        OL.ora_free_buf(o); OL.ora_free(oh)
        ora["config5_destroy_sweep"] = {"seconds": dt, "deletes": deleted.value, "ops_per_s (Nodes + Delete calls)": (deleted.value + C) / dt,
                                        "final_Bytes_s": tbytes, "final_document_bytes": len(ofinal)}
        out["oracle"] = ora
        out["config5_final_documents_equal"] = final == ofinal
        assert final == ofinal
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
