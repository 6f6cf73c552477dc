"""ge_flatten_edges_static (csrc/graph.cu): the allocation-free variant of the adjacency flattening —
sort buffers in a file-scope device array, caller-owned device inputs and outputs — against the
existing entry point ge_set_graph_edges / ge_graph_export on the same edge lists, bit for bit (the
bar of tests/test_gpu_graph.py), on consecutive calls that share the scratch.  Collected last on
purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from graph_embeddings_b200 import BCA, Configuration, Opt, _lib, synth
from graph_embeddings_b200.api import _Native, _ptr

pytestmark = pytest.mark.gpu

MAX_E = 1 << 20  # GE_STATIC_MAX_EDGES (include/graphemb.h)


def cfg():
    return Configuration(method="glove", dim=8, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0, 1))


def edge_list(V, E, seed, dup_frac=0.3):
    rng = np.random.default_rng(seed)
    n0 = E - int(E * dup_frac)
    s = rng.integers(0, V, n0).astype(np.int32)
    d = rng.integers(0, V, n0).astype(np.int32)
    pick = rng.integers(0, n0, E - n0)            # parallel edges with other weights
    s = np.concatenate([s, s[pick]]); d = np.concatenate([d, d[pick]])
    w = rng.uniform(0.05, 1.0, E).astype(np.float32)
    p = rng.permutation(E)
    return s[p], d[p], w[p]


def existing_entry_point(n, V, s, d, w):
    """ge_set_graph_edges + ge_graph_export: (off, nbr, w) of the out- and the in-lists"""
    L = n.L
    n.ck(L.ge_set_graph_edges(n.h, V, len(s), _ptr(s), _ptr(d), _ptr(w)))
    eo, ei, sec = C.c_int64(), C.c_int64(), C.c_double()
    n.ck(L.ge_graph_sizes(n.h, C.byref(eo), C.byref(ei), C.byref(sec)))
    oo, io = np.zeros(V + 1, np.int64), np.zeros(V + 1, np.int64)
    on, inn = np.zeros(eo.value, np.int32), np.zeros(ei.value, np.int32)
    ow, iw = np.zeros(eo.value, np.float32), np.zeros(ei.value, np.float32)
    n.ck(L.ge_graph_export(n.h, _ptr(oo), _ptr(on), _ptr(ow), _ptr(io), _ptr(inn), _ptr(iw)))
    return (oo, on, ow), (io, inn, iw)


def static_entry_point(n, V, s, d, w, by_dst, poison):
    """one direction through ge_flatten_edges_static; device buffers come from torch (the caller's
    allocator), outputs pre-filled with `poison` so stale or unwritten entries show"""
    import torch
    E = len(s)
    dev = torch.device("cuda", 0)
    ds, dd, dw = (torch.from_numpy(a).to(dev) for a in (s, d, w))
    off = torch.full((V + 1,), poison, dtype=torch.int64, device=dev)
    nbr = torch.full((max(E, 1),), poison, dtype=torch.int32, device=dev)
    wout = torch.full((max(E, 1),), float(poison), dtype=torch.float32, device=dev)
    info = torch.full((2,), poison, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()                      # the handle queues on its own non-blocking stream
    n.ck(n.L.ge_flatten_edges_static(n.h, V, E, by_dst, ds.data_ptr(), dd.data_ptr(), dw.data_ptr(),
                                     off.data_ptr(), nbr.data_ptr(), wout.data_ptr(), info.data_ptr()))
    torch.cuda.synchronize()
    m, bad = (int(x) for x in info.cpu())
    return off.cpu().numpy(), nbr.cpu().numpy()[:m], wout.cpu().numpy()[:m], m, bad


def same(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


def test_static_flatten_matches_existing_entry_point_on_consecutive_calls():
    n = _Native(cfg())
    try:
        # warm-up form (E = 0): zero offsets, no sort buffer touched
        z = np.zeros(0, np.int32)
        off, nbr, wout, m, bad = static_entry_point(n, 7, z, z, np.zeros(0, np.float32), 0, -3)
        assert m == 0 and bad == 0 and not off.any()
        # a large call, then smaller ones on the same scratch: nothing of the earlier call may leak,
        # then the large one again (the brief's "two calls in a row", both directions each time)
        cases = [(70000, 300000, 11), (300, 2000, 12), (3, 100000, 13), (1, 3, 14), (50000, MAX_E, 15), (70000, 300000, 11)]
        for k, (V, E, seed) in enumerate(cases):
            s, d, w = edge_list(V, E, seed)
            ref_out, ref_in = existing_entry_point(n, V, s, d, w)
            for by_dst, ref in ((0, ref_out), (1, ref_in)):
                off, nbr, wout, m, bad = static_entry_point(n, V, s, d, w, by_dst, -7 - k)
                assert bad == 0 and m == len(ref[1]), (V, E, by_dst, m, len(ref[1]))
                assert same(off, ref[0]) and same(nbr, ref[1]) and same(wout, ref[2]), (V, E, by_dst)
    finally:
        n.close()


def test_static_flatten_rejects_sizes_beyond_the_scratch_and_flags_bad_ids():
    import torch
    n = _Native(cfg())
    try:
        one = torch.zeros(4, dtype=torch.int64, device="cuda")
        p = one.data_ptr()                         # never dereferenced: the size check comes first
        rc = n.L.ge_flatten_edges_static(n.h, 10, MAX_E + 1, 0, p, p, p, p, p, p, p)
        assert rc == -1
        msg = n.L.ge_last_error(n.h).decode()
        assert "GE_STATIC_MAX_EDGES" in msg and str(MAX_E) in msg
        s = np.array([1, 2, 10], np.int32); d = np.array([0, 1, 2], np.int32); w = np.ones(3, np.float32)
        *_, bad = static_entry_point(n, 10, s, d, w, 0, -1)
        assert bad == 1                            # reported in d_info[1]; the call itself never synchronises
        # and the flag does not stick: the next call on the same scratch is clean
        s2, d2, w2 = edge_list(10, 50, 3)
        ref_out, _ = existing_entry_point(n, 10, s2, d2, w2)
        off, nbr, wout, m, bad = static_entry_point(n, 10, s2, d2, w2, 0, -1)
        assert bad == 0 and same(off, ref_out[0]) and same(nbr, ref_out[1]) and same(wout, ref_out[2])
    finally:
        n.close()
