"""The drop-in boundary used from plain C (tests/c/abi_driver.c): the header compiles as C99, the
program links against libgraphemb.so, fails loudly without a device (CPU), and on a GPU reproduces
the oracle's co-occurrence matrix bit for bit and the single-thread reference trajectory."""
import os
import struct
import subprocess

import numpy as np
import pytest

from graph_embeddings_b200 import _lib, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def driver(tmp_path_factory):
    _lib.build()
    exe = str(tmp_path_factory.mktemp("abi") / "abi_driver")
    libdir = os.path.dirname(_lib.LIB_PATH)
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c", "abi_driver.c"), "-o", exe,
                           "-L", libdir, "-lgraphemb", "-Wl,-rpath," + libdir])
    return exe


def write_graph(path, g):
    with open(path, "wb") as f:
        f.write(struct.pack("<iqq", g.V, len(g.out_nbr), len(g.in_nbr)))
        for a, t in ((g.out_off, np.int64), (g.out_nbr, np.int32), (g.out_w, np.float32),
                     (g.in_off, np.int64), (g.in_nbr, np.int32), (g.in_w, np.float32)):
            f.write(np.ascontiguousarray(a, t).tobytes())


def test_c_caller_fails_loudly_without_a_device(driver, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present: covered by the gpu test")
    g = synth.rmat(64, 256, 1)
    write_graph(tmp_path / "g.bin", g)
    r = subprocess.run([driver, str(tmp_path / "g.bin"), str(tmp_path / "o.bin"), "8", "2", "5"], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr
    assert not (tmp_path / "o.bin").exists()


@pytest.mark.gpu
def test_c_caller_reproduces_the_reference_trajectory(driver, tmp_path):
    import oracle
    g = synth.rmat(300, 1500, 2)
    dim, epochs, seed = 12, 3, 77
    write_graph(tmp_path / "g.bin", g)
    r = subprocess.run([driver, str(tmp_path / "g.bin"), str(tmp_path / "o.bin"), str(dim), str(epochs), str(seed)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    raw = open(tmp_path / "o.bin", "rb").read()
    nnz, xmax = struct.unpack_from("<qd", raw, 0)
    at = 16
    I = np.frombuffer(raw, np.int32, nnz, at); at += 4 * nnz
    J = np.frombuffer(raw, np.int32, nnz, at); at += 4 * nnz
    X = np.frombuffer(raw, np.float32, nnz, at); at += 4 * nnz
    ran, final_cost = struct.unpack_from("<id", raw, at); at += 12
    hist = np.frombuffer(raw, np.float64, ran, at); at += 8 * ran
    emb = np.frombuffer(raw, np.float64, g.V * dim, at).reshape(g.V, dim)
    ref = oracle.cooc_build(g, 0.1, 1e-3, True, "none", threads=2)
    assert nnz == ref.nnz and xmax == ref.max
    a, b = np.lexsort((J, I)), np.lexsort((ref.J, ref.I))
    assert np.array_equal(I[a], ref.I[b]) and np.array_equal(J[a], ref.J[b])
    assert np.array_equal(X[a].view(np.uint32), ref.X[b].view(np.uint32))
    # the trainer is fed the library's own COO order, as the reference's trainer is fed BookmarkColoring's lists
    t = oracle.Trainer(g.V, I, J, X, xmax, dim, "glove", seed)
    want, _ = t.optimize(epochs, 0.0)
    assert ran == epochs and final_cost == 0.0                       # Optimizer.java:73: maxiter reached -> finalCost stays 0
    np.testing.assert_allclose(hist, want, rtol=1e-6)
    np.testing.assert_allclose(emb, t.extract().reshape(g.V, dim), rtol=1e-5, atol=1e-7)
