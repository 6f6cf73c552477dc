"""Committed fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py).

not gpu: the oracle reproduces its own committed outputs bit for bit (guards the checker against
         compiler / flag / platform drift) and java.util.Random matches the public JDK answers.
gpu    : the CUDA path, through the C ABI, reproduces the same bytes (BCA: bit-exact) or tracks
         them (serial JAVA-math trainer: last-bit differences of CUDA's fp64 log/pow vs libm only)."""
import os

import numpy as np
import pytest

import oracle
from graph_embeddings_b200 import synth

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BCA = np.load(os.path.join(G, "bca.npz"))
TRAIN = np.load(os.path.join(G, "train.npz"))
BCA_NAMES = sorted({k.split("/")[0] for k in BCA.files})
TRAIN_NAMES = sorted({k.split("/")[0] for k in TRAIN.files})


def _graph(name):
    return synth.CsrGraph(int(BCA[f"{name}/V"]), *(BCA[f"{name}/{k}"] for k in
                          ("out_off", "out_nbr", "out_w", "in_off", "in_nbr", "in_w")))


def test_java_random_known_answers():
    kat = np.load(os.path.join(G, "java_random.npz"))
    assert np.array_equal(oracle.java_ints(42, 2), kat["seed42_nextInt"])
    assert np.array_equal(oracle.java_ints(0, 2), kat["seed0_nextInt"])


@pytest.mark.parametrize("name", BCA_NAMES)
def test_oracle_reproduces_bca_fixture(name):
    alpha, eps, directed = BCA[f"{name}/params"]
    r = oracle.cooc_build(_graph(name), alpha, eps, bool(directed), str(BCA[f"{name}/normalize"]), threads=3)
    assert np.array_equal(r.I, BCA[f"{name}/I"]) and np.array_equal(r.J, BCA[f"{name}/J"])
    assert np.array_equal(r.X.view(np.uint32), BCA[f"{name}/Xbits"])
    assert r.max == float(BCA[f"{name}/max"]) and r.pops == int(BCA[f"{name}/pops"])


@pytest.mark.parametrize("name", TRAIN_NAMES)
def test_oracle_reproduces_train_fixture(name):
    optimizer, method, _ = name.split("_")
    V, nnz, dim, seed, epochs = (int(x) for x in TRAIN[f"{name}/meta"])
    I, J, X = TRAIN[f"{name}/I"], TRAIN[f"{name}/J"], TRAIN[f"{name}/X"]
    t = oracle.Trainer(V, I, J, X, float(X.max()), dim, method, seed, optimizer=optimizer)
    assert np.array_equal(t.store("focus")[0].view(np.uint32), TRAIN[f"{name}/init_focus_row0"])
    hist = np.array([t.epoch() for _ in range(epochs)])
    # libm log/pow may differ in the last bit across glibc versions: allow 1e-12 relative
    assert np.allclose(hist, TRAIN[f"{name}/history"], rtol=1e-12, atol=0)
    assert np.array_equal(t.get_perm(), TRAIN[f"{name}/perm_last"])
    assert np.allclose(t.extract().astype(np.float32), TRAIN[f"{name}/result"], rtol=1e-6, atol=1e-9)


# ------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", BCA_NAMES)
def test_gpu_bca_matches_fixture(name):
    from graph_embeddings_b200 import BCA as BcaCfg, BookmarkColoring, Configuration, Opt
    alpha, eps, directed = BCA[f"{name}/params"]
    norm = str(BCA[f"{name}/normalize"])
    c = Configuration(method="glove", dim=8, bca=BcaCfg(float(alpha), float(eps), bool(directed), norm),
                      opt=Opt("adagrad", 0, 1))
    b = BookmarkColoring(_graph(name), c)
    I, J, X = b.coo()
    assert b.coOccurrenceCount() == len(BCA[f"{name}/I"])
    assert np.array_equal(I, BCA[f"{name}/I"]) and np.array_equal(J, BCA[f"{name}/J"])
    assert np.array_equal(X.view(np.uint32), BCA[f"{name}/Xbits"])
    assert b.max() == float(BCA[f"{name}/max"]) and b.stats()["pops"] == int(BCA[f"{name}/pops"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", TRAIN_NAMES)
def test_gpu_serial_trainer_tracks_fixture(name):
    from graph_embeddings_b200 import BCA as BcaCfg, Configuration, ImportedMatrix, Opt, createOptimizer
    optimizer, method, _ = name.split("_")
    V, nnz, dim, seed, epochs = (int(x) for x in TRAIN[f"{name}/meta"])
    I, J, X = TRAIN[f"{name}/I"], TRAIN[f"{name}/J"], TRAIN[f"{name}/X"]
    c = Configuration(method=method, dim=dim, bca=BcaCfg(0.1, 1e-3, True), opt=Opt(optimizer, 0.0, epochs),
                      seed=seed, order="java", math="java", serial=True)
    opt = createOptimizer(c, ImportedMatrix(V, I, J, X, c, xmax=float(X.max())))
    assert np.array_equal(opt.store("focus")[0].view(np.uint32), TRAIN[f"{name}/init_focus_row0"])
    assert np.array_equal(np.stack([opt.store("fBias"), opt.store("cBias")]).view(np.uint32), TRAIN[f"{name}/init_bias"])
    hist = np.array([opt.epoch() for _ in range(epochs)])
    assert np.allclose(hist, TRAIN[f"{name}/history"], rtol=1e-6, atol=0), (hist, TRAIN[f"{name}/history"])
    assert np.array_equal(opt.epoch_order(epochs - 1), TRAIN[f"{name}/perm_last"])
    assert np.allclose(opt.extractResultF32(), TRAIN[f"{name}/result"], rtol=1e-5, atol=1e-7)
