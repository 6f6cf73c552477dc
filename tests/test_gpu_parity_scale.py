"""The throughput path (FAST f32 arithmetic, red.add row updates, thousands of Hogwild warps) against
the single-thread reference trajectory AT THE DIMENSIONS AND SIZES THE CONFIGS USE — north_star bars,
un-relaxed: per-epoch loss within 1 % for EVERY epoch including the first, top-10 nearest-neighbour
overlap >= 0.9 on a 10 000-row sample.

Reference loop: J/opt/grad/Adagrad.java:49-95 with J/opt/GloveCost.java:7-20 / PGloveCost.java:7-20;
the three shipped configs run `method: pglove`, `dim: 300` (dblp.config.yml:2-3), BASELINE configs use
d=100 and d=128.  Kernel instantiations covered: NT=1 TP=1 (d=128, the benchmarked one), NT=1 TP=0 at
d=100 (25 of 32 lanes live), NT=3 (d=300).

The single-thread oracle needs 5-25 minutes per case, so its trajectories are committed fixtures
(tests/golden/parity_*.npz, written by tests/golden/make_parity_curves.py from oracle/oracle.cpp); the
test rebuilds the same graph and co-occurrence matrix on the GPU and first proves, by checksum, that
it feeds the trainer the very matrix the oracle was fed.

Loss reference = `hist_exact`: the oracle trajectory's per-epoch cost summed in fp64.  What the
reference itself REPORTS (`hist`) runs the sum in a float (Adagrad.java:60) and loses increments
beyond ~10 M nonzeros per thread (C1: 14.6 % low); the GPU sums in fp64.  At test sizes where the
float sum is still accurate both are checked.
"""
import os
import sys

import numpy as np
import pytest

from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, ImportedMatrix, Opt, createOptimizer

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLD)
import make_parity_curves as mp  # noqa: E402  (case table + graph factories + checksums; imports no oracle at module level)

CASES = [n for n in mp.CASES if os.path.exists(os.path.join(GOLD, f"parity_{n}.npz"))]


def _topk_gpu(E, sample, k=10):
    import torch
    E = torch.as_tensor(np.asarray(E, np.float32), device="cuda")
    En = E / E.norm(dim=1, keepdim=True).clamp_min(1e-12)
    s = torch.as_tensor(sample.astype(np.int64), device="cuda")
    out = []
    for a in range(0, len(s), 2000):
        rows = s[a:a + 2000]
        S = En[rows] @ En.T
        S[torch.arange(len(rows), device="cuda"), rows] = -float("inf")
        out.append(S.topk(k, dim=1).indices.to(torch.int32).cpu())
    return torch.cat(out).numpy()


def _overlap(a, b):
    return float(np.mean([len(set(x) & set(y)) / a.shape[1] for x, y in zip(a, b)]))


def test_fixtures_present():
    assert len(CASES) >= 6, "run tests/golden/make_parity_curves.py"


@pytest.mark.parametrize("name", CASES)
def test_throughput_path_tracks_single_thread_reference(name):
    F = np.load(os.path.join(GOLD, f"parity_{name}.npz"))
    kind, args, method, dim, epochs = mp.CASES[name]
    g = mp.make_graph(kind, args)
    base = dict(method=method, dim=dim, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, epochs), seed=int(F["seed"]))
    c = Configuration(**base)                                   # FAST math, red.add updates, device epoch order
    bca = BookmarkColoring(g, c)
    I, J, X = bca.coo()
    assert np.array_equal(mp.coo_checksums(I, J, X), F["coo"]) and bca.max() == float(F["xmax"])   # same matrix as the oracle's
    ref, sample = F["hist_exact"], F["sample"]
    out = {}
    for tag, kw in (("device", {}), ("java", dict(order="java"))):
        cc = Configuration(**base, **kw)
        m = bca if tag == "device" else ImportedMatrix(g.V, I, J, X, cc, xmax=bca.max())
        opt = createOptimizer(cc, m)
        hist = np.array([opt.epoch() for _ in range(epochs)])
        rel = np.abs(hist - ref) / ref
        ov = _overlap(_topk_gpu(opt.extractResultF32(), sample), F["nn"])
        out[tag] = (rel, ov)
        print(f"[{name}/{tag}] max rel loss diff {rel.max():.5f} per epoch {np.round(rel, 5)}  top-10 overlap {ov:.4f} "
              f"(oracle vs oracle under another epoch order: {float(F['self_ov']):.4f}, "
              f"loss {np.round(np.abs(F['hist_other'] - ref) / ref, 5).max():.5f})")
        del opt
    self_ov = float(F["self_ov"])
    for tag, (rel, ov) in out.items():
        assert rel.max() < 0.01, (tag, rel)                      # every epoch, the first included
    # order="java" replays the reference's own epoch order: the north_star bar as written
    assert out["java"][1] >= 0.9, out["java"]
    # device order = another uniformly random order.  Where the REFERENCE ITSELF, single-threaded, only
    # reaches self_ov < 0.9 against its own run under another shuffle (glove d=128 after 8 epochs is
    # still far from converged: 0.79 on the R-MAT case), no implementation that changes the visiting
    # order can do better: then the bar is the reference's own self-overlap (minus 0.01 of sampling noise).
    assert out["device"][1] >= min(0.9, self_ov - 0.01), (out["device"], self_ov)
