#!/usr/bin/env python
"""Regenerates tests/golden/parity_*.npz — single-thread reference trajectories at realistic sizes.

The north_star bars for the trainer are statistical: per-epoch GloVe loss within 1 % of the
reference's SINGLE-THREADED run and top-10 nearest-neighbour overlap >= 0.9 on the final
embeddings.  A single-thread epoch of the oracle (oracle/oracle.cpp, the C++ restatement of
J/opt/grad/Adagrad.java:49-95 with GloveCost / PGloveCost) at these sizes takes 40-170 s, i.e.
tens of minutes per case: far too long to run beside the GPU tests.  So the trajectories are
computed once here, on the CPU, and committed; the `-m gpu` tests rebuild the same graph and
matrix on the device (checksums below prove the inputs are identical), train with the
throughput path and compare.

Per case the file holds
  hist        per-epoch loss, 1 thread, the reference's own epoch order (java.util.Random
              Fisher-Yates continued from the init draws, seed SEED) — as the reference REPORTS
              it: Adagrad.java:60 keeps the running sum in a float, which stops absorbing
              increments after some 10-30 M nonzeros (C1, 32 M nonzeros: 14.6 % too low)
  hist_exact  the same trajectory's loss with the running sum kept in fp64 (oracle-only
              quantity, oracle.cpp Trainer::job): what the GPU's fp64 loss is compared with
  nn          top-10 cosine neighbours of `sample` rows in (W+C)/2 after the last epoch
  hist_other  the SAME single-thread arithmetic under another epoch order (numpy permutations)
  self_ov     top-10 overlap of those two single-thread runs: what ANY implementation that
              changes the visiting order can reach at this size
  coo_*       order-independent checksums of the co-occurrence matrix the run was fed

PARITY UNPINNED with respect to the Java original (no JDK here): these are
oracle outputs, like tests/golden/make_golden.py's.

    python tests/golden/make_parity_curves.py [case ...]      # rewrites the fixtures in place
"""
import os
import sys
import time
from multiprocessing import Pool

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

SEED = 31
SAMPLE = 10_000

# name: (graph factory name, args, method, dim, epochs)
CASES = {
    "rmat100k_glove_d128": ("rmat", (100_000, 800_000, 21), "glove", 128, 8),
    "rmat100k_pglove_d300": ("rmat", (100_000, 800_000, 21), "pglove", 300, 8),
    "rmat100k_pglove_d100": ("rmat", (100_000, 800_000, 21), "pglove", 100, 8),
    "dblp03_glove_d128": ("dblp_like", (0.3, 1), "glove", 128, 8),
    "dblp03_pglove_d300": ("dblp_like", (0.3, 1), "pglove", 300, 8),
    "dblp03_pglove_d100": ("dblp_like", (0.3, 1), "pglove", 100, 8),
    # BASELINE configs[0] at full size (V = 320 k): dblp.config.yml as shipped (pglove, d=300) and d=100
    "c1_pglove_d300": ("dblp_like", (1.0, 1), "pglove", 300, 6),
    "c1_pglove_d100": ("dblp_like", (1.0, 1), "pglove", 100, 6),
}


def make_graph(kind, args):
    from graph_embeddings_b200 import synth
    return getattr(synth, kind)(*args)


def coo_checksums(I, J, X):
    """order-independent: the GPU builder emits rows in another order than the oracle"""
    xb = X.view(np.uint32).astype(np.uint64)
    k = (I.astype(np.uint64) << np.uint64(32)) | J.astype(np.uint64)
    with np.errstate(over="ignore"):
        mixed = (k * np.uint64(0x9E3779B97F4A7C15)) ^ xb
        return np.array([len(I), int(xb.sum(dtype=np.uint64)), int(mixed.sum(dtype=np.uint64))], np.uint64)


def topk(E, sample, k=10, chunk=None):
    E = np.asarray(E, np.float32)
    chunk = chunk or max(64, 50_000_000 // len(E))
    En = E / np.maximum(np.linalg.norm(E, axis=1, keepdims=True), 1e-12)
    out = np.empty((len(sample), k), np.int32)
    for a in range(0, len(sample), chunk):
        rows = sample[a:a + chunk]
        S = En[rows] @ En.T
        S[np.arange(len(rows)), rows] = -np.inf
        part = np.argpartition(-S, k, axis=1)[:, :k]
        sc = np.take_along_axis(S, part, axis=1)
        out[a:a + chunk] = np.take_along_axis(part, np.argsort(-sc, axis=1), axis=1)
    return out


def overlap(a, b):
    return float(np.mean([len(set(x) & set(y)) / a.shape[1] for x, y in zip(a, b)]))


def run_one(job):
    """one single-thread trajectory: (case, which) with which in {'java', 'other'}"""
    name, which = job
    import oracle
    kind, args, method, dim, epochs = CASES[name]
    g = make_graph(kind, args)
    c = oracle.cooc_build(g, 0.1, 1e-3, True, "none", threads=2)
    t = oracle.Trainer(g.V, c.I, c.J, c.X, c.max, dim, method, SEED)
    t0 = time.time()
    if which == "java":
        hist, _ = t.optimize(epochs, 0.0)
    else:
        rng = np.random.default_rng(5)
        hist = []
        for _ in range(epochs):
            t.set_perm(rng.permutation(c.nnz).astype(np.int32))
            hist.append(t.epoch(False))
        hist = np.array(hist)
    hist_exact = t.history_exact()
    sample = np.sort(np.random.default_rng(0).choice(g.V, min(SAMPLE, g.V), replace=False)).astype(np.int32)
    nn = topk(t.extract(), sample)
    t.close()
    print(f"[{name}/{which}] V={g.V} nnz={c.nnz} {epochs} epochs in {time.time() - t0:.0f}s hist={np.round(hist, 6)} exact={np.round(hist_exact, 6)}", flush=True)
    return name, which, dict(hist=np.asarray(hist, np.float64), hist_exact=hist_exact, nn=nn, sample=sample,
                             coo=coo_checksums(c.I, c.J, c.X), xmax=np.float64(c.max), V=np.int64(g.V))


def main():
    names = sys.argv[1:] or list(CASES)
    jobs = [(n, w) for n in names for w in ("java", "other")]
    # longest first (d=300 full-size), 6 workers: the rest of the cores stay free
    jobs.sort(key=lambda j: -CASES[j[0]][3] * (10 if j[0].startswith("c1") else 3))
    res = {}
    with Pool(int(os.environ.get("PARITY_WORKERS", "6"))) as pool:
        for name, which, d in pool.imap_unordered(run_one, jobs):
            res.setdefault(name, {})[which] = d
            if len(res[name]) == 2:
                a, b = res[name]["java"], res[name]["other"]
                kind, args, method, dim, epochs = CASES[name]
                self_ov = overlap(a["nn"], b["nn"])
                rel = np.abs(b["hist_exact"] - a["hist_exact"]) / a["hist_exact"]
                print(f"== {name}: self_ov {self_ov:.4f}  rel loss (other order vs java order) {np.round(rel, 5)}", flush=True)
                np.savez_compressed(os.path.join(HERE, f"parity_{name}.npz"), hist=a["hist"], hist_exact=a["hist_exact"], nn=a["nn"],
                                    sample=a["sample"], hist_other=b["hist_exact"], self_ov=np.float64(self_ov), coo=a["coo"], xmax=a["xmax"],
                                    V=a["V"], dim=np.int64(dim), epochs=np.int64(epochs), seed=np.int64(SEED),
                                    method=np.array(method), graph=np.array(f"{kind}{args}"))


if __name__ == "__main__":
    main()
