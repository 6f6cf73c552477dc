#!/usr/bin/env python
"""Regenerates tests/golden/*.npz.

The reference (Java 8 + un-vendored grph/Jena jars) cannot be built or run in this image and
ships no golden vectors of its own, so these fixtures do NOT come from the reference: they are
outputs of the CPU oracle (oracle/oracle.cpp, the line-by-line C++ restatement of the Java
path) on seeded inputs, plus java.util.Random known answers that are public facts about the
JDK.  Their job: (1) pin the oracle against drift (compiler, flags, platform) — `-m "not gpu"`;
(2) give the GPU parity tests bytes that were not produced in the same process — `-m gpu`.
PARITY UNPINNED with respect to the Java original.

    python tests/golden/make_golden.py       # rewrites the fixtures in place
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle  # noqa: E402
from graph_embeddings_b200 import synth  # noqa: E402

HAND = {
    "isolated": ([], 1), "edge": ([(0, 1, 1.0)], None), "two_cycle": ([(0, 1, 1.0), (1, 0, 1.0)], None),
    "self_loop": ([(0, 0, 1.0)], None), "star": ([(0, 1, 1.0), (0, 2, 0.1)], None),
    "both_lists": ([(0, 1, 1.0), (1, 0, 0.5)], None), "dangling": ([(0, 1, 1.0), (2, 1, 0.1)], 5),
}

# (name, graph factory, alpha, eps, directed, normalize)
BCA_CASES = [(f"hand_{n}_{'d' if d else 'u'}", (lambda e=e, V=V: synth.tiny(e, V)), 0.1, 1e-3, d, "none")
             for n, (e, V) in sorted(HAND.items()) for d in (True, False)]
BCA_CASES += [
    ("rmat400_d", lambda: synth.rmat(400, 2400, 21), 0.1, 1e-3, True, "none"),
    ("rmat400_u", lambda: synth.rmat(400, 2400, 21), 0.1, 1e-3, False, "none"),
    ("rmat400_d_counts", lambda: synth.rmat(400, 2400, 21), 0.1, 1e-3, True, "counts"),
    ("rmat400_d_unity", lambda: synth.rmat(400, 2400, 21), 0.1, 1e-3, True, "unity"),
    ("rmat300_d_a02_e1e4", lambda: synth.rmat(300, 3000, 22), 0.2, 1e-4, True, "none"),
    ("dblp_like_d", lambda: synth.dblp_like(scale=0.002, seed=1), 0.1, 1e-3, True, "none"),
    ("onstage_like_d", lambda: synth.onstage_like(scale=0.003, seed=2), 0.1, 1e-3, True, "none"),
    ("saa_like_d", lambda: synth.saa_like(scale=0.002, seed=3), 0.1, 1e-3, True, "none"),
]

# (name, V, nnz, dim, method, optimizer, seed, epochs)
TRAIN_CASES = [(f"{o}_{m}_d{d}", 60, 900, d, m, o, 1000 + k, 3)
               for k, (o, m, d) in enumerate((o, m, d) for o in ("adagrad", "adam", "amsgrad")
                                             for m in ("glove", "pglove") for d in (6, 128))]

# java.util.Random known answers (JDK-specified LCG; public facts, not oracle output)
JAVA_RANDOM_KAT = {"seed42_nextInt": [-1170105035, 234785527], "seed0_nextInt": [-1155484576, -723955400]}


def graph_arrays(g):
    return dict(V=np.int64(g.V), out_off=g.out_off, out_nbr=g.out_nbr, out_w=g.out_w,
                in_off=g.in_off, in_nbr=g.in_nbr, in_w=g.in_w)


def main():
    bca = {}
    for name, mk, alpha, eps, directed, norm in BCA_CASES:
        g = mk()
        r = oracle.cooc_build(g, alpha, eps, directed, norm, threads=1)
        for k, v in graph_arrays(g).items():
            bca[f"{name}/{k}"] = v
        bca[f"{name}/params"] = np.array([alpha, eps, float(directed)], np.float64)
        bca[f"{name}/normalize"] = np.array(norm)
        bca[f"{name}/I"], bca[f"{name}/J"] = r.I, r.J
        bca[f"{name}/Xbits"] = r.X.view(np.uint32)          # bit patterns: exact compare
        bca[f"{name}/max"] = np.float64(r.max)
        bca[f"{name}/pops"] = np.int64(r.pops)
    np.savez_compressed(os.path.join(HERE, "bca.npz"), **bca)

    tr = {}
    for name, V, nnz, dim, method, optimizer, seed, epochs in TRAIN_CASES:
        rng = np.random.default_rng(seed)
        I = rng.integers(0, V, nnz).astype(np.int32)
        J = rng.integers(0, V, nnz).astype(np.int32)
        X = rng.uniform(1e-3, 0.4, nnz).astype(np.float32)
        t = oracle.Trainer(V, I, J, X, float(X.max()), dim, method, seed, optimizer=optimizer)
        init = {s: t.store(s) for s in ("focus", "context", "fBias", "cBias")}
        hist = [t.epoch() for _ in range(epochs)]
        tr[f"{name}/meta"] = np.array([V, nnz, dim, seed, epochs], np.int64)
        tr[f"{name}/I"], tr[f"{name}/J"], tr[f"{name}/X"] = I, J, X
        tr[f"{name}/perm_last"] = t.get_perm()
        tr[f"{name}/history"] = np.array(hist, np.float64)
        tr[f"{name}/init_focus_row0"] = init["focus"][0].view(np.uint32)
        tr[f"{name}/init_bias"] = np.stack([init["fBias"], init["cBias"]]).view(np.uint32)
        tr[f"{name}/result"] = t.extract().astype(np.float32)   # (focus+context)/2 is exactly f32
        tr[f"{name}/gsFBias"] = t.store("gsFBias")
    np.savez_compressed(os.path.join(HERE, "train.npz"), **tr)

    kat = {k: np.array(v, np.int32) for k, v in JAVA_RANDOM_KAT.items()}
    np.savez_compressed(os.path.join(HERE, "java_random.npz"), **kat)
    for f in ("bca.npz", "train.npz", "java_random.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
