#!/usr/bin/env python
"""Round-2 GPU probe: dumps what the throughput path produces on the parity-fixture cases
(tests/golden/make_parity_curves.py) so that it can be compared with the single-thread oracle
trajectories offline, plus the hub (Zipf) case against a live oracle run.

    python scripts/r02_probe.py parity [case ...]   -> gpurun_out/probe_<case>.npz
    python scripts/r02_probe.py zipf                -> gpurun_out/probe_zipf.json
    python scripts/r02_probe.py blocks              -> gpurun_out/probe_blocks.json  (C4, block orders on one GPU)
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)


def topk_gpu(E, sample, k=10):
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


def parity(names):
    import make_parity_curves as mp
    from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, ImportedMatrix, Opt, createOptimizer
    for name in names or list(mp.CASES):
        kind, args, method, dim, epochs = mp.CASES[name]
        g = mp.make_graph(kind, args)
        base = dict(method=method, dim=dim, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, epochs), seed=mp.SEED)
        c0 = Configuration(**base)
        t0 = time.time()
        bca = BookmarkColoring(g, c0)
        I, J, X = bca.coo()
        res = dict(coo=mp.coo_checksums(I, J, X), xmax=np.float64(bca.max()), build_s=np.float64(bca.stats()["seconds"]))
        sample = np.sort(np.random.default_rng(0).choice(g.V, min(mp.SAMPLE, g.V), replace=False)).astype(np.int32)
        for tag, kw in (("device", {}), ("java", dict(order="java")), ("store", dict(update="store")),
                        ("javamath", dict(math="java"))):
            c = Configuration(**base, **kw)
            m = bca if tag == "device" else ImportedMatrix(g.V, I, J, X, c, xmax=bca.max())
            opt = createOptimizer(c, m)
            t1 = time.time()
            hist = [opt.epoch() for _ in range(epochs)]
            dt = time.time() - t1
            st = opt.stats()
            res[f"hist_{tag}"] = np.array(hist)
            res[f"nn_{tag}"] = topk_gpu(opt.extractResultF32(), sample)
            res[f"upd_s_{tag}"] = np.float64(st["update_seconds"] / epochs)
            print(f"[{name}/{tag}] nnz={len(I)} {epochs} epochs {dt:.2f}s kernel {st['update_seconds'] / epochs * 1e3:.2f} ms/epoch "
                  f"hist={np.round(hist, 6)}", flush=True)
            del opt, m
        np.savez_compressed(os.path.join(OUT, f"probe_{name}.npz"), sample=sample, **res)
        print(f"[{name}] done in {time.time() - t0:.1f}s", flush=True)


def zipf():
    """hub columns: GPU update modes vs the single-thread oracle, per epoch"""
    import oracle
    from graph_embeddings_b200 import BCA, Configuration, ImportedMatrix, Opt, createOptimizer, synth
    from graph_embeddings_b200.dist import DistTrainer
    V, nnz, dim, epochs = 300_000, 6_000_000, 32, 6
    I, J, X = synth.synthetic_coo_torch(V, nnz, 0, V, seed=6, col_zipf=1.0)
    out = {"top_col_share": float(np.bincount(J, minlength=V).max() / nnz), "hot_writers": os.environ.get("GE_HOT_WRITERS", "16")}
    t0 = time.time()
    t = oracle.Trainer(V, I, J, X, 0.2, dim, "glove", 31)
    out["oracle_f32sum"] = [t.epoch() for _ in range(epochs)]
    out["oracle"] = t.history_exact().tolist()          # fp64-summed loss of the same trajectory
    out["oracle_s"] = time.time() - t0
    t.close()
    for upd in ("atomic", "store"):
        c = Configuration(method="glove", dim=dim, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, epochs), seed=31, update=upd)
        opt = createOptimizer(c, ImportedMatrix(V, I, J, X, c, xmax=0.2))
        out[upd] = [opt.epoch() for _ in range(epochs)]
        out[upd + "_ms"] = opt.stats()["update_seconds"] / epochs * 1e3
    c = Configuration(method="glove", dim=dim, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, epochs), seed=31)
    opt = createOptimizer(c, ImportedMatrix(V, I, J, X, c, xmax=0.2))
    opt.epoch()
    out["hub"] = opt.stats()["hub"]
    del opt
    tr = DistTrainer(c, 0, 1, 8, None)
    tr.import_rows(V, I, J, X, xmax=0.2)
    tr.prepare()
    out["dist8x8"] = [tr.epoch() for _ in range(epochs)]
    out["dist_hub"] = tr.stats()["hub"]
    out["dist8x8_ms"] = tr.stats()["update_seconds"] / epochs * 1e3
    # the oracle walking the same block order (sub-epoch s: half blocks (k, (k+s) mod 8, p), random inside)
    Bs = (V + 7) // 8
    Hs = (Bs + 1) // 2
    bucket = ((I // Bs) * 8 + J // Bs) * 2 + (J % Bs >= Hs)
    srt = np.argsort(bucket, kind="stable")
    start = np.searchsorted(bucket[srt], np.arange(129))
    rng = np.random.default_rng(1)
    t = oracle.Trainer(V, I, J, X, 0.2, dim, "glove", 31)
    for _ in range(epochs):
        order = []
        for s_ in range(8):
            for k in range(8):
                for p_ in range(2):
                    b = (k * 8 + (k + s_) % 8) * 2 + p_
                    order.append(rng.permutation(srt[start[b]:start[b + 1]]))
        t.set_perm(np.concatenate(order).astype(np.int32))
        t.epoch(False)
    out["oracle_blockwalk"] = t.history_exact().tolist()
    t.close()
    o = np.array(out["oracle"])
    for k in ("atomic", "store", "dist8x8"):
        out[k + "_rel"] = (np.abs(np.array(out[k]) - o) / o).tolist()
    ob = np.array(out["oracle_blockwalk"])
    out["dist8x8_rel_blockwalk"] = (np.abs(np.array(out["dist8x8"]) - ob) / ob).tolist()
    print(json.dumps(out), flush=True)
    tag = "w%s_hs%s_on%s" % (out["hot_writers"], os.environ.get("GE_HUB_SERIAL", "16"), os.environ.get("GE_HUB_ONSET", "64"))
    with open(os.path.join(OUT, f"probe_zipf_{tag}.json"), "w") as f:
        json.dump(out, f)


def blocks():
    """C4 on one GPU: uniformly shuffled order vs nb x nb block orders (DistTrainer world=1)"""
    import torch
    from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, Opt, createOptimizer, synth
    from graph_embeddings_b200.dist import DistTrainer
    wl = os.environ.get("PROBE_WL", "c4")
    V, E = {"c4": (10_000_000, 100_000_000), "c3": (2_000_000, 20_000_000)}[wl]
    g = synth.rmat_torch(V, E, 5, device="cuda")
    out = {"workload": wl}
    epochs = 4
    c = Configuration(method="glove", dim=128, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, epochs), seed=12345)
    bca = BookmarkColoring(g, c)
    opt = createOptimizer(c, bca)
    out["random"] = [opt.epoch() for _ in range(epochs)]
    out["random_ms"] = opt.stats()["update_seconds"] / epochs * 1e3
    out["nnz"] = bca.coOccurrenceCount()
    del opt, bca
    torch.cuda.empty_cache()
    for nb in (8, 16, 32):
        tr = DistTrainer(c, 0, 1, nb, None)
        tr.build(g)
        tr.prepare()
        out[f"nb{nb}"] = [tr.epoch() for _ in range(epochs)]
        out[f"nb{nb}_ms"] = tr.stats()["update_seconds"] / epochs * 1e3
        del tr
        print(json.dumps(out), flush=True)
    with open(os.path.join(OUT, f"probe_blocks_{wl}.json"), "w") as f:
        json.dump(out, f)


def keep():
    """C4 (or PROBE_WL) on one GPU: L2-resident hot-row budget sweep, same matrix, fresh init each time"""
    import torch
    from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, Opt, createOptimizer, synth
    wl = os.environ.get("PROBE_WL", "c4")
    V, E = {"c4": (10_000_000, 100_000_000), "c3": (2_000_000, 20_000_000)}[wl]
    dim = int(os.environ.get("PROBE_DIM", "128"))
    g = synth.rmat_torch(V, E, 5, device="cuda")
    epochs = 4
    c = Configuration(method="glove", dim=dim, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, epochs), seed=12345)
    bca = BookmarkColoring(g, c)
    opt = createOptimizer(c, bca)
    out = {"workload": wl, "dim": dim, "nnz": bca.coOccurrenceCount()}
    for mb in os.environ.get("PROBE_KEEP", "0,32,64,96").split(","):
        os.environ["GE_L2_KEEP_MB"] = mb
        opt.reset()
        s0 = opt.stats()["update_seconds"]
        hist = [opt.epoch() for _ in range(epochs)]
        ms = (opt.stats()["update_seconds"] - s0) / epochs * 1e3
        out[f"keep{mb}"] = {"ms": ms, "hist": hist}
        print(f"keep {mb} MB: {ms:.1f} ms/epoch hist {np.round(hist, 6)}", flush=True)
    with open(os.path.join(OUT, f"probe_keep_{wl}_d{dim}.json"), "w") as f:
        json.dump(out, f)


if __name__ == "__main__":
    {"parity": lambda: parity(sys.argv[2:]), "zipf": zipf, "blocks": blocks, "keep": keep}[sys.argv[1]]()
