"""Similarity-edge compare: device pairs/s on a synthetic name group, the CPU oracle on a bounded
sample beside it (dev measurement for DESIGN.md; writes gpurun_out/sim_bench.json)."""
import json, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from graph_embeddings_b200 import BCA, Configuration, Opt, similarity as S
from graph_embeddings_b200.api import _Native
from oracle import similarity as O
from test_gpu_similarity import name_labels

native = _Native(Configuration(method="glove", dim=8, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, 1)))
out = []
for method, thr, n in [("ngram_jaccard", 0.6, 30000), ("ngram_cosine", 0.7, 30000), ("token_jaccard", 0.5, 30000), ("numeric", 0.5, 30000)]:
    rng = np.random.default_rng(1)
    if method == "numeric":
        labels = [str(int(v)) + "^^http://www.w3.org/2001/XMLSchema#integer" for v in rng.integers(1000, 9999, n)]
    else:
        labels = [s + " " + str(int(rng.integers(0, 50))) for s in name_labels(rng, n)]
    cfg = {"predicate": "p", "method": method, "threshold": thr, "smooth": 0.5}
    g = S.CompareGroup(cfg)
    for v in range(n):
        g.addToSource(v); g.addToTarget(v)
    lab = dict(enumerate(labels))
    t0 = time.time(); res = g.compare(native, lab); wall = time.time() - t0
    res = g.compare(native, lab)
    pairs = n * (n - 1) // 2
    m = 2500
    t0 = time.time(); exp = O.pairs(method.upper(), labels[:m], np.arange(m), g.source[:m], g.target[:m], thr, smooth=0.5, upper_triangle=True); cpu = time.time() - t0
    rec = {"method": method, "n": n, "pairs": pairs, "kept": int(len(res[0])), "device_seconds": g.seconds, "device_pairs_per_s": pairs / g.seconds,
           "host_prep_plus_device_wall_s": wall, "cpu_oracle_sample_n": m, "cpu_oracle_pairs_per_s_1thread": (m * (m - 1) // 2) / cpu}
    print(rec, flush=True); out.append(rec)
json.dump(out, open("gpurun_out/sim_bench.json", "w"), indent=1)
