#!/usr/bin/env python
"""hub report (hottest id's concurrency, damped share, capped launches) of a workload: single launch and 8x8 buckets"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, Opt, createOptimizer, synth
from graph_embeddings_b200.dist import DistTrainer
wl = sys.argv[1] if len(sys.argv) > 1 else "c4"
V, E = {"c4": (10_000_000, 100_000_000), "c3": (2_000_000, 20_000_000)}[wl]
g = synth.rmat_torch(V, E, 5, device="cuda")
c = Configuration(method="glove", dim=128, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, 1), seed=12345)
out = {"workload": wl}
bca = BookmarkColoring(g, c)
opt = createOptimizer(c, bca)
opt.epoch()
out["single"] = opt.stats()["hub"]
del opt, bca
tr = DistTrainer(c, 0, 1, 8, None)
tr.build(g)
tr.prepare()
out["buckets_8x8x2"] = tr.stats()["hub"]
print(json.dumps(out))
