#!/usr/bin/env python
"""BCA builder: tier-1 table capacity sweep (GE_BCA_T1_CAP) on a workload; COO checksum must not move"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, Opt, synth
wl = sys.argv[1] if len(sys.argv) > 1 else "c4"
if wl == "c1":
    g = synth.dblp_like(1.0)
else:
    V, E = {"c4": (10_000_000, 100_000_000), "c3": (2_000_000, 20_000_000)}[wl]
    g = synth.rmat_torch(V, E, 5, device="cuda")
c = Configuration(method="glove", dim=128, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, 1), seed=12345)
out = {"workload": wl, "caps": {}}
ref = None
for cap in sys.argv[2:] or ["472", "232", "152", "104"]:
    os.environ["GE_BCA_T1_CAP"] = cap
    best = None
    for rep in range(2):
        b = BookmarkColoring(g, c)
        st = b.stats()
        if best is None or st["seconds"] < best["seconds"]:
            best = st
        if rep == 0:
            I, J, X = b.coo()
            chk = (int(I.astype(np.int64).sum()), int(J.astype(np.int64).sum()), float(X.astype(np.float64).sum()), len(I))
            assert ref is None or chk == ref, (chk, ref)
            ref = chk
            del I, J, X
        del b
        torch.cuda.empty_cache()
    out["caps"][cap] = {k: best[k] for k in ("seconds", "pops", "tier2_rows", "fetched")}
    print(cap, out["caps"][cap], flush=True)
out["checksum"] = ref
print(json.dumps(out))
