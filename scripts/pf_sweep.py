"""L2 row-prefetch distance sweep for the update kernel (development aid): one matrix, epochs timed
per GE_PF_DIST value.  usage: pf_sweep.py V E dim"""
import os, sys
sys.path.insert(0, ".")
import torch
from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, Opt, createOptimizer, synth

V, E, dim = int(float(sys.argv[1])), int(float(sys.argv[2])), int(sys.argv[3])
g = synth.rmat_torch(V, E, 5, device="cuda")
c = Configuration(method="glove", dim=dim, bca=BCA(.1, 1e-3, True), opt=Opt("adagrad", 0, 1), init="counter")
b = BookmarkColoring(g, c); nnz = b.coOccurrenceCount()
opt = createOptimizer(c, b)
B = 32 * dim + 44
for _ in range(2):
    opt.epoch()
for pf in [int(x) for x in os.environ.get('PF_LIST', '0').split(',')]:
    os.environ["GE_PF_DIST"] = str(pf)
    ts = []
    for _ in range(3):
        s0 = opt.stats()["update_seconds"]; cost = opt.epoch(); ts.append(opt.stats()["update_seconds"] - s0)
    dt = min(ts)
    print(f"pf={pf:2d} best {dt*1e3:8.2f} ms  {nnz/dt/1e9:.3f} Gnnz/s  {nnz*B/dt/1e9:.0f} GB/s alg ({nnz*B/dt/1e9/6548.2:.3f})  cost {cost:.5f}", flush=True)
