"""world=1 block-rotation trainer on an R-MAT graph: epoch time with the current hub threshold (dev aid)"""
import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
from graph_embeddings_b200 import BCA, Configuration, Opt, synth
from graph_embeddings_b200.dist import DistTrainer
V, E = int(float(sys.argv[1])), int(float(sys.argv[2]))
g = synth.rmat_torch(V, E, 5, device="cuda")
c = Configuration(method="glove", dim=128, bca=BCA(.1, 1e-3, True), opt=Opt("adagrad", 0, 1), init="counter")
tr = DistTrainer(c, 0, 1, 8, None)
nnz, _ = tr.build(g); tr.prepare()
for ep in range(4):
    s0 = tr.stats()["update_seconds"]; cost = tr.epoch(); dt = tr.stats()["update_seconds"] - s0
    print(f"epoch {ep} cost {cost:.5f} {dt*1e3:.1f} ms {nnz/dt/1e9:.3f} Gnnz/s", flush=True)
