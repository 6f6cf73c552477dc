"""Quick device-timed update-kernel throughput on a synthetic Zipf COO (development aid)."""
import sys, time
sys.path.insert(0, ".")
import numpy as np
from graph_embeddings_b200 import BCA, Configuration, ImportedMatrix, Opt, createOptimizer, synth

V = int(float(sys.argv[1])) if len(sys.argv) > 1 else 2_000_000
nnz = int(float(sys.argv[2])) if len(sys.argv) > 2 else 50_000_000
dim = int(sys.argv[3]) if len(sys.argv) > 3 else 128
zipf = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
t = time.time()
if zipf > 0:
    I, J, X = synth.synthetic_coo(V, nnz, 6)
else:
    rng = np.random.default_rng(1)
    I = rng.integers(0, V, nnz).astype(np.int32); J = rng.integers(0, V, nnz).astype(np.int32)
    X = rng.uniform(1e-4, .2, nnz).astype(np.float32)
print("gen", time.time() - t, flush=True)
c = Configuration(method="glove", dim=dim, bca=BCA(.1, 1e-3, True), opt=Opt("adagrad", 0, 1), init="counter")
m = ImportedMatrix(V, I, J, X, c)
opt = createOptimizer(c, m)
for ep in range(6):
    s0 = opt.stats()["update_seconds"]
    cost = opt.epoch()
    dt = opt.stats()["update_seconds"] - s0
    B = 32 * dim + 44
    print(f"epoch {ep} cost {cost:.6f} {dt*1e3:.2f} ms  {nnz/dt/1e9:.3f} Gnnz/s  {nnz*B/dt/1e9:.0f} GB/s algorithmic "
          f"({nnz*B/dt/1e9/6538.9:.3f} of measured HBM)", flush=True)
