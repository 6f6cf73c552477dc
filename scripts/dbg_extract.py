"""dev aid: D2H time of the result (V x dim f32) through ge_extract_f32"""
import sys, time
sys.path.insert(0, ".")
import numpy as np
from graph_embeddings_b200 import BCA, Configuration, ImportedMatrix, Opt, createOptimizer
V, dim = int(float(sys.argv[1])), int(sys.argv[2])
c = Configuration(method="glove", dim=dim, bca=BCA(.1, 1e-3, True), opt=Opt("adagrad", 0, 1), init="counter")
I = np.arange(1000, dtype=np.int32); J = I[::-1].copy(); X = np.full(1000, 0.5, np.float32)
opt = createOptimizer(c, ImportedMatrix(V, I, J, X, c))
for k in range(3):
    t0 = time.time(); e = opt.extractResultF32(); dt = time.time() - t0
    print(f"extract f32 {e.nbytes/1e9:.2f} GB in {dt:.3f} s = {e.nbytes/dt/1e9:.1f} GB/s", flush=True)
    f = opt.store("focus"); cc = opt.store("context")
    assert np.array_equal(e[::1000], ((f[::1000] + cc[::1000]) / np.float32(2))); del e, f, cc
t0 = time.time(); e = opt.extractResult(); dt = time.time() - t0
print(f"extract f64 {e.nbytes/1e9:.2f} GB in {dt:.3f} s = {e.nbytes/dt/1e9:.1f} GB/s")
