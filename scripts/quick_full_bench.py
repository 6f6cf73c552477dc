"""R-MAT graph -> GPU BCA build -> GPU training epochs; prints device-timed throughput (dev aid)."""
import sys, time
sys.path.insert(0, ".")
import numpy as np
from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, Opt, createOptimizer, synth

V = int(float(sys.argv[1])) if len(sys.argv) > 1 else 2_000_000
E = int(float(sys.argv[2])) if len(sys.argv) > 2 else 20_000_000
dim = int(sys.argv[3]) if len(sys.argv) > 3 else 128
epochs = int(sys.argv[4]) if len(sys.argv) > 4 else 5
t = time.time(); g = synth.rmat(V, E, 4); print("gen %.1fs V=%d E=%d" % (time.time() - t, g.V, g.E), flush=True)
upd = sys.argv[5] if len(sys.argv) > 5 else "atomic"
c = Configuration(method="glove", dim=dim, bca=BCA(.1, 1e-3, True), opt=Opt("adagrad", 0, 1), init="counter", update=upd)
t = time.time(); b = BookmarkColoring(g, c); wall = time.time() - t
s = b.stats(); nnz = s["nnz"]
bytes_build = 8 * s["scanned"] + 8 * s["pops"] + 12 * nnz
print("BCA: nnz %d (%.1f/row) pops %d scanned %d  device %.3fs wall %.2fs  %.1f Mpops/s %.1f Mnnz/s  alg %.0f GB/s max %.5f"
      % (nnz, nnz / V, s["pops"], s["scanned"], s["seconds"], wall, s["pops"] / s["seconds"] / 1e6, nnz / s["seconds"] / 1e6,
         bytes_build / s["seconds"] / 1e9, b.max()), flush=True)
I, J, X = b.coo()
cnt = np.sort(np.bincount(J, minlength=V))[::-1] / nnz
print("col share top1 %.4f top10 %.4f top100 %.4f top1000 %.4f" % (cnt[0], cnt[:10].sum(), cnt[:100].sum(), cnt[:1000].sum()), flush=True)
opt = createOptimizer(c, b)
B = 32 * dim + 44
for ep in range(epochs):
    s0 = opt.stats()["update_seconds"]; cost = opt.epoch(); dt = opt.stats()["update_seconds"] - s0
    print(f"epoch {ep} cost {cost:.6f} {dt*1e3:.2f} ms {nnz/dt/1e9:.3f} Gnnz/s {nnz*B/dt/1e9:.0f} GB/s ({nnz*B/dt/1e9/6538.9:.3f} of measured HBM)", flush=True)
