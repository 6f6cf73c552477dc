import sys
sys.path.insert(0, ".")
import numpy as np
from graph_embeddings_b200 import BCA, Configuration, ImportedMatrix, Opt, createOptimizer, synth
from graph_embeddings_b200.dist import DistTrainer
V, nnz, dim = int(float(sys.argv[1])), int(float(sys.argv[2])), int(sys.argv[3])
zipf = float(sys.argv[4]) if len(sys.argv) > 4 else 0.8
I, J, X = synth.synthetic_coo_torch(V, nnz, 0, V, seed=6, col_zipf=zipf)
cnt = np.sort(np.bincount(J, minlength=V))[::-1] / nnz
print("col share top1 %.4f top10 %.4f ; row max %d" % (cnt[0], cnt[:10].sum(), np.bincount(I, minlength=V).max()))
for upd in ("atomic", "store"):
    c = Configuration(method="glove", dim=dim, bca=BCA(.1, 1e-3, True), opt=Opt("adagrad", 0, 1), init="java", update=upd)
    opt = createOptimizer(c, ImportedMatrix(V, I, J, X, c, xmax=0.2))
    print("single", upd, [round(opt.epoch(), 5) for _ in range(6)], flush=True)
c = Configuration(method="glove", dim=dim, bca=BCA(.1, 1e-3, True), opt=Opt("adagrad", 0, 1), init="java")
tr = DistTrainer(c, 0, 1, 8, None)
tr.import_rows(V, I, J, X, xmax=0.2); tr.prepare()
print("dist world=1", [round(tr.epoch(), 5) for _ in range(6)], np.isfinite(tr.extract_local()).all(), flush=True)
