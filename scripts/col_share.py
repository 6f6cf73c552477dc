import sys
sys.path.insert(0, ".")
import numpy as np, torch
from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, Opt, synth
V, E = int(float(sys.argv[1])), int(float(sys.argv[2]))
g = synth.rmat_torch(V, E, 5, device="cuda")
c = Configuration(method="glove", dim=8, bca=BCA(.1, 1e-3, True), opt=Opt("adagrad", 0, 1))
b = BookmarkColoring(g, c); I, J, X = b.coo(); nnz = len(J)
cnt = np.sort(np.bincount(J, minlength=V))[::-1] / nnz
print("V %d nnz %d col share top1 %.5f top10 %.5f top100 %.5f top1000 %.5f; row max %d; X min %.2e max %.3f" % (
    V, nnz, cnt[0], cnt[:10].sum(), cnt[:100].sum(), cnt[:1000].sum(), np.bincount(I, minlength=V).max(), X.min(), X.max()))
Bs = (V + 7) // 8
blk = J // Bs
for k in range(8):
    m = blk == k
    cj = np.bincount(J[m] - k * Bs, minlength=Bs)
    print(" col-block %d: nnz share %.3f, hottest column holds %.4f of the block" % (k, m.mean(), cj.max() / m.sum()))
