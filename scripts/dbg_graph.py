import sys
sys.path.insert(0, ".")
import numpy as np
from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, Opt, synth
c = Configuration(method="glove", dim=8, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0, 1))
V, E = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(0)
el = synth.EdgeList(V, rng.integers(0, V, E).astype(np.int32), rng.integers(0, V, E).astype(np.int32), rng.uniform(.1, 1, E).astype(np.float32))
b = BookmarkColoring(el, c)
print("ok", b.coOccurrenceCount())
