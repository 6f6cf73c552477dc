"""A/B of the host drain threads of ge_extract (f64 result, fresh destination each time) on a C4-sized
store (V = 10 M, d = 128: 5.1 GB over PCIe, 10.2 GB written on the host).  One JSON line."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_embeddings_b200 import BCA, Adagrad, Configuration, ImportedMatrix, Opt  # noqa: E402

V, d = 10_000_000, 128
cfg = Configuration(method="glove", dim=d, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0, 1))
I = np.arange(1000, dtype=np.int32); J = (I * 7 + 3) % V
m = ImportedMatrix(V, I, J.astype(np.int32), np.full(1000, 0.5, np.float32), cfg)
opt = Adagrad(m, cfg)
res = {}
for thr in (8, 16, 8, 16, 12):
    os.environ["GE_HOST_THREADS"] = str(thr)
    t0 = time.perf_counter()
    out = opt.extractResult()
    dt = time.perf_counter() - t0
    res.setdefault(str(thr), []).append(round(dt, 4))
    chk = float(out[::100003, 0].sum())
    del out
del os.environ["GE_HOST_THREADS"]
t0 = time.perf_counter(); out = opt.extractResult(); res["default"] = [round(time.perf_counter() - t0, 4)]
print(json.dumps({"V": V, "dim": d, "cores": os.cpu_count(), "extract_f64_seconds_by_threads": res, "checksum": chk}))
