"""Times ge_flatten_edges_static (no allocation, static scratch) beside ge_set_graph_edges on the same
edge list: device seconds for both directions, CUDA events on the stream the kernels run on.
Usage: python scripts/r02_static_flatten_bench.py  ->  one JSON line."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_embeddings_b200 import BCA, Configuration, Opt  # noqa: E402
from graph_embeddings_b200.api import _Native, _ptr  # noqa: E402

V, E = 200_000, 1 << 20
rng = np.random.default_rng(5)
s = rng.integers(0, V, E).astype(np.int32); d = rng.integers(0, V, E).astype(np.int32)
w = rng.uniform(0.05, 1, E).astype(np.float32)
n = _Native(Configuration(method="glove", dim=8, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0, 1)))
L = n.L
dyn = []
for _ in range(5):
    n.ck(L.ge_set_graph_edges(n.h, V, E, _ptr(s), _ptr(d), _ptr(w)))
    sec = C.c_double(); eo = C.c_int64(); ei = C.c_int64()
    n.ck(L.ge_graph_sizes(n.h, C.byref(eo), C.byref(ei), C.byref(sec)))
    dyn.append(sec.value)
dev = torch.device("cuda", 0)
ds, dd, dw = (torch.from_numpy(a).to(dev) for a in (s, d, w))
off = torch.empty(V + 1, dtype=torch.int64, device=dev); nbr = torch.empty(E, dtype=torch.int32, device=dev)
wo = torch.empty(E, dtype=torch.float32, device=dev); info = torch.empty(2, dtype=torch.int64, device=dev)
n.set_stream(torch.cuda.current_stream().cuda_stream)
args = lambda by: (n.h, V, E, by, ds.data_ptr(), dd.data_ptr(), dw.data_ptr(), off.data_ptr(), nbr.data_ptr(), wo.data_ptr(), info.data_ptr())
free0 = torch.cuda.mem_get_info()[0]
st = []
for it in range(8):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    n.ck(L.ge_flatten_edges_static(*args(0))); n.ck(L.ge_flatten_edges_static(*args(1)))
    e1.record(); torch.cuda.synchronize()
    st.append(e0.elapsed_time(e1) * 1e-3)
free1 = torch.cuda.mem_get_info()[0]
n.set_stream(None)
print(json.dumps({"V": V, "E": E, "unique_in": int(info[0]), "bad": int(info[1]),
                  "ge_set_graph_edges_flatten_s": sorted(dyn)[len(dyn) // 2], "ge_flatten_edges_static_both_s": sorted(st[3:])[len(st[3:]) // 2],
                  "kernels_per_direction": 1 + 5 * ((2 * 18 + 7) // 8) + 6,
                  "device_free_bytes_delta_over_calls": free0 - free1,
                  "algorithmic_bytes_both": 2 * (12 + 8) * E}))
