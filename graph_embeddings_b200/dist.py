"""One-process-per-GPU front end of the block-rotation trainer (csrc/dist.cu).  torch.distributed
is used only to hand the NCCL unique id to the other ranks; the data path is ncclSend/ncclRecv
inside libgraphemb.so."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .api import Configuration, _Native, _ptr


class DistTrainer:
    def __init__(self, config: Configuration, rank: int = 0, world: int = 1, nblocks: int = 8, id_bytes: bytes | None = None,
                 layout: str = "contiguous"):
        """layout: "contiguous" — block b = ids [b*Bs, (b+1)*Bs); "cyclic" — block b = ids b, b+nblocks, ...
        (balanced buckets on id-ordered graphs such as the DBLP-shaped configs; profiles/r02_block_balance.md)."""
        self.config, self.rank, self.world, self.nb = config, rank, world, nblocks
        self._n = _Native(config)
        buf = C.create_string_buffer(id_bytes, _lib.NCCL_ID_BYTES) if id_bytes else None
        self._n.ck(self._n.L.ge_dist_init(self._n.h, rank, world, buf))
        self.layout = {"contiguous": 0, "cyclic": 1}[layout]
        self._n.ck(self._n.L.ge_dist_set_layout(self._n.h, self.layout))
        self.V = 0

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(_lib.NCCL_ID_BYTES)
        rc = _lib.lib().ge_dist_unique_id(buf)
        if rc:
            raise _lib.GraphEmbError(rc, "ncclGetUniqueId failed (libnccl not loadable?)")
        return buf.raw

    def row_range(self, V: int):
        """contiguous layout: this rank's ids [first, first + count).  Cyclic: the same numbers are VIRTUAL ids
        (whole blocks, padding included) — use owned_ids(V) for the vertices."""
        f, c = C.c_int32(0), C.c_int32(0)
        self._n.ck(self._n.L.ge_dist_row_range(self._n.h, V, self.nb, C.byref(f), C.byref(c)))
        return f.value, c.value

    def build(self, graph, root: int | None = None):
        """hand the CSR to every rank, build this rank's co-occurrence rows (bookmarks are independent).
        root = None: every rank uploads its own host copy; root = r: only rank r's arrays are read (one
        H2D there + ncclBroadcast over NVLink), the other ranks may pass graph = None."""
        n, L, g = self._n, self._n.L, graph
        if root is None or self.world == 1:
            arrs = [np.ascontiguousarray(g.out_off, np.int64), np.ascontiguousarray(g.out_nbr, np.int32),
                    np.ascontiguousarray(g.out_w, np.float32), np.ascontiguousarray(g.in_off, np.int64),
                    np.ascontiguousarray(g.in_nbr, np.int32), np.ascontiguousarray(g.in_w, np.float32)]
            n.ck(L.ge_set_graph_csr(n.h, g.V, *map(_ptr, arrs)))
        elif self.rank == root:
            arrs = [np.ascontiguousarray(g.out_off, np.int64), np.ascontiguousarray(g.out_nbr, np.int32),
                    np.ascontiguousarray(g.out_w, np.float32), np.ascontiguousarray(g.in_off, np.int64),
                    np.ascontiguousarray(g.in_nbr, np.int32), np.ascontiguousarray(g.in_w, np.float32)]
            n.ck(L.ge_dist_set_graph_csr(n.h, root, g.V, *map(_ptr, arrs)))
        else:
            n.ck(L.ge_dist_set_graph_csr(n.h, root, 0, None, None, None, None, None, None))
        self.V = L.ge_vocab_size(n.h)
        first, count = self.row_range(self.V)
        nnz, xmax = C.c_int64(0), C.c_double(0)
        if self.layout:
            n.ck(L.ge_cooc_build_rows_cyclic(n.h, self.nb, first, count, C.byref(nnz), C.byref(xmax)))
        else:
            n.ck(L.ge_cooc_build_rows(n.h, first, count, C.byref(nnz), C.byref(xmax)))
        return nnz.value, xmax.value

    def import_rows(self, V, I, J, X, xmax=0.0):
        """feed nonzeros whose rows already lie in this rank's range."""
        I = np.ascontiguousarray(I, np.int32); J = np.ascontiguousarray(J, np.int32); X = np.ascontiguousarray(X, np.float32)
        self._n.ck(self._n.L.ge_cooc_import(self._n.h, V, len(I), _ptr(I), _ptr(J), _ptr(X), float(xmax)))
        self.V = V

    def cooc_stats(self):
        pops, scanned, sec = C.c_int64(0), C.c_int64(0), C.c_double(0)
        self._n.ck(self._n.L.ge_cooc_stats(self._n.h, C.byref(pops), C.byref(scanned), C.byref(sec)))
        fetched, tier2 = C.c_int64(0), C.c_int64(0)
        self._n.ck(self._n.L.ge_cooc_fetched(self._n.h, C.byref(fetched), C.byref(tier2)))
        return {"pops": pops.value, "scanned": scanned.value, "fetched": fetched.value, "tier2_rows": tier2.value, "seconds": sec.value}

    def prepare(self):
        self._n.ck(self._n.L.ge_dist_prepare(self._n.h, self.nb))

    def owned_ids(self, V: int) -> np.ndarray:
        """vertex ids whose rows this rank builds / imports / trains, in extraction order (needs no prepare)"""
        first, count = self.row_range(V)
        if not self.layout:
            return np.arange(first, first + count, dtype=np.int32)
        Bs = (V + self.nb - 1) // self.nb
        vid = np.arange(first, first + count, dtype=np.int64)
        ids = vid // Bs + self.nb * (vid % Bs)
        return ids[ids < V].astype(np.int32)

    def local_ids(self) -> np.ndarray:
        """the same from the prepared trainer (ge_dist_local_ids)"""
        out = np.empty(len(self.owned_ids(self.V)), np.int32)
        n = self._n.L.ge_dist_local_ids(self._n.h, _ptr(out), out.size)
        if n < 0:
            self._n.ck(int(n))
        assert n == out.size
        return out

    def epoch(self) -> float:
        cost = C.c_double(0)
        self._n.ck(self._n.L.ge_dist_train_epoch(self._n.h, C.byref(cost)))
        return cost.value

    def extract_local(self) -> np.ndarray:
        """(W + C) / 2 of this rank's rows, f32, row k = vertex owned_ids(V)[k]"""
        count = len(self.owned_ids(self.V))
        out = np.empty((count, self.config.dim), np.float32)
        self._n.ck(self._n.L.ge_dist_extract_f32(self._n.h, _ptr(out), out.size))
        return out

    def extract_local_f64(self) -> np.ndarray:
        """this rank's rows of Optimum.getResult(): double[count * dim] as the seam returns it"""
        count = len(self.owned_ids(self.V))
        out = np.empty((count, self.config.dim), np.float64)
        self._n.ck(self._n.L.ge_dist_extract(self._n.h, _ptr(out), out.size))
        return out

    def stats(self):
        b = np.zeros(1024, np.int64)
        comm, nnzg = C.c_double(0), C.c_int64(0)
        n = self._n.ck(self._n.L.ge_dist_stats(self._n.h, _ptr(b), 1024, C.byref(comm), C.byref(nnzg)))
        s, l, u = C.c_double(0), C.c_int64(0), C.c_int64(0)
        self._n.ck(self._n.L.ge_train_stats(self._n.h, C.byref(s), C.byref(l), C.byref(u)))
        rot = {-1: "pointer shift (1 GPU)", 0: "nccl send/recv in-stream", 1: "nccl half blocks, second stream",
               2: "copy engine into the neighbour's IPC-mapped ring"}.get(self._n.L.ge_dist_rotation_mode(self._n.h), "?")
        k, sh, nc, w = C.c_double(0), C.c_double(0), C.c_int32(0), C.c_int64(0)
        self._n.ck(self._n.L.ge_train_hub_info(self._n.h, C.byref(k), C.byref(sh), C.byref(nc), C.byref(w)))
        return {"bucket_nnz": b[:n].copy(), "comm_seconds": comm.value, "nnz_global": nnzg.value, "rotation": rot,
                "update_seconds": s.value, "launches": l.value, "updates": u.value,
                "hub": {"max_concurrency": k.value, "damped_share": sh.value, "capped_launches": nc.value, "warp_cap": w.value}}

    def set_stream(self, s):
        self._n.set_stream(s)
