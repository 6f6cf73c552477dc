"""CPU oracle — TEST INFRASTRUCTURE ONLY (parity unpinned; see oracle.cpp header).

ctypes front-end for ``liboracle.so`` (the C++ restatement of the reference's Java hot path).
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import this package.  The product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

NORM = {"none": 0, "unity": 1, "counts": 2}
METHOD = {"glove": 0, "pglove": 1}
OPTIMIZER = {"adagrad": 0, "adam": 1, "amsgrad": 2}


def build(force: bool = False) -> str:
    """Compile liboracle.so (gcc only; no GPU, no reference sources)."""
    if force or not os.path.exists(_LIB_PATH) or (
        os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "oracle.cpp"))
    ):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        p, i32, i64, f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
        L.orc_jr_ints.argtypes = [i64, i32, p]
        L.orc_jr_bounded.argtypes = [i64, i32, i32, p]
        L.orc_jr_floats.argtypes = [i64, i32, p]
        L.orc_jr_shuffle.argtypes = [i64, i64, i32, i32, p]
        L.orc_cooc_build.argtypes = [i32, p, p, p, p, p, p, f64, f64, i32, i32, i32, i32, i32]
        L.orc_cooc_build.restype = p
        for n in ("orc_cooc_nnz", "orc_cooc_pops", "orc_cooc_scanned"):
            getattr(L, n).argtypes = [p]
            getattr(L, n).restype = i64
        L.orc_cooc_max.argtypes = [p]
        L.orc_cooc_max.restype = f64
        L.orc_cooc_get.argtypes = [p, p, p, p, p]
        L.orc_cooc_free.argtypes = [p]
        L.orc_train_create.argtypes = [i32, i64, p, p, p, f64, i32, i32, i64, i32]
        L.orc_train_create.restype = p
        L.orc_train_create_opt.argtypes = [i32, i64, p, p, p, f64, i32, i32, i64, i32, i32]
        L.orc_train_create_opt.restype = p
        L.orc_train_epoch.argtypes = [p, i32]
        L.orc_train_epoch.restype = f64
        L.orc_train_set_perm.argtypes = [p, p]
        L.orc_train_get_perm.argtypes = [p, p]
        L.orc_train_optimize.argtypes = [p, i32, f64, p, p]
        L.orc_train_optimize.restype = i32
        L.orc_train_history_exact.argtypes = [p, p, i32]
        L.orc_train_history_exact.restype = i32
        L.orc_train_extract.argtypes = [p, p]
        L.orc_train_get_store.argtypes = [p, i32, p]
        L.orc_train_free.argtypes = [p]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


# ---- java.util.Random ------------------------------------------------------------------
def java_ints(seed: int, n: int) -> np.ndarray:
    out = np.empty(n, np.int32)
    lib().orc_jr_ints(seed, n, _ptr(out))
    return out


def java_bounded(seed: int, bound: int, n: int) -> np.ndarray:
    out = np.empty(n, np.int32)
    lib().orc_jr_bounded(seed, bound, n, _ptr(out))
    return out


def java_floats(seed: int, n: int) -> np.ndarray:
    out = np.empty(n, np.float32)
    lib().orc_jr_floats(seed, n, _ptr(out))
    return out


def java_shuffle(seed: int, skip: int, n: int, rounds: int = 1) -> np.ndarray:
    out = np.empty(n, np.int32)
    lib().orc_jr_shuffle(seed, skip, n, rounds, _ptr(out))
    return out


# ---- co-occurrence ----------------------------------------------------------------------
class Cooc:
    """Result of the BCA build: COO (I, J, X), row offsets, global max, event counters."""

    def __init__(self, row_off, I, J, X, xmax, pops, scanned):
        self.row_off, self.I, self.J, self.X = row_off, I, J, X
        self.max, self.pops, self.scanned = xmax, pops, scanned

    @property
    def nnz(self):
        return int(self.I.shape[0])

    def sorted(self):
        """(I, J, X) sorted by (i, j) — the set-compare form."""
        order = np.lexsort((self.J, self.I))
        return self.I[order], self.J[order], self.X[order]


def cooc_build(csr, alpha=0.1, epsilon=1e-3, directed=True, normalize="none", threads=1,
               first=0, count=-1) -> Cooc:
    """csr: object with V, out_off/out_nbr/out_w, in_off/in_nbr/in_w (numpy, i64/i32/f32)."""
    L = lib()
    h = L.orc_cooc_build(csr.V, _ptr(csr.out_off), _ptr(csr.out_nbr), _ptr(csr.out_w),
                         _ptr(csr.in_off), _ptr(csr.in_nbr), _ptr(csr.in_w), alpha, epsilon,
                         int(directed), NORM[normalize], threads, first, count)
    try:
        nnz = L.orc_cooc_nnz(h)
        row_off = np.empty(csr.V + 1, np.int64)
        I = np.empty(nnz, np.int32)
        J = np.empty(nnz, np.int32)
        X = np.empty(nnz, np.float32)
        L.orc_cooc_get(h, _ptr(row_off), _ptr(I), _ptr(J), _ptr(X))
        return Cooc(row_off, I, J, X, L.orc_cooc_max(h), L.orc_cooc_pops(h), L.orc_cooc_scanned(h))
    finally:
        L.orc_cooc_free(h)


# ---- trainer ----------------------------------------------------------------------------
class Trainer:
    """Restatement of Optimizer + Adagrad / Adam / AMSGrad (J/opt/Optimizer.java, J/opt/grad/*.java).
    For adam / amsgrad the gs* names are the second moments M2*, m1* the first moments."""

    STORE = {"focus": 0, "context": 1, "fBias": 2, "cBias": 3,
             "gsFocus": 4, "gsContext": 5, "gsFBias": 6, "gsCBias": 7,
             "m1Focus": 8, "m1Context": 9, "m1FBias": 10, "m1CBias": 11}

    def __init__(self, V, I, J, X, xmax, dim, method="glove", seed=12345, threads=1, optimizer="adagrad"):
        self.V, self.dim, self.nnz = int(V), int(dim), int(len(I))
        self._keep = (np.ascontiguousarray(I, np.int32), np.ascontiguousarray(J, np.int32),
                      np.ascontiguousarray(X, np.float32))
        self.optimizer = optimizer
        self._h = lib().orc_train_create_opt(self.V, self.nnz, *map(_ptr, self._keep), float(xmax),
                                             self.dim, METHOD[method], seed, threads, OPTIMIZER[optimizer])

    def epoch(self, shuffle=True) -> float:
        return lib().orc_train_epoch(self._h, int(shuffle))

    def set_perm(self, perm):
        perm = np.ascontiguousarray(perm, np.int32)
        assert perm.shape[0] == self.nnz
        lib().orc_train_set_perm(self._h, _ptr(perm))

    def get_perm(self):
        out = np.empty(self.nnz, np.int32)
        lib().orc_train_get_perm(self._h, _ptr(out))
        return out

    def optimize(self, maxiter, tolerance):
        hist = np.zeros(maxiter, np.float64)
        fc = C.c_double(0)
        n = lib().orc_train_optimize(self._h, maxiter, tolerance, _ptr(hist), C.byref(fc))
        return hist[:n].copy(), fc.value

    def history_exact(self) -> np.ndarray:
        """Per-epoch cost of the epochs run so far with the job sums kept in fp64.  NOT what the
        reference reports: Adagrad.java:60 accumulates in a float, which stops absorbing increments
        after some 10-30 M nonzeros per thread (see oracle.cpp Trainer::job)."""
        out = np.zeros(65536, np.float64)
        n = lib().orc_train_history_exact(self._h, _ptr(out), len(out))
        return out[:n].copy()

    def extract(self) -> np.ndarray:
        out = np.empty((self.V, self.dim), np.float64)
        lib().orc_train_extract(self._h, _ptr(out))
        return out

    def store(self, name) -> np.ndarray:
        w = self.STORE[name]
        shape = (self.V,) if w in (2, 3, 6, 7, 10, 11) else (self.V, self.dim)
        out = np.empty(shape, np.float32)
        lib().orc_train_get_store(self._h, w, _ptr(out))
        return out

    def close(self):
        if self._h:
            lib().orc_train_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
