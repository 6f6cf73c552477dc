"""Host-side mirror of the reference's seam (J/Main.java:70-77), backed by libgraphemb.so.

    bca = BookmarkColoring(graph, config)        # J/bca/BookmarkColoring.java:32
    optimizer = createOptimizer(config, bca)     # J/Main.java:107-131
    optimum = optimizer.optimize()               # J/opt/Optimizer.java:67-120
    optimum.getResult()                          # J/opt/Optimum.java:26

Same names, argument meaning and error behaviour as the Java classes; all compute happens in
the CUDA library through the C ABI (include/graphemb.h).  No torch types, no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import GeConfig, GraphEmbError


class InvalidConfigurationException(Exception):
    """Configuration.InvalidConfigurationException (Configuration.java:496-500)."""

    def __init__(self, message):
        super().__init__("Invalid configuration: " + message)


@dataclass
class BCA:  # Configuration.BCA (:320-363)
    alpha: float = 0.0
    epsilon: float = 0.0
    directed: bool = False
    normalize: str | None = None
    reverse: bool | None = None       # accepted and ignored (stale key of the shipped yml)
    predicates: bool | None = None    # accepted and ignored

    def getNormalize(self):
        return "none" if self.normalize is None else self.normalize


@dataclass
class Opt:  # Configuration.Opt (:365-398)
    method: str = "adagrad"
    tolerance: float = 0.0
    maxiter: int = 0


@dataclass
class Configuration:
    """The fields of the reference's YAML bean that the hot path reads."""

    graph: str | None = None
    method: str | None = None
    dim: int = 0
    threads: int = 0
    bca: BCA | None = None
    opt: Opt | None = None
    output: dict | None = None
    weights: dict | None = None
    similarity: list | None = None
    pca: dict | None = None
    # native-path knobs (not in the reference schema; defaults keep the drop-in behaviour)
    seed: int = 12345
    math: str = "fast"
    order: str = "device"
    init: str = "java"
    update: str = "atomic"
    serial: bool = False
    device: int = 0
    extra: dict = field(default_factory=dict)

    @staticmethod
    def check(config: "Configuration", require_io: bool = True):
        """Configuration.check (:478-494)."""
        if not config.dim > 0:
            raise InvalidConfigurationException("No dimension specified")
        if require_io and not config.graph:
            raise InvalidConfigurationException("No input graph specified")
        if not config.method:
            raise InvalidConfigurationException("Invalid method, choose one of: glove, pglove")
        if config.bca is None or not (config.bca.alpha > 0 and config.bca.epsilon > 0):
            raise InvalidConfigurationException("Invalid BCA parameters, alpha and epsilon are mandatory")
        if require_io and not (config.output and any(k in config.output for k in ("uri", "blank", "predicate", "literal"))):
            raise InvalidConfigurationException("Invalid output parameters, specify at least one group")

    def to_native(self) -> GeConfig:
        m = (self.method or "").lower()
        if m not in _lib.METHOD:
            raise ValueError("Invalid cost function")             # Main.java:111-112
        om = (self.opt.method if self.opt else "adagrad").lower()
        if om not in _lib.OPT:
            raise ValueError("Invalid optimization method")        # Main.java:122-123
        c = GeConfig()
        _lib.lib().ge_config_default(C.byref(c))
        c.dim, c.method, c.opt_method = self.dim, _lib.METHOD[m], _lib.OPT[om]
        c.directed = int(self.bca.directed)
        c.normalize = _lib.NORM[self.bca.getNormalize().lower()]
        c.alpha, c.epsilon = self.bca.alpha, self.bca.epsilon
        c.tolerance = self.opt.tolerance if self.opt else 1e-4
        c.maxiter = self.opt.maxiter if self.opt else 1000
        c.seed = self.seed
        c.math, c.order, c.init = _lib.MATH[self.math], _lib.ORDER[self.order], _lib.INIT[self.init]
        c.serial, c.device = int(self.serial), self.device
        c.update = _lib.UPDATE[self.update]
        return c


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class _Native:
    """Owns one ge_ctx."""

    def __init__(self, config: Configuration):
        self.L = _lib.lib()
        self.config = config
        h = C.c_void_p()
        cfg = config.to_native()
        rc = self.L.ge_create(C.byref(cfg), C.byref(h))
        if rc:
            raise GraphEmbError(rc, self.L.ge_last_error(None).decode())
        self.h = h

    def ck(self, rc):
        if rc < 0:
            raise GraphEmbError(rc, self.L.ge_last_error(self.h).decode())
        return rc

    def set_stream(self, cuda_stream: int | None):
        self.ck(self.L.ge_set_stream(self.h, C.c_void_p(cuda_stream or 0)))

    def close(self):
        if getattr(self, "h", None):
            self.L.ge_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CoOccurrenceMatrix:
    """J/util/CoOccurrenceMatrix.java:6-16."""

    def vocabSize(self) -> int: raise NotImplementedError
    def max(self) -> float: raise NotImplementedError
    def coOccurrenceCount(self) -> int: raise NotImplementedError
    def cIdx_I(self, i: int) -> int: raise NotImplementedError
    def cIdx_J(self, j: int) -> int: raise NotImplementedError
    def cIdx_C(self, i: int) -> float: raise NotImplementedError
    def shuffle(self) -> None: raise NotImplementedError
    def getKey(self, index: int) -> str: raise NotImplementedError
    def getType(self, index: int) -> int: raise NotImplementedError


class BookmarkColoring(CoOccurrenceMatrix):
    """new BookmarkColoring(graph, config) — builds the COO on the GPU.  `graph` is the CSR
    hand-off (synth.CsrGraph or anything with the same fields), i.e. what the Java class gets
    from getIn/OutNeighborhoods + In/OutEdgeNeighborhoodAlgorithm (BookmarkColoring.java:49-52)."""

    def __init__(self, graph, config: Configuration, keys=None, types=None, rows=None):
        Configuration.check(config, require_io=False)
        self.config = config
        self._n = _Native(config)
        n, L = self._n, self._n.L
        g = graph
        if hasattr(g, "src"):       # raw edge list: adjacency flattening on the device (ge_set_graph_edges)
            arrs = [np.ascontiguousarray(g.src, np.int32), np.ascontiguousarray(g.dst, np.int32),
                    np.ascontiguousarray(g.w, np.float32)]
            n.ck(L.ge_set_graph_edges(n.h, g.V, len(arrs[0]), *map(_ptr, arrs)))
        else:
            arrs = [np.ascontiguousarray(g.out_off, np.int64), np.ascontiguousarray(g.out_nbr, np.int32),
                    np.ascontiguousarray(g.out_w, np.float32), np.ascontiguousarray(g.in_off, np.int64),
                    np.ascontiguousarray(g.in_nbr, np.int32), np.ascontiguousarray(g.in_w, np.float32)]
            n.ck(L.ge_set_graph_csr(n.h, g.V, *map(_ptr, arrs)))
        nnz, xmax = C.c_int64(0), C.c_double(0)
        if rows is None:
            n.ck(L.ge_cooc_build(n.h, C.byref(nnz), C.byref(xmax)))
        else:
            n.ck(L.ge_cooc_build_rows(n.h, rows[0], rows[1], C.byref(nnz), C.byref(xmax)))
        self._nnz, self._max, self._V = nnz.value, xmax.value, g.V
        self._keys, self._types = keys, types
        self._coo = None
        self._perm = None

    # --- CoOccurrenceMatrix ---
    def vocabSize(self): return self._V
    def max(self): return self._max
    def coOccurrenceCount(self): return self._nnz

    def _host(self):
        if self._coo is None:
            I = np.empty(self._nnz, np.int32); J = np.empty(self._nnz, np.int32); X = np.empty(self._nnz, np.float32)
            self._n.ck(self._n.L.ge_cooc_export(self._n.h, _ptr(I), _ptr(J), _ptr(X), self._nnz))
            self._coo = (I, J, X)
        return self._coo

    def coo(self):
        """(I, J, X) host copies in storage order (row-major by bookmark, discovery order)."""
        return self._host()

    def _p(self, k):
        return k if self._perm is None else int(self._perm[k])

    def cIdx_I(self, i): return int(self._host()[0][self._p(i)])
    def cIdx_J(self, j): return int(self._host()[1][self._p(j)])
    def cIdx_C(self, i): return float(self._host()[2][self._p(i)])

    def shuffle(self):
        """The native trainer owns the epoch order (ge_train_epoch shuffles); kept for interface
        parity: refreshes the host view of the current order."""
        pass

    def getKey(self, index): return str(index) if self._keys is None else self._keys[index]
    def getType(self, index): return 0 if self._types is None else int(self._types[index])

    def stats(self):
        pops, scanned, sec = C.c_int64(0), C.c_int64(0), C.c_double(0)
        self._n.ck(self._n.L.ge_cooc_stats(self._n.h, C.byref(pops), C.byref(scanned), C.byref(sec)))
        fetched, tier2 = C.c_int64(0), C.c_int64(0)
        self._n.ck(self._n.L.ge_cooc_fetched(self._n.h, C.byref(fetched), C.byref(tier2)))
        return {"pops": pops.value, "scanned": scanned.value, "fetched": fetched.value, "tier2_rows": tier2.value, "seconds": sec.value, "nnz": self._nnz}

    def graph_csr(self):
        """The hand-off CSR as it sits on the device (validation of ge_set_graph_edges):
        (out_off, out_nbr, out_w, in_off, in_nbr, in_w, flatten_seconds)."""
        n, L = self._n, self._n.L
        eo, ei, sec = C.c_int64(0), C.c_int64(0), C.c_double(0)
        n.ck(L.ge_graph_sizes(n.h, C.byref(eo), C.byref(ei), C.byref(sec)))
        a = [np.empty(self._V + 1, np.int64), np.empty(eo.value, np.int32), np.empty(eo.value, np.float32),
             np.empty(self._V + 1, np.int64), np.empty(ei.value, np.int32), np.empty(ei.value, np.float32)]
        n.ck(L.ge_graph_export(n.h, *map(_ptr, a)))
        return (*a, sec.value)


class ImportedMatrix(BookmarkColoring):
    """A CoOccurrenceMatrix fed from host arrays (ge_cooc_import) — validation and C5-style COO."""

    def __init__(self, V, I, J, X, config: Configuration, xmax: float = 0.0):
        Configuration.check(config, require_io=False)
        self.config = config
        self._n = _Native(config)
        I = np.ascontiguousarray(I, np.int32); J = np.ascontiguousarray(J, np.int32); X = np.ascontiguousarray(X, np.float32)
        self._n.ck(self._n.L.ge_cooc_import(self._n.h, V, len(I), _ptr(I), _ptr(J), _ptr(X), float(xmax)))
        self._nnz, self._V = len(I), V
        self._max = self._n.L.ge_cooc_max(self._n.h)
        self._keys = self._types = None
        self._coo = (I, J, X)
        self._perm = None


class Optimum:
    """J/opt/Optimum.java:9-41."""

    def __init__(self):
        self._finalCost = 0.0
        self._result = None
        self.costHistory = []

    def addIntermediaryResult(self, r): self.costHistory.append(r)
    def getResult(self): return self._result
    def setResult(self, r): self._result = r
    def getFinalCost(self): return self._finalCost
    def setFinalCost(self, c): self._finalCost = c


class IOptimizer:
    """J/opt/IOptimizer.java:6-11."""

    def optimize(self) -> Optimum: raise NotImplementedError
    def getName(self) -> str: raise NotImplementedError
    def extractResult(self) -> np.ndarray: raise NotImplementedError


class Adagrad(IOptimizer):
    """new Adagrad(coMatrix, config, costFunction) (J/opt/grad/Adagrad.java:19-34); the cost
    function (GloveCost / PGloveCost) is selected by config.method as in Main.createOptimizer."""

    def __init__(self, coMatrix: BookmarkColoring, config: Configuration, on_epoch=None):
        self.coMatrix, self.config, self.on_epoch = coMatrix, config, on_epoch
        self._n = coMatrix._n
        self._n.ck(self._n.L.ge_store_init(self._n.h))

    def getName(self): return "Adagrad"

    def epoch(self) -> float:
        cost = C.c_double(0)
        self._n.ck(self._n.L.ge_train_epoch(self._n.h, C.byref(cost)))
        return cost.value

    def optimize(self) -> Optimum:
        n, L = self._n, self._n.L
        opt = Optimum()
        user_cb = self.on_epoch

        def _cb(_u, epoch, cost, seconds):
            return int(bool(user_cb(epoch, cost, seconds))) if user_cb else 0

        cb = _lib.EPOCH_CB(_cb)
        fc, ne = C.c_double(0), C.c_int32(0)
        n.ck(L.ge_train(n.h, cb, None, C.byref(fc), C.byref(ne)))
        hist = np.zeros(max(ne.value, 1) + len(opt.costHistory) + 4096, np.float64)
        cnt = n.ck(L.ge_cost_history(n.h, _ptr(hist), len(hist)))
        for c in hist[:cnt][-ne.value:] if ne.value else []:
            opt.addIntermediaryResult(float(c))
        opt.setResult(self.extractResult())
        opt.setFinalCost(fc.value)
        return opt

    def extractResult(self) -> np.ndarray:
        out = np.empty((self.coMatrix.vocabSize(), self.config.dim), np.float64)
        self._n.ck(self._n.L.ge_extract(self._n.h, _ptr(out)))
        return out

    def extractResultF32(self) -> np.ndarray:
        out = np.empty((self.coMatrix.vocabSize(), self.config.dim), np.float32)
        self._n.ck(self._n.L.ge_extract_f32(self._n.h, _ptr(out)))
        return out

    def store(self, name) -> np.ndarray:
        which = {"focus": 0, "context": 1, "fBias": 2, "cBias": 3, "gsFocus": 4, "gsContext": 5,
                 "gsFBias": 6, "gsCBias": 7, "m1Focus": 8, "m1Context": 9, "m1FBias": 10, "m1CBias": 11}[name]
        V, d = self.coMatrix.vocabSize(), self.config.dim
        out = np.empty((V,) if (which >> 1) & 1 else (V, d), np.float32)
        self._n.ck(self._n.L.ge_store_export(self._n.h, which, _ptr(out)))
        return out

    def epoch_order(self, epoch) -> np.ndarray:
        out = np.empty(self.coMatrix.coOccurrenceCount(), np.int32)
        self._n.ck(self._n.L.ge_epoch_order(self._n.h, epoch, _ptr(out), len(out)))
        return out

    def stats(self):
        s, l, u = C.c_double(0), C.c_int64(0), C.c_int64(0)
        self._n.ck(self._n.L.ge_train_stats(self._n.h, C.byref(s), C.byref(l), C.byref(u)))
        k, sh, n, w = C.c_double(0), C.c_double(0), C.c_int32(0), C.c_int64(0)
        self._n.ck(self._n.L.ge_train_hub_info(self._n.h, C.byref(k), C.byref(sh), C.byref(n), C.byref(w)))
        return {"update_seconds": s.value, "launches": l.value, "updates": u.value,
                "hub": {"max_concurrency": k.value, "damped_share": sh.value, "capped_launches": n.value, "warp_cap": w.value}}

    def reset(self):
        """A fresh optimizer on the same matrix (new Adagrad(coMatrix, ...), Main.java:72): store
        re-initialised, nonzeros re-packed."""
        self._n.ck(self._n.L.ge_train_reset(self._n.h))
        self._n.ck(self._n.L.ge_store_init(self._n.h))


class Adam(Adagrad):
    """new Adam(coMatrix, config, costFunction) (J/opt/grad/Adam.java:55-66); same epoch loop
    (Optimizer.optimize), update kernel = adam_update.  store("gs*") are the second moments."""

    def getName(self): return "Adam"


class AMSGrad(Adagrad):
    """new AMSGrad(coMatrix, config, costFunction) (J/opt/grad/AMSGrad.java:72-83)."""

    def getName(self): return "AMSGrad"


def createOptimizer(config: Configuration, coMatrix: BookmarkColoring, on_epoch=None) -> IOptimizer:
    """Main.createOptimizer (J/Main.java:107-131).  The matrix handle was created with
    config.opt.method, which fixes the store layout; the class only names it."""
    if (config.method or "").lower() not in ("glove", "pglove"):
        raise ValueError("Invalid cost function")
    om = (config.opt.method if config.opt else "adagrad").lower()
    if om not in _lib.OPT:
        raise ValueError("Invalid optimization method")
    if coMatrix.config is not config and (coMatrix.config.opt.method if coMatrix.config.opt else "adagrad").lower() != om:
        raise ValueError("the matrix was built for a different opt.method")
    return {"adagrad": Adagrad, "adam": Adam, "amsgrad": AMSGrad}[om](coMatrix, config, on_epoch)
