"""CPU oracle for the similarity-edge compare — TEST INFRASTRUCTURE ONLY (parity unpinned; see
similarity.cpp).  ctypes front-end for ``liboracle_sim.so``; only ``tests/`` may import it."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle_sim.so")

METHOD = {"NGRAM_COSINE": 0, "NGRAM_JACCARD": 1, "TOKEN_COSINE": 2, "TOKEN_JACCARD": 3, "NUMERIC": 4,
          "DATE_DAYS": 5, "DATE_MONTHS": 6, "DATE_YEARS": 7, "JAROWINKLER": 8, "LEVENSHTEIN": 9}
TIME = {"bidirectional": 0, "backwards": 1, "forwards": 2}

_lib = None


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "similarity.cpp")
        if not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle_sim.so"])
        L = C.CDLL(_LIB_PATH)
        p, i32, i64, f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
        L.orc_sim_pairs.argtypes = [i32, i32, f64, f64, f64, i32, i32, i32, p, p, p, i32, p, i32, p, i64, p, p, p]
        L.orc_sim_pairs.restype = i64
        L.orc_sim_parse_int.argtypes = [p, i32, p]
        L.orc_sim_parse_int.restype = i32
        L.orc_sim_parse_iso.argtypes = [p, i32, p, p]
        L.orc_sim_parse_iso.restype = i32
        L.orc_sim_profile.argtypes = [i32, i32, p, i32, i32, i32, p, p, p]
        L.orc_sim_profile.restype = i32
        _lib = L
    return _lib


def utf16(s: str) -> np.ndarray:
    """Java's view of a string: UTF-16 code units"""
    return np.frombuffer(s.encode("utf-16-le", "surrogatepass"), dtype=np.uint16).copy()


def _pack(labels):
    units = [utf16(s) for s in labels]
    off = np.zeros(len(labels) + 1, np.int64)
    np.cumsum([len(u) for u in units], out=off[1:])
    chars = np.concatenate(units) if units and off[-1] else np.zeros(1, np.uint16)
    return np.ascontiguousarray(chars, np.uint16), off


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def pairs(method, labels, vert, source, target, threshold, *, ngram=3, smooth=1.0, distance=0.0,
          time="bidirectional", upper_triangle=False):
    """CompareJob over every source literal: (vert, other, similarity f32) ordered by (source, target) index."""
    chars, off = _pack(labels)
    vert = np.ascontiguousarray(vert, np.int32)
    source = np.ascontiguousarray(source, np.int32)
    target = np.ascontiguousarray(target, np.int32)
    args = [METHOD[method.upper()], ngram, float(threshold), float(smooth), float(distance), TIME[time.lower()],
            1 if upper_triangle else 0, len(labels), _p(chars), _p(off), _p(vert), len(source), _p(source), len(target), _p(target)]
    cap = max(1, lib().orc_sim_pairs(*args, 0, None, None, None))
    v, o, s = np.zeros(cap, np.int32), np.zeros(cap, np.int32), np.zeros(cap, np.float32)
    n = lib().orc_sim_pairs(*args, cap, _p(v), _p(o), _p(s))
    return v[:n], o[:n], s[:n]


def parse_int(s: str):
    u = utf16(s) if s else np.zeros(1, np.uint16)
    out = C.c_int32(0)
    ok = lib().orc_sim_parse_int(_p(u), len(s.encode("utf-16-le", "surrogatepass")) // 2, C.byref(out))
    return out.value if ok else None


def parse_iso(s: str):
    u = utf16(s) if s else np.zeros(1, np.uint16)
    e, m = C.c_int64(0), C.c_int64(0)
    ok = lib().orc_sim_parse_iso(_p(u), len(s.encode("utf-16-le", "surrogatepass")) // 2, C.byref(e), C.byref(m))
    return (e.value, m.value) if ok else None


def profile(method, s: str, ngram=3) -> dict:
    u = utf16(s) if s else np.zeros(1, np.uint16)
    n = len(s.encode("utf-16-le", "surrogatepass")) // 2
    cap_e, cap_c = n + 2, (n + 2) * max(ngram, n + 1)
    kc, ko, cnt = np.zeros(cap_c, np.uint16), np.zeros(cap_e + 1, np.int32), np.zeros(cap_e, np.int32)
    k = lib().orc_sim_profile(METHOD[method.upper()], ngram, _p(u), n, cap_e, cap_c, _p(kc), _p(ko), _p(cnt))
    assert k >= 0
    return {kc[ko[i]:ko[i + 1]].tobytes().decode("utf-16-le", "surrogatepass"): int(cnt[i]) for i in range(k)}
