"""Host mirror of the reference's output stage, served from the device store.

    writer = EmbeddingTextWriter(createFileName(config), config)     # J/Main.java:76, :80-105
    writer.write(optimizer, bca, out_dir)                            # J/util/write/EmbeddingTextWriter.java:73-149

The row formatting and file writing happen in libgraphemb.so (ge_write_embeddings); this module
only builds the '#' settings block (writeConfig, :40-70) and the file names — string glue, no
compute.  Java's implicit double/float -> String conversions are reproduced by java_double /
java_float (Double.toString switches to computerized scientific notation outside [1e-3, 1e7)).
"""
from __future__ import annotations

import ctypes as C
import os
from decimal import Decimal

import numpy as np

from . import _lib

NODE_TYPES = ("URI", "BLANK", "LITERAL")        # NodeInfo ids 0, 1, 2 (J/convert/util/NodeInfo.java:8-34)


def _java_decimal(r: str) -> str:
    """shortest round-trip digits (Python repr) -> Double.toString / Float.toString layout"""
    if r in ("nan", "inf", "-inf"):
        return {"nan": "NaN", "inf": "Infinity", "-inf": "-Infinity"}[r]
    sign, dig, exp = Decimal(r).as_tuple()
    sgn = "-" if sign else ""
    digits = "".join(map(str, dig)).lstrip("0")
    if not digits:
        return sgn + "0.0"
    stripped = digits.rstrip("0")
    exp += len(digits) - len(stripped)
    digits = stripped
    e10 = exp + len(digits) - 1                            # decimal exponent of the leading digit
    if -3 <= e10 < 7:                                      # plain decimal, at least one fraction digit
        if e10 >= 0:
            whole = digits[:e10 + 1].ljust(e10 + 1, "0")
            frac = digits[e10 + 1:] or "0"
        else:
            whole, frac = "0", "0" * (-e10 - 1) + digits
        return f"{sgn}{whole}.{frac}"
    return f"{sgn}{digits[0]}.{digits[1:] or '0'}E{e10}"   # computerized scientific notation


def java_double(x: float) -> str:
    """Double.toString(x)"""
    return _java_decimal(repr(float(x)))


def java_float(x: float) -> str:
    """Float.toString(x)"""
    return _java_decimal(str(np.float32(x)))


def createFileName(config) -> str:
    """Main.createFileName (J/Main.java:80-105)."""
    name = os.path.basename(config.graph or "graph").lower()
    if "." in name:
        name = name[:name.rindex(".")]
    name += "_" + config.method.lower()
    name += "_partial" if config.similarity else "_exact"
    name += "_directed" if config.bca.directed else "_undirected"
    name += "_" + java_double(config.bca.alpha) + "_" + java_double(config.bca.epsilon)
    name += "_" + config.opt.method
    name += ("_pca_" if config.pca is not None else "_") + str(config.dim)
    return name


def _similarity_line(s: dict) -> str:
    """Configuration.SimilarityGroup.toString (Configuration.java:230-241)."""
    src = s.get("sourcePredicate", s.get("predicate"))
    dst = s.get("targetPredicate", s.get("predicate"))
    out = f"{src} -> {dst}\n method:{s.get('method')}, threshold: {java_double(s.get('threshold', 0.0))}"
    m = str(s.get("method", "")).upper()
    if m in ("NGRAM_COSINE", "NGRAM_JACCARD"):
        return out + f", ngram: {s.get('ngram') or 3}"
    if m == "NUMERIC":
        return out + f", smooth: {java_double(s.get('smooth') or 1.0)}"
    if m in ("DATE_DAYS", "DATE_MONTHS", "DATE_YEARS"):
        return out + f", pattern:{s.get('pattern') or 'iso'}, smooth: {java_double(s.get('smooth') or 1.0)}, time: {s.get('time') or 'bidirectional'}"
    return out


def config_header(config) -> str:
    """EmbeddingTextWriter.writeConfig (:40-70)."""
    L = ["# Starting the embedding creation process with following settings:",
         f"# Graph File: {config.graph}",
         f"# Embedding dimensions: {config.dim}",
         f"# Threads: {config.threads}",
         f"# BCA Alpha: {java_double(config.bca.alpha)}",
         f"# BCA Epsilon: {java_double(config.bca.epsilon)}",
         f"# BCA Directed: {'true' if config.bca.directed else 'false'}",
         f"# BCA normalize: {config.bca.getNormalize()}",
         f"# Gradient Descent Algorithm: {config.opt.method}",
         f"# {config.method} Tolerance: {java_double(config.opt.tolerance)}",
         f"# {config.method} Maximum Iterations: {config.opt.maxiter}"]
    if config.pca is not None:
        L.append(f"# PCA Minimum Variance: {java_double((config.pca or {}).get('variance', 0.0))}")
    else:
        L.append("# No PCA will be performed")
    if config.weights:
        L.append("# Using weights, predicates that are not listed are ignored:")
        L += [f"# {k}: {java_float(v)}" for k, v in config.weights.items()]
    else:
        L.append("# No weights specified, using linear weight")
    if config.similarity:
        L.append("# Using the following similarity metrics:")
        L += ["# " + _similarity_line(s) for s in config.similarity]
    else:
        L.append("# No similarity matching will be performed")
    return "\n".join(L) + "\n"


def format_e(v: float) -> str:
    """String.format("%11.6E", v) through the library's host-side formatter (no device needed)."""
    buf = C.create_string_buffer(48)
    n = _lib.lib().ge_format_e(float(v), buf, 48)
    assert n > 0
    return buf.value.decode()


class EmbeddingTextWriter:
    """J/util/write/EmbeddingTextWriter.java: <name>.vectors.tsv + <name>.dict.tsv."""

    def __init__(self, fileName: str, config):
        self.VECTORS_FILE = fileName + ".vectors.tsv"
        self.DICT_FILE = fileName + ".dict.tsv"
        self.config = config
        out = config.output or {}
        self.writeNodeTypes = [out.get("uri") is not None, out.get("blank") is not None, out.get("literal") is not None]

    def write(self, optimizer, coMatrix, outputFolder: str) -> int:
        """`optimizer` owns the device store (the reference passes the Optimum's double[] instead;
        the rows written are the same (focus+context)/2 values).  Returns the vertices emitted."""
        os.makedirs(outputFolder, exist_ok=True)
        n = optimizer._n
        V = coMatrix.vocabSize()
        keys = types = None
        keep = []
        if coMatrix._keys is not None:
            enc = [str(k).encode("utf-8") for k in coMatrix._keys]
            keep.append(enc)
            keys = (C.c_char_p * V)(*enc)
        if coMatrix._types is not None:
            types = np.ascontiguousarray(coMatrix._types, np.uint8)
        enabled = np.array(self.writeNodeTypes, np.uint8)
        out = self.config.output or {}
        plist = [list(out.get(k) or []) for k in ("uri", "blank", "literal")]
        flat = [p.encode("utf-8") for l in plist for p in l]
        nprefix = np.array([len(l) for l in plist], np.int32)
        parr = (C.c_char_p * max(1, len(flat)))(*flat)
        written = C.c_int64(0)
        n.ck(n.L.ge_write_embeddings(
            n.h, os.path.join(outputFolder, self.VECTORS_FILE).encode(), os.path.join(outputFolder, self.DICT_FILE).encode(),
            config_header(self.config).encode("utf-8"), keys, None if types is None else types.ctypes.data_as(C.c_void_p),
            enabled.ctypes.data_as(C.c_void_p), parr, nprefix.ctypes.data_as(C.c_void_p), C.byref(written)))
        return written.value
