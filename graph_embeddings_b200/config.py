"""Reader for the reference's ``*.config.yml`` files (dblp / onstage / saa), schema unchanged.

The reference loads them with SnakeYAML ``loadAs(stream, Configuration.class)``
(J/util/read/ConfigReader.java:15-21) into the JavaBean of J/util/config/Configuration.java.
The shipped files carry keys the bean has no setter for: ``bca.reverse``,
``bca.predicates``, ``similarity[].predicate`` and the similarity methods ``token`` / ``jaccard``.
This reader accepts the UNION (bean properties + those extras), ignores what the hot path does
not read, and rejects anything else the way a strict bean loader would.  Scalars are coerced
through the bean's property types (``alpha: 1e-1`` is a YAML *string* without a dot; the bean's
``double`` setter parses it), enums through ``valueOf(x.toUpperCase())``
(Configuration.java:79-81,279-281,327-329,371-373).

Host logic only: no compute, no device.
"""
from __future__ import annotations

import os

import yaml

from .api import BCA, Configuration, InvalidConfigurationException, Opt

# bean properties of Configuration (:35-67) + nested beans; values = coercion
_TOP = {"graph", "method", "dim", "threads", "weights", "similarity", "bca", "opt", "pca", "output"}
_BCA = {"alpha": float, "epsilon": float, "directed": bool, "normalize": str,
        "reverse": bool, "predicates": bool}          # last two: stale keys of the shipped files, ignored
_OPT = {"method": str, "tolerance": float, "maxiter": int}
_SIM = {"sourcePredicate", "targetPredicate", "predicate", "method", "threshold", "ngram", "smooth",
        "pattern", "time", "offset", "distance", "thresholdDistance"}
_OUT = {"name", "uri", "blank", "predicate", "literal"}
_METHODS = {"GLOVE", "PGLOVE"}                          # EmbeddingMethod (:19-21)
_OPT_METHODS = {"ADAGRAD", "ADAM", "AMSGRAD"}           # OptimizationMethod (:23-25)
_NORMS = {"NONE", "UNITY", "COUNTS"}                    # BCANormalization (:31-33)


def _bool(v):
    if isinstance(v, bool):
        return v
    if isinstance(v, str) and v.lower() in ("true", "false"):
        return v.lower() == "true"
    raise InvalidConfigurationException(f"expected a boolean, got {v!r}")


def _coerce(kind, v, where):
    try:
        if kind is bool:
            return _bool(v)
        if kind is int:
            if isinstance(v, bool):
                raise ValueError
            if isinstance(v, float) and v != int(v):
                raise ValueError
            return int(v)
        if kind is float:
            if isinstance(v, bool):
                raise ValueError
            return float(v)
        return str(v)
    except (TypeError, ValueError):
        raise InvalidConfigurationException(f"{where}: cannot convert {v!r} to {kind.__name__}") from None


def _bean(d, allowed, where):
    if d is None:
        return {}
    if not isinstance(d, dict):
        raise InvalidConfigurationException(f"{where}: expected a mapping")
    unknown = set(d) - set(allowed)
    if unknown:
        # SnakeYAML: "Unable to find property 'x' on class ..."
        raise InvalidConfigurationException(f"{where}: unable to find property {sorted(unknown)[0]!r}")
    return d


def parse_config(text: str, check: bool = True, **native) -> Configuration:
    """YAML text -> Configuration.  `native` = knobs outside the reference schema (seed, math,
    order, init, update, serial, device)."""
    doc = yaml.safe_load(text)
    doc = _bean(doc, _TOP, "Configuration")
    b = _bean(doc.get("bca"), _BCA, "bca")
    o = _bean(doc.get("opt"), _OPT, "opt")
    for k, s in enumerate(doc.get("similarity") or []):
        _bean(s, _SIM, f"similarity[{k}]")
    _bean(doc.get("output"), _OUT, "output")

    bca = None
    if doc.get("bca") is not None:
        bb = {k: _coerce(_BCA[k], v, f"bca.{k}") for k, v in b.items()}
        norm = bb.get("normalize")
        if norm is not None and norm.upper() not in _NORMS:
            raise InvalidConfigurationException(f"bca.normalize: no enum constant {norm.upper()}")
        bca = BCA(alpha=bb.get("alpha", 0.0), epsilon=bb.get("epsilon", 0.0), directed=bb.get("directed", False),
                  normalize=None if norm is None else norm.lower(), reverse=bb.get("reverse"),
                  predicates=bb.get("predicates"))
    opt = None
    if doc.get("opt") is not None:
        oo = {k: _coerce(_OPT[k], v, f"opt.{k}") for k, v in o.items()}
        om = oo.get("method", "adagrad")
        if om.upper() not in _OPT_METHODS:
            raise InvalidConfigurationException(f"opt.method: no enum constant {om.upper()}")
        opt = Opt(method=om.lower(), tolerance=oo.get("tolerance", 0.0), maxiter=oo.get("maxiter", 0))
    method = doc.get("method")
    if method is not None and str(method).upper() not in _METHODS:
        raise InvalidConfigurationException("Invalid method, choose one of: glove, pglove")
    cfg = Configuration(graph=doc.get("graph"), method=None if method is None else str(method).lower(),
                        dim=_coerce(int, doc.get("dim", 0), "dim"),
                        threads=_coerce(int, doc.get("threads", 0), "threads"),
                        bca=bca, opt=opt, output=doc.get("output"), weights=doc.get("weights"),
                        similarity=doc.get("similarity"), pca=doc.get("pca"), **native)
    if cfg.threads == 0:                                   # Configuration.getThreads :71-73
        cfg.threads = max(1, (os.cpu_count() or 2) - 1)
    if check:
        Configuration.check(cfg)                           # :478-494
    return cfg


def load_config(path: str, check: bool = True, **native) -> Configuration:
    """ConfigReader.load(file) (J/util/read/ConfigReader.java:15-21)."""
    with open(path, "r", encoding="utf-8") as f:
        return parse_config(f.read(), check=check, **native)
