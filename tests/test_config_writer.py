"""Host logic around the hot path: *.config.yml reader (union schema), output file
naming (Main.createFileName), the '#' settings header and the "%11.6E" value format of
EmbeddingTextWriter — checked against an independent Python statement of Java's rules."""
import os
from decimal import ROUND_HALF_UP, Decimal

import numpy as np
import pytest

from graph_embeddings_b200 import InvalidConfigurationException, createFileName, parse_config
from graph_embeddings_b200.writer import config_header, format_e, java_double, java_float

DBLP = """
graph: data/DBLP/dblp-2015-2017.ttl
method: pglove
dim: 300
#threads: 1
weights:
  http://xmlns.com/foaf/0.1/name: 1
  http://purl.org/dc/terms/references: 1
similarity:
  - predicate: http://xmlns.com/foaf/0.1/name
    method: jarowinkler
    threshold: 0.95
  - predicate: http://purl.org/dc/elements/1.1/title
    method: token
    threshold: 0.5
bca:
  alpha: 1e-1
  epsilon: 1e-3
  reverse: true
  directed: true
  predicates: true
opt:
  method: adagrad
  tolerance: 1e-4
  maxiter: 1000
pca:
  variance: 0.95
output:
  uri: [ http://dblp.l3s.de/d2r/resource/authors/ ]
"""


def java_format_e(v: float) -> str:
    """java.util.Formatter %11.6E: shortest round-trip digits, rounded HALF_UP to 6 decimals."""
    d = Decimal(repr(float(v)))
    if d == 0:
        return "0.000000E+00"
    sign, dig, exp = d.as_tuple()
    e10 = exp + len(dig) - 1
    m = Decimal((0, dig, -(len(dig) - 1))).quantize(Decimal("0.000001"), rounding=ROUND_HALF_UP)
    if m >= 10:
        m, e10 = (m / 10).quantize(Decimal("0.000001")), e10 + 1
    return f"{'-' if sign else ''}{m}E{'-' if e10 < 0 else '+'}{abs(e10):02d}".rjust(11)


def test_shipped_schema_is_accepted_and_ignored_keys_are_ignored():
    c = parse_config(DBLP)
    assert (c.method, c.dim, c.graph) == ("pglove", 300, "data/DBLP/dblp-2015-2017.ttl")
    assert c.bca.alpha == 0.1 and c.bca.epsilon == 1e-3 and c.bca.directed is True
    assert c.bca.reverse is True and c.bca.predicates is True and c.bca.getNormalize() == "none"
    assert (c.opt.method, c.opt.tolerance, c.opt.maxiter) == ("adagrad", 1e-4, 1000)
    assert c.threads == max(1, (os.cpu_count() or 2) - 1)            # Configuration.getThreads :71-73
    assert c.similarity[1]["method"] == "token"                       # not in SimilarityMethod: upstream only


@pytest.mark.parametrize("opt", ["adagrad", "adam", "amsgrad", "AdaGrad"])
def test_enums_are_case_insensitive(opt):
    c = parse_config(DBLP.replace("method: adagrad", f"method: {opt}").replace("method: pglove", "method: GloVe"))
    assert c.opt.method == opt.lower() and c.method == "glove"


@pytest.mark.parametrize("edit,msg", [
    (("dim: 300", "dim: 0"), "No dimension specified"),
    (("graph: data/DBLP/dblp-2015-2017.ttl", ""), "No input graph specified"),
    (("method: pglove", "method: word2vec"), "Invalid method"),
    (("  alpha: 1e-1", "  alpha: 0"), "Invalid BCA parameters"),
    (("  uri: [ http://dblp.l3s.de/d2r/resource/authors/ ]", "  name: x"), "Invalid output parameters"),
    (("dim: 300", "dim: 300\nwindow: 10"), "unable to find property 'window'"),
    (("  directed: true", "  directed: true\n  hops: 3"), "unable to find property 'hops'"),
    (("  maxiter: 1000", "  maxiter: lots"), "cannot convert"),
    (("  method: adagrad", "  method: sgd"), "no enum constant SGD"),
])
def test_rejections_mirror_the_reference(edit, msg):
    with pytest.raises(InvalidConfigurationException) as e:
        parse_config(DBLP.replace(*edit))
    assert msg in str(e.value)


def test_file_name_and_header_follow_java_string_conversions():
    c = parse_config(DBLP)
    assert createFileName(c) == "dblp-2015-2017_pglove_partial_directed_0.1_0.001_adagrad_pca_300"
    h = config_header(c).splitlines()
    assert h[0] == "# Starting the embedding creation process with following settings:"
    assert "# BCA Alpha: 0.1" in h and "# BCA Epsilon: 0.001" in h and "# BCA Directed: true" in h
    assert "# pglove Tolerance: 1.0E-4" in h                          # Double.toString(1e-4)
    assert "# PCA Minimum Variance: 0.95" in h
    assert "# http://xmlns.com/foaf/0.1/name: 1.0" in h               # Map<String, Float>
    assert "# http://xmlns.com/foaf/0.1/name -> http://xmlns.com/foaf/0.1/name" in h
    assert " method:jarowinkler, threshold: 0.95" in h
    c2 = parse_config("graph: g.nt\nmethod: glove\ndim: 50\nbca: {alpha: 0.25, epsilon: 0.00001}\n"
                      "opt: {method: adam, tolerance: 0, maxiter: 5}\noutput: {uri: []}\n")
    assert createFileName(c2) == "g_glove_exact_undirected_0.25_1.0E-5_adam_50"
    assert "# No PCA will be performed" in config_header(c2)


def test_double_and_float_to_string():
    for x, s in [(0.1, "0.1"), (1e-3, "0.001"), (9.999e-4, "9.999E-4"), (1e-4, "1.0E-4"), (1.0, "1.0"), (1e7, "1.0E7"),
                 (9999999.0, "9999999.0"), (12345678.9, "1.23456789E7"), (0.0, "0.0"), (-2.5e-5, "-2.5E-5"), (100.0, "100.0")]:
        assert java_double(x) == s, (x, java_double(x), s)
    assert java_float(0.1) == "0.1" and java_float(1) == "1.0" and java_float(np.float32(0.95)) == "0.95"


def test_format_e_matches_java_formatter_rules():
    rng = np.random.default_rng(0)
    vals = [0.0, 1.0, -1.0, 0.5, 1.2345675, 0.12345675, 1.2345665, 9.9999995, 9.9999994, -9.9999995e-5, 1e-10, 123456.75,
            1e100, 2.5e-300, float(np.float32(0.1)), float(np.float32(-3.3333333e-3))]
    vals += [float(np.float32(x)) for x in rng.normal(0, 1, 20000)]
    vals += [float(np.float32(x)) for x in rng.normal(0, 1e-3, 5000)]
    vals += [float(x) for x in rng.normal(0, 1, 5000)]
    vals += [k / 1e7 + 5e-8 for k in range(10000000, 10002000)]        # decimal ties at the 7th digit
    for v in vals:
        assert format_e(v) == java_format_e(v), (repr(v), format_e(v), java_format_e(v))
    assert format_e(float("nan")).strip() == "NaN" and format_e(float("inf")).strip() == "Infinity"
