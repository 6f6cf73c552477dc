"""ge_write_embeddings (EmbeddingTextWriter.write served from the device store) against a Python
restatement of J/util/write/EmbeddingTextWriter.java:73-149 fed with extractResult()."""
import os
from decimal import ROUND_HALF_UP, Decimal

import numpy as np
import pytest

from graph_embeddings_b200 import BookmarkColoring, EmbeddingTextWriter, createFileName, createOptimizer, parse_config, synth
from graph_embeddings_b200.writer import config_header

pytestmark = pytest.mark.gpu

CFG = """
graph: data/Test/Tiny.nt
method: glove
dim: {dim}
bca: {{alpha: 1e-1, epsilon: 1e-3, directed: true}}
opt: {{method: adagrad, tolerance: 0, maxiter: 2}}
output:
  uri: [ "http://a/", "http://b/" ]
  literal: []
"""


def java_format_e(v):
    d = Decimal(repr(float(v)))
    if d == 0:
        return "0.000000E+00"
    sign, dig, exp = d.as_tuple()
    e10 = exp + len(dig) - 1
    m = Decimal((0, dig, -(len(dig) - 1))).quantize(Decimal("0.000001"), rounding=ROUND_HALF_UP)
    if m >= 10:
        m, e10 = (m / 10).quantize(Decimal("0.000001")), e10 + 1
    return f"{'-' if sign else ''}{m}E{'-' if e10 < 0 else '+'}{abs(e10):02d}"


@pytest.mark.parametrize("dim", [7, 128])
def test_files_match_reference_writer(tmp_path, dim):
    g = synth.rmat(3000, 20000, 8)
    c = parse_config(CFG.format(dim=dim))
    rng = np.random.default_rng(1)
    types = rng.integers(0, 3, g.V).astype(np.uint8)                  # URI / BLANK / LITERAL
    keys = [("http://a/n%d" if i % 3 == 0 else "http://b/x\t%d" if i % 3 == 1 else "http://c/%d") % i for i in range(g.V)]
    bca = BookmarkColoring(g, c, keys=keys, types=types)
    opt = createOptimizer(c, bca)
    optimum = opt.optimize()
    name = createFileName(c)
    assert name == "tiny_glove_exact_directed_0.1_0.001_adagrad_%d" % dim
    n = EmbeddingTextWriter(name, c).write(opt, bca, str(tmp_path))

    # --- reference restatement (EmbeddingTextWriter.java:94-147) on Optimum.getResult()
    res = optimum.getResult()
    hdr = config_header(c)
    vect, dic = [hdr], [hdr, "key\ttype\n"]
    enabled = [True, False, True]                                     # uri given, blank absent, literal given (empty list)
    prefixes = [["http://a/", "http://b/"], [], []]
    emitted = 0
    for i in range(g.V):
        t = int(types[i])
        if not enabled[t]:
            continue
        if prefixes[t] and not any(keys[i].startswith(p) for p in prefixes[t]):
            continue
        vect.append("\t".join(java_format_e(v) for v in res[i]) + "\n")
        dic.append(keys[i].replace("\n", "").replace("\r", "").replace("\t", "") + "\t" + ("URI", "BLANK", "LITERAL")[t] + "\n")
        emitted += 1
    assert n == emitted and 0 < emitted < g.V
    with open(os.path.join(tmp_path, name + ".vectors.tsv"), "rb") as f:
        assert f.read() == "".join(vect).encode()
    with open(os.path.join(tmp_path, name + ".dict.tsv"), "rb") as f:
        assert f.read() == "".join(dic).encode()
