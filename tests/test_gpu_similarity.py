"""Similarity edges on the device against the CPU oracle, through the C ABI:
same pairs in the same order; similarities bit-equal for the profile and string metrics (integer work
plus a few correctly rounded IEEE operations), within one f32 ulp for NUMERIC / DATE_* (Math.pow vs CUDA pow)."""
import numpy as np
import pytest

from graph_embeddings_b200 import BCA, Configuration, Opt, similarity as S, synth
from graph_embeddings_b200.api import _Native
from oracle import similarity as O

pytestmark = pytest.mark.gpu

XI = "^^http://www.w3.org/2001/XMLSchema#integer"
XD = "^^http://www.w3.org/2001/XMLSchema#date"


@pytest.fixture(scope="module")
def native():
    n = _Native(Configuration(method="glove", dim=8, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, 1)))
    yield n
    n.close()


def name_labels(rng, n):
    first = ["anna", "anne", "hans", "johan", "johann", "maria", "marie", "pieter", "peter", "the", "of"]
    last = ["de vries", "devries", "jansen", "janssen", "van dijk", "van  dyk", "bakker", "visser", "a", ""]
    out = []
    for _ in range(n):
        s = rng.choice(first) + " " + rng.choice(last)
        if rng.random() < 0.3:                                       # a typo
            k = rng.integers(0, len(s) + 1)
            s = s[:k] + rng.choice(list("aeio \t")) + s[k:]
        if rng.random() < 0.02:
            s = rng.choice(["", "a", "ab", "\U0001F600x"])
        out.append(s)
    return out


def run_group(native, cfg, labels, vert, src_items, dst_items):
    g = S.CompareGroup(cfg)
    vert = np.asarray(vert)
    for k in src_items:
        g.addToSource(int(vert[k]))
    for k in dst_items:
        g.addToTarget(int(vert[k]))
    lab = {int(v): labels[k] for k, v in enumerate(vert)}
    got = g.compare(native, lab)
    item = {int(v): k for k, v in enumerate(vert)}
    exp = O.pairs(g.method, labels, vert, [item[v] for v in g.source], [item[v] for v in g.target], g.threshold,
                  ngram=g.ngram, smooth=g.smooth, distance=g.distance, time=g.time.lower(), upper_triangle=g.upperTriangle)
    return got, exp, g


def check(got, exp, exact):
    assert got[0].tolist() == exp[0].tolist() and got[1].tolist() == exp[1].tolist()
    if exact:
        assert np.array_equal(got[2].view(np.uint32), exp[2].view(np.uint32))
    else:
        ulp = np.abs(got[2].view(np.int32).astype(np.int64) - exp[2].view(np.int32).astype(np.int64))
        assert ulp.max(initial=0) <= 1


@pytest.mark.parametrize("method,thr", [("NGRAM_JACCARD", 0.45), ("NGRAM_COSINE", 0.6), ("TOKEN_JACCARD", 0.3), ("TOKEN_COSINE", 0.45)])
@pytest.mark.parametrize("upper", [True, False])
def test_profile_metrics_match_oracle(native, method, thr, upper):
    rng = np.random.default_rng(3)
    n = 700
    labels = name_labels(rng, n)
    vert = rng.permutation(5000)[:n] + 10
    cfg = {"sourcePredicate": "p", "targetPredicate": "p" if upper else "q", "method": method, "threshold": thr, "ngram": 3}
    if upper:
        src = dst = range(n)
    else:
        src, dst = range(0, 400), range(300, n)             # overlapping sets: the own-vertex skip matters
    got, exp, _ = run_group(native, cfg, labels, vert, src, dst)
    assert len(exp[0]) > 100
    check(got, exp, exact=True)


def test_many_segments_and_scan_tiles(native):
    """|target| > 1024 (several segments per row) and rows x segments > 1024 (several scan tiles)"""
    rng = np.random.default_rng(5)
    n = 2600
    labels = name_labels(rng, n)
    vert = np.arange(n) * 3 + 1
    cfg = {"predicate": "p", "method": "ngram_jaccard", "threshold": 0.55, "ngram": 2}
    got, exp, g = run_group(native, cfg, labels, vert, range(n), range(n))
    assert len(exp[0]) > 5000 and g.seconds > 0
    check(got, exp, exact=True)


def test_threshold_zero_keeps_everything_but_nan(native):
    labels = ["abcd", "abce", "ab", "a", "abcd", ""]
    vert = [1, 2, 3, 4, 5, 6]
    cfg = {"sourcePredicate": "p", "targetPredicate": "q", "method": "NGRAM_JACCARD", "threshold": 0.0}
    got, exp, _ = run_group(native, cfg, labels, vert, range(6), range(6))
    check(got, exp, exact=True)
    pairs = set(zip(got[0].tolist(), got[1].tolist()))
    assert (1, 5) in pairs and (1, 3) in pairs and (3, 4) not in pairs      # equal strings 1; 0/2 = 0 kept; 0/0 NaN dropped


@pytest.mark.parametrize("upper", [True, False])
def test_numeric_matches_oracle_including_the_cut_quirk(native, upper):
    rng = np.random.default_rng(7)
    n = 900
    labels = []
    for _ in range(n):
        r = rng.random()
        v = int(rng.integers(-50, 3000))
        if r < 0.85:
            labels.append(str(v) + XI)
        elif r < 0.9:
            labels.append(str(v) + "^^xsd:int")            # same cut position, other suffix
        else:
            labels.append(rng.choice(["", "12a" + XI, "99999999999" + XI, "+5" + XI, "-" + XI, "12345" + XI]))
    # labels WITHOUT a datatype suffix: s1 is then parsed whole; as s2 they are cut at s1's '^' position,
    # and a label shorter than that position kills s1's whole job (uncaught substring exception)
    labels[50], labels[650], labels[700] = "7", "123", "55^x"
    vert = rng.permutation(4000)[:n]
    cfg = {"sourcePredicate": "p", "targetPredicate": "p" if upper else "q", "method": "numeric", "threshold": 0.3,
           "smooth": 0.5, "distance": 2.0}
    src, dst = (range(n), range(n)) if upper else (range(0, 500), range(400, n))
    got, exp, _ = run_group(native, cfg, labels, vert, src, dst)
    assert len(exp[0]) > 100
    check(got, exp, exact=False)
    if not upper:
        # target 650 is "123": every source whose own cut position is past 3 ("1234^^..", "-123^^..") loses its job
        lost = {int(vert[k]) for k in src if labels[k].find("^") >= 4}
        kept = {int(vert[k]) for k in src if 0 < labels[k].find("^") <= 3}
        assert lost and not lost & set(exp[0].tolist()) and kept & set(exp[0].tolist())


@pytest.mark.parametrize("method", ["DATE_DAYS", "DATE_MONTHS", "DATE_YEARS"])
@pytest.mark.parametrize("time", ["bidirectional", "backwards", "forwards"])
def test_dates_match_oracle(native, method, time):
    rng = np.random.default_rng(11)
    n = 600
    labels = []
    for _ in range(n):
        y, m, d = int(rng.integers(1995, 2006)), int(rng.integers(1, 13)), int(rng.integers(1, 32))
        s = "%04d%02d%02d" % (y, m, d)                      # some are impossible dates (strict -> 0)
        s += rng.choice(["", XD, "Z" + XD, "+0100", "junk"], p=[0.2, 0.6, 0.1, 0.05, 0.05])
        labels.append(s)
    vert = np.arange(n) + 100
    thr = {"DATE_DAYS": 0.12, "DATE_MONTHS": 0.3, "DATE_YEARS": 0.55}[method]
    cfg = {"predicate": "born", "method": method, "threshold": thr, "smooth": 0.5, "distance": 1.0, "time": time}
    got, exp, _ = run_group(native, cfg, labels, vert, range(n), range(n))
    assert len(exp[0]) > 50
    check(got, exp, exact=False)


@pytest.mark.parametrize("method,thr", [("JAROWINKLER", 0.9), ("LEVENSHTEIN", 0.75)])
@pytest.mark.parametrize("upper", [True, False])
def test_string_metrics_match_oracle(native, method, thr, upper):
    """character-level metrics: labels go to the device as UTF-16; long labels exercise the scratch rows"""
    rng = np.random.default_rng(17)
    n = 600
    labels = name_labels(rng, n)
    for k in range(0, n, 97):
        labels[k] = labels[k] * 9 + "x" * int(rng.integers(0, 40))          # up to ~200 code units
    vert = rng.permutation(3000)[:n]
    cfg = {"sourcePredicate": "p", "targetPredicate": "p" if upper else "q", "method": method, "threshold": thr}
    src, dst = (range(n), range(n)) if upper else (range(0, 350), range(250, n))
    got, exp, _ = run_group(native, cfg, labels, vert, src, dst)
    assert len(exp[0]) > 100
    check(got, exp, exact=True)


def test_empty_sides_and_errors(native):
    g = S.CompareGroup({"predicate": "p", "method": "token_cosine", "threshold": 0.5})
    v, o, s = g.compare(native, {})
    assert len(v) == 0
    g.addToSource(3)
    v, o, s = g.compare(native, {3: "only one"})
    assert len(v) == 0
    import ctypes as C
    from graph_embeddings_b200 import GraphEmbError, _lib
    bad = _lib.GeSimGroup()
    bad.method = 10                                          # past LEVENSHTEIN
    with pytest.raises(GraphEmbError):
        native.ck(native.L.ge_similarity_pairs(native.h, C.byref(bad), None, None))


def test_similarity_edges_feed_the_builder(native):
    """pairs -> two directed edges each -> device adjacency flattening -> co-occurrence build"""
    rng = np.random.default_rng(13)
    base = synth.rmat(400, 2400, 9)
    src = np.repeat(np.arange(base.V), np.diff(base.out_off)).astype(np.int32)
    edges = synth.EdgeList(base.V, src, base.out_nbr.astype(np.int32), base.out_w.astype(np.float32))
    lit = rng.permutation(base.V)[:120]
    labels = name_labels(rng, len(lit))
    got, exp, _ = run_group(native, {"predicate": "name", "method": "ngram_cosine", "threshold": 0.7}, labels, lit, range(120), range(120))
    check(got, exp, exact=True)
    more = S.add_similarity_edges(edges, *got)
    assert len(more.src) == len(edges.src) + 2 * len(got[0])
    k = len(edges.src)
    assert more.src[k] == got[0][0] and more.dst[k] == got[1][0] and more.src[k + 1] == got[1][0] and more.w[k + 1] == got[2][0]
    from graph_embeddings_b200 import BookmarkColoring
    c = Configuration(method="glove", dim=8, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0.0, 1))
    a, b = BookmarkColoring(edges, c), BookmarkColoring(more, c)
    assert b.coOccurrenceCount() >= a.coOccurrenceCount() and b.vocabSize() == a.vocabSize()
