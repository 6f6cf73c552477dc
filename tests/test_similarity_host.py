"""Similarity-edge step, CPU side: the oracle restatement against cases derived
by hand from the Java source, and the product's host-side label -> numbers code against the oracle's
independent C++ versions.  No device work here (the compare itself is tested in test_gpu_similarity.py)."""
import math

import numpy as np
import pytest

from graph_embeddings_b200 import similarity as S
from oracle import similarity as O


def one(method, s1, s2, **kw):
    """metric.similarity(s1, s2) through the oracle's CompareJob loop (threshold -inf keeps everything but NaN)"""
    v, o, s = O.pairs(method, [s1, s2], [10, 11], [0], [1], -1e300, **kw)
    return None if len(s) == 0 else float(s[0])


# ---- profiles ------------------------------------------------------------------------------------
def test_ngram_profile_by_hand():
    # "ab  abc" -> whitespace run collapses -> "ab abc": shingles ab_, b_a, _ab, abc
    assert O.profile("NGRAM_JACCARD", "ab \t abc", 3) == {"ab ": 1, "b a": 1, " ab": 1, "abc": 1}
    assert O.profile("NGRAM_COSINE", "aaaa", 2) == {"aa": 3}
    assert O.profile("NGRAM_COSINE", "ab", 3) == {}                      # shorter than k: empty profile


def test_token_profile_by_hand():
    # stop words and one-letter tokens are dropped; the last token ends at the last character
    assert O.profile("TOKEN_JACCARD", "the quick brown fox a x fox") == {"quick": 1, "brown": 1, "fox": 2}
    assert O.profile("TOKEN_COSINE", "  lead  and trail  ") == {"lead": 1, "trail": 1}
    assert O.profile("TOKEN_COSINE", "x") == {}


@pytest.mark.parametrize("seed", [0, 1])
def test_host_profiles_match_oracle(seed):
    rng = np.random.default_rng(seed)
    alphabet = list("ab c\tdé中") + ["\U0001F600", "the", " of ", "  "]
    for _ in range(300):
        s = "".join(rng.choice(alphabet, size=rng.integers(0, 14)))
        for k in (1, 2, 3):
            assert S.ngram_profile(s, k) == {S._units(a): c for a, c in O.profile("NGRAM_JACCARD", s, k).items()}, repr(s)
        assert S.token_profile(s) == {S._units(a): c for a, c in O.profile("TOKEN_JACCARD", s).items()}, repr(s)


# ---- metrics, by hand ----------------------------------------------------------------------------
def test_jaccard_and_cosine_by_hand():
    # "abcd" {abc, bcd}, "abce" {abc, bce}: inter 1, union 3
    assert one("NGRAM_JACCARD", "abcd", "abce", ngram=3) == np.float32(1 / 3)
    # cosine: profiles {aa:2, ab:1} ("aaab") and {aa:1, ab:1, bb:1} ("aabb"): dot 3, norms sqrt5, sqrt3
    assert one("NGRAM_COSINE", "aaab", "aabb", ngram=2) == np.float32(3 / (math.sqrt(5) * math.sqrt(3)))
    assert one("NGRAM_COSINE", "ab", "abab", ngram=3) == 0.0            # s1 shorter than k -> 0 (:27-31)
    assert one("NGRAM_JACCARD", "ab", "a", ngram=3) is None             # 0/0 = NaN never passes a threshold
    assert one("NGRAM_JACCARD", "same", "same") == 1.0
    assert one("TOKEN_JACCARD", "red big car", "big blue car") == np.float32(2 / 4)
    assert one("TOKEN_COSINE", "fox fox dog", "fox cat") == np.float32(2 / (math.sqrt(5) * math.sqrt(2)))


def test_numeric_by_hand():
    x = "^^http://www.w3.org/2001/XMLSchema#integer"
    # |a-b| = 3, distance 1, alpha 0.5 -> (|3-1|+1)^(-0.5)
    assert one("NUMERIC", "10" + x, "13" + x, smooth=0.5, distance=1.0) == np.float32(3 ** -0.5)
    assert one("NUMERIC", "7" + x, "7" + x, smooth=0.5) == 1.0
    assert one("NUMERIC", "", "7" + x, smooth=0.5) == 0.0
    # Numeric.java:30 cuts s2 at S1's '^' position: s1 = "5^^.." cuts "123^^.." to "1" -> |5-1| = 4
    assert one("NUMERIC", "5" + x, "123" + x, smooth=0.0) == np.float32(1 / 5)
    # ... and the other way round s2 = "5^^" -> "5^^"[:3] is not a number -> 0
    assert one("NUMERIC", "123" + x, "5" + x, smooth=0.0) == 0.0
    assert one("NUMERIC", "12x" + x, "5" + x, smooth=0.0) == 0.0           # NumberFormatException -> 0
    assert one("NUMERIC", "2147483648", "5", smooth=0.0) == 0.0           # overflow is a NumberFormatException
    # int overflow of a - b wraps: 2147483647 - (-2) = -2147483647 -> abs 2147483647
    assert one("NUMERIC", "2147483647", "-2", smooth=0.0) == np.float32(1 / 2147483648.0)
    # s2 shorter than s1's cut: substring throws outside the try -> the whole job of s1 is lost
    v, o, s = O.pairs("NUMERIC", ["12345" + x, "7", "12346" + x], [1, 2, 3], [0], [2, 1], 0.0, smooth=0.0)
    assert len(v) == 0
    v, o, s = O.pairs("NUMERIC", ["12345" + x, "7", "12346" + x], [1, 2, 3], [0], [2], 0.0, smooth=0.0)
    assert o.tolist() == [3] and s[0] == np.float32(0.5)


def test_date_by_hand():
    x = "^^http://www.w3.org/2001/XMLSchema#date"
    kw = dict(smooth=0.0, distance=0.0)
    assert one("DATE_DAYS", "20200301" + x, "20200228" + x, **kw) == np.float32(1 / 3)      # leap year: 2 days
    assert one("DATE_MONTHS", "20200131" + x, "20200301" + x, **kw) == np.float32(1 / 2)    # 1 full month
    assert one("DATE_MONTHS", "20200301" + x, "20200131" + x, **kw) == np.float32(1 / 2)    # truncation toward zero
    assert one("DATE_YEARS", "20000615" + x, "20030614" + x, **kw) == np.float32(1 / 3)     # 2 full years
    assert one("DATE_DAYS", "20200230" + x, "20200228" + x, **kw) == 0.0                     # strict: no 30 Feb
    assert one("DATE_DAYS", "20200101Z", "20200103+0100", **kw) == np.float32(1 / 3)
    assert one("DATE_DAYS", "2020-01-01", "20200103", **kw) == 0.0
    assert one("DATE_DAYS", "20200105", "20200101", time="backwards", **kw) == 0.0          # d1 after d2
    assert one("DATE_DAYS", "20200101", "20200105", time="backwards", **kw) == np.float32(1 / 5)
    assert one("DATE_DAYS", "20200101", "20200105", time="forwards", **kw) == 0.0           # d1 before d2


def test_compare_job_loop_rules():
    labels = ["abcd", "abce", "abcd", "zzzz"]
    vert = [5, 6, 7, 5]                                  # item 3 is the same vertex as item 0
    # upper triangle: j starts at i+1; own vertex skipped; equal strings -> 1
    v, o, s = O.pairs("NGRAM_JACCARD", labels, vert, [0, 1, 2, 3], [0, 1, 2, 3], 0.3, upper_triangle=True)
    assert list(zip(v.tolist(), o.tolist())) == [(5, 6), (5, 7), (6, 7)]
    assert s.tolist() == [np.float32(1 / 3), 1.0, np.float32(1 / 3)]
    v, o, s = O.pairs("NGRAM_JACCARD", labels, vert, [0], [0, 1, 2, 3], 0.0)
    assert o.tolist() == [6, 7]                         # targets 0 and 3 are vertex 5 itself


# ---- host-side parsing against the oracle's ------------------------------------------------------
def test_parse_int_matches_oracle():
    for s in ["0", "-0", "+7", "2147483647", "2147483648", "-2147483648", "-2147483649", "", "-", "+", "12a", " 1", "1 ",
              "007", "1_0", "٣"]:
        assert S.parse_int(s) == O.parse_int(s), s


def test_parse_date_matches_oracle_and_patterns():
    for s in ["20200229", "20190229", "20201301", "20200100", "20200131Z", "20200131z", "20200131+01", "20200131+0130",
              "20200131-013015", "20200131+1", "20200131+1900", "20200131+1800", "20200131+1801", "2020013", "2020-01-31", "00010101"]:
        got = S.parse_local_date(s, "iso")
        exp = O.parse_iso(s)
        assert (got is None) == (exp is None), s
        if got is not None:
            assert (S.epoch_day(*got), (got[0] * 12 + got[1] - 1) * 32 + got[2]) == exp, s
    assert S.epoch_day(1970, 1, 1) == 0 and S.epoch_day(2000, 3, 1) == 11017
    assert S.parse_local_date("2021-02-30", "yyyy-MM-dd") == (2021, 2, 28)        # SMART clamps
    assert S.parse_local_date("2021-02-32", "yyyy-MM-dd") is None
    assert S.parse_local_date("31/1/99", "d/M/yy") == (2099, 1, 31)
    assert S.parse_local_date("20210315", "yyyyMMdd") == (2021, 3, 15)
    assert S.parse_local_date("on 2021.03.15", "'on' yyyy.MM.dd") == (2021, 3, 15)
    with pytest.raises(ValueError):
        S.parse_local_date("15 Mar 2021", "dd MMM yyyy")


def test_java_hashset_order():
    assert S.java_hashset_order([3, 1, 2, 1]) == [1, 2, 3]
    assert S.java_hashset_order([17, 1]) == [17, 1]                  # same bucket of a 16-table: insertion order
    assert S.java_hashset_order([33, 2, 1]) == [33, 1, 2]
    assert S.java_hashset_order(range(100, 0, -1)) == list(range(1, 101))   # 100 ints in a 256-table: ascending
    assert S.java_hashset_order([65536, 1, 0]) == [0, 65536, 1]      # h ^ (h >>> 16) puts 65536 in bucket 1, ahead of the later 1


def test_compare_group_config_defaults():
    g = S.CompareGroup({"sourcePredicate": "p", "targetPredicate": "p", "method": "ngram_jaccard", "threshold": 0.5})
    assert g.upperTriangle and g.ngram == 3 and g.smooth == 1.0 and g.pattern == "iso" and g.time == "BIDIRECTIONAL"
    g = S.CompareGroup({"sourcePredicate": "p", "targetPredicate": "q", "method": "DATE_DAYS", "smooth": 0, "time": "forwards"})
    assert not g.upperTriangle and g.smooth == 1.0 and g.time == "FORWARDS"
    assert S.CompareGroup({"predicate": "p", "method": "jarowinkler"}).method == "JAROWINKLER"
    with pytest.raises(ValueError):
        S.CompareGroup({"predicate": "p", "method": "soundex"})


def test_string_metrics_known_answers():
    """Jaro-Winkler / normalized Levenshtein of info.debatty java-string-similarity (not under
    /root/reference): the classic published values, in the library's float-then-double arithmetic."""
    f = np.float32
    m, t = f(6), 1                                           # MARTHA / MARHTA: 6 matches, 1 transposition, prefix 3
    j = float((m / f(6) + m / f(6) + (m - f(t)) / m) / f(3))
    assert one("JAROWINKLER", "MARTHA", "MARHTA") == f(j + min(0.1, 1 / 6) * 3 * (1 - j))
    assert abs(one("JAROWINKLER", "DWAYNE", "DUANE") - 0.84) < 1e-6
    assert abs(one("JAROWINKLER", "DIXON", "DICKSONX") - 0.813333) < 1e-6
    assert one("JAROWINKLER", "abc", "xyz") == 0.0 and one("JAROWINKLER", "", "xyz") == 0.0
    assert one("JAROWINKLER", "same", "same") == 1.0
    assert one("LEVENSHTEIN", "kitten", "sitting") == f(1 - 3 / 7)
    assert one("LEVENSHTEIN", "", "abc") == 0.0 and one("LEVENSHTEIN", "abc", "abc") == 1.0 and one("LEVENSHTEIN", "", "") is not None
    assert one("LEVENSHTEIN", "flaw", "lawn") == f(0.5)
