"""Hand-derived Bookmark-Coloring cases pinning the oracle's restatement of
J/bca/jobs/DirectedWeighted.java:31-101, UndirectedWeighted.java:31-114, J/bca/util/BCV.java.
Expected values are written out from the rules (fp64 paint, f32 accumulate in pop order)."""
import numpy as np

import oracle
from graph_embeddings_b200 import synth

f32 = np.float32
A, E = 0.1, 1e-3


def coo(c):
    i, j, x = c.sorted()
    return {(int(a), int(b)): f32(v) for a, b, v in zip(i, j, x)}


def test_isolated_node_directed():
    c = oracle.cooc_build(synth.tiny([], V=1), A, E, True)
    # forward pop: 0f + (float)(0.1*1.0); reverse the same; merge = Float::sum
    assert coo(c) == {(0, 0): f32(0.1) + f32(0.1)}
    assert c.nnz == 1 and c.pops == 2 and c.scanned == 0
    assert c.max == float(f32(0.1) + f32(0.1))


def test_single_edge_directed():
    c = coo(oracle.cooc_build(synth.tiny([(0, 1, 1.0)]), A, E, True))
    p1 = (1 - A) * 1.0 * (1.0 / 1.0)          # paint arriving at the neighbour
    assert c == {
        (0, 0): f32(0.1) + f32(0.1),         # fwd root + rev root
        (0, 1): f32(A * p1),                  # fwd only
        (1, 1): f32(0.1) + f32(0.1),
        (1, 0): f32(A * p1),                  # rev only (follows the in-edge)
    }


def _geometric(alpha, eps):
    """2-cycle / self-loop: the paint bounces with factor (1-alpha); returns the f32 sums at
    the root (even pops) and at the other node (odd pops)."""
    paint, k = 1.0, 0
    acc = [None, None]
    while True:
        v = f32(alpha * paint)
        acc[k & 1] = v if acc[k & 1] is None else f32(acc[k & 1] + v)
        if paint < eps:
            break
        nxt = (1 - alpha) * paint * (1.0 / 1.0)
        if nxt < eps:
            break
        paint, k = nxt, k + 1
    return acc, k + 1


def test_two_cycle_directed():
    c = oracle.cooc_build(synth.tiny([(0, 1, 1.0), (1, 0, 1.0)]), A, E, True)
    (root, other), pops = _geometric(A, E)
    assert pops == 66                         # 0.9^65 >= 1e-3 > 0.9^66
    got = coo(c)
    # forward and reverse passes are identical on a symmetric 2-cycle
    assert got == {(0, 0): f32(root + root), (0, 1): f32(other + other),
                   (1, 1): f32(root + root), (1, 0): f32(other + other)}
    assert c.pops == 4 * pops


def test_self_loop_directed():
    c = coo(oracle.cooc_build(synth.tiny([(0, 0, 1.0)]), A, E, True))
    # every pop is node 0: one f32 running sum over all 66 pops, forward == reverse
    paint, s = 1.0, None
    for _ in range(66):
        v = f32(A * paint)
        s = v if s is None else f32(s + v)
        paint = (1 - A) * paint * (1.0 / 1.0)
    assert c == {(0, 0): f32(s + s)}


def test_weighted_star_and_epsilon_cut():
    g = synth.tiny([(0, 1, 1.0), (0, 2, 0.1)])
    total = 1.0 + float(f32(0.1))             # fp64 sum of the f32 weights
    p1 = (1 - A) * 1.0 * (1.0 / total)
    p2 = (1 - A) * 1.0 * (float(f32(0.1)) / total)
    c = coo(oracle.cooc_build(g, A, E, True))
    assert c[(0, 1)] == f32(A * p1) and c[(0, 2)] == f32(A * p2)
    assert c[(1, 0)] == f32(A * 0.9) and c[(2, 0)] == f32(A * 0.9)   # reverse: single in-edge
    # with epsilon = 0.1 the light edge's share (0.0818) is dropped before insertion
    c = coo(oracle.cooc_build(g, A, 0.1, True))
    assert (0, 2) not in c and c[(0, 1)] == f32(A * p1)
    # epsilon > 1: the root itself does not distribute
    c = coo(oracle.cooc_build(g, A, 2.0, True))
    assert set(c) == {(0, 0), (1, 1), (2, 2)}


def test_undirected_follows_both_directions():
    # single edge 0->1, undirected: bookmark 1 reaches 0 through the in-edge, then bounces
    c = oracle.cooc_build(synth.tiny([(0, 1, 1.0)]), A, E, False)
    (root, other), pops = _geometric(A, E)
    assert coo(c) == {(0, 0): root, (0, 1): other, (1, 1): root, (1, 0): other}
    assert c.pops == 2 * pops


def test_undirected_node_in_both_lists_gets_out_then_in():
    # 0->1 (w 1) and 1->0 (w .5), undirected, eps=.25: from 0, node 1 is out- and
    # in-neighbour, so it receives two adds in the order out, in (UndirectedWeighted.java:73-111)
    g = synth.tiny([(0, 1, 1.0), (1, 0, 0.5)])
    eps = 0.25
    c = coo(oracle.cooc_build(g, A, eps, False))
    t0 = 1.0 + 0.5                                   # node 0: out {1: 1.0} then in {1: 0.5}
    t1 = 0.5 + 1.0                                   # node 1: out {0: 0.5} then in {0: 1.0}
    k = 1 - A
    n1 = k * 1.0 * (1.0 / t0) + k * 1.0 * (0.5 / t0)          # .6 then +.3, both >= eps
    n0 = k * n1 * (0.5 / t1) + k * n1 * (1.0 / t1)            # .27 then +.54
    n1b = k * n0 * (1.0 / t0)                                  # .486 ; in-share .243 < eps dropped
    assert k * n0 * (0.5 / t0) < eps <= n1b
    n0b = k * n1b * (1.0 / t1)                                 # .2916 ; out-share .1458 dropped
    assert k * n1b * (0.5 / t1) < eps <= n0b
    assert k * n0b * (1.0 / t0) < eps                          # nothing left to distribute
    assert c[(0, 0)] == f32(f32(f32(A * 1.0) + f32(A * n0)) + f32(A * n0b))
    assert c[(0, 1)] == f32(f32(A * n1) + f32(A * n1b))


def test_normalise_counts_and_unity():
    g = synth.tiny([(0, 1, 1.0), (0, 2, 0.1)])
    raw = oracle.cooc_build(g, A, E, True)
    r = coo(raw)
    row0 = [r[(0, 0)], r[(0, 1)], r[(0, 2)]]
    mx, mn = max(row0), min(row0)
    cnt = coo(oracle.cooc_build(g, A, E, True, "counts"))
    assert (0, 0) not in cnt                                  # root removed (BCV.java:58)
    for j in (1, 2):
        assert cnt[(0, j)] == f32(f32(r[(0, j)] / f32(f32(mx - mn) / f32(999.0))) + f32(1.0))
    un = oracle.cooc_build(g, A, E, True, "unity")
    u = coo(un)
    s = f32(r[(0, 1)] + r[(0, 2)])                            # discovery order: 1 then 2
    assert u[(0, 1)] == f32(f32(r[(0, 1)] / s) - f32(1e-6))
    assert u[(0, 2)] == f32(f32(r[(0, 2)] / s) - f32(1e-6))
    assert (0, 0) not in u
    # rows 1 and 2 keep one entry (the reverse hit on node 0)
    assert u[(1, 0)] == f32(f32(1.0) - f32(1e-6))


def test_empty_row_after_root_removal_sets_max_to_one():
    c = oracle.cooc_build(synth.tiny([], V=2), A, E, True, "unity")
    assert c.nnz == 0 and c.max == 1.0                        # BCV.max() of empty -> 1f


def test_thread_count_does_not_change_result():
    g = synth.rmat(3000, 20000, 9)
    a = oracle.cooc_build(g, threads=1)
    b = oracle.cooc_build(g, threads=4)
    assert np.array_equal(a.I, b.I) and np.array_equal(a.J, b.J) and np.array_equal(a.X, b.X)
    assert a.max == b.max and a.pops == b.pops
    # sample window == corresponding rows of the full build
    s = oracle.cooc_build(g, first=100, count=50)
    lo, hi = a.row_off[100], a.row_off[150]
    assert np.array_equal(s.J, a.J[lo:hi]) and np.array_equal(s.X, a.X[lo:hi])


# ---- independent re-derivation ------------------------------------------------------------------
# A second, pure-Python reading of the Java source (dict frontier, Python floats = fp64, numpy f32
# scalars for the BCV), written without looking at oracle.cpp: both restatements must agree bit for
# bit on random weighted graphs.  This is the strongest pin available without a JVM.
def _py_pass(g, bookmark, alpha, eps, mode):
    """mode 'out' / 'in': DirectedWeighted.doWork(graph, reverse = mode == 'in') (:31-101);
    mode 'both': UndirectedWeighted.doWork (:31-114)."""
    def nbrs(v, which):
        off, nbr, w = (g.out_off, g.out_nbr, g.out_w) if which == "out" else (g.in_off, g.in_nbr, g.in_w)
        return [(int(nbr[k]), f32(w[k])) for k in range(off[v], off[v + 1])]
    frontier = {bookmark: 1.0}                                   # TreeMap<Integer, PaintedNode>
    bcv = {}
    while frontier:
        focus = min(frontier)                                    # pollFirstEntry: smallest node id
        wet = frontier.pop(focus)
        bcv[focus] = f32(bcv.get(focus, f32(0)) + f32(alpha * wet))       # BCV.add(int, double) -> add(int, float)
        if wet < eps:
            continue
        if mode == "both":
            lists = [nbrs(focus, "out"), nbrs(focus, "in")]
        else:
            lists = [nbrs(focus, mode)]
            if not lists[0]:                                     # :66 (directed only)
                continue
        total = 0.0
        for lst in lists:
            for _, w in lst:
                total += float(w)                                # double += float
        if mode != "both" and total == 0:                        # :77 (directed only)
            continue
        for lst in lists:
            for n, w in lst:
                partial = (1 - alpha) * wet * (float(w) / total)
                if partial < eps:
                    continue
                frontier[n] = frontier[n] + partial if n in frontier else partial
    return bcv


def _py_cooc(g, alpha, eps, directed, normalize):
    out, gmax = {}, None
    for b in range(g.V):
        if directed:                                             # BCAJob.call (:32-36): forward, then merge(reverse)
            bcv = _py_pass(g, b, alpha, eps, "out")
            for k, v in _py_pass(g, b, alpha, eps, "in").items():
                bcv[k] = f32(bcv[k] + v) if k in bcv else v      # HashMap.merge(key, value, Float::sum)
        else:
            bcv = _py_pass(g, b, alpha, eps, "both")
        if normalize == "counts":                                # BCV.toCounts (:52-59), scale (:89-91)
            amax, amin = max(bcv.values()), min(bcv.values())
            bcv = {k: f32(f32(v / f32(f32(amax - amin) / f32(f32(1000) - f32(1)))) + f32(1)) for k, v in bcv.items()}
            del bcv[b]
        m = max(bcv.values()) if bcv else f32(1)                 # BCV.max(): orElse(1f)
        gmax = float(m) if gmax is None else max(gmax, float(m))
        for k, v in bcv.items():
            out[(b, k)] = v
    return out, gmax


def _random_graph(seed, V=70, E=420, zero_weight=False):
    rng = np.random.default_rng(seed)
    edges = {}
    while len(edges) < E:
        s, d = int(rng.integers(V)), int(rng.integers(V))
        edges.setdefault((s, d), float(f32(rng.choice([1.0, 0.1, 0.5, 0.95, 0.75]))))     # self loops allowed
    return synth.tiny([(s, d, w) for (s, d), w in edges.items()], V=V)


import pytest  # noqa: E402


@pytest.mark.parametrize("seed", [1, 2, 3])
@pytest.mark.parametrize("directed", [True, False])
@pytest.mark.parametrize("normalize", ["none", "counts"])
def test_independent_python_rederivation_agrees_bit_for_bit(seed, directed, normalize):
    g = _random_graph(seed)
    alpha, eps = (0.1, 1e-3) if seed != 3 else (0.25, 5e-3)
    c = oracle.cooc_build(g, alpha, eps, directed, normalize, threads=3)
    want, gmax = _py_cooc(g, alpha, eps, directed, normalize)
    got = coo(c)
    assert set(got) == set(want)
    bad = [k for k in want if got[k].view(np.uint32) != f32(want[k]).view(np.uint32)]
    assert not bad, (bad[:5], [(got[k], want[k]) for k in bad[:5]])
    assert c.max == gmax


# ---- BCV.toUnity: how far is the reference's HashMap-order f32 sum from the creation-order sum? -----------
# The oracle and the kernel sum a row in creation order; Java sums values().stream() in HashMap iteration
# order (BCV.java:64-70,96-98).  tests/java_hashmap.py restates that order (OpenJDK 8: tail insertion on
# put, head insertion on merge, order-preserving doubling); the cases below are worked out by hand from it.
from java_hashmap import JavaHashMap  # noqa: E402


def test_hashmap_iteration_order_hand_cases():
    m = JavaHashMap()
    for k in (16, 0, 32, 5):
        m.put(k, 0)
    assert m.keys() == [16, 0, 32, 5]                  # 16 buckets: 16, 0, 32 chain in bucket 0 in arrival order
    for k in range(100, 109):                          # 13 entries > threshold 12 -> 32 buckets
        m.put(k, 0)
    assert len(m.table) == 32
    # bucket 0 splits into low {0, 32} and high {16}, each in the old relative order; 5 and 101 share
    # bucket 5 (101 & 31 = 5) in arrival order; 100 -> 4, 102..108 -> 6..12
    assert m.keys() == [0, 32, 100, 5, 101, 102, 103, 104, 105, 106, 107, 108, 16]
    m.merge(64, 1, lambda a, b: a + b)                 # merge links a NEW key at the head of its bin
    assert m.keys()[:3] == [64, 0, 32]
    m.merge(32, 1, lambda a, b: a + b)
    assert m.get(32) == 1 and m.keys()[:3] == [64, 0, 32]
    m.remove(0)
    assert m.keys()[:2] == [64, 32] and m.size == 13
    big = JavaHashMap()
    for k in (70000, 4464, 65536 + 7, 6):              # h ^ h >>> 16: 70000 -> 70001, 65543 -> 65542
        big.put(k, 0)
    # buckets: 70001 & 15 = 1, 4464 & 15 = 0, 65542 & 15 = 6, 6 -> walked by index, bin 6 in arrival order
    assert big.keys() == [4464, 70000, 65543, 6]


def _java_bcv(g, b, alpha, eps, directed):
    """BCAJob.call (:32-36) on real HashMaps: the forward vector (put per first visit, in pop order), then
    merge(reverse vector) walking the reverse map in ITS iteration order."""
    def as_map(d):
        m = JavaHashMap()
        for k, v in d.items():                                   # dict order = first-visit order; later adds only
            m.put(k, v)                                          # overwrite a value, the structure is already fixed
        return m
    if not directed:
        return as_map(_py_pass(g, b, alpha, eps, "both"))
    fwd = as_map(_py_pass(g, b, alpha, eps, "out"))
    for k, v in as_map(_py_pass(g, b, alpha, eps, "in")).items():
        fwd.merge(k, v, lambda x, y: f32(x + y))
    return fwd


@pytest.mark.parametrize("directed", [True, False])
def test_unity_sum_order_gap_between_reference_and_creation_order(directed):
    """UNITY is oracle-consistent, not reference-bit-exact: this measures what the order of the f32 sum
    changes.  Same key set and the same per-key values before normalisation; the two sequential f32 sums of
    an n-entry row each sit within (n-1) * 2^-24 of the exact sum (asserted per row), in practice ~sqrt(n)
    ulps apart, and the normalised values inherit that relative gap (asserted <= 1e-5 for rows up to ~1500
    entries; the observed maxima are printed)."""
    worst, worst_sum, rows, reordered, longest = 0.0, 0.0, 0, 0, 0
    for seed, (V, E) in enumerate([(70, 420), (400, 4000), (1500, 9000)]):
        g = _random_graph(seed + 20, V, E)
        u = oracle.cooc_build(g, A, 1e-3, directed, "unity")
        got = coo(u)
        for b in range(0, g.V, max(1, g.V // 60)):
            m = _java_bcv(g, b, A, 1e-3, directed)
            m.remove(b)                                          # toUnity: remove(rootNode)
            raw = dict(m.items())
            created = [int(j) for i, j in zip(u.I, u.J) if i == b]   # the oracle's (and the kernel's) row order
            assert sorted(raw) == sorted(created)
            if not raw:
                continue

            def fsum(keys):                                      # values().stream().reduce(Float::sum)
                keys = list(keys)
                t = raw[keys[0]]
                for k in keys[1:]:
                    t = f32(t + raw[k])
                return t
            java_sum, our_sum = fsum(m.keys()), fsum(created)
            exact = sum(float(v) for v in raw.values())
            n = len(raw)
            for t in (java_sum, our_sum):
                assert abs(float(t) - exact) <= max(n - 1, 1) * 2.0 ** -24 * exact * 1.01
            rows += 1
            longest = max(longest, n)
            reordered += m.keys() != created
            worst_sum = max(worst_sum, abs(float(java_sum) - float(our_sum)) / float(np.spacing(our_sum)))
            for k, v in raw.items():
                assert got[(b, k)] == f32(f32(v / our_sum) - f32(1e-6))          # the oracle, restated
                want = f32(f32(v / java_sum) - f32(1e-6))                         # the reference
                worst = max(worst, abs(float(want) - float(got[(b, k)])) / abs(float(want)))
    print(f"UNITY: {rows} rows (longest {longest}), {reordered} iterate in another order than creation order; "
          f"sums up to {worst_sum:.0f} ulp apart, final values up to {worst:.2e} relative")
    assert rows > 100 and reordered > rows // 2
    assert worst <= 1e-5
