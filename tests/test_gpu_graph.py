"""Device adjacency flattening (ge_set_graph_edges, csrc/graph.cu) vs the host CSR builder:
unique sorted neighbour lists, first-parallel-edge-wins weights, both directions; then the BCA
build from the edge list equals the build from the host CSR bit for bit."""
import numpy as np
import pytest

from graph_embeddings_b200 import BCA, BookmarkColoring, Configuration, GraphEmbError, Opt, synth

pytestmark = pytest.mark.gpu


def cfg():
    return Configuration(method="glove", dim=8, bca=BCA(0.1, 1e-3, True), opt=Opt("adagrad", 0, 1))


def edges(V, E, seed, dup_frac=0.3):
    rng = np.random.default_rng(seed)
    s = rng.integers(0, V, E).astype(np.int32)
    d = rng.integers(0, V, E).astype(np.int32)
    nd = int(E * dup_frac)                       # parallel edges with different weights + self loops
    pick = rng.integers(0, E, nd)
    s = np.concatenate([s, s[pick]]); d = np.concatenate([d, d[pick]])
    w = rng.uniform(0.05, 1.0, len(s)).astype(np.float32)
    p = rng.permutation(len(s))
    return synth.EdgeList(V, s[p], d[p], w[p])


@pytest.mark.parametrize("V,E", [(1, 0), (1, 3), (5, 1), (300, 2000), (5000, 4096), (70000, 300000), (3, 100000)])
def test_flattened_csr_matches_host_builder(V, E):
    el = edges(V, E, V + E) if E else synth.EdgeList(V, np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.float32))
    ref = synth.csr_from_edges(V, el.src, el.dst, el.w)      # np.unique(return_index): first edge wins
    b = BookmarkColoring(el, cfg())
    oo, on, ow, io, inn, iw, sec = b.graph_csr()
    assert np.array_equal(oo, ref.out_off) and np.array_equal(on, ref.out_nbr) and np.array_equal(ow, ref.out_w)
    assert np.array_equal(io, ref.in_off) and np.array_equal(inn, ref.in_nbr) and np.array_equal(iw, ref.in_w)
    b2 = BookmarkColoring(ref, cfg())
    for x, y in zip(b.coo(), b2.coo()):
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32))
    assert b.max() == b2.max()


def test_out_of_range_vertex_is_rejected():
    el = synth.EdgeList(10, np.array([1, 2, 10], np.int32), np.array([0, 1, 2], np.int32), np.ones(3, np.float32))
    with pytest.raises(GraphEmbError) as e:
        BookmarkColoring(el, cfg())
    assert "out of range" in str(e.value)
