"""Seeded synthetic graphs in the hand-off layout of the hot path.

The reference ingests RDF (JenaReader / Rdf2GrphConverter — out of scope) and hands the
builder four jagged arrays: unique in-/out-neighbours per vertex plus the aligned edge whose
f32 weight is used (J/bca/BookmarkColoring.java:49-52).  Here the same information is a pair
of CSR structures with neighbour lists sorted by id and no duplicates (simple digraphs).

All generators are pure numpy and deterministic for a given seed, so fixtures and parity
cases are reproducible on any machine.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class CsrGraph:
    """out_* : out-neighbours of v are out_nbr[out_off[v]:out_off[v+1]] with weights out_w.
    in_*  : same for in-neighbours; in_w[k] is the weight of edge in_nbr[k] -> v."""

    V: int
    out_off: np.ndarray
    out_nbr: np.ndarray
    out_w: np.ndarray
    in_off: np.ndarray
    in_nbr: np.ndarray
    in_w: np.ndarray

    @property
    def E(self) -> int:
        return int(self.out_nbr.shape[0])


@dataclass
class EdgeList:
    """Raw directed edge list (parallel edges allowed) — the input of the device-side adjacency
    flattening (ge_set_graph_edges), i.e. the graph BEFORE getIn/OutNeighborhoods and
    In/OutEdgeNeighborhoodAlgorithm.compute (J/bca/BookmarkColoring.java:49-52)."""

    V: int
    src: np.ndarray
    dst: np.ndarray
    w: np.ndarray


def csr_from_edges(V: int, src, dst, w) -> CsrGraph:
    """Build both CSRs from an edge list; duplicates (s,d) keep the first weight, order by id."""
    src = np.asarray(src, np.int64)
    dst = np.asarray(dst, np.int64)
    w = np.asarray(w, np.float32)
    key = (src << 32) | dst
    key, first = np.unique(key, return_index=True)
    src, dst, w = (key >> 32), (key & 0xFFFFFFFF), w[first]
    out_off = np.zeros(V + 1, np.int64)
    np.cumsum(np.bincount(src, minlength=V), out=out_off[1:])
    out_nbr = dst.astype(np.int32)
    out_w = w.copy()
    order = np.argsort((dst << 32) | src, kind="stable")
    in_off = np.zeros(V + 1, np.int64)
    np.cumsum(np.bincount(dst, minlength=V), out=in_off[1:])
    in_nbr = src[order].astype(np.int32)
    in_w = w[order]
    return CsrGraph(V, out_off, out_nbr, out_w, in_off, in_nbr, in_w)


def _mix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser (vectorised) — used for label scrambling and edge-id hashing."""
    x = x.astype(np.uint64)
    with np.errstate(over="ignore"):
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return x


def rmat(V: int, E: int, seed: int, a=0.57, b=0.19, c=0.19, light_frac=0.1) -> CsrGraph:
    """R-MAT (a,b,c,d) power-law digraph: self-loops and duplicate edges removed, labels
    scrambled, weight 1.0f for ~90 % of edges and 0.1f for ~10 % (by edge hash) — C3/C4."""
    rng = np.random.default_rng(seed)
    bits = max(1, int(np.ceil(np.log2(V))))
    src = np.zeros(E, np.int64)
    dst = np.zeros(E, np.int64)
    for _ in range(bits):
        r = rng.random(E, dtype=np.float32)
        sbit = r >= (a + b)                              # quadrants c, d -> source bit 1
        dbit = ((r >= a) & (r < a + b)) | (r >= a + b + c)  # quadrants b, d -> dest bit 1
        src = (src << 1) | sbit
        dst = (dst << 1) | dbit
    salt = np.uint64(seed * 0x9E3779B97F4A7C15 & 0xFFFFFFFFFFFFFFFF)
    src = (_mix64(src.astype(np.uint64) ^ salt) % np.uint64(V)).astype(np.int64)
    dst = (_mix64(dst.astype(np.uint64) ^ salt) % np.uint64(V)).astype(np.int64)
    keep = src != dst
    src, dst = src[keep], dst[keep]
    h = _mix64(((src.astype(np.uint64) << np.uint64(32)) | dst.astype(np.uint64)) + np.uint64(seed))
    w = np.where((h % np.uint64(1000)) < np.uint64(int(light_frac * 1000)), np.float32(0.1), np.float32(1.0))
    return csr_from_edges(V, src, dst, w.astype(np.float32))


def _zipf_pick(rng, n_items: int, size: int, s: float = 1.0) -> np.ndarray:
    ranks = np.arange(1, n_items + 1, dtype=np.float64)
    p = ranks ** (-s)
    p /= p.sum()
    return rng.choice(n_items, size=size, p=p)


def _similarity_edges(rng, lo: int, n: int, pairs: int, wlo: float, whi: float = 1.0):
    """Bidirectional literal<->literal edges with weight = similarity in [wlo, whi)
    (J/convert/Rdf2GrphConverter.java:158-171)."""
    a = lo + rng.integers(0, n, pairs)
    b = lo + rng.integers(0, n, pairs)
    ok = a != b
    a, b = a[ok], b[ok]
    w = rng.uniform(wlo, whi, a.shape[0]).astype(np.float32)
    return np.concatenate([a, b]), np.concatenate([b, a]), np.concatenate([w, w])


def dblp_like(scale: float = 1.0, seed: int = 1) -> CsrGraph:
    """C1: DBLP-shaped co-authorship graph (dblp.config.yml:6-16).  scale=1 -> V=320 k, E~1 M."""
    rng = np.random.default_rng(seed)
    nP, nA = int(100_000 * scale), int(60_000 * scale)
    nT, nN = nP, nA
    P0, A0, T0, N0 = 0, nP, nP + nA, nP + nA + nT
    V = nP + nA + nT + nN
    S, D, W = [], [], []

    def add(s, d, w=1.0):
        S.append(np.asarray(s, np.int64)); D.append(np.asarray(d, np.int64))
        W.append(np.full(len(s), w, np.float32) if np.isscalar(w) else np.asarray(w, np.float32))

    add(P0 + np.arange(nP), T0 + np.arange(nT))                   # paper -> title literal
    add(A0 + np.arange(nA), N0 + np.arange(nN))                   # author -> name literal
    add(P0 + rng.integers(0, nP, 3 * nP), A0 + _zipf_pick(rng, nA, 3 * nP))   # dc:creator
    ref_src = rng.integers(1, nP, int(5.4 * nP))                  # dcterms:references (older papers, skewed)
    ref_dst = (ref_src * rng.random(ref_src.shape[0]) ** 2).astype(np.int64)
    add(P0 + ref_src, P0 + ref_dst)
    npairs = max(1, int(2_000 * scale))
    add(*_similarity_edges(rng, N0, nN, npairs, 0.95))            # name ~ name  (threshold .95)
    add(*_similarity_edges(rng, T0, nT, npairs, 0.5))             # title ~ title (threshold .5)
    s, d, w = np.concatenate(S), np.concatenate(D), np.concatenate(W)
    keep = s != d
    return csr_from_edges(V, s[keep], d[keep], w[keep])


def onstage_like(scale: float = 1.0, seed: int = 2) -> CsrGraph:
    """C2a: onstage-shaped — 11 predicates, nine with weight 1.0f and two with 0.1f
    (onstage.config.yml:10-22); shows -> plays/persons/dates/literals; V~200 k, E~800 k."""
    rng = np.random.default_rng(seed)
    nShow, nPlay, nPers = int(60_000 * scale), int(15_000 * scale), int(25_000 * scale)
    nLit = int(100_000 * scale)
    S0, L0 = 0, nShow + nPlay + nPers
    PL0, PE0 = nShow, nShow + nPlay
    V = nShow + nPlay + nPers + nLit
    S, D, W = [], [], []

    def add(s, d, w):
        S.append(np.asarray(s, np.int64)); D.append(np.asarray(d, np.int64))
        W.append(np.full(len(s), w, np.float32) if np.isscalar(w) else np.asarray(w, np.float32))

    shows = S0 + np.arange(nShow)
    add(shows, PL0 + _zipf_pick(rng, nPlay, nShow), 0.1)                         # workPerformed 0.1
    add(S0 + rng.integers(0, nShow, nShow), S0 + rng.integers(0, nShow, nShow), 0.1)  # subEvent 0.1
    add(shows, L0 + rng.integers(0, nLit // 4, nShow), 1.0)                      # startDate (shared dates)
    add(PL0 + np.arange(nPlay), L0 + nLit // 4 + np.arange(nPlay) % (nLit // 4), 1.0)  # name
    add(PL0 + rng.integers(0, nPlay, 2 * nPlay), PE0 + _zipf_pick(rng, nPers, 2 * nPlay), 1.0)  # creator
    add(PL0 + rng.integers(0, nPlay, nPlay), PL0 + rng.integers(0, nPlay, nPlay), 1.0)  # isBasedOn
    add(PL0 + np.arange(nPlay), L0 + nLit // 2 + rng.integers(0, 40, nPlay), 1.0)       # inLanguage
    add(PE0 + np.arange(nPers), L0 + nLit // 2 + 64 + np.arange(nPers) % (nLit // 4), 1.0)  # person name
    add(shows, PL0 + rng.integers(0, nPlay, nShow), 1.0)                         # exampleOfWork
    add(PL0 + np.arange(nPlay), L0 + rng.integers(0, nLit // 4, nPlay), 1.0)     # dateCreated
    add(shows, L0 + 3 * (nLit // 4) + np.arange(nShow) % (nLit // 4), 1.0)       # headline
    add(*_similarity_edges(rng, L0 + nLit // 2 + 64, nLit // 4, int(3_000 * scale), 0.8))
    add(*_similarity_edges(rng, L0 + 3 * (nLit // 4), nLit // 4, int(3_000 * scale), 0.75))
    s, d, w = np.concatenate(S), np.concatenate(D), np.concatenate(W)
    keep = s != d
    return csr_from_edges(V, s[keep], d[keep], w[keep])


def saa_like(scale: float = 1.0, seed: int = 3) -> CsrGraph:
    """C2b: saa-shaped — 12 predicates all weight 1 (saa.config.yml:8-33): record -> person
    stars, person -> full_name literal, name ~ name similarity >= .95; V~500 k, E~1.2 M."""
    rng = np.random.default_rng(seed)
    nRec, nPers = int(120_000 * scale), int(190_000 * scale)
    nName = nPers
    R0, P0, N0 = 0, nRec, nRec + nPers
    V = nRec + nPers + nName
    S, D, W = [], [], []

    def add(s, d, w=1.0):
        S.append(np.asarray(s, np.int64)); D.append(np.asarray(d, np.int64))
        W.append(np.full(len(s), w, np.float32) if np.isscalar(w) else np.asarray(w, np.float32))

    for _ in range(4):                                             # mentionsGroom/Bride/Mother/Father...
        add(R0 + np.arange(nRec), P0 + rng.integers(0, nPers, nRec))
    add(R0 + np.arange(nRec), P0 + _zipf_pick(rng, min(nPers, 2_000), nRec))  # religion/church hubs
    add(P0 + np.arange(nPers), N0 + np.arange(nName))             # full_name
    add(*_similarity_edges(rng, N0, nName, int(150_000 * scale), 0.95))
    s, d, w = np.concatenate(S), np.concatenate(D), np.concatenate(W)
    keep = s != d
    return csr_from_edges(V, s[keep], d[keep], w[keep])


def tiny(edges, V=None) -> CsrGraph:
    """Hand-written graph from [(src, dst, weight), ...] — for hand-derived oracle cases."""
    if len(edges) == 0:
        assert V is not None
        z = np.zeros(0, np.int64)
        return csr_from_edges(V, z, z, np.zeros(0, np.float32))
    s, d, w = zip(*edges)
    V = V if V is not None else int(max(max(s), max(d))) + 1
    return csr_from_edges(V, np.array(s), np.array(d), np.array(w, np.float32))


def synthetic_coo(V: int, nnz: int, seed: int = 6, s: float = 1.0):
    """C5-style direct COO: i, j ~ Zipf(s) over V ids (hashed to decorrelate), X ~ U(1e-4, .2)."""
    rng = np.random.default_rng(seed)
    u = rng.random(nnz)
    # inverse-CDF of a continuous Zipf(1) approximation on [1, V]
    i = np.floor(np.exp(u * np.log(V))).astype(np.int64) - 1
    u = rng.random(nnz)
    j = np.floor(np.exp(u * np.log(V))).astype(np.int64) - 1
    i = (_mix64(i.astype(np.uint64) + np.uint64(seed)) % np.uint64(V)).astype(np.int32)
    j = (_mix64(j.astype(np.uint64) + np.uint64(seed + 1)) % np.uint64(V)).astype(np.int32)
    x = rng.uniform(1e-4, 0.2, nnz).astype(np.float32)
    return i, j, x


def synthetic_coo_torch(V: int, nnz: int, row_first: int, row_count: int, seed: int = 6, device="cuda",
                         col_zipf: float = 0.7):
    """C5-style direct COO for ONE rank of the block-rotation trainer, generated with torch on
    `device`.  Shaped like a BCA-built matrix: rows uniform over the rank's own id range (a BCA row
    is bounded by 2/(alpha*eps) entries, so no row is a hub), columns Zipf(col_zipf) over all V ids
    (hashed to decorrelate id and popularity; s = 0.7 puts ~0.15 % of all nonzeros on the hottest
    column at V = 50 M, the order of the hub columns of the R-MAT matrices), X ~ U(1e-4, 0.2).
    Zipf(1.0) on BOTH sides (6 % of all nonzeros on one id = hundreds of concurrent writers on one
    row) makes the atomic-add Hogwild update overshoot and diverge — see DESIGN.md §5.
    Returns host numpy arrays (ge_cooc_import takes host pointers)."""
    import torch

    gen = torch.Generator(device=device)
    gen.manual_seed(seed * 1000003 + row_first)

    def lsr(x, n):
        return (x >> n) & ((1 << (64 - n)) - 1)

    def mix(x):
        x = (x ^ lsr(x, 30)) * (-4658895280553007687)
        x = (x ^ lsr(x, 27)) * (-7723592293110705685)
        return x ^ lsr(x, 31)

    i = (row_first + torch.randint(0, row_count, (nnz,), device=device, generator=gen)).to(torch.int32).cpu().numpy()
    u = torch.rand(nnz, device=device, generator=gen, dtype=torch.float64)
    if abs(col_zipf - 1.0) < 1e-9:
        r = torch.exp(u * float(np.log(V)))
    else:                                   # inverse CDF of the continuous Zipf(s) on [1, V]
        a = 1.0 - col_zipf
        r = ((float(V) ** a - 1.0) * u + 1.0) ** (1.0 / a)
    j = (lsr(mix(torch.floor(r).to(torch.int64) - 1 + (seed + 1)), 1) % V).to(torch.int32).cpu().numpy()
    del u, r
    x = (1e-4 + (0.2 - 1e-4) * torch.rand(nnz, device=device, generator=gen)).to(torch.float32).cpu().numpy()
    return i, j, x


def rmat_torch(V: int, E: int, seed: int, device="cuda", a=0.57, b=0.19, c=0.19, light_frac=0.1) -> CsrGraph:
    """Same construction as rmat() but generated with torch on `device` (bench-scale graphs:
    100 M edges in seconds).  The random stream differs from the numpy version; parity tests
    use rmat(), the bench only needs a graph of the configured shape."""
    import torch

    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    bits = max(1, int(np.ceil(np.log2(V))))
    src = torch.zeros(E, dtype=torch.int64, device=device)
    dst = torch.zeros(E, dtype=torch.int64, device=device)
    for _ in range(bits):
        r = torch.rand(E, device=device, generator=gen)
        sbit = (r >= (a + b)).to(torch.int64)
        dbit = (((r >= a) & (r < a + b)) | (r >= a + b + c)).to(torch.int64)
        src = (src << 1) | sbit
        dst = (dst << 1) | dbit
    del r, sbit, dbit

    def lsr(x, n):  # logical shift right on int64
        return (x >> n) & ((1 << (64 - n)) - 1)

    def mix(x):
        x = (x ^ lsr(x, 30)) * (-4658895280553007687)   # 0xBF58476D1CE4E5B9 as int64
        x = (x ^ lsr(x, 27)) * (-7723592293110705685)   # 0x94D049BB133111EB as int64
        return x ^ lsr(x, 31)

    salt = (seed * 0x9E3779B97F4A7C15) & 0x7FFFFFFFFFFFFFFF
    src = lsr(mix(src ^ salt), 1) % V
    dst = lsr(mix(dst ^ salt), 1) % V
    keep = src != dst
    src, dst = src[keep], dst[keep]
    key = torch.unique((src << 32) | dst)           # sorted, duplicates removed
    src, dst = key >> 32, key & 0xFFFFFFFF
    h = lsr(mix(key + seed), 1) % 1000
    w = torch.where(h < int(light_frac * 1000), 0.1, 1.0).to(torch.float32)
    out_off = torch.zeros(V + 1, dtype=torch.int64, device=device)
    out_off[1:] = torch.cumsum(torch.bincount(src, minlength=V), 0)
    order = torch.argsort((dst << 32) | src)
    in_off = torch.zeros(V + 1, dtype=torch.int64, device=device)
    in_off[1:] = torch.cumsum(torch.bincount(dst, minlength=V), 0)
    g = CsrGraph(V, out_off.cpu().numpy(), dst.to(torch.int32).cpu().numpy(), w.cpu().numpy(),
                 in_off.cpu().numpy(), src[order].to(torch.int32).cpu().numpy(), w[order].cpu().numpy())
    del key, src, dst, h, w, order
    torch.cuda.empty_cache() if str(device).startswith("cuda") else None
    return g
