"""Oracle trainer restatement checked against an independent numpy re-derivation of
J/opt/Optimizer.java:50-57,67-120, J/opt/grad/Adagrad.java:49-95, GloveCost/PGloveCost."""
import math

import numpy as np
import pytest

import oracle

f32 = np.float32


def _init_store(V, dim, seed):
    fl = oracle.java_floats(seed, (2 + 2 * dim) * V).astype(np.float64)
    vals = ((fl - 0.5).astype(np.float32) / f32(dim)).reshape(V, 2 + 2 * dim)
    fb, cb = vals[:, 0].copy(), vals[:, 1].copy()
    focus, context = vals[:, 2::2].copy(), vals[:, 3::2].copy()
    return focus, context, fb, cb


def _py_epoch(st, I, J, X, perm, xmax, method):
    focus, context, fb, cb, gf, gc, gfb, gcb = st
    dim = focus.shape[1]
    lr = float(f32(0.05))
    cost = f32(0)
    for k in perm:
        u, v, x = int(I[k]), int(J[k]), f32(X[k])
        inner = f32(0)
        for d in range(dim):
            inner = f32(inner + f32(focus[u, d] * context[v, d]))
        if method == "glove":
            tgt = math.log(float(x))
        else:
            tgt = math.log(float(f32(x / f32(f32(1.0) - x))))
        inner = f32(float(inner) + (float(f32(fb[u] + cb[v])) - tgt))
        if method == "glove":
            wc = inner if float(x) > xmax else f32(f32(math.pow(float(x) / xmax, 0.75)) * inner)
        else:
            wc = f32(x * inner)
        cost = f32(float(cost) + 0.5 * float(wc) * float(inner))
        for d in range(dim):
            g1, g2 = f32(wc * context[v, d]), f32(wc * focus[u, d])
            focus[u, d] = f32(float(focus[u, d]) - float(g1) / math.sqrt(float(gf[u, d])) * lr)
            context[v, d] = f32(float(context[v, d]) - float(g2) / math.sqrt(float(gc[v, d])) * lr)
            gf[u, d] = f32(gf[u, d] + f32(g1 * g1))
            gc[v, d] = f32(gc[v, d] + f32(g2 * g2))
        fb[u] = f32(float(fb[u]) - float(wc) / math.sqrt(float(gfb[u])))
        cb[v] = f32(float(cb[v]) - float(wc) / math.sqrt(float(gcb[v])))
        wc2 = f32(wc * wc)
        gfb[u] = f32(gfb[u] + wc2)
        gcb[v] = f32(gcb[v] + wc2)
    return float(cost) / len(perm)


@pytest.mark.parametrize("method", ["glove", "pglove"])
def test_epochs_match_python_rederivation(method):
    rng = np.random.default_rng(3)
    V, dim, nnz, seed = 12, 5, 60, 777
    I = rng.integers(0, V, nnz).astype(np.int32)
    J = rng.integers(0, V, nnz).astype(np.int32)
    X = rng.uniform(1e-3, 0.4, nnz).astype(np.float32)
    xmax = float(X.max())
    t = oracle.Trainer(V, I, J, X, xmax, dim, method, seed)
    focus, context, fb, cb = _init_store(V, dim, seed)
    assert np.array_equal(t.store("focus"), focus) and np.array_equal(t.store("cBias"), cb)
    st = [focus, context, fb, cb, np.ones_like(focus), np.ones_like(context),
          np.ones_like(fb), np.ones_like(cb)]
    skip = (2 + 2 * dim) * V          # init draws precede the first shuffle on the same stream
    for ep in range(1, 4):
        perm = oracle.java_shuffle(seed, skip, nnz, ep)       # cumulative Fisher-Yates
        c_py = _py_epoch(st, I, J, X, perm, xmax, method)
        c_or = t.epoch()
        assert np.array_equal(t.get_perm(), perm)
        assert c_or == c_py
    for name, ref in zip(["focus", "context", "fBias", "cBias", "gsFocus", "gsContext",
                          "gsFBias", "gsCBias"], st):
        assert np.array_equal(t.store(name), ref), name
    assert np.array_equal(t.extract(), ((focus + context) / f32(2)).astype(np.float64))


def test_optimize_stop_rule_and_final_cost():
    rng = np.random.default_rng(4)
    V, dim, nnz = 20, 4, 150
    I = rng.integers(0, V, nnz).astype(np.int32)
    J = rng.integers(0, V, nnz).astype(np.int32)
    X = rng.uniform(1e-3, 0.4, nnz).astype(np.float32)
    # maxiter exhausted -> finalCost stays 0 (Optimizer.java:73,104-107)
    h, fc = oracle.Trainer(V, I, J, X, float(X.max()), dim, "glove", 1).optimize(5, 0.0)
    assert len(h) == 5 and fc == 0.0 and h[0] > h[-1]
    # huge tolerance -> stops after the first epoch with finalCost = that cost
    h, fc = oracle.Trainer(V, I, J, X, float(X.max()), dim, "glove", 1).optimize(5, 1e9)
    assert len(h) == 1 and fc == h[0]


def test_hogwild_threads_slice_like_reference():
    rng = np.random.default_rng(5)
    V, dim, nnz = 2000, 8, 20000
    I = rng.integers(0, V, nnz).astype(np.int32)
    J = rng.integers(0, V, nnz).astype(np.int32)
    X = rng.uniform(1e-3, 0.4, nnz).astype(np.float32)
    a = oracle.Trainer(V, I, J, X, float(X.max()), dim, "glove", 9, threads=1)
    b = oracle.Trainer(V, I, J, X, float(X.max()), dim, "glove", 9, threads=4)
    ca = [a.epoch() for _ in range(3)]
    cb = [b.epoch() for _ in range(3)]
    assert np.allclose(ca, cb, rtol=0.02)


def _py_epoch_adam(st, I, J, X, perm, xmax, method, iteration, amsgrad):
    """J/opt/grad/Adam.java:75-146 / AMSGrad.java:91-162 re-derived with numpy scalar types."""
    focus, context, fb, cb, v_f, v_c, v_fb, v_cb, m_f, m_c, m_fb, m_cb = st
    dim = focus.shape[1]
    b1, b2, eps, lr = f32(0.9), f32(0.999), float(f32(1e-7)), float(f32(0.05))
    omb1, omb2 = f32(f32(1) - b1), f32(f32(1) - b2)
    corr = lr * math.sqrt(1 - math.pow(float(b2), iteration + 1)) / (1 - math.pow(float(b1), iteration + 1))

    def second(old, g2):
        nv = f32(f32(b2 * old) + f32(omb2 * g2))
        return (nv if old <= nv else old) if amsgrad else nv

    def step(p, m, v):
        if amsgrad:
            return f32(float(p) - lr / (math.sqrt(float(v)) + eps) * float(m))
        return f32(float(p) - corr * float(m) / (math.sqrt(float(v)) + eps))

    cost = f32(0)
    for k in perm:
        u, v, x = int(I[k]), int(J[k]), f32(X[k])
        inner = f32(0)
        for d in range(dim):
            inner = f32(inner + f32(focus[u, d] * context[v, d]))
        tgt = math.log(float(x)) if method == "glove" else math.log(float(f32(x / f32(f32(1.0) - x))))
        inner = f32(float(inner) + (float(f32(fb[u] + cb[v])) - tgt))
        if method == "glove":
            wc = inner if float(x) > xmax else f32(f32(math.pow(float(x) / xmax, 0.75)) * inner)
        else:
            wc = f32(x * inner)
        cost = f32(float(cost) + 0.5 * float(wc) * float(inner))
        for d in range(dim):
            gu, gv = f32(wc * context[v, d]), f32(wc * focus[u, d])
            m1 = f32(f32(b1 * m_f[u, d]) + f32(omb1 * gu))
            m2 = f32(f32(b1 * m_c[v, d]) + f32(omb1 * gv))
            v1, v2 = second(v_f[u, d], f32(gu * gu)), second(v_c[v, d], f32(gv * gv))
            focus[u, d], context[v, d] = step(focus[u, d], m1, v1), step(context[v, d], m2, v2)
            m_f[u, d], m_c[v, d], v_f[u, d], v_c[v, d] = m1, m2, v1, v2
        m1 = f32(f32(b1 * m_fb[u]) + f32(omb1 * wc))
        m2 = f32(f32(b1 * m_cb[v]) + f32(omb1 * wc))
        w2 = f32(wc * wc)
        v1, v2 = second(v_fb[u], w2), second(v_cb[v], w2)
        fb[u], cb[v] = step(fb[u], m1, v1), step(cb[v], m2, v2)
        m_fb[u], m_cb[v], v_fb[u], v_cb[v] = m1, m2, v1, v2
    return float(cost) / len(perm)


@pytest.mark.parametrize("optimizer", ["adam", "amsgrad"])
@pytest.mark.parametrize("method", ["glove", "pglove"])
def test_adam_amsgrad_match_python_rederivation(method, optimizer):
    rng = np.random.default_rng(6)
    V, dim, nnz, seed = 10, 4, 50, 4242
    I = rng.integers(0, V, nnz).astype(np.int32)
    J = rng.integers(0, V, nnz).astype(np.int32)
    X = rng.uniform(1e-3, 0.4, nnz).astype(np.float32)
    xmax = float(X.max())
    t = oracle.Trainer(V, I, J, X, xmax, dim, method, seed, optimizer=optimizer)
    focus, context, fb, cb = _init_store(V, dim, seed)
    z = np.zeros_like
    st = [focus, context, fb, cb, z(focus), z(context), z(fb), z(cb), z(focus), z(context), z(fb), z(cb)]
    skip = (2 + 2 * dim) * V
    for ep in range(1, 4):
        perm = oracle.java_shuffle(seed, skip, nnz, ep)
        c_py = _py_epoch_adam(st, I, J, X, perm, xmax, method, ep - 1, optimizer == "amsgrad")
        assert t.epoch() == c_py
    names = ["focus", "context", "fBias", "cBias", "gsFocus", "gsContext", "gsFBias", "gsCBias",
             "m1Focus", "m1Context", "m1FBias", "m1CBias"]
    for name, ref in zip(names, st):
        assert np.array_equal(t.store(name), ref), name
