"""Known-answer tests that pin the oracle; the reference has no vectors for the
path (documentation only, k8s-operator.md:6), so these hand-checkable cases are the pin."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st


def test_shard_map_kat(oracle, nporacle):
    assert oracle.shard_map([10], 3, 1) == (10, [0], [(0, 4), (4, 8), (8, 10)])
    assert nporacle.shard_map([10], 3, 1) == (10, [0], [(0, 4), (4, 8), (8, 10)])
    # padding: 3 variables of 5 elements, align 4 -> offsets 0, 8, 16; P_pad 24; 2 shards of 12
    assert oracle.shard_map([5, 5, 5], 2, 4) == (24, [0, 8, 16], [(0, 12), (12, 24)])
    # more shards than aligned chunks: trailing slices are empty
    assert oracle.shard_map([4], 3, 4)[2] == [(0, 4), (4, 4), (4, 4)]


def test_div_rows_kat(oracle, nporacle):
    assert [oracle.row_range(10, 4, s) for s in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert [nporacle.row_range(10, 4, s) for s in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert [oracle.row_owner(10, 4, i) for i in range(10)] == [0, 0, 0, 1, 1, 1, 2, 2, 3, 3]


def test_sgd_kat(oracle, nporacle):
    p = np.array([1, 2, 3, 4], np.float32)
    g0, g1 = np.ones(4, np.float32), 3 * np.ones(4, np.float32)
    want = np.array([0, 1, 2, 3], np.float32)
    assert np.array_equal(oracle.dense_sgd(p.copy(), [g0, g1], 0.5), want)
    assert np.array_equal(nporacle.dense_sgd(p, [g0, g1], 0.5), want)


def test_adam_step1_closed_form(oracle, nporacle):
    """With m = v = 0, step 1 is p - lr*g/(|g| + eps/sqrt(1-b2)) ~ p - lr*sign(g)."""
    rng = np.random.default_rng(7)
    n = 4096
    p = rng.standard_normal(n).astype(np.float32)
    g = rng.uniform(-1, 1, n).astype(np.float32)
    g[np.abs(g) < 1e-3] = 0.5
    lr, b1, b2, eps = 1e-3, 0.9, 0.999, 1e-8
    a = oracle.adam_scalars(lr, b1, b2, eps, 1)
    pc, m, v = oracle.dense_adam(p.copy(), np.zeros(n, np.float32), np.zeros(n, np.float32), [g], a)
    closed = p.astype(np.float64) - lr * g.astype(np.float64) / (np.abs(g.astype(np.float64)) + eps / np.sqrt(1 - b2))
    # the update has magnitude ~lr, so 1 ulp of p (~1e-7 relative of |p| <= 5) bounds the difference
    assert np.max(np.abs(pc.astype(np.float64) - closed)) <= 2 * np.spacing(np.float32(5.0))
    pn, mn, vn = nporacle.dense_adam(p, np.zeros(n, np.float32), np.zeros(n, np.float32), [g],
                                     nporacle.adam_scalars(lr, b1, b2, eps, 1))
    assert np.array_equal(pc, pn) and np.array_equal(m, mn) and np.array_equal(v, vn)


def test_sparse_kat(oracle, nporacle):
    ids = np.array([5, 2, 5, 9])
    for o in (oracle, nporacle):
        s, perm, seg, uniq, off = o.sort_segment(ids)
        assert s.tolist() == [2, 5, 5, 9]
        assert perm.tolist() == [1, 0, 2, 3]  # duplicates keep (worker, position) order
        assert seg.tolist() == [0, 1, 1, 2]
        assert uniq.tolist() == [2, 5, 9]
        assert np.diff(off).tolist() == [1, 2, 1]
    # duplicate rows are summed in sorted order, divided by W, applied once
    table = np.zeros((10, 2), np.float32)
    grad = np.array([[1, 1], [2, 2], [3, 3], [4, 4]], np.float32)
    oracle.sparse_apply_sgd(table, 0, grad, ids, np.float32(0.5), np.float32(1.0))
    assert table[5].tolist() == [-2.0, -2.0] and table[2].tolist() == [-1.0, -1.0] and table[9].tolist() == [-2.0, -2.0]
    assert np.count_nonzero(table.any(axis=1)) == 3  # only touched rows change


def test_adagrad_kat(oracle, nporacle):
    table = np.ones((4, 2), np.float32)
    accum = np.full((4, 2), 0.1, np.float32)  # TF's initial accumulator value
    ids = np.array([3, 3])
    grad = np.array([[1, 2], [1, 2]], np.float32)
    oracle.sparse_apply_adagrad(table, accum, 0, grad, ids, np.float32(0.5), np.float32(0.1))
    g = np.array([1, 2], np.float32)
    a2 = np.float32(0.1) + g * g
    assert np.array_equal(accum[3], a2)
    assert np.array_equal(table[3], np.float32(1) - (np.float32(0.1) * g) / np.sqrt(a2))
    assert np.array_equal(table[:3], np.ones((3, 2), np.float32))


def test_hash_kat(oracle, nporacle):
    # fixed vectors of the synthetic-input hash (generated once, committed; guards accidental edits)
    from oracle import c_oracle

    L = c_oracle.lib()
    assert L.ps_hash64(0, 0, 0, 0) == 0xE220A8397B1DCDAF  # splitmix64's first output for state 0
    a = oracle.fill_uniform(8, 2, 3, 1, first=5)
    assert np.array_equal(a, nporacle.uniform(2, 3, 1, 5, 8))
    assert (a >= -1).all() and (a < 1).all()


@settings(max_examples=40, deadline=None)
@given(n=st.integers(1, 3000), W=st.integers(1, 9), steps=st.integers(1, 3), seed=st.integers(0, 2 ** 31), threads=st.sampled_from([1, 3]))
def test_dense_c_vs_numpy(oracle, nporacle, n, W, steps, seed, threads):
    p = oracle.fill_uniform(n, 1, 0, 0, first=seed)
    ps, pa = p.copy(), p.copy()
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    qs, qa, qm, qv = p.copy(), p.copy(), m.copy(), v.copy()
    for t in range(1, steps + 1):
        gs = [oracle.fill_uniform(n, seed, t, w) for w in range(W)]
        oracle.dense_sgd(ps, gs, 0.1, threads)
        oracle.dense_adam(pa, m, v, gs, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t), threads)
        qs = nporacle.dense_sgd(qs, gs, 0.1)
        qa, qm, qv = nporacle.dense_adam(qa, qm, qv, gs, nporacle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t))
    assert np.array_equal(ps, qs)
    assert np.array_equal(pa, qa) and np.array_equal(m, qm) and np.array_equal(v, qv)


@settings(max_examples=40, deadline=None)
@given(ids=st.lists(st.integers(0, 49), min_size=0, max_size=200), dim=st.sampled_from([1, 4, 64]), seed=st.integers(0, 1000))
def test_sparse_c_vs_numpy(oracle, nporacle, ids, dim, seed):
    ids = np.array(ids, np.int64)
    a, b = oracle.sort_segment(ids), nporacle.sort_segment(ids)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    rows = 50
    table = oracle.fill_uniform(rows * dim, 1, 0, 0, first=seed).reshape(rows, dim)
    grad = oracle.fill_uniform(len(ids) * dim, 2, 1, 0, first=seed).reshape(len(ids), dim)
    t1 = oracle.sparse_apply_sgd(table.copy(), 0, grad, ids, np.float32(0.25), np.float32(0.1))
    t2 = nporacle.sparse_apply_sgd(table.copy(), 0, grad, ids, np.float32(0.25), np.float32(0.1))
    assert np.array_equal(t1, t2)
    acc = np.full((rows, dim), 0.1, np.float32)
    t3, a3 = oracle.sparse_apply_adagrad(table.copy(), acc.copy(), 0, grad, ids, np.float32(0.25), np.float32(0.1))
    t4, a4 = nporacle.sparse_apply_adagrad(table.copy(), acc.copy(), 0, grad, ids, np.float32(0.25), np.float32(0.1))
    assert np.array_equal(t3, t4) and np.array_equal(a3, a4)
    if len(ids):
        assert np.array_equal(oracle.gather(table, ids), table[ids])


@settings(max_examples=60, deadline=None)
@given(rows=st.integers(1, 500), S=st.integers(1, 9))
def test_div_partition_properties(oracle, nporacle, rows, S):
    ranges = [oracle.row_range(rows, S, s) for s in range(S)]
    assert ranges[0][0] == 0 and ranges[-1][1] == rows
    assert all(ranges[s][1] == ranges[s + 1][0] for s in range(S - 1))
    sizes = [e - b for b, e in ranges]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    own = nporacle.row_owner(rows, S, np.arange(rows))
    for i in range(rows):
        o = oracle.row_owner(rows, S, i)
        assert o == own[i] and ranges[o][0] <= i < ranges[o][1]


def test_bucket_by_owner(oracle):
    ids = np.array([9, 0, 5, 3, 9, 1, 7])
    perm, off = oracle.bucket_by_owner(ids, 10, 4)
    assert off.tolist() == [0, 2, 4, 5, 7]
    assert perm.tolist() == [1, 5, 2, 3, 6, 0, 4]  # stable inside each bucket


def test_long_segment_chunked_fold(oracle, nporacle):
    """Hot ids: a segment longer than 256 entries is folded chunk-wise; C and NumPy agree, and the
    result differs from a naive left fold only in rounding."""
    rng = np.random.default_rng(3)
    n, dim = 1000, 4
    ids = np.concatenate([np.full(700, 7), rng.integers(0, 20, n - 700)])
    rng.shuffle(ids)
    grad = rng.uniform(-1, 1, (n, dim)).astype(np.float32)
    t0 = rng.uniform(-1, 1, (20, dim)).astype(np.float32)
    a = oracle.sparse_apply_sgd(t0.copy(), 0, grad, ids, np.float32(0.125), np.float32(0.5))
    b = nporacle.sparse_apply_sgd(t0.copy(), 0, grad, ids, np.float32(0.125), np.float32(0.5))
    assert np.array_equal(a, b)
    naive = t0[7].astype(np.float64) - 0.5 * 0.125 * grad[ids == 7].astype(np.float64).sum(0)
    assert np.allclose(a[7], naive, rtol=1e-4, atol=1e-5)


def test_momentum_kat_and_c_vs_numpy(oracle, nporacle):
    """TF ApplyMomentum: accum = momentum*accum + g; p -= lr*accum.  p=1, g=2, lr=.5, momentum=.5:
    step 1 accum 2 -> p 0; step 2 accum 3 -> p -1.5 (all exact in binary)."""
    p, acc = np.array([1.0], np.float32), np.zeros(1, np.float32)
    g = [np.array([2.0], np.float32)]
    oracle.dense_momentum(p, acc, g, 0.5, 0.5)
    assert (p[0], acc[0]) == (0.0, 2.0)
    oracle.dense_momentum(p, acc, g, 0.5, 0.5)
    assert (p[0], acc[0]) == (-1.5, 3.0)
    n, W = 2777, 5
    p = oracle.fill_uniform(n, 1, 0, 0)
    a = np.zeros(n, np.float32)
    q, b = p.copy(), a.copy()
    for t in (1, 2, 3):
        gs = [oracle.fill_uniform(n, 2, t, w) for w in range(W)]
        oracle.dense_momentum(p, a, gs, 0.05, 0.9, threads=3)
        q, b = nporacle.dense_momentum(q, b, gs, 0.05, 0.9)
    assert np.array_equal(p, q) and np.array_equal(a, b)


def test_chunked_fold_of_a_130k_entry_hot_row_against_the_plain_left_fold(oracle):
    """SURVEY 8(a) a12 wrote "sum in sorted order"; the oracle folds a segment longer than 256
    entries chunk-wise (256-entry chunks left to right, then the chunk sums left to right).  For the
    hottest id of C2 under Zipf(1.05) -- 130 108 entries -- this measures what that changes.

    No two fp32 summation orders of 130 k values of magnitude <= 1 agree to 1e-6 relative: the plain
    left fold is itself ~1.6e-5 (absolute, in the updated parameter) from the exact sum, because every
    add rounds a partial sum of magnitude up to ~400.  So the pin is: (1) the chunked result is at
    least as close to the exact (float64) answer as the plain left fold is, (2) the two differ by less
    than the plain fold's own rounding-error bound, and (3) up to 256 entries the two are the same bits."""
    n, dim = 130_108, 64
    g = oracle.fill_uniform(n * dim, 2, 1, 0, threads=4).reshape(n, dim)
    t0 = oracle.fill_uniform(dim, 1, 0, 0).reshape(1, dim)
    inv_w, lr = np.float32(1.0) / np.float32(8), np.float32(0.05)
    chunked = oracle.sparse_apply_sgd(t0.copy(), 0, g, np.zeros(n, np.int64), inv_w, lr)[0]
    plain_sum = np.cumsum(g, axis=0, dtype=np.float32)[-1]  # cumsum is the sequential left fold
    plain = t0[0] - lr * (plain_sum * inv_w)
    exact = t0[0].astype(np.float64) - 0.05 * (g.astype(np.float64).sum(0) / 8.0)
    err_chunked, err_plain = np.abs(chunked - exact).max(), np.abs(plain - exact).max()
    assert err_chunked <= err_plain                      # measured: 1.7e-6 vs 1.6e-5
    assert err_chunked <= 4e-6
    # standard bound for recursive summation: n * u * max|partial sum|, scaled by lr / W
    bound = n * 2.0 ** -24 * np.abs(np.cumsum(g, axis=0, dtype=np.float64)).max() * 0.05 / 8.0
    assert np.abs(chunked - plain).max() <= bound        # measured 1.7e-5, bound ~2e-2
    assert np.abs(chunked - plain).max() <= 5e-5
    # (3) a segment of exactly 256 entries is the plain left fold, bit for bit
    k = 256
    c = oracle.sparse_apply_sgd(t0.copy(), 0, g[:k].copy(), np.zeros(k, np.int64), inv_w, lr)[0]
    q = t0[0] - lr * (np.cumsum(g[:k], axis=0, dtype=np.float32)[-1] * inv_w)
    assert np.array_equal(c.view(np.uint32), q.view(np.uint32))


def test_ftrl_kat_and_c_vs_numpy(oracle, nporacle):
    """TF ApplyFtrl, lr_power = -0.5.  p = 0, accum = 0.1, linear = 0, g = 1, lr = 1, l1 = l2 = 0:
    na = 1.1, linear = 1, p = -1 / sqrt(1.1).  With l1 >= |linear| the weight is clipped to exactly 0."""
    p, a, l = np.zeros(1, np.float32), np.full(1, 0.1, np.float32), np.zeros(1, np.float32)
    oracle.dense_ftrl(p, a, l, [np.ones(1, np.float32)], 1.0, 0.0, 0.0)
    assert p[0] == np.float32(-1.0) / np.sqrt(np.float32(1.1)) and a[0] == np.float32(1.1) and l[0] == 1.0
    p, a, l = np.zeros(1, np.float32), np.full(1, 0.1, np.float32), np.zeros(1, np.float32)
    oracle.dense_ftrl(p, a, l, [np.ones(1, np.float32)], 1.0, 2.0, 0.0)  # l1 = 2 > |linear| = 1
    assert p[0] == 0.0 and l[0] == 1.0
    n, W = 5003, 3
    p, a, l = oracle.fill_uniform(n, 1, 0, 0), np.full(n, 0.1, np.float32), np.zeros(n, np.float32)
    q, b, m = p.copy(), a.copy(), l.copy()
    for t in (1, 2, 3, 4):
        gs = [oracle.fill_uniform(n, 2, t, w) for w in range(W)]
        oracle.dense_ftrl(p, a, l, gs, 0.05, 0.001, 0.01, threads=3)
        q, b, m = nporacle.dense_ftrl(q, b, m, gs, 0.05, 0.001, 0.01)
    assert np.array_equal(p, q) and np.array_equal(a, b) and np.array_equal(l, m)
    assert (p == 0).any() and (p != 0).any()  # both branches of the l1 clip were taken
