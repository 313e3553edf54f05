"""The committed fixtures of tests/golden/ps_golden.npz (made by tests/golden/make_golden.py from
the NumPy restatement) are reproduced bit for bit by the C oracle -- and, in the gpu-marked test,
by the CUDA kernels through the C ABI."""
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ps_golden.npz"))


def eq(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a,
                                                 b.view(np.uint32) if b.dtype == np.float32 else b)


def test_c_oracle_reproduces_the_golden_fixtures(oracle):
    n, W = int(G["dense_n"]), int(G["dense_W"])
    p0 = oracle.fill_uniform(n, 1, 0, 0)
    ps, pa = p0.copy(), p0.copy()
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    for t in range(1, int(G["dense_steps"]) + 1):
        g = [oracle.fill_uniform(n, 2, t, w) for w in range(W)]
        oracle.dense_sgd(ps, g, 0.05)
        oracle.dense_adam(pa, m, v, g, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t))
    assert eq(ps, G["dense_sgd_p"]) and eq(pa, G["dense_adam_p"]) and eq(m, G["dense_adam_m"]) and eq(v, G["dense_adam_v"])
    g = [oracle.fill_uniform(n, 2, 9, w) for w in range(6)]
    assert eq(oracle.dense_sgd(p0.copy(), oracle.two_level_partials(g, 2), 0.05, inv_w=np.float32(1.0) / np.float32(6)), G["two_level_sgd_p"])
    ids = G["sparse_ids"]
    s, perm, seg, uniq, off = oracle.sort_segment(ids.reshape(-1))
    assert eq(s, G["sparse_sorted"]) and eq(perm, G["sparse_perm"]) and eq(seg, G["sparse_seg"])
    assert eq(uniq, G["sparse_unique"]) and eq(off, G["sparse_off"])
    rows, dim = 50, 8
    grad = oracle.fill_uniform(ids.size * dim, 2, 1, 0).reshape(ids.size, dim)
    table = oracle.fill_uniform(rows * dim, 1, 0, 0).reshape(rows, dim)
    inv_w = np.float32(1.0) / np.float32(ids.shape[0])
    assert eq(oracle.sparse_apply_sgd(table.copy(), 0, grad, ids.reshape(-1), inv_w, np.float32(0.05)), G["sparse_sgd_table"])
    t2, a2 = oracle.sparse_apply_adagrad(table.copy(), np.full((rows, dim), 0.1, np.float32), 0, grad, ids.reshape(-1), inv_w, np.float32(0.05))
    assert eq(t2, G["sparse_adagrad_table"]) and eq(a2, G["sparse_adagrad_accum"])
    hg = oracle.fill_uniform(1000 * 4, 2, 3, 0).reshape(1000, 4)
    ht = oracle.fill_uniform(20 * 4, 1, 0, 0).reshape(20, 4)
    assert eq(oracle.sparse_apply_sgd(ht, 0, hg, G["hot_ids"], np.float32(0.125), np.float32(0.5)), G["hot_sgd_table"])


@pytest.mark.gpu
def test_cuda_kernels_reproduce_the_golden_fixtures():
    torch = pytest.importorskip("torch")
    import b200ps
    from b200ps import kernels as K

    n, W = int(G["dense_n"]), int(G["dense_W"])
    new = lambda k: torch.empty(k, dtype=torch.float32, device="cuda")
    ps = K.fill_uniform(new(n), 1, 0, 0)
    pa, p0 = ps.clone(), ps.clone()
    m, v = torch.zeros_like(ps), torch.zeros_like(ps)
    g = [new(n) for _ in range(W)]
    for t in range(1, int(G["dense_steps"]) + 1):
        for w in range(W):
            K.fill_uniform(g[w], 2, t, w)
        K.dense_apply_sgd(ps, g, 0.05)
        K.dense_apply_adam(pa, m, v, g, K.adam_scalars(1e-3, t=t))
    host = lambda t: t.cpu().numpy()
    assert eq(host(ps), G["dense_sgd_p"]) and eq(host(pa), G["dense_adam_p"]) and eq(host(m), G["dense_adam_m"]) and eq(host(v), G["dense_adam_v"])
    g6 = [K.fill_uniform(new(n), 2, 9, w) for w in range(6)]
    parts = [K.dense_fold(new(n), g6[k::2]) for k in range(2)]
    K.dense_apply_sgd(p0, parts, 0.05, inv_w=K.inv_w_f32(6))
    assert eq(host(p0), G["two_level_sgd_p"])
    ids = G["sparse_ids"]
    Wk, B, rows, dim = ids.shape[0], ids.shape[1], 50, 8
    geom, ws = K.table_geom(rows, dim, 3), K.sparse_workspace(ids.size, dim)
    dids = [torch.from_numpy(ids[w]).cuda() for w in range(Wk)]
    K.sparse_sort_segment(dids, B, geom, ws)
    plan = K.sparse_plan_read(ws, ids.size, rows)
    assert eq(plan["sorted_ids"], G["sparse_sorted"]) and eq(plan["perm"], G["sparse_perm"]) and eq(plan["seg_id"], G["sparse_seg"])
    assert eq(plan["unique_ids"], G["sparse_unique"]) and eq(plan["seg_offset"], G["sparse_off"])
    grad = K.fill_uniform(new(ids.size * dim), 2, 1, 0).view(Wk, B, dim)
    table = K.fill_uniform(new(rows * dim), 1, 0, 0).view(rows, dim)
    shards = lambda t: [t[slice(*b200ps.row_range(rows, 3, s))] for s in range(3)]
    t1 = table.clone()
    K.sparse_apply_sgd(shards(t1), [grad[w] for w in range(Wk)], B, geom, 0.05, ws)
    assert eq(host(t1), G["sparse_sgd_table"])
    t2, a2 = table.clone(), torch.full((rows, dim), 0.1, device="cuda")
    K.sparse_apply_adagrad(shards(t2), shards(a2), [grad[w] for w in range(Wk)], B, geom, 0.05, ws)
    assert eq(host(t2), G["sparse_adagrad_table"]) and eq(host(a2), G["sparse_adagrad_accum"])
    hot = torch.from_numpy(G["hot_ids"]).cuda()
    hg = K.fill_uniform(new(1000 * 4), 2, 3, 0).view(1000, 4)
    ht = K.fill_uniform(new(20 * 4), 1, 0, 0).view(20, 4)
    g1, w1 = K.table_geom(20, 4, 1), K.sparse_workspace(1000, 4)
    for W8 in (8,):  # 8 workers x 125 ids: inv_w = 1/8 = 0.125, entries in the same (worker, position) order
        K.sparse_sort_segment([hot[i * 125:(i + 1) * 125] for i in range(W8)], 125, g1, w1)
        K.sparse_apply_sgd([ht], [hg[i * 125:(i + 1) * 125] for i in range(W8)], 125, g1, 0.5, w1)
    assert eq(host(ht), G["hot_sgd_table"])
