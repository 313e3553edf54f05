"""Generates tests/golden/ps_golden.npz from the NumPy restatement (oracle/np_oracle.py).

The reference has no vectors for this path (documentation only; its one sentence on it is
k8s-operator.md:6), so these are SELF-generated regression fixtures: they freeze today's oracle
semantics so that neither the C oracle, nor the NumPy one, nor the CUDA kernels
can drift unnoticed.  Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import np_oracle as N  # noqa: E402


def main():
    out = {}
    # dense: P = 3000 (not a multiple of 4), W = 3, 4 steps of SGD and of Adam
    n, W = 3000, 3
    p0 = N.uniform(1, 0, 0, 0, n)
    ps, pa = p0.copy(), p0.copy()
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    for t in range(1, 5):
        g = [N.uniform(2, t, w, 0, n) for w in range(W)]
        ps = N.dense_sgd(ps, g, 0.05)
        pa, m, v = N.dense_adam(pa, m, v, g, N.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t))
    out.update(dense_n=n, dense_W=W, dense_steps=4, dense_sgd_p=ps, dense_adam_p=pa, dense_adam_m=m, dense_adam_v=v)
    # two-level fold of 6 workers over 2 GPUs, one SGD step
    g = [N.uniform(2, 9, w, 0, n) for w in range(6)]
    parts = [N.fold(g[k::2]) for k in range(2)]
    out["two_level_sgd_p"] = N.dense_sgd(p0.copy(), parts, 0.05, inv_w=np.float32(1.0) / np.float32(6))
    # sparse: 50 rows x 8, 3 workers x 40 ids, one hot id with a segment longer than 256? -> separate case
    rows, dim, Wk, B = 50, 8, 3, 40
    rng = np.random.default_rng(2026)
    ids = rng.integers(0, rows, (Wk, B)).astype(np.int64)
    grad = N.uniform(2, 1, 0, 0, Wk * B * dim).reshape(Wk * B, dim)
    table = N.uniform(1, 0, 0, 0, rows * dim).reshape(rows, dim)
    s, perm, seg, uniq, off = N.sort_segment(ids.reshape(-1))
    out.update(sparse_ids=ids, sparse_sorted=s, sparse_perm=perm, sparse_seg=seg, sparse_unique=uniq, sparse_off=off)
    inv_w = np.float32(1.0) / np.float32(Wk)
    out["sparse_sgd_table"] = N.sparse_apply_sgd(table.copy(), 0, grad, ids.reshape(-1), inv_w, 0.05)
    t2, a2 = N.sparse_apply_adagrad(table.copy(), np.full((rows, dim), 0.1, np.float32), 0, grad, ids.reshape(-1), inv_w, 0.05)
    out.update(sparse_adagrad_table=t2, sparse_adagrad_accum=a2)
    # hot id: 700 of 1000 entries hit row 7 (chunked fold beyond 256 entries)
    hot = np.concatenate([np.full(700, 7), rng.integers(0, 20, 300)]).astype(np.int64)
    rng.shuffle(hot)
    hg = N.uniform(2, 3, 0, 0, 1000 * 4).reshape(1000, 4)
    ht = N.uniform(1, 0, 0, 0, 20 * 4).reshape(20, 4)
    out.update(hot_ids=hot, hot_sgd_table=N.sparse_apply_sgd(ht.copy(), 0, hg, hot, np.float32(0.125), 0.5))
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ps_golden.npz"), **out)
    print("wrote ps_golden.npz:", {k: getattr(val, "shape", val) for k, val in out.items()})


if __name__ == "__main__":
    main()
