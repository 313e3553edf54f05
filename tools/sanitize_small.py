"""A small pass over every kernel family, meant to run under compute-sanitizer
(`--tool memcheck` or `--tool racecheck`, one tool per gpurun call)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import kernels as K  # noqa: E402


def main():
    new = lambda n: torch.empty(n, dtype=torch.float32, device="cuda")
    for n, W in ((4099, 3), (65536, 4), (1023, 9)):
        p = K.fill_uniform(new(n), 1, 0, 0)
        m, v = torch.zeros_like(p), torch.zeros_like(p)
        g = [K.fill_uniform(new(n), 2, 1, w) for w in range(W)]
        K.dense_apply_sgd(p, g, 0.1, mirrors=[new(n)])
        K.dense_apply_adam(p, m, v, g, K.adam_scalars(1e-3, t=1))
        K.dense_fold(new(n), g)
    rng = np.random.default_rng(0)
    for rows, S, dim, W, B, hot in ((1000, 3, 64, 4, 257, False), (300, 2, 8, 2, 2500, True), (70000, 5, 128, 3, 700, False)):
        ids = rng.integers(0, rows, (W, B))
        if hot:
            ids[:, : B // 2] = 5
        dids = [torch.from_numpy(ids[w].astype(np.int64)).cuda() for w in range(W)]
        grads = [K.fill_uniform(new(B * dim), 2, 1, w).view(B, dim) for w in range(W)]
        table = K.fill_uniform(new(rows * dim), 1, 0, 0).view(rows, dim)
        acc = torch.full((rows, dim), 0.1, device="cuda")
        sh = lambda t: [t[slice(*b200ps.row_range(rows, S, s))] for s in range(S)]
        geom, ws = K.table_geom(rows, dim, S), K.sparse_workspace(W * B, dim)
        K.sparse_sort_segment(dids, B, geom, ws)
        K.sparse_apply_sgd(sh(table), grads, B, geom, 0.05, ws)
        K.sparse_apply_adagrad(sh(table), sh(acc), grads, B, geom, 0.05, ws)
        out = new(B * dim).view(B, dim)
        K.sparse_gather(out, sh(table), geom, dids[0])
        K.sparse_push_sgd_noalloc(sh(table), dids, grads, B, geom, 0.05)
        K.sparse_plan_read(ws, W * B, rows)
    tab = b200ps.EmbeddingTable(5000, 16, 4, 4, 333, optimizer="adagrad")
    tab.init_uniform(1)
    for w in range(4):
        tab.ids[w].copy_(torch.from_numpy(rng.integers(0, 5000, 333)))
        K.fill_uniform(tab.grad_rows[w].view(-1), 2, 1, w)
    tab.pull_step()
    tab.push_step()
    tab.close()
    torch.cuda.synchronize()
    print("sanitize_small: done")


if __name__ == "__main__":
    main()
