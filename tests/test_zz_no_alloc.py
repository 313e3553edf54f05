"""The allocation-free sparse push (b200ps_sparse_push_sgd_noalloc) against the ordinary two-call
push with a caller-owned workspace: the updated tables must be identical bit for bit, on two
calls in a row (the second call starts from the scratch the first one left behind).

This file sorts last on purpose, so that a run stopped at the first failure has already run every
other test.  It passes alone in a fresh process (first use of the static scratch) and after the
rest of the suite."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

import b200ps
from b200ps import kernels as K

pytestmark = pytest.mark.gpu


def _push_inputs(rng, rows, dim, W, B, step, zipf):
    ids = (np.minimum(rng.zipf(1.05, (W, B)) - 1, rows - 1) if zipf else rng.integers(0, rows, (W, B))).astype(np.int64)
    grad = torch.empty(W * B * dim, dtype=torch.float32, device="cuda")
    K.fill_uniform(grad, 2, step, 0)
    grad = grad.view(W, B, dim)
    return [torch.from_numpy(ids[w]).cuda() for w in range(W)], [grad[w].contiguous() for w in range(W)]


def _shards(table, rows, S):
    return [table[slice(*b200ps.row_range(rows, S, s))] for s in range(S)]


@pytest.mark.parametrize("rows,S,dim,W,B,zipf", [(50_000, 8, 64, 8, 4096, True), (3_000, 3, 16, 2, 999, False)])
def test_noalloc_push_matches_workspace_push_twice_in_a_row(rows, S, dim, W, B, zipf):
    rng = np.random.default_rng(rows)
    geom = K.table_geom(rows, dim, S)
    ref = K.fill_uniform(torch.empty(rows * dim, dtype=torch.float32, device="cuda"), 1, 0, 0).view(rows, dim)
    got = ref.clone()
    ws = K.sparse_workspace(W * B, dim)
    for call in (1, 2):  # the second call reuses the scratch left by the first
        ids, grads = _push_inputs(rng, rows, dim, W, B, call, zipf)
        K.sparse_sort_segment(ids, B, geom, ws)
        K.sparse_apply_sgd(_shards(ref, rows, S), grads, B, geom, 0.05, ws)
        K.sparse_push_sgd_noalloc(_shards(got, rows, S), ids, grads, B, geom, 0.05)
        torch.cuda.synchronize()
        # tolerance of the sparse tests: none -- the fold order is fixed, so the bits must agree
        assert torch.equal(got.view(torch.int32), ref.view(torch.int32)), f"call {call} differs"
    assert not torch.equal(ref, K.fill_uniform(torch.empty_like(ref.view(-1)), 1, 0, 0).view(rows, dim))  # it did update


def test_noalloc_push_rejects_more_than_the_scratch_holds():
    rows, dim, S, W = 1000, 4, 1, 2
    B = K.NOALLOC_MAX_ENTRIES // W + 1
    ids = [torch.zeros(B, dtype=torch.int64, device="cuda") for _ in range(W)]
    grads = [torch.zeros(B, dim, dtype=torch.float32, device="cuda") for _ in range(W)]
    table = torch.zeros(rows, dim, dtype=torch.float32, device="cuda")
    with pytest.raises(b200ps.B200PSError, match="exceeds the 262144") as ei:
        K.sparse_push_sgd_noalloc([table], ids, grads, B, K.table_geom(rows, dim, S), 0.1)
    assert ei.value.status == b200ps._lib.ERR_RANGE
    # exactly at the limit is accepted
    B = K.NOALLOC_MAX_ENTRIES // W
    K.sparse_push_sgd_noalloc([table], [i[:B] for i in ids], [g[:B] for g in grads], B, K.table_geom(rows, dim, S), 0.1)
    torch.cuda.synchronize()
