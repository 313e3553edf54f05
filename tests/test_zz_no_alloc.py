"""The allocation-free sparse push (b200ps_sparse_push_sgd_noalloc) against the ordinary two-call
push with a caller-owned workspace AND against the CPU oracle: the updated tables must be
identical bit for bit, on two calls in a row (the second call starts from the scratch the first
one left behind).  Both limits of the fixed scratch are exercised: the entry count and the row
width.

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
def test_noalloc_push_matches_workspace_push_and_oracle_twice_in_a_row(oracle, rows, S, dim, W, B, zipf):
    rng = np.random.default_rng(rows)
    geom = K.table_geom(rows, dim, S)
    ref = K.fill_uniform(torch.empty(rows * dim, dtype=torch.float32, device="cuda"), 1, 0, 0).view(rows, dim)
    got = ref.clone()
    host = oracle.fill_uniform(rows * dim, 1, 0, 0).reshape(rows, dim)  # the oracle's table
    ws = K.sparse_workspace(W * B, dim)
    for call in (1, 2):  # the second call reuses the scratch left by the first
        ids, grads = _push_inputs(rng, rows, dim, W, B, call, zipf)
        K.sparse_sort_segment(ids, B, geom, ws)
        K.sparse_apply_sgd(_shards(ref, rows, S), grads, B, geom, 0.05, ws)
        K.sparse_push_sgd_noalloc(_shards(got, rows, S), ids, grads, B, geom, 0.05)
        torch.cuda.synchronize()
        # tolerance of the sparse tests: none -- the fold order is fixed, so the bits must agree
        assert torch.equal(got.view(torch.int32), ref.view(torch.int32)), f"call {call} differs from the two-call push"
        h_ids = np.concatenate([i.cpu().numpy() for i in ids])
        h_grad = np.concatenate([g.cpu().numpy() for g in grads]).reshape(W * B, dim)
        oracle.sparse_apply_sgd(host, 0, h_grad, h_ids, np.float32(1.0) / np.float32(W), np.float32(0.05))
        assert np.array_equal(got.cpu().numpy().view(np.uint32), host.view(np.uint32)), f"call {call} differs from the oracle"
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


def test_noalloc_push_rejects_rows_too_wide_for_the_scratch():
    """The scratch also holds one partial row per 256-entry chunk of a hot id, so its capacity depends
    on the row width: at the full 262 144 entries dim 512 fits the 12 MiB array, dim 640 does not."""
    rows, S, W = 64, 1, 8
    B = K.NOALLOC_MAX_ENTRIES // W
    ids = [torch.zeros(B, dtype=torch.int64, device="cuda") for _ in range(W)]
    for dim, fits in ((512, True), (640, False)):
        table = torch.zeros(rows, dim, dtype=torch.float32, device="cuda")
        grad = torch.zeros(B, dim, dtype=torch.float32, device="cuda")  # one buffer stands in for every worker's rows
        assert (int(b200ps._lib.lib.b200ps_sparse_workspace_bytes(W * B, dim)) <= 12 << 20) == fits
        if fits:
            K.sparse_push_sgd_noalloc([table], ids, [grad] * W, B, K.table_geom(rows, dim, S), 0.1)
            torch.cuda.synchronize()
            assert not table.any()  # zero gradients: the (hot, chunk-folded) row stays zero
        else:
            with pytest.raises(b200ps.B200PSError, match="cannot hold") as ei:
                K.sparse_push_sgd_noalloc([table], ids, [grad] * W, B, K.table_geom(rows, dim, S), 0.1)
            assert ei.value.status == b200ps._lib.ERR_RANGE
