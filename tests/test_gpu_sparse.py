"""GPU parity of the sparse PS path (a10..a12) against the oracle, through the C ABI.
Integer outputs (sorted ids, permutation, segment ids, unique ids, offsets) and the updated rows
are BIT-EXACT."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

import b200ps
from b200ps import kernels as K

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def make_ids(rng, kind, W, B, rows):
    if kind == "uniform":
        return rng.integers(0, rows, (W, B))
    if kind == "zipf":
        return np.minimum(rng.zipf(1.05, (W, B)) - 1, rows - 1)
    if kind == "one_hot_row":
        return np.full((W, B), rows // 2)
    if kind == "distinct":
        return rng.permutation(rows)[: W * B].reshape(W, B)
    raise ValueError(kind)


def shard_views(table, rows, S):
    """Per-shard tensors over one contiguous table (div partition)."""
    return [table[slice(*b200ps.row_range(rows, S, s))] for s in range(S)]


@pytest.mark.parametrize("rows,S,dim,W,B,kind", [
    (10, 1, 4, 1, 4, "uniform"),
    (1000, 3, 64, 4, 257, "uniform"),
    (100_000, 8, 64, 8, 4096, "uniform"),
    (100_000, 8, 64, 8, 4096, "zipf"),
    (5_000, 4, 8, 2, 3000, "one_hot_row"),
    (70_000, 5, 128, 3, 5000, "distinct"),
    (300, 7, 16, 5, 1, "uniform"),
    (1 << 20, 2, 32, 2, 50_000, "zipf"),
])
def test_sort_segment_bit_exact(oracle, rows, S, dim, W, B, kind):
    rng = np.random.default_rng(rows + W + B)
    ids = make_ids(rng, kind, W, B, rows).astype(np.int64)
    dids = [dev(ids[w]) for w in range(W)]
    ws = K.sparse_workspace(W * B, dim)
    geom = K.table_geom(rows, dim, S)
    K.sparse_sort_segment(dids, B, geom, ws)
    plan = K.sparse_plan_read(ws, W * B, rows)
    s, perm, seg, uniq, off = oracle.sort_segment(ids.reshape(-1))
    assert plan["n_kept"] == W * B and plan["n_unique"] == len(uniq) and plan["n_bad"] == 0
    assert np.array_equal(plan["sorted_ids"], s)
    assert np.array_equal(plan["perm"], perm)
    assert np.array_equal(plan["seg_id"], seg)
    assert np.array_equal(plan["unique_ids"], uniq)
    assert np.array_equal(plan["seg_offset"], off)


@pytest.mark.parametrize("opt", ["sgd", "adagrad"])
@pytest.mark.parametrize("rows,S,dim,W,B,kind", [
    (1000, 3, 64, 4, 257, "uniform"),
    (50_000, 8, 64, 8, 2048, "zipf"),
    (2_000, 4, 8, 2, 1500, "one_hot_row"),
    (20_000, 5, 128, 3, 2000, "distinct"),
    (64, 2, 4, 2, 100, "uniform"),
])
def test_push_apply_bit_exact_over_steps(oracle, opt, rows, S, dim, W, B, kind):
    rng = np.random.default_rng(7 * rows + B)
    table = oracle.fill_uniform(rows * dim, 1, 0, 0).reshape(rows, dim)
    accum = np.full((rows, dim), 0.1, np.float32)
    dt, da = dev(table), dev(accum)
    geom = K.table_geom(rows, dim, S)
    ws = K.sparse_workspace(W * B, dim)
    inv_w = np.float32(1.0) / np.float32(W)
    for step in range(3):
        ids = make_ids(rng, kind, W, B, rows).astype(np.int64)
        grad = oracle.fill_uniform(W * B * dim, 2, step + 1, 0).reshape(W * B, dim)
        dids = [dev(ids[w]) for w in range(W)]
        dgr = [dev(grad[w * B:(w + 1) * B]) for w in range(W)]
        K.sparse_sort_segment(dids, B, geom, ws)
        if opt == "sgd":
            K.sparse_apply_sgd(shard_views(dt, rows, S), dgr, B, geom, 0.05, ws)
            oracle.sparse_apply_sgd(table, 0, grad, ids.reshape(-1), inv_w, np.float32(0.05))
        else:
            K.sparse_apply_adagrad(shard_views(dt, rows, S), shard_views(da, rows, S), dgr, B, geom, 0.05, ws)
            oracle.sparse_apply_adagrad(table, accum, 0, grad, ids.reshape(-1), inv_w, np.float32(0.05))
    assert np.array_equal(dt.cpu().numpy().view(np.uint32), table.view(np.uint32))
    if opt == "adagrad":
        assert np.array_equal(da.cpu().numpy().view(np.uint32), accum.view(np.uint32))


@pytest.mark.parametrize("rows,S,dim,n", [(10, 1, 4, 7), (1000, 3, 64, 5000), (100_003, 8, 64, 100_000), (777, 5, 256, 300)])
def test_gather_exact_copies(oracle, rows, S, dim, n):
    rng = np.random.default_rng(n)
    table = oracle.fill_uniform(rows * dim, 1, 0, 0).reshape(rows, dim)
    ids = rng.integers(0, rows, n).astype(np.int64)
    dt = dev(table)
    out = torch.empty(n, dim, dtype=torch.float32, device="cuda")
    K.sparse_gather(out, shard_views(dt, rows, S), K.table_geom(rows, dim, S), dev(ids))
    assert np.array_equal(out.cpu().numpy().view(np.uint32), oracle.gather(table, ids).view(np.uint32))


def test_partition_keeps_only_this_gpus_shards(oracle):
    """world > 1 on one device: each 'rank' keeps exactly the ids of its shards, in order; the
    union of the ranks' updates equals the single-rank update."""
    rows, S, dim, W, B, world = 10_000, 8, 16, 4, 3000, 4
    rng = np.random.default_rng(11)
    ids = rng.integers(0, rows, (W, B)).astype(np.int64)
    grad = oracle.fill_uniform(W * B * dim, 2, 1, 0).reshape(W * B, dim)
    table = oracle.fill_uniform(rows * dim, 1, 0, 0).reshape(rows, dim)
    dt = dev(table)
    dids, dgr = [dev(ids[w]) for w in range(W)], [dev(grad[w * B:(w + 1) * B]) for w in range(W)]
    ws = K.sparse_workspace(W * B, dim)
    owners = np.array([oracle.row_owner(rows, S, int(i)) for i in ids.reshape(-1)])
    total_kept = 0
    for rank in range(world):
        geom = K.table_geom(rows, dim, S, rank, world)
        K.sparse_sort_segment(dids, B, geom, ws)
        plan = K.sparse_plan_read(ws, W * B, rows)
        mine = np.flatnonzero(owners % world == rank)
        s, perm, seg, uniq, off = oracle.sort_segment(ids.reshape(-1)[mine])
        assert plan["n_kept"] == len(mine)
        assert np.array_equal(plan["sorted_ids"], s) and np.array_equal(plan["perm"], mine[perm])
        assert np.array_equal(plan["seg_id"], seg) and np.array_equal(plan["seg_offset"], off)
        total_kept += plan["n_kept"]
        K.sparse_apply_sgd(shard_views(dt, rows, S), dgr, B, geom, 0.1, ws)
    assert total_kept == W * B
    oracle.sparse_apply_sgd(table, 0, grad, ids.reshape(-1), np.float32(0.25), np.float32(0.1))
    assert np.array_equal(dt.cpu().numpy().view(np.uint32), table.view(np.uint32))


def test_out_of_range_ids_are_dropped_and_counted(oracle):
    rows, S, dim = 100, 2, 4
    ids = np.array([[5, -1, 100, 7, 5, 1 << 40]], np.int64)
    ws = K.sparse_workspace(ids.size, dim)
    K.sparse_sort_segment([dev(ids[0])], ids.shape[1], K.table_geom(rows, dim, S), ws)
    plan = K.sparse_plan_read(ws, ids.size, rows)
    assert plan["n_bad"] == 3 and plan["sorted_ids"].tolist() == [5, 5, 7] and plan["perm"].tolist() == [0, 4, 3]
    out = torch.full((6, dim), 9.0, device="cuda")
    t = torch.ones(rows, dim, device="cuda")
    K.sparse_gather(out, shard_views(t, rows, S), K.table_geom(rows, dim, S), dev(ids[0]))
    assert out.sum(1).tolist() == [4.0, 0.0, 0.0, 4.0, 4.0, 0.0]


def test_argument_errors():
    ws = K.sparse_workspace(8, 4)
    ids = torch.zeros(8, dtype=torch.int64, device="cuda")
    with pytest.raises(b200ps.B200PSError):
        K.sparse_sort_segment([ids], 8, K.table_geom(100, 6, 1), ws)  # dim not a multiple of 4
    with pytest.raises(b200ps.B200PSError):
        K.sparse_sort_segment([ids], 8, K.table_geom(1 << 33, 4, 1), ws)  # rows >= 2^32
    big = torch.zeros(4096, dtype=torch.int64, device="cuda")
    with pytest.raises(b200ps.B200PSError, match="too small"):
        K.sparse_sort_segment([big] * 4, 4096, K.table_geom(100, 4, 1), ws)
    with pytest.raises(b200ps.B200PSError):
        K.sparse_sort_segment([ids], 8, K.table_geom(100, 4, 2, 2, 2), ws)  # rank outside world


def test_c2_shapes_full_batch(oracle):
    """BASELINE configs[2] geometry at a table size the oracle sorts in seconds: 8 shards, 8
    workers, 262 144 ids per worker, dim 64, Zipf(1.05) ids; sortedness / conservation properties
    at full batch plus bit-exact rows."""
    rows, S, dim, W, B = 1_000_000, 8, 64, 8, 262_144
    rng = np.random.default_rng(5)
    ids = np.minimum(rng.zipf(1.05, (W, B)) - 1, rows - 1).astype(np.int64)
    dids = [dev(ids[w]) for w in range(W)]
    geom = K.table_geom(rows, dim, S)
    ws = K.sparse_workspace(W * B, dim)
    K.sparse_sort_segment(dids, B, geom, ws)
    plan = K.sparse_plan_read(ws, W * B, rows)
    s = plan["sorted_ids"]
    assert plan["n_kept"] == W * B and (np.diff(s) >= 0).all()                      # sorted
    assert np.array_equal(np.sort(plan["perm"]), np.arange(W * B, dtype=np.uint32))   # a permutation
    assert np.array_equal(ids.reshape(-1)[plan["perm"]], s)                           # of the input
    same = s[1:] == s[:-1]
    assert (np.diff(plan["perm"].astype(np.int64))[same] > 0).all()                   # stable
    assert plan["n_unique"] == len(np.unique(ids)) and plan["seg_offset"][-1] == W * B
    table = oracle.fill_uniform(rows * dim, 1, 0, 0, threads=8).reshape(rows, dim)
    grad = oracle.fill_uniform(W * B * dim, 2, 1, 0, threads=8).reshape(W * B, dim)
    dt = dev(table)
    dgr = [dev(grad[w * B:(w + 1) * B]) for w in range(W)]
    K.sparse_apply_sgd(shard_views(dt, rows, S), dgr, B, geom, 0.05, ws)
    oracle.sparse_apply_sgd(table, 0, grad, ids.reshape(-1), np.float32(1.0) / np.float32(W), np.float32(0.05))
    assert np.array_equal(dt.cpu().numpy().view(np.uint32), table.view(np.uint32))


@pytest.mark.parametrize("kind", ["uniform", "zipf"])
def test_c2_at_baseline_size(oracle, kind):
    """BASELINE configs[2] at its real size -- 10 M x 64 table over 8 PS shards, 8 workers, 262 144
    ids per worker per step, the ids bench.py uses (synth.make_ids) -- one pull and one push:
    every integer of the plan, the pulled rows of a worker and the whole updated table are
    bit-exact against the oracle."""
    from b200ps import synth

    rows, S, dim, W, B = 10_000_000, 8, 64, 8, 262_144
    ids = synth.make_ids(kind, rows, W, B, 1)
    flat = ids.reshape(-1)
    geom = K.table_geom(rows, dim, S)
    dids = [dev(ids[w]) for w in range(W)]
    nt = min(16, len(__import__("os").sched_getaffinity(0)))
    table = oracle.fill_uniform(rows * dim, 1, 0, 0, threads=nt).reshape(rows, dim)
    dt = K.fill_uniform(torch.empty(rows * dim, dtype=torch.float32, device="cuda"), 1, 0, 0).view(rows, dim)
    shards = shard_views(dt, rows, S)
    # pull (a10): worker 3's rows
    out = torch.empty(B, dim, dtype=torch.float32, device="cuda")
    K.sparse_gather(out, shards, geom, dids[3])
    assert np.array_equal(out.cpu().numpy().view(np.uint32), oracle.gather(table, ids[3]).view(np.uint32))
    # push (a11): the plan
    ws = K.sparse_workspace(W * B, dim)
    K.sparse_sort_segment(dids, B, geom, ws)
    plan = K.sparse_plan_read(ws, W * B, rows)
    s, perm, seg, uniq, off = oracle.sort_segment(flat)
    assert plan["n_kept"] == W * B and plan["n_bad"] == 0 and plan["n_unique"] == len(uniq)
    assert np.array_equal(plan["sorted_ids"], s) and np.array_equal(plan["perm"], perm)
    assert np.array_equal(plan["seg_id"], seg) and np.array_equal(plan["unique_ids"], uniq)
    assert np.array_equal(plan["seg_offset"], off)
    # push (a12): the updated table
    grad = oracle.fill_uniform(W * B * dim, 2, 1, 0, threads=nt).reshape(W * B, dim)
    dg = K.fill_uniform(torch.empty(W * B * dim, dtype=torch.float32, device="cuda"), 2, 1, 0).view(W, B, dim)
    K.sparse_apply_sgd(shards, [dg[w] for w in range(W)], B, geom, 0.05, ws)
    oracle.sparse_apply_sgd(table, 0, grad, flat, np.float32(1.0) / np.float32(W), np.float32(0.05))
    got = dt.cpu().numpy()
    assert np.array_equal(got.view(np.uint32), table.view(np.uint32))


@pytest.mark.parametrize("opt,kind", [("sgd", "zipf"), ("adagrad", "uniform")])
def test_embedding_table_host_api(oracle, opt, kind):
    """b200ps.EmbeddingTable (world = 1): pull then push for 2 steps equals the oracle bit for bit."""
    rows, dim, S, W, B = 40_000, 64, 8, 8, 2048
    tab = b200ps.EmbeddingTable(rows, dim, S, W, B, optimizer=opt, lr=0.05)
    tab.init_uniform(1)
    table = oracle.fill_uniform(rows * dim, 1, 0, 0).reshape(rows, dim)
    accum = np.full((rows, dim), 0.1, np.float32)
    out = torch.empty(B, dim, dtype=torch.float32, device="cuda")
    for step in (1, 2):
        ids = b200ps.synth.make_ids(kind, rows, W, B, step)
        grad = oracle.fill_uniform(W * B * dim, 2, step, 0).reshape(W, B, dim)
        for w in range(W):
            tab.ids[w].copy_(torch.from_numpy(ids[w]))
            tab.grad_rows[w].copy_(torch.from_numpy(grad[w]))
        for w in (0, W - 1):
            tab.pull(w, out)
            assert np.array_equal(out.cpu().numpy().view(np.uint32), oracle.gather(table, ids[w]).view(np.uint32))
        tab.pull_step()  # all workers in one launch
        for w in range(W):
            assert np.array_equal(tab.pulled[w].cpu().numpy().view(np.uint32), oracle.gather(table, ids[w]).view(np.uint32))
        tab.push_step()
        if opt == "sgd":
            oracle.sparse_apply_sgd(table, 0, grad.reshape(-1, dim), ids.reshape(-1), np.float32(1.0) / np.float32(W), np.float32(0.05))
        else:
            oracle.sparse_apply_adagrad(table, accum, 0, grad.reshape(-1, dim), ids.reshape(-1), np.float32(1.0) / np.float32(W),
                                        np.float32(0.05))
    got = np.concatenate([v for _, (_, v) in sorted(tab.dump().items())])
    assert np.array_equal(got.view(np.uint32), table.view(np.uint32))
    assert tab.plan()["n_kept"] == W * B
    tab.close()


def test_embedding_table_graph_replay_matches(oracle):
    """One captured pull+push step replayed on new ids / gradient rows gives the oracle's bits."""
    rows, dim, S, W, B = 20_000, 64, 4, 4, 1024
    tab = b200ps.EmbeddingTable(rows, dim, S, W, B, optimizer="sgd", lr=0.05)
    tab.init_uniform(1)
    table = oracle.fill_uniform(rows * dim, 1, 0, 0).reshape(rows, dim)

    def load(step):
        ids = b200ps.synth.make_ids("zipf", rows, W, B, step)
        grad = oracle.fill_uniform(W * B * dim, 2, step, 0).reshape(W, B, dim)
        for w in range(W):
            tab.ids[w].copy_(torch.from_numpy(ids[w]))
            tab.grad_rows[w].copy_(torch.from_numpy(grad[w]))
        return ids, grad

    ids, grad = load(1)
    g = tab.capture_step()  # runs one real step (warm-up), then records one
    oracle.sparse_apply_sgd(table, 0, grad.reshape(-1, dim), ids.reshape(-1), np.float32(0.25), np.float32(0.05))
    for step in (2, 3):
        ids, grad = load(step)
        torch.cuda.synchronize()
        g.replay()
        torch.cuda.synchronize()
        for w in range(W):
            assert np.array_equal(tab.pulled[w].cpu().numpy().view(np.uint32), oracle.gather(table, ids[w]).view(np.uint32))
        oracle.sparse_apply_sgd(table, 0, grad.reshape(-1, dim), ids.reshape(-1), np.float32(0.25), np.float32(0.05))
    got = np.concatenate([v for _, (_, v) in sorted(tab.dump().items())])
    assert np.array_equal(got.view(np.uint32), table.view(np.uint32))
    tab.close()


@pytest.mark.parametrize("dim,world", [(4, 2), (8, 3), (32, 2), (128, 3), (256, 4)])
def test_serve_and_route_with_emulated_ranks(oracle, dim, world):
    """The multi-GPU routing on ONE device: every 'rank' runs serve over all workers' ids and the
    union fills every worker's rows; every worker's 'GPU' routes the rows other ranks own into their
    staging buffers and each rank's sort + apply reads local-or-staged rows only.  All lane-group
    widths (dim 4 .. 256) and an out-of-range id are exercised; result = the oracle's full update."""
    rows, S, W, B = 5000, 6, 4, 700
    rng = np.random.default_rng(dim * 10 + world)
    ids = rng.integers(0, rows, (W, B)).astype(np.int64)
    ids[1, 5] = -3  # out of range: pulled as zeros, dropped by the push
    table = oracle.fill_uniform(rows * dim, 1, 0, 0).reshape(rows, dim)
    grad = oracle.fill_uniform(W * B * dim, 2, 1, 0).reshape(W, B, dim)
    dt = dev(table)
    dids, dgr = [dev(ids[w]) for w in range(W)], [dev(grad[w]) for w in range(W)]
    my_shards = lambda r: [t if s % world == r else None for s, t in enumerate(shard_views(dt, rows, S))]
    # pull
    outs = [torch.full((B, dim), float("nan"), device="cuda") for _ in range(W)]
    for r in range(world):
        K.sparse_serve(outs, my_shards(r), dids, B, K.table_geom(rows, dim, S, r, world))
    for w in range(W):
        want = oracle.gather(table, np.where(ids[w] < 0, 0, ids[w]))
        want[ids[w] < 0] = 0.0
        assert np.array_equal(outs[w].cpu().numpy().view(np.uint32), want.view(np.uint32))
    # push
    staging = {(o, w): torch.full((B, dim), float("nan"), device="cuda") for o in range(world) for w in range(W) if w % world != o}
    for w in range(W):
        rw = w % world
        K.sparse_route(dgr[w], dids[w], K.table_geom(rows, dim, S, rw, world), [None if o == rw else staging[(o, w)] for o in range(world)])
    ws = K.sparse_workspace(W * B, dim)
    for r in range(world):
        geom = K.table_geom(rows, dim, S, r, world)
        rows_for = [dgr[w] if w % world == r else staging[(r, w)] for w in range(W)]
        K.sparse_sort_segment(dids, B, geom, ws)
        K.sparse_apply_sgd(my_shards(r), rows_for, B, geom, 0.05, ws)
    keep = ids.reshape(-1) >= 0
    oracle.sparse_apply_sgd(table, 0, np.ascontiguousarray(grad.reshape(-1, dim)[keep]), ids.reshape(-1)[keep],
                            np.float32(1.0) / np.float32(W), np.float32(0.05))
    assert np.array_equal(dt.cpu().numpy().view(np.uint32), table.view(np.uint32))


def test_embedding_table_checkpoint_resume(tmp_path):
    """save -> new table -> load -> one more push equals the uninterrupted run (rows and accumulators)."""
    rows, dim, S, W, B = 9_000, 16, 3, 2, 500

    def feed(tab, step):
        ids = b200ps.synth.make_ids("uniform", rows, W, B, step)
        for w in range(W):
            tab.ids[w].copy_(torch.from_numpy(ids[w]))
            K.fill_uniform(tab.grad_rows[w].view(-1), 2, step, w)

    a = b200ps.EmbeddingTable(rows, dim, S, W, B, optimizer="adagrad")
    a.init_uniform(1)
    feed(a, 1)
    a.push_step()
    a.save(str(tmp_path))
    feed(a, 2)
    a.push_step()
    b = b200ps.EmbeddingTable(rows, dim, S, W, B, optimizer="adagrad")
    b.load(str(tmp_path))
    feed(b, 2)
    b.push_step()
    for s in range(S):
        assert torch.equal(a.shards[s], b.shards[s]) and torch.equal(a.accums[s], b.accums[s])
    c = b200ps.EmbeddingTable(rows, dim, S, W, B, optimizer="sgd")
    with pytest.raises(ValueError):
        c.load(str(tmp_path))
    for t in (a, b, c):
        t.close()


def test_apply_refuses_a_workspace_that_holds_another_push(oracle):
    """Advisor finding (round 1): apply carves the workspace from its own (W, ids_per_worker) -- given the
    plan of a different push it used to read garbage offsets.  sort_segment now tags the plan and apply
    checks the tag on the device: a mismatched call updates nothing and the next plan_read reports it."""
    rows, S, dim, W, B = 5000, 2, 16, 2, 300
    rng = np.random.default_rng(11)
    ids = [dev(rng.integers(0, rows, B)) for _ in range(W)]
    grads = [torch.ones(B, dim, dtype=torch.float32, device="cuda") for _ in range(W)]
    table = K.fill_uniform(torch.empty(rows * dim, dtype=torch.float32, device="cuda"), 1, 0, 0).view(rows, dim)
    before = table.clone()
    geom = K.table_geom(rows, dim, S)
    ws = K.sparse_workspace(W * B, dim)
    K.sparse_sort_segment(ids, B, geom, ws)
    K.sparse_apply_sgd(shard_views(table, rows, S), [g[:B - 1] for g in grads], B - 1, geom, 0.05, ws)  # another push's shape
    torch.cuda.synchronize()
    assert torch.equal(table, before)
    with pytest.raises(b200ps.B200PSError) as ei:
        K.sparse_plan_read(ws, W * B, rows)
    assert ei.value.status == b200ps._lib.ERR_STATE
    K.sparse_sort_segment(ids, B, geom, ws)  # a fresh plan clears the mark; the matching apply works
    K.sparse_apply_sgd(shard_views(table, rows, S), grads, B, geom, 0.05, ws)
    K.sparse_plan_read(ws, W * B, rows)
    assert not torch.equal(table, before)
