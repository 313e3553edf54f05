"""world_size-2 (and 4) runs of the N>1 host logic on CPU with gloo: placement, arena layout,
per-rank launch plans and the push / pull data movement they imply.  The device kernel is stood
in for by the oracle here (test infrastructure), the 'NVLink' by gloo all_gather; what is under
test is b200ps.build_plan -- the same plan ParameterServer executes with peer pointers."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import b200ps


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, sizes, S, W, opt, steps, out_dir, aggregate="flat"):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import c_oracle as O

    smap = b200ps.ShardMap(sizes, S)
    plan = b200ps.build_plan(smap, W, rank, world, aggregate)
    P, g0 = plan.P, plan.first_grad_slot
    inv_w = np.float32(1.0) / np.float32(W)
    arena = np.zeros(plan.arena_elems, np.float32)
    arena[:P] = O.fill_uniform(P, 1, 0, 0)                         # replica
    state = {L.shard: (np.zeros(L.end - L.begin, np.float32), np.zeros(L.end - L.begin, np.float32)) for L in plan.launches}
    for t in range(1, steps + 1):
        for i, w in enumerate(plan.local_workers):                 # local workers write their gradients
            arena[P * (g0 + i): P * (g0 + i + 1)] = O.fill_uniform(P, 2, t, w)
        if plan.two_level and plan.local_workers:                  # local half of the aggregation
            arena[P:2 * P] = O.dense_fold([arena[P * (g0 + i): P * (g0 + i + 1)].copy() for i in range(len(plan.local_workers))])
        # "peer memory": every rank can read every arena
        sizes_all = [None] * world
        dist.all_gather_object(sizes_all, arena.size)
        arenas = [torch.zeros(n, dtype=torch.float32) for n in sizes_all]
        dist.all_gather(arenas, torch.from_numpy(arena)) if len(set(sizes_all)) == 1 else None
        if len(set(sizes_all)) != 1:                               # uneven worker counts per rank
            objs = [None] * world
            dist.all_gather_object(objs, arena)
            arenas = [torch.from_numpy(o) for o in objs]
        arenas = [a.numpy() for a in arenas]
        updated = {}
        for L in plan.launches:                                    # this rank's kernels
            n = L.end - L.begin
            grads = [np.ascontiguousarray(arenas[r][off + L.begin: off + L.end]) for r, off in L.grad_src]
            p = arena[L.begin:L.end]
            if opt == "adam":
                m, v = state[L.shard]
                O.dense_adam(p, m, v, grads, O.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t), inv_w=inv_w) if n else None
            else:
                O.dense_sgd(p, grads, 0.05, inv_w=inv_w) if n else None
            updated[(L.begin, L.end)] = (p.copy(), L.mirror_ranks)
            assert sorted(L.mirror_ranks + [rank]) == list(range(world))
        allup = [None] * world                                     # mirror stores
        dist.all_gather_object(allup, updated)
        for r, up in enumerate(allup):
            for (b, e), (vals, mirrors) in up.items():
                if rank in mirrors:
                    arena[b:e] = vals
    np.save(os.path.join(out_dir, f"replica{rank}.npy"), arena[:P])
    dist.destroy_process_group()


@pytest.mark.parametrize("world,S,W,opt,aggregate", [(2, 2, 4, "adam", "flat"), (2, 1, 2, "sgd", "flat"), (2, 3, 3, "adam", "flat"),
                                                     (4, 2, 4, "sgd", "flat"), (2, 2, 8, "adam", "two_level"),
                                                     (4, 3, 6, "sgd", "two_level")])
def test_plan_executes_to_the_single_process_result(tmp_path, oracle, world, S, W, opt, aggregate):
    sizes, steps = [5000, 77, 3000], 3
    mp.spawn(_worker, args=(world, _free_port(), sizes, S, W, opt, steps, str(tmp_path), aggregate), nprocs=world, join=True)
    P = b200ps.ShardMap(sizes, S).total
    p = oracle.fill_uniform(P, 1, 0, 0)
    m, v = np.zeros(P, np.float32), np.zeros(P, np.float32)
    inv_w = np.float32(1.0) / np.float32(W)
    for t in range(1, steps + 1):
        g = [oracle.fill_uniform(P, 2, t, w) for w in range(W)]
        if aggregate == "two_level":
            g = oracle.two_level_partials(g, world)
        if opt == "adam":
            oracle.dense_adam(p, m, v, g, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t), inv_w=inv_w)
        else:
            oracle.dense_sgd(p, g, 0.05, inv_w=inv_w)
    for r in range(world):
        got = np.load(tmp_path / f"replica{r}.npy")
        assert np.array_equal(got.view(np.uint32), p.view(np.uint32)), f"rank {r} replica differs"


def test_plan_layout():
    smap = b200ps.ShardMap(b200ps.synth.mlp_c1_variables(), 2)
    P = smap.total
    plans = [b200ps.build_plan(smap, 4, r, 2) for r in range(2)]
    assert [p.local_workers for p in plans] == [[0, 2], [1, 3]]
    assert [p.local_shards for p in plans] == [[0], [1]]
    assert all(p.arena_elems == 3 * P for p in plans)
    # every worker's gradient is read exactly once per slice, from the rank that holds it
    for p in plans:
        (L,) = p.launches
        assert L.grad_src == [(0, P), (1, P), (0, 2 * P), (1, 2 * P)]
    # a single GPU holds everything
    solo = b200ps.build_plan(smap, 4, 0, 1)
    assert solo.local_workers == [0, 1, 2, 3] and solo.local_shards == [0, 1] and solo.launches[0].mirror_ranks == []
    # C1 on 4 GPUs: shards only on GPUs 0 and 1; GPUs 2, 3 are pure workers
    p4 = [b200ps.build_plan(smap, 4, r, 4) for r in range(4)]
    assert [len(p.launches) for p in p4] == [1, 1, 0, 0]
    with pytest.raises(ValueError):
        b200ps.build_plan(smap, 4, 2, 2)
    # ranks may host different numbers of workers: every rank must know every arena's size (barrier flags sit behind it)
    odd = [b200ps.build_plan(smap, 5, r, 2) for r in range(2)]
    assert [p.arena_elems for p in odd] == [4 * P, 3 * P] and all(p.arena_elems_of(r) == odd[r].arena_elems for p in odd for r in range(2))
    # two-level: one partial sum per GPU is folded, arena gains the partial slot
    t2 = [b200ps.build_plan(smap, 8, r, 2, "two_level") for r in range(2)]
    assert all(p.arena_elems == (2 + 4) * P for p in t2) and t2[0].launches[0].grad_src == [(0, P), (1, P)]
    assert t2[1].grad_offset(3) == (1, 3 * P)
    # on one GPU two-level IS the flat fold
    one = b200ps.build_plan(smap, 4, 0, 1, "two_level")
    assert not one.two_level and one.arena_elems == solo.arena_elems and one.launches[0].grad_src == solo.launches[0].grad_src
