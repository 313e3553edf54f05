"""Arena layout of the sparse path: pure arithmetic every rank must agree on (no GPU)."""
import b200ps


def test_layout_c2_two_gpus():
    L = b200ps.TableLayout(10_000_000, 64, 8, 8, 262_144, 2, adagrad=True)
    shard_bytes = 1_250_000 * 64 * 4
    assert L.shard_range[0] == (0, 1_250_000) and L.shard_range[7] == (8_750_000, 10_000_000)
    assert [L.shard_rank(s) for s in range(8)] == [0, 1] * 4 and [L.worker_rank(w) for w in range(8)] == [0, 1] * 4
    # per rank: 4 shards, 4 accumulators, then 4 x (ids, gradient rows); everything 256-byte aligned
    assert L.shard_off[0] == 0 and L.shard_off[2] == shard_bytes and L.shard_off[1] == 0
    assert L.accum_off[0] == 4 * shard_bytes and L.ids_off[0] == 8 * shard_bytes
    assert L.grad_off[0] == L.ids_off[0] + 262_144 * 8
    rows_bytes = 262_144 * 64 * 4
    assert L.out_off[0] == L.grad_off[0] + rows_bytes and L.ids_off[2] == L.out_off[0] + rows_bytes
    # after the 4 local workers: one staging slot per remote worker (1, 3, 5, 7 on rank 0)
    assert L.stage_off[(0, 1)] == L.out_off[6] + rows_bytes and sorted(w for r, w in L.stage_off if r == 0) == [1, 3, 5, 7]
    assert L.flag_off[0] == L.stage_off[(0, 7)] + rows_bytes and L.arena_bytes[0] == L.arena_bytes[1] == L.flag_off[0] + 256
    assert all(v % 256 == 0 for d in (L.shard_off, L.accum_off, L.ids_off, L.grad_off, L.out_off, L.stage_off) for v in d.values())


def test_layout_uneven():
    L = b200ps.TableLayout(10, 4, 3, 5, 7, 2, adagrad=False)
    assert L.shard_range == [(0, 4), (4, 7), (7, 10)] and not L.accum_off
    assert sorted(L.ids_off) == [0, 1, 2, 3, 4] and L.arena_bytes[0] > L.arena_bytes[1]  # rank 0: 2 shards, 3 workers


def test_id_generator_is_reproducible_and_in_range():
    a = b200ps.synth.make_ids("zipf", 1000, 3, 500)
    b = b200ps.synth.make_ids("zipf", 1000, 3, 500)
    assert (a == b).all() and a.min() >= 0 and a.max() < 1000 and a.shape == (3, 500)
    u = b200ps.synth.make_ids("uniform", 1000, 3, 500, step=2)
    assert u.min() >= 0 and u.max() < 1000 and not (u == b200ps.synth.make_ids("uniform", 1000, 3, 500, step=3)).all()
    # Zipf is skewed: id 0 is the most frequent
    import numpy as np

    assert np.bincount(a.reshape(-1)).argmax() == 0
