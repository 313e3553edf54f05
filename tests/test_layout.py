"""Shard maps and row partitioning are bit-exact against the oracle."""
import pytest
from hypothesis import given, settings, strategies as st

from b200ps import B200PSError, ShardMap, row_owner, row_range, synth


def test_kats():
    m = ShardMap([10], 3, align=1)
    assert (m.total, m.slices(), m.var_offsets()) == (10, [(0, 4), (4, 8), (8, 10)], [0])
    assert [row_range(10, 4, s) for s in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert [row_owner(10, 4, i) for i in range(10)] == [0, 0, 0, 1, 1, 1, 2, 2, 3, 3]


@settings(max_examples=80, deadline=None)
@given(sizes=st.lists(st.integers(0, 5000), min_size=0, max_size=40), S=st.integers(1, 9), align=st.sampled_from([1, 4, 256, 1024]))
def test_shard_map_matches_oracle(oracle, nporacle, sizes, S, align):
    m = ShardMap(sizes, S, align)
    got = (m.total, m.var_offsets(), m.slices())
    assert got == oracle.shard_map(sizes, S, align) if sizes else got[0] == 0
    if sizes:
        assert got == nporacle.shard_map(sizes, S, align)
    sl = m.slices()
    assert sl[0][0] == 0 and sl[-1][1] == m.total
    assert all(a[1] == b[0] for a, b in zip(sl, sl[1:]))          # contiguous, disjoint, covering
    assert all(b % align == 0 for b, _ in sl) and all(o % align == 0 for o in m.var_offsets())


@settings(max_examples=60, deadline=None)
@given(rows=st.integers(1, 10 ** 7), S=st.integers(1, 16), data=st.data())
def test_rows_match_oracle(oracle, rows, S, data):
    for s in range(S):
        assert row_range(rows, S, s) == oracle.row_range(rows, S, s)
    for _ in range(20):
        i = data.draw(st.integers(0, rows - 1))
        assert row_owner(rows, S, i) == oracle.row_owner(rows, S, i)


def test_baseline_configs():
    # C1: 100 685 824 params over 2 shards; C3: 161 ResNet-50 variables; C2: 10 M rows over 8 shards
    c1 = synth.mlp_c1_variables()
    assert sum(c1) == 100_685_824 and len(c1) == 8
    assert ShardMap(c1, 2).slices() == [(0, 50_342_912), (50_342_912, 100_685_824)]
    c3 = synth.resnet50_c3_variables()
    assert (len(c3), sum(c3), min(c3), max(c3)) == (161, 25_557_032, 64, 2_359_296)
    m = ShardMap(c3, 8)
    assert m.total >= sum(c3) and m.total % 1024 == 0 and m.slices()[-1][1] == m.total
    assert [row_range(10_000_000, 8, s) for s in (0, 7)] == [(0, 1_250_000), (8_750_000, 10_000_000)]
    assert synth.sweep_c4_sizes()[0] * 4 == 1024 and synth.sweep_c4_sizes()[-1] * 4 == 1 << 30


def test_resnet50_shapes_match_torchvision():
    tv = pytest.importorskip("torchvision")
    net = tv.models.resnet50(weights=None)
    assert [p.numel() for p in net.parameters()] == synth.resnet50_c3_variables()


def test_errors():
    with pytest.raises(B200PSError):
        ShardMap([1], 0)
    with pytest.raises(B200PSError):
        row_owner(10, 4, 10)
    with pytest.raises(B200PSError):
        row_range(10, 4, 4)
    m = ShardMap([1, 2], 2)
    with pytest.raises(B200PSError):
        m.slice(2)
    with pytest.raises(B200PSError):
        m.var_offset(2)
