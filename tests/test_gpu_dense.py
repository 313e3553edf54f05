"""GPU parity of the dense PS step (a5..a9) against the oracle: BIT-EXACT (fixed worker order,
one rounding per operation), through the C ABI (ctypes -> libb200ps.so)."""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")

import b200ps
from b200ps import kernels as K

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def bits(t):
    return t.detach().cpu().numpy().view(np.uint32)


def same(t, a):
    return np.array_equal(bits(t), np.ascontiguousarray(a).view(np.uint32))


def test_library_is_the_cuda_one():
    assert torch.cuda.is_available()
    assert b200ps.LIB_PATH.endswith("libb200ps.so")
    assert torch.cuda.get_device_capability(0)[0] >= 10, "built for sm_100a only"


@pytest.mark.parametrize("n,first", [(1, 0), (1000, 5), (1 << 20, 1 << 40), (4097, 123456789)])
def test_fill_uniform_bit_exact(oracle, n, first):
    d = torch.empty(n, dtype=torch.float32, device="cuda")
    K.fill_uniform(d, seed=2, step=7, worker=3, first=first)
    assert same(d, oracle.fill_uniform(n, 2, 7, 3, first=first, threads=4))


@pytest.mark.parametrize("W", [1, 2, 3, 4, 5, 8, 9, 16, 64])
@pytest.mark.parametrize("n", [1, 3, 4, 1024, 4099, 262144 + 5])
def test_sgd_bit_exact(oracle, W, n):
    p = oracle.fill_uniform(n, 1, 0, 0)
    grads = [oracle.fill_uniform(n, 2, 1, w) for w in range(W)]
    dp, dg = dev(p), [dev(g) for g in grads]
    for step in range(3):
        K.dense_apply_sgd(dp, dg, 0.05)
        oracle.dense_sgd(p, grads, 0.05, threads=4)
    assert same(dp, p)


@pytest.mark.parametrize("W", [1, 2, 3, 4, 7, 8, 12])
@pytest.mark.parametrize("n", [1, 5, 2048, 4099, 300001])
def test_adam_bit_exact_10_steps(oracle, W, n):
    p = oracle.fill_uniform(n, 1, 0, 0)
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    dp, dm, dv = dev(p), dev(m), dev(v)
    dg = [torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(W)]
    for t in range(1, 11):
        grads = [oracle.fill_uniform(n, 2, t, w, threads=4) for w in range(W)]
        for w in range(W):
            K.fill_uniform(dg[w], 2, t, w)  # gradients produced on the device by the same hash
        K.dense_apply_adam(dp, dm, dv, dg, K.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t))
        oracle.dense_adam(p, m, v, grads, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t), threads=4)
    assert same(dp, p) and same(dm, m) and same(dv, v)
    # the north-star tolerance (1e-6 relative) holds a fortiori
    assert np.allclose(dp.cpu().numpy(), p, rtol=1e-6, atol=0)


@pytest.mark.parametrize("threads,ctas,unroll", [(128, 8, 1), (256, 4, 2), (512, 2, 4), (1024, 1, 2), (256, 16, 4)])
def test_every_launch_geometry_gives_the_same_bits(oracle, threads, ctas, unroll):
    n, W = 1_000_003, 4
    p = oracle.fill_uniform(n, 1, 0, 0)
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    grads = [oracle.fill_uniform(n, 2, 1, w, threads=4) for w in range(W)]
    dp, dm, dv, dg = dev(p), dev(m), dev(v), [dev(g) for g in grads]
    K.set_dense_tuning(threads, ctas, unroll)
    try:
        K.dense_apply_adam(dp, dm, dv, dg, K.adam_scalars(1e-3, t=1))
        ds = dev(p)
        K.dense_apply_sgd(ds, dg, 0.1)
    finally:
        K.set_dense_tuning(256, 8, 2)
    oracle.dense_adam(p, m, v, grads, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, 1), threads=4)
    assert same(dp, p) and same(dm, m) and same(dv, v)
    assert same(ds, oracle.dense_sgd(oracle.fill_uniform(n, 1, 0, 0), grads, 0.1, threads=4))


def test_unaligned_pointers_take_the_scalar_path(oracle):
    n, W = 10_001, 3
    p = oracle.fill_uniform(n, 1, 0, 0)
    grads = [oracle.fill_uniform(n, 2, 1, w) for w in range(W)]
    base = torch.empty(n + 1, dtype=torch.float32, device="cuda")
    dp = base[1:]  # 4-byte aligned only
    dp.copy_(torch.from_numpy(p))
    dg = [dev(g) for g in grads]
    K.dense_apply_sgd(dp, dg, 0.25)
    assert same(dp, oracle.dense_sgd(p, grads, 0.25))


def test_mirrors_receive_the_updated_slice(oracle):
    # pull fused into the step: the new parameters land in every worker copy
    n, W = 50_000, 2
    p = oracle.fill_uniform(n, 1, 0, 0)
    grads = [oracle.fill_uniform(n, 2, 1, w) for w in range(W)]
    dp, dg = dev(p), [dev(g) for g in grads]
    mirrors = [torch.zeros(n, dtype=torch.float32, device="cuda") for _ in range(3)]
    m, v = torch.zeros_like(dp), torch.zeros_like(dp)
    K.dense_apply_adam(dp, m, v, dg, K.adam_scalars(1e-3, t=1), mirrors=mirrors)
    for mi in mirrors:
        assert torch.equal(mi, dp)
    oracle.dense_adam(p, np.zeros(n, np.float32), np.zeros(n, np.float32), grads, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, 1))
    assert same(dp, p)


def test_sharded_application_equals_whole(oracle):
    """Size-independent property: applying the step slice by slice (any shard count) gives the
    bits of applying it to the whole buffer -- the PS shards are independent."""
    sizes = b200ps.synth.resnet50_c3_variables()
    W = 4
    whole = b200ps.ShardMap(sizes, 1)
    P = whole.total
    p0 = torch.empty(P, dtype=torch.float32, device="cuda")
    K.fill_uniform(p0, 1, 0, 0)
    dg = [K.fill_uniform(torch.empty(P, dtype=torch.float32, device="cuda"), 2, 1, w) for w in range(W)]
    ref = p0.clone()
    rm, rv = torch.zeros_like(ref), torch.zeros_like(ref)
    opt = K.adam_scalars(1e-3, t=1)
    K.dense_apply_adam(ref, rm, rv, dg, opt)
    for S in (2, 4, 8):
        sm = b200ps.ShardMap(sizes, S)
        p, m, v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
        for b, e in sm.slices():
            K.dense_apply_adam(p[b:e], m[b:e], v[b:e], [g[b:e] for g in dg], opt)
        assert torch.equal(p, ref) and torch.equal(m, rm) and torch.equal(v, rv)
    # and the whole thing against the oracle at C3's full size
    hp = oracle.fill_uniform(P, 1, 0, 0, threads=8)
    hg = [oracle.fill_uniform(P, 2, 1, w, threads=8) for w in range(W)]
    hm, hv = np.zeros(P, np.float32), np.zeros(P, np.float32)
    oracle.dense_adam(hp, hm, hv, hg, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, 1), threads=8)
    assert same(ref, hp) and same(rm, hm) and same(rv, hv)


def test_c1_full_size_two_shards_four_workers(oracle):
    """BASELINE configs[1] at full size: 100 685 824 params, 2 PS shards, 4 workers, Adam, 2 steps."""
    sm = b200ps.ShardMap(b200ps.synth.mlp_c1_variables(), 2)
    P, W = sm.total, 4
    nt = oracle.max_threads()
    dp = K.fill_uniform(torch.empty(P, dtype=torch.float32, device="cuda"), 1, 0, 0)
    dm, dv = torch.zeros_like(dp), torch.zeros_like(dp)
    dg = [torch.empty(P, dtype=torch.float32, device="cuda") for _ in range(W)]
    hp = oracle.fill_uniform(P, 1, 0, 0, threads=nt)
    hm, hv = np.zeros(P, np.float32), np.zeros(P, np.float32)
    for t in (1, 2):
        for w in range(W):
            K.fill_uniform(dg[w], 2, t, w)
        hg = [oracle.fill_uniform(P, 2, t, w, threads=nt) for w in range(W)]
        opt = K.adam_scalars(1e-3, t=t)
        for b, e in sm.slices():
            K.dense_apply_adam(dp[b:e], dm[b:e], dv[b:e], [g[b:e] for g in dg], opt)
        oracle.dense_adam(hp, hm, hv, hg, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t), threads=nt)
    assert same(dp, hp) and same(dm, hm) and same(dv, hv)


def test_argument_errors_are_statuses_not_crashes():
    p = torch.zeros(8, dtype=torch.float32, device="cuda")
    with pytest.raises(b200ps.B200PSError):
        K.dense_apply_sgd(p, [], 0.1)  # W = 0
    with pytest.raises(b200ps.B200PSError):
        K.dense_apply_sgd(p, [p] * 65, 0.1)  # W > B200PS_MAX_WORKERS
    with pytest.raises(b200ps.B200PSError):
        K.dense_apply_sgd(p, [p], 0.1, mirrors=[p] * 9)
    with pytest.raises(TypeError):
        K.dense_apply_sgd(torch.zeros(8), [p], 0.1)  # CPU tensor: there is no fallback
    K.dense_apply_sgd(p, [p], 0.1, n=0)  # empty slice is a no-op
    torch.cuda.synchronize()


def test_launch_counter_counts():
    p = torch.zeros(1024, dtype=torch.float32, device="cuda")
    before = K.launch_count()
    K.dense_apply_sgd(p, [p], 0.1)
    K.fill_uniform(p, 1, 1, 1)
    assert K.launch_count() == before + 2


def test_parameter_server_host_api_and_graph_replay(oracle):
    """b200ps.ParameterServer on one GPU (C0 shape: 1 PS + 2 workers, SGD): step(), a captured
    CUDA-graph replay of it, and step_host() with pinned host buffers all match the oracle bit for bit."""
    n, W = 1_000_000, 2
    ps = b200ps.ParameterServer([n], 1, W, optimizer="sgd", lr=0.05)
    ps.init_uniform(1)
    P = ps.plan.P
    hp = oracle.fill_uniform(P, 1, 0, 0, threads=4)
    hg = [oracle.fill_uniform(P, 2, 1, w, threads=4) for w in range(W)]
    for w in range(W):
        K.fill_uniform(ps.grad_buffer(w), 2, 1, w)
    ps.step()
    oracle.dense_sgd(hp, hg, 0.05, threads=4)
    g = ps.capture_step()          # runs one warm-up step, then records one (recording does not execute)
    oracle.dense_sgd(hp, hg, 0.05, threads=4)
    torch.cuda.synchronize()
    g.replay()
    oracle.dense_sgd(hp, hg, 0.05, threads=4)
    torch.cuda.synchronize()
    assert same(ps.replica, hp)
    host_g = {w: torch.from_numpy(hg[w]).pin_memory() for w in range(W)}
    host_o = {0: torch.empty(P, dtype=torch.float32).pin_memory()}
    for _ in range(3):  # back-to-back host steps: copy-out of one overlaps copy-in of the next
        ps.step_host(host_g, host_o)
        oracle.dense_sgd(hp, hg, 0.05, threads=4)
    ps.host_sync()
    assert np.array_equal(host_o[0].numpy().view(np.uint32), hp.view(np.uint32))
    torch.cuda.synchronize()
    assert same(ps.replica, hp)
    ps.close()


def test_parameter_server_adam_two_shards(oracle):
    sizes, S, W = [300_000, 77, 50_000], 2, 4
    ps = b200ps.ParameterServer(sizes, S, W, optimizer="adam", lr=1e-3)
    ps.init_uniform(1)
    P = ps.plan.P
    hp = oracle.fill_uniform(P, 1, 0, 0)
    hm, hv = np.zeros(P, np.float32), np.zeros(P, np.float32)
    for t in (1, 2, 3):
        hg = [oracle.fill_uniform(P, 2, t, w) for w in range(W)]
        for w in range(W):
            K.fill_uniform(ps.grad_buffer(w), 2, t, w)
        ps.step()
        oracle.dense_adam(hp, hm, hv, hg, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t))
    d = ps.dump()
    for s, x in d.items():
        b, e = x["range"]
        assert same(torch.from_numpy(x["p"]), hp[b:e]) and same(torch.from_numpy(x["m"]), hm[b:e]) and same(torch.from_numpy(x["v"]), hv[b:e])
    # Adam replays as a CUDA graph too: the bias-corrected step size is read from device memory
    for w in range(W):
        K.fill_uniform(ps.grad_buffer(w), 2, 4, w)
    g = ps.capture_step()  # step 4 (captured and run)
    hg = [oracle.fill_uniform(P, 2, 4, w) for w in range(W)]
    oracle.dense_adam(hp, hm, hv, hg, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, 4))
    g.replay()             # step 5, same gradients
    oracle.dense_adam(hp, hm, hv, hg, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, 5))
    assert ps.t == 5
    for s, x in ps.dump().items():
        b, e = x["range"]
        assert same(torch.from_numpy(x["p"]), hp[b:e]) and same(torch.from_numpy(x["m"]), hm[b:e]) and same(torch.from_numpy(x["v"]), hv[b:e])
    ps.close()


def test_checkpoint_resume_continues_bit_exactly(tmp_path):
    """save -> new ParameterServer -> load -> keep training gives the bits of never having stopped
    (parameters, m, v and the Adam step counter all travel)."""
    sizes, S, W = [40_000, 123, 9_000], 2, 3

    def grads(ps, t):
        for w in range(W):
            K.fill_uniform(ps.grad_buffer(w), 2, t, w)

    a = b200ps.ParameterServer(sizes, S, W, optimizer="adam")
    a.init_uniform(1)
    for t in (1, 2):
        grads(a, t)
        a.step()
    a.save(str(tmp_path))
    for t in (3, 4):
        grads(a, t)
        a.step()
    b = b200ps.ParameterServer(sizes, S, W, optimizer="adam")
    b.load(str(tmp_path))
    assert b.t == 2
    for t in (3, 4):
        grads(b, t)
        b.step()
    da, db = a.dump(), b.dump()
    for s in da:
        for k in ("p", "m", "v"):
            assert np.array_equal(da[s][k].view(np.uint32), db[s][k].view(np.uint32))
    c = b200ps.ParameterServer(sizes, 1, W, optimizer="adam")
    with pytest.raises((ValueError, FileNotFoundError)):
        c.load(str(tmp_path))  # a different shard map is refused
    for x in (a, b, c):
        x.close()


@pytest.mark.parametrize("W,n", [(1, 5), (4, 4099), (6, 300_001)])
def test_momentum_bit_exact(oracle, W, n):
    p = oracle.fill_uniform(n, 1, 0, 0)
    acc = np.zeros(n, np.float32)
    dp, da = dev(p), dev(acc)
    dg = [torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(W)]
    for t in (1, 2, 3, 4):
        grads = [oracle.fill_uniform(n, 2, t, w, threads=4) for w in range(W)]
        for w in range(W):
            K.fill_uniform(dg[w], 2, t, w)
        K.dense_apply_momentum(dp, da, dg, 0.05, 0.9)
        oracle.dense_momentum(p, acc, grads, 0.05, 0.9, threads=4)
    assert same(dp, p) and same(da, acc)


def test_parameter_server_momentum_with_checkpoint(oracle, tmp_path):
    sizes, S, W = [50_000, 3, 7_000], 2, 3
    ps = b200ps.ParameterServer(sizes, S, W, optimizer="momentum", lr=0.05, momentum=0.9)
    ps.init_uniform(1)
    P = ps.plan.P
    hp, ha = oracle.fill_uniform(P, 1, 0, 0), np.zeros(P, np.float32)
    for t in (1, 2, 3):
        hg = [oracle.fill_uniform(P, 2, t, w) for w in range(W)]
        for w in range(W):
            K.fill_uniform(ps.grad_buffer(w), 2, t, w)
        ps.step()
        oracle.dense_momentum(hp, ha, hg, 0.05, 0.9)
        if t == 2:
            ps.save(str(tmp_path))
    for s, x in ps.dump().items():
        b, e = x["range"]
        assert same(torch.from_numpy(x["p"]), hp[b:e]) and same(torch.from_numpy(x["m"]), ha[b:e]) and "v" not in x
    q = b200ps.ParameterServer(sizes, S, W, optimizer="momentum", lr=0.05, momentum=0.9)
    q.load(str(tmp_path))
    for w in range(W):
        K.fill_uniform(q.grad_buffer(w), 2, 3, w)
    q.step()
    assert torch.equal(q.replica, ps.replica)
    ps.close()
    q.close()


@pytest.mark.parametrize("opt", ["sgd", "adam"])
def test_ordered_async_mode(oracle, opt):
    """Fixed round-robin 'asynchronous' mode: each worker's gradient applied on its own, in worker
    order -- W optimizer steps per call, bit-identical to the same loop on the oracle."""
    sizes, S, W = [30_000, 5, 4_000], 2, 3
    ps = b200ps.ParameterServer(sizes, S, W, optimizer=opt, lr=1e-3 if opt == "adam" else 0.05)
    ps.init_uniform(1)
    P = ps.plan.P
    hp, hm, hv = oracle.fill_uniform(P, 1, 0, 0), np.zeros(P, np.float32), np.zeros(P, np.float32)
    t = 0
    for step in (1, 2):
        hg = [oracle.fill_uniform(P, 2, step, w) for w in range(W)]
        for w in range(W):
            K.fill_uniform(ps.grad_buffer(w), 2, step, w)
        ps.step_async_ordered()
        for w in range(W):
            t += 1
            if opt == "adam":
                oracle.dense_adam(hp, hm, hv, [hg[w]], oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t))
            else:
                oracle.dense_sgd(hp, [hg[w]], 0.05)
    assert ps.t == t and same(ps.replica, hp)
    ps.close()


@pytest.mark.parametrize("W,n", [(1, 5), (4, 4099), (5, 300_001)])
def test_ftrl_bit_exact(oracle, W, n):
    p = oracle.fill_uniform(n, 1, 0, 0)
    acc, lin = np.full(n, 0.1, np.float32), np.zeros(n, np.float32)
    dp, da, dl = dev(p), dev(acc), dev(lin)
    dg = [torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(W)]
    for t in (1, 2, 3, 4):
        grads = [oracle.fill_uniform(n, 2, t, w, threads=4) for w in range(W)]
        for w in range(W):
            K.fill_uniform(dg[w], 2, t, w)
        K.dense_apply_ftrl(dp, da, dl, dg, 0.05, 0.001, 0.01)
        oracle.dense_ftrl(p, acc, lin, grads, 0.05, 0.001, 0.01, threads=4)
    assert same(dp, p) and same(da, acc) and same(dl, lin)
    assert n < 1000 or ((p == 0).any() and (p != 0).any())  # both branches of the l1 clip were taken


def test_parameter_server_ftrl_with_checkpoint(oracle, tmp_path):
    sizes, S, W = [50_000, 3, 7_000], 2, 3
    kw = dict(optimizer="ftrl", lr=0.05, l1=0.001, l2=0.01)
    ps = b200ps.ParameterServer(sizes, S, W, **kw)
    ps.init_uniform(1)
    P = ps.plan.P
    hp, ha, hl = oracle.fill_uniform(P, 1, 0, 0), np.full(P, 0.1, np.float32), np.zeros(P, np.float32)
    for t in (1, 2, 3):
        hg = [oracle.fill_uniform(P, 2, t, w) for w in range(W)]
        for w in range(W):
            K.fill_uniform(ps.grad_buffer(w), 2, t, w)
        ps.step()
        oracle.dense_ftrl(hp, ha, hl, hg, 0.05, 0.001, 0.01)
        if t == 2:
            ps.save(str(tmp_path))
    for s, x in ps.dump().items():
        b, e = x["range"]
        assert same(torch.from_numpy(x["p"]), hp[b:e]) and same(torch.from_numpy(x["m"]), ha[b:e]) and same(torch.from_numpy(x["v"]), hl[b:e])
    q = b200ps.ParameterServer(sizes, S, W, **kw)
    q.load(str(tmp_path))
    for w in range(W):
        K.fill_uniform(q.grad_buffer(w), 2, 3, w)
    q.step()
    assert torch.equal(q.replica, ps.replica)
    ps.close()
    q.close()


def test_step_host_adam_matches_oracle_and_per_context_tuning(oracle):
    """The pipelined host step (several chunks per slice) with Adam, on a context with its own
    launch geometry: bit-exact; the process-wide tuning of the stateless entry points is untouched."""
    sizes, S, W = [300_000, 77, 50_000], 2, 4
    ps = b200ps.ParameterServer(sizes, S, W, optimizer="adam", lr=1e-3, chunk_elems=65_536, tuning=(128, 4, 1))
    ps.init_uniform(1)
    P = ps.plan.P
    hp, hm, hv = oracle.fill_uniform(P, 1, 0, 0), np.zeros(P, np.float32), np.zeros(P, np.float32)
    out = torch.empty(P, dtype=torch.float32).pin_memory()
    for t in (1, 2, 3):
        hg = [oracle.fill_uniform(P, 2, t, w) for w in range(W)]
        host_g = {w: torch.from_numpy(hg[w]).pin_memory() for w in range(W)}
        ps.step_host(host_g, {0: out})
        ps.host_sync()
        oracle.dense_adam(hp, hm, hv, hg, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t))
        assert np.array_equal(out.numpy().view(np.uint32), hp.view(np.uint32)), f"step {t}"
    ps.step()  # a device step right after a host step is ordered behind its copy-out
    oracle.dense_adam(hp, hm, hv, hg, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, 4))
    torch.cuda.synchronize()
    assert same(ps.replica, hp) and ps.t == 4
    ps.close()
