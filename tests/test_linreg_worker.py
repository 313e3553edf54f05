"""BASELINE configs[0] with a REAL worker: the linear-regression gradient is
computed from synthetic features, pushed through the PS path, and the parameters converge to the
generating weights.  CPU: C oracle vs NumPy.  GPU: kernels vs oracle, bit-exact."""
import numpy as np
import pytest

SEED_X, SEED_W = 4, 5


def test_oracle_c_vs_numpy_and_convergence(oracle, nporacle):
    n = 3001  # not a multiple of the 256-element block
    p = oracle.fill_uniform(n, 1, 0, 0)
    for batch, step, worker in ((1, 1, 0), (16, 3, 1), (64, 7, 2)):
        a = oracle.linreg_grad(p, batch, SEED_X, SEED_W, step, worker, threads=3)
        assert np.array_equal(a.view(np.uint32), nporacle.linreg_grad(p, batch, SEED_X, SEED_W, step, worker).view(np.uint32))
    # at the generating weights the gradient vanishes exactly
    wt = nporacle.uniform(SEED_W, 0, 0, 0, n)
    assert not oracle.linreg_grad(wt, 8, SEED_X, SEED_W, 1, 0).any()
    # 1 PS + 2 workers, SGD: the parameters converge to the generating weights
    n = 512
    p, wt = np.zeros(n, np.float32), nporacle.uniform(SEED_W, 0, 0, 0, n)
    for t in range(1, 201):
        oracle.dense_sgd(p, [oracle.linreg_grad(p, 64, SEED_X, SEED_W, t, w) for w in range(2)], 1.0)
    assert np.linalg.norm(p - wt) / np.linalg.norm(wt) < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("n,batch", [(3001, 16), (100_000, 64), (255, 1)])
def test_gpu_gradient_bit_exact(oracle, n, batch):
    torch = pytest.importorskip("torch")
    from b200ps import kernels as K

    p = oracle.fill_uniform(n, 1, 0, 0)
    g = K.linreg_grad(torch.empty(n, device="cuda"), torch.from_numpy(p).cuda(), batch, SEED_X, SEED_W, 5, 3)
    want = oracle.linreg_grad(p, batch, SEED_X, SEED_W, 5, 3, threads=4)
    assert np.array_equal(g.cpu().numpy().view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_c0_training_through_the_parameter_server(oracle, nporacle):
    """configs[0]: 1 PS + 2 workers, SGD.  Workers read the replica (pull), write their gradient
    into their buffer of the arena (push), the PS steps; 150 steps converge, and the first 10 are
    bit-identical to the same loop on the oracle."""
    torch = pytest.importorskip("torch")
    import b200ps
    from b200ps import kernels as K

    n, W, B = 512, 2, 64  # lr = 1 is stable while n / (3 * W * B) < 2
    ps = b200ps.ParameterServer([n], 1, W, optimizer="sgd", lr=1.0)
    ps.init_params(np.zeros(ps.plan.P, np.float32))
    hp = np.zeros(n, np.float32)
    for t in range(1, 151):
        for w in range(W):
            K.linreg_grad(ps.grad_buffer(w), ps.replica, B, SEED_X, SEED_W, t, w, n=n)
        ps.step()
        if t <= 10:
            oracle.dense_sgd(hp, [oracle.linreg_grad(hp, B, SEED_X, SEED_W, t, w) for w in range(W)], 1.0)
            assert np.array_equal(ps.replica[:n].cpu().numpy().view(np.uint32), hp.view(np.uint32)), f"step {t}"
    wt = nporacle.uniform(SEED_W, 0, 0, 0, n)
    err = np.linalg.norm(ps.replica[:n].cpu().numpy() - wt) / np.linalg.norm(wt)
    assert err < 1e-3, err
    ps.close()
