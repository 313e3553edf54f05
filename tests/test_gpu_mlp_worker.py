"""The MLP worker: the hand-written tcgen05 GEMM and the forward / backward it
drives, against a plain PyTorch fp32 reference of the same op.  This is the one floating-point kernel
of the project, so the reference is torch fp32 and the tolerance is stated here: the GEMM's fp32
accumulation of bf16 products is compared at 2^-8 relative (one bf16 rounding of the output); the
gradients of the mixed-precision step (bf16 activations between layers) at 2 % of their norm."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

import b200ps
from b200ps import kernels as K
from b200ps.worker import MlpWorker, gemm_bf16, mlp_variables

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 256, 128), (512, 384, 1024), (2048, 8192, 1024), (1024, 2048, 8192)])
def test_gemm_matches_fp32_reference(M, N, K):
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) * 0.5).bfloat16()
    ct = torch.empty(N, M, dtype=torch.bfloat16, device="cuda")
    c = gemm_bf16(a, b, out_t=ct)
    ref = a.float() @ b.float().t()
    torch.cuda.synchronize()
    err = (c.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err <= 2.0 ** -8 * scale + 1e-3, (err, scale)   # one bf16 rounding of an fp32-accumulated result
    assert torch.equal(ct, c.t().contiguous())               # the transposed copy is the same values
    # exactness property: with operands whose products and sums are exactly representable the result is exact
    ai = torch.randint(-2, 3, (M, K), device="cuda", generator=g).bfloat16()
    bi = torch.randint(-2, 3, (N, K), device="cuda", generator=g).bfloat16()
    ci = gemm_bf16(ai, bi)
    refi = ai.float() @ bi.float().t()
    ok = refi.abs() < 256  # integers below 2^8 are exact in bf16
    assert torch.equal(ci.float()[ok], refi[ok])


def _reference_step(params, offs, dims, x, t, batch, tasks):
    """fp32 PyTorch reference of the worker's mixed-precision step: bf16-rounded weights and activations,
    fp32 math; returns per-task (loss, {offset: gradient})."""
    out = []
    L = len(dims) - 1
    for w in range(tasks):
        xs, ts = x[w * batch:(w + 1) * batch].float(), t[w * batch:(w + 1) * batch].float()
        Ws = [params[offs[2 * l]: offs[2 * l] + dims[l] * dims[l + 1]].view(dims[l], dims[l + 1]).bfloat16().float() for l in range(L)]
        bs = [params[offs[2 * l + 1]: offs[2 * l + 1] + dims[l + 1]] for l in range(L)]
        hs = [xs]
        for l in range(L):
            z = hs[-1] @ Ws[l] + bs[l]
            if l + 1 < L:
                z = torch.relu(z)
            hs.append(z.bfloat16().float())
        diff = hs[-1] - ts
        loss = (diff * diff).sum().item() / (2 * batch)
        dh = (diff / batch).bfloat16().float()
        grads = {}
        for l in range(L - 1, -1, -1):
            grads[offs[2 * l]] = hs[l].t() @ dh
            grads[offs[2 * l + 1]] = dh.sum(0)
            if l >= 1:
                dh = ((dh @ Ws[l].t()) * (hs[l] > 0)).bfloat16().float()
        out.append((loss, grads))
    return out


@pytest.mark.parametrize("dims,batch,S,W", [((256, 512, 256), 128, 1, 2), ((128, 384, 128, 256), 64, 2, 4), ((1024, 8192, 8192, 2048, 4096), 128, 2, 1)])
def test_mlp_step_matches_fp32_reference(dims, batch, S, W):
    ps = b200ps.ParameterServer(mlp_variables(dims), S, W, optimizer="adam")
    # parameters of a sensible scale: N(0, 1/in) weights, small biases
    P = ps.plan.P
    offs = ps.map.var_offsets()
    g = torch.Generator().manual_seed(1)
    hp = torch.zeros(P)
    for l in range(len(dims) - 1):
        n = dims[l] * dims[l + 1]
        hp[offs[2 * l]: offs[2 * l] + n] = torch.randn(n, generator=g) / dims[l] ** 0.5
        hp[offs[2 * l + 1]: offs[2 * l + 1] + dims[l + 1]] = torch.randn(dims[l + 1], generator=g) * 0.1
    ps.init_params(hp.numpy())
    wk = MlpWorker(ps, dims, batch)
    rows = batch * W
    x = torch.randn(rows, dims[0], generator=g).bfloat16().pin_memory()
    t = torch.randn(rows, dims[-1], generator=g).bfloat16().pin_memory()
    loss = wk.step(x, t)
    torch.cuda.synchronize()
    ref = _reference_step(ps.replica.clone(), offs, dims, x.cuda(), t.cuda(), batch, W)
    for w in range(W):
        want_loss, grads = ref[w]
        assert abs(loss[w].item() - want_loss) <= 2e-2 * abs(want_loss) + 1e-6, (w, loss[w].item(), want_loss)
        got = ps.grad_buffer(w)
        for l in range(len(dims) - 1):
            gw = got[offs[2 * l]: offs[2 * l] + dims[l] * dims[l + 1]].view(dims[l], dims[l + 1])
            gb = got[offs[2 * l + 1]: offs[2 * l + 1] + dims[l + 1]]
            for name, a, b in ((f"dW{l}", gw, grads[offs[2 * l]]), (f"db{l}", gb, grads[offs[2 * l + 1]])):
                rel = ((a - b).norm() / (b.norm() + 1e-30)).item()
                assert rel <= 2e-2, (w, name, rel)
    # and the gradients drive the PS: one Adam step through the fused kernel changes the parameters
    before = ps.replica.clone()
    ps.step()
    torch.cuda.synchronize()
    assert not torch.equal(before, ps.replica)
    wk.close()
    ps.close()


def test_mlp_training_reduces_the_loss():
    """A few real steps -- worker forward/backward -> zero-copy push -> fused aggregate + Adam -> zero-copy
    pull -- on a fixed mini-batch: the loss goes down."""
    dims, batch, S, W = (256, 512, 256), 128, 2, 2
    ps = b200ps.ParameterServer(mlp_variables(dims), S, W, optimizer="adam", lr=1e-3)
    offs = ps.map.var_offsets()
    g = torch.Generator().manual_seed(2)
    hp = torch.zeros(ps.plan.P)
    for l in range(len(dims) - 1):
        n = dims[l] * dims[l + 1]
        hp[offs[2 * l]: offs[2 * l] + n] = torch.randn(n, generator=g) / dims[l] ** 0.5
    ps.init_params(hp.numpy())
    wk = MlpWorker(ps, dims, batch)
    x = torch.randn(batch * W, dims[0], generator=g).bfloat16().pin_memory()
    t = torch.randn(batch * W, dims[-1], generator=g).bfloat16().pin_memory()
    losses = []
    for _ in range(30):
        wk.step(x, t)
        ps.step()
        torch.cuda.synchronize()
        losses.append(wk.loss.mean().item())
    assert losses[-1] < 0.7 * losses[0], losses
    wk.close()
    ps.close()
