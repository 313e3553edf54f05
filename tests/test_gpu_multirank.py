"""N > 1 on real GPUs: one process per GPU, gradients read from and parameters written to peer
memory inside the step kernel.  Skipped on a 1-GPU box (run with `gpurun --gpus 2`)."""
import os
import socket
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(n, *args, script="_mr_worker.py"):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(_port()), os.path.join(HERE, script)] + [str(a) for a in args]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "MULTIRANK_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("S,W,opt", [(2, 4, "adam"), (2, 2, "sgd"), (3, 5, "adam"), (1, 2, "adam")])
def test_peer_fused_step_is_bit_exact_on_every_replica(S, W, opt):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    _run(2, S, W, opt, "peer")


def test_all_gpus_of_the_box():
    n = torch.cuda.device_count()
    if n < 4:
        pytest.skip("needs >= 4 GPUs")
    _run(n, 2 * n, 4 * n, "adam", "peer")


def test_nccl_transport_matches_within_tolerance():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    _run(2, 2, 2, "adam", "nccl")


@pytest.mark.parametrize("opt,kind", [("sgd", "zipf"), ("adagrad", "uniform")])
def test_sparse_pull_and_push_across_gpus(opt, kind):
    """Gather from peer shards; every GPU sorts its shards' ids out of all workers' buffers and folds
    gradient rows from peer memory: bit-exact against the oracle after 3 steps."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    _run(2, opt, kind, script="_mr_sparse_worker.py")


@pytest.mark.parametrize("S,W,opt", [(4, 8, "adam"), (2, 5, "sgd"), (2, 1, "adam")])
def test_two_level_aggregation_is_bit_exact(S, W, opt):
    """Each GPU folds its own workers first, one partial per GPU crosses NVLink; the oracle does
    the same two-level fold, so every replica still matches bit for bit."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    _run(2, S, W, opt, "peer", "two_level")


@pytest.mark.parametrize("S,W,opt", [(4, 8, "adam"), (3, 5, "sgd"), (2, 1, "adam")])
def test_pipelined_two_level_step_is_bit_exact(S, W, opt):
    """Fold stream + apply stream with per-slice ready flags in peer memory: same bits."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    _run(2, S, W, opt, "peer", "two_level", "overlap", "peer")


def test_peer_memory_barrier():
    """The flag barrier through peer memory instead of the NCCL all-reduce: same bits, uneven arenas included."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    _run(2, 3, 5, "adam", "peer", "flat", "sequential", "peer")


@pytest.mark.parametrize("S,W,opt,mode", [(4, 8, "adam", "sequential"), (2, 2, "sgd", "sequential"), (4, 8, "adam", "overlap")])
def test_in_switch_reduction_matches_within_tolerance(S, W, opt, mode):
    """B200PS_AGG_MULTIMEM: the per-GPU partial sums are added inside the NVSwitch (multimem.ld_reduce) and
    the updated slices are multicast to every replica (multimem.st).  The switch picks the summation
    order, so the check is the north-star tolerance (1e-6 relative), not bit-exactness."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    import b200ps

    torch.cuda.set_device(0)
    if not b200ps._lib.lib.b200ps_multicast_supported():
        pytest.skip("the device reports no multicast support")
    _run(2, S, W, opt, "peer", "multimem", mode)


@pytest.mark.parametrize("aggregate,mode", [("two_level", "sequential"), ("two_level", "overlap"), ("multimem", "overlap")])
def test_whole_rank_step_replayed_as_one_cuda_graph(aggregate, mode):
    """fold -> barrier -> applies -> barrier (or the pipelined form) of every rank captured once and replayed:
    the barrier epoch lives in device memory and Adam's step size is re-armed on the apply nodes."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    if aggregate == "multimem":
        import b200ps

        torch.cuda.set_device(0)
        if not b200ps._lib.lib.b200ps_multicast_supported():
            pytest.skip("the device reports no multicast support")
    _run(2, 4, 8, "adam", "peer", aggregate, mode, "graph")


@pytest.mark.parametrize("S,W,opt", [(3, 5, "adam"), (2, 4, "sgd")])
def test_ordered_async_mode_across_gpus(S, W, opt):
    """The reproducible stand-in for asynchronous PS training at world = 2: every worker's gradient is
    applied on its own in worker order (the owner reads it over NVLink), mirrors written with the last one."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    _run(2, S, W, opt, "peer", "flat", "async")
