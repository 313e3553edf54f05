"""Rank body of tests/test_gpu_multirank.py (launched by torch.distributed.run, one rank per GPU)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import b200ps  # noqa: E402
from b200ps import kernels as K  # noqa: E402
from oracle import c_oracle as O  # noqa: E402


def main():
    S, W, opt, transport, steps = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4], 3
    aggregate = sys.argv[5] if len(sys.argv) > 5 else "flat"
    overlap = (sys.argv[6] == "overlap") if len(sys.argv) > 6 else False
    ordered_async = (sys.argv[6] == "async") if len(sys.argv) > 6 else False
    graph = (sys.argv[7] == "graph") if len(sys.argv) > 7 else False
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sizes = [700_000, 333, 123_456, 2_000_000]
    ps = b200ps.ParameterServer(sizes, S, W, optimizer=opt, lr=1e-3 if opt == "adam" else 0.05, rank=rank, world=world,
                                transport=transport, aggregate=aggregate, overlap=overlap, use_graph=graph)
    ps.init_uniform(1)
    P = ps.plan.P
    hp = O.fill_uniform(P, 1, 0, 0, threads=4)
    hm, hv = np.zeros(P, np.float32), np.zeros(P, np.float32)
    for t in range(1, steps + 1):
        for w in ps.plan.local_workers:
            K.fill_uniform(ps.grad_buffer(w), 2, t, w)
        hg = [O.fill_uniform(P, 2, t, w, threads=4) for w in range(W)]
        if ordered_async:  # every worker's gradient applied on its own, in worker order: W optimizer steps per call
            ps.step_async_ordered()
            for w in range(W):
                k = (t - 1) * W + w + 1
                if opt == "adam":
                    O.dense_adam(hp, hm, hv, [hg[w]], O.adam_scalars(1e-3, 0.9, 0.999, 1e-8, k), threads=4, inv_w=np.float32(1.0))
                else:
                    O.dense_sgd(hp, [hg[w]], 0.05, threads=4, inv_w=np.float32(1.0))
            continue
        ps.step()
        inv_w = np.float32(1.0) / np.float32(W)
        if aggregate in ("two_level", "multimem"):  # the oracle folds per-GPU partials, then the partials
            hg = O.two_level_partials(hg, world, threads=4)
        if opt == "adam":
            O.dense_adam(hp, hm, hv, hg, O.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t), threads=4, inv_w=inv_w)
        else:
            O.dense_sgd(hp, hg, 0.05, threads=4, inv_w=inv_w)
    torch.cuda.synchronize()
    got = ps.replica.cpu().numpy()
    if aggregate == "multimem" and not b200ps._lib.lib.b200ps_multicast_supported():
        raise SystemExit("multicast is not supported on this box")
    if transport == "peer" and aggregate != "multimem":  # fixed worker order: bit-exact on every rank's replica
        ok = np.array_equal(got.view(np.uint32), hp.view(np.uint32))
    else:  # NCCL picks its own summation order: the north-star tolerance, 1e-6 relative
        ok = np.allclose(got, hp, rtol=1e-6, atol=1e-7)
    for s, d in ps.dump().items():
        b, e = d["range"]
        if "m" in d and transport == "peer" and aggregate != "multimem":
            ok = ok and np.array_equal(d["m"].view(np.uint32), hm[b:e].view(np.uint32))
            ok = ok and np.array_equal(d["v"].view(np.uint32), hv[b:e].view(np.uint32))
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ps.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MULTIRANK_OK" if flag.item() == 1 else "MULTIRANK_MISMATCH", flush=True)
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
