"""Rank body of the sparse multi-GPU test (launched by torch.distributed.run, one rank per GPU)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import b200ps  # noqa: E402
from oracle import c_oracle as O  # noqa: E402


def main():
    opt, kind = sys.argv[1], sys.argv[2]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rows, dim, S, W, B = 60_000, 64, 8, 8, 4096
    tab = b200ps.EmbeddingTable(rows, dim, S, W, B, optimizer=opt, lr=0.05, rank=rank, world=world)
    tab.init_uniform(1)
    table = O.fill_uniform(rows * dim, 1, 0, 0, threads=4).reshape(rows, dim)
    accum = np.full((rows, dim), 0.1, np.float32)
    out = torch.empty(B, dim, dtype=torch.float32, device="cuda")
    ok = True
    for step in (1, 2, 3):
        ids = b200ps.synth.make_ids(kind, rows, W, B, step)
        grad = O.fill_uniform(W * B * dim, 2, step, 0, threads=4).reshape(W, B, dim)
        for w in tab.local_workers:
            tab.ids[w].copy_(torch.from_numpy(ids[w]))
            tab.grad_rows[w].copy_(torch.from_numpy(grad[w]))
        if step == 1:
            tab.pull_step()  # owners serve rows into the workers' buffers (posted NVLink stores)
            for w in tab.local_workers:
                tab.pull(w, out)  # the single-worker form reads peer shards directly
                ok = ok and np.array_equal(out.cpu().numpy().view(np.uint32), O.gather(table, ids[w]).view(np.uint32))
            pulled = {w: tab.pulled[w].cpu().numpy() for w in tab.local_workers}
            tab.push_step()
        elif step == 2:
            tab.step()  # pull + push sharing one barrier (b200ps_sparse_step)
            pulled = {w: tab.pulled[w].cpu().numpy() for w in tab.local_workers}
        else:
            tab.capture_step()  # the same, captured and replayed as one CUDA graph per rank (barriers included)
            pulled = {w: tab.pulled[w].cpu().numpy() for w in tab.local_workers}
        for w in tab.local_workers:
            ok = ok and np.array_equal(pulled[w].view(np.uint32), O.gather(table, ids[w]).view(np.uint32))
        inv_w = np.float32(1.0) / np.float32(W)
        if opt == "sgd":
            O.sparse_apply_sgd(table, 0, grad.reshape(-1, dim), ids.reshape(-1), inv_w, np.float32(0.05))
        else:
            O.sparse_apply_adagrad(table, accum, 0, grad.reshape(-1, dim), ids.reshape(-1), inv_w, np.float32(0.05))
    for s, ((b, e), vals) in tab.dump().items():
        ok = ok and np.array_equal(vals.view(np.uint32), table[b:e].view(np.uint32))
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    tab.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MULTIRANK_OK" if flag.item() == 1 else "MULTIRANK_MISMATCH", flush=True)
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
