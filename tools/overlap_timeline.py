"""Timeline of the pipelined multi-GPU dense step (torchrun): when does each fold / wait / apply
finish, relative to the start of the step, on the first and last rank?"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import kernels as K, synth  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ps = b200ps.ParameterServer(synth.mlp_c1_variables(), 2 * world, 4 * world, rank=rank, world=world, aggregate="two_level")
    ps.init_uniform(1)
    for w in ps.plan.local_workers:
        K.fill_uniform(ps.grad_buffer(w), 2, 1, w)
    for _ in range(10):
        ps.step()
    torch.cuda.synchronize()
    dist.barrier()
    for it in range(2):
        ps.trace = []
        ps.step()
        torch.cuda.synchronize()
        t0 = ps.trace[0][1]
        line = {"rank": rank, "iter": it, "ms": {lab: round(t0.elapsed_time(ev), 3) for lab, ev in ps.trace[1:]}}
        ps.trace = None
        dist.barrier()
        for r in (0, world - 1):
            if rank == r:
                print(json.dumps(line), flush=True)
            dist.barrier()
    ps.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
