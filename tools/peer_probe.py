"""NVLink probe (torchrun, 2+ ranks): what do peer loads / peer stores reach inside the dense
kernels?  Rank 0 prints GB/s for: local fold, fold from a peer buffer (pure peer read), SGD with a
peer mirror (pure peer write), and both, for a few launch geometries."""
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import kernels as K  # noqa: E402
from b200ps._lib import check, lib  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = 64 << 20  # 256 MB per buffer
    ps = b200ps.ParameterServer([n], world, world, optimizer="sgd", rank=rank, world=world)
    P = ps.plan.P
    st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    me = ps.arena_base(rank)
    peers = [ps.arena_base(r) for r in range(world) if r != rank]
    out = torch.empty(P, dtype=torch.float32, device="cuda")
    pa = lambda v: (C.c_void_p * max(len(v), 1))(*v)

    def timed(fn, iters=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        dist.barrier()
        return a.elapsed_time(b) / iters

    cases = {
        "fold local (1 buf)": lambda: check(lib.b200ps_dense_fold(out.data_ptr(), pa([me + 4 * P]), 1, P, st())),
        "fold 1 peer buf (pure peer read)": lambda: check(lib.b200ps_dense_fold(out.data_ptr(), pa([peers[0] + 4 * P]), 1, P, st())),
        "fold all peers' bufs": lambda: check(lib.b200ps_dense_fold(out.data_ptr(), pa([p + 4 * P for p in peers]), len(peers), P, st())),
        "sgd local grad + 1 peer mirror (pure peer write)": lambda: check(lib.b200ps_dense_apply_sgd(
            out.data_ptr(), pa([me + 4 * P]), 1, P, 1.0, 0.0, pa([peers[0]]), 1, st())),
        "sgd peer grads + peer mirrors (both)": lambda: check(lib.b200ps_dense_apply_sgd(
            out.data_ptr(), pa([p + 4 * P for p in peers]), len(peers), P, 1.0, 0.0, pa(peers[:8]), min(len(peers), 8), st())),
    }
    for geo in ((256, 8, 2), (256, 8, 4), (512, 4, 4), (1024, 2, 4), (256, 16, 1)):
        K.set_dense_tuning(*geo)
        for name, fn in cases.items():
            ms = timed(fn)
            if rank == 0:
                nb = {"fold local (1 buf)": 1, "fold 1 peer buf (pure peer read)": 1, "fold all peers' bufs": len(peers),
                      "sgd local grad + 1 peer mirror (pure peer write)": 1, "sgd peer grads + peer mirrors (both)": len(peers)}[name]
                print(json.dumps({"geo": geo, "case": name, "ms": round(ms, 4), "link_GBps_per_dir": round(nb * P * 4 / ms / 1e6, 1),
                                  "load_kind": os.environ.get("B200PS_DENSE_LOAD", "0")}), flush=True)
    ps.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
