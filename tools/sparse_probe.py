"""Library reference points for the sparse kernels on one B200 (run under gpurun): what does
torch's own index_select / index_add reach on the C2 shapes?  Context for the roofline fractions
of sparse_gather_kernel / sparse_apply_kernel -- not a product path."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import kernels as K  # noqa: E402


def time_ms(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    rows, dim, S, W, B = 10_000_000, 64, 8, 8, 262_144
    table = torch.empty(rows, dim, device="cuda")
    K.fill_uniform(table.view(-1), 1, 0, 0)
    shards = [table[slice(*b200ps.row_range(rows, S, s))] for s in range(S)]
    geom = K.table_geom(rows, dim, S)
    for kind in ("uniform", "zipf"):
        ids = torch.from_numpy(b200ps.synth.make_ids(kind, rows, W, B, 1)).cuda()
        flat = ids.reshape(-1)
        out = torch.empty(W * B, dim, device="cuda")
        bytes_g = W * B * (8 + 2 * 4 * dim)
        ms = time_ms(lambda: torch.index_select(table, 0, flat, out=out))
        print(json.dumps({"what": f"torch.index_select {kind}", "ms": round(ms, 4), "GBps": round(bytes_g / ms / 1e6, 1)}))
        ms = time_ms(lambda: K.sparse_gather(out, shards, geom, flat))
        print(json.dumps({"what": f"b200ps gather one launch of W*B ids {kind}", "ms": round(ms, 4), "GBps": round(bytes_g / ms / 1e6, 1)}))
        o1 = torch.empty(B, dim, device="cuda")

        def per_worker():
            for w in range(W):
                K.sparse_gather(o1, shards, geom, ids[w])

        ms = time_ms(per_worker)
        print(json.dumps({"what": f"b200ps gather W launches of B ids {kind}", "ms": round(ms, 4), "GBps": round(bytes_g / ms / 1e6, 1)}))
        ms = time_ms(lambda: torch.sort(flat, stable=True))
        print(json.dumps({"what": f"torch.sort(stable) of W*B int64 ids {kind}", "ms": round(ms, 4)}))
        ws = K.sparse_workspace(W * B, dim)
        idl = [ids[w] for w in range(W)]
        ms = time_ms(lambda: K.sparse_sort_segment(idl, B, geom, ws))
        print(json.dumps({"what": f"b200ps sort_segment {kind}", "ms": round(ms, 4)}))
    # streaming copy of the same volume for scale
    x = torch.empty(W * B * dim, device="cuda")
    y = torch.empty_like(x)
    ms = time_ms(lambda: y.copy_(x))
    print(json.dumps({"what": "torch copy_ of W*B rows (0.54 GB)", "ms": round(ms, 4), "GBps": round(2 * x.numel() * 4 / ms / 1e6, 1)}))


if __name__ == "__main__":
    main()
