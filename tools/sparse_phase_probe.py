"""Times the stateless sparse kernels of one C2 step on one B200, phase by phase (run under gpurun):
serve (pull), sort_segment, apply_sgd / apply_adagrad, for uniform and Zipf ids.  Prints one JSON
line per measurement; `--tag` labels the build being measured (kernel variants are compared by
swapping libb200ps.so between runs).  Not a product path."""
import argparse
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
    ap = argparse.ArgumentParser()
    ap.add_argument("--tag", default="default")
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--per-worker", type=int, default=262_144)
    args = ap.parse_args()
    rows, dim, S, W, B = args.rows, args.dim, 8, 8, args.per_worker
    table = torch.empty(rows, dim, device="cuda")
    K.fill_uniform(table.view(-1), 1, 0, 0)
    accum = torch.full((rows, dim), 0.1, device="cuda")
    shards = [table[slice(*b200ps.row_range(rows, S, s))] for s in range(S)]
    ashards = [accum[slice(*b200ps.row_range(rows, S, s))] for s in range(S)]
    geom = K.table_geom(rows, dim, S)
    grads = [torch.empty(B, dim, device="cuda") for _ in range(W)]
    for w in range(W):
        K.fill_uniform(grads[w].view(-1), 2, 1, w)
    peak = 6551.7
    for kind in ("uniform", "zipf"):
        ids = torch.from_numpy(b200ps.synth.make_ids(kind, rows, W, B, 1)).cuda()
        idl = [ids[w] for w in range(W)]
        outs = [torch.empty(B, dim, device="cuda") for _ in range(W)]
        ws = K.sparse_workspace(W * B, dim)
        K.sparse_sort_segment(idl, B, geom, ws)
        plan = K.sparse_plan_read(ws, W * B, rows)
        uniq = int(plan["n_unique"])
        res = {"tag": args.tag, "ids": kind, "unique": uniq}
        ms = time_ms(lambda: K.sparse_serve(outs, shards, idl, B, geom))
        res["serve_ms"] = round(ms, 4)
        res["serve_frac"] = round(W * B * (8 + 8 * dim) / ms / 1e6 / peak, 3)
        ms = time_ms(lambda: K.sparse_sort_segment(idl, B, geom, ws))
        res["sort_segment_ms"] = round(ms, 4)
        ms = time_ms(lambda: K.sparse_apply_sgd(shards, grads, B, geom, 0.01, ws))
        res["apply_sgd_ms"] = round(ms, 4)
        res["apply_sgd_frac"] = round((W * B * (8 + 4 * dim) + uniq * 8 * dim) / ms / 1e6 / peak, 3)
        ms = time_ms(lambda: K.sparse_apply_adagrad(shards, ashards, grads, B, geom, 0.01, ws))
        res["apply_adagrad_ms"] = round(ms, 4)
        res["apply_adagrad_frac"] = round((W * B * (8 + 4 * dim) + uniq * 16 * dim) / ms / 1e6 / peak, 3)
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
