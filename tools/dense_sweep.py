"""Tuning sweep of the dense aggregate+apply kernel on one B200 (run under gpurun).
Prints one JSON line per launch geometry: algorithmic GB/s = (4W + 8|24) * n / kernel time."""
import itertools
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import kernels as K  # noqa: E402


def time_ms(fn, iters=10, warm=3):
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
    n = 50_342_912  # one C1 shard
    out = []
    x = torch.empty(1 << 29, dtype=torch.float32, device="cuda")
    y = torch.empty_like(x)
    ms = time_ms(lambda: y.copy_(x))
    print(json.dumps({"what": "torch copy_ 2 GiB", "GBps": round(2 * x.numel() * 4 / ms / 1e6, 1)}), flush=True)
    del x, y
    for opt, W in (("adam", 4), ("sgd", 4), ("adam", 1), ("adam", 2), ("adam", 8), ("sgd", 1), ("adam", 5)):
        p = K.fill_uniform(torch.empty(n, dtype=torch.float32, device="cuda"), 1, 0, 0)
        m, v = torch.zeros_like(p), torch.zeros_like(p)
        g = [K.fill_uniform(torch.empty(n, dtype=torch.float32, device="cuda"), 2, 1, w) for w in range(W)]
        sc = K.adam_scalars(1e-3, t=1)
        bytes_ = (4 * W + (24 if opt == "adam" else 8)) * n
        geos = list(itertools.product((128, 256, 512, 1024), (1, 2, 4, 8, 16), (1, 2, 4)))
        if not (opt == "adam" and W == 4):
            geos = [(256, 4, 2), (256, 8, 1), (512, 2, 2), (512, 4, 1), (256, 8, 2), (1024, 2, 1), (128, 16, 1), (256, 4, 4)]
        best = None
        for th, ct, un in geos:
            if th * ct > 2048:
                continue
            K.set_dense_tuning(th, ct, un)
            fn = (lambda: K.dense_apply_adam(p, m, v, g, sc)) if opt == "adam" else (lambda: K.dense_apply_sgd(p, g, 0.01))
            ms = time_ms(fn)
            rec = {"opt": opt, "W": W, "threads": th, "ctas_per_sm": ct, "unroll": un, "ms": round(ms, 4),
                   "GBps": round(bytes_ / ms / 1e6, 1)}
            print(json.dumps(rec), flush=True)
            if best is None or rec["GBps"] > best["GBps"]:
                best = rec
        print(json.dumps({"best": best}), flush=True)
        out.append(best)
        del p, m, v, g
    print(json.dumps({"summary": out}))


if __name__ == "__main__":
    main()
