"""BASELINE configs[4]: sweep of variable sizes 1 KB .. 1 GB x PS shard counts 1/2/4/8 (W = S), SGD
and Adam, on one B200: the roofline curve of the aggregate+apply kernel.  One JSON line per point:
algorithmic GB/s = (4W + 8|24) * P / step time, step = S launches (one per shard)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import kernels as K  # noqa: E402


def time_ms(fn, iters, warm=5):
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
    graph = "--graph" in sys.argv
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > L2: written between timed steps
    for opt in ("sgd", "adam"):
        for S in (1, 2, 4, 8):
            for n in b200ps.synth.sweep_c4_sizes():
                W = S
                if (W + 3) * n * 4 > 60e9:
                    continue
                ps = b200ps.ParameterServer([n], S, W, optimizer=opt)
                ps.init_uniform(1)
                for w in range(W):
                    K.fill_uniform(ps.grad_buffer(w), 2, 1, w)
                iters = 200 if n <= (1 << 22) else 20
                step = ps.step
                if graph:  # launch-bound sizes: the S launches of a step replayed as one CUDA graph
                    ps.step()
                    torch.cuda.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        ps.step()
                    step = g.replay
                small = ps.plan.P * 4 * (W + 3) < (126 << 20)

                def one():
                    if small:
                        flush.fill_(1)  # evict the working set so every step reads HBM
                    step()

                ms = time_ms(one, iters)
                if small:
                    ms -= time_ms(lambda: flush.fill_(1), iters)
                nbytes = ps.algorithmic_bytes_per_step()
                print(json.dumps({"opt": opt, "S": S, "W": W, "var_bytes": n * 4, "ms_per_step": round(ms, 5),
                                  "GBps": round(nbytes / ms / 1e6, 1), "graph": graph, "l2_flushed": small}), flush=True)
                ps.close()


if __name__ == "__main__":
    main()
