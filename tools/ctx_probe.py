"""Where does a C1 step spend its time?  Stateless kernel launches vs the context's step (direct and
graph replay), CUDA events, one GPU."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import kernels as K, synth  # noqa: E402


def timed(fn, iters=50):
    for _ in range(5):
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
    torch.cuda.set_device(0)
    sizes, S, W = synth.mlp_c1_variables(), 2, 4
    sm = b200ps.ShardMap(sizes, S)
    P = sm.total
    p = K.fill_uniform(torch.empty(P, dtype=torch.float32, device="cuda"), 1, 0, 0)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    g = [K.fill_uniform(torch.empty(P, dtype=torch.float32, device="cuda"), 2, 1, w) for w in range(W)]
    opt = K.adam_scalars(1e-3, t=1)

    def stateless():
        for b, e in sm.slices():
            K.dense_apply_adam(p[b:e], m[b:e], v[b:e], [x[b:e] for x in g], opt)

    out = {"stateless_2_launches_ms": round(timed(stateless), 4)}
    del p, m, v, g
    for graph in (False, True):
        ps = b200ps.ParameterServer(sizes, S, W, optimizer="adam", use_graph=graph)
        ps.init_uniform(1)
        for w in range(W):
            K.fill_uniform(ps.grad_buffer(w), 2, 1, w)
        with torch.cuda.stream(torch.cuda.Stream()):
            out["ctx_step_graph_ms" if graph else "ctx_step_ms"] = round(timed(ps.step), 4)
        ps.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
