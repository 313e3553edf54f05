"""Experiment (one B200, under gpurun): how much of the push plan (sort_segment, needs the ids only) hides
behind the pull (serve) when b200ps_sparse_step runs it on a second, high-priority stream?  Times the
library's table step launched eagerly, captured by torch, and as the library's own graph.  Prints JSON
lines.  Finding (round 2): 0.59 -> 0.535 ms eager for C2 uniform; inside a graph both branches start
together, serve's CTAs take every slot and the step is as long as the sequential one.  Not a product path."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import kernels as K  # noqa: E402


def table_path():
    """The same question through the library's own table step (b200ps_sparse_step), eager and as its graph."""
    rows, dim, S, W, B = 10_000_000, 64, 8, 8, 262_144
    for kind in ("uniform", "zipf"):
        tab = b200ps.EmbeddingTable(rows, dim, S, W, B, optimizer="sgd", lr=0.01, rank=0, world=1)
        tab.init_uniform(1)
        ids = b200ps.synth.make_ids(kind, rows, W, B, 1)
        for w in tab.local_workers:
            tab.ids[w].copy_(torch.from_numpy(ids[w]))
            K.fill_uniform(tab.grad_rows[w].view(-1), 2, 1, 0, first=w * B * dim)
        res = {"table_step": kind}
        for name in ("eager", "torch_graph", "graph"):
            if name == "graph":
                tab.capture_step()
            if name == "torch_graph":  # the library's eager step (its own fork / join) captured by torch
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    tab.step()
                for _ in range(3):
                    g.replay()
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(20):
                    g.replay()
                b.record()
                torch.cuda.synchronize()
                res[name + "_ms"] = round(a.elapsed_time(b) / 20, 4)
                continue
            for _ in range(3):
                tab.step()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(20):
                tab.step()
            b.record()
            torch.cuda.synchronize()
            res[name + "_ms"] = round(a.elapsed_time(b) / 20, 4)
        print(json.dumps(res), flush=True)
        tab.close()
        del tab
        torch.cuda.empty_cache()


if __name__ == "__main__":
    table_path()
