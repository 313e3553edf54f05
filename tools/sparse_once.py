"""One C2-sized round of the stateless sparse kernels (uniform ids), twice: sort_segment, serve, apply_sgd,
apply_adagrad.  The second round is what `ncu --set full -k regex:... -s 8 -c 8` captures.  Not a product path."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import kernels as K  # noqa: E402

rows, dim, S, W, B = 10_000_000, 64, 8, 8, 262_144
table = torch.empty(rows, dim, device="cuda")
K.fill_uniform(table.view(-1), 1, 0, 0)
accum = torch.full((rows, dim), 0.1, device="cuda")
shards = [table[slice(*b200ps.row_range(rows, S, s))] for s in range(S)]
ashards = [accum[slice(*b200ps.row_range(rows, S, s))] for s in range(S)]
geom = K.table_geom(rows, dim, S)
grads = [torch.empty(B, dim, device="cuda") for _ in range(W)]
for w in range(W):
    K.fill_uniform(grads[w].view(-1), 2, 1, w)
ids = torch.from_numpy(b200ps.synth.make_ids("uniform", rows, W, B, 1)).cuda()
idl = [ids[w] for w in range(W)]
outs = [torch.empty(B, dim, device="cuda") for _ in range(W)]
ws = K.sparse_workspace(W * B, dim)
for _ in range(2):
    K.sparse_sort_segment(idl, B, geom, ws)
    K.sparse_serve(outs, shards, idl, B, geom)
    K.sparse_apply_sgd(shards, grads, B, geom, 0.01, ws)
    K.sparse_apply_adagrad(shards, ashards, grads, B, geom, 0.01, ws)
    torch.cuda.synchronize()
print("ok")
