"""Host side of the sparse PS path for the tasks on THIS GPU.

An embedding table of `rows` x `dim` fp32 is partitioned by TF "div" over S PS shards; shard s
lives on GPU s mod G, compute task w on GPU w mod G.  Each rank keeps one exported arena:

    [ rows of local shard 0 | local shard 1 | ... | (Adagrad accumulators, same order) |
      per local worker: ids int64[B] | gradient rows fp32[B, dim] | pulled rows fp32[B, dim] |
      per REMOTE worker: staging fp32[B, dim] for the gradient rows it routes here ]

Every rank can compute every other rank's layout, so after the one-time handle exchange a kernel
is given plain base pointers, and the routing all-to-alls of a classic PS are fused into kernels:
ids are streamed from wherever the workers wrote them, and rows cross NVLink only as posted
stores -- `pull_step`: each owner serves the rows of its shards into the requesting workers'
buffers; `push_step`: each worker's GPU routes the rows other GPUs own into their staging slots,
then every GPU sorts the ids of ITS shards and folds rows out of local memory only.

Reference: none (TensorFlow's embedding_lookup / sparse apply ops inside the containers a TFJob
launches); roles per k8s-operator.md:6.  torch is plumbing (streams, distributed rendezvous).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from ._lib import TableGeom, check, lib
from .contract import placement_gpu
from .layout import row_range

IPC_BYTES = 64
BARRIER_WORDS = 16  # B200PS_BARRIER_WORDS


def _up(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


class _ByteArray:
    def __init__(self, ptr: int, nbytes: int, typestr: str, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}


class TableLayout:
    """Pure arithmetic: where everything lives, for every rank (testable without a GPU)."""

    def __init__(self, rows: int, dim: int, S: int, W: int, B: int, world: int, adagrad: bool):
        self.rows, self.dim, self.S, self.W, self.B, self.world, self.adagrad = rows, dim, S, W, B, world, adagrad
        self.shard_range = [row_range(rows, S, s) for s in range(S)]
        self.shard_off: Dict[int, int] = {}
        self.accum_off: Dict[int, int] = {}
        self.ids_off: Dict[int, int] = {}
        self.grad_off: Dict[int, int] = {}
        self.out_off: Dict[int, int] = {}
        self.stage_off: Dict[Tuple[int, int], int] = {}
        self.flag_off: Dict[int, int] = {}
        self.arena_bytes: List[int] = []
        for r in range(world):
            off = 0
            mine = [s for s in range(S) if placement_gpu(s, world) == r]
            for s in mine:
                self.shard_off[s] = off
                off += _up((self.shard_range[s][1] - self.shard_range[s][0]) * dim * 4)
            if adagrad:
                for s in mine:
                    self.accum_off[s] = off
                    off += _up((self.shard_range[s][1] - self.shard_range[s][0]) * dim * 4)
            for w in (w for w in range(W) if placement_gpu(w, world) == r):
                self.ids_off[w] = off
                off += _up(B * 8)
                self.grad_off[w] = off
                off += _up(B * dim * 4)
                self.out_off[w] = off
                off += _up(B * dim * 4)
            for w in (w for w in range(W) if placement_gpu(w, world) != r):
                self.stage_off[(r, w)] = off  # rows worker w routes to GPU r
                off += _up(B * dim * 4)
            self.flag_off[r] = off  # barrier flags behind everything else
            off += _up(BARRIER_WORDS * 4)
            self.arena_bytes.append(off)

    def shard_rank(self, s: int) -> int:
        return placement_gpu(s, self.world)

    def worker_rank(self, w: int) -> int:
        return placement_gpu(w, self.world)


class EmbeddingTable:
    def __init__(self, rows: int, dim: int, num_shards: int, num_workers: int, ids_per_worker: int,
                 optimizer: str = "sgd", lr: float = 0.05, rank: int = 0, world: int = 1, group=None,
                 accum_init: float = 0.1, barrier: str = "nccl"):
        if barrier not in ("peer", "nccl"):
            raise ValueError("barrier must be 'peer' or 'nccl'")
        if optimizer not in ("sgd", "adagrad"):
            raise ValueError("optimizer must be 'sgd' or 'adagrad'")
        if not torch.cuda.is_available():
            raise RuntimeError("b200ps.EmbeddingTable needs a CUDA device: there is no CPU fallback")
        self.lay = TableLayout(rows, dim, num_shards, num_workers, ids_per_worker, world, optimizer == "adagrad")
        self.optimizer, self.lr, self.rank, self.world, self.group = optimizer, lr, rank, world, group
        self.geom = TableGeom(rows, dim, num_shards, rank, world)
        nbytes = self.lay.arena_bytes[rank]
        p = C.c_void_p()
        check(lib.b200ps_dev_alloc(nbytes, C.byref(p)))
        self.base: Dict[int, int] = {rank: p.value}
        self._own = p.value
        self.local_shards = [s for s in range(num_shards) if self.lay.shard_rank(s) == rank]
        self.local_workers = [w for w in range(num_workers) if self.lay.worker_rank(w) == rank]
        dev = torch.device("cuda", torch.cuda.current_device())
        B = ids_per_worker
        self.shards: Dict[int, torch.Tensor] = {}
        self.accums: Dict[int, torch.Tensor] = {}
        for s in self.local_shards:
            n = self.lay.shard_range[s][1] - self.lay.shard_range[s][0]
            self.shards[s] = torch.as_tensor(_ByteArray(p.value + self.lay.shard_off[s], n * dim * 4, "<f4", (n, dim)), device=dev)
            if self.lay.adagrad:
                self.accums[s] = torch.as_tensor(_ByteArray(p.value + self.lay.accum_off[s], n * dim * 4, "<f4", (n, dim)), device=dev)
                self.accums[s].fill_(accum_init)
        self.ids: Dict[int, torch.Tensor] = {}
        self.grad_rows: Dict[int, torch.Tensor] = {}
        self.pulled: Dict[int, torch.Tensor] = {}
        for w in self.local_workers:
            self.ids[w] = torch.as_tensor(_ByteArray(p.value + self.lay.ids_off[w], B * 8, "<i8", (B,)), device=dev)
            self.grad_rows[w] = torch.as_tensor(_ByteArray(p.value + self.lay.grad_off[w], B * dim * 4, "<f4", (B, dim)), device=dev)
            self.pulled[w] = torch.as_tensor(_ByteArray(p.value + self.lay.out_off[w], B * dim * 4, "<f4", (B, dim)), device=dev)
        self.workspace = torch.empty(int(lib.b200ps_sparse_workspace_bytes(num_workers * B, dim)), dtype=torch.uint8, device=dev)
        self._flag = torch.zeros(1, dtype=torch.int32, device=dev)
        self.barrier_kind, self._epoch = barrier, 0
        torch.as_tensor(_ByteArray(p.value + self.lay.flag_off[rank], BARRIER_WORDS * 4, "<i4", (BARRIER_WORDS,)), device=dev).zero_()
        if world > 1:
            self._exchange_handles()
            self._flag_ptrs = (C.c_void_p * world)(*[self.base[r] + self.lay.flag_off[r] for r in range(world)])
            import torch.distributed as dist

            torch.cuda.synchronize()
            dist.barrier(group=group)  # flags zeroed and arenas mapped everywhere
        L = self.lay
        self._shard_ptrs = (C.c_void_p * num_shards)(*[self.base[L.shard_rank(s)] + L.shard_off[s] for s in range(num_shards)])
        self._accum_ptrs = (C.c_void_p * num_shards)(*[(self.base[L.shard_rank(s)] + L.accum_off[s]) if L.adagrad else 0
                                                       for s in range(num_shards)])
        self._ids_ptrs = (C.c_void_p * num_workers)(*[self.base[L.worker_rank(w)] + L.ids_off[w] for w in range(num_workers)])
        # apply reads local memory only: a local worker's own buffer, or this GPU's staging slot for a remote one
        self._grad_ptrs = (C.c_void_p * num_workers)(*[
            self.base[rank] + (L.grad_off[w] if L.worker_rank(w) == rank else L.stage_off[(rank, w)]) for w in range(num_workers)])
        self._out_ptrs = (C.c_void_p * num_workers)(*[self.base[L.worker_rank(w)] + L.out_off[w] for w in range(num_workers)])
        self._route_dst = {w: (C.c_void_p * world)(*[0 if o == rank else self.base[o] + L.stage_off[(o, w)] for o in range(world)])
                           for w in self.local_workers}

    def _exchange_handles(self):
        import torch.distributed as dist

        buf = C.create_string_buffer(IPC_BYTES)
        check(lib.b200ps_ipc_export(C.c_void_p(self._own), buf))
        blobs: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(blobs, buf.raw, group=self.group)
        for r, blob in enumerate(blobs):
            if r != self.rank:
                q = C.c_void_p()
                check(lib.b200ps_ipc_open(C.create_string_buffer(blob, IPC_BYTES), C.byref(q)))
                self.base[r] = q.value

    def close(self):
        torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier(group=self.group)
        for r, b in list(self.base.items()):
            if r != self.rank:
                lib.b200ps_ipc_close(C.c_void_p(b))
        self.shards, self.accums, self.ids, self.grad_rows, self.pulled = {}, {}, {}, {}, {}
        if self._own:
            check(lib.b200ps_dev_free(C.c_void_p(self._own)))
            self._own = 0

    @staticmethod
    def _stream() -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def _barrier(self):
        if self.world > 1:
            if self.barrier_kind == "peer":
                self._epoch += 1
                check(lib.b200ps_peer_barrier(self._flag_ptrs, self.rank, self.world, self._epoch, self._stream()))
            else:
                import torch.distributed as dist

                dist.all_reduce(self._flag, group=self.group)

    def check_barriers(self):
        if self.world > 1 and self.barrier_kind == "peer":
            check(lib.b200ps_peer_barrier_error(C.c_void_p(self._own + self.lay.flag_off[self.rank]), self._stream()))

    # ---- table -------------------------------------------------------------------------------
    def init_uniform(self, seed: int = 1):
        """table[i, d] = u(seed, 0, 0, i*dim + d) on the shards of this GPU."""
        for s in self.local_shards:
            b, _ = self.lay.shard_range[s]
            t = self.shards[s]
            check(lib.b200ps_fill_uniform(C.c_void_p(t.data_ptr()), t.numel(), seed, 0, 0, b * self.lay.dim, self._stream()))

    # ---- worker-facing -----------------------------------------------------------------------
    def pull(self, worker: int, out: torch.Tensor) -> torch.Tensor:
        """a10: out[k, :] = table[ids_w[k], :] for a local worker, rows read from the owning GPUs."""
        check(lib.b200ps_sparse_gather(C.c_void_p(out.data_ptr()), self._shard_ptrs, C.byref(self.geom),
                                       C.c_void_p(self.ids[worker].data_ptr()), self.lay.B, self._stream()))
        return out

    def serve(self):
        """Owner side of pull for every worker: rows of this GPU's shards -> the workers' `pulled` buffers."""
        check(lib.b200ps_sparse_serve(self._out_ptrs, self._shard_ptrs, self._ids_ptrs, self.lay.W, self.lay.B,
                                      C.byref(self.geom), self._stream()))

    def pull_step(self):
        """a10 for all workers of the job at once (collective): afterwards `pulled[w]` holds
        table[ids_w] for every local worker w."""
        self._barrier()  # every worker's ids are written
        self.serve()
        self._barrier()  # every owner has delivered its rows

    def route(self):
        """Worker side of push: gradient rows owned by other GPUs go to their staging slots."""
        for w in self.local_workers:
            check(lib.b200ps_sparse_route(C.c_void_p(self.grad_rows[w].data_ptr()), C.c_void_p(self.ids[w].data_ptr()),
                                          self.lay.B, C.byref(self.geom), self._route_dst[w], self._stream()))

    def sort_segment(self):
        """a11: the plan (sorted ids, permutation, segments) of this GPU's shards, into the workspace."""
        check(lib.b200ps_sparse_sort_segment(self._ids_ptrs, self.lay.W, self.lay.B, C.byref(self.geom),
                                             C.c_void_p(self.workspace.data_ptr()), self.workspace.numel(), self._stream()))

    def apply(self):
        """a12: fold the gradient rows of every unique id of the plan and update its row."""
        if self.optimizer == "sgd":
            check(lib.b200ps_sparse_apply_sgd(self._shard_ptrs, self._grad_ptrs, self.lay.W, self.lay.B, C.byref(self.geom),
                                              self.lr, C.c_void_p(self.workspace.data_ptr()), self._stream()))
        else:
            check(lib.b200ps_sparse_apply_adagrad(self._shard_ptrs, self._accum_ptrs, self._grad_ptrs, self.lay.W, self.lay.B,
                                                  C.byref(self.geom), self.lr, C.c_void_p(self.workspace.data_ptr()),
                                                  self._stream()))

    def push_step(self):
        """a11 + a12 for the shards of this GPU over every worker's (ids, gradient rows)."""
        if self.world > 1:
            self.route()
        self._barrier()  # all workers' ids are written and every routed row has landed
        self.sort_segment()
        self.apply()
        self._barrier()  # every shard is updated before anyone pulls again

    def capture_step(self) -> "torch.cuda.CUDAGraph":
        """pull_step + push_step of one GPU as a CUDA graph: ~25 small launches (radix passes, scans,
        segments) replayed without per-launch host work.  All counts live on the device and every
        grid is fixed, so the graph is valid for any ids / gradient rows written into the same
        buffers.  One GPU only (the cross-rank barriers are not captured)."""
        if self.world != 1:
            raise ValueError("capture_step supports one GPU")
        self.pull_step()
        self.push_step()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.pull_step()
            self.push_step()
        return g

    def plan(self) -> dict:
        """The plan of the last sort_segment as NumPy arrays (debug accessor; synchronises)."""
        from .kernels import sparse_plan_read

        return sparse_plan_read(self.workspace, self.lay.W * self.lay.B, self.lay.rows)

    def dump(self) -> Dict[int, Tuple[Tuple[int, int], np.ndarray]]:
        self.check_barriers()
        torch.cuda.synchronize()
        return {s: (self.lay.shard_range[s], self.shards[s].cpu().numpy()) for s in self.local_shards}

    # ---- checkpoint / resume of the table shards ---------------------------
    def save(self, directory: str):
        """Every rank writes its shards (and Adagrad accumulators) with the row partition as header:
        <directory>/table-rank<r>-of-<G>.npz."""
        import json
        import os

        os.makedirs(directory, exist_ok=True)
        self.check_barriers()
        torch.cuda.synchronize()
        L = self.lay
        header = {"format": "b200ps-table-1", "rows": L.rows, "dim": L.dim, "shards": L.S, "optimizer": self.optimizer,
                  "rank": self.rank, "world": self.world, "ranges": {str(s): list(L.shard_range[s]) for s in self.local_shards}}
        arrays = {"header": np.frombuffer(json.dumps(header).encode(), dtype=np.uint8)}
        for s in self.local_shards:
            arrays[f"rows{s}"] = self.shards[s].cpu().numpy()
            if L.adagrad:
                arrays[f"accum{s}"] = self.accums[s].cpu().numpy()
        np.savez(os.path.join(directory, f"table-rank{self.rank}-of-{self.world}.npz"), **arrays)

    def load(self, directory: str):
        """Restores this rank's shards from a checkpoint written with the same geometry and world size."""
        import json
        import os

        z = np.load(os.path.join(directory, f"table-rank{self.rank}-of-{self.world}.npz"))
        h = json.loads(bytes(z["header"]).decode())
        L = self.lay
        if (h["rows"], h["dim"], h["shards"], h["optimizer"], h["world"]) != (L.rows, L.dim, L.S, self.optimizer, self.world):
            raise ValueError(f"table checkpoint was written for a different geometry: {h}")
        for s in self.local_shards:
            self.shards[s].copy_(torch.from_numpy(z[f"rows{s}"]))
            if L.adagrad:
                self.accums[s].copy_(torch.from_numpy(z[f"accum{s}"]))
        torch.cuda.synchronize()

    def bytes_per_step(self, n_unique_local: int) -> Dict[str, int]:
        """Algorithmic bytes: gather B*(8 + 2*4*dim) per worker; push B*(8 + 4*dim) per
        entry kept + unique * 8*dim (SGD) or 16*dim (Adagrad)."""
        d4 = 4 * self.lay.dim
        return {"gather_per_worker": self.lay.B * (8 + 2 * d4),
                "push_per_entry": 8 + d4, "push_per_unique": (4 if self.lay.adagrad else 2) * d4}
