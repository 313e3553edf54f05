"""Host mirror of the table group of include/b200ps.h: `EmbeddingTable` is a thin ctypes wrapper of
one `b200ps_table_t` (csrc/table.cu) -- the shards of an embedding table and the buffers of the
compute tasks that live on THIS process's GPU.

The arena layout, the CUDA-IPC mapping of the other ranks' arenas, the pointer tables the kernels
take, the peer-memory barriers, the pull / push sequences, graph replay and checkpoints live behind
the C ABI.  Here: torch tensor views of the library's device memory and `torch.distributed` to move
the 64-byte handles between the ranks (one process per GPU).

Reference: none (TensorFlow's embedding_lookup / sparse apply ops inside the containers a TFJob
launches); roles per k8s-operator.md:6.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import TableOptions, check, lib
from .contract import Cluster, placement_gpu
from .layout import row_range

IPC_BYTES = 64
_NONE = (1 << 64) - 1


class TableLayout:
    """Where everything lives, for every rank: the library's arithmetic (b200ps_table_layout), readable
    without a GPU."""

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
            ab, fo = C.c_uint64(), C.c_uint64()
            so, ao = (C.c_uint64 * S)(), (C.c_uint64 * S)()
            io, go, oo, st = (C.c_uint64 * W)(), (C.c_uint64 * W)(), (C.c_uint64 * W)(), (C.c_uint64 * W)()
            rc = lib.b200ps_table_layout(rows, dim, S, W, B, world, int(adagrad), r, C.byref(ab), C.byref(fo), so, ao, io, go, oo, st)
            if rc == _lib.ERR_RANGE:
                raise ValueError(_lib.last_error())
            check(rc)
            self.arena_bytes.append(ab.value)
            self.flag_off[r] = fo.value
            for s in range(S):
                if so[s] != _NONE:
                    self.shard_off[s] = so[s]
                if ao[s] != _NONE:
                    self.accum_off[s] = ao[s]
            for w in range(W):
                if io[w] != _NONE:
                    self.ids_off[w], self.grad_off[w], self.out_off[w] = io[w], go[w], oo[w]
                if st[w] != _NONE:
                    self.stage_off[(r, w)] = st[w]  # rows worker w routes to GPU r

    def shard_rank(self, s: int) -> int:
        return placement_gpu(s, self.world)

    def worker_rank(self, w: int) -> int:
        return placement_gpu(w, self.world)


class _DevArray:
    def __init__(self, ptr: int, typestr: str, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}


class _Replay:
    def __init__(self, tab):
        self._tab = tab

    def replay(self):
        self._tab.step()


class EmbeddingTable:
    def __init__(self, rows: int, dim: int, num_shards: int, num_workers: int, ids_per_worker: int,
                 optimizer: str = "sgd", lr: float = 0.05, rank: int = 0, world: int = 1, group=None,
                 accum_init: float = 0.1, use_graph: bool = False, cluster: Optional[Cluster] = None):
        if optimizer not in ("sgd", "adagrad"):
            raise ValueError("optimizer must be 'sgd' or 'adagrad'")
        if not torch.cuda.is_available():
            raise RuntimeError("b200ps.EmbeddingTable needs a CUDA device: there is no CPU fallback")
        from .ps import synthetic_cluster

        self.cluster = cluster if cluster is not None else synthetic_cluster(num_shards, num_workers)
        if (self.cluster.num_shards, self.cluster.num_workers) != (num_shards, num_workers):
            raise ValueError("the cluster's ps / compute task counts differ from num_shards / num_workers")
        self.lay = TableLayout(rows, dim, num_shards, num_workers, ids_per_worker, world, optimizer == "adagrad")
        self.optimizer, self.lr, self.rank, self.world, self.group = optimizer, lr, rank, world, group
        o = TableOptions()
        check(lib.b200ps_table_options_default(C.byref(o)))
        o.optimizer = _lib.OPT_ADAGRAD if optimizer == "adagrad" else _lib.OPT_SGD
        o.lr, o.accum_init, o.use_graph = lr, accum_init, int(use_graph)
        h = C.c_void_p()
        check(lib.b200ps_table_create(self.cluster._h, rows, dim, ids_per_worker, rank, world, C.byref(o), C.byref(h)))
        self._h = h
        self.local_shards = [s for s in range(num_shards) if self.lay.shard_rank(s) == rank]
        self.local_workers = [w for w in range(num_workers) if self.lay.worker_rank(w) == rank]
        dev = torch.device("cuda", torch.cuda.current_device())
        B = ids_per_worker
        self.shards: Dict[int, torch.Tensor] = {}
        self.accums: Dict[int, torch.Tensor] = {}
        for s in self.local_shards:
            p, a, b, e = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_uint64()
            check(lib.b200ps_table_shard(h, s, C.byref(p), C.byref(a), C.byref(b), C.byref(e)))
            n = e.value - b.value
            self.shards[s] = torch.as_tensor(_DevArray(p.value, "<f4", (max(n, 1), dim)), device=dev)[:n]
            if a.value:
                self.accums[s] = torch.as_tensor(_DevArray(a.value, "<f4", (max(n, 1), dim)), device=dev)[:n]
        self.ids: Dict[int, torch.Tensor] = {}
        self.grad_rows: Dict[int, torch.Tensor] = {}
        self.pulled: Dict[int, torch.Tensor] = {}
        for w in self.local_workers:
            i, g, p = C.c_void_p(), C.c_void_p(), C.c_void_p()
            check(lib.b200ps_table_worker_buffers(h, w, C.byref(i), C.byref(g), C.byref(p)))
            self.ids[w] = torch.as_tensor(_DevArray(i.value, "<i8", (B,)), device=dev)
            self.grad_rows[w] = torch.as_tensor(_DevArray(g.value, "<f4", (B, dim)), device=dev)
            self.pulled[w] = torch.as_tensor(_DevArray(p.value, "<f4", (B, dim)), device=dev)
        if world > 1:
            self._exchange_handles()

    def _exchange_handles(self):
        import torch.distributed as dist

        mine = C.create_string_buffer(IPC_BYTES)
        check(lib.b200ps_table_export_handle(self._h, mine))
        blobs: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(blobs, mine.raw, group=self.group)
        check(lib.b200ps_table_import_handles(self._h, C.create_string_buffer(b"".join(blobs), IPC_BYTES * self.world)))
        torch.cuda.synchronize()
        dist.barrier(group=self.group)  # flags zeroed and arenas mapped everywhere

    def close(self):
        if not getattr(self, "_h", None):
            return
        torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier(group=self.group)
        self.shards, self.accums, self.ids, self.grad_rows, self.pulled = {}, {}, {}, {}, {}
        h, self._h = self._h, None
        check(lib.b200ps_table_destroy(h))

    def __del__(self):
        try:
            if getattr(self, "_h", None) and self.world == 1:
                lib.b200ps_table_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @staticmethod
    def _stream() -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def check_barriers(self):
        check(lib.b200ps_table_sync(self._h, self._stream()))

    # ---- table -------------------------------------------------------------------------------
    def init_uniform(self, seed: int = 1):
        """table[i, d] = u(seed, 0, 0, i*dim + d) on the shards of this GPU."""
        check(lib.b200ps_table_init_uniform(self._h, seed, self._stream()))

    # ---- worker-facing -----------------------------------------------------------------------
    def pull(self, worker: int, out: torch.Tensor) -> torch.Tensor:
        """a10: out[k, :] = table[ids_w[k], :] for a local worker, rows read from the owning GPUs."""
        check(lib.b200ps_sparse_pull_worker(self._h, worker, C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def pull_step(self):
        """a10 for all workers of the job at once (collective): afterwards `pulled[w]` holds
        table[ids_w] for every local worker w."""
        check(lib.b200ps_sparse_pull(self._h, self._stream()))

    def push_step(self):
        """a11 + a12 for the shards of this GPU over every worker's (ids, gradient rows)."""
        check(lib.b200ps_sparse_push(self._h, self._stream()))

    def step(self):
        """pull of every worker + push, sharing one barrier (what a training step does)."""
        check(lib.b200ps_sparse_step(self._h, self._stream()))

    def capture_step(self) -> _Replay:
        """Switches `step` to CUDA-graph replay (about ten launches -- digit counts, radix passes, segment scan, applies --
        replayed without per-launch host work; all counts live on the device and every grid is fixed,
        so the graph is valid for any ids / gradient rows written into the same buffers) and runs one
        step, which captures it.  Works across ranks too: the barriers' epoch lives in device memory."""
        check(lib.b200ps_table_set_graph(self._h, 1))
        self.step()
        return _Replay(self)

    @property
    def profile(self) -> bool:
        return getattr(self, "_profile", False)

    @profile.setter
    def profile(self, on: bool):
        self._profile = bool(on)
        check(lib.b200ps_table_set_profile(self._h, int(self._profile)))

    def phase_ms(self) -> Dict[str, float]:
        a, b, c = C.c_float(), C.c_float(), C.c_float()
        check(lib.b200ps_table_phase_ms(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return {"serve_pull": a.value, "route+barrier+sort_segment": b.value, "apply": c.value}

    def plan(self) -> dict:
        """The plan of the last push as NumPy arrays (debug accessor; synchronises)."""
        n_total = self.lay.W * self.lay.B
        kept, uniq, bad = C.c_uint64(), C.c_uint64(), C.c_uint64()
        s = np.empty(n_total, np.int64)
        perm = np.empty(n_total, np.uint32)
        seg = np.empty(n_total, np.int32)
        un = np.empty(n_total, np.int64)
        off = np.empty(n_total + 1, np.uint32)
        check(lib.b200ps_table_plan_read(self._h, C.byref(kept), C.byref(uniq), C.byref(bad), s.ctypes.data, perm.ctypes.data,
                                         seg.ctypes.data, un.ctypes.data, off.ctypes.data, self._stream()))
        k, u = kept.value, uniq.value
        return {"n_kept": k, "n_unique": u, "n_bad": bad.value, "sorted_ids": s[:k], "perm": perm[:k], "seg_id": seg[:k],
                "unique_ids": un[:u], "seg_offset": off[:u + 1]}

    def dump(self) -> Dict[int, Tuple[Tuple[int, int], np.ndarray]]:
        self.check_barriers()
        torch.cuda.synchronize()
        return {s: (self.lay.shard_range[s], self.shards[s].cpu().numpy()) for s in self.local_shards}

    # ---- checkpoint / resume of the table shards ---------------------------
    def save(self, directory: str):
        """Every rank writes its shards (and Adagrad accumulators) with the geometry as header:
        <directory>/table-rank<r>-of-<G>.b200ps (temporary file + rename)."""
        check(lib.b200ps_table_save(self._h, directory.encode(), self._stream()))
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier(group=self.group)

    def load(self, directory: str):
        """Restores this rank's shards from a checkpoint written with the same geometry and world size."""
        rc = lib.b200ps_table_load(self._h, directory.encode(), self._stream())
        if rc == _lib.ERR_INVALID:
            msg = _lib.last_error()
            raise (FileNotFoundError if "cannot read" in msg else ValueError)(msg)
        check(rc)
        torch.cuda.synchronize()

    def bytes_per_step(self, n_unique_local: int) -> Dict[str, int]:
        """Algorithmic bytes: gather B*(8 + 2*4*dim) per worker; push B*(8 + 4*dim) per
        entry kept + unique * 8*dim (SGD) or 16*dim (Adagrad)."""
        d4 = 4 * self.lay.dim
        return {"gather_per_worker": self.lay.B * (8 + 2 * d4),
                "push_per_entry": 8 + d4, "push_per_unique": (4 if self.lay.adagrad else 2) * d4}
