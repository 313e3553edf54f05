"""Host mirror of the context group of include/b200ps.h: `ParameterServer` is a thin ctypes
wrapper of one `b200ps_ctx_t` -- the PS tasks and compute-task buffers of THIS process's GPU.

Everything that *is* the parameter server lives behind the C ABI (csrc/ctx.cu, csrc/plan.cpp): the
arena and its CUDA-IPC export, the launch plan, the peer-memory barriers, the fused step kernels,
the Adam step counter, the pipelined host step, checkpoints.  What is left here is plumbing: torch
tensor views of the library's device memory, the current torch stream, and `torch.distributed` to
move the 64-byte IPC handles between the ranks (one process per GPU).

transport="nccl" (S = W = G only) is the plain library version of the same step --
reduce-scatter(avg) -> apply -> all-gather with torch.distributed collectives around the stateless
apply kernel -- kept as the comparison point and checked at 1e-6 (NCCL picks the summation order).

Reference: none -- this is what TensorFlow's PS runtime does inside the containers a TFJob
launches; the reference states only the roles (k8s-operator.md:6).
"""
from __future__ import annotations

import ctypes as C
import json
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import AdamScalars, DenseOptions, check, lib
from .contract import Cluster
from .layout import ShardMap

IPC_BYTES = 64
_OPT = {"sgd": _lib.OPT_SGD, "adam": _lib.OPT_ADAM, "momentum": _lib.OPT_MOMENTUM, "ftrl": _lib.OPT_FTRL}
_AGG = {"flat": _lib.AGG_FLAT, "two_level": _lib.AGG_TWO_LEVEL, "multimem": _lib.AGG_MULTIMEM}


# --------------------------------------------------------------------------------------------------
# The plan (b200ps_plan_*): pure host arithmetic in the library, readable without a GPU, so that
# world_size > 1 can be tested on CPU with gloo.
# --------------------------------------------------------------------------------------------------
@dataclass
class Launch:
    shard: int
    begin: int  # element range of the slice in the flat buffer
    end: int
    grad_src: List[Tuple[int, int]]  # buffers the kernel folds, in fold order: (rank, arena offset in elements)
    mirror_ranks: List[int]  # ranks whose replica receives the updated slice (all but this one)


@dataclass
class Plan:
    rank: int
    world: int
    P: int  # padded parameter count
    W: int
    S: int
    local_workers: List[int]
    local_shards: List[int]
    arena_elems: int  # replica [+ partial sum] + one gradient buffer per local worker
    aggregate: str = "flat"
    launches: List[Launch] = field(default_factory=list)
    two_level: bool = False
    first_grad_slot: int = 1
    _arena_of: List[int] = field(default_factory=list)
    _grad_off: List[Tuple[int, int]] = field(default_factory=list)

    def arena_elems_of(self, rank: int) -> int:
        """Size of rank `rank`'s arena (ranks can host different numbers of workers)."""
        return self._arena_of[rank]

    def grad_offset(self, worker: int) -> Tuple[int, int]:
        """(rank, element offset in that rank's arena) of a worker's gradient buffer."""
        return self._grad_off[worker]


def _read_plan(h, S: int, W: int, rank: int, world: int, aggregate: str) -> Plan:
    n = check(lib.b200ps_plan_num_launches(h))
    launches = []
    for i in range(n):
        shard, nsrc, nmir = C.c_int(), C.c_int(), C.c_int()
        b, e = C.c_uint64(), C.c_uint64()
        sr, so, mr = (C.c_int * 64)(), (C.c_uint64 * 64)(), (C.c_int * 8)()
        check(lib.b200ps_plan_launch(h, i, C.byref(shard), C.byref(b), C.byref(e), C.byref(nsrc), sr, so, C.byref(nmir), mr))
        launches.append(Launch(shard.value, b.value, e.value, [(sr[k], so[k]) for k in range(nsrc.value)],
                               [mr[k] for k in range(nmir.value)]))
    goff = []
    for w in range(W):
        r, o = C.c_int(), C.c_uint64()
        check(lib.b200ps_plan_grad_offset(h, w, C.byref(r), C.byref(o)))
        goff.append((r.value, o.value))
    lw = [check(lib.b200ps_plan_local_worker(h, i)) for i in range(check(lib.b200ps_plan_num_local_workers(h)))]
    return Plan(rank, world, int(lib.b200ps_plan_params_padded(h)), W, S, lw, [L.shard for L in launches],
                int(lib.b200ps_plan_arena_elems(h, rank)), aggregate, launches, bool(check(lib.b200ps_plan_two_level(h))),
                check(lib.b200ps_plan_first_grad_slot(h)), [int(lib.b200ps_plan_arena_elems(h, r)) for r in range(world)], goff)


def build_plan(smap: ShardMap, num_workers: int, rank: int, world: int, aggregate: str = "flat") -> Plan:
    """aggregate="flat": g = (1/W) * left fold over the workers in worker order.
    aggregate="two_level": every GPU first folds its own workers (worker order) into one partial
    sum; the owner of a slice folds the partials in GPU order and scales by 1/W.  Same bits as
    "flat" on one GPU; on G GPUs only one buffer per GPU crosses NVLink instead of one per worker."""
    if aggregate not in _AGG:
        raise ValueError("aggregate must be 'flat', 'two_level' or 'multimem'")
    out = C.c_void_p()
    rc = lib.b200ps_plan_build(smap._h, num_workers, rank, world, _AGG[aggregate], C.byref(out))
    if rc == _lib.ERR_RANGE:
        raise ValueError(_lib.last_error())
    check(rc)
    try:
        return _read_plan(out, smap.nshards, num_workers, rank, world, aggregate)
    finally:
        lib.b200ps_plan_destroy(out)


def synthetic_cluster(num_shards: int, num_workers: int) -> Cluster:
    """A TFJob with `num_shards` PS replicas and `num_workers` WORKER replicas (v1alpha1 list form)."""
    return Cluster.from_tfjob({"apiVersion": "tensorflow.org/v1alpha1", "kind": "TFJob", "metadata": {"name": "b200ps"},
                               "spec": {"replicaSpecs": [{"replicas": num_shards, "tfReplicaType": "PS"},
                                                         {"replicas": num_workers, "tfReplicaType": "WORKER"}]}})


# --------------------------------------------------------------------------------------------------
class _CudaArray:
    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def _view(ptr: int, n: int) -> torch.Tensor:
    return torch.as_tensor(_CudaArray(ptr, max(n, 1)), device=torch.device("cuda", torch.cuda.current_device()))[:n]


class _Replay:
    def __init__(self, ps):
        self._ps = ps

    def replay(self):
        self._ps.step()


class ParameterServer:
    """The PS tasks and compute-task buffers of one GPU.  See the module docstring."""

    def __init__(self, var_numel: Sequence[int], num_shards: int, num_workers: int, optimizer: str = "adam",
                 lr: float = 1e-3, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, rank: int = 0,
                 world: int = 1, group=None, transport: str = "peer", chunk_elems: int = 4 << 20,
                 aggregate: str = "flat", overlap: bool = False, momentum: float = 0.9, use_graph: bool = False,
                 l1: float = 0.0, l2: float = 0.0, cluster: Optional[Cluster] = None, tuning: Tuple[int, int, int] = (0, 0, 0)):
        if optimizer not in _OPT:
            raise ValueError("optimizer must be 'sgd', 'momentum', 'adam' or 'ftrl'")
        if transport not in ("peer", "nccl"):
            raise ValueError("transport must be 'peer' or 'nccl'")
        if transport == "nccl" and optimizer not in ("sgd", "adam"):
            raise ValueError("transport='nccl' supports 'sgd' and 'adam'")
        if aggregate not in _AGG:
            raise ValueError("aggregate must be 'flat', 'two_level' or 'multimem'")
        if not torch.cuda.is_available():
            raise RuntimeError("b200ps.ParameterServer needs a CUDA device: there is no CPU fallback")
        self.map = ShardMap(var_numel, num_shards)
        self.cluster = cluster if cluster is not None else synthetic_cluster(num_shards, num_workers)
        if (self.cluster.num_shards, self.cluster.num_workers) != (num_shards, num_workers):
            raise ValueError("the cluster's ps / compute task counts differ from num_shards / num_workers")
        self.optimizer, self.lr, self.betas, self.eps = optimizer, lr, (beta1, beta2), eps
        self.rank, self.world, self.group, self.transport = rank, world, group, transport
        o = DenseOptions()
        check(lib.b200ps_dense_options_default(C.byref(o)))
        o.optimizer, o.aggregate, o.lr, o.beta1, o.beta2, o.eps, o.momentum = _OPT[optimizer], _AGG[aggregate], lr, beta1, beta2, eps, momentum
        o.ftrl_l1, o.ftrl_l2 = l1, l2
        o.threads, o.ctas_per_sm, o.unroll = tuning
        o.use_graph, o.pipeline_folds, o.host_chunk_elems = int(use_graph), int(overlap), chunk_elems
        h = C.c_void_p()
        check(lib.b200ps_ctx_create(self.cluster._h, self.map._h, rank, world, C.byref(o), C.byref(h)))
        self._h = h
        self.plan = _read_plan(C.c_void_p(lib.b200ps_ctx_plan(h)), num_shards, num_workers, rank, world, aggregate)
        self.overlap = bool(overlap) and self.plan.two_level
        self._profile = False
        P = self.plan.P
        p = C.c_void_p()
        check(lib.b200ps_dense_replica(h, C.byref(p)))
        self.replica = _view(p.value, P)
        self.grads: Dict[int, torch.Tensor] = {}
        for w in self.plan.local_workers:
            check(lib.b200ps_dense_grad_buffer(h, w, C.byref(p)))
            self.grads[w] = _view(p.value, P)
        if world > 1:
            self._exchange_handles()
            if transport == "nccl":
                self._init_nccl()

    # ---- set-up -------------------------------------------------------------------------------
    def _exchange_handles(self):
        import torch.distributed as dist

        mine = C.create_string_buffer(IPC_BYTES)
        check(lib.b200ps_ctx_export_handle(self._h, mine))
        blobs: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(blobs, mine.raw, group=self.group)
        check(lib.b200ps_ctx_import_handles(self._h, C.create_string_buffer(b"".join(blobs), IPC_BYTES * self.world)))
        torch.cuda.synchronize()
        dist.barrier(group=self.group)  # every rank has zeroed its flags and mapped every arena

    def _init_nccl(self):
        if not (self.plan.S == self.world == self.plan.W):
            raise ValueError("transport='nccl' is the one-shard-and-one-worker-per-GPU case (S = W = G)")
        sizes = {e - b for b, e in self.map.slices()}
        if len(sizes) != 1:
            raise ValueError("transport='nccl' needs equal slices (P_pad divisible by S * align)")
        self._gsum = torch.empty(sizes.pop(), dtype=torch.float32, device="cuda")
        self._nccl_t = 0

    def close(self):
        if not getattr(self, "_h", None):
            return
        torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier(group=self.group)  # nobody unmaps while a peer may still touch the arena
        self.replica, self.grads = None, {}
        h, self._h = self._h, None
        check(lib.b200ps_ctx_destroy(h))

    def __del__(self):
        try:
            if getattr(self, "_h", None) and self.world == 1:
                lib.b200ps_ctx_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # ---- parameters ---------------------------------------------------------------------------
    @staticmethod
    def _stream() -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def init_uniform(self, seed: int = 1):
        """p[i] = u(seed, 0, 0, i): every rank fills its whole replica with the same values."""
        check(lib.b200ps_dense_init_uniform(self._h, seed, self._stream()))

    def init_params(self, host_params: np.ndarray):
        hp = np.ascontiguousarray(host_params, np.float32)
        if hp.size != self.plan.P:
            raise ValueError(f"expected {self.plan.P} padded parameters, got {hp.size}")
        check(lib.b200ps_dense_init(self._h, hp.ctypes.data, self._stream()))
        torch.cuda.current_stream().synchronize()  # hp may be a temporary

    @property
    def t(self) -> int:
        """Optimizer steps taken (Adam's t)."""
        return int(lib.b200ps_ctx_step_count(self._h)) if self.transport == "peer" or self.world == 1 else self._nccl_t

    def arena_base(self, rank: int) -> int:
        """Device address of rank `rank`'s arena as mapped in this process."""
        p = C.c_void_p()
        check(lib.b200ps_ctx_arena_base(self._h, rank, C.byref(p)))
        return p.value

    def state(self, shard: int) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
        """Device views of a local shard's optimizer state (Adam m, v | momentum accumulator | FTRL accum, linear)."""
        a, b, n = C.c_void_p(), C.c_void_p(), C.c_uint64()
        check(lib.b200ps_dense_state(self._h, shard, C.byref(a), C.byref(b), C.byref(n)))
        return (_view(a.value, n.value) if a.value else None, _view(b.value, n.value) if b.value else None)

    # ---- the step -----------------------------------------------------------------------------
    def step(self):
        """push + aggregate + apply + pull for every local shard; gradients are already in the
        workers' device buffers (`grad_buffer`)."""
        if self.transport == "nccl" and self.world > 1:
            return self._step_nccl()
        check(lib.b200ps_dense_step(self._h, self._stream()))

    def capture_step(self) -> _Replay:
        """Switches the step to CUDA-graph replay (the whole step of this rank -- fold, barrier,
        applies, barrier -- is one graph; Adam's step size and the barrier epoch live in device
        memory) and runs ONE step, which captures it.  `.replay()` of the result runs one more step."""
        check(lib.b200ps_ctx_set_graph(self._h, 1))
        self.step()
        return _Replay(self)

    @property
    def profile(self) -> bool:
        return self._profile

    @profile.setter
    def profile(self, on: bool):
        self._profile = bool(on)
        check(lib.b200ps_ctx_set_profile(self._h, int(self._profile)))

    def phase_ms(self, reset: bool = True) -> Dict[str, float]:
        """Mean device time per step of the phases recorded while `profile` was on (synchronises)."""
        f, a, b = C.c_float(), C.c_float(), C.c_float()
        check(lib.b200ps_ctx_phase_ms(self._h, C.byref(f), C.byref(a), C.byref(b)))
        return {"fold": f.value, "apply": a.value, "barrier": b.value}

    def set_tuning(self, threads: int = 0, ctas_per_sm: int = 0, unroll: int = 0):
        check(lib.b200ps_ctx_set_tuning(self._h, threads, ctas_per_sm, unroll))

    def step_async_ordered(self):
        """A reproducible stand-in for TensorFlow's default ASYNCHRONOUS PS training (SURVEY.md 8(f)
        rank 4): every worker's gradient is applied on its own, one after the other in worker order
        (no averaging), so W optimizer steps happen per call and the result does not depend on timing."""
        if self.transport == "nccl":
            raise ValueError("step_async_ordered needs transport='peer'")
        rc = lib.b200ps_dense_step_async_ordered(self._h, self._stream())
        if rc == _lib.ERR_STATE and "B200PS_AGG_FLAT" in _lib.last_error():
            raise ValueError("step_async_ordered needs aggregate='flat'")
        check(rc)

    def _step_nccl(self):
        import torch.distributed as dist

        self._nccl_t += 1
        L = self.plan.launches[0]
        n = L.end - L.begin
        g = self.grads[self.plan.local_workers[0]]
        dist.reduce_scatter_tensor(self._gsum, g, op=dist.ReduceOp.AVG, group=self.group)
        gp = (C.c_void_p * 1)(self._gsum.data_ptr())
        none = (C.c_void_p * 1)()
        p = C.c_void_p(self.replica.data_ptr() + 4 * L.begin)
        if self.optimizer == "sgd":  # NCCL already averaged: inv_w = 1
            check(lib.b200ps_dense_apply_sgd(p, gp, 1, n, 1.0, self.lr, none, 0, self._stream()))
        else:
            a = AdamScalars()
            check(lib.b200ps_adam_scalars(self.lr, self.betas[0], self.betas[1], self.eps, self._nccl_t, C.byref(a)))
            m, v = self.state(L.shard)
            check(lib.b200ps_dense_apply_adam(p, C.c_void_p(m.data_ptr()), C.c_void_p(v.data_ptr()), gp, 1, n, 1.0,
                                              C.byref(a), none, 0, self._stream()))
        dist.all_gather_into_tensor(self.replica, self.replica[L.begin:L.end], group=self.group)

    # ---- worker-facing ------------------------------------------------------------------------
    def grad_buffer(self, worker: int) -> torch.Tensor:
        """Device buffer a local worker writes its gradient into (zero-copy push)."""
        return self.grads[worker]

    def push(self, worker: int, host_grad: torch.Tensor):
        """H2D of one local worker's gradient (pinned host memory), asynchronous."""
        check(lib.b200ps_dense_push(self._h, worker, C.c_void_p(host_grad.data_ptr()), self._stream()))

    def pull(self, worker: int, host_out: torch.Tensor):
        """D2H of the replica into a local worker's host buffer, asynchronous."""
        if worker not in self.grads:
            raise KeyError(f"worker {worker} is not placed on rank {self.rank}")
        check(lib.b200ps_dense_pull(self._h, worker, C.c_void_p(host_out.data_ptr()), self._stream()))

    def step_host(self, host_grads: Dict[int, torch.Tensor], host_out: Dict[int, torch.Tensor]):
        """A whole step with HOST buffers (pinned): every local worker's gradient goes host ->
        device, the step runs, and the updated parameters go device -> host into ONE buffer (the
        workers of a GPU share the host replica as they share the device one).  Copies and kernels
        are pipelined chunk by chunk inside the library, and the copy-out of one step overlaps the
        copy-in of the next: call host_sync() before reading the host buffer."""
        if self.transport == "nccl" and self.world > 1:
            for w, h in host_grads.items():
                self.grads[w].copy_(h, non_blocking=True)
            self.step()
            for h in host_out.values():
                h.copy_(self.replica, non_blocking=True)
            return
        hg = (C.c_void_p * max(len(self.plan.local_workers), 1))(*[host_grads[w].data_ptr() for w in self.plan.local_workers])
        outs = list(host_out.values())
        check(lib.b200ps_dense_step_host(self._h, hg, C.c_void_p(outs[0].data_ptr()) if outs else None, self._stream()))
        for extra in outs[1:]:  # more than one host copy asked for: plain D2H of the replica
            self.host_sync()
            extra.copy_(outs[0])

    def host_sync(self):
        """Blocks until the host buffers of the last step_host() hold the pulled parameters."""
        check(lib.b200ps_dense_host_sync(self._h))

    def check_barriers(self):
        """Raises if a peer barrier timed out (some rank never arrived).  Synchronises."""
        check(lib.b200ps_ctx_sync(self._h, self._stream()))

    # ---- inspection ---------------------------------------------------------------------------
    def dump(self) -> Dict[int, Dict[str, np.ndarray]]:
        """Local slices: {shard: {"range": (b, e), "p": ..., "m": ..., "v": ...}} as NumPy arrays."""
        P = self.plan.P
        p, a, b = np.empty(P, np.float32), np.zeros(P, np.float32), np.zeros(P, np.float32)
        check(lib.b200ps_dense_dump(self._h, p.ctypes.data, a.ctypes.data, b.ctypes.data, self._stream()))
        out = {}
        nstate = {"sgd": 0, "momentum": 1}.get(self.optimizer, 2)
        for L in self.plan.launches:
            d = {"range": (L.begin, L.end), "p": p[L.begin:L.end].copy()}
            if nstate >= 1:
                d["m"] = a[L.begin:L.end].copy()  # Adam's m, the momentum accumulator, or FTRL's accum
            if nstate >= 2:
                d["v"] = b[L.begin:L.end].copy()  # Adam's v or FTRL's linear
            out[L.shard] = d
        return out

    # ---- checkpoint / resume ("storage" is a stated goal, k8s-operator.md:2) ----
    def save(self, directory: str):
        """Writes this rank's shards -- parameters, optimizer state, step counter -- with the shard
        map as header: <directory>/ps-rank<r>-of-<G>.b200ps (temporary file + rename).  Every rank
        calls it; ranks are joined afterwards so that a following load() sees every file."""
        check(lib.b200ps_dense_save(self._h, directory.encode(), self._stream()))
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier(group=self.group)

    def load(self, directory: str):
        """Restores a job saved with the same shard map, worker count, optimizer and world size."""
        rc = lib.b200ps_dense_load(self._h, directory.encode(), self._stream())
        if rc == _lib.ERR_INVALID:
            msg = _lib.last_error()
            raise (FileNotFoundError if "cannot read" in msg else ValueError)(msg)
        check(rc)
        torch.cuda.synchronize()

    def algorithmic_bytes_per_step(self) -> int:
        """(4W + 8) n for SGD, (4W + 24) n for Adam, summed over local slices.
        (Two-level aggregation moves more local bytes -- the partial sum is written and re-read --
        but the algorithmic figure is the same: W gradients in, one update out.)"""
        return int(lib.b200ps_ctx_algorithmic_bytes_per_step(self._h))
