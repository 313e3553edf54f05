"""Host side of the parameter-server step for the tasks placed on THIS process's GPU.

One process per GPU (rank r of G, `torch.distributed`).  The TFJob's PS shards and compute tasks
are placed on GPUs as SURVEY.md 8(e) says (shard s -> GPU s mod G, worker w -> GPU w mod G); a
rank owns the gradient buffers of its workers, one replica of the whole parameter vector (what
its workers read) and the optimizer state of its shards.  The master copy of slice s is the
slice-s range of its owner's replica.

A step is ONE kernel per local shard (csrc/dense.cu): it loads the W gradient slices straight
from the buffers of the ranks that hold them (NVLink peer loads = push), folds them in worker
order, applies the optimizer and stores the new slice into every rank's replica (peer stores =
pull).  The only other traffic is two stream-ordered barriers (all gradients written / all
replicas written).  transport="nccl" is the plain library version of the same step
(reduce-scatter -> apply -> all-gather) kept as the comparison point.

Reference: none -- this is what TensorFlow's PS runtime does inside the containers a TFJob
launches; the reference states only the roles (k8s-operator.md:6).  torch is plumbing here
(streams, pinned host memory, torch.distributed); all arithmetic goes through the C ABI.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ._lib import AdamScalars, check, lib
from .contract import placement_gpu
from .layout import ShardMap

IPC_BYTES = 64
BARRIER_WORDS = 16  # B200PS_BARRIER_WORDS
SLICE_FLAG_STRIDE = 16  # words per slice in the slice-ready flags: one per producing rank
MAX_SLICES = 64
FLAG_WORDS = BARRIER_WORDS + MAX_SLICES * SLICE_FLAG_STRIDE  # behind every arena


# --------------------------------------------------------------------------------------------------
# The plan: pure host logic (no CUDA), so that world_size > 1 can be tested on CPU with gloo.
# --------------------------------------------------------------------------------------------------
@dataclass
class Launch:
    shard: int
    begin: int  # element range of the slice in the flat buffer
    end: int
    # buffers the kernel folds, in fold order: (rank, arena offset in elements).  "flat": one per
    # worker, worker order.  "two_level": one per GPU that hosts workers (its partial sum), GPU order.
    grad_src: List[Tuple[int, int]]
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

    @property
    def two_level(self) -> bool:
        """Two-level aggregation only differs from the flat fold when there are several GPUs."""
        return self.aggregate == "two_level" and self.world > 1

    @property
    def first_grad_slot(self) -> int:
        return 2 if self.two_level else 1  # arena slot 0 = replica, slot 1 = partial sum (two-level)

    def arena_elems_of(self, rank: int) -> int:
        """Size of rank `rank`'s arena (ranks can host different numbers of workers)."""
        n_local = sum(1 for w in range(self.W) if placement_gpu(w, self.world) == rank)
        return self.P * (self.first_grad_slot + n_local)

    def grad_offset(self, worker: int) -> Tuple[int, int]:
        """(rank, element offset in that rank's arena) of a worker's gradient buffer."""
        r = placement_gpu(worker, self.world)
        return r, self.P * (self.first_grad_slot + worker // self.world)


def build_plan(smap: ShardMap, num_workers: int, rank: int, world: int, aggregate: str = "flat") -> Plan:
    """aggregate="flat": g = (1/W) * left fold over the workers in worker order.
    aggregate="two_level": every GPU first folds its own workers (worker order) into one partial
    sum; the owner of a slice folds the partials in GPU order and scales by 1/W.  Same bits as
    "flat" on one GPU; on G GPUs only one buffer per GPU crosses NVLink instead of one per worker."""
    S, P = smap.nshards, smap.total
    if aggregate not in ("flat", "two_level"):
        raise ValueError("aggregate must be 'flat' or 'two_level'")
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} is outside 0..{world - 1}")
    if world - 1 > 8:
        raise ValueError("one launch writes at most 8 peer replicas (B200PS_MAX_MIRRORS)")
    lw = [w for w in range(num_workers) if placement_gpu(w, world) == rank]
    ls = [s for s in range(S) if placement_gpu(s, world) == rank]
    plan = Plan(rank, world, P, num_workers, S, lw, ls, 0, aggregate)
    plan.arena_elems = P * (plan.first_grad_slot + len(lw))
    if plan.two_level:
        src = [(r, P) for r in range(world) if any(placement_gpu(w, world) == r for w in range(num_workers))]
    else:
        src = [plan.grad_offset(w) for w in range(num_workers)]
    for s in ls:
        b, e = smap.slice(s)
        plan.launches.append(Launch(s, b, e, list(src), [r for r in range(world) if r != rank]))
    return plan


# --------------------------------------------------------------------------------------------------
class _CudaArray:
    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


class Arena:
    """A cudaMalloc'ed fp32 block of this rank, exportable to the other ranks of the box."""

    def __init__(self, nelems: int):
        p = C.c_void_p()
        check(lib.b200ps_dev_alloc(nelems * 4, C.byref(p)))
        self.ptr, self.n = p.value, nelems
        self.tensor = torch.as_tensor(_CudaArray(self.ptr, nelems), device=torch.device("cuda", torch.cuda.current_device()))

    def export(self) -> bytes:
        buf = C.create_string_buffer(IPC_BYTES)
        check(lib.b200ps_ipc_export(C.c_void_p(self.ptr), buf))
        return buf.raw

    def free(self):
        if self.ptr:
            self.tensor = None
            check(lib.b200ps_dev_free(C.c_void_p(self.ptr)))
            self.ptr = 0


def _ptrs(vals: Sequence[int]):
    return (C.c_void_p * max(len(vals), 1))(*vals)


class ParameterServer:
    """The PS tasks and compute-task buffers of one GPU.  See the module docstring."""

    def __init__(self, var_numel: Sequence[int], num_shards: int, num_workers: int, optimizer: str = "adam",
                 lr: float = 1e-3, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, rank: int = 0,
                 world: int = 1, group=None, transport: str = "peer", chunk_elems: int = 4 << 20,
                 aggregate: str = "flat", barrier: str = "nccl", overlap: bool = False, momentum: float = 0.9):
        if barrier not in ("peer", "nccl"):
            raise ValueError("barrier must be 'peer' (flags in peer memory) or 'nccl' (a 4-byte all-reduce)")
        if optimizer not in ("sgd", "adam", "momentum"):
            raise ValueError("optimizer must be 'sgd', 'momentum' or 'adam'")
        if optimizer == "momentum" and transport == "nccl":
            raise ValueError("transport='nccl' supports 'sgd' and 'adam'")
        if transport not in ("peer", "nccl"):
            raise ValueError("transport must be 'peer' or 'nccl'")
        if not torch.cuda.is_available():
            raise RuntimeError("b200ps.ParameterServer needs a CUDA device: there is no CPU fallback")
        self.map = ShardMap(var_numel, num_shards)
        self.plan = build_plan(self.map, num_workers, rank, world, aggregate)
        self.optimizer, self.lr, self.betas, self.eps = optimizer, lr, (beta1, beta2), eps
        self.rank, self.world, self.group, self.transport = rank, world, group, transport
        import os as _os0

        chunk_elems = int(_os0.environ.get("B200PS_HOST_CHUNK_ELEMS", chunk_elems))  # pipeline granularity (experiments)
        self.chunk = max(1024, chunk_elems // 1024 * 1024)
        self.t = 0
        import os as _os

        self._debug_no_mirrors = bool(_os.environ.get("B200PS_DEBUG_NO_MIRRORS"))
        self.trace = None  # a list: the pipelined step appends (label, event) pairs (tools/overlap_timeline.py)
        self.profile = False  # True: step() brackets its phases with CUDA events (see phase_ms)
        self._events: Dict[str, List[Tuple[torch.cuda.Event, torch.cuda.Event]]] = {"fold": [], "apply": []}
        P = self.plan.P
        self.arena = Arena(self.plan.arena_elems + FLAG_WORDS)  # barrier + slice-ready flags live behind the buffers
        self.overlap = overlap and self.plan.two_level and num_shards <= MAX_SLICES
        import os

        geo = os.environ.get("B200PS_OVERLAP_APPLY_GEOMETRY")  # "threads,ctas_per_sm,unroll" -- for experiments
        self.overlap_apply_geometry = tuple(int(x) for x in geo.split(",")) if geo else (256, 2, 2)
        self.barrier_kind, self._epoch, self._step_no = barrier, 0, 0
        self.arena.tensor[self.plan.arena_elems:].zero_()
        self.replica = self.arena.tensor[:P]
        g0 = self.plan.first_grad_slot
        self.partial = self.arena.tensor[P: 2 * P] if self.plan.two_level else None
        self.grads: Dict[int, torch.Tensor] = {w: self.arena.tensor[P * (g0 + i): P * (g0 + i + 1)]
                                               for i, w in enumerate(self.plan.local_workers)}
        self.inv_w = float(np.float32(1.0) / np.float32(num_workers))
        self.state: Dict[int, Tuple[torch.Tensor, torch.Tensor]] = {}
        self.momentum = momentum
        if optimizer in ("adam", "momentum"):  # Adam: (m, v); momentum: (accumulator, nothing)
            for L in self.plan.launches:
                n = L.end - L.begin
                self.state[L.shard] = (torch.zeros(n, dtype=torch.float32, device="cuda"),
                                       torch.zeros(n if optimizer == "adam" else 0, dtype=torch.float32, device="cuda"))
        self.peer_base: Dict[int, int] = {rank: self.arena.ptr}
        self._flag = torch.zeros(1, dtype=torch.int32, device="cuda")
        self._s_in, self._s_out = torch.cuda.Stream(), torch.cuda.Stream()
        self._d2h_done: Dict[int, torch.cuda.Event] = {}  # per chunk: the last step_host's copy-out
        if world > 1:
            self._exchange_handles()
            self._flag_ptrs = _ptrs([self.peer_base[r] + 4 * self.plan.arena_elems_of(r) for r in range(world)])
            if transport == "nccl":
                self._init_nccl()
            import torch.distributed as dist

            torch.cuda.synchronize()
            dist.barrier(group=group)  # every rank has zeroed its flags and mapped every arena

    # ---- set-up -------------------------------------------------------------------------------
    def _exchange_handles(self):
        import torch.distributed as dist

        blobs: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(blobs, self.arena.export(), group=self.group)
        for r, blob in enumerate(blobs):
            if r == self.rank:
                continue
            p = C.c_void_p()
            check(lib.b200ps_ipc_open(C.create_string_buffer(blob, IPC_BYTES), C.byref(p)))
            self.peer_base[r] = p.value

    def _init_nccl(self):
        if not (self.plan.S == self.world == self.plan.W):
            raise ValueError("transport='nccl' is the one-shard-and-one-worker-per-GPU case (S = W = G)")
        sizes = {e - b for b, e in self.map.slices()}
        if len(sizes) != 1:
            raise ValueError("transport='nccl' needs equal slices (P_pad divisible by S * align)")
        self._gsum = torch.empty(sizes.pop(), dtype=torch.float32, device="cuda")

    def close(self):
        torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier(group=self.group)  # nobody unmaps while a peer may still touch the arena
        for r, base in list(self.peer_base.items()):
            if r != self.rank:
                lib.b200ps_ipc_close(C.c_void_p(base))
        self.peer_base = {}
        self.replica, self.grads = None, {}
        self.arena.free()

    # ---- parameters ---------------------------------------------------------------------------
    def init_uniform(self, seed: int = 1):
        """p[i] = u(seed, 0, 0, i): every rank fills its whole replica with the same values."""
        check(lib.b200ps_fill_uniform(C.c_void_p(self.arena.ptr), self.plan.P, seed, 0, 0, 0, self._stream()))
        self._reset_state()

    def init_params(self, host_params: np.ndarray):
        hp = np.ascontiguousarray(host_params, np.float32)
        if hp.size != self.plan.P:
            raise ValueError(f"expected {self.plan.P} padded parameters, got {hp.size}")
        self.replica.copy_(torch.from_numpy(hp))
        self._reset_state()

    def _reset_state(self):
        self.t = 0
        for m, v in self.state.values():
            m.zero_()
            v.zero_()

    # ---- the step -----------------------------------------------------------------------------
    @staticmethod
    def _stream() -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def _barrier(self):
        """Stream-ordered cross-rank barrier: work queued before it on every rank completes before
        work queued after it on any rank starts."""
        if self.world > 1:
            if self.barrier_kind == "peer":
                self._epoch += 1
                check(lib.b200ps_peer_barrier(self._flag_ptrs, self.rank, self.world, self._epoch, self._stream()))
            else:
                import torch.distributed as dist

                dist.all_reduce(self._flag, group=self.group)

    def check_barriers(self):
        """Raises if a peer barrier timed out (some rank never arrived).  Synchronises."""
        if self.world > 1 and self.barrier_kind == "peer":
            check(lib.b200ps_peer_barrier_error(C.c_void_p(self.arena.ptr + 4 * self.plan.arena_elems), self._stream()))

    def _opt_scalars(self) -> Optional[AdamScalars]:
        if self.optimizer != "adam":
            return None
        a = AdamScalars()
        check(lib.b200ps_adam_scalars(self.lr, self.betas[0], self.betas[1], self.eps, self.t, C.byref(a)))
        return a

    def _launch(self, L: Launch, lo: int, hi: int, a: Optional[AdamScalars]):
        """One kernel over elements [lo, hi) of slice L (lo/hi in flat-buffer coordinates)."""
        n = hi - lo
        g = _ptrs([self.peer_base[r] + 4 * (off + lo) for r, off in L.grad_src])
        mirror_ranks = [] if self._debug_no_mirrors else L.mirror_ranks  # timing experiments only: results are wrong
        mir = _ptrs([self.peer_base[r] + 4 * lo for r in mirror_ranks])
        p = C.c_void_p(self.arena.ptr + 4 * lo)
        nbuf = len(L.grad_src)
        if self.optimizer == "momentum":
            acc, _ = self.state[L.shard]
            check(lib.b200ps_dense_apply_momentum(p, C.c_void_p(acc.data_ptr() + 4 * (lo - L.begin)), g, nbuf, n, self.inv_w,
                                                  self.lr, self.momentum, mir, len(mirror_ranks), self._stream()))
        elif a is None:
            check(lib.b200ps_dense_apply_sgd(p, g, nbuf, n, self.inv_w, self.lr, mir, len(mirror_ranks), self._stream()))
        else:
            m, v = self.state[L.shard]
            o = 4 * (lo - L.begin)
            check(lib.b200ps_dense_apply_adam(p, C.c_void_p(m.data_ptr() + o), C.c_void_p(v.data_ptr() + o), g, nbuf, n,
                                              self.inv_w, C.byref(a), mir, len(mirror_ranks), self._stream()))

    def _step_overlapped(self, a):
        """Two-level aggregation pipelined slice by slice over two streams: the fold stream produces
        this GPU's partial sum of slice 0, 1, 2, ... and signals each slice's owner as soon as it is
        done; the apply stream of an owner waits for all producers of a slice, then runs the fused
        apply (NVLink-bound) while the local folds (HBM-bound) of later slices are still running."""
        cur = torch.cuda.current_stream()
        self._step_no += 1
        e, P, S = self._step_no, self.plan.P, self.plan.S
        trace = self.trace  # None, or a list that receives (label, event) pairs of this step

        def tick(label):
            if trace is not None:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                trace.append((label, ev))

        tick("start")
        self._s_in.wait_stream(cur)   # fold stream
        self._s_out.wait_stream(cur)  # apply stream
        flag_base = lambda r: self.peer_base[r] + 4 * (self.plan.arena_elems_of(r) + BARRIER_WORDS)
        with torch.cuda.stream(self._s_in):
            if self.plan.local_workers:
                check(lib.b200ps_set_dense_tuning(256, 8, 2))
                for s in range(S):
                    b, en = self.map.slice(s)
                    if en > b:
                        bufs = _ptrs([self.grads[w].data_ptr() + 4 * b for w in self.plan.local_workers])
                        check(lib.b200ps_dense_fold(C.c_void_p(self.partial.data_ptr() + 4 * b), bufs,
                                                    len(self.plan.local_workers), en - b, self._stream()))
                    owner = placement_gpu(s, self.world)
                    check(lib.b200ps_peer_signal(C.c_void_p(flag_base(owner) + 4 * (s * SLICE_FLAG_STRIDE + self.rank)), e,
                                                 self._stream()))
                    tick(f"fold+signal {s}")
        with torch.cuda.stream(self._s_out):
            # few CTAs: the apply is NVLink-latency-bound and must leave SM slots to the folds
            check(lib.b200ps_set_dense_tuning(*self.overlap_apply_geometry))
            try:
                err = C.c_void_p(self.arena.ptr + 4 * (self.plan.arena_elems + 9))
                for L in self.plan.launches:
                    mask = 0
                    for r, _ in L.grad_src:
                        mask |= 1 << r
                    check(lib.b200ps_peer_wait(C.c_void_p(flag_base(self.rank) + 4 * L.shard * SLICE_FLAG_STRIDE), mask, e, err,
                                               self._stream()))
                    tick(f"wait {L.shard}")
                    if L.end > L.begin:
                        self._launch(L, L.begin, L.end, a)
                    tick(f"apply {L.shard}")
            finally:
                check(lib.b200ps_set_dense_tuning(256, 8, 2))  # the process-wide default geometry
        cur.wait_stream(self._s_in)
        cur.wait_stream(self._s_out)
        tick("joined")
        self._barrier()  # every replica holds every updated slice; partial sums may be overwritten
        tick("barrier")

    def step(self):
        """push + aggregate + apply + pull for every local shard; gradients are already in the
        workers' device buffers (`grad_buffer`)."""
        self.t += 1
        a = self._opt_scalars()
        if self.transport == "nccl" and self.world > 1:
            return self._step_nccl(a)
        if self.overlap and not self.profile:
            return self._step_overlapped(a)
        if self.plan.two_level and self.plan.local_workers:
            # local half of the aggregation: one partial sum per GPU, before anything crosses NVLink
            bufs = _ptrs([self.grads[w].data_ptr() for w in self.plan.local_workers])
            e0 = self._mark()
            check(lib.b200ps_dense_fold(C.c_void_p(self.partial.data_ptr()), bufs, len(self.plan.local_workers),
                                        self.plan.P, self._stream()))
            self._mark("fold", e0)
        self._barrier()  # every worker's gradient (or every GPU's partial sum) is in place
        e0 = self._mark()
        for L in self.plan.launches:
            if L.end > L.begin:
                self._launch(L, L.begin, L.end, a)
        self._mark("apply", e0)
        self._barrier()  # every replica holds every updated slice

    def capture_step(self) -> "torch.cuda.CUDAGraph":
        """The launches of one step as a CUDA graph (replay it instead of calling step()).  For the
        launch-bound end of the size range (BASELINE configs[0], the small end of configs[4]): a
        1 M-parameter step is ~3 us of kernel time behind ~25 us of host-side launch work.  SGD on
        one GPU only: Adam's bias-corrected step size is a launch parameter that changes every
        step, and the cross-rank barriers are not captured."""
        if self.optimizer not in ("sgd", "momentum") or self.world != 1:
            raise ValueError("capture_step supports SGD / momentum on one GPU")
        self.step()  # warm up outside the capture
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.step()
        return g

    def _mark(self, phase: Optional[str] = None, start=None):
        if not self.profile:
            return None
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        if phase is not None:
            self._events[phase].append((start, e))
        return e

    def phase_ms(self, reset: bool = True) -> Dict[str, float]:
        """Mean device time per step of the phases recorded while `profile` was on (synchronises)."""
        torch.cuda.synchronize()
        out = {k: (sum(a.elapsed_time(b) for a, b in v) / len(v) if v else 0.0) for k, v in self._events.items()}
        if reset:
            self._events = {k: [] for k in self._events}
        return out

    def step_async_ordered(self):
        """A reproducible stand-in for TensorFlow's default ASYNCHRONOUS PS training (SURVEY.md 8(f)
        rank 4): every worker's gradient is applied on its own, one after the other in worker order
        (no averaging), so W optimizer steps happen per call and the result does not depend on
        timing.  True asynchrony cannot be parity-checked; this fixed round-robin can.  "flat"
        aggregation only (each launch folds exactly one worker's buffer)."""
        if self.plan.two_level or self.transport == "nccl":
            raise ValueError("step_async_ordered needs aggregate='flat' and transport='peer'")
        self._barrier()
        inv_w, self.inv_w = self.inv_w, 1.0
        try:
            for w in range(self.plan.W):
                self.t += 1
                a = self._opt_scalars()
                for L in self.plan.launches:
                    if L.end > L.begin:
                        one = Launch(L.shard, L.begin, L.end, [L.grad_src[w]], L.mirror_ranks if w == self.plan.W - 1 else [])
                        self._launch(one, L.begin, L.end, a)
        finally:
            self.inv_w = inv_w
        self._barrier()

    def _step_nccl(self, a):
        import torch.distributed as dist

        L = self.plan.launches[0]
        n = L.end - L.begin
        g = self.grads[self.plan.local_workers[0]]
        dist.reduce_scatter_tensor(self._gsum, g, op=dist.ReduceOp.AVG, group=self.group)
        gp = _ptrs([self._gsum.data_ptr()])
        p = C.c_void_p(self.arena.ptr + 4 * L.begin)
        if a is None:  # NCCL already averaged: inv_w = 1
            check(lib.b200ps_dense_apply_sgd(p, gp, 1, n, 1.0, self.lr, _ptrs([]), 0, self._stream()))
        else:
            m, v = self.state[L.shard]
            check(lib.b200ps_dense_apply_adam(p, C.c_void_p(m.data_ptr()), C.c_void_p(v.data_ptr()), gp, 1, n, 1.0,
                                              C.byref(a), _ptrs([]), 0, self._stream()))
        dist.all_gather_into_tensor(self.replica, self.replica[L.begin:L.end], group=self.group)

    # ---- worker-facing ------------------------------------------------------------------------
    def grad_buffer(self, worker: int) -> torch.Tensor:
        """Device buffer a local worker writes its gradient into (zero-copy push)."""
        return self.grads[worker]

    def push(self, worker: int, host_grad: torch.Tensor):
        """H2D of one local worker's gradient (pinned host memory), asynchronous."""
        self.grads[worker].copy_(host_grad, non_blocking=True)

    def pull(self, worker: int, host_out: torch.Tensor):
        """D2H of the replica into a local worker's host buffer, asynchronous."""
        if worker not in self.grads:
            raise KeyError(f"worker {worker} is not placed on rank {self.rank}")
        host_out.copy_(self.replica, non_blocking=True)

    def step_host(self, host_grads: Dict[int, torch.Tensor], host_out: Dict[int, torch.Tensor]):
        """A whole step with HOST buffers: every local worker's gradient goes host -> device, the
        step runs, and the updated parameters go device -> host into every local worker's buffer.
        With one GPU the three phases are pipelined chunk by chunk over three streams, and the
        copy-out of one step overlaps the copy-in of the next; call host_sync() (or synchronise the
        device) before reading the host buffers."""
        if self.world > 1:
            for w, h in host_grads.items():
                self.push(w, h)
            self.step()
            for w, h in host_out.items():
                self.pull(w, h)
            return
        self.t += 1
        a = self._opt_scalars()
        cur = torch.cuda.current_stream()
        # this step's H2D may start once the previous step's kernels are done with the gradient
        # buffers; it does NOT wait for the previous step's D2H (other PCIe direction)
        self._s_in.wait_stream(cur)
        ci = 0
        for L in self.plan.launches:
            for lo in range(L.begin, L.end, self.chunk):
                hi = min(lo + self.chunk, L.end)
                with torch.cuda.stream(self._s_in):
                    for w, h in host_grads.items():
                        self.grads[w][lo:hi].copy_(h[lo:hi], non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(self._s_in)
                cur.wait_event(ev)
                prev = self._d2h_done.get(ci)
                if prev is not None:
                    cur.wait_event(prev)  # the previous step's copy-out of this chunk has left the replica
                self._launch(L, lo, hi, a)
                ev2 = torch.cuda.Event()
                ev2.record(cur)
                self._s_out.wait_event(ev2)
                with torch.cuda.stream(self._s_out):
                    for w, h in host_out.items():
                        h[lo:hi].copy_(self.replica[lo:hi], non_blocking=True)
                    ev3 = torch.cuda.Event()
                    ev3.record(self._s_out)
                self._d2h_done[ci] = ev3
                ci += 1

    def host_sync(self):
        """Blocks until the host buffers of the last step_host() hold the pulled parameters."""
        self._s_out.synchronize()

    # ---- inspection ---------------------------------------------------------------------------
    def dump(self) -> Dict[int, Dict[str, np.ndarray]]:
        """Local slices: {shard: {"range": (b, e), "p": ..., "m": ..., "v": ...}} as NumPy arrays."""
        self.check_barriers()
        torch.cuda.synchronize()
        out = {}
        for L in self.plan.launches:
            d = {"range": (L.begin, L.end), "p": self.replica[L.begin:L.end].cpu().numpy()}
            if L.shard in self.state:
                m, v = self.state[L.shard]
                d["m"] = m.cpu().numpy()  # Adam's m, or the momentum accumulator
                if v.numel():
                    d["v"] = v.cpu().numpy()
            out[L.shard] = d
        return out

    # ---- checkpoint / resume ("storage" is a stated goal, k8s-operator.md:2) ----
    def save(self, directory: str):
        """Writes this rank's shards -- parameters, optimizer state, step counter -- with the shard
        map as header: <directory>/ps-rank<r>-of-<G>.npz.  Every rank calls it."""
        import json
        import os

        os.makedirs(directory, exist_ok=True)
        d = self.dump()
        header = {"format": "b200ps-dense-1", "P_pad": self.plan.P, "shards": self.plan.S, "workers": self.plan.W,
                  "optimizer": self.optimizer, "t": self.t, "rank": self.rank, "world": self.world,
                  "slices": {str(s): list(x["range"]) for s, x in d.items()}}
        arrays = {"header": np.frombuffer(json.dumps(header).encode(), dtype=np.uint8)}
        for s, x in d.items():
            for k in ("p", "m", "v"):
                if k in x:
                    arrays[f"{k}{s}"] = x[k]
        np.savez(os.path.join(directory, f"ps-rank{self.rank}-of-{self.world}.npz"), **arrays)

    def load(self, directory: str):
        """Restores a job saved with the same shard map and world size: every rank reads every
        rank's parameters (its replica holds all slices) and its own optimizer state."""
        import json
        import os

        for r in range(self.world):
            z = np.load(os.path.join(directory, f"ps-rank{r}-of-{self.world}.npz"))
            h = json.loads(bytes(z["header"]).decode())
            if (h["P_pad"], h["shards"], h["optimizer"], h["world"]) != (self.plan.P, self.plan.S, self.optimizer, self.world):
                raise ValueError(f"checkpoint of rank {r} was written for a different job layout: {h}")
            for s, (b, e) in h["slices"].items():
                self.replica[b:e].copy_(torch.from_numpy(z[f"p{s}"]))
                if r == self.rank and int(s) in self.state:
                    m, v = self.state[int(s)]
                    m.copy_(torch.from_numpy(z[f"m{s}"]))
                    if v.numel():
                        v.copy_(torch.from_numpy(z[f"v{s}"]))
            if r == self.rank:
                self.t = int(h["t"])
        torch.cuda.synchronize()

    def algorithmic_bytes_per_step(self) -> int:
        """(4W + 8) n for SGD, (4W + 24) n for Adam, summed over local slices.
        (Two-level aggregation moves more local bytes -- the partial sum is written and re-read --
        but the algorithmic figure is the same: W gradients in, one update out.)"""
        k = 4 * self.plan.W + {"adam": 24, "momentum": 16}.get(self.optimizer, 8)
        return sum(k * (L.end - L.begin) for L in self.plan.launches)
