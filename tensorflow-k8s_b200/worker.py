"""Host mirror of the MLP worker group of include/b200ps.h: the compute tasks of BASELINE
configs[1] that live on this GPU, as one `b200ps_mlp_t` (csrc/mlp_worker.cu, csrc/gemm_sm100.cuh).

`MlpWorker.step` takes the mini-batches of the local compute tasks in pinned host memory, runs
forward and backward for all of them on the GPU and leaves each task's fp32 gradient in its
gradient buffer of the ParameterServer's arena (zero-copy push); the parameters are read from the
replica (zero-copy pull).  Reference: none -- "WORKER computes" (k8s-operator.md:6) is the user's
TensorFlow program; SURVEY.md 8(f) rank 1.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from ._lib import check, lib

C1_DIMS = (1024, 8192, 8192, 2048, 4096)


def mlp_variables(dims: Sequence[int]) -> List[int]:
    """Variable sizes in declaration order: W_0 [d_0][d_1], b_0 [d_1], W_1, b_1, ..."""
    out: List[int] = []
    for a, b in zip(dims[:-1], dims[1:]):
        out += [a * b, b]
    return out


class MlpWorker:
    def __init__(self, ps, dims: Sequence[int] = C1_DIMS, batch_per_task: int = 512):
        self.ps, self.dims, self.batch = ps, tuple(dims), batch_per_task
        self.tasks = list(ps.plan.local_workers)
        offs = ps.map.var_offsets()
        if len(offs) != 2 * (len(dims) - 1):
            raise ValueError("the parameter server's variable list is not this MLP's (W_0, b_0, W_1, b_1, ...)")
        d = (C.c_int * len(dims))(*dims)
        vo = (C.c_uint64 * len(offs))(*offs)
        h = C.c_void_p()
        check(lib.b200ps_mlp_create(d, len(dims), batch_per_task, len(self.tasks), vo, C.byref(h)))
        self._h = h
        self._grads = (C.c_void_p * len(self.tasks))(*[ps.grad_buffer(w).data_ptr() for w in self.tasks])
        self.rows = batch_per_task * len(self.tasks)
        self.loss = torch.zeros(len(self.tasks), dtype=torch.float32).pin_memory()

    def step(self, host_x: torch.Tensor, host_t: torch.Tensor, want_loss: bool = True) -> torch.Tensor:
        """host_x: bf16 [tasks * batch, d_0], host_t: bf16 [tasks * batch, d_L], pinned.  Asynchronous
        on the current stream; `self.loss` (pinned) holds one loss per task once the stream has passed."""
        for t, cols in ((host_x, self.dims[0]), (host_t, self.dims[-1])):
            if t.dtype != torch.bfloat16 or tuple(t.shape) != (self.rows, cols) or not t.is_contiguous():
                raise TypeError(f"expected a contiguous bf16 host tensor of shape {(self.rows, cols)}")
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        check(lib.b200ps_mlp_step(self._h, C.c_void_p(self.ps.replica.data_ptr()), self._grads, C.c_void_p(host_x.data_ptr()),
                                  C.c_void_p(host_t.data_ptr()), C.c_void_p(self.loss.data_ptr()) if want_loss else None, st))
        return self.loss

    def flops_per_step(self) -> int:
        """2 * rows * sum(in * out) for forward, the same for the weight gradient, and for the data
        gradient of every layer but the first."""
        mac = [a * b for a, b in zip(self.dims[:-1], self.dims[1:])]
        return 2 * self.rows * (2 * sum(mac) + sum(mac[1:]))

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            check(lib.b200ps_mlp_destroy(h))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, out_t: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[M][N] = a[M][K] @ b[N][K]^T through the worker's tcgen05 kernel (bf16 in, fp32 accumulate, bf16 out)."""
    M, K = a.shape
    N = b.shape[0]
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=a.device)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib.b200ps_gemm_bf16(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_void_p(out.data_ptr()),
                               C.c_void_p(out_t.data_ptr()) if out_t is not None else None, M, N, K, st))
    return out
