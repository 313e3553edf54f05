"""Tensor-level wrappers of the kernel group of include/b200ps.h.

torch is plumbing here: it owns the device memory and the stream; every call below goes through
the C ABI with raw device pointers.  There is no eager / CPU fallback: a CPU tensor is an error.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from ._lib import AdamScalars, check, lib


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dev_f32(t: torch.Tensor, name: str) -> int:
    if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
        raise TypeError(f"{name} must be a contiguous float32 CUDA tensor (no CPU fallback exists)")
    return t.data_ptr()


def _ptr_array(ptrs: Sequence[int]):
    return (C.c_void_p * max(len(ptrs), 1))(*ptrs)


def adam_scalars(lr: float, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, t: int = 1) -> AdamScalars:
    a = AdamScalars()
    check(lib.b200ps_adam_scalars(lr, beta1, beta2, eps, t, C.byref(a)))
    return a


def dense_apply_sgd(params: torch.Tensor, grads: Sequence[torch.Tensor], lr: float,
                    mirrors: Sequence[torch.Tensor] = (), n: Optional[int] = None) -> None:
    """params -= lr * mean_w(grads[w]) (fixed worker order), then copy params to every mirror."""
    n = params.numel() if n is None else n
    g = _ptr_array([_dev_f32(t, "grad") for t in grads])
    m = _ptr_array([_dev_f32(t, "mirror") for t in mirrors])
    check(lib.b200ps_dense_apply_sgd(_dev_f32(params, "params"), g, len(grads), n, lr, m, len(mirrors), _stream()))


def dense_apply_adam(params: torch.Tensor, m: torch.Tensor, v: torch.Tensor, grads: Sequence[torch.Tensor],
                     opt: AdamScalars, mirrors: Sequence[torch.Tensor] = (), n: Optional[int] = None) -> None:
    n = params.numel() if n is None else n
    g = _ptr_array([_dev_f32(t, "grad") for t in grads])
    mi = _ptr_array([_dev_f32(t, "mirror") for t in mirrors])
    check(lib.b200ps_dense_apply_adam(_dev_f32(params, "params"), _dev_f32(m, "m"), _dev_f32(v, "v"), g, len(grads), n,
                                      C.byref(opt), mi, len(mirrors), _stream()))


def fill_uniform(dst: torch.Tensor, seed: int, step: int, worker: int, first: int = 0) -> torch.Tensor:
    check(lib.b200ps_fill_uniform(_dev_f32(dst, "dst"), dst.numel(), seed, step, worker, first, _stream()))
    return dst


def set_dense_tuning(threads: int = 0, ctas_per_sm: int = 0, unroll: int = 0) -> None:
    check(lib.b200ps_set_dense_tuning(threads, ctas_per_sm, unroll))


def launch_count() -> int:
    return int(lib.b200ps_launch_count())
