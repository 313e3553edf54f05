"""Tensor-level wrappers of the kernel group of include/b200ps.h.

torch is plumbing here: it owns the device memory and the stream; every call below goes through
the C ABI with raw device pointers.  There is no eager / CPU fallback: a CPU tensor is an error.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
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


def inv_w_f32(total_workers: int) -> float:
    """The fp32 constant 1/W exactly as the oracle forms it (one fp32 division)."""
    return float(np.float32(1.0) / np.float32(total_workers))


def dense_apply_sgd(params: torch.Tensor, grads: Sequence[torch.Tensor], lr: float,
                    mirrors: Sequence[torch.Tensor] = (), n: Optional[int] = None, inv_w: Optional[float] = None) -> None:
    """params -= lr * inv_w * fold(grads) (fixed buffer order), then copy params to every mirror.
    inv_w defaults to 1/len(grads): the mean over the workers."""
    n = params.numel() if n is None else n
    inv_w = inv_w_f32(max(len(grads), 1)) if inv_w is None else inv_w
    g = _ptr_array([_dev_f32(t, "grad") for t in grads])
    m = _ptr_array([_dev_f32(t, "mirror") for t in mirrors])
    check(lib.b200ps_dense_apply_sgd(_dev_f32(params, "params"), g, len(grads), n, inv_w, lr, m, len(mirrors), _stream()))


def dense_apply_adam(params: torch.Tensor, m: torch.Tensor, v: torch.Tensor, grads: Sequence[torch.Tensor],
                     opt: AdamScalars, mirrors: Sequence[torch.Tensor] = (), n: Optional[int] = None,
                     inv_w: Optional[float] = None) -> None:
    n = params.numel() if n is None else n
    inv_w = inv_w_f32(max(len(grads), 1)) if inv_w is None else inv_w
    g = _ptr_array([_dev_f32(t, "grad") for t in grads])
    mi = _ptr_array([_dev_f32(t, "mirror") for t in mirrors])
    check(lib.b200ps_dense_apply_adam(_dev_f32(params, "params"), _dev_f32(m, "m"), _dev_f32(v, "v"), g, len(grads), n,
                                      inv_w, C.byref(opt), mi, len(mirrors), _stream()))


def dense_apply_momentum(params: torch.Tensor, accum: torch.Tensor, grads: Sequence[torch.Tensor], lr: float, momentum: float,
                         mirrors: Sequence[torch.Tensor] = (), n: Optional[int] = None, inv_w: Optional[float] = None) -> None:
    """accum = momentum*accum + g; params -= lr*accum (TensorFlow's ApplyMomentum, no Nesterov)."""
    n = params.numel() if n is None else n
    inv_w = inv_w_f32(max(len(grads), 1)) if inv_w is None else inv_w
    g = _ptr_array([_dev_f32(t, "grad") for t in grads])
    mi = _ptr_array([_dev_f32(t, "mirror") for t in mirrors])
    check(lib.b200ps_dense_apply_momentum(_dev_f32(params, "params"), _dev_f32(accum, "accum"), g, len(grads), n, inv_w, lr,
                                          momentum, mi, len(mirrors), _stream()))


def dense_apply_ftrl(params: torch.Tensor, accum: torch.Tensor, linear: torch.Tensor, grads: Sequence[torch.Tensor], lr: float,
                     l1: float = 0.0, l2: float = 0.0, mirrors: Sequence[torch.Tensor] = (), n: Optional[int] = None,
                     inv_w: Optional[float] = None) -> None:
    """TensorFlow's ApplyFtrl with lr_power = -0.5 (accum starts at 0.1, linear at 0)."""
    n = params.numel() if n is None else n
    inv_w = inv_w_f32(max(len(grads), 1)) if inv_w is None else inv_w
    g = _ptr_array([_dev_f32(t, "grad") for t in grads])
    mi = _ptr_array([_dev_f32(t, "mirror") for t in mirrors])
    check(lib.b200ps_dense_apply_ftrl(_dev_f32(params, "params"), _dev_f32(accum, "accum"), _dev_f32(linear, "linear"), g,
                                      len(grads), n, inv_w, lr, l1, l2, mi, len(mirrors), _stream()))


def dense_fold(out: torch.Tensor, bufs: Sequence[torch.Tensor], n: Optional[int] = None) -> torch.Tensor:
    """out = ((b0 + b1) + b2) + ... : the per-GPU partial sum of the two-level aggregation."""
    n = out.numel() if n is None else n
    b = _ptr_array([_dev_f32(t, "buffer") for t in bufs])
    check(lib.b200ps_dense_fold(_dev_f32(out, "out"), b, len(bufs), n, _stream()))
    return out


def fill_uniform(dst: torch.Tensor, seed: int, step: int, worker: int, first: int = 0) -> torch.Tensor:
    check(lib.b200ps_fill_uniform(_dev_f32(dst, "dst"), dst.numel(), seed, step, worker, first, _stream()))
    return dst


def set_dense_tuning(threads: int = 0, ctas_per_sm: int = 0, unroll: int = 0) -> None:
    check(lib.b200ps_set_dense_tuning(threads, ctas_per_sm, unroll))


def launch_count() -> int:
    return int(lib.b200ps_launch_count())


# ---- sparse group (a10..a12) ---------------------------------------------------------------------
from ._lib import TableGeom  # noqa: E402

import numpy as np  # noqa: E402


def table_geom(rows: int, dim: int, nshards: int, rank: int = 0, world: int = 1) -> TableGeom:
    return TableGeom(rows, dim, nshards, rank, world)


def _dev(t: torch.Tensor, dtype, name: str) -> int:
    if not (t.is_cuda and t.dtype == dtype and t.is_contiguous()):
        raise TypeError(f"{name} must be a contiguous {dtype} CUDA tensor (no CPU fallback exists)")
    return t.data_ptr()


def _ptr_or_int(x, dtype, name):
    return x if isinstance(x, int) else _dev(x, dtype, name)


def sparse_gather(out: torch.Tensor, shards: Sequence, geom: TableGeom, ids: torch.Tensor) -> torch.Tensor:
    """out[k, :] = table[ids[k], :]; `shards[s]` is shard s (tensor or raw peer pointer)."""
    sb = _ptr_array([_ptr_or_int(t, torch.float32, "shard") for t in shards])
    check(lib.b200ps_sparse_gather(_dev(out, torch.float32, "out"), sb, C.byref(geom), _dev(ids, torch.int64, "ids"),
                                   ids.numel(), _stream()))
    return out


def sparse_workspace(n_total: int, dim: int) -> torch.Tensor:
    """Scratch for one push of n_total = W * ids_per_worker entries of a dim-wide table."""
    return torch.empty(int(lib.b200ps_sparse_workspace_bytes(n_total, dim)), dtype=torch.uint8, device="cuda")


def sparse_sort_segment(ids: Sequence, ids_per_worker: int, geom: TableGeom, workspace: torch.Tensor) -> None:
    ip = _ptr_array([_ptr_or_int(t, torch.int64, "ids") for t in ids])
    check(lib.b200ps_sparse_sort_segment(ip, len(ids), ids_per_worker, C.byref(geom), workspace.data_ptr(),
                                         workspace.numel(), _stream()))


def sparse_apply_sgd(shards: Sequence, grad_rows: Sequence, ids_per_worker: int, geom: TableGeom, lr: float,
                     workspace: torch.Tensor) -> None:
    sb = _ptr_array([_ptr_or_int(t, torch.float32, "shard") if t is not None else 0 for t in shards])
    gp = _ptr_array([_ptr_or_int(t, torch.float32, "grad_rows") for t in grad_rows])
    check(lib.b200ps_sparse_apply_sgd(sb, gp, len(grad_rows), ids_per_worker, C.byref(geom), lr, workspace.data_ptr(),
                                      _stream()))


def sparse_apply_adagrad(shards: Sequence, accums: Sequence, grad_rows: Sequence, ids_per_worker: int, geom: TableGeom,
                         lr: float, workspace: torch.Tensor) -> None:
    sb = _ptr_array([_ptr_or_int(t, torch.float32, "shard") if t is not None else 0 for t in shards])
    ab = _ptr_array([_ptr_or_int(t, torch.float32, "accum") if t is not None else 0 for t in accums])
    gp = _ptr_array([_ptr_or_int(t, torch.float32, "grad_rows") for t in grad_rows])
    check(lib.b200ps_sparse_apply_adagrad(sb, ab, gp, len(grad_rows), ids_per_worker, C.byref(geom), lr,
                                          workspace.data_ptr(), _stream()))


def sparse_plan_read(workspace: torch.Tensor, n_total: int, rows: int) -> dict:
    """The plan of the last sort_segment as NumPy arrays (debug accessor; synchronises)."""
    kept, uniq, bad = C.c_uint64(), C.c_uint64(), C.c_uint64()
    s = np.empty(n_total, np.int64)
    perm = np.empty(n_total, np.uint32)
    seg = np.empty(n_total, np.int32)
    un = np.empty(n_total, np.int64)
    off = np.empty(n_total + 1, np.uint32)
    check(lib.b200ps_sparse_plan_read(workspace.data_ptr(), n_total, rows, C.byref(kept), C.byref(uniq), C.byref(bad),
                                      s.ctypes.data, perm.ctypes.data, seg.ctypes.data, un.ctypes.data, off.ctypes.data,
                                      _stream()))
    k, u = kept.value, uniq.value
    return {"n_kept": k, "n_unique": u, "n_bad": bad.value, "sorted_ids": s[:k], "perm": perm[:k], "seg_id": seg[:k],
            "unique_ids": un[:u], "seg_offset": off[:u + 1]}


NOALLOC_MAX_ENTRIES = 262144  # B200PS_NOALLOC_MAX_ENTRIES


def sparse_push_sgd_noalloc(shards: Sequence, ids: Sequence, grad_rows: Sequence, ids_per_worker: int, geom: TableGeom,
                            lr: float) -> None:
    """A whole sparse SGD push (sort_segment + apply_sgd) that allocates nothing on the device:
    for tasks whose HBM is fully handed out at launch and for step loops that cannot afford the
    device-wide stall of cudaMalloc.  The scratch is a fixed array inside the library, so W *
    ids_per_worker is limited to NOALLOC_MAX_ENTRIES, and calls that may overlap on the GPU must
    be ordered by the caller (they share the array).  Make the first call before any stream
    capture."""
    sb = _ptr_array([_ptr_or_int(t, torch.float32, "shard") if t is not None else 0 for t in shards])
    ip = _ptr_array([_ptr_or_int(t, torch.int64, "ids") for t in ids])
    gp = _ptr_array([_ptr_or_int(t, torch.float32, "grad_rows") for t in grad_rows])
    check(lib.b200ps_sparse_push_sgd_noalloc(sb, ip, gp, len(ids), ids_per_worker, C.byref(geom), lr, _stream()))


# ---- worker group: a real worker for BASELINE configs[0] ------------------------------------------
def linreg_grad(grad: torch.Tensor, params: torch.Tensor, batch: int, seed_x: int, seed_w: int, step: int, worker: int,
                scratch: Optional[torch.Tensor] = None, n: Optional[int] = None) -> torch.Tensor:
    """grad = d/dp of (1/2B) sum_b (x_b . p - y_b)^2 on the synthetic mini-batch of (step, worker)."""
    n = params.numel() if n is None else n
    need = int(lib.b200ps_linreg_scratch_bytes(n, batch))
    if scratch is None:
        scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
    check(lib.b200ps_linreg_grad(_dev_f32(grad, "grad"), _dev_f32(params, "params"), n, batch, seed_x, seed_w, step, worker,
                                 scratch.data_ptr(), scratch.numel(), _stream()))
    return grad


def sparse_serve(outs: Sequence, shards: Sequence, ids: Sequence, ids_per_worker: int, geom: TableGeom) -> None:
    """Owner side of pull for all workers: rows of this GPU's shards -> outs[w] (tensor or raw peer pointer)."""
    op = _ptr_array([_ptr_or_int(t, torch.float32, "out") for t in outs])
    sb = _ptr_array([_ptr_or_int(t, torch.float32, "shard") if t is not None else 0 for t in shards])
    ip = _ptr_array([_ptr_or_int(t, torch.int64, "ids") for t in ids])
    check(lib.b200ps_sparse_serve(op, sb, ip, len(ids), ids_per_worker, C.byref(geom), _stream()))


def sparse_route(grad_rows: torch.Tensor, ids: torch.Tensor, geom: TableGeom, dst_by_gpu: Sequence) -> None:
    """Worker side of push: rows owned by GPU o go to dst_by_gpu[o] (None for this GPU)."""
    dp = _ptr_array([_ptr_or_int(t, torch.float32, "staging") if t is not None else 0 for t in dst_by_gpu])
    check(lib.b200ps_sparse_route(_dev(grad_rows, torch.float32, "grad_rows"), _dev(ids, torch.int64, "ids"), ids.numel(),
                                  C.byref(geom), dp, _stream()))
