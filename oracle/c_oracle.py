"""ctypes binding of oracle/libps_oracle.so (the C oracle).  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED -- see oracle/ps_oracle.h.  Imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py; never by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libps_oracle.so")


def build(force: bool = False) -> str:
    src = [os.path.join(_HERE, f) for f in ("ps_oracle.c", "ps_oracle.h", "Makefile")]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


class AdamT(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("lr_t", "beta1", "beta2", "omb1", "omb2", "eps")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        u64, i32, f32, vp = C.c_uint64, C.c_int, C.c_float, C.c_void_p
        L.ps_hash64.restype = u64
        L.ps_hash64.argtypes = [u64] * 4
        L.ps_uniform.restype = f32
        L.ps_uniform.argtypes = [u64] * 4
        L.ps_fill_uniform.argtypes = [vp, u64, u64, u64, u64, u64, i32]
        L.ps_shard_map.restype = u64
        L.ps_shard_map.argtypes = [vp, i32, i32, C.c_uint32, vp, vp, vp]
        L.ps_row_range.argtypes = [u64, i32, i32, vp, vp]
        L.ps_row_owner.restype = i32
        L.ps_row_owner.argtypes = [u64, i32, u64]
        L.ps_dense_sgd.argtypes = [vp, vp, i32, u64, f32, f32, i32]
        L.ps_dense_fold.argtypes = [vp, vp, i32, u64, i32]
        L.ps_dense_momentum.argtypes = [vp, vp, vp, i32, u64, f32, f32, f32, i32]
        L.ps_dense_ftrl.argtypes = [vp, vp, vp, vp, i32, u64, f32, f32, f32, f32, i32]
        L.ps_adam_scalars.restype = AdamT
        L.ps_adam_scalars.argtypes = [C.c_double] * 4 + [u64]
        L.ps_dense_adam.argtypes = [vp, vp, vp, vp, i32, u64, f32, C.POINTER(AdamT), i32]
        L.ps_gather.argtypes = [vp, vp, vp, u64, i32, C.c_int64]
        L.ps_bucket_by_owner.argtypes = [vp, u64, u64, i32, vp, vp]
        L.ps_sort_segment.restype = u64
        L.ps_sort_segment.argtypes = [vp, u64, vp, vp, vp, vp, vp]
        L.ps_sparse_apply_sgd.argtypes = [vp, i32, C.c_int64, vp, vp, vp, vp, u64, f32, f32]
        L.ps_sparse_apply_adagrad.argtypes = [vp, vp, i32, C.c_int64, vp, vp, vp, vp, u64, f32, f32]
        L.ps_linreg_grad.argtypes = [vp, vp, u64, i32, u64, u64, u64, u64, i32]
        L.ps_oracle_max_threads.restype = i32
        L.ps_oracle_set_threads.argtypes = [i32]
        _lib = L
    return _lib


def _p(a: np.ndarray):
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


def max_threads() -> int:
    return int(lib().ps_oracle_max_threads())


def use_all_cores() -> int:
    """Use every core this process may run on, whatever OMP_NUM_THREADS says (torchrun exports 1)."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().ps_oracle_set_threads(n)
    return n


def fill_uniform(n, seed, step, worker, first=0, threads=1) -> np.ndarray:
    out = np.empty(n, np.float32)
    lib().ps_fill_uniform(_p(out), n, seed, step, worker, first, threads)
    return out


def shard_map(var_numel, nshards, align):
    vn = np.asarray(var_numel, np.uint64)
    off = np.zeros(len(vn), np.uint64)
    b = np.zeros(nshards, np.uint64)
    e = np.zeros(nshards, np.uint64)
    ppad = lib().ps_shard_map(_p(vn), len(vn), nshards, align, _p(off), _p(b), _p(e))
    return int(ppad), [int(x) for x in off], [(int(x), int(y)) for x, y in zip(b, e)]


def row_range(rows, nshards, shard):
    b, e = C.c_uint64(), C.c_uint64()
    lib().ps_row_range(rows, nshards, shard, C.addressof(b), C.addressof(e))
    return b.value, e.value


def row_owner(rows, nshards, idv) -> int:
    return int(lib().ps_row_owner(rows, nshards, idv))


def _gradptrs(grads):
    arr = (C.c_void_p * len(grads))(*[_p(g) for g in grads])
    return arr


def dense_fold(bufs, threads=1):
    out = np.empty_like(bufs[0])
    gp = _gradptrs(bufs)
    lib().ps_dense_fold(_p(out), C.addressof(gp), len(bufs), out.size, threads)
    return out


def two_level_partials(grads, world, threads=1):
    """Per-GPU partial sums for the placement worker w -> GPU w mod world (GPUs without workers are skipped)."""
    return [dense_fold(grads[k::world], threads) for k in range(world) if grads[k::world]]


def dense_sgd(p, grads, lr, threads=1, inv_w=None):
    """In place on p (np.float32)."""
    gp = _gradptrs(grads)
    inv_w = np.float32(1.0) / np.float32(len(grads)) if inv_w is None else np.float32(inv_w)
    lib().ps_dense_sgd(_p(p), C.addressof(gp), len(grads), p.size, inv_w, lr, threads)
    return p


def dense_momentum(p, accum, grads, lr, momentum, threads=1, inv_w=None):
    gp = _gradptrs(grads)
    inv_w = np.float32(1.0) / np.float32(len(grads)) if inv_w is None else np.float32(inv_w)
    lib().ps_dense_momentum(_p(p), _p(accum), C.addressof(gp), len(grads), p.size, inv_w, lr, momentum, threads)
    return p, accum


def dense_ftrl(p, accum, linear, grads, lr, l1, l2, threads=1, inv_w=None):
    gp = _gradptrs(grads)
    inv_w = np.float32(1.0) / np.float32(len(grads)) if inv_w is None else np.float32(inv_w)
    lib().ps_dense_ftrl(_p(p), _p(accum), _p(linear), C.addressof(gp), len(grads), p.size, inv_w, lr, l1, l2, threads)
    return p, accum, linear


def adam_scalars(lr, beta1, beta2, eps, t) -> AdamT:
    return lib().ps_adam_scalars(lr, beta1, beta2, eps, t)


def dense_adam(p, m, v, grads, a: AdamT, threads=1, inv_w=None):
    gp = _gradptrs(grads)
    inv_w = np.float32(1.0) / np.float32(len(grads)) if inv_w is None else np.float32(inv_w)
    lib().ps_dense_adam(_p(p), _p(m), _p(v), C.addressof(gp), len(grads), p.size, inv_w, C.byref(a), threads)
    return p, m, v


def linreg_grad(params, batch, seed_x, seed_w, step, worker, threads=1):
    """Gradient of the mean squared error of one synthetic mini-batch (BASELINE configs[0] worker)."""
    g = np.empty_like(params)
    lib().ps_linreg_grad(_p(g), _p(params), params.size, batch, seed_x, seed_w, step, worker, threads)
    return g


def gather(table, ids, row_base=0):
    ids = np.ascontiguousarray(ids, np.int64)
    out = np.empty((len(ids), table.shape[1]), np.float32)
    lib().ps_gather(_p(out), _p(table), _p(ids), len(ids), table.shape[1], row_base)
    return out


def bucket_by_owner(ids, rows, nshards):
    ids = np.ascontiguousarray(ids, np.int64)
    perm = np.empty(len(ids), np.uint32)
    off = np.zeros(nshards + 1, np.uint64)
    lib().ps_bucket_by_owner(_p(ids), len(ids), rows, nshards, _p(perm), _p(off))
    return perm, off


def sort_segment(ids):
    ids = np.ascontiguousarray(ids, np.int64)
    n = len(ids)
    s = np.empty(n, np.int64)
    perm = np.empty(n, np.uint32)
    seg = np.empty(n, np.int32)
    uniq = np.empty(n, np.int64)
    off = np.zeros(n + 1, np.uint32)
    nu = int(lib().ps_sort_segment(_p(ids), n, _p(s), _p(perm), _p(seg), _p(uniq), _p(off)))
    return s, perm, seg, uniq[:nu].copy(), off[: nu + 1].copy()


def sparse_apply_sgd(table, row_base, grad_rows, ids, inv_w, lr):
    _, perm, _, uniq, off = sort_segment(ids)
    lib().ps_sparse_apply_sgd(_p(table), table.shape[1], row_base, _p(grad_rows), _p(perm), _p(uniq), _p(off),
                              len(uniq), inv_w, lr)
    return table


def sparse_apply_adagrad(table, accum, row_base, grad_rows, ids, inv_w, lr):
    _, perm, _, uniq, off = sort_segment(ids)
    lib().ps_sparse_apply_adagrad(_p(table), _p(accum), table.shape[1], row_base, _p(grad_rows), _p(perm),
                                  _p(uniq), _p(off), len(uniq), inv_w, lr)
    return table, accum
