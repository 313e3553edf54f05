"""NumPy restatement of the parameter-server step.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the reference (yph152/tensorflow-k8s) is documentation only -- its one sentence
about this path is k8s-operator.md:6 -- so there is no reference output to pin against.  This file
is an independent second statement of SURVEY.md section 8(a) rows a4..a12, written array-at-a-time
in np.float32 (one IEEE rounding per operation, no FMA), and is compared element-wise with the C
oracle (oracle/ps_oracle.c) by tests/test_oracle_*.py.  The product never imports it.
"""
from __future__ import annotations

import numpy as np

_M64 = (1 << 64) - 1
_U = np.uint64


def hash64(seed: int, step: int, worker: int, index) -> np.ndarray:
    """SURVEY 8(d) synthetic inputs: packed counter -> splitmix64 finaliser (vectorised)."""
    idx = np.asarray(index, dtype=np.uint64)
    key = ((seed * 0xD6E8FEB86659FD93) ^ (step * 0xA24BAED4963EE407) ^ (worker * 0x9FB21C651E98DF25)) & _M64
    with np.errstate(over="ignore"):
        z = idx ^ _U(key)
        z = z + _U(0x9E3779B97F4A7C15)
        z = (z ^ (z >> _U(30))) * _U(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> _U(27))) * _U(0x94D049BB133111EB)
        return z ^ (z >> _U(31))


def uniform(seed: int, step: int, worker: int, first: int, n: int) -> np.ndarray:
    k = (hash64(seed, step, worker, np.arange(first, first + n, dtype=np.uint64)) >> _U(40)).astype(np.float32)
    return k * np.float32(2.0 ** -23) - np.float32(1.0)


# ---- a4 -----------------------------------------------------------------------------------------

def shard_map(var_numel, nshards: int, align: int):
    """Returns (P_pad, var_offset[], [(begin, end)] per shard)."""
    ru = lambda x: -(-x // align) * align
    offs, off = [], 0
    for n in var_numel:
        offs.append(off)
        off += ru(int(n))
    c = ru(-(-off // nshards))
    return off, offs, [(min(s * c, off), min((s + 1) * c, off)) for s in range(nshards)]


def row_range(rows: int, nshards: int, shard: int):
    q, r = divmod(rows, nshards)
    b = shard * q + min(shard, r)
    return b, b + q + (1 if shard < r else 0)


def row_owner(rows: int, nshards: int, ids) -> np.ndarray:
    ids = np.asarray(ids, dtype=np.int64)
    q, r = divmod(rows, nshards)
    t = r * (q + 1)
    lo = ids // (q + 1)
    hi = r + (ids - t) // max(q, 1)
    return np.where(ids < t, lo, hi).astype(np.int32)


# ---- a6..a8 -------------------------------------------------------------------------------------

def aggregate(grads, inv_w) -> np.ndarray:
    acc = np.array(grads[0], dtype=np.float32, copy=True)
    for g in grads[1:]:
        acc = acc + np.asarray(g, dtype=np.float32)
    return acc * np.float32(inv_w)


def fold(bufs) -> np.ndarray:
    """((b0 + b1) + b2) + ... without scaling (the per-GPU partial of the two-level aggregation)."""
    return aggregate(bufs, np.float32(1.0))


def dense_sgd(p, grads, lr, inv_w=None):
    g = aggregate(grads, np.float32(1.0) / np.float32(len(grads)) if inv_w is None else inv_w)
    return p - np.float32(lr) * g


def dense_momentum(p, accum, grads, lr, momentum, inv_w=None):
    g = aggregate(grads, np.float32(1.0) / np.float32(len(grads)) if inv_w is None else inv_w)
    a2 = np.float32(momentum) * accum + g
    return p - np.float32(lr) * a2, a2


def dense_ftrl(p, accum, linear, grads, lr, l1, l2, inv_w=None):
    """TF ApplyFtrl, lr_power = -0.5 (same operation order as ps_dense_ftrl)."""
    f = np.float32
    g = aggregate(grads, f(1.0) / f(len(grads)) if inv_w is None else inv_w)
    lr, l1, two_l2 = f(lr), f(l1), f(2.0) * f(l2)
    na = accum + g * g
    sn = np.sqrt(na)
    sigma = (sn - np.sqrt(accum)) / lr
    li = linear + (g - sigma * p)
    x = np.where(li > 0, l1, np.where(li < 0, -l1, f(0))).astype(f) - li
    y = sn / lr + two_l2
    p2 = np.where(np.abs(li) > l1, x / y, f(0)).astype(f)
    return p2, na.astype(f), li.astype(f)


def adam_scalars(lr: float, beta1: float, beta2: float, eps: float, t: int):
    lr_t = np.float32(lr * np.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t))
    return dict(lr_t=lr_t, beta1=np.float32(beta1), beta2=np.float32(beta2),
                omb1=np.float32(1.0 - beta1), omb2=np.float32(1.0 - beta2), eps=np.float32(eps))


def dense_adam(p, m, v, grads, a, inv_w=None):
    g = aggregate(grads, np.float32(1.0) / np.float32(len(grads)) if inv_w is None else inv_w)
    m2 = a["beta1"] * m + a["omb1"] * g
    v2 = a["beta2"] * v + a["omb2"] * (g * g)
    p2 = p - (a["lr_t"] * m2) / (np.sqrt(v2) + a["eps"])
    return p2, m2, v2


# ---- 8(f).1: linear-regression worker -----------------------------------------------------------

def linreg_grad(params, batch, seed_x, seed_w, step, worker, block=256):
    """Same fixed-order double arithmetic as ps_linreg_grad, array-at-a-time."""
    n = len(params)
    d = params.astype(np.float64) - uniform(seed_w, 0, 0, 0, n).astype(np.float64)
    nblk = -(-n // block)
    X = np.stack([uniform(seed_x, step, worker * batch + b, 0, n) for b in range(batch)]).astype(np.float64)
    prod = np.zeros((batch, nblk * block))
    prod[:, :n] = X * d
    blocks = prod.reshape(batch, nblk, block)
    acc = np.zeros((batch, nblk))
    for j in range(block):  # left fold inside each block (padding adds +0.0), then over the blocks
        acc = acc + blocks[:, :, j]
    r = np.zeros(batch)
    for k in range(nblk):
        r = r + acc[:, k]
    g = np.zeros(n)
    for b in range(batch):
        g = g + X[b] * r[b]
    return (g * (1.0 / batch)).astype(np.float32)


# ---- a10..a12 -----------------------------------------------------------------------------------

def sort_segment(ids):
    """Stable sort by id; returns sorted_ids, perm, seg_id, unique_ids, seg_offset."""
    ids = np.asarray(ids, dtype=np.int64)
    perm = np.argsort(ids, kind="stable").astype(np.uint32)
    s = ids[perm]
    if len(s) == 0:
        z = np.zeros(0, np.int64)
        return z, perm, np.zeros(0, np.int32), z, np.zeros(1, np.uint32)
    head = np.concatenate(([True], s[1:] != s[:-1]))
    seg_id = (np.cumsum(head) - 1).astype(np.int32)
    starts = np.flatnonzero(head).astype(np.uint32)
    return s, perm, seg_id, s[head], np.concatenate((starts, [np.uint32(len(s))])).astype(np.uint32)


SEG_CHUNK = 256


def segment_mean_rows(grad_rows, perm, seg_offset, inv_w):
    """Fixed-order fold of the rows of each segment in sorted order, then one multiply by inv_w.
    Up to SEG_CHUNK entries: a plain left fold.  Longer: each chunk of SEG_CHUNK consecutive
    entries left to right, then the chunk sums left to right."""
    nu = len(seg_offset) - 1
    out = np.empty((nu, grad_rows.shape[1]), np.float32)
    for u in range(nu):
        a, b = int(seg_offset[u]), int(seg_offset[u + 1])
        total = None
        for c0 in range(a, b, SEG_CHUNK):
            acc = grad_rows[perm[c0]].astype(np.float32, copy=True)
            for j in range(c0 + 1, min(c0 + SEG_CHUNK, b)):
                acc = acc + grad_rows[perm[j]]
            total = acc if total is None else total + acc
        out[u] = total * np.float32(inv_w)
    return out


def sparse_apply_sgd(table, row_base, grad_rows, ids, inv_w, lr):
    _, perm, _, uniq, off = sort_segment(ids)
    g = segment_mean_rows(grad_rows, perm, off, inv_w)
    r = uniq - row_base
    table[r] = table[r] - np.float32(lr) * g
    return table


def sparse_apply_adagrad(table, accum, row_base, grad_rows, ids, inv_w, lr):
    _, perm, _, uniq, off = sort_segment(ids)
    g = segment_mean_rows(grad_rows, perm, off, inv_w)
    r = uniq - row_base
    a2 = accum[r] + g * g
    accum[r] = a2
    table[r] = table[r] - (np.float32(lr) * g) / np.sqrt(a2)
    return table, accum
