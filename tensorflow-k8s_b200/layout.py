"""PS variable sharding -- host mirror of the layout group of include/b200ps.h (csrc/layout.cpp).

Reference: none (TensorFlow's partitioners inside the user container); roles per
k8s-operator.md:6; semantics SURVEY.md 8(a) row a4.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence, Tuple

from ._lib import check, lib

ALIGN_ELEMS = 1024


class ShardMap:
    """Flat layout of a variable list, range-partitioned over `nshards` PS shards."""

    def __init__(self, var_numel: Sequence[int], nshards: int, align: int = ALIGN_ELEMS):
        arr = (C.c_uint64 * max(len(var_numel), 1))(*[int(x) for x in var_numel])
        out = C.c_void_p()
        check(lib.b200ps_shard_map_build(arr, len(var_numel), nshards, align, C.byref(out)))
        self._h = out
        self.align = align

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib.b200ps_shard_map_destroy(h)

    @property
    def total(self) -> int:
        return int(lib.b200ps_shard_map_total(self._h))

    @property
    def nshards(self) -> int:
        return check(lib.b200ps_shard_map_nshards(self._h))

    @property
    def nvars(self) -> int:
        return check(lib.b200ps_shard_map_nvars(self._h))

    def slice(self, shard: int) -> Tuple[int, int]:
        b, e = C.c_uint64(), C.c_uint64()
        check(lib.b200ps_shard_map_slice(self._h, shard, C.byref(b), C.byref(e)))
        return b.value, e.value

    def slices(self) -> List[Tuple[int, int]]:
        return [self.slice(s) for s in range(self.nshards)]

    def var_offset(self, var: int) -> int:
        o = C.c_uint64()
        check(lib.b200ps_shard_map_var_offset(self._h, var, C.byref(o)))
        return o.value

    def var_offsets(self) -> List[int]:
        return [self.var_offset(v) for v in range(self.nvars)]


def row_range(rows: int, nshards: int, shard: int) -> Tuple[int, int]:
    b, e = C.c_uint64(), C.c_uint64()
    check(lib.b200ps_row_range(rows, nshards, shard, C.byref(b), C.byref(e)))
    return b.value, e.value


def row_owner(rows: int, nshards: int, row_id: int) -> int:
    return check(lib.b200ps_row_owner(rows, nshards, row_id))
