"""ctypes binding of libb200ps.so -- every symbol include/b200ps.h declares.

There is no fallback: if the library has not been built (`python -c "import __graft_entry__ as g;
g.build()"` or `make -C tensorflow-k8s_b200/csrc`) importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200ps.so")


class B200PSError(RuntimeError):
    """A libb200ps call returned a negative status; `.status` holds it."""

    def __init__(self, status: int, message: str):
        super().__init__(f"b200ps status {status}: {message}")
        self.status = status


class TableGeom(C.Structure):
    _fields_ = [("rows", C.c_uint64), ("dim", C.c_int), ("nshards", C.c_int), ("rank", C.c_int), ("world", C.c_int)]


class AdamScalars(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("lr_t", "beta1", "beta2", "omb1", "omb2", "eps")]


class DenseOptions(C.Structure):
    """b200ps_dense_options_t"""
    _fields_ = [("optimizer", C.c_int), ("aggregate", C.c_int), ("lr", C.c_double), ("beta1", C.c_double),
                ("beta2", C.c_double), ("eps", C.c_double), ("momentum", C.c_double), ("ftrl_l1", C.c_double),
                ("ftrl_l2", C.c_double), ("ftrl_accum_init", C.c_double), ("threads", C.c_int), ("ctas_per_sm", C.c_int),
                ("unroll", C.c_int), ("use_graph", C.c_int), ("pipeline_folds", C.c_int), ("host_chunk_elems", C.c_uint64)]


OPT_SGD, OPT_ADAM, OPT_MOMENTUM, OPT_FTRL, OPT_ADAGRAD = 0, 1, 2, 3, 4


class TableOptions(C.Structure):
    """b200ps_table_options_t"""
    _fields_ = [("optimizer", C.c_int), ("use_graph", C.c_int), ("lr", C.c_double), ("accum_init", C.c_double)]

AGG_FLAT, AGG_TWO_LEVEL, AGG_MULTIMEM = 0, 1, 2

OK, ERR_INVALID, ERR_PARSE, ERR_VALIDATION, ERR_CUDA, ERR_RANGE, ERR_STATE = 0, -1, -2, -3, -4, -5, -6

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make -C {os.path.join(_HERE, 'csrc')}` "
        "(there is no CPU or PyTorch fallback for the parameter-server kernels)")

lib = C.CDLL(LIB_PATH)

_u64, _i32, _u32, _f32, _vp, _cp, _sz = C.c_uint64, C.c_int, C.c_uint32, C.c_float, C.c_void_p, C.c_char_p, C.c_size_t
_pp = C.POINTER(C.c_void_p)

# name -> (restype, argtypes); mirrors include/b200ps.h one to one (tests/test_abi.py checks it)
# (TableGeom / AdamScalars mirror the structs of the header)
SIGNATURES = {
    "b200ps_last_error": (_cp, []),
    "b200ps_version": (_i32, []),
    "b200ps_cluster_from_tfjob": (_i32, [_cp, _pp]),
    "b200ps_cluster_from_tf_config": (_i32, [_cp, _pp, _vp, _sz, C.POINTER(_i32)]),
    "b200ps_cluster_destroy": (None, [_vp]),
    "b200ps_cluster_num_tasks": (_i32, [_vp, _cp]),
    "b200ps_cluster_task_address": (_i32, [_vp, _cp, _i32, _vp, _sz]),
    "b200ps_tf_config_for": (_i32, [_vp, _cp, _i32, _vp, _sz]),
    "b200ps_cluster_num_shards": (_i32, [_vp]),
    "b200ps_cluster_num_workers": (_i32, [_vp]),
    "b200ps_cluster_worker_rank": (_i32, [_vp, _cp, _i32]),
    "b200ps_placement_gpu": (_i32, [_i32, _i32]),
    "b200ps_shard_map_build": (_i32, [_vp, _i32, _i32, _u32, _pp]),
    "b200ps_shard_map_destroy": (None, [_vp]),
    "b200ps_shard_map_total": (_u64, [_vp]),
    "b200ps_shard_map_nshards": (_i32, [_vp]),
    "b200ps_shard_map_nvars": (_i32, [_vp]),
    "b200ps_shard_map_slice": (_i32, [_vp, _i32, C.POINTER(_u64), C.POINTER(_u64)]),
    "b200ps_shard_map_var_offset": (_i32, [_vp, _i32, C.POINTER(_u64)]),
    "b200ps_row_range": (_i32, [_u64, _i32, _i32, C.POINTER(_u64), C.POINTER(_u64)]),
    "b200ps_row_owner": (_i32, [_u64, _i32, _u64]),
    "b200ps_adam_scalars": (_i32, [C.c_double] * 4 + [_u64, C.POINTER(AdamScalars)]),
    "b200ps_dense_apply_sgd": (_i32, [_vp, _vp, _i32, _u64, _f32, _f32, _vp, _i32, _vp]),
    "b200ps_dense_apply_adam": (_i32, [_vp, _vp, _vp, _vp, _i32, _u64, _f32, C.POINTER(AdamScalars), _vp, _i32, _vp]),
    "b200ps_dense_fold": (_i32, [_vp, _vp, _i32, _u64, _vp]),
    "b200ps_dense_apply_momentum": (_i32, [_vp, _vp, _vp, _i32, _u64, _f32, _f32, _f32, _vp, _i32, _vp]),
    "b200ps_dense_apply_ftrl": (_i32, [_vp, _vp, _vp, _vp, _i32, _u64, _f32, _f32, _f32, _f32, _vp, _i32, _vp]),
    "b200ps_fill_uniform": (_i32, [_vp, _u64, _u64, _u64, _u64, _u64, _vp]),
    "b200ps_set_dense_tuning": (_i32, [_i32, _i32, _i32]),
    "b200ps_launch_count": (_u64, []),
    "b200ps_linreg_scratch_bytes": (_sz, [_u64, _i32]),
    "b200ps_linreg_grad": (_i32, [_vp, _vp, _u64, _i32, _u64, _u64, _u64, _u64, _vp, _sz, _vp]),
    "b200ps_mlp_create": (_i32, [_vp, _i32, _i32, _i32, _vp, _pp]),
    "b200ps_mlp_destroy": (_i32, [_vp]),
    "b200ps_mlp_step": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200ps_gemm_bf16": (_i32, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp]),
    "b200ps_dev_alloc": (_i32, [_sz, _pp]),
    "b200ps_dev_free": (_i32, [_vp]),
    "b200ps_ipc_export": (_i32, [_vp, _vp]),
    "b200ps_ipc_open": (_i32, [_vp, _pp]),
    "b200ps_ipc_close": (_i32, [_vp]),
    "b200ps_peer_barrier": (_i32, [_vp, _i32, _i32, _u32, _vp]),
    "b200ps_peer_barrier_error": (_i32, [_vp, _vp]),
    "b200ps_peer_signal": (_i32, [_vp, _u32, _vp]),
    "b200ps_peer_wait": (_i32, [_vp, _u32, _u32, _vp, _vp]),
    "b200ps_sparse_gather": (_i32, [_vp, _vp, C.POINTER(TableGeom), _vp, _u64, _vp]),
    "b200ps_sparse_serve": (_i32, [_vp, _vp, _vp, _i32, _u64, C.POINTER(TableGeom), _vp]),
    "b200ps_sparse_route": (_i32, [_vp, _vp, _u64, C.POINTER(TableGeom), _vp, _vp]),
    "b200ps_sparse_workspace_bytes": (_sz, [_u64, _i32]),
    "b200ps_sparse_sort_segment": (_i32, [_vp, _i32, _u64, C.POINTER(TableGeom), _vp, _sz, _vp]),
    "b200ps_sparse_apply_sgd": (_i32, [_vp, _vp, _i32, _u64, C.POINTER(TableGeom), _f32, _vp, _vp]),
    "b200ps_sparse_apply_adagrad": (_i32, [_vp, _vp, _vp, _i32, _u64, C.POINTER(TableGeom), _f32, _vp, _vp]),
    "b200ps_sparse_push_sgd_noalloc": (_i32, [_vp, _vp, _vp, _i32, _u64, C.POINTER(TableGeom), _f32, _vp]),
    "b200ps_sparse_plan_read": (_i32, [_vp, _u64, _u64, C.POINTER(_u64), C.POINTER(_u64), C.POINTER(_u64), _vp, _vp, _vp,
                                        _vp, _vp, _vp]),
    # context group
    "b200ps_dense_options_default": (_i32, [C.POINTER(DenseOptions)]),
    "b200ps_ctx_create": (_i32, [_vp, _vp, _i32, _i32, C.POINTER(DenseOptions), _pp]),
    "b200ps_ctx_destroy": (_i32, [_vp]),
    "b200ps_ctx_export_handle": (_i32, [_vp, _vp]),
    "b200ps_ctx_import_handles": (_i32, [_vp, _vp]),
    "b200ps_ctx_sync": (_i32, [_vp, _vp]),
    "b200ps_plan_build": (_i32, [_vp, _i32, _i32, _i32, _i32, _pp]),
    "b200ps_plan_destroy": (None, [_vp]),
    "b200ps_plan_params_padded": (_u64, [_vp]),
    "b200ps_plan_two_level": (_i32, [_vp]),
    "b200ps_plan_first_grad_slot": (_i32, [_vp]),
    "b200ps_plan_arena_elems": (_u64, [_vp, _i32]),
    "b200ps_plan_num_local_workers": (_i32, [_vp]),
    "b200ps_plan_local_worker": (_i32, [_vp, _i32]),
    "b200ps_plan_grad_offset": (_i32, [_vp, _i32, C.POINTER(_i32), C.POINTER(_u64)]),
    "b200ps_plan_num_launches": (_i32, [_vp]),
    "b200ps_plan_launch": (_i32, [_vp, _i32, C.POINTER(_i32), C.POINTER(_u64), C.POINTER(_u64), C.POINTER(_i32), _vp, _vp,
                                   C.POINTER(_i32), _vp]),
    "b200ps_ctx_plan": (_vp, [_vp]),
    "b200ps_ctx_step_count": (_u64, [_vp]),
    "b200ps_ctx_algorithmic_bytes_per_step": (_u64, [_vp]),
    "b200ps_dense_init": (_i32, [_vp, _vp, _vp]),
    "b200ps_dense_init_uniform": (_i32, [_vp, _u64, _vp]),
    "b200ps_dense_replica": (_i32, [_vp, _pp]),
    "b200ps_dense_grad_buffer": (_i32, [_vp, _i32, _pp]),
    "b200ps_ctx_arena_base": (_i32, [_vp, _i32, _pp]),
    "b200ps_dense_state": (_i32, [_vp, _i32, _pp, _pp, C.POINTER(_u64)]),
    "b200ps_dense_push": (_i32, [_vp, _i32, _vp, _vp]),
    "b200ps_dense_pull": (_i32, [_vp, _i32, _vp, _vp]),
    "b200ps_dense_step": (_i32, [_vp, _vp]),
    "b200ps_dense_step_async_ordered": (_i32, [_vp, _vp]),
    "b200ps_dense_step_host": (_i32, [_vp, _vp, _vp, _vp]),
    "b200ps_dense_host_sync": (_i32, [_vp]),
    "b200ps_dense_dump": (_i32, [_vp, _vp, _vp, _vp, _vp]),
    "b200ps_dense_save": (_i32, [_vp, _cp, _vp]),
    "b200ps_dense_load": (_i32, [_vp, _cp, _vp]),
    "b200ps_ctx_set_tuning": (_i32, [_vp, _i32, _i32, _i32]),
    "b200ps_ctx_set_graph": (_i32, [_vp, _i32]),
    "b200ps_ctx_set_profile": (_i32, [_vp, _i32]),
    "b200ps_ctx_phase_ms": (_i32, [_vp, C.POINTER(_f32), C.POINTER(_f32), C.POINTER(_f32)]),
    "b200ps_multicast_supported": (_i32, []),
    # table group
    "b200ps_table_options_default": (_i32, [C.POINTER(TableOptions)]),
    "b200ps_table_layout": (_i32, [_u64, _i32, _i32, _i32, _u64, _i32, _i32, _i32, C.POINTER(_u64), C.POINTER(_u64), _vp, _vp, _vp, _vp,
                                    _vp, _vp]),
    "b200ps_table_create": (_i32, [_vp, _u64, _i32, _u64, _i32, _i32, C.POINTER(TableOptions), _pp]),
    "b200ps_table_destroy": (_i32, [_vp]),
    "b200ps_table_export_handle": (_i32, [_vp, _vp]),
    "b200ps_table_import_handles": (_i32, [_vp, _vp]),
    "b200ps_table_sync": (_i32, [_vp, _vp]),
    "b200ps_table_init_uniform": (_i32, [_vp, _u64, _vp]),
    "b200ps_table_worker_buffers": (_i32, [_vp, _i32, _pp, _pp, _pp]),
    "b200ps_table_shard": (_i32, [_vp, _i32, _pp, _pp, C.POINTER(_u64), C.POINTER(_u64)]),
    "b200ps_sparse_pull": (_i32, [_vp, _vp]),
    "b200ps_sparse_pull_worker": (_i32, [_vp, _i32, _vp, _vp]),
    "b200ps_sparse_push": (_i32, [_vp, _vp]),
    "b200ps_sparse_step": (_i32, [_vp, _vp]),
    "b200ps_table_set_graph": (_i32, [_vp, _i32]),
    "b200ps_table_set_profile": (_i32, [_vp, _i32]),
    "b200ps_table_phase_ms": (_i32, [_vp, C.POINTER(_f32), C.POINTER(_f32), C.POINTER(_f32)]),
    "b200ps_table_plan_read": (_i32, [_vp, C.POINTER(_u64), C.POINTER(_u64), C.POINTER(_u64), _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200ps_table_save": (_i32, [_vp, _cp, _vp]),
    "b200ps_table_load": (_i32, [_vp, _cp, _vp]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = the built library is older than the header
    _fn.restype = _res
    _fn.argtypes = _args


def last_error() -> str:
    return lib.b200ps_last_error().decode("utf-8", "replace")


def check(status: int) -> int:
    """Raise B200PSError for a negative status; pass non-negative values (lengths, counts) through."""
    if status < 0:
        raise B200PSError(status, last_error())
    return status
