"""TFJob replica specs in, TF_CONFIG per task out -- host mirror of the contract group of
include/b200ps.h (the C++ in csrc/contract.cpp does the work).

Stands in for what the reference only names in images/tf3.PNG (pkg/apis/tensorflow/v1alpha1/
types.go, defaults.go, validation/validation.go, pkg/trainer/replicas.go); the cluster / job /
task role model is k8s-operator.md:6.
"""
from __future__ import annotations

import ctypes as C
import json
from typing import Iterator, Tuple

from . import _lib
from ._lib import check, lib

JOB_ORDER = ("chief", "master", "ps", "worker")  # sorted, as Go marshals a map


class Cluster:
    """A cluster spec: job name -> task addresses."""

    def __init__(self, handle: int):
        self._h = C.c_void_p(handle)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib.b200ps_cluster_destroy(h)

    # ---- constructors -------------------------------------------------------------------------
    @classmethod
    def from_tfjob(cls, tfjob) -> "Cluster":
        """`tfjob`: a dict, a JSON string or a YAML string holding one TFJob object."""
        if isinstance(tfjob, dict):
            text = json.dumps(tfjob)
        else:
            text = str(tfjob)
            if not text.lstrip().startswith(("{", "[")):
                import yaml  # the operator's API server hands Go JSON; YAML is converted first

                text = json.dumps(yaml.safe_load(text))
        out = C.c_void_p()
        check(lib.b200ps_cluster_from_tfjob(text.encode(), C.byref(out)))
        return cls(out.value)

    @classmethod
    def from_tf_config(cls, tf_config: str) -> Tuple["Cluster", str, int]:
        """Returns (cluster, task_type, task_index) of one TF_CONFIG value."""
        out, buf, idx = C.c_void_p(), C.create_string_buffer(64), C.c_int(-1)
        check(lib.b200ps_cluster_from_tf_config(tf_config.encode(), C.byref(out), buf, len(buf), C.byref(idx)))
        return cls(out.value), buf.value.decode(), idx.value

    # ---- accessors ----------------------------------------------------------------------------
    def num_tasks(self, job: str) -> int:
        return check(lib.b200ps_cluster_num_tasks(self._h, job.encode()))

    def task_address(self, job: str, index: int) -> str:
        buf = C.create_string_buffer(512)
        check(lib.b200ps_cluster_task_address(self._h, job.encode(), index, buf, len(buf)))
        return buf.value.decode()

    def tf_config_for(self, job: str, index: int) -> str:
        cap = 4096
        while True:
            buf = C.create_string_buffer(cap)
            rc = lib.b200ps_tf_config_for(self._h, job.encode(), index, buf, cap)
            if rc == _lib.ERR_RANGE and cap < (1 << 24) and "too small" in _lib.last_error():
                cap *= 4
                continue
            check(rc)
            return buf.value.decode()

    @property
    def num_shards(self) -> int:
        return check(lib.b200ps_cluster_num_shards(self._h))

    @property
    def num_workers(self) -> int:
        return check(lib.b200ps_cluster_num_workers(self._h))

    def worker_rank(self, job: str, index: int) -> int:
        return check(lib.b200ps_cluster_worker_rank(self._h, job.encode(), index))

    def tasks(self) -> Iterator[Tuple[str, int]]:
        for job in JOB_ORDER:
            for i in range(self.num_tasks(job)):
                yield job, i

    def as_dict(self) -> dict:
        return {j: [self.task_address(j, i) for i in range(self.num_tasks(j))] for j in JOB_ORDER if self.num_tasks(j)}


def placement_gpu(task_rank: int, ngpus: int) -> int:
    """ps shard s -> GPU s mod G; compute task w -> GPU w mod G."""
    return check(lib.b200ps_placement_gpu(task_rank, ngpus))
