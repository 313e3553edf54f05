"""b200ps -- the parameter-server step of a TFJob (push / aggregate / apply / pull) on B200.

Importable as `b200ps` (repo-root shim) because the directory name required by the build
contract, `tensorflow-k8s_b200`, is not a Python identifier.
"""
from ._lib import B200PSError, AdamScalars, LIB_PATH  # noqa: F401  (raises if the library is not built)
from .contract import Cluster, placement_gpu  # noqa: F401
from .layout import ShardMap, row_owner, row_range  # noqa: F401
from . import synth, kernels, ps, sparse_ps, launcher, rank, worker  # noqa: F401  (eager: the b200ps shim aliases loaded submodules)
from .ps import ParameterServer, build_plan  # noqa: F401
from .sparse_ps import EmbeddingTable, TableLayout  # noqa: F401
from .worker import MlpWorker  # noqa: F401

__all__ = ["B200PSError", "AdamScalars", "Cluster", "ShardMap", "placement_gpu", "row_owner", "row_range", "synth",
           "LIB_PATH", "ParameterServer", "build_plan", "EmbeddingTable", "TableLayout"]
