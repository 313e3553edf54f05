"""Synthetic workloads of BASELINE.json `configs` (shapes only -- no datasets, no checkpoints)."""
from __future__ import annotations

from typing import List

SEED_PARAMS, SEED_GRADS, SEED_IDS = 1, 2, 3


def mlp_c1_variables() -> List[int]:
    """C1: dense MLP 1024 -> 8192 -> 8192 -> 2048 -> 4096 (weights and biases): 100 685 824 params."""
    dims = [1024, 8192, 8192, 2048, 4096]
    out: List[int] = []
    for a, b in zip(dims[:-1], dims[1:]):
        out += [a * b, b]
    return out


def linreg_c0_variables() -> List[int]:
    """C0: linear regression, 1 M fp32 parameters in one variable."""
    return [1_000_000]


def resnet50_c3_variables() -> List[int]:
    """C3: the 161 trainable variables of ResNet-50 (torchvision layout), 25 557 032 params."""
    out: List[int] = [64 * 3 * 7 * 7, 64, 64]  # conv1, bn1 (weight, bias)
    inplanes = 64
    for planes, blocks in ((64, 3), (128, 4), (256, 6), (512, 3)):
        for b in range(blocks):
            out += [planes * inplanes, planes, planes]          # conv1 1x1, bn1
            out += [planes * planes * 9, planes, planes]        # conv2 3x3, bn2
            out += [planes * 4 * planes, planes * 4, planes * 4]  # conv3 1x1, bn3
            if b == 0:
                out += [planes * 4 * inplanes, planes * 4, planes * 4]  # downsample conv + bn
            inplanes = planes * 4
    out += [1000 * 2048, 1000]  # fc
    return out


def sweep_c4_sizes() -> List[int]:
    """C4: 1 KB ... 1 GB in powers of 4 (fp32 element counts)."""
    return [256 * 4 ** k for k in range(11)]


def make_ids(kind: str, rows: int, workers: int, per_worker: int, step: int = 1, seed: int = SEED_IDS):
    """Embedding ids int64 [workers, per_worker]: 'uniform' over the rows, or
    'zipf' = Zipf(1.05) by the inverse CDF of the continuous power law.  Driven by the same
    splitmix64 counter hash as the fp32 generator, so the arrays are reproducible everywhere;
    they are INPUTS (fed identically to the GPU path and to the checker)."""
    import numpy as np

    u64 = np.uint64
    idx = np.arange(workers * per_worker, dtype=np.uint64)
    key = ((seed * 0xD6E8FEB86659FD93) ^ (step * 0xA24BAED4963EE407)) & ((1 << 64) - 1)
    with np.errstate(over="ignore"):
        z = (idx ^ u64(key)) + u64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> u64(30))) * u64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> u64(27))) * u64(0x94D049BB133111EB)
        z = z ^ (z >> u64(31))
    if kind == "uniform":
        ids = (z % u64(rows)).astype(np.int64)
    elif kind == "zipf":
        alpha = 1.05
        u = (z >> u64(11)).astype(np.float64) * 2.0 ** -53
        x = ((float(rows) ** (1.0 - alpha) - 1.0) * u + 1.0) ** (1.0 / (1.0 - alpha))
        ids = np.clip(np.floor(x).astype(np.int64) - 1, 0, rows - 1)
    else:
        raise ValueError("kind must be 'uniform' or 'zipf'")
    return ids.reshape(workers, per_worker)
