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
