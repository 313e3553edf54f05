"""A C99 program compiled with gcc and linked against libb200ps.so drives contract -> shard map ->
context -> two Adam steps -> dump: a real through-the-C-ABI caller, what a cgo binding would be.
The program is compiled in both cases; it runs (and is compared with the oracle) on a GPU."""
import os
import subprocess

import numpy as np
import pytest

import b200ps

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "native", "abi_c_caller.c")


def _build(tmp_path):
    exe = str(tmp_path / "abi_c_caller")
    libdir = os.path.dirname(b200ps.LIB_PATH)
    subprocess.run(["/usr/bin/gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"), SRC,
                    "-o", exe, "-L", libdir, "-lb200ps", f"-Wl,-rpath,{libdir}"], check=True)
    return exe


def test_c_caller_compiles_and_links(tmp_path):
    exe = _build(tmp_path)
    out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    assert "libb200ps.so" in out and "oracle" not in out and "torch" not in out


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["direct", "graph"])
def test_c_caller_matches_the_oracle(oracle, tmp_path, mode):
    exe = _build(tmp_path)
    outf = str(tmp_path / "dump.bin")
    r = subprocess.run([exe, outf] + (["graph"] if mode == "graph" else []), capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert '"task":{"type":"worker","index":1}' in r.stdout
    raw = np.fromfile(outf, np.uint8)
    P, t = np.frombuffer(raw[:16].tobytes(), np.uint64)
    P = int(P)
    assert t == 2 and P == b200ps.ShardMap([70000, 300, 33000], 2).total
    p, m, v = np.frombuffer(raw[16:].tobytes(), np.float32).reshape(3, P)
    hp, hm, hv = oracle.fill_uniform(P, 1, 0, 0), np.zeros(P, np.float32), np.zeros(P, np.float32)
    for step in (1, 2):
        hg = [oracle.fill_uniform(P, 2, step, w) for w in range(3)]
        oracle.dense_adam(hp, hm, hv, hg, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, step))
    for got, want in ((p, hp), (m, hm), (v, hv)):
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
