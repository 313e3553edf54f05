"""The C-ABI library loads and exports every symbol include/b200ps.h declares; the product never
touches the oracle.  No compute calls here (no GPU)."""
import ctypes
import os
import re
import subprocess

import b200ps

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200ps.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200ps_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    names = declared_functions()
    assert len(names) >= 25
    lib = ctypes.CDLL(b200ps.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in b200ps.h but not exported by libb200ps.so"
        assert n in b200ps._lib.SIGNATURES, f"{n} has no ctypes signature in _lib.py"
    assert sorted(b200ps._lib.SIGNATURES) == names


def test_only_the_abi_is_exported():
    out = subprocess.check_output(["nm", "-D", "--defined-only", b200ps.LIB_PATH], text=True)
    syms = [l.split()[-1] for l in out.splitlines() if " T " in l]
    assert syms and all(s.startswith("b200ps_") for s in syms)


def test_header_is_plain_c():
    src = '#include "b200ps.h"\nint main(void){return b200ps_version() < 0;}\n'
    subprocess.run(["/usr/bin/gcc", "-std=c99", "-pedantic", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"),
                    "-x", "c", "-"], input=src, text=True, check=True)


def test_product_never_uses_the_oracle():
    pkg = os.path.join(ROOT, "tensorflow-k8s_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                # comments may cite the oracle; code may not include, import, link or dlopen it
                assert not re.search(r'#\s*include\s*[<"][^>"]*oracle', text), f
                assert not re.search(r"^\s*(from|import)\s+[\w.]*oracle", text, flags=re.M), f
                assert "libps_oracle" not in text and "dlopen" not in text, f
    out = subprocess.check_output(["ldd", b200ps.LIB_PATH], text=True)
    assert "oracle" not in out


def test_errors_are_reported_not_fatal():
    lib = b200ps._lib.lib
    assert lib.b200ps_version() >= 1
    rc = lib.b200ps_set_dense_tuning(100, 0, 0)  # not a multiple of 32
    assert rc == b200ps._lib.ERR_RANGE and "multiple of 32" in b200ps._lib.last_error()
    a = b200ps.AdamScalars()
    assert lib.b200ps_adam_scalars(1e-3, 0.9, 0.999, 1e-8, 0, ctypes.byref(a)) == b200ps._lib.ERR_RANGE
    assert lib.b200ps_adam_scalars(1e-3, 0.9, 0.999, 1e-8, 1, ctypes.byref(a)) == 0
    assert abs(a.lr_t - 1e-3 * (1 - 0.999) ** 0.5 / (1 - 0.9)) < 1e-9


def test_adam_scalars_match_oracle(oracle):
    for t in (1, 2, 10, 1000, 100000):
        a = b200ps.kernels.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t)
        b = oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t)
        assert [getattr(a, k) for k, _ in a._fields_] == [getattr(b, k) for k, _ in b._fields_]
