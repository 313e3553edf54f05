"""The host-side C++ of libb200ps (TFJob / TF_CONFIG JSON reader, cluster model, shard map) under
AddressSanitizer + UBSan: tests/native/host_fuzz.cpp feeds it well-formed inputs, malformed JSON
shapes and a few thousand random mutations.  These files take untrusted text (a TFJob object, an
environment variable), so they are the part of the library worth fuzzing; no GPU is involved."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "tensorflow-k8s_b200", "csrc")
GXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")


@pytest.mark.skipif(GXX is None, reason="no g++")
def test_contract_and_layout_under_asan_ubsan(tmp_path):
    exe = str(tmp_path / "host_fuzz")
    cmd = [GXX, "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined",
           "-I", os.path.join(ROOT, "include"), "-I", CSRC, os.path.join(ROOT, "tests", "native", "host_fuzz.cpp"),
           os.path.join(CSRC, "contract.cpp"), os.path.join(CSRC, "layout.cpp"), os.path.join(CSRC, "common.cpp"), "-o", exe]
    build = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    if build.returncode != 0 and "sanitize" in build.stderr and "cannot find" in build.stderr:
        pytest.skip("sanitizer runtimes are not installed")
    assert build.returncode == 0, build.stderr[-3000:]
    env = {**os.environ, "ASAN_OPTIONS": "detect_leaks=1:abort_on_error=0", "UBSAN_OPTIONS": "print_stacktrace=1"}
    env.pop("LD_PRELOAD", None)
    run = subprocess.run([exe], capture_output=True, text=True, timeout=300, env=env)
    assert run.returncode == 0 and "host_fuzz: 0 problems" in run.stdout, (run.stdout + run.stderr)[-3000:]
