"""bench.py's reference arm (the CPU oracle, the only arm that runs without a GPU) prints ONE JSON
line with the contract's keys; the GPU arm's helpers that need no device behave."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, env=None):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=300,
                         env={**os.environ, **(env or {})})
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, out.stdout
    return json.loads(lines[0])


def test_reference_arm_line_dense():
    d = _run("--impl", "reference", "--workload", "c0", "--steps", "3", "--warmup", "1", env={"OMP_NUM_THREADS": "1"})
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["metric"] == "ps_steps_per_sec" and d["unit"] == "worker-steps/s"
    assert d["vs_baseline"] is None and d["data"] == "synthetic" and "workload" in d["config"]
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    # torchrun exports OMP_NUM_THREADS=1; the reference must still use every core it may run on
    assert d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))


def test_reference_arm_other_ranks_stay_silent():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c0", "--steps", "1",
                          "--gpus", "2"], capture_output=True, text=True, timeout=120, env={**os.environ, "RANK": "1", "WORLD_SIZE": "2"})
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_reference_arm_line_sparse():
    d = _run("--impl", "reference", "--workload", "c2", "--rows", "200000", "--steps", "1")
    assert d["impl"] == "reference" and d["scaling"] == "strong" and d["cpu_baseline"]["cores"] >= 1 and d["value"] > 0
