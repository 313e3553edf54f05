"""The TF_CONFIG-driven rank program (b200ps.rank): what the launcher hands a GPU rank is enough to
build the parameter server -- S, W and the local tasks come from TF_CONFIG / B200PS_TASKS alone
(k8s-operator.md:6 roles; pkg/trainer/replicas.go in images/tf3.PNG owns TF_CONFIG upstream)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import b200ps
from b200ps import launcher, rank as rankmod

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
C1 = {"metadata": {"name": "mlp"}, "spec": {"runtimeId": "ab12", "replicaSpecs": [
    {"tfReplicaType": "PS", "replicas": 2}, {"tfReplicaType": "WORKER", "replicas": 4}]}}
MASTER = {"metadata": {"name": "wd"}, "spec": {"tfReplicaSpecs": {"Chief": {"replicas": 1}, "PS": {"replicas": 3},
                                                                   "Worker": {"replicas": 2}}}}


def test_job_from_env_matches_the_placement():
    c = b200ps.Cluster.from_tfjob(C1)
    for G in (1, 2, 4, 8):
        by_gpu = launcher.plan_tasks(c, G)
        seen_s, seen_w = [], []
        for g in range(G):
            env = launcher.rank_env(by_gpu, g, G, 29500)
            if not by_gpu[g]:  # 2 ps + 4 workers on 8 GPUs: GPUs 4..7 host nothing
                with pytest.raises(ValueError):
                    rankmod.job_from_env(env)
                continue
            job = rankmod.job_from_env(env)
            assert (job.shards, job.workers, job.rank, job.world) == (2, 4, g, G)
            assert job.is_chief == (g == 0)
            seen_s += job.local_shards
            seen_w += job.local_workers
        assert sorted(seen_s) == [0, 1] and sorted(seen_w) == [0, 1, 2, 3]
    # chief + workers: the chief is compute task 0
    c = b200ps.Cluster.from_tfjob(MASTER)
    job = rankmod.job_from_env(launcher.rank_env(launcher.plan_tasks(c, 2), 0, 2, 1))
    assert (job.shards, job.workers) == (3, 3) and job.local_workers == [0, 2] and job.local_shards == [0, 2] and job.is_chief
    # a lone TF_CONFIG (what an operator hands one container) is accepted for one task
    one = rankmod.job_from_env({"TF_CONFIG": c.tf_config_for("worker", 1), "RANK": "0", "WORLD_SIZE": "1"})
    assert (one.shards, one.workers) == (3, 3)
    # a rank that was handed the wrong tasks refuses to run
    env = launcher.rank_env(launcher.plan_tasks(b200ps.Cluster.from_tfjob(C1), 2), 0, 2, 1)
    env["RANK"] = "1"
    with pytest.raises(ValueError, match="placement"):
        rankmod.job_from_env(env)
    with pytest.raises(ValueError):
        rankmod.job_from_env({"RANK": "0"})


def _oracle_replica(oracle, sizes, S, W, steps, world=1, aggregate="two_level"):
    P = b200ps.ShardMap(sizes, S).total
    p, m, v = oracle.fill_uniform(P, 1, 0, 0), np.zeros(P, np.float32), np.zeros(P, np.float32)
    for t in range(1, steps + 1):
        g = [oracle.fill_uniform(P, 2, t, w) for w in range(W)]
        if aggregate == "two_level" and world > 1:
            g = oracle.two_level_partials(g, world)
        oracle.dense_adam(p, m, v, g, oracle.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t), inv_w=np.float32(1.0) / np.float32(W))
    return p


def _launch(tmp_path, gpus, extra, policy="Never"):
    job = tmp_path / "job.json"
    job.write_text(json.dumps(C1))
    env = {**os.environ, "PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")}
    cmd = [sys.executable, "-m", "b200ps.launcher", "--tfjob", str(job), "--gpus", str(gpus), "--restart-policy", policy, "--",
           sys.executable, "-m", "b200ps.rank", "--workload", "70000,300,33000", "--dump", str(tmp_path)] + extra
    return subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=600)


@pytest.mark.gpu
def test_tf_config_driven_job_on_one_gpu(oracle, tmp_path):
    """All six tasks of the C1 job (2 ps + 4 workers) on one GPU, started by the launcher, driven only
    by TF_CONFIG / B200PS_TASKS: 3 Adam steps, replica bit-exact against the oracle."""
    r = _launch(tmp_path, 1, ["--steps", "3", "--graph"])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["tasks"] == ["ps:0", "ps:1", "worker:0", "worker:1", "worker:2", "worker:3"] and line["steps"] == 3
    want = _oracle_replica(oracle, [70000, 300, 33000], 2, 4, 3)
    got = np.load(tmp_path / "replica-rank0.npy")
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_onfailure_restart_resumes_from_the_last_checkpoint(oracle, tmp_path):
    """The rank crashes before step 4 on the first attempt; restartPolicy OnFailure restarts the job,
    which resumes from the checkpoint of step 2 and ends bit-identical to an uninterrupted run."""
    ck = tmp_path / "ckpt"
    r = _launch(tmp_path, 1, ["--steps", "5", "--checkpoint-dir", str(ck), "--save-every", "2", "--fail-at-step", "4"], policy="OnFailure")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "restarting the job" in r.stderr
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["resumed_from"] == 2 and line["steps"] == 5
    want = _oracle_replica(oracle, [70000, 300, 33000], 2, 4, 5)
    assert np.array_equal(np.load(tmp_path / "replica-rank0.npy").view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_tf_config_driven_job_on_every_gpu(oracle, tmp_path):
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    G = 2 if n < 4 else 4
    r = _launch(tmp_path, G, ["--steps", "3"])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    want = _oracle_replica(oracle, [70000, 300, 33000], 2, 4, 3, world=G)
    for g in range(G):
        assert np.array_equal(np.load(tmp_path / f"replica-rank{g}.npy").view(np.uint32), want.view(np.uint32)), g
