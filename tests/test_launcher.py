"""Single-box launcher: TFJob -> task placement -> per-rank environment (TF_CONFIG et al.), and
the restart policies of k8s-operator.md:47-52 (OnFailure restarts in place, Never fails the job)."""
import json
import os
import sys

import b200ps
from b200ps import launcher

C1 = {"metadata": {"name": "mlp"}, "spec": {"runtimeId": "ab12", "replicaSpecs": [
    {"tfReplicaType": "PS", "replicas": 2}, {"tfReplicaType": "WORKER", "replicas": 4}]}}


def test_placement_and_env():
    c = b200ps.Cluster.from_tfjob(C1)
    by_gpu = launcher.plan_tasks(c, 4)
    names = {g: [f"{t['type']}:{t['index']}" for t in ts] for g, ts in by_gpu.items()}
    assert names == {0: ["ps:0", "worker:0"], 1: ["ps:1", "worker:1"], 2: ["worker:2"], 3: ["worker:3"]}
    env = launcher.rank_env(by_gpu, 1, 4, 29500)
    assert env["RANK"] == "1" and env["WORLD_SIZE"] == "4" and env["MASTER_ADDR"] == "127.0.0.1"
    assert env["TF_CONFIG"] == c.tf_config_for("ps", 1)
    tasks = json.loads(env["B200PS_TASKS"])
    assert [t["type"] for t in tasks] == ["ps", "worker"] and json.loads(tasks[1]["tf_config"])["task"] == {"type": "worker", "index": 1}
    # one GPU hosts everything
    assert len(launcher.plan_tasks(c, 1)[0]) == 6


def test_dry_run(tmp_path, capsys):
    f = tmp_path / "job.json"
    f.write_text(json.dumps(C1))
    assert launcher.main(["--tfjob", str(f), "--gpus", "2", "--dry-run"]) == 0
    out = json.loads(capsys.readouterr().out)
    assert out["shards"] == 2 and out["workers"] == 4 and out["placement"]["1"] == ["ps:1", "worker:1", "worker:3"]


def test_launch_runs_one_process_per_gpu_with_its_tf_config(tmp_path):
    c = b200ps.Cluster.from_tfjob(C1)
    prog = ("import os, json; t = json.loads(os.environ['TF_CONFIG'])['task']; "
            f"open(os.path.join({str(tmp_path)!r}, 'rank' + os.environ['RANK']), 'w').write(t['type'] + str(t['index']))")
    assert launcher.launch(c, 2, [sys.executable, "-c", prog]) == 0
    assert (tmp_path / "rank0").read_text() == "ps0" and (tmp_path / "rank1").read_text() == "ps1"


def test_restart_policies(tmp_path):
    c = b200ps.Cluster.from_tfjob(C1)
    marker = tmp_path / "attempt"
    # rank 0 fails on its first attempt only
    prog = (f"import os, sys; p = {str(marker)!r}\n"
            "if os.environ['RANK'] == '0' and not os.path.exists(p):\n"
            "    open(p, 'w').write('x'); sys.exit(3)\n")
    assert launcher.launch(c, 2, [sys.executable, "-c", prog], restart_policy="OnFailure") == 0
    os.remove(marker)
    assert launcher.launch(c, 2, [sys.executable, "-c", prog], restart_policy="Never") == 3
