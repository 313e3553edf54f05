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


def test_onfailure_gives_up_after_max_restarts_and_returns_the_failing_code(tmp_path):
    """Rank 0 always fails: the gang is restarted max_restarts times (never a lone rank, never after
    the job has failed for good) and the launcher returns rank 0's code, not a stopped peer's -15."""
    c = b200ps.Cluster.from_tfjob(C1)
    log = tmp_path / "starts"
    prog = (f"import os, sys, time; open({str(log)!r}, 'a').write(os.environ['RANK'] + ':' + os.environ['B200PS_RESTART_COUNT'] + ' ')\n"
            "if os.environ['RANK'] == '0': sys.exit(7)\n"
            "time.sleep(30)\n")
    assert launcher.launch(c, 2, [sys.executable, "-c", prog], restart_policy="OnFailure", max_restarts=2) == 7
    starts = sorted(log.read_text().split())
    assert starts == ["0:0", "0:1", "0:2", "1:0", "1:1", "1:2"]  # three gang starts, each with its restart count


def test_chief_decides_completion_and_cleanup_hook_runs(tmp_path):
    c = b200ps.Cluster.from_tfjob(C1)
    done = tmp_path / "cleaned"
    # rank 1 (a pure PS / worker rank) never exits on its own; the chief's rank finishes
    prog = "import os, time\nif os.environ['RANK'] != '0': time.sleep(60)\n"
    hook = [sys.executable, "-c", f"import os; open({str(done)!r}, 'w').write(os.environ['B200PS_JOB_STATUS'])"]
    assert launcher.launch(c, 2, [sys.executable, "-c", prog], completion="chief", cleanup=hook, chief_grace=0.5) == 0
    assert done.read_text() == "succeeded"
    bad = "import sys; sys.exit(4)"
    assert launcher.launch(c, 2, [sys.executable, "-c", bad], cleanup=hook) == 4
    assert done.read_text() == "failed"
