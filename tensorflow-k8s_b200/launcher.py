"""Single-box launcher: a TFJob in, one process per GPU out.

Stands where the operator's reconcile step stood: upstream, `pkg/trainer/replicas.go` (named in
images/tf3.PNG) creates one Job + Service per replica and injects the `TF_CONFIG` environment
variable into its container; the roles it wires together are those of k8s-operator.md:6 (PS tasks
serve parameters, WORKER tasks compute).  Here every task of the TFJob is mapped onto a GPU of the
box (PS shard s -> GPU s mod G, compute task w -> GPU w mod G, `b200ps_placement_gpu`), one
process is started per GPU, and each process receives:

    TF_CONFIG            the exact string of its first task (what a TensorFlow program would read)
    B200PS_TASKS         JSON list of every task placed on this GPU: [{"type", "index", "tf_config"}]
    RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT   for torch.distributed

Restart policy (k8s-operator.md:47-52): "OnFailure" restarts the job (the whole gang, with a fresh
rendezvous; up to --max-restarts), "Never" fails the job with the first failing rank's exit code.
--completion chief: the chief's rank decides completion.  --cleanup: a finalizer-style hook
(k8s-operator.md:35-43) run once when the job ends.

    python -m b200ps.launcher --tfjob job.yaml --gpus 8 -- python train.py
    python -m b200ps.launcher --tfjob job.yaml --gpus 2 --dry-run        # print the plan only
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import subprocess
import sys
import time
from typing import Dict, List

from .contract import Cluster, placement_gpu


def plan_tasks(cluster: Cluster, ngpus: int) -> Dict[int, List[dict]]:
    """GPU -> tasks placed on it, each with its exact TF_CONFIG."""
    by_gpu: Dict[int, List[dict]] = {g: [] for g in range(ngpus)}
    for job, idx in cluster.tasks():
        rank = idx if job == "ps" else cluster.worker_rank(job, idx)
        gpu = placement_gpu(rank, ngpus)
        by_gpu[gpu].append({"type": job, "index": idx, "tf_config": cluster.tf_config_for(job, idx)})
    return by_gpu


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def rank_env(by_gpu: Dict[int, List[dict]], gpu: int, ngpus: int, port: int) -> Dict[str, str]:
    tasks = by_gpu[gpu]
    env = {"RANK": str(gpu), "LOCAL_RANK": str(gpu), "WORLD_SIZE": str(ngpus), "MASTER_ADDR": "127.0.0.1",
           "MASTER_PORT": str(port), "B200PS_TASKS": json.dumps(tasks)}
    if tasks:
        env["TF_CONFIG"] = tasks[0]["tf_config"]
    return env


def chief_rank(cluster: Cluster, ngpus: int) -> int:
    """The rank whose exit decides the job (upstream: the MASTER / chief replica's exit code is the
    TFJob's; PS replicas never exit on their own [RECALL]).  Compute task 0 is the master / chief when
    the job has one, else worker 0 -- both are placed on GPU 0."""
    return placement_gpu(0, ngpus)


def _terminate(procs, grace: float = 5.0):
    for q in procs:
        if q.poll() is None:
            q.terminate()
    for q in procs:
        try:
            q.wait(timeout=grace)
        except subprocess.TimeoutExpired:
            q.kill()
            q.wait()


def launch(cluster: Cluster, ngpus: int, command: List[str], restart_policy: str = "Never", max_restarts: int = 3,
           completion: str = "all", cleanup: List[str] = None, chief_grace: float = 10.0) -> int:
    """Starts one process per GPU and supervises the gang.

    restart_policy (k8s-operator.md:47-52): "Never" -- the first rank that fails fails the job; the
    others are stopped and the failing rank's exit code is returned.  "OnFailure" -- the job is
    restarted, up to `max_restarts` times.  A k8s Job restarts one container in place; here the
    ranks of a job share NCCL communicators and CUDA-IPC mappings that a lone re-executed rank cannot
    rejoin, so the whole gang is stopped and started again with a fresh rendezvous port and
    B200PS_RESTART_COUNT set (the rank program resumes from its last checkpoint, b200ps.rank does).
    completion: "all" -- the job succeeds when every rank exits 0.  "chief" -- when the chief's rank
    (GPU 0) exits 0 the job is complete; ranks still running after `chief_grace` seconds are stopped.
    cleanup (the finalizer of k8s-operator.md:35-43: a hook that runs before the object is gone) --
    a command run once when the job ends, whatever the outcome, with B200PS_JOB_STATUS=succeeded|failed
    and B200PS_JOB_EXIT_CODE in its environment; the launcher returns only after it finished."""
    if restart_policy not in ("Never", "OnFailure"):
        raise ValueError("restartPolicy must be 'Never' or 'OnFailure'")
    if completion not in ("all", "chief"):
        raise ValueError("completion must be 'all' or 'chief'")
    by_gpu = plan_tasks(cluster, ngpus)
    chief = chief_rank(cluster, ngpus)
    attempt, code = 0, 0
    while True:
        port = _free_port()
        extra = {"B200PS_RESTART_COUNT": str(attempt)}
        procs = {g: subprocess.Popen(command, env={**os.environ, **rank_env(by_gpu, g, ngpus, port), **extra}) for g in range(ngpus)}
        failed_rank, code = None, 0
        chief_done_at = None
        while procs:
            for g, p in list(procs.items()):
                try:
                    rc = p.wait(timeout=0.1)
                except subprocess.TimeoutExpired:
                    continue
                del procs[g]
                if rc != 0:  # the FIRST failing rank decides the code; peers we stop ourselves exit -15
                    failed_rank, code = g, rc
                    break
                if g == chief and completion == "chief":
                    chief_done_at = time.monotonic()
            if failed_rank is not None:
                _terminate(list(procs.values()))
                procs = {}
            elif chief_done_at is not None and procs and time.monotonic() - chief_done_at > chief_grace:
                print(f"[launcher] chief (rank {chief}) completed; stopping {sorted(procs)}", file=sys.stderr)
                _terminate(list(procs.values()))
                procs = {}
        if failed_rank is None:
            break
        if restart_policy == "OnFailure" and attempt < max_restarts:
            attempt += 1
            print(f"[launcher] rank {failed_rank} exited {code}; restarting the job, attempt {attempt}/{max_restarts} (OnFailure)",
                  file=sys.stderr)
            continue
        print(f"[launcher] rank {failed_rank} exited {code}; the job has failed ({restart_policy})", file=sys.stderr)
        break
    if cleanup:
        env = {**os.environ, "B200PS_JOB_STATUS": "succeeded" if code == 0 else "failed", "B200PS_JOB_EXIT_CODE": str(code),
               "B200PS_RESTART_COUNT": str(attempt)}
        rc = subprocess.call(cleanup, env=env)
        if rc != 0:
            print(f"[launcher] cleanup hook exited {rc}", file=sys.stderr)
            code = code or rc
    return code


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="b200ps.launcher", description=__doc__.split("\n\n")[0])
    ap.add_argument("--tfjob", required=True, help="TFJob object as YAML or JSON (file path, or '-' for stdin)")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--restart-policy", default="Never", choices=["Never", "OnFailure"])
    ap.add_argument("--max-restarts", type=int, default=3)
    ap.add_argument("--completion", default="all", choices=["all", "chief"],
                    help="all: every rank must exit 0; chief: the job is complete when the chief's rank (GPU 0) exits 0")
    ap.add_argument("--cleanup", default=None, help="finalizer-style hook: a shell command run once when the job ends")
    ap.add_argument("--dry-run", action="store_true", help="print the placement and every TF_CONFIG, start nothing")
    ap.add_argument("command", nargs=argparse.REMAINDER, help="-- program and arguments run once per GPU")
    args = ap.parse_args(argv)
    text = sys.stdin.read() if args.tfjob == "-" else open(args.tfjob).read()
    cluster = Cluster.from_tfjob(text)
    by_gpu = plan_tasks(cluster, args.gpus)
    if args.dry_run:
        print(json.dumps({"shards": cluster.num_shards, "workers": cluster.num_workers,
                          "placement": {str(g): [f"{t['type']}:{t['index']}" for t in ts] for g, ts in by_gpu.items()},
                          "tf_config": {f"{t['type']}:{t['index']}": t["tf_config"] for ts in by_gpu.values() for t in ts}},
                         indent=1))
        return 0
    cmd = args.command[1:] if args.command[:1] == ["--"] else args.command
    if not cmd:
        ap.error("no command given (use: -- program args...)")
    return launch(cluster, args.gpus, cmd, args.restart_policy, args.max_restarts, args.completion,
                  ["/bin/sh", "-c", args.cleanup] if args.cleanup else None)


if __name__ == "__main__":
    sys.exit(main())
