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

Restart policy (k8s-operator.md:47-52): "OnFailure" re-executes a failed rank in place (up to
--max-restarts), "Never" fails the job and leaves the other ranks' exit codes to be inspected.

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


def launch(cluster: Cluster, ngpus: int, command: List[str], restart_policy: str = "Never", max_restarts: int = 3) -> int:
    if restart_policy not in ("Never", "OnFailure"):
        raise ValueError("restartPolicy must be 'Never' or 'OnFailure'")
    by_gpu, port = plan_tasks(cluster, ngpus), _free_port()

    def start(g):
        return subprocess.Popen(command, env={**os.environ, **rank_env(by_gpu, g, ngpus, port)})

    procs = {g: start(g) for g in range(ngpus)}
    restarts = {g: 0 for g in range(ngpus)}
    codes: Dict[int, int] = {}
    while procs:
        for g, p in list(procs.items()):
            try:
                rc = p.wait(timeout=0.2)
            except subprocess.TimeoutExpired:
                continue
            if rc != 0 and restart_policy == "OnFailure" and restarts[g] < max_restarts:
                restarts[g] += 1
                print(f"[launcher] rank {g} exited {rc}; restart {restarts[g]}/{max_restarts} (OnFailure)", file=sys.stderr)
                procs[g] = start(g)
                continue
            codes[g] = rc
            del procs[g]
            if rc != 0:  # Never (or restarts exhausted): the job has failed; stop the others
                for q in procs.values():
                    q.terminate()
    return max(codes.values(), key=abs) if codes else 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="b200ps.launcher", description=__doc__.split("\n\n")[0])
    ap.add_argument("--tfjob", required=True, help="TFJob object as YAML or JSON (file path, or '-' for stdin)")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--restart-policy", default="Never", choices=["Never", "OnFailure"])
    ap.add_argument("--max-restarts", type=int, default=3)
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
    return launch(cluster, args.gpus, cmd, args.restart_policy, args.max_restarts)


if __name__ == "__main__":
    sys.exit(main())
