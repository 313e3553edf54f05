"""The program one GPU rank of a TFJob runs: everything it needs comes from the environment the
launcher (or an operator) prepared -- `TF_CONFIG`, `B200PS_TASKS`, RANK / WORLD_SIZE / MASTER_*.

    python -m b200ps.launcher --tfjob job.yaml --gpus 8 -- python -m b200ps.rank --workload c1 --steps 100

Upstream, `pkg/trainer/replicas.go` (named in images/tf3.PNG) injects `TF_CONFIG` into every replica's
container and the TensorFlow program inside derives its role from it: PS tasks serve parameters,
WORKER tasks compute (k8s-operator.md:6).  Here the tasks of the job that were placed on this GPU are
listed in `B200PS_TASKS` (each with its own exact TF_CONFIG string); the cluster spec inside any of
them gives S = #ps tasks and W = #compute tasks, `b200ps_placement_gpu` says which of them live here,
and one `ParameterServer` context serves them all.  `job_from_env` is that derivation; `main` drives
synthetic PS steps through it (gradients u(seed 2, step, worker, i)), with
checkpoints so that a job restarted by the launcher (restartPolicy OnFailure, k8s-operator.md:47-52)
resumes from the last saved step.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
from dataclasses import dataclass
from typing import Dict, List, Mapping, Tuple

from . import synth
from .contract import Cluster, placement_gpu


@dataclass
class Job:
    cluster: Cluster
    shards: int            # S = #ps tasks
    workers: int           # W = #compute tasks (master / chief first, then workers)
    rank: int
    world: int
    tasks: List[dict]      # the tasks placed on this GPU: {"type", "index", "tf_config"}
    local_shards: List[int]
    local_workers: List[int]
    is_chief: bool         # compute task 0 (master / chief, else worker 0) lives here


def job_from_env(env: Mapping[str, str] = os.environ) -> Job:
    """Derives this rank's part of the job from TF_CONFIG / B200PS_TASKS and checks it against the
    placement rule.  Raises ValueError when the environment is inconsistent."""
    rank, world = int(env.get("RANK", "0")), int(env.get("WORLD_SIZE", "1"))
    if "B200PS_TASKS" in env:
        tasks = json.loads(env["B200PS_TASKS"])
    elif "TF_CONFIG" in env:  # a lone TF_CONFIG (what an operator gives one container): one task
        t = json.loads(env["TF_CONFIG"])["task"]
        tasks = [{"type": t["type"], "index": int(t["index"]), "tf_config": env["TF_CONFIG"]}]
    else:
        raise ValueError("neither B200PS_TASKS nor TF_CONFIG is set: start this program through b200ps.launcher")
    tf_config = tasks[0]["tf_config"] if tasks else env.get("TF_CONFIG")
    if not tf_config:
        raise ValueError(f"rank {rank} hosts no task and has no TF_CONFIG to read the cluster from")
    cluster, ttype, tindex = Cluster.from_tf_config(tf_config)
    if tasks and (ttype, tindex) != (tasks[0]["type"], int(tasks[0]["index"])):
        raise ValueError("TF_CONFIG's task differs from the first entry of B200PS_TASKS")
    S, W = cluster.num_shards, cluster.num_workers
    want_shards = [s for s in range(S) if placement_gpu(s, world) == rank]
    want_workers = [w for w in range(W) if placement_gpu(w, world) == rank]
    got_shards = sorted(int(t["index"]) for t in tasks if t["type"] == "ps")
    got_workers = sorted(cluster.worker_rank(t["type"], int(t["index"])) for t in tasks if t["type"] != "ps")
    if "B200PS_TASKS" in env and (got_shards, got_workers) != (want_shards, want_workers):
        raise ValueError(f"rank {rank} of {world} was handed ps {got_shards} / compute tasks {got_workers}, "
                         f"but placement (task mod GPUs) puts ps {want_shards} / compute tasks {want_workers} here")
    for t in tasks[1:]:  # every task of the job sees the same cluster
        c2, _, _ = Cluster.from_tf_config(t["tf_config"])
        if c2.as_dict() != cluster.as_dict():
            raise ValueError("the tasks placed on this GPU disagree about the cluster spec")
    return Job(cluster, S, W, rank, world, tasks, want_shards, want_workers, 0 in want_workers)


def workload_variables(name: str) -> List[int]:
    if name == "c1":
        return synth.mlp_c1_variables()
    if name == "c3":
        return synth.resnet50_c3_variables()
    if name == "c0":
        return synth.linreg_c0_variables()
    return [int(x) for x in name.split(",")]  # explicit sizes: "70000,300,33000"


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="b200ps.rank", description="one GPU rank of a TF_CONFIG-driven PS job")
    ap.add_argument("--workload", default="c1", help="c0 | c1 | c3 | comma-separated variable sizes")
    ap.add_argument("--optimizer", default="adam", choices=["sgd", "adam", "momentum", "ftrl"])
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--aggregate", default="two_level", choices=["flat", "two_level", "multimem"])
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--graph", action="store_true", help="replay each step as one CUDA graph")
    ap.add_argument("--checkpoint-dir", default=None, help="save every --save-every steps; resume from it after a restart")
    ap.add_argument("--save-every", type=int, default=0)
    ap.add_argument("--dump", default=None, help="directory: write this rank's final replica (replica-rank<r>.npy) for checks")
    ap.add_argument("--fail-at-step", type=int, default=0, help="fault injection: exit 17 before this step on the first attempt")
    ap.add_argument("--fail-rank", type=int, default=0)
    args = ap.parse_args(argv)

    import numpy as np
    import torch
    import torch.distributed as dist

    from . import kernels as K
    from .ps import ParameterServer

    job = job_from_env()
    local = int(os.environ.get("LOCAL_RANK", str(job.rank)))
    attempt = int(os.environ.get("B200PS_RESTART_COUNT", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("b200ps.rank needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local % torch.cuda.device_count())
    if job.world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    ps = ParameterServer(workload_variables(args.workload), job.shards, job.workers, optimizer=args.optimizer, lr=args.lr,
                         rank=job.rank, world=job.world, aggregate=args.aggregate, use_graph=args.graph, cluster=job.cluster)
    ps.init_uniform(synth.SEED_PARAMS)
    start = 0
    if args.checkpoint_dir and attempt > 0 and os.path.exists(os.path.join(args.checkpoint_dir, f"ps-rank{job.rank}-of-{job.world}.b200ps")):
        ps.load(args.checkpoint_dir)
        start = ps.t  # one optimizer step per job step
    for t in range(start + 1, args.steps + 1):
        if args.fail_at_step == t and attempt == 0 and job.rank == args.fail_rank:
            os._exit(17)  # a crashed container: no clean-up, peers are left waiting
        for w in ps.plan.local_workers:  # the compute tasks placed here "compute": synthetic gradients
            K.fill_uniform(ps.grad_buffer(w), synth.SEED_GRADS, t, w)
        ps.step()
        if args.checkpoint_dir and args.save_every and t % args.save_every == 0:
            ps.save(args.checkpoint_dir)
    ps.check_barriers()
    replica = ps.replica.cpu().numpy()
    if args.dump:
        os.makedirs(args.dump, exist_ok=True)
        np.save(os.path.join(args.dump, f"replica-rank{job.rank}.npy"), replica)
    print(json.dumps({"rank": job.rank, "world": job.world, "shards": job.shards, "workers": job.workers,
                      "tasks": [f"{t['type']}:{t['index']}" for t in job.tasks], "local_shards": job.local_shards,
                      "local_workers": job.local_workers, "chief": job.is_chief, "steps": ps.t, "resumed_from": start,
                      "params_padded": ps.plan.P, "replica_sha256": hashlib.sha256(replica.tobytes()).hexdigest()}), flush=True)
    ps.close()
    if job.world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
