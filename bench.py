#!/usr/bin/env python
"""bench.py -- PS steps/sec and aggregate+apply HBM GB/s (BASELINE.json `metric`).

A step = push + aggregate + apply + pull of one synthetic batch of gradients through the dense PS
path.  Workload = BASELINE.json configs[1] (dense MLP, 100 685 824 fp32 params, Adam): at N=1
exactly 2 PS shards + 4 workers on one GPU; at N GPUs the same per-GPU density (2N shards, 4N
workers), i.e. weak scaling, one process per GPU.  The parameter server of a rank is built the way a
launched rank builds it: TFJob spec -> placement -> TF_CONFIG / B200PS_TASKS -> b200ps.rank.job_from_env.

  value   worker-steps/s = W x synchronous PS steps/s, gradients resident in HBM (whole job)
  parity_checked  two steps of the same plan on shrunk variables against the oracle, every rank
  e2e     the training step a user runs: mini-batches H2D from pinned memory, the compute tasks'
          forward/backward on the GPU (b200ps_mlp_step, tcgen05 GEMMs), the fused PS step, loss D2H;
          e2e.host_gradients = b200ps_dense_step_host with W host gradient vectors in, parameters out
  roofline  N=1: the aggregate+apply kernel against the measured HBM copy peak;
            N>1: the busiest GPU's inbound NVLink bytes against the measured peer-copy rate
  also    C3, C0 and C2 (uniform / Zipf) measured briefly in the same process
  cpu_baseline / --impl reference  the CPU oracle (a SUBSTITUTE reference: the upstream repo is
          documentation only) on the box's host cores, all threads; same `config` as this arm

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FALLBACK_HBM_GBS = 6650.0  # B200_PROFILING.md, used only if MEASURED_PEAKS.json is absent


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200ps", choices=["b200ps", "reference"])
    ap.add_argument("--workload", default="c1", choices=["c0", "c1", "c2", "c3"])
    ap.add_argument("--ids", default="zipf", choices=["zipf", "uniform"], help="id distribution of --workload c2")
    ap.add_argument("--rows", type=int, default=10_000_000, help="table rows of --workload c2")
    ap.add_argument("--transport", default="peer", choices=["peer", "nccl"])
    ap.add_argument("--overlap", default="auto", choices=["auto", "on", "off"],
                    help="N > 1: the local fold publishes slice after slice and the fused applies start as their slices arrive")
    ap.add_argument("--aggregate", default="auto", choices=["auto", "flat", "two_level", "multimem"],
                    help="N > 1: how the W gradients are folded.  flat / two_level: fixed order, bit-exact against the oracle.  "
                         "multimem: the per-GPU partial sums are added inside the NVSwitch (1e-6).  auto: two_level at N = 2, "
                         "multimem from N = 4 up when the device supports multicast (measured fastest, profiles/r2_summary.md)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline leg")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-profile", action="store_true", help="N > 1: do not bracket the phases of a step with events")
    ap.add_argument("--no-graph", action="store_true", help="SGD workloads on one GPU: launch every step instead of replaying a CUDA graph")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--graph", default="auto", choices=["auto", "on", "off"], help="replay each rank's step as one CUDA graph")
    ap.add_argument("--no-also", action="store_true", help="skip the brief C3 / C0 / C2 measurements that ride along")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = same as --steps, capped at 30")
    ap.add_argument("--batch", type=int, default=512, help="mini-batch rows per compute task of the C1 worker (e2e leg)")
    return ap.parse_args()


def _synth():
    """tensorflow-k8s_b200/synth.py loaded by path: pure-Python shape lists and the id generator, WITHOUT
    importing the package (which would dlopen libb200ps.so) -- the reference arm stays oracle-only."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("_b200ps_synth", os.path.join(ROOT, "tensorflow-k8s_b200", "synth.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def workload(name: str, n_gpus: int):
    """(label, variable sizes, S, W, optimizer) -- per-GPU density held fixed as N grows."""
    synth = _synth()
    if name == "c1":
        return ("C1 dense MLP 1024-8192-8192-2048-4096, 100685824 fp32 params, Adam, "
                f"{2 * n_gpus} PS shards + {4 * n_gpus} workers", synth.mlp_c1_variables(), 2 * n_gpus, 4 * n_gpus, "adam")
    if name == "c3":
        return (f"C3 ResNet-50-sized set, 25557032 params in 161 variables, Adam, {n_gpus} shards + {n_gpus} workers",
                synth.resnet50_c3_variables(), n_gpus, n_gpus, "adam")
    return (f"C0 linear regression 1M params, SGD, {n_gpus} PS + {2 * n_gpus} workers", synth.linreg_c0_variables(),
            n_gpus, 2 * n_gpus, "sgd")


def tfjob_spec(name: str, S: int, W: int) -> dict:
    """The TFJob the bench drives (v1alpha1 list form): S PS replicas + W WORKER replicas."""
    return {"apiVersion": "tensorflow.org/v1alpha1", "kind": "TFJob", "metadata": {"name": f"bench-{name}"},
            "spec": {"replicaSpecs": [{"replicas": S, "tfReplicaType": "PS"}, {"replicas": W, "tfReplicaType": "WORKER"}]}}


def dense_config(label, sizes, S, W, opt):
    """`config` of the JSON line -- identical in both arms (it names the workload, not the implementation)."""
    P = sum((n + 1023) // 1024 * 1024 for n in sizes)
    step_bytes = (4 * W + (24 if opt == "adam" else 8)) * P
    return {"workload": label, "shards": S, "workers": W, "optimizer": opt, "params_padded": P,
            "l2": (f"inputs exceed L2 (a step streams {step_bytes / 1e9:.2f} GB >> 126 MB), no flush needed") if step_bytes > (512 << 20)
            else (f"working set ({step_bytes / 1e6:.0f} MB per step) fits the 126 MB L2 and is NOT flushed between steps: "
                  "an L2-resident, launch-bound figure, not an HBM one")}


def hbm_peak():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full`
    capture of this kernel (profiles/ncu_summary.json), or None if that shape was not captured."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_summary.json")))[kernel]["traffic_bytes_per_launch"]
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of one GPU, sampled through NVML while a timed region runs.

    Every NVML query stalls its GPU for a moment -- the first queries of a process by far the most -- and with N
    ranks in lock-step behind peer barriers a stall of any GPU is everyone's: sampling each GPU every 5 ms from the
    first timed step on cost an 8-GPU C1 step 4-12 % (profiles/r2_n8_timeline.md).  So the thread is started
    early (`start()` then `wait_ready()`: its first two queries are over before any timed work), `mark()` at the
    start of a timed region drops what was seen so far and asks for one sample AT ONCE (so that even a 20-step
    region holds one), and further samples follow every `interval` seconds."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, device_index: int, interval: float = 0.01, nv=None, handle=None):
        super().__init__(daemon=True)
        self.interval = interval
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt, self._wake, self._ready, self._lock = threading.Event(), threading.Event(), threading.Event(), threading.Lock()
        self._taken = 0
        self.err = None
        if nv is not None:  # tests hand in a stand-in for pynvml
            self.nv, self.h = nv, handle
            return
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            import torch

            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(device_index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # no NVML: report it, do not fake numbers
            self.nv, self.err = None, repr(e)

    def run(self):
        if not self.nv:
            self._ready.set()
            return
        while not self._halt.is_set():
            try:
                mhz = self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)
                bits = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception as e:
                self.err = repr(e)
                self._ready.set()
                return
            with self._lock:
                self.samples.append(mhz)
                for b, name in self.REASONS.items():
                    if bits & b:
                        self.reasons.add(name)
            self._taken += 1
            if self._taken >= 2:
                self._ready.set()
            self._wake.wait(self.interval if self._taken >= 2 else 0.0)
            self._wake.clear()

    def wait_ready(self, timeout: float = 5.0):
        """Returns once the first two queries are done (or NVML is absent)."""
        self._ready.wait(timeout)
        return self

    def mark(self):
        """A timed region starts here: forget what was sampled before it and take one sample now."""
        with self._lock:
            self.samples, self.reasons = [], set()
        self._wake.set()

    def finish(self):
        self._halt.set()
        self._wake.set()
        self.join(timeout=2)
        with self._lock:
            s = sorted(self.samples)
            reasons = sorted(self.reasons)
        out = {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(s)}
        if self.err:
            out["error"] = self.err
        return out


# ------------------------------------------------------------------------------------------------ CPU legs
def cpu_run(label, sizes, S, W, opt, steps, warmup, seconds=None, threads=None):
    """The oracle (oracle/ps_oracle.c, OpenMP, all host threads unless `threads` says otherwise) on
    the same workload.  Returns (worker_steps_per_s, steps_done, ms_per_step, threads, algorithmic GB/s)."""
    import numpy as np

    from oracle import c_oracle as O, np_oracle

    nt = O.use_all_cores()
    fill_threads = nt
    if threads:
        nt = threads
    P, _, slices = np_oracle.shard_map(sizes, S, 1024)
    p = O.fill_uniform(P, 1, 0, 0, threads=fill_threads)
    m, v = np.zeros(P, np.float32), np.zeros(P, np.float32)
    g = [O.fill_uniform(P, 2, 1, w, threads=fill_threads) for w in range(W)]

    def one(t):
        a = O.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t)
        for b, e in slices:
            if opt == "adam":
                O.dense_adam(p[b:e], m[b:e], v[b:e], [x[b:e] for x in g], a, threads=nt)
            else:
                O.dense_sgd(p[b:e], [x[b:e] for x in g], 0.01, threads=nt)

    t = 0
    for _ in range(warmup):
        t += 1
        one(t)
    t0 = time.perf_counter()
    done = 0
    while done < steps:
        t += 1
        one(t)
        done += 1
        if seconds is not None and time.perf_counter() - t0 > seconds:
            break
    dt = time.perf_counter() - t0
    bytes_step = (4 * W + (24 if opt == "adam" else 8)) * P
    return W * done / dt, done, 1e3 * dt / done, nt, bytes_step * done / dt / 1e9


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    label, sizes, S, W, opt = workload(args.workload, args.gpus)
    warm = max(args.warmup, 3)
    val, done, ms, nt, gbs = cpu_run(label, sizes, S, W, opt, args.steps, warm)
    print(json.dumps({
        "impl": "reference", "metric": "ps_steps_per_sec", "value": round(val, 3), "unit": "worker-steps/s",
        "n_gpus": args.gpus, "steps": done, "warmup": warm, "ms_per_step": round(ms, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dense_config(label, sizes, S, W, opt),
        "ps_steps_per_sec": round(val / W, 3), "algorithmic_GBps": round(gbs, 1),
        "cpu_baseline": {"value": round(val, 3), "unit": "worker-steps/s", "cores": nt, "kind": "port",
                         "sample": f"{done} full steps of the workload (every slice, every worker); substitute CPU "
                                   "reference (oracle/ps_oracle.c, OpenMP) -- the upstream repo has no implementation"},
        "e2e": {"value": round(val, 3), "unit": "worker-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------ sparse (C2)
C2 = dict(dim=64, S=8, W=8, B=262_144, lr=0.05)


def ncores():
    from oracle import c_oracle as O

    return O.use_all_cores()


def sparse_cpu(rows, ids_kind, steps, seconds):
    """Oracle on the host: per step W gathers + sort_segment + SGD apply.  Gather and apply run on
    all host threads (OpenMP over ids / unique rows); the stable sort is a serial qsort."""
    import numpy as np

    from oracle import c_oracle as O

    synth = _synth()
    dim, W, B = C2["dim"], C2["W"], C2["B"]
    nt = O.use_all_cores()
    table = O.fill_uniform(rows * dim, 1, 0, 0, threads=nt).reshape(rows, dim)
    grad = O.fill_uniform(W * B * dim, 2, 1, 0, threads=nt).reshape(W * B, dim)
    ids = synth.make_ids(ids_kind, rows, W, B, 1)
    t0, done = time.perf_counter(), 0
    while done < steps:
        for w in range(W):
            O.gather(table, ids[w])
        O.sparse_apply_sgd(table, 0, grad, ids.reshape(-1), np.float32(1.0) / np.float32(W), np.float32(C2["lr"]))
        done += 1
        if time.perf_counter() - t0 > seconds:
            break
    dt = time.perf_counter() - t0
    return W * done / dt, done, 1e3 * dt / done


def sparse_arm(args):
    """--workload c2: Wide&Deep embedding table 10M x 64 over 8 PS shards, 8 workers, 262144 ids per
    worker per step; a step = every worker pulls its rows (a10) + one push (a11 + a12)."""
    rank = int(os.environ.get("RANK", "0"))
    rows, dim, S, W, B = args.rows, C2["dim"], C2["S"], C2["W"], C2["B"]
    label = f"C2 Wide&Deep embedding {rows}x{dim}, {S} PS shards, {W} workers, {B} ids/worker/step, {args.ids} ids, SGD"
    if args.impl != "reference":
        import torch
        import torch.distributed as dist

        world, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        if world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        out = sparse_measure(args, args.ids, args.steps, max(args.warmup, 3), e2e=not args.no_e2e, cpu=not args.no_cpu,
                             graph=not args.no_graph)
        if rank == 0:
            print(json.dumps(out))
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    if args.impl == "reference":
        if rank == 0:
            val, done, ms = sparse_cpu(rows, args.ids, args.steps, 120.0)
            print(json.dumps({"impl": "reference", "metric": "ps_steps_per_sec", "value": round(val, 3), "unit": "worker-steps/s",
                              "n_gpus": args.gpus, "steps": done, "warmup": 0, "ms_per_step": round(ms, 2),
                              "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32+i64",
                              "data": "synthetic", "config": {"workload": label},
                              "cpu_baseline": {"value": round(val, 3), "unit": "worker-steps/s", "cores": ncores(), "kind": "port",
                                               "sample": f"{done} full steps; substitute CPU reference (oracle/ps_oracle.c: "
                                                         "OpenMP gather/apply, serial sort)"},
                              "e2e": {"value": round(val, 3), "unit": "worker-steps/s", "h2d_bytes_per_step": 0,
                                      "d2h_bytes_per_step": 0}}))
        return


def sparse_measure(args, ids_kind, steps, warmup, e2e=False, cpu=False, graph=True):
    """One C2 measurement on the current process group (every rank calls it); returns the JSON object."""
    import torch
    import torch.distributed as dist

    import b200ps
    from b200ps import kernels as K, synth

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = torch.cuda.current_device()
    rows, dim, S, W, B = args.rows, C2["dim"], C2["S"], C2["W"], C2["B"]
    label = f"C2 Wide&Deep embedding {rows}x{dim}, {S} PS shards, {W} workers, {B} ids/worker/step, {ids_kind} ids, SGD"
    tab = b200ps.EmbeddingTable(rows, dim, S, W, B, optimizer="sgd", lr=C2["lr"], rank=rank, world=world)
    tab.init_uniform(1)
    ids = synth.make_ids(ids_kind, rows, W, B, 1)
    outs = {}
    for w in tab.local_workers:
        tab.ids[w].copy_(torch.from_numpy(ids[w]))
        K.fill_uniform(tab.grad_rows[w].view(-1), 2, 1, 0, first=w * B * dim)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def reduce_max(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            tab.step()  # every task pulls its rows (a10), one push (a11 + a12): b200ps_sparse_step
        b.record()
        sync()
        return reduce_max(a.elapsed_time(b)) / n

    stream = torch.cuda.Stream()
    sync()
    with torch.cuda.stream(stream):
        sampler = ClockSampler(local, 0.01 if world <= 2 else 0.03)
        sampler.start()
        sampler.wait_ready()
        for _ in range(warmup):
            tab.step()
        sync()
        plan = tab.plan()
        sampler.mark()
        l0 = K.launch_count()
        launched_ms = timed(steps)
        launches = K.launch_count() - l0
        clocks = sampler.finish()
        tab.profile = True  # a shorter pass with events around serve / route+barrier+sort / apply
        for _ in range(min(steps, 20)):
            tab.step()
        sync()
        ph = tab.phase_ms()
        tab.profile = False
        g_ms, s_ms, p_ms = ph["serve_pull"], ph["route+barrier+sort_segment"], ph["apply"]
        graph_ms = None
        if graph:  # the same step replayed as one CUDA graph per rank (no per-launch host work)
            tab.capture_step()
            for _ in range(3):
                tab.step()
            sync()
            graph_ms = timed(steps)
    ms = min(launched_ms, graph_ms) if graph_ms is not None else launched_ms
    d4 = 4 * dim
    gather_bytes = W * B * 8 + (plan["n_kept"] + len(tab.local_workers) * B) * d4  # ids scanned, rows served (read) + landed (write)
    apply_bytes = plan["n_kept"] * (4 + d4) + plan["n_unique"] * (2 * d4 + 4)  # perm + rows; unique id + row R/W
    peak, peak_src = hbm_peak()
    e2e_d = None
    if e2e:
        # host buffers: ids + gradient rows of every local worker go H2D, the pulled rows come back D2H
        h_ids = {w: torch.from_numpy(ids[w]).pin_memory() for w in tab.local_workers}
        h_gr = {w: torch.empty(B, dim, dtype=torch.float32).pin_memory() for w in tab.local_workers}
        h_out = {w: torch.empty(B, dim, dtype=torch.float32).pin_memory() for w in tab.local_workers}
        for w in tab.local_workers:
            h_gr[w].copy_(tab.grad_rows[w])

        def host_step():
            for w in tab.local_workers:
                tab.ids[w].copy_(h_ids[w], non_blocking=True)
                tab.grad_rows[w].copy_(h_gr[w], non_blocking=True)
            tab.pull_step()
            for w in tab.local_workers:
                h_out[w].copy_(tab.pulled[w], non_blocking=True)
            tab.push_step()

        for _ in range(3):
            host_step()
        sync()
        n_e = min(steps, 30)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n_e):
            host_step()
        b.record()
        sync()
        e_ms = reduce_max(a.elapsed_time(b)) / n_e
        e2e_d = {"value": round(W * 1e3 / e_ms, 3), "unit": "worker-steps/s", "h2d_bytes_per_step": W * B * (8 + d4),
               "d2h_bytes_per_step": W * B * d4, "ms_per_step": round(e_ms, 3), "steps": n_e,
               "api": "b200ps_sparse_pull / b200ps_sparse_push via b200ps.EmbeddingTable: ids + gradient rows H2D, pull -> D2H, push"}
    longest = int(plan["seg_offset"][1:].astype("int64").__sub__(plan["seg_offset"][:-1]).max())
    tab.close()
    del tab
    torch.cuda.empty_cache()
    cpu_d = None
    if world == 1 and cpu:
        val, done, cms = sparse_cpu(rows, ids_kind, 10 ** 6, args.cpu_seconds if cpu is True else cpu)
        cpu_d = {"value": round(val, 3), "unit": "worker-steps/s", "cores": ncores(), "kind": "port",
               "sample": f"{done} full steps ({cms:.0f} ms each); substitute CPU reference (oracle/ps_oracle.c: OpenMP "
                         "gather/apply, serial sort)"}
    return {
            "metric": "ps_steps_per_sec", "value": round(W * 1e3 / ms, 3), "unit": "worker-steps/s", "n_gpus": world,
            "steps": steps, "warmup": warmup, "ms_per_step": round(ms, 4), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32+i64", "data": "synthetic",
            "config": {"workload": label, "cuda_graph": graph_ms is not None and graph_ms <= launched_ms,
                       "ms_per_step_launched": round(launched_ms, 4),
                       "ms_per_step_graph": round(graph_ms, 4) if graph_ms is not None else None, "n_kept": int(plan["n_kept"]), "n_unique": int(plan["n_unique"]),
                       "longest_segment": longest,
                       "l2": "table 2.56 GB and 0.54 GB of gradient rows per step exceed L2; no flush"},
            "ps_steps_per_sec": round(1e3 / ms, 3),
            "phases_ms": {"serve_pull": round(g_ms, 4), "route+barrier+sort_segment": round(s_ms, 4), "apply": round(p_ms, 4),
                          "note": "timed one after the other in a separate pass; in the timed step sort_segment runs on a second "
                                  "stream beside serve + route (it needs the ids only), so the step is shorter than the sum"},
            "roofline": {"bound": "hbm", "achieved": round(gather_bytes / g_ms / 1e6, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(gather_bytes / g_ms / 1e6 / peak, 4), "traffic": ncu_traffic("sparse_serve_kernel<16>"),
                         "kernel": "sparse_serve_kernel<16>",
                         "peak_source": peak_src, "algorithmic_bytes_per_step": gather_bytes},
            "roofline_apply": {"bound": "hbm", "achieved": round(apply_bytes / p_ms / 1e6, 1), "peak": peak, "unit": "GB/s",
                               "frac": round(apply_bytes / p_ms / 1e6 / peak, 4), "traffic": ncu_traffic("sparse_apply_kernel<16,sgd>"),
                               "kernel": "sparse_apply_kernel<16,sgd>",
                               "algorithmic_bytes_per_step": int(apply_bytes)},
            "cpu_baseline": cpu_d, "e2e": e2e_d, "gpu_launches": int(launches), "clocks": clocks}


# ------------------------------------------------------------------------------------------------ GPU arm
NVLINK_PEAK_GBS = 770.0  # measured peer copy per direction per GPU on this pool (B200_PROFILING.md); 900 nominal


def make_ps(name, args, rank, world, graph=False, scale=1, aggregate=None):
    """The TF_CONFIG-driven path: TFJob spec -> launcher placement -> this rank's TF_CONFIG / B200PS_TASKS ->
    b200ps.rank.job_from_env -> ParameterServer.  `scale` > 1 shrinks every variable (parity pre-check)."""
    import b200ps
    from b200ps import launcher, rank as rankmod

    label, sizes, S, W, opt = workload(name, world)
    full_params = sum(sizes)  # the policy below looks at the real workload, so the shrunk parity run uses the same plan
    if scale > 1:
        sizes = [max(1, n // scale) for n in sizes]
    agg, ov = aggregate or args.aggregate, args.overlap
    if agg == "auto":
        agg = "multimem" if world >= 4 and b200ps._lib.lib.b200ps_multicast_supported() else "two_level"
    if ov == "auto":  # pipelining pays when the fold is long enough to hide something (C1: 0.32 ms); not for C3-sized sets
        ov = getattr(args, "overlap_tuned", {}).get(name)  # measured on this box by tune_overlap(), when it ran
        if ov is None:
            ov = "on" if world == 2 and full_params >= 50_000_000 else "off"
    cluster = b200ps.Cluster.from_tfjob(tfjob_spec(name, S, W))
    env = launcher.rank_env(launcher.plan_tasks(cluster, world), rank, world, int(os.environ.get("MASTER_PORT", "29500")))
    job = rankmod.job_from_env({**env, "RANK": str(rank), "WORLD_SIZE": str(world)})
    assert (job.shards, job.workers) == (S, W)
    ps = b200ps.ParameterServer(sizes, job.shards, job.workers, optimizer=opt, lr=1e-3 if opt == "adam" else 0.01, rank=rank,
                                world=world, transport=args.transport, aggregate=agg, overlap=ov == "on",
                                use_graph=graph, cluster=job.cluster)
    return ps, label, sizes, S, W, opt


def tune_overlap(name, args, rank, world, steps=20):
    """N > 1 and --overlap auto: which step form is faster on THIS box?  The sequential form (fold -> barrier ->
    applies -> barrier) is steady (C1 multimem: 1.24 / 1.20 ms at N = 4 / 8); the pipelined form (fold publishes
    unit after unit, applies wait in their prologue) depends on how the two kernels share the SMs and was measured
    anywhere between 1.11 and 1.51 ms at N = 4 on different boxes.  So both are timed (10 warm-up + `steps` steps,
    CUDA events, max over ranks -- every rank sees the same numbers and takes the same decision) and the faster one
    is used for the parity pre-check, the timed run and the e2e leg.  Returns what it measured, or None."""
    import torch
    import torch.distributed as dist

    from b200ps import kernels as K

    _, sizes, _, _, _ = workload(name, world)
    if world == 1 or args.overlap != "auto" or args.transport != "peer" or sum(sizes) < 50_000_000:
        return None
    seen = {}
    for ov in ("off", "on"):
        args.overlap_tuned = {name: ov}
        ps, *_ = make_ps(name, args, rank, world)
        ps.init_uniform(1)
        for w in ps.plan.local_workers:
            K.fill_uniform(ps.grad_buffer(w), 2, 1, w)
        with torch.cuda.stream(torch.cuda.Stream()):
            for _ in range(10):
                ps.step()
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(steps):
                ps.step()
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b) / steps], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        seen[ov] = float(t.item())
        ps.check_barriers()
        ps.close()
        del ps
        torch.cuda.empty_cache()
        dist.barrier()
    best = "on" if seen["on"] < 0.97 * seen["off"] else "off"  # the pipelined form must win clearly: it is the less steady one
    args.overlap_tuned = {name: best}
    return {"sequential_ms": round(seen["off"], 4), "pipelined_ms": round(seen["on"], 4), "chosen": "pipelined" if best == "on" else "sequential",
            "steps": steps}


def parity_check(name, args, rank, world, steps=2, scale=64):
    """Two steps of the SAME plan (same S, W, aggregation, transport, barriers, graph replay) on variables
    shrunk by `scale`, every rank's replica and optimizer state compared bit for bit with the oracle
    (the checker); the verdict is all-reduced.  Exits non-zero on a mismatch."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from b200ps import kernels as K
    from oracle import c_oracle as O

    ps, label, sizes, S, W, opt = make_ps(name, args, rank, world, graph=not args.no_graph, scale=scale)
    ps.init_uniform(1)
    P = ps.plan.P
    hp, hm, hv = O.fill_uniform(P, 1, 0, 0, threads=4), np.zeros(P, np.float32), np.zeros(P, np.float32)
    inv_w = np.float32(1.0) / np.float32(W)
    for t in range(1, steps + 1):
        for w in ps.plan.local_workers:
            K.fill_uniform(ps.grad_buffer(w), 2, t, w)
        ps.step()
        hg = [O.fill_uniform(P, 2, t, w, threads=4) for w in range(W)]
        if ps.plan.two_level:
            hg = O.two_level_partials(hg, world, threads=4)
        if opt == "adam":
            O.dense_adam(hp, hm, hv, hg, O.adam_scalars(1e-3, 0.9, 0.999, 1e-8, t), threads=4, inv_w=inv_w)
        else:
            O.dense_sgd(hp, hg, 0.01, threads=4, inv_w=inv_w)
    exact = args.transport == "peer" and not ps.plan.aggregate == "multimem"
    detail = {}
    d = ps.dump()
    got = ps.replica.cpu().numpy()
    if exact:
        ok = np.array_equal(got.view(np.uint32), hp.view(np.uint32))
        for sh, x in d.items():
            b, e = x["range"]
            if "m" in x:
                ok = ok and np.array_equal(x["m"].view(np.uint32), hm[b:e].view(np.uint32))
                ok = ok and np.array_equal(x["v"].view(np.uint32), hv[b:e].view(np.uint32))
    else:
        # Library / in-switch summation order.  What the mode changes is the gradient SUM, so that is what
        # the north-star tolerance is applied to: Adam's first moment m is linear in it and must match to
        # 1e-6 relative (plus the rounding bound of a W-term fp32 sum, W * 2^-24 * max|g|, times 1 - beta1).
        # The parameters themselves cannot match element-wise under ANY reordering: Adam normalises by
        # sqrt(v), so an element whose gradient sum is ~0 moves by up to lr per step whichever sign the last
        # ulp gives it.  They are required to be within 1e-6 relative for >= 99.9 % of the elements and
        # within the Adam step bound (steps * 2 lr) everywhere.
        rel = np.abs(got - hp) <= 1e-6 * np.abs(hp) + 1e-7
        detail["params_within_1e-6"] = float(rel.mean())
        detail["params_max_abs_diff"] = float(np.abs(got - hp).max())
        ok = rel.mean() >= 0.999 and detail["params_max_abs_diff"] <= steps * 2e-3
        if opt == "adam":
            atol_m = W * 2.0 ** -24 * 1.0 * 0.1 * steps
            worst = 0.0
            for sh, x in d.items():
                b, e = x["range"]
                err = np.abs(x["m"] - hm[b:e]) - 1e-6 * np.abs(hm[b:e])
                worst = max(worst, float(err.max()) if err.size else 0.0)
            detail["m_max_abs_excess_over_1e-6_rel"] = worst
            ok = ok and worst <= atol_m
    ps.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok = bool(flag.item())
    res = {"n": world, "ok": ok, "elements": int(P), "steps": steps, "shards": S, "workers": W,
           "compared": "every rank's replica + optimizer state vs oracle, " +
                       ("bit-exact" if exact else "gradient sum (Adam m) at 1e-6 relative; parameters 1e-6 relative on >= 99.9 %"),
           "detail": detail or None,
           "same_plan_as_timed_run": True, "variables_scaled_by": f"1/{scale}"}
    if not ok:
        if rank == 0:
            print(json.dumps({"parity_checked": res}), flush=True)
        raise SystemExit(3)
    return res


def measure_dense(name, args, rank, world, local, steps, warmup, e2e=False, cpu=False, graph=None, profile=True):
    """Device-resident step timing (+ optional host-buffer e2e and CPU baseline) of one dense workload."""
    import torch
    import torch.distributed as dist

    from b200ps import kernels as K

    label, sizes, S, W, opt = workload(name, world)
    cfg = dense_config(label, sizes, S, W, opt)
    small = cfg["params_padded"] * (4 * W + 24) < (512 << 20)
    graphed = (small if graph is None else graph) and not args.no_graph  # launch-bound workloads replay one CUDA graph per step
    if args.graph != "auto":
        graphed = args.graph == "on"
    ps, *_ = make_ps(name, args, rank, world, graph=graphed)
    ps.init_uniform(1)
    for w in ps.plan.local_workers:
        K.fill_uniform(ps.grad_buffer(w), 2, 1, w)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def reduce_max(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    stream = torch.cuda.Stream()  # a real stream: the step is ordered on it (the legacy default stream serialises)
    sync()
    with torch.cuda.stream(stream):
        sampler = ClockSampler(local, 0.01 if world <= 2 else 0.03)
        sampler.start()
        sampler.wait_ready()
        for _ in range(warmup):
            ps.step()
        sync()
        sampler.mark()
        launches0 = K.launch_count()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        host_t0 = time.perf_counter()
        ev[0].record()
        for i in range(steps):
            ps.step()
        ev[1].record()
        host_ms = 1e3 * (time.perf_counter() - host_t0) / steps  # time the host needed to ENQUEUE a step
        sync()
        clocks = sampler.finish()
        launches = K.launch_count() - launches0
        ms_step = reduce_max(ev[0].elapsed_time(ev[1])) / steps
        # phases: a second, shorter pass with events around fold / apply / barriers (adds ~10 % at N = 8,
        # so it is not the timed run)
        phases, per_rank = None, None
        if profile and not args.no_profile:
            ps.profile = True
            for _ in range(2):  # first launches of the sequential kernel variants (lazy module loading) stay out of the means
                ps.step()
            sync()
            ps.phase_ms()
            pe = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            pe[0].record()
            for _ in range(min(steps, 20)):
                ps.step()
            pe[1].record()
            sync()
            phases = {k: round(v, 4) for k, v in ps.phase_ms().items()}
            phases["step_in_profile_pass"] = round(pe[0].elapsed_time(pe[1]) / min(steps, 20), 4)
            ps.profile = False
            if world > 1:
                per_rank = [None] * world
                dist.all_gather_object(per_rank, phases)
    n_launch = len([L for L in ps.plan.launches if L.end > L.begin])
    step_bytes = ps.algorithmic_bytes_per_step()
    peak, peak_src = hbm_peak()
    kname = f"dense_apply_kernel<{opt},W={len(ps.plan.launches[0].grad_src) if ps.plan.launches else W}>"
    if world == 1:
        # one GPU: the step IS its apply launches (back to back on one stream), so the average launch
        # duration is step time / launches; phases["apply"] (events around the launches) agrees
        k_ms = (phases["apply"] if phases else ms_step) / max(n_launch, 1)
        ach = step_bytes / max(n_launch, 1) / k_ms / 1e6
        roofline = {"bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                    "traffic": ncu_traffic(f"dense_apply_kernel<{opt},W={W}>"), "kernel": kname, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": int(step_bytes / max(n_launch, 1)), "avg_launch_ms": round(k_ms, 5),
                    "launches_per_step": n_launch}
    else:
        # N > 1: per step a GPU receives the partial sums (or gradients) of its slices from the other GPUs
        # plus every other owner's updated slices into its replica -- all inbound on its NVLink ports
        mine = sum(L.end - L.begin for L in ps.plan.launches)
        inbound = 4 * (sum((L.end - L.begin) * sum(1 for r, _ in L.grad_src if r != rank) for L in ps.plan.launches)
                       + (ps.plan.P - mine))
        inbound = reduce_max(float(inbound))
        ach = inbound / ms_step / 1e6
        roofline = {"bound": "nvlink", "achieved": round(ach, 1), "peak": NVLINK_PEAK_GBS, "unit": "GB/s",
                    "frac": round(ach / NVLINK_PEAK_GBS, 4), "traffic": None,
                    "kernel": kname + " (peer loads = push, peer stores = pull) + dense_fold + 2 peer barriers",
                    "peak_source": "measured peer copy per direction per GPU (B200_PROFILING.md); 900 nominal",
                    "inbound_bytes_per_gpu_per_step": int(inbound),
                    "note": "achieved = inbound NVLink bytes of the busiest GPU / whole step time (local fold and barriers included)"}
    out = {"metric": "ps_steps_per_sec", "value": round(W * 1e3 / ms_step, 3), "unit": "worker-steps/s", "n_gpus": world,
           "steps": steps, "warmup": warmup, "ms_per_step": round(ms_step, 5), "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
           "impl_config": {"transport": args.transport if world > 1 else "local",
                           "aggregate": {"multimem": "per-GPU partial sums added in the NVSwitch (multimem.ld_reduce), updated slices "
                                                     "multicast (multimem.st); 1e-6, not bit-exact",
                                         "two_level": "two-level fixed-order fold (per-GPU partial sums, then GPU order); bit-exact",
                                         "flat": "flat fixed-order fold; bit-exact"}[ps.plan.aggregate if world > 1 else "flat"],
                           "barrier": "peer-memory flags (device-side epoch)" if world > 1 else "none",
                           "pipelined": bool(ps.overlap), "cuda_graph": bool(graphed),
                           "driven_by": "TFJob spec -> launcher placement -> TF_CONFIG / B200PS_TASKS -> b200ps.rank.job_from_env"},
           "ps_steps_per_sec": round(1e3 / ms_step, 3), "host_enqueue_ms_per_step": round(host_ms, 4),
           "algorithmic_GBps_per_gpu": round(step_bytes / ms_step / 1e6, 1), "roofline": roofline,
           "phases_ms": phases, "phases_ms_per_rank": per_rank, "gpu_launches": int(launches), "clocks": clocks}

    # ---- end to end with HOST buffers ----
    if e2e:
        P = ps.plan.P
        n_e2e = args.e2e_steps or min(steps, 30)

        def timed_host_loop(fn, tail, n=None, sampler=None):
            n = n or n_e2e
            with torch.cuda.stream(stream):
                for _ in range(5):
                    fn()
                sync()
                if sampler:
                    sampler.mark()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(n):
                    fn()
                b.record()
                tail()
                sync()
            return reduce_max(a.elapsed_time(b)) / n

        # (1) host gradients: every worker's pinned host gradient in, pulled parameters out to one pinned host
        #     replica per GPU (the workers of a GPU share it as they share the device replica)
        hg = {w: torch.empty(P, dtype=torch.float32).pin_memory() for w in ps.plan.local_workers}
        ho = {ps.plan.local_workers[0]: torch.empty(P, dtype=torch.float32).pin_memory()} if ps.plan.local_workers else {}
        for w in hg:
            hg[w].copy_(ps.grad_buffer(w))
        sync()
        e_ms = timed_host_loop(lambda: ps.step_host(hg, ho), ps.host_sync)
        host_grad = {"value": round(W * 1e3 / e_ms, 3), "unit": "worker-steps/s", "h2d_bytes_per_step": 4 * P * W,
                     "d2h_bytes_per_step": 4 * P * world, "ms_per_step": round(e_ms, 3), "steps": n_e2e,
                     "ps_steps_per_sec": round(1e3 / e_ms, 3),
                     "pcie_bound_ms": round(4 * P * len(ps.plan.local_workers) / 55e6, 2),
                     "api": "b200ps_dense_step_host via b200ps.ParameterServer.step_host (every worker's pinned host gradient in, "
                            "pulled parameters out to one pinned host replica per GPU; chunk-pipelined inside the library)"}
        del hg, ho
        out["e2e"] = host_grad
        if name == "c1":
            # (2) the training step a user runs: the compute tasks are GPU tasks (b200ps_mlp_step: hand-written
            #     tcgen05 GEMMs), so the host input of a step is the mini-batch; gradients never cross PCIe
            from b200ps.worker import C1_DIMS, MlpWorker

            ps.init_uniform(1)
            ps.replica.mul_(0.02)  # weights of a trainable scale (u(-1, 1) * 0.02)
            wk = MlpWorker(ps, C1_DIMS, args.batch)
            gen = torch.Generator().manual_seed(7 + rank)
            hx = torch.randn(wk.rows, C1_DIMS[0], generator=gen).bfloat16().pin_memory()
            ht = torch.randn(wk.rows, C1_DIMS[-1], generator=gen).bfloat16().pin_memory()

            def train_step():
                wk.step(hx, ht)  # H2D of the mini-batches, forward/backward of the local tasks, loss D2H
                ps.step()        # push (zero-copy) + aggregate + apply + pull (zero-copy)

            smp = ClockSampler(local, 0.01 if world <= 2 else 0.03)
            smp.start()
            smp.wait_ready()
            n_train = args.e2e_steps or min(steps, 100)
            t_ms = timed_host_loop(train_step, lambda: None, n_train, sampler=smp)
            e2e_clocks = smp.finish()
            w_ms = timed_host_loop(lambda: wk.step(hx, ht), lambda: None)  # the worker half alone, for the breakdown
            losses = [float(v) for v in wk.loss]
            out["e2e"] = {"value": round(W * 1e3 / t_ms, 3), "unit": "worker-steps/s",
                          "h2d_bytes_per_step": 2 * wk.rows * (C1_DIMS[0] + C1_DIMS[-1]) * world,
                          "d2h_bytes_per_step": 4 * len(wk.tasks) * world, "ms_per_step": round(t_ms, 3), "steps": n_train,
                          "ps_steps_per_sec": round(1e3 / t_ms, 3), "batch_per_task": args.batch,
                          "worker_TFLOPs_per_gpu": round(wk.flops_per_step() / t_ms / 1e9, 1),
                          "loss_of_last_step": [round(v, 3) for v in losses], "worker_only_ms": round(w_ms, 3),
                          "clocks": e2e_clocks,
                          "api": "b200ps_mlp_step + b200ps_dense_step (b200ps.MlpWorker.step, ParameterServer.step): per step the "
                                 "mini-batches of the local compute tasks go H2D from pinned memory, the tasks' forward/backward "
                                 "runs on the GPU (bf16 tcgen05 GEMMs, fp32 gradients written into the PS arena), the fused PS "
                                 "step follows, one loss per task comes back D2H",
                          "host_gradients": host_grad}
            wk.close()
    ps.close()
    del ps
    torch.cuda.empty_cache()
    # ---- CPU baseline (rank 0, N = 1 only) ----
    if world == 1 and cpu:
        secs = args.cpu_seconds if cpu is True else cpu
        val, done, ms, nt, gbs = cpu_run(label, sizes, S, W, opt, 10 ** 6, 1, seconds=secs)
        out["cpu_baseline"] = {"value": round(val, 3), "unit": "worker-steps/s", "cores": nt, "kind": "port",
                               "sample": f"{done} full steps ({ms:.1f} ms each, {gbs:.0f} GB/s algorithmic) of the same workload within "
                                         f"{secs:.0f} s; substitute CPU reference (oracle/ps_oracle.c, OpenMP)"}
        if cpu is True:
            v1, d1, ms1, _, g1 = cpu_run(label, sizes, S, W, opt, 10 ** 6, 1, seconds=secs / 4, threads=1)
            out["cpu_baseline"]["single_thread"] = {"value": round(v1, 3), "ms_per_step": round(ms1, 1),
                                                    "algorithmic_GBps": round(g1, 1), "steps": d1}
    return out


def brief(d, keys=("value", "unit", "ms_per_step", "ps_steps_per_sec", "roofline", "roofline_apply", "phases_ms", "cpu_baseline",
                   "e2e", "gpu_launches", "steps")):
    out = {k: d[k] for k in keys if d.get(k) is not None}
    out["workload"] = d["config"]["workload"]
    return out


def main():
    args = parse()
    if args.workload == "c2":
        return sparse_arm(args)
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    warm = max(args.warmup, 3)
    tuned = tune_overlap(args.workload, args, rank, world)
    parity = parity_check(args.workload, args, rank, world)
    out = measure_dense(args.workload, args, rank, world, local, args.steps, warm, e2e=not args.no_e2e, cpu=not args.no_cpu)
    out["parity_checked"] = parity
    out["impl_config"]["step_form_autotune"] = tuned
    out.setdefault("cpu_baseline", None)
    out.setdefault("e2e", None)
    # ---- the other BASELINE configs, briefly, in the same process (not the headline) ----
    also = {}
    if not args.no_also:
        n_also = max(10, min(args.steps, 30))
        if args.workload != "c3":  # BASELINE configs[3]: ResNet-50-sized set at W = S = G
            also["c3"] = brief(measure_dense("c3", args, rank, world, local, n_also, 3, cpu=(3.0 if world == 1 else False)))
        if world == 1 and args.workload != "c0":  # BASELINE configs[0] (launch-bound: graph replay)
            also["c0"] = brief(measure_dense("c0", args, rank, world, local, 200, 10, cpu=2.0, profile=False))
        for kind in (("uniform", "zipf") if world == 1 else ("uniform",)):  # BASELINE configs[2] on N GPUs
            also[f"c2_{kind}"] = brief(sparse_measure(args, kind, n_also, 3, e2e=False, cpu=(3.0 if world == 1 else False)),
                                       keys=("value", "unit", "ms_per_step", "ps_steps_per_sec", "phases_ms", "roofline",
                                             "roofline_apply", "cpu_baseline", "gpu_launches", "steps", "scaling"))
    out["also"] = also or None
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
