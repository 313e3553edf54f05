"""Experiment (N GPUs of one box, under torchrun): variants of the multimem C1 step inside ONE process
group, so that a comparison costs one rendezvous.  Each variant builds its own ParameterServer through
bench.make_ps, runs 10 warm-up and 50 timed steps (CUDA events, max over ranks) and rank 0 prints a JSON
line.  Variants are environment knobs of the library (B200PS_STEP_EVENTS, B200PS_MC_ROTATE) and the
pipelined / sequential step form.  Not a product path."""
import importlib.util
import json
import os
import sys
import types

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from b200ps import kernels as K
    # (name, pipelined, environment knobs, on a torch side stream, NVML clock sampler running)
    variants = [
        ("sequential, default stream", "off", {}, False, False),
        ("sequential, side stream", "off", {}, True, False),
        ("sequential, side stream, clock sampler", "off", {}, True, True),
        ("sequential, default stream, clock sampler", "off", {}, False, True),
        ("pipelined, side stream", "on", {}, True, False),
        ("pipelined, default stream", "on", {}, False, False),
    ]
    only = os.environ.get("VARIANTS")
    side = torch.cuda.Stream()
    for name, ov, env, on_side, sample in variants:
        if only and not any(x in name for x in only.split("|")):
            continue
        for k in ("B200PS_STEP_EVENTS", "B200PS_MC_ROTATE"):
            os.environ.pop(k, None)
        os.environ.update(env)
        args = types.SimpleNamespace(aggregate="multimem", overlap=ov, transport="peer")
        ps, label, sizes, S, W, opt = bench.make_ps("c1", args, rank, world)
        ps.init_uniform(1)
        for w in ps.plan.local_workers:
            K.fill_uniform(ps.grad_buffer(w), 2, 1, w)
        torch.cuda.synchronize()
        with torch.cuda.stream(side if on_side else torch.cuda.default_stream()):
            for _ in range(10):
                ps.step()
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            sampler = bench.ClockSampler(torch.cuda.current_device()) if sample else None
            if sampler:
                sampler.start()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(100):
                ps.step()
            b.record()
            torch.cuda.synchronize()
            if sampler:
                sampler.finish()
            dist.barrier()
            t = torch.tensor([a.elapsed_time(b) / 100], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ps.check_barriers()
        if rank == 0:
            print(json.dumps({"variant": name, "n_gpus": world, "ms_per_step": round(float(t.item()), 4)}), flush=True)
        ps.close()
        del ps
        torch.cuda.empty_cache()
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
