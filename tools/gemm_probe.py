"""Throughput of the worker's tcgen05 GEMM on the C1 shapes, next to torch.matmul (cuBLAS) on the same
operands, and the whole MLP worker step (4 tasks x batch) + PS step.  One GPU."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ps  # noqa: E402
from b200ps import synth  # noqa: E402
from b200ps.worker import C1_DIMS, MlpWorker, gemm_bf16, mlp_variables  # noqa: E402


def timed(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    torch.cuda.set_device(0)
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    M = 4 * batch
    shapes = [("fwd L0", M, 8192, 1024), ("fwd L1", M, 8192, 8192), ("fwd L2", M, 2048, 8192), ("fwd L3", M, 4096, 2048),
              ("wgrad L1 (1 task)", 8192, 8192, batch), ("big", 8192, 8192, 8192)]
    for name, m, n, k in ([] if "mlp-only" in sys.argv else shapes):
        a = torch.randn(m, k, device="cuda").bfloat16()
        b = torch.randn(n, k, device="cuda").bfloat16()
        c = torch.empty(m, n, dtype=torch.bfloat16, device="cuda")
        ct = torch.empty(n, m, dtype=torch.bfloat16, device="cuda")
        ms = timed(lambda: gemm_bf16(a, b, out=c))
        ms_t = timed(lambda: gemm_bf16(a, b, out=c, out_t=ct))
        ms_ref = timed(lambda: torch.matmul(a, b.t(), out=c))
        fl = 2.0 * m * n * k
        print(json.dumps({"gemm": name, "M": m, "N": n, "K": k, "ms": round(ms, 4), "TFLOPs": round(fl / ms / 1e9, 1),
                          "ms_with_transposed_copy": round(ms_t, 4), "torch_matmul_ms": round(ms_ref, 4),
                          "torch_TFLOPs": round(fl / ms_ref / 1e9, 1)}), flush=True)
    ps = b200ps.ParameterServer(synth.mlp_c1_variables(), 2, 4, optimizer="adam")
    ps.init_uniform(1)
    ps.replica.mul_(0.02)
    wk = MlpWorker(ps, C1_DIMS, batch)
    x = torch.randn(M, C1_DIMS[0]).bfloat16().pin_memory()
    t = torch.randn(M, C1_DIMS[-1]).bfloat16().pin_memory()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        ms_w = timed(lambda: wk.step(x, t))
        ms_both = timed(lambda: (wk.step(x, t), ps.step()))
    print(json.dumps({"mlp_worker_step_ms": round(ms_w, 4), "TFLOPs": round(wk.flops_per_step() / ms_w / 1e9, 1),
                      "worker_plus_ps_step_ms": round(ms_both, 4), "batch_per_task": batch, "tasks": 4,
                      "loss": [round(float(v), 4) for v in wk.loss]}))


if __name__ == "__main__":
    main()
