"""K8 timing: device time per call (host queued ahead behind a spin kernel) and host enqueue time per call, for the
fused o_proj + all-reduce, the NCCL pair it replaces and the bare GEMM.  One process per GPU:

    python scripts/oproj_bench.py                                   # world 1 (GEMM + local reduce only)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
        scripts/oproj_bench.py
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from baby_sglang_b200 import DROPIN_DIR  # noqa: E402

sys.path.insert(0, DROPIN_DIR)
import tp_runtime  # noqa: E402


def main():
    multi = "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1
    if multi:
        rank, world, group = tp_runtime.init_tp()
    else:
        rank, world, group = 0, 1, None
        torch.cuda.set_device(0)
    dev = torch.device("cuda", torch.cuda.current_device())
    shapes = [(1024, 5120, 5120), (256, 5120, 5120), (64, 5120, 5120), (1024, 3584, 3584)]
    n = 40
    for T, hidden, Kfull in shapes:
        K = Kfull // world
        if K % 64:
            continue
        comm = tp_runtime.FusedOProjAllReduce(rank, world, group, hidden=hidden, max_tokens=1024)
        a = torch.randn(T, K, device=dev, dtype=torch.bfloat16)
        w = torch.randn(hidden, K, device=dev, dtype=torch.bfloat16) * 0.05

        def fused():
            return comm(a, w)

        def gemm():
            return torch.nn.functional.linear(a, w)

        def pair():
            o = torch.nn.functional.linear(a, w)
            if world > 1:
                dist.all_reduce(o, group=group)
            return o

        def timed(fn):
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier(group=group)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(int(8e6))                    # ~4 ms: the host queues all n calls behind it
            e0.record()
            t0 = time.perf_counter()
            for _ in range(n):
                fn()
            host_us = (time.perf_counter() - t0) / n * 1e6
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / n * 1e3, host_us], device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            return [round(float(x), 1) for x in t]

        line = {"T": T, "hidden": hidden, "K_per_rank": K, "world": world}
        flops = 2.0 * T * hidden * K
        for mode in ("single", "pair"):                    # K8's GEMM form: 128 x 128 single-CTA tiles / 256 x 256 CTA pairs
            os.environ["BSGL_K8_GEMM"] = mode
            d, h = timed(fused)
            line[f"fused_{mode}_device_us"], line[f"fused_{mode}_tflops_per_rank"] = d, round(flops / (d * 1e-6) / 1e12, 1)
        os.environ.pop("BSGL_K8_GEMM", None)
        for name, fn in (("fused", fused), ("nccl_pair", pair), ("gemm_only", gemm)):
            d, h = timed(fn)
            line[name + "_device_us"], line[name + "_host_us"] = d, h
        line["fused_tflops_per_rank"] = round(flops / (line["fused_device_us"] * 1e-6) / 1e12, 1)
        line["gemm_only_tflops"] = round(flops / (line["gemm_only_device_us"] * 1e-6) / 1e12, 1)
        if rank == 0:
            print(json.dumps(line), flush=True)
        comm.close()
    if world > 1:
        dist.barrier(group=group)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
