"""K8 (fused o_proj + all-reduce) under torchrun: time per call and the per-CTA phase timestamps the kernel leaves
behind (entry / last partial tile published / last owned tile delivered / exit), for both exchange paths.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29551 \
        scripts/k8_trace.py
Development aid; bench.py is the contract."""
import json
import os
import sys

import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from baby_sglang_b200 import DROPIN_DIR  # noqa: E402

sys.path.insert(0, DROPIN_DIR)
import tp_runtime  # noqa: E402


def timed(fn, group, dev, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    dist.barrier(group=group)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return round(float(t) * 1e3, 2)


def main():
    rank, world, group = tp_runtime.init_tp()
    dev = torch.device("cuda", torch.cuda.current_device())
    for T, hidden, Kfull in ((1024, 5120, 5120), (256, 3584, 3584), (256, 5120, 5120), (64, 3584, 3584)):
        if Kfull % (world * 64):
            continue
        K = Kfull // world
        a = torch.randn(T, K, device=dev, dtype=torch.bfloat16)
        w = torch.randn(hidden, K, device=dev, dtype=torch.bfloat16) * 0.05
        line = {"T": T, "hidden": hidden, "K_per_rank": K, "world": world}

        def pair():
            o = torch.nn.functional.linear(a, w)
            dist.all_reduce(o, group=group)
            return o

        line["nccl_pair_us"] = timed(pair, group, dev)
        line["gemm_only_us"] = timed(lambda: torch.nn.functional.linear(a, w), group, dev)
        for mode in ("nvls", "symm-peer", "peer"):
            os.environ["BSGL_K8_EXCHANGE"] = mode
            comm = tp_runtime.FusedOProjAllReduce(rank, world, group, hidden=hidden, max_tokens=1024)
            us = timed(lambda: comm(a, w), group, dev)
            dist.barrier(group=group)
            comm(a, w)
            tr = comm.trace()
            def stat(col):
                v = tr[:, col][tr[:, col] > 0]
                return None if v.numel() == 0 else [round(float(v.median()), 1), round(float(v.max()), 1)]
            line[mode] = {"exchange": comm.last_exchange, "us": us, "ctas": int(tr.shape[0]),
                          "entry_spread_us": round(float(tr[:, 0].max()), 1),
                          "published_med_max_us": stat(1), "delivered_med_max_us": stat(2), "exit_med_max_us": stat(3)}
            comm.close()
        os.environ["BSGL_K8_EXCHANGE"] = ""
        if rank == 0:
            print(json.dumps(line), flush=True)
    dist.barrier(group=group)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
