"""Small-batch decode step (K4 + L x (K1, K2, K2c)): eager launches vs one CUDA-graph replay.

At small batch the kernels are short (tens of microseconds) and the Python/driver launch path dominates an eager
step; the step makes no host sync, so it captures into one graph (programmatic dependent launches included) and a
replay costs the kernels only.  Development aid; bench.py is the contract.

    python scripts/graph_decode_bench.py [--batch 16] [--ctx 4096] [--layers 28]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from baby_sglang_b200 import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--ctx", type=int, default=4096)
    ap.add_argument("--layers", type=int, default=28)
    ap.add_argument("--iters", type=int, default=30)
    a = ap.parse_args()
    Hq, Hkv, D = 28, 4, 128                                   # Qwen2.5-7B
    dev = torch.device("cuda")
    B, L = a.batch, a.layers
    g = torch.Generator(device="cuda").manual_seed(0)
    seq = torch.full((B,), a.ctx, device=dev, dtype=torch.int64)
    S = B * a.ctx + 64
    C = a.ctx + 8
    perm = torch.randperm(S, generator=g, device=dev).to(torch.int32) + 1
    r2t = torch.zeros(B, C, dtype=torch.int32, device=dev)
    r2t[:, :a.ctx] = perm[:B * a.ctx].view(B, a.ctx)
    rpi = torch.arange(B, device=dev)
    loc = r2t[rpi, seq - 1].contiguous()
    slabs = [(torch.randn(S + 1, Hkv, D, device=dev, dtype=torch.bfloat16),
              torch.randn(S + 1, Hkv, D, device=dev, dtype=torch.bfloat16)) for _ in range(L)]
    q = torch.randn(B, Hq, D, device=dev, dtype=torch.bfloat16)
    k = torch.randn(B, Hkv, D, device=dev, dtype=torch.bfloat16)
    v = torch.randn(B, Hkv, D, device=dev, dtype=torch.bfloat16)
    outs = [torch.empty(B, Hq * D, device=dev, dtype=torch.bfloat16) for _ in range(L)]
    scale = D ** -0.5

    def step():
        md = ops.build_decode_metadata(r2t, rpi, seq, Hkv)
        for l in range(L):
            ops.kv_store(slabs[l][0], slabs[l][1], loc, k, v)
            ops.paged_attention(q, slabs[l][0], slabs[l][1], r2t, md.plan, scale, out=outs[l])

    def timed(fn):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / a.iters

    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        step()
    side.synchronize()
    eager_ms = timed(step)
    ref = [o.clone() for o in outs]
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        step()
    for o in outs:
        o.zero_()
    graph_ms = timed(graph.replay)
    same = all(torch.equal(x, y) for x, y in zip(ref, outs))
    bytes_ = L * (B * a.ctx * Hkv * D * 2 * 2 + 4 * B * a.ctx + 2 * B * Hq * D * 2)
    print(json.dumps({"batch": B, "ctx": a.ctx, "layers": L, "eager_ms_per_step": round(eager_ms, 4),
                      "graph_ms_per_step": round(graph_ms, 4), "speedup": round(eager_ms / graph_ms, 2),
                      "graph_us_per_layer": round(graph_ms / L * 1e3, 1),
                      "graph_GBps": round(bytes_ / graph_ms / 1e6, 1), "replay_bit_identical": same}))


if __name__ == "__main__":
    main()
