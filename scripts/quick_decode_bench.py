"""Ad-hoc kernel timing of the paged decode kernel (development aid; bench.py is the contract)."""
import argparse
import json
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from baby_sglang_b200 import ops  # noqa: E402

GEOM = {"1.5b": (12, 2, 128), "7b": (28, 4, 128), "32b": (40, 8, 128), "32b_tp8": (5, 1, 128)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geom", default="7b")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--ctx", type=int, default=4096)
    ap.add_argument("--ragged", type=int, default=0, help="if >0: seq ~ U[ctx, ragged]")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--n-ctas", type=int, default=0)
    ap.add_argument("--no-merge", action="store_true", help="time K2 alone (skip the K2c merge launch)")
    ap.add_argument("--sequential", action="store_true", help="unfragmented page table (slots in order)")
    a = ap.parse_args()
    Hq, Hkv, D = GEOM[a.geom]
    dev = torch.device("cuda")
    g = torch.Generator(device="cuda").manual_seed(0)
    if a.ragged:
        seq = torch.randint(a.ctx, a.ragged + 1, (a.batch,), generator=g, device=dev)
    else:
        seq = torch.full((a.batch,), a.ctx, device=dev, dtype=torch.int64)
    C = int(seq.max().item()) + 64
    S = int(seq.sum().item()) + 1024
    perm = torch.randperm(S, generator=g, device=dev).to(torch.int32) + 1
    if a.sequential:
        perm = torch.arange(1, S + 1, device=dev, dtype=torch.int32)
    r2t = torch.zeros(a.batch, C, dtype=torch.int32, device=dev)
    off = 0
    for i, n in enumerate(seq.tolist()):
        r2t[i, :n] = perm[off:off + n]
        off += n
    rpi = torch.arange(a.batch, device=dev)
    slabs = [(torch.randn(S + 1, Hkv, D, device=dev, dtype=torch.bfloat16),
              torch.randn(S + 1, Hkv, D, device=dev, dtype=torch.bfloat16)) for _ in range(a.layers)]
    q = torch.randn(a.batch, Hq, D, device=dev, dtype=torch.bfloat16)
    md = ops.build_decode_metadata(r2t, rpi, seq, Hkv, n_ctas=a.n_ctas or None)
    ws = ops.partial_buffers(dev, md.plan.n_ctas, Hq, Hkv, D)
    scale = D ** -0.5
    for i in range(3):
        ops.paged_attention(q, *slabs[i % a.layers], r2t, md.plan, scale, workspace=ws, merge=not a.no_merge)
    torch.cuda.synchronize()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(a.iters)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(a.iters)]
    for i in range(a.iters):
        e0[i].record()
        ops.paged_attention(q, *slabs[i % a.layers], r2t, md.plan, scale, workspace=ws, merge=not a.no_merge)
        e1[i].record()
    torch.cuda.synchronize()
    ms = sorted(x.elapsed_time(y) for x, y in zip(e0, e1))
    tot = int(seq.sum().item())
    bytes_ = tot * Hkv * D * 2 * 2 + 4 * tot + 2 * a.batch * Hq * D * 2
    med = ms[len(ms) // 2]
    print(json.dumps({"geom": a.geom, "batch": a.batch, "ctx": a.ctx, "ragged": a.ragged, "sequential": a.sequential, "merge": not a.no_merge, "ms_med": med,
                      "ms_min": ms[0], "GBps_med": bytes_ / med / 1e6, "GBps_best": bytes_ / ms[0] / 1e6,
                      "frac_of_6524.6": bytes_ / med / 1e6 / 6524.6, "bytes": bytes_,
                      "plan_hdr": md.plan.plan_hdr.tolist()}))


if __name__ == "__main__":
    main()
