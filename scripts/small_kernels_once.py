"""Development aid: a few launches of K1 / K5 / K6 at their bench.py sizes, for an `ncu --set full` capture
(`ncu -k regex:"kv_store|greedy" ... python scripts/small_kernels_once.py`).  bench.py is the contract."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from baby_sglang_b200 import ops  # noqa: E402

dev = torch.device("cuda")
Hq, Hkv, D, vocab, T, S = 28, 4, 128, 152064, 16384, 1 << 20
dt = torch.bfloat16
ks = torch.empty(S + 1, Hkv, D, device=dev, dtype=dt)
vs = torch.empty(S + 1, Hkv, D, device=dev, dtype=dt)
inv = 1.0 / (1e6 ** (torch.arange(0, D, 2, device=dev).float() / D))
sets = []
for i in range(3):
    loc = (torch.randperm(S, device=dev)[:T] + 1).to(torch.int32)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, device=dev, dtype=dt)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    sets.append((loc, qkv, k.contiguous().view(T, Hkv, D), v.view(T, Hkv, D), torch.randint(0, 4096, (T,), device=dev),
                 torch.randn(256, vocab, device=dev, dtype=dt)))
for loc, qkv, k, v, pos, logits in sets:
    ops.kv_store(ks, vs, loc, k, v)
    ops.rope_kv_store(qkv, pos, inv, Hq, Hkv, D, ks, vs, loc)
    ops.greedy_sample(logits)
torch.cuda.synchronize()
print("done")
