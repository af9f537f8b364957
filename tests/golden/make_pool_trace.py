"""Generate tests/golden/pool_trace.json: a seeded alloc/free trace replayed on the UNMODIFIED reference pools
(`ReqToTokenPool`, `BaseTokenToKVPool` via `MHATokenToKVPool`; memory_pool.py:7-102), recording every result.
KV-slot indexing must be bit-exact with the reference, and the allocators' FIFO order is where that lives.

    python tests/golden/make_pool_trace.py        (build container only)
"""
import json
import os
import random
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from baseline import ref_loader  # noqa: E402


def trace(ref, seed, size, n_ops):
    rnd = random.Random(seed)
    kv = ref.memory_pool.MHATokenToKVPool(size=size, dtype=torch.float32, head_num=1, head_dim=8, layer_num=1,
                                          device="cpu")
    rq = ref.memory_pool.ReqToTokenPool(size=16, max_context_len=8, device="cpu")
    held, held_rows, ops = [], [], []
    for _ in range(n_ops):
        kind = rnd.choice(["alloc", "alloc", "free", "ralloc", "rfree", "partial_free"])
        if kind == "alloc":
            n = rnd.randint(0, size // 3)
            got = kv.alloc(n)
            res = None if got is None else got.tolist()
            if res:
                held.append(res)
            ops.append(["alloc", n, res, kv.available_size()])
        elif kind == "free" and held:
            blk = held.pop(rnd.randrange(len(held)))
            kv.free(torch.tensor(blk, dtype=torch.int32))
            ops.append(["free", blk, None, kv.available_size()])
        elif kind == "partial_free" and held:
            blk = held.pop(rnd.randrange(len(held)))
            k = rnd.randint(0, len(blk))
            part, rest = blk[:k][::-1], blk[k:]           # out-of-order return
            if rest:
                held.append(rest)
            kv.free(torch.tensor(part, dtype=torch.int32))
            ops.append(["free", part, None, kv.available_size()])
        elif kind == "ralloc":
            n = rnd.randint(1, 6)
            got = rq.alloc(n)
            if got:
                held_rows.append(list(got))
            ops.append(["ralloc", n, None if got is None else list(got), rq.available_size()])
        elif kind == "rfree" and held_rows:
            blk = held_rows.pop(rnd.randrange(len(held_rows)))
            arg = blk[0] if len(blk) == 1 and rnd.random() < 0.5 else blk
            rq.free(arg)
            ops.append(["rfree", arg, None, rq.available_size()])
    return {"size": size, "ops": ops, "final_free_slots": kv.free_slots.tolist(), "final_rows": list(rq.free_slots)}


if __name__ == "__main__":
    ref = ref_loader.load_stack(names=("memory_pool",), optional=())
    out = {f"seed{s}": trace(ref, s, size, 400) for s, size in ((1, 64), (2, 257), (3, 1000))}
    with open(os.path.join(HERE, "pool_trace.json"), "w") as f:
        json.dump(out, f)
    print({k: len(v["ops"]) for k, v in out.items()})
