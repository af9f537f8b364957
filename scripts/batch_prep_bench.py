"""Ad-hoc timing of the K7 batch-prep hooks against the reference's pattern (development aid).

Decode prep at NS size (B=256, C=4160): the reference issues B scalar device writes + three `torch.tensor`
uploads (scheduler.py:486-497); the hook issues one upload + one launch.  Prefill prep (B=32 x 512 new tokens,
prefix 1024): zeros([B, C]) + 2 slice writes and one `.item()` per request + a whole-row index-put
(scheduler.py:423-442) vs one upload + one launch.  Wall clock including a final synchronize.
"""
import os
import sys
import time
from types import SimpleNamespace

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from baby_sglang_b200 import DROPIN_DIR, model_hooks  # noqa: E402

sys.path.insert(0, DROPIN_DIR)
import memory_pool  # noqa: E402

dev = "cuda"
B, C = 256, 4160
kv = memory_pool.MHATokenToKVPool(size=B * C, dtype=torch.bfloat16, head_num=1, head_dim=16, layer_num=1, device=dev)
rq = memory_pool.ReqToTokenPool(size=B, max_context_len=C, device=dev)
sched = SimpleNamespace(req_to_token_pool=rq, token_to_kv_pool=kv, device=dev,
                        tree_cache=SimpleNamespace(evict=lambda n, cb: None, evictable_size=lambda: 0))
reqs = [SimpleNamespace(output_ids=[1], req_pool_idx=i, seq_len=2000, overlap_next_input_id=None, is_retracted=False,
                        input_ids=torch.zeros(1536, dtype=torch.int64), num_cached_tokens=1024,
                        prefix_indices=torch.arange(1, 1025, dtype=torch.int32, device=dev)) for i in range(B)]


def reference_decode_prep():
    slots = kv.alloc(B)
    rows, ids, lens = [], [], []
    for i, r in enumerate(reqs):
        rows.append(r.req_pool_idx)
        ids.append(r.output_ids[-1])
        rq.req_to_token[r.req_pool_idx, r.seq_len] = slots[i]
        r.seq_len += 1
        lens.append(r.seq_len)
    return (torch.tensor(rows, dtype=torch.long, device=dev), torch.tensor(lens, dtype=torch.long, device=dev),
            torch.tensor(ids, dtype=torch.long, device=dev), slots)


def reference_prefill_prep(batch):
    n = len(batch)
    rows = torch.tensor(list(range(n)), dtype=torch.long, device=dev)
    seq = [len(r.input_ids) for r in batch]
    total = sum(s - r.num_cached_tokens for s, r in zip(seq, batch))
    loc = kv.alloc(total)
    mapping = torch.zeros((n, C), dtype=torch.int32, device=dev)
    off = 0
    for i, s in enumerate(seq):
        r = batch[i]
        mapping[i, :r.num_cached_tokens] = r.prefix_indices
        mapping[i, r.num_cached_tokens:s] = loc[off:off + s - r.num_cached_tokens]
        off += s - r.num_cached_tokens
        r.req_pool_idx = rows[i].item()
    rq.write(rows, mapping)
    return loc


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / n * 1e3


def hook_decode():
    fb = model_hooks._device_prepare_decode_batch(sched, reqs)
    kv.free(fb.out_cache_loc)


def ref_decode():
    out = reference_decode_prep()
    kv.free(out[3])


def hook_prefill():
    rq.clear()
    fb = model_hooks._device_prepare_prefill_batch(sched, reqs[:32])
    kv.free(fb.out_cache_loc)


def ref_prefill():
    kv.free(reference_prefill_prep(reqs[:32]))


for r in reqs:
    r.seq_len = 2000
print("decode prep  B=256:   reference pattern %.3f ms   K7 hook (+K4 metadata) %.3f ms" % (timeit(ref_decode), timeit(hook_decode)))
for r in reqs:
    r.seq_len = 0
print("prefill prep B=32:    reference pattern %.3f ms   K7 hook (+K4 metadata) %.3f ms" % (timeit(ref_prefill), timeit(hook_prefill)))
