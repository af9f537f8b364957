"""Generate tests/golden/batch_prep_trace.json with the UNMODIFIED reference scheduler (build container only).

    python tests/golden/make_batch_prep_golden.py

A scripted sequence of `Scheduler._prepare_prefill_batch` / `_prepare_decode_batch` calls
(scheduler.py:381-507) and `token_to_kv_pool.free` calls on the reference's own pools, CPU.  Requests get
hand-made prefix hits (`prefix_indices` = slots taken from the pool beforehand, as a radix match would return
them, scheduler.py:264-268).  After every call the trace records the batch tensors, the page-table rows of every
live request and the free-slot FIFO, which the K7 hooks must reproduce bit for bit.
"""
from __future__ import annotations

import json
import os
import random
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from baseline import ref_loader  # noqa: E402

TOTAL, REQS = 640, 8          # C = 80


def main():
    os.environ["BABYSGL_MAX_TOTAL_TOKENS"] = str(TOTAL)
    os.environ["BABYSGL_MAX_REQUESTS"] = str(REQS)
    ref = ref_loader.load_stack()
    cfg = ref.model_loader.BabyQwenConfig(hidden_size=64, num_attention_heads=2, num_key_value_heads=1,
                                          intermediate_size=64, num_hidden_layers=1, vocab_size=128)
    model = ref.qwen2.BabyQwen2ForCausalLM(cfg).float().eval()
    sched = ref.scheduler.Scheduler(model, tokenizer=None)
    rnd = random.Random(7)
    ops, live, next_id = [], {}, [0]

    def new_request(n_prompt, prefix):
        rid = next_id[0]
        next_id[0] += 1
        req = ref.request.Request(input_ids=torch.tensor([rnd.randrange(128) for _ in range(n_prompt)]),
                                  output_ids=[], max_new_tokens=8, eos_id=-1, temperature=1.0, top_k=0, top_p=1.0,
                                  do_sample=False, repetition_penalty=1.0)
        req.prefix_indices = torch.tensor(prefix, dtype=torch.int32)
        req.num_cached_tokens = len(prefix)
        live[rid] = req
        return rid

    def snapshot():
        return {"rows": {str(rid): sched.req_to_token_pool.req_to_token[r.req_pool_idx].tolist()
                         for rid, r in live.items() if r.req_pool_idx is not None},
                "free_slots": sched.token_to_kv_pool.free_slots.tolist(),
                "free_rows": list(sched.req_to_token_pool.free_slots)}

    def batch_fields(fb):
        return {"input_ids": fb.input_ids.tolist(), "req_pool_indices": fb.req_pool_indices.tolist(),
                "seq_lens": fb.seq_lens.tolist(), "out_cache_loc": fb.out_cache_loc.tolist(),
                "positions": fb.positions.tolist(),
                "prefix_lens": None if fb.prefix_lens is None else fb.prefix_lens.tolist(),
                "extended_lens": None if fb.extended_lens is None else fb.extended_lens.tolist()}

    def prefill(specs):
        rids = []
        for n_prompt, n_prefix in specs:
            prefix = sched.token_to_kv_pool.alloc(n_prefix).tolist() if n_prefix else []
            if n_prefix:
                ops.append({"op": "alloc_prefix", "n": n_prefix, "got": prefix})
            rids.append(new_request(n_prompt, prefix))
        reqs = [live[r] for r in rids]
        fb = sched._prepare_prefill_batch(reqs)
        ops.append({"op": "prefill", "rids": rids,
                    "requests": [{"input_ids": r.input_ids.tolist(), "prefix_indices": r.prefix_indices.tolist()} for r in reqs],
                    "batch": batch_fields(fb), "req_pool_idx": [r.req_pool_idx for r in reqs],
                    "seq_len": [r.seq_len for r in reqs], "after": snapshot()})
        for r in reqs:
            r.output_ids.append(rnd.randrange(128))

    def decode(rids):
        reqs = [live[r] for r in rids]
        fb = sched._prepare_decode_batch(reqs)
        ops.append({"op": "decode", "rids": rids, "last_ids": [r.output_ids[-1] for r in reqs],
                    "batch": batch_fields(fb), "seq_len": [r.seq_len for r in reqs], "after": snapshot()})
        for r in reqs:
            r.output_ids.append(rnd.randrange(128))

    def finish(rid, keep_prefix):
        """Free a request the way the scheduler's retraction does (scheduler.py:333-337): the slots after
        `keep_prefix` go back to the FIFO straight from the page-table row, then the row."""
        r = live.pop(rid)
        view = sched.req_to_token_pool.req_to_token[r.req_pool_idx, keep_prefix:r.seq_len]
        sched.token_to_kv_pool.free(view)
        sched.req_to_token_pool.free(r.req_pool_idx)
        ops.append({"op": "finish", "rid": rid, "keep_prefix": keep_prefix, "seq_len": r.seq_len, "after": snapshot()})

    prefill([(37, 0), (5, 0), (64, 16)])
    for _ in range(3):
        decode([0, 1, 2])
    finish(1, 0)
    prefill([(70, 33), (9, 8)])
    for _ in range(4):
        decode([0, 2, 3, 4])
    finish(0, 0)
    finish(3, 33)
    prefill([(80 - 9, 0), (1, 0), (30, 29)])     # a full-width row (C = 80 after 8 decodes) and a 1-token prompt
    for _ in range(8):
        decode([2, 4, 5, 6, 7])
    finish(5, 0)
    finish(2, 16)
    prefill([(60, 0), (60, 10), (60, 0)])         # drives the FIFO around its ring
    decode([4, 6, 7, 8, 9, 10])
    out = {"total_tokens": TOTAL, "max_requests": REQS, "C": TOTAL // REQS, "ops": ops}
    with open(os.path.join(HERE, "batch_prep_trace.json"), "w") as f:
        json.dump(out, f)
    print(len(ops), "ops;", sum(1 for o in ops if o["op"] == "prefill"), "prefills,",
          sum(1 for o in ops if o["op"] == "decode"), "decodes")


if __name__ == "__main__":
    main()
