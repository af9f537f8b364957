"""Generate tests/golden/sample_greedy.json by running the UNMODIFIED reference's
`Scheduler._sample_next_ids` (scheduler.py:509-546, which calls sample.py:39-60) on seeded logits with
do_sample=False requests, with and without repetition penalty, in prefill and decode mode.
Build container only; the .json is committed.

    python tests/golden/make_sample_golden.py
"""
from __future__ import annotations

import json
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from baseline import ref_loader  # noqa: E402


def main():
    ref = ref_loader.load_stack()
    assert ref.scheduler is not None, getattr(ref, "scheduler_error", "")
    sample_next = ref.scheduler.Scheduler._sample_next_ids          # unbound: uses no scheduler state
    cases = {}
    for name, dtype, V, mode, seed in (("decode_bf16", torch.bfloat16, 1531, "decode", 1),
                                       ("prefill_bf16", torch.bfloat16, 1024, "prefill", 2),
                                       ("decode_f32", torch.float32, 777, "decode", 3),
                                       ("prefill_f16", torch.float16, 2048, "prefill", 4)):
        g = torch.Generator().manual_seed(seed)
        B = 6
        ext = [int(x) for x in torch.randint(1, 9, (B,), generator=g)] if mode == "prefill" else [1] * B
        T = sum(ext)
        logits = (torch.randn(T, V, generator=g) * 3).to(dtype)
        # make ties and penalised maxima likely: duplicate the row maximum at a later index
        for r in range(T):
            j = int(torch.argmax(logits[r]))
            logits[r, (j + 17) % V] = logits[r, j]
        reqs, prevs, pens = [], [], []
        ends = np.cumsum(ext) - 1
        for i in range(B):
            row = logits[ends[i]] if mode == "prefill" else logits[i]
            top = torch.topk(row.float(), 5).indices.tolist()
            prev = ([] if i == 0 else top[: 1 + i % 3] + [int(x) for x in torch.randint(0, V, (4,), generator=g)] + top[:1])
            pen = [1.0, 1.3, 2.0, 0.5, 1.0001, 5.0][i]
            n_in = ext[i] + 2
            reqs.append(SimpleNamespace(input_ids=list(range(n_in)), prefix_indices=list(range(n_in - ext[i])),
                                        constraint_state=None, do_sample=False, temperature=1.0, top_k=0, top_p=1.0,
                                        output_ids=prev, repetition_penalty=pen))
            prevs.append(prev)
            pens.append(pen)
        out = sample_next(None, logits.clone(), reqs, mode)
        bits = logits.view(torch.int16 if dtype != torch.float32 else torch.int32).numpy()
        cases[name] = dict(dtype=str(dtype).split(".")[-1], V=V, mode=mode, extended_lens=ext, prev_ids=prevs,
                           penalties=pens, logits_bits=bits.astype(np.int64).tolist(),
                           next_ids=[int(x) for x in out])
        print(name, cases[name]["next_ids"])
    with open(os.path.join(HERE, "sample_greedy.json"), "w") as f:
        json.dump(cases, f, separators=(",", ":"))


if __name__ == "__main__":
    main()
