"""TEST INFRASTRUCTURE ONLY — full-size parity by sampling.

At BASELINE.json's real sizes (2-20 GB of KV per layer) the oracle cannot recompute every row in seconds, so the
checker picks requests (always including the longest and the shortest) and, for extends, row ranges inside them,
gathers exactly the K/V rows those need from the device slabs, and runs the oracle's own arithmetic
(`paged_attention_oracle._gqa_attend`, fp32, CPU) on them.  Nothing here is timed or shipped; callers are
`tests/` and the `parity_check` block `bench.py` emits after each timed region.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import paged_attention_oracle as O


def pick_requests(seq_lens: torch.Tensor, n: int, seed: int = 0) -> List[int]:
    """`n` request indices: the longest, the shortest, the rest seeded-random (no duplicates)."""
    B = int(seq_lens.shape[0])
    s = seq_lens.detach().cpu()
    chosen = [int(torch.argmax(s)), int(torch.argmin(s))]
    g = torch.Generator().manual_seed(seed)
    for i in torch.randperm(B, generator=g).tolist():
        if len(chosen) >= min(n, B):
            break
        if i not in chosen:
            chosen.append(i)
    return sorted(set(chosen))


def _kv_rows(k_slab, v_slab, req_to_token, row: int, n_tokens: int) -> Tuple[torch.Tensor, torch.Tensor]:
    idx = req_to_token[row, :n_tokens].to(torch.int64)              # attn_backend.py:55-57,122-125
    return k_slab[idx].float().cpu(), v_slab[idx].float().cpu()


def decode_rows(q, k_slab, v_slab, req_to_token, req_pool_indices, seq_lens, out, requests: Sequence[int]) -> dict:
    """max |out - oracle| over the sampled requests of a decode batch.  q [B,Hq,D], out [B,Hq*D] (device)."""
    worst, where = 0.0, None
    rpi = req_pool_indices.detach().cpu().tolist()
    sl = seq_lens.detach().cpu().tolist()
    for i in requests:
        n = int(sl[i])
        K, V = _kv_rows(k_slab, v_slab, req_to_token, int(rpi[i]), n)
        ref = O._gqa_attend(q[i:i + 1].float().cpu(), K, V, first_row_limit=n - 1).reshape(-1)
        err = float((out[i].float().cpu().reshape(-1) - ref).abs().max())
        if err >= worst:
            worst, where = err, (int(i), n)
    return {"rows": len(list(requests)), "max_abs_err": worst, "worst_request_and_len": where}


def extend_rows(q, k_slab, v_slab, req_to_token, req_pool_indices, seq_lens, prefix_lens, out,
                requests: Sequence[int], rows_per_range: int = 8) -> dict:
    """max |out - oracle| over row ranges (first rows, middle rows, last rows) of the sampled requests of an extend
    batch.  q [T,Hq,D] packed in request order, out [T,Hq*D]."""
    rpi = req_pool_indices.detach().cpu().tolist()
    sl = seq_lens.detach().cpu().tolist()
    pl = prefix_lens.detach().cpu().tolist()
    ext = [s - p for s, p in zip(sl, pl)]
    offs = [0]
    for e in ext:
        offs.append(offs[-1] + e)
    worst, where, n_rows = 0.0, None, 0
    for i in requests:
        e = ext[i]
        if e == 0:
            continue
        starts = sorted({0, max(0, e // 2 - rows_per_range // 2), max(0, e - rows_per_range)})
        for r0 in starts:
            n = min(rows_per_range, e - r0)
            need = pl[i] + r0 + n                                   # row r0+n-1 sees columns 0..prefix+r0+n-1
            K, V = _kv_rows(k_slab, v_slab, req_to_token, int(rpi[i]), need)
            qs = q[offs[i] + r0: offs[i] + r0 + n].float().cpu()
            ref = O._gqa_attend(qs, K, V, first_row_limit=pl[i] + r0).reshape(n, -1)
            got = out[offs[i] + r0: offs[i] + r0 + n].float().cpu().reshape(n, -1)
            err = float((got - ref).abs().max())
            n_rows += n
            if err >= worst:
                worst, where = err, (int(i), int(pl[i]), int(r0))
    return {"rows": n_rows, "max_abs_err": worst, "worst_request_prefix_row": where}
