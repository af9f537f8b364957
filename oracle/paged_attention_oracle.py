"""TEST INFRASTRUCTURE ONLY — the parity oracle for the paged-attention hot path.

An independent restatement, in plain fp32 tensor arithmetic, of what the
reference (baby-llm/baby-sglang, pure Python/PyTorch) computes on the path

    SimpleAttentionBackend.forward        attn_backend.py:8-32
      MHATokenToKVPool.set_kv_buffer      memory_pool.py:152-171
      _prefill_attention                  attn_backend.py:34-110
      _decode_attention                   attn_backend.py:112-149
    SimplifiedForwardBatch factories      forward_batch.py:24-84

The arithmetic itself lives in a third-party dependency of the reference,
PyTorch (`requirements.txt:1`, `torch>=2.0.0`, not vendored; 2.11.0 here):
`F.scaled_dot_product_attention`, whose published semantics are
`softmax(Q K^T / sqrt(D) + bias) V` with the softmax and accumulation done in
fp32 for half inputs.  This file restates exactly that, with explicit loops over
requests and no call to SDPA, so a bug shared with the reference's call chain
would show up.

Pinning.  The reference ships no tests, golden vectors or fixtures for this
path, so the pin is the reference itself: `tests/golden/` holds
outputs of the reference's own modules run in the build container
(`tests/golden/make_golden.py` generated them by importing `/root/reference`),
and `tests/test_oracle_vs_golden.py` checks every function here against them.
`tests/test_oracle_vs_reference.py` repeats the check against the live
reference whenever a checkout is reachable.

Index/byte work (kv_store, metadata, plans) is bit-exact by construction;
floating-point work is fp32 and is compared with a tolerance stated in the tests.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional, Tuple

import torch

# --------------------------------------------------------------------------------------
# K1  kv_store   (memory_pool.py:152-171)
# --------------------------------------------------------------------------------------


def kv_store(k_slab: torch.Tensor, v_slab: torch.Tensor, loc: torch.Tensor,
             cache_k: torch.Tensor, cache_v: torch.Tensor) -> None:
    """slab[loc[t]] = row t, for K and V, after a round-to-nearest cast to the slab dtype.

    Follows memory_pool.py:159-171: `loc` is widened to int64 (:159-160), the sources
    are cast to the pool dtype when they differ (:163-165), then two index-puts (:170-171).
    Written here as an explicit per-token loop so duplicates resolve "last writer wins"
    deterministically (the reference leaves duplicates undefined; callers never produce
    them: slots come from a free list).
    """
    T = int(loc.shape[0])
    loc64 = loc.to(torch.int64).tolist()
    ck = cache_k.to(k_slab.dtype)
    cv = cache_v.to(v_slab.dtype)
    for t in range(T):
        k_slab[loc64[t]] = ck[t]
        v_slab[loc64[t]] = cv[t]


# --------------------------------------------------------------------------------------
# the attention arithmetic (what SDPA is documented to do), fp32
# --------------------------------------------------------------------------------------


def _gqa_attend(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, first_row_limit: int,
                acc_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """q [n, Hq, D], k/v [full, Hkv, D].  Row i of q sees columns 0..first_row_limit+i (inclusive).

    q head h reads kv head h // (Hq // Hkv)  — `repeat_interleave(G, dim=1)` at
    attn_backend.py:61-66 / :130-135.  Scale is 1/sqrt(D): the reference accepts
    `scaling` but never forwards it (attn_backend.py:18,94-101,141-143), so SDPA's
    default applies.
    """
    n, Hq, D = q.shape
    full, Hkv, _ = k.shape
    G = Hq // Hkv
    qf = q.to(acc_dtype)
    kf = k.to(acc_dtype)
    vf = v.to(acc_dtype)
    scale = 1.0 / math.sqrt(D)
    out = torch.empty((n, Hq, D), dtype=acc_dtype, device=q.device)
    cols = torch.arange(full, device=q.device)
    rows = torch.arange(n, device=q.device)
    # True where masked: col > prefix + row   (attn_backend.py:83-85)
    masked = cols[None, :] > (first_row_limit + rows)[:, None]
    for h in range(Hq):
        s = (qf[:, h, :] @ kf[:, h // G, :].T) * scale           # [n, full]
        s = s.masked_fill(masked, float("-inf"))
        m = s.max(dim=-1, keepdim=True).values
        p = torch.exp(s - m)
        p = p / p.sum(dim=-1, keepdim=True)
        out[:, h, :] = p @ vf[:, h // G, :]
    return out


def decode_attention(q: torch.Tensor, k_slab: torch.Tensor, v_slab: torch.Tensor,
                     req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                     seq_lens: torch.Tensor, out_dtype: Optional[torch.dtype] = None,
                     acc_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """attn_backend.py:112-149.  q [B, Hq, D] -> [B, Hq*D].

    `seq_lens[i]` already counts the token being decoded (scheduler.py:489-490); its K/V
    row was stored before this call (attn_backend.py:20-23), so the new token attends to
    itself through `req_to_token[r, seq-1]`.
    """
    B, Hq, D = q.shape
    out_dtype = out_dtype or q.dtype
    outs = []
    for i in range(B):
        n = int(seq_lens[i])
        r = int(req_pool_indices[i])
        idx = req_to_token[r, :n].to(torch.int64)                 # :122-125
        o = _gqa_attend(q[i:i + 1], k_slab[idx], v_slab[idx], first_row_limit=n - 1,
                        acc_dtype=acc_dtype)
        outs.append(o.reshape(1, Hq * D))
    if not outs:
        return q.new_zeros((0, Hq * D), dtype=out_dtype)
    return torch.cat(outs, dim=0).to(out_dtype)


def extend_attention(q: torch.Tensor, k_slab: torch.Tensor, v_slab: torch.Tensor,
                     req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                     seq_lens: torch.Tensor, prefix_lens: torch.Tensor,
                     extended_lens: torch.Tensor, out_dtype: Optional[torch.dtype] = None,
                     acc_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """attn_backend.py:34-110.  q [sum(ext), Hq, D] packed in request order -> [sum(ext), Hq*D].

    Request i owns q rows [off, off+ext_i); row j sees the whole prefix plus new tokens
    0..j (mask `col > prefix + row`, :83-85).  K/V for prefix AND new tokens are read back
    from the pool (:55-59).  ext_i == 0 contributes nothing (:44-46); an empty batch gives
    a [0, Hq*D] tensor (:108-109).
    """
    T, Hq, D = q.shape
    out_dtype = out_dtype or q.dtype
    outs = []
    off = 0
    for i in range(int(req_pool_indices.shape[0])):
        ext = int(extended_lens[i])
        if ext == 0:
            continue
        pre = int(prefix_lens[i])
        full = int(seq_lens[i])
        r = int(req_pool_indices[i])
        idx = req_to_token[r, :full].to(torch.int64)
        o = _gqa_attend(q[off:off + ext], k_slab[idx], v_slab[idx], first_row_limit=pre,
                        acc_dtype=acc_dtype)
        outs.append(o.reshape(ext, Hq * D))
        off += ext
    if not outs:
        return q.new_zeros((0, Hq * D), dtype=out_dtype)
    return torch.cat(outs, dim=0).to(out_dtype)


def backend_forward(q, k, v, k_slab, v_slab, req_to_token, req_pool_indices, seq_lens,
                    out_cache_loc, is_prefill: bool, prefix_lens=None, extended_lens=None,
                    save_kv_cache: bool = True, out_dtype=None, acc_dtype=torch.float32):
    """attn_backend.py:8-32: store first (:20-23), then attend (:25-32)."""
    if save_kv_cache and k is not None and v is not None:
        kv_store(k_slab, v_slab, out_cache_loc, k, v)
    if is_prefill:
        return extend_attention(q, k_slab, v_slab, req_to_token, req_pool_indices, seq_lens,
                                prefix_lens, extended_lens, out_dtype, acc_dtype)
    return decode_attention(q, k_slab, v_slab, req_to_token, req_pool_indices, seq_lens,
                            out_dtype, acc_dtype)


# --------------------------------------------------------------------------------------
# K4  batch metadata  (forward_batch.py:24-84 plus the device-built additions)
# --------------------------------------------------------------------------------------


@dataclass
class OracleMetadata:
    extended_lens: Optional[torch.Tensor]   # int64 [B]   (prefill only)
    positions: torch.Tensor                 # int64 [T]
    qo_indptr: torch.Tensor                 # int32 [B+1]
    kv_indptr: torch.Tensor                 # int32 [B+1]
    kv_indices: torch.Tensor                # int32 [sum(seq)]


def build_metadata(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                   seq_lens: torch.Tensor, prefix_lens: Optional[torch.Tensor]) -> OracleMetadata:
    """Decode (prefix_lens None): positions = seq_lens - 1 (forward_batch.py:72), one q row each.
    Prefill: extended = seq - prefix (:36-38); positions = prefix..prefix+ext-1 per request,
    concatenated (:40-45).  kv_indptr / kv_indices are the CSR flattening of
    `req_to_token[r, :seq]` the attention loops read (attn_backend.py:55-57,123-125).
    """
    B = int(req_pool_indices.shape[0])
    seq = [int(x) for x in seq_lens]
    if prefix_lens is None:
        ext_list = [1] * B
        pos = [s - 1 for s in seq]
        extended = None
    else:
        pre = [int(x) for x in prefix_lens]
        ext_list = [s - p for s, p in zip(seq, pre)]
        pos = []
        for p, e in zip(pre, ext_list):
            pos.extend(range(p, p + e))
        extended = torch.tensor(ext_list, dtype=torch.int64)
    qo = [0]
    kv = [0]
    for e, s in zip(ext_list, seq):
        qo.append(qo[-1] + e)
        kv.append(kv[-1] + s)
    rows = [req_to_token[int(req_pool_indices[i]), :seq[i]].to(torch.int32).cpu() for i in range(B)]
    kv_indices = torch.cat(rows) if rows else torch.zeros(0, dtype=torch.int32)
    return OracleMetadata(
        extended_lens=extended,
        positions=torch.tensor(pos, dtype=torch.int64),
        qo_indptr=torch.tensor(qo, dtype=torch.int32),
        kv_indptr=torch.tensor(kv, dtype=torch.int32),
        kv_indices=kv_indices,
    )


# --------------------------------------------------------------------------------------
# split-KV plan (no reference counterpart: the reference never splits; this restates the
# partition the CUDA decode kernel uses so the plan kernel can be checked bit-exactly)
# --------------------------------------------------------------------------------------


def split_plan(kv_lens, n_head_groups: int, tile_tokens: int, n_ctas: int):
    """Flatten (row, head_group) "sequences" into `tile_tokens`-token tiles and cut the tile
    space into `n_ctas` equal contiguous ranges.

    Returns (tile_indptr [NS+1], cta_tile_begin [n_ctas+1], cta_first_seq [n_ctas+1]).
    Sequence j = row * n_head_groups + head_group owns tiles [tile_indptr[j], tile_indptr[j+1]).
    CTA c owns tiles [floor(c*NT/n_ctas), floor((c+1)*NT/n_ctas)); cta_first_seq[c] is the
    sequence holding its first tile (NS when the CTA has no tiles to its right).
    """
    tile_indptr = [0]
    for n in kv_lens:
        nt = (int(n) + tile_tokens - 1) // tile_tokens
        for _ in range(n_head_groups):
            tile_indptr.append(tile_indptr[-1] + nt)
    NT = tile_indptr[-1]
    NS = len(tile_indptr) - 1
    begins = [(c * NT) // n_ctas for c in range(n_ctas + 1)]
    first = []
    j = 0
    for c in range(n_ctas + 1):
        t = begins[c]
        while j < NS and tile_indptr[j + 1] <= t:
            j += 1
        first.append(j)
    return tile_indptr, begins, first


def merge_partials(o_parts: torch.Tensor, lse2_parts: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """LSE merge of split-KV partials.  o_parts [P, H, D] normalised partial outputs,
    lse2_parts [P, H] base-2 log-sum-exp of each part's scores.  Returns (o [H, D], lse2 [H])."""
    m = lse2_parts.max(dim=0).values
    w = torch.exp2(lse2_parts - m[None, :])
    den = w.sum(dim=0)
    o = (w[:, :, None] * o_parts).sum(dim=0) / den[:, None]
    return o, m + torch.log2(den)


# --------------------------------------------------------------------------------------
# K5  RoPE + split + store   (qwen2.py:249-268, 285-290; memory_pool.py:152-171)
# --------------------------------------------------------------------------------------


def _round_to(x: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """One elementwise torch op on a 16-bit tensor: compute in fp32, round to nearest even."""
    return x.to(dtype).to(torch.float32)


def rope(positions: torch.Tensor, x: torch.Tensor, inv_freq: torch.Tensor) -> torch.Tensor:
    """RotaryEmbedding.forward + _apply_rope for one tensor x [T, H, D] (qwen2.py:249-268).

    freqs = outer(positions.float(), inv_freq) in fp32 (:251); cos/sin in fp32 (:252-253) and then
    cast to x's dtype (:262-263); the rotation `cat([x1*cos - x2*sin, x1*sin + x2*cos])` (:266-268)
    is evaluated in x's dtype, i.e. every product and every sum is rounded to it.  Restated with
    explicit fp32 arithmetic and explicit roundings; returns x's dtype."""
    dt = x.dtype
    T, H, D = x.shape
    half = D // 2
    freqs = positions.to(torch.float32)[:, None] * inv_freq.to(torch.float32)[None, :]
    cos = _round_to(torch.cos(freqs), dt)[:, None, :]
    sin = _round_to(torch.sin(freqs), dt)[:, None, :]
    xf = x.to(torch.float32)
    x1, x2 = xf[..., :half], xf[..., half:]
    o1 = _round_to(_round_to(x1 * cos, dt) - _round_to(x2 * sin, dt), dt)
    o2 = _round_to(_round_to(x1 * sin, dt) + _round_to(x2 * cos, dt), dt)
    out = torch.empty_like(xf)
    out[..., :half] = o1
    out[..., half:] = o2
    return out.to(dt)


def rope_kv_store(qkv: torch.Tensor, positions: torch.Tensor, inv_freq: torch.Tensor, Hq: int, Hkv: int,
                  D: int, k_slab: torch.Tensor, v_slab: torch.Tensor, loc: torch.Tensor) -> torch.Tensor:
    """What the model does between the fused projection and attention: split (qwen2.py:285-290),
    rotate q and k (:185-191), store rotated k and v at `loc` (memory_pool.py:152-171).
    Returns the rotated q [T, Hq, D]; the slabs are updated in place."""
    T = qkv.shape[0]
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    q_rot = rope(positions, q.reshape(T, Hq, D), inv_freq)
    k_rot = rope(positions, k.reshape(T, Hkv, D), inv_freq)
    kv_store(k_slab, v_slab, loc, k_rot, v.reshape(T, Hkv, D))
    return q_rot


# --------------------------------------------------------------------------------------
# K6  batched greedy sampling   (scheduler.py:509-546, sample.py:19-60)
# --------------------------------------------------------------------------------------


def greedy_sample(logits: torch.Tensor, extended_lens, prev_ids_per_request, penalties) -> torch.Tensor:
    """Next token ids for requests decoded with do_sample=False.

    * `extended_lens` (list of ints) not None = prefill: request i samples from row
      cumsum(extended_lens)[i] - 1 of the packed logits (scheduler.py:513-521); None = row i.
    * repetition penalty (sample.py:19-36): skipped when penalty <= 1 or the history is empty; otherwise
      for every DISTINCT previous id `logit = logit / penalty if logit > 0 else logit * penalty`, computed
      in the logits' dtype.
    * argmax (sample.py:59-60): first index of the maximum.
    Explicit per-request, per-id loops over a private copy of each row."""
    B = len(prev_ids_per_request)
    if extended_lens is not None:
        rows, acc = [], 0
        for n in extended_lens:
            acc += int(n)
            rows.append(acc - 1)
    else:
        rows = list(range(B))
    out = []
    for i in range(B):
        row = logits[rows[i]].clone()
        pen = float(penalties[i])
        prev = [int(t) for t in prev_ids_per_request[i]]
        if pen > 1.0 and len(prev) > 0:
            for tid in sorted(set(prev)):
                val = row[tid]
                row[tid] = val / pen if float(val) > 0 else val * pen
        best, best_i = None, 0
        vals = row.to(torch.float32).tolist()
        for j, v in enumerate(vals):
            if v != v:                     # NaN counts as the maximum (torch.argmax)
                best_i = j
                break
            if best is None or v > best:
                best, best_i = v, j
        out.append(best_i)
    return torch.tensor(out, dtype=torch.int64)


# ------------------------------------------------------------------------------------------------------
# K7: page-table writes and the free-slot FIFO (integer work, bit-exact)
# ------------------------------------------------------------------------------------------------------
def prefill_page_table_rows(req_to_token: torch.Tensor, rows, prefix_indices, new_slots, seq_lens) -> None:
    """Reference scheduler.py:423-442 in place on `req_to_token` (int32 [R, C]): row rows[i] becomes
    [prefix_indices[i] | the request's share of new_slots, in batch order | 0 ...] over ALL C cells
    (the reference builds zeros([B, C]) and writes whole rows, so stale cells past seq_len are cleared)."""
    C = int(req_to_token.shape[1])
    off = 0
    for i, row in enumerate(rows):
        pre = [int(v) for v in prefix_indices[i]]
        n_new = int(seq_lens[i]) - len(pre)
        line = [0] * C
        for c, v in enumerate(pre):
            line[c] = v
        for c in range(n_new):
            line[len(pre) + c] = int(new_slots[off + c])
        off += n_new
        req_to_token[int(row)] = torch.tensor(line, dtype=torch.int32)


def decode_page_table_append(req_to_token: torch.Tensor, rows, seq_lens_before, slots) -> None:
    """Reference scheduler.py:485-489: the new token's slot goes to cell `seq_len` (the length BEFORE the
    increment) of the request's row."""
    for row, pos, slot in zip(rows, seq_lens_before, slots):
        req_to_token[int(row), int(pos)] = int(slot)


class SlotFifo:
    """Reference BaseTokenToKVPool free list (memory_pool.py:46-84) as a plain list: slots 1..size, `alloc`
    takes from the head (None when short), `free` appends at the tail in the given order, no sort, no de-dup."""

    def __init__(self, size: int):
        self.slots = list(range(1, size + 1))

    def available_size(self) -> int:
        return len(self.slots)

    def alloc(self, n: int):
        if n > len(self.slots):
            return None
        out, self.slots = self.slots[:n], self.slots[n:]
        return out

    def free(self, vals) -> None:
        self.slots.extend(int(v) for v in vals)
