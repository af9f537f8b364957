"""Shared test helpers: golden-fixture loading and synthetic paged-KV case generation."""
from __future__ import annotations

import glob
import os
from types import SimpleNamespace

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_names(prefix=""):
    """Attention fixtures (decode_*/extend_*); the RoPE fixtures are listed by rope_golden_names()."""
    names = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, prefix + "*.npz")))
    return [n for n in names if not n.startswith("rope_")]


def rope_golden_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "rope_*.npz")))


def load_rope_golden(name: str) -> SimpleNamespace:
    """Outputs of the reference's RotaryEmbedding (tests/golden/make_rope_golden.py)."""
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    dt = torch.bfloat16 if int(z["is_bf16"]) else torch.float16
    bits = lambda a: torch.from_numpy(a.astype(np.int16)).view(dt)   # noqa: E731
    g = SimpleNamespace(name=name, dtype=dt, Hq=int(z["Hq"]), Hkv=int(z["Hkv"]), D=int(z["D"]))
    g.positions = torch.from_numpy(z["positions"])
    g.inv_freq = torch.from_numpy(z["inv_freq"])
    g.qkv = bits(z["qkv"])
    g.q_rot = bits(z["q_rot"])
    g.k_rot = bits(z["k_rot"])
    return g


def from_bf16_bits(a: np.ndarray) -> torch.Tensor:
    """uint16 bf16 bit patterns -> torch.bfloat16 tensor (exact)."""
    return torch.from_numpy(a.astype(np.int16)).view(torch.bfloat16)


def load_golden(name: str) -> SimpleNamespace:
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    g = SimpleNamespace(name=name)
    for k in ("Hq", "Hkv", "D", "S", "C", "layer_id", "is_prefill"):
        setattr(g, k, int(z[k]))
    g.k_slab0 = from_bf16_bits(z["k_slab0"])
    g.v_slab0 = from_bf16_bits(z["v_slab0"])
    g.req_to_token = torch.from_numpy(z["req_to_token"])
    g.req_pool_indices = torch.from_numpy(z["req_pool_indices"])
    g.seq_lens = torch.from_numpy(z["seq_lens"])
    g.out_cache_loc = torch.from_numpy(z["out_cache_loc"])
    g.qkv = from_bf16_bits(z["qkv"])
    g.positions = torch.from_numpy(z["positions"])
    g.out = torch.from_numpy(z["out"])
    g.k_rows1 = from_bf16_bits(z["k_rows1"])
    g.v_rows1 = from_bf16_bits(z["v_rows1"])
    g.k_slab1_sum = int(z["k_slab1_sum"])
    g.v_slab1_sum = int(z["v_slab1_sum"])
    if g.is_prefill:
        g.prefix_lens = torch.from_numpy(z["prefix_lens"])
        g.extended_lens = torch.from_numpy(z["extended_lens"])
    else:
        g.prefix_lens = None
        g.extended_lens = None
    T = g.qkv.shape[0]
    q, k, v = g.qkv.split([g.Hq * g.D, g.Hkv * g.D, g.Hkv * g.D], dim=-1)
    g.q = q.contiguous().view(T, g.Hq, g.D)
    g.k = k.contiguous().view(T, g.Hkv, g.D)
    g.v = v.view(T, g.Hkv, g.D)           # strided view of the fused buffer, as qwen2.py:285-290 yields
    return g


def slab_checksum(slab: torch.Tensor) -> int:
    bits = slab.contiguous().view(torch.int16).cpu().numpy().astype(np.uint16).astype(np.uint64)
    return int(bits.sum())


def make_paged_case(*, B, Hq, Hkv, D, seq_lens, prefix_lens=None, seed=0, dtype=torch.bfloat16,
                    device="cpu", extra_slots=7, C=None, alias_prefix_of=None):
    """Seeded synthetic case: slabs filled with N(0,1), fragmented page table (random permutation of
    slots), fused-qkv q/k/v.  Mirrors tests/golden/make_golden.py but without the reference."""
    gen = torch.Generator().manual_seed(seed)
    seq_lens = [int(x) for x in seq_lens]
    is_prefill = prefix_lens is not None
    total = sum(seq_lens)
    S = total + extra_slots
    C = C or (max(seq_lens) + 5)
    k_slab = torch.randn(S + 1, Hkv, D, generator=gen).to(dtype)
    v_slab = torch.randn(S + 1, Hkv, D, generator=gen).to(dtype)
    perm = (torch.randperm(S, generator=gen) + 1).to(torch.int32)
    R = B + 2
    rows = torch.randperm(R, generator=gen)[:B].to(torch.int64)
    r2t = torch.zeros(R, C, dtype=torch.int32)
    cur = 0
    locs = []
    for i in range(B):
        n = seq_lens[i]
        if is_prefill and alias_prefix_of is not None and i != alias_prefix_of:
            p = min(prefix_lens[i], prefix_lens[alias_prefix_of])
            r2t[rows[i], :p] = r2t[rows[alias_prefix_of], :p]
            r2t[rows[i], p:n] = perm[cur:cur + n - p]
            cur += n - p
        else:
            r2t[rows[i], :n] = perm[cur:cur + n]
            cur += n
        lo = prefix_lens[i] if is_prefill else n - 1
        locs.append(r2t[rows[i], lo:n].clone())
    loc = torch.cat(locs).to(torch.int32) if locs else torch.zeros(0, dtype=torch.int32)
    T = int(loc.shape[0])
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=gen).to(dtype)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    c = SimpleNamespace(
        B=B, Hq=Hq, Hkv=Hkv, D=D, S=S, C=C, T=T, is_prefill=is_prefill,
        k_slab=k_slab, v_slab=v_slab, req_to_token=r2t, req_pool_indices=rows,
        seq_lens=torch.tensor(seq_lens, dtype=torch.int64),
        prefix_lens=torch.tensor(prefix_lens, dtype=torch.int64) if is_prefill else None,
        out_cache_loc=loc, qkv=qkv,
        q=q.contiguous().view(T, Hq, D), k=k.contiguous().view(T, Hkv, D), v=v.view(T, Hkv, D),
    )
    if is_prefill:
        c.extended_lens = c.seq_lens - c.prefix_lens
    else:
        c.extended_lens = None
    if device != "cpu":
        for name in ("k_slab", "v_slab", "req_to_token", "req_pool_indices", "seq_lens", "prefix_lens",
                     "extended_lens", "out_cache_loc", "qkv"):
            t = getattr(c, name)
            if t is not None:
                setattr(c, name, t.to(device))
        q, k, v = c.qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        c.q = q.contiguous().view(T, Hq, D)
        c.k = k.contiguous().view(T, Hkv, D)
        c.v = v.view(T, Hkv, D)
    return c


def load_sample_golden():
    """Outputs of the reference's Scheduler._sample_next_ids (tests/golden/make_sample_golden.py)."""
    import json
    with open(os.path.join(GOLDEN_DIR, "sample_greedy.json")) as f:
        raw = json.load(f)
    cases = {}
    for name, c in raw.items():
        dt = getattr(torch, c["dtype"])
        bits = np.asarray(c["logits_bits"], dtype=np.int64)
        if dt == torch.float32:
            logits = torch.from_numpy(bits.astype(np.int32)).view(torch.float32)
        else:
            logits = torch.from_numpy(bits.astype(np.int16)).view(dt)
        cases[name] = SimpleNamespace(name=name, dtype=dt, V=c["V"], mode=c["mode"], logits=logits,
                                      extended_lens=c["extended_lens"], prev_ids=c["prev_ids"],
                                      penalties=c["penalties"], next_ids=torch.tensor(c["next_ids"], dtype=torch.int64))
    return cases


def prev_ids_csr(prev_ids, device="cpu"):
    """Per-request histories -> (prev_ids int64, prev_indptr int32)."""
    flat = [t for p in prev_ids for t in p]
    indptr = np.concatenate([[0], np.cumsum([len(p) for p in prev_ids])]).astype(np.int32)
    return (torch.tensor(flat, dtype=torch.int64, device=device), torch.from_numpy(indptr).to(device))
