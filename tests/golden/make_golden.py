"""Generate the golden fixtures in this directory by running the UNMODIFIED reference
(/root/reference, imported through baseline/ref_loader.py) on seeded synthetic inputs.

    python tests/golden/make_golden.py            # rewrites tests/golden/*.npz

Run in the build container only (the reference does not travel to the GPU box); the
.npz files it writes are committed.  Everything is computed by the reference's own
classes: `MHATokenToKVPool`, `ReqToTokenPool`, `SimplifiedForwardBatch.create_*_batch`
and `SimpleAttentionBackend.forward`, on CPU in fp32 (the reference's CPU dtype policy,
model_loader.py:56-64).  Inputs are rounded to bf16 first so that the same bits can be fed
to the bf16 CUDA kernels; they are stored as uint16 bf16 bit patterns.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from baseline import ref_loader  # noqa: E402


def bf16_bits(x: torch.Tensor) -> np.ndarray:
    return x.to(torch.bfloat16).view(torch.int16).numpy().astype(np.uint16)


def rnd(gen, *shape):
    """N(0,1) rounded to bf16, returned as fp32."""
    return torch.randn(*shape, generator=gen).to(torch.bfloat16).to(torch.float32)


class _Layer:
    def __init__(self, layer_id, D):
        self.layer_id = layer_id
        self.scaling = D ** -0.5


def make_attention_case(ref, name, *, Hq, Hkv, D, seq_lens, prefix_lens, S, C, seed, n_layers=2,
                        layer_id=1, alias_prefix=False, strided_v=True):
    """One backend.forward call.  prefix_lens None -> decode; else extend (prefill)."""
    gen = torch.Generator().manual_seed(seed)
    B = len(seq_lens)
    is_prefill = prefix_lens is not None
    pool = ref.memory_pool.MHATokenToKVPool(size=S, dtype=torch.float32, head_num=Hkv, head_dim=D,
                                            layer_num=n_layers, device="cpu")
    r2t = ref.memory_pool.ReqToTokenPool(size=B + 3, max_context_len=C, device="cpu")
    # fill the slabs of the layer under test with finite data (torch.empty is garbage)
    k0 = rnd(gen, S + 1, Hkv, D)
    v0 = rnd(gen, S + 1, Hkv, D)
    pool.k_buffer[layer_id].copy_(k0)
    pool.v_buffer[layer_id].copy_(v0)
    # fragmented page table: a seeded permutation of the slots, handed out request by request
    perm = (torch.randperm(S, generator=gen) + 1).to(torch.int32)
    req_rows = torch.randperm(B + 3, generator=gen)[:B].to(torch.int64)
    cursor = 0
    out_loc = []
    for i in range(B):
        n = seq_lens[i]
        if alias_prefix and i > 0 and is_prefill:
            # share request 0's prefix slots (radix-cache style aliasing, radix_cache.py:182-199)
            p = min(prefix_lens[i], prefix_lens[0])
            r2t.req_to_token[req_rows[i], :p] = r2t.req_to_token[req_rows[0], :p]
            fresh = perm[cursor:cursor + n - p]
            r2t.req_to_token[req_rows[i], p:n] = fresh
            cursor += n - p
        else:
            r2t.req_to_token[req_rows[i], :n] = perm[cursor:cursor + n]
            cursor += n
        if is_prefill:
            out_loc.append(r2t.req_to_token[req_rows[i], prefix_lens[i]:n].clone())
        else:
            out_loc.append(r2t.req_to_token[req_rows[i], n - 1:n].clone())
    assert cursor <= S
    out_cache_loc = torch.cat(out_loc).to(torch.int32)
    T = int(out_cache_loc.shape[0])
    # q contiguous; k and v are views of a fused qkv buffer exactly as qwen2.py:285-290 produces
    qkv = rnd(gen, T, (Hq + 2 * Hkv) * D)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    q = q.contiguous().view(T, Hq, D)
    k = k.contiguous().view(T, Hkv, D)
    v = v.view(T, Hkv, D) if strided_v else v.contiguous().view(T, Hkv, D)
    seq_t = torch.tensor(seq_lens, dtype=torch.int64)
    ids = torch.zeros(T, dtype=torch.int64)
    if is_prefill:
        pre_t = torch.tensor(prefix_lens, dtype=torch.int64)
        fb = ref.forward_batch.SimplifiedForwardBatch.create_prefill_batch(
            input_ids=ids, req_pool_indices=req_rows, seq_lens=seq_t, out_cache_loc=out_cache_loc,
            req_to_token_pool=r2t, token_to_kv_pool=pool, prefix_lens=pre_t)
    else:
        fb = ref.forward_batch.SimplifiedForwardBatch.create_decode_batch(
            input_ids=ids, req_pool_indices=req_rows, seq_lens=seq_t, out_cache_loc=out_cache_loc,
            req_to_token_pool=r2t, token_to_kv_pool=pool)
    out = ref.attn_backend.SimpleAttentionBackend.forward(q, k, v, _Layer(layer_id, D), fb, True)
    rec = dict(
        Hq=Hq, Hkv=Hkv, D=D, S=S, C=C, layer_id=layer_id, is_prefill=int(is_prefill),
        k_slab0=bf16_bits(k0), v_slab0=bf16_bits(v0),
        req_to_token=r2t.req_to_token.numpy().copy(),
        req_pool_indices=req_rows.numpy(), seq_lens=seq_t.numpy(),
        out_cache_loc=out_cache_loc.numpy(), qkv=bf16_bits(qkv),
        positions=fb.positions.numpy(),
        out=out.numpy().astype(np.float32),
        # the slabs after the store: the rows the reference wrote, plus whole-slab checksums
        k_rows1=bf16_bits(pool.k_buffer[layer_id][out_cache_loc.long()]),
        v_rows1=bf16_bits(pool.v_buffer[layer_id][out_cache_loc.long()]),
        k_slab1_sum=np.uint64(bf16_bits(pool.k_buffer[layer_id]).astype(np.uint64).sum()),
        v_slab1_sum=np.uint64(bf16_bits(pool.v_buffer[layer_id]).astype(np.uint64).sum()),
    )
    if is_prefill:
        rec["prefix_lens"] = np.asarray(prefix_lens, dtype=np.int64)
        rec["extended_lens"] = fb.extended_lens.numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **rec)
    print(f"{name}: T={T} out={tuple(out.shape)} |out|max={out.abs().max():.3f}")


def make_pool_trace(ref, name, seed):
    """Allocator FIFO order (memory_pool.py:29-41,68-84): a seeded sequence of alloc/free
    operations against the reference pools; records every result."""
    rng = np.random.default_rng(seed)
    S, R = 97, 7
    kv = ref.memory_pool.BaseTokenToKVPool(size=S, dtype=torch.float16, device="cpu")
    rq = ref.memory_pool.ReqToTokenPool(size=R, max_context_len=8, device="cpu")
    ops, held_kv, held_rq = [], [], []
    for step in range(200):
        kind = rng.integers(0, 4)
        if kind == 0:
            n = int(rng.integers(0, 20))
            got = kv.alloc(n)
            ops.append(("kv_alloc", n, None if got is None else got.numpy().copy()))
            if got is not None and n:
                held_kv.append(got)
        elif kind == 1 and held_kv:
            j = int(rng.integers(0, len(held_kv)))
            t = held_kv.pop(j)
            # free a shuffled sub-slice first, then the rest (order must be preserved verbatim)
            cut = int(rng.integers(0, len(t) + 1))
            kv.free(t[:cut])
            kv.free(t[cut:])
            ops.append(("kv_free", cut, t.numpy().copy()))
        elif kind == 2:
            n = int(rng.integers(0, 4))
            got = rq.alloc(n)
            ops.append(("rq_alloc", n, None if got is None else np.asarray(got, dtype=np.int64)))
            if got:
                held_rq.extend(got)
        elif kind == 3 and held_rq:
            j = int(rng.integers(0, len(held_rq)))
            x = held_rq.pop(j)
            rq.free(x)
            ops.append(("rq_free", 1, np.asarray([x], dtype=np.int64)))
        else:
            continue
        ops.append(("state", kv.available_size(), kv.free_slots.numpy().copy()))
        ops.append(("rq_state", rq.available_size(), np.asarray(rq.free_slots, dtype=np.int64)))
    import json
    recs = [{"kind": kind, "n": int(n), "none": arr is None,
             "arr": [] if arr is None else [int(x) for x in np.asarray(arr).ravel()]} for kind, n, arr in ops]
    with open(os.path.join(HERE, name + ".json"), "w") as f:
        json.dump({"S": S, "R": R, "ops": recs}, f, separators=(",", ":"))
    print(f"{name}: {len(ops)} records")


def main():
    torch.set_num_threads(4)
    ref = ref_loader.load_stack(names=("memory_pool", "forward_batch", "attn_backend"))
    # decode, Qwen2.5-1.5B geometry (G=6), ragged, fragmented
    make_attention_case(ref, "decode_g6_d128", Hq=12, Hkv=2, D=128, seq_lens=[1, 17, 64, 130, 16, 33],
                        prefix_lens=None, S=300, C=160, seed=101)
    # decode, 7B geometry (G=7)
    make_attention_case(ref, "decode_g7_d128", Hq=28, Hkv=4, D=128, seq_lens=[95, 2, 48], prefix_lens=None,
                        S=160, C=128, seed=102)
    # decode, 0.5B geometry (D=64, G=7) and MHA (G=1)
    make_attention_case(ref, "decode_g7_d64", Hq=14, Hkv=2, D=64, seq_lens=[31, 32, 33, 200], prefix_lens=None,
                        S=320, C=256, seed=103)
    make_attention_case(ref, "decode_mha_d128", Hq=4, Hkv=4, D=128, seq_lens=[40, 7], prefix_lens=None,
                        S=64, C=64, seed=104)
    # extend: no prefix / ragged prefix hits / a request with nothing new / aliased prefix slots
    make_attention_case(ref, "extend_noprefix_g6", Hq=6, Hkv=1, D=128, seq_lens=[37, 5, 64],
                        prefix_lens=[0, 0, 0], S=128, C=80, seed=201)
    make_attention_case(ref, "extend_prefix_g7", Hq=7, Hkv=1, D=128, seq_lens=[70, 20, 33, 81],
                        prefix_lens=[50, 0, 33, 1], S=300, C=160, seed=202)
    make_attention_case(ref, "extend_alias_g5", Hq=10, Hkv=2, D=128, seq_lens=[48, 60, 41],
                        prefix_lens=[32, 32, 20], S=160, C=64, seed=203, alias_prefix=True)
    make_attention_case(ref, "extend_d64", Hq=14, Hkv=2, D=64, seq_lens=[150, 9], prefix_lens=[100, 8],
                        S=200, C=160, seed=204, strided_v=False)
    make_pool_trace(ref, "pool_fifo_trace", seed=7)


if __name__ == "__main__":
    main()
