"""Parity at the BASELINE configurations' REAL sizes (VERDICT r1 "missing" item 4).

The small parity cases cannot see what only shows at full load: slab element offsets past 2^31, 32 K-token rows,
hundreds of sequences per CTA, 68 KV tiles of lazy-rescale state per extend row, the programmatic-launch chain
K1 -> K2 -> K2c -> consumer under back-to-back launches.  The oracle cannot recompute 20 GB of attention in
seconds, so each case checks a seeded SAMPLE of requests (always the longest and the shortest) against the oracle's
own arithmetic (oracle/sampled_check.py), plus size-independent properties: bit-exact KV rows after the store,
split-count invariance, run-to-run bit identity.  Tolerance: bf16 output vs fp32 oracle <= 2e-2 (north_star).
"""
import pytest
import torch

from oracle import sampled_check as SC

pytestmark = pytest.mark.gpu
TOL = 2e-2


@pytest.fixture(scope="module")
def ops():
    from baby_sglang_b200 import ops as _ops
    return _ops


def _big_case(Hq, Hkv, D, seq, dtype=torch.bfloat16, seed=0, pad_slots=4096):
    """Device-built paged case: fragmented page table (seeded permutation of the slots), N(0,1) slabs."""
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(seed)
    seq = seq.to(dev)
    B = int(seq.shape[0])
    S = int(seq.sum().item()) + pad_slots
    C = int(seq.max().item()) + 8
    k_slab = torch.empty(S + 1, Hkv, D, device=dev, dtype=dtype).normal_(generator=g)
    v_slab = torch.empty(S + 1, Hkv, D, device=dev, dtype=dtype).normal_(generator=g)
    perm = (torch.randperm(S, generator=g, device=dev) + 1).to(torch.int32)
    ends = torch.cumsum(seq, 0)
    starts = ends - seq
    r2t = torch.zeros(B, C, dtype=torch.int32, device=dev)
    chunk = max(1, (1 << 26) // C)            # build the table in row blocks (bounded temporaries)
    col = torch.arange(C, device=dev)[None, :]
    for b0 in range(0, B, chunk):
        b1 = min(B, b0 + chunk)
        src = (starts[b0:b1, None] + col).clamp_(max=S - 1)
        blk = perm[src]
        blk[col >= seq[b0:b1, None]] = 0
        r2t[b0:b1] = blk
    rpi = torch.arange(B, dtype=torch.int64, device=dev)
    return dict(B=B, S=S, C=C, Hq=Hq, Hkv=Hkv, D=D, seq=seq, k_slab=k_slab, v_slab=v_slab, r2t=r2t, rpi=rpi, gen=g)


def _decode(ops, c, q, k_new=None, v_new=None, n_ctas=None):
    if k_new is not None:
        loc = c["r2t"][c["rpi"], c["seq"] - 1].contiguous()
        ops.kv_store(c["k_slab"], c["v_slab"], loc, k_new, v_new)
    md = ops.build_decode_metadata(c["r2t"], c["rpi"], c["seq"], c["Hkv"], n_ctas=n_ctas)
    return ops.paged_attention(q, c["k_slab"], c["v_slab"], c["r2t"], md.plan, c["D"] ** -0.5)


def _check_decode(ops, c, n_sample=16, n_ctas=None, store=True):
    dev, dt = c["k_slab"].device, c["k_slab"].dtype
    B, Hq, Hkv, D = c["B"], c["Hq"], c["Hkv"], c["D"]
    q = torch.randn(B, Hq, D, device=dev, dtype=dt, generator=c["gen"])
    k_new = v_new = None
    if store:
        k_new = torch.randn(B, Hkv, D, device=dev, dtype=dt, generator=c["gen"])
        fused = torch.randn(B, 3 * Hkv * D, device=dev, dtype=dt, generator=c["gen"])
        v_new = fused.view(B, 3 * Hkv, D)[:, Hkv:2 * Hkv]       # strided view, like the fused QKV output
    out = _decode(ops, c, q, k_new, v_new, n_ctas)
    torch.cuda.synchronize()
    if store:   # K1 bit-exact at full size (incl. slab offsets past 2^31 elements)
        loc = c["r2t"][c["rpi"], c["seq"] - 1].long()
        assert torch.equal(c["k_slab"][loc], k_new) and torch.equal(c["v_slab"][loc], v_new)
    sample = SC.pick_requests(c["seq"], n_sample, seed=1)
    res = SC.decode_rows(q, c["k_slab"], c["v_slab"], c["r2t"], c["rpi"], c["seq"], out, sample)
    assert res["max_abs_err"] <= TOL, res
    assert torch.isfinite(out.float()).all()
    return q, out, res


def test_ns_full_size_decode(ops):
    """BASELINE metric config: Qwen2.5-7B heads, B=256, every context 4096 (2.1 GB of KV)."""
    c = _big_case(28, 4, 128, torch.full((256,), 4096, dtype=torch.int64))
    q, out, res = _check_decode(ops, c)
    again = _decode(ops, c, q)
    assert torch.equal(out, again), "run-to-run bit identity"
    print("[fullsize] NS", res)


def test_c2_full_size_decode(ops):
    """Config 2: Qwen2.5-1.5B heads (12q/2kv), B=256, ctx 2048."""
    c = _big_case(12, 2, 128, torch.full((256,), 2048, dtype=torch.int64), seed=2)
    print("[fullsize] C2", _check_decode(ops, c)[2])


def test_c4_full_size_split_kv_decode_32k_rows_and_big_offsets(ops):
    """Config 4: B=512, contexts U[4096, 32768] (~9.4 M tokens: each slab ~9.7 GB, element offsets past 2^31),
    rows split over many CTAs; split-count invariance at full size."""
    g = torch.Generator().manual_seed(4)
    seq = torch.randint(4096, 32769, (512,), generator=g)
    seq[0], seq[1] = 32768, 4096
    c = _big_case(28, 4, 128, seq, seed=4)
    assert c["k_slab"].numel() > 2 ** 31
    q, out, res = _check_decode(ops, c)
    print("[fullsize] C4", res)
    for n_ctas in (37, 147):
        other = _decode(ops, c, q, n_ctas=n_ctas)
        assert (other.float() - out.float()).abs().max().item() <= 2 ** -6
    del c
    torch.cuda.empty_cache()


def test_c5_full_size_ragged_decode_and_tp8_shard(ops):
    """Config 5: Qwen2.5-32B heads (40q/8kv), B=1024, contexts U[512, 8192]; and one rank's shard under KV-head
    TP8 (5q/1kv: 256-byte rows, the warp-per-tile producer path)."""
    g = torch.Generator().manual_seed(5)
    seq = torch.randint(512, 8193, (1024,), generator=g)
    c = _big_case(40, 8, 128, seq, seed=5)
    print("[fullsize] C5 tp1", _check_decode(ops, c)[2])
    del c
    torch.cuda.empty_cache()
    c = _big_case(5, 1, 128, seq, seed=6)
    print("[fullsize] C5 tp8 shard", _check_decode(ops, c)[2])
    del c
    torch.cuda.empty_cache()


@pytest.mark.parametrize("Hq,Hkv", [(40, 8), (5, 1), (12, 2)])
def test_many_sequences_per_cta(ops, Hq, Hkv):
    """B=6000 short ragged rows: 81+ sequences per CTA at 148 CTAs, i.e. more than one 64-item pass per CTA."""
    g = torch.Generator().manual_seed(Hq)
    seq = torch.randint(1, 65, (6000,), generator=g)
    c = _big_case(Hq, Hkv, 128, seq, seed=Hq, pad_slots=64)
    q, out, res = _check_decode(ops, c, n_sample=48)
    one = _decode(ops, c, q, n_ctas=1)                       # a single CTA walks all 6000+ sequences
    assert (one.float() - out.float()).abs().max().item() <= 2 ** -6
    print("[fullsize] many-seqs", Hq, Hkv, res)


def test_c3_full_size_extend(ops):
    """Config 3: Qwen2.5-7B heads, B=16, prefix U[0, 8192] (0 and 8192 forced), 512 new tokens each: up to 68 KV
    tiles per q chunk.  K1 stores the new tokens first (store-before-attend), K3 attends; sampled rows vs oracle,
    K3 vs row-wise K2 on one request, run-to-run bit identity."""
    Hq, Hkv, D, B, ext = 28, 4, 128, 16, 512
    g = torch.Generator().manual_seed(0)
    prefix = torch.randint(0, 8193, (B,), generator=g)
    prefix[0], prefix[1] = 0, 8192
    seq = prefix + ext
    c = _big_case(Hq, Hkv, D, seq, seed=3)
    dev, dt = torch.device("cuda"), torch.bfloat16
    T = B * ext
    q = torch.randn(T, Hq, D, device=dev, dtype=dt, generator=c["gen"])
    k_new = torch.randn(T, Hkv, D, device=dev, dtype=dt, generator=c["gen"])
    v_new = torch.randn(T, Hkv, D, device=dev, dtype=dt, generator=c["gen"])
    pre_d, seq_d = prefix.to(dev), seq.to(dev)
    cols = torch.arange(ext, device=dev)[None, :] + pre_d[:, None]
    loc = torch.gather(c["r2t"], 1, cols).reshape(-1).contiguous()
    ops.kv_store(c["k_slab"], c["v_slab"], loc, k_new, v_new)
    md = ops.build_extend_metadata(c["r2t"], c["rpi"], seq_d, pre_d, T, Hkv, row_plan=False)
    out = ops.extend_attention_tc(q, c["k_slab"], c["v_slab"], c["r2t"], c["rpi"], seq_d, pre_d, md.qo_indptr,
                                  D ** -0.5, req_order=md.heavy_order)
    torch.cuda.synchronize()
    res = SC.extend_rows(q, c["k_slab"], c["v_slab"], c["r2t"], c["rpi"], seq_d, pre_d, out,
                         SC.pick_requests(seq, 6, seed=2), rows_per_range=8)
    assert res["max_abs_err"] <= TOL, res
    assert torch.isfinite(out.float()).all()
    again = ops.extend_attention_tc(q, c["k_slab"], c["v_slab"], c["r2t"], c["rpi"], seq_d, pre_d, md.qo_indptr,
                                    D ** -0.5, req_order=md.heavy_order)
    assert torch.equal(out, again)
    # the longest request again, row-wise on K2 (an independent kernel over the same page table)
    i = 1
    mdr = ops.build_extend_metadata(c["r2t"], c["rpi"][i:i + 1], seq_d[i:i + 1], pre_d[i:i + 1], ext, Hkv,
                                    row_plan=True)
    rw = ops.paged_attention(q[i * ext:(i + 1) * ext].contiguous(), c["k_slab"], c["v_slab"], c["r2t"], mdr.plan,
                             D ** -0.5)
    assert (rw.float() - out[i * ext:(i + 1) * ext].float()).abs().max().item() <= 2 ** -6
    print("[fullsize] C3", res)


@pytest.mark.parametrize("B,n_ctas", [(148, None), (148, 1), (7, 1), (296, None)])
def test_unsplit_plans_followed_by_a_plain_consumer(ops, B, n_ctas):
    """with no sequence split across a CTA boundary (B*nhg == n_eff with equal lengths, or n_eff == 1)
    no thread of the merge kernel used to execute griddepcontrol.wait, so the merge grid -- and with it the next,
    normally launched kernel on the stream -- could run ahead of the attention kernel.  Here the consumer is a
    plain copy of `out` launched right behind K2/K2c, 200 times back to back, against one synchronised result."""
    Hq, Hkv, D = 28, 4, 128                    # hpb = 4 -> one sequence per request
    c = _big_case(Hq, Hkv, D, torch.full((B,), 1024, dtype=torch.int64), seed=B, pad_slots=64)
    dev, dt = torch.device("cuda"), torch.bfloat16
    q = torch.randn(B, Hq, D, device=dev, dtype=dt, generator=c["gen"])
    md = ops.build_decode_metadata(c["r2t"], c["rpi"], c["seq"], Hkv, n_ctas=n_ctas)
    want = ops.paged_attention(q, c["k_slab"], c["v_slab"], c["r2t"], md.plan, D ** -0.5)
    torch.cuda.synchronize()
    want = want.clone()
    out = torch.empty_like(want)
    copies = []
    for _ in range(200):
        out.fill_(float("nan"))
        ops.paged_attention(q, c["k_slab"], c["v_slab"], c["r2t"], md.plan, D ** -0.5, out=out)
        copies.append(out.clone())             # normal launch, ordered after the merge grid only
    torch.cuda.synchronize()
    for i, cp in enumerate(copies):
        assert torch.equal(cp, want), f"iteration {i}: consumer saw an incomplete attention output"
