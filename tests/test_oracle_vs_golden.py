"""Pin the oracle: every function in oracle/paged_attention_oracle.py against the golden
fixtures the UNMODIFIED reference produced (tests/golden/make_golden.py)."""
import pytest
import torch

from oracle import paged_attention_oracle as O
from tests.helpers import golden_names, load_golden, load_rope_golden, rope_golden_names, slab_checksum

# fp32 oracle vs fp32 reference: both are softmax(QK^T/sqrt(D))V in fp32; only summation order differs
FP32_TOL = 2e-5


@pytest.mark.parametrize("name", golden_names())
def test_backend_forward_matches_reference(name):
    g = load_golden(name)
    k_slab = g.k_slab0.float().clone()
    v_slab = g.v_slab0.float().clone()
    out = O.backend_forward(
        g.q.float(), g.k.float(), g.v.float(), k_slab, v_slab, g.req_to_token, g.req_pool_indices,
        g.seq_lens, g.out_cache_loc, bool(g.is_prefill), g.prefix_lens, g.extended_lens)
    assert out.shape == g.out.shape
    assert out.dtype == torch.float32
    err = (out - g.out).abs().max().item()
    assert err <= FP32_TOL, err
    # kv_store: bit-exact rows + whole-slab checksum (nothing else was touched)
    loc = g.out_cache_loc.long()
    assert torch.equal(k_slab[loc].to(torch.bfloat16).view(torch.int16), g.k_rows1.view(torch.int16))
    assert torch.equal(v_slab[loc].to(torch.bfloat16).view(torch.int16), g.v_rows1.view(torch.int16))
    assert slab_checksum(k_slab.to(torch.bfloat16)) == g.k_slab1_sum
    assert slab_checksum(v_slab.to(torch.bfloat16)) == g.v_slab1_sum


@pytest.mark.parametrize("name", golden_names())
def test_metadata_matches_reference(name):
    g = load_golden(name)
    md = O.build_metadata(g.req_to_token, g.req_pool_indices, g.seq_lens, g.prefix_lens)
    assert torch.equal(md.positions, g.positions)
    if g.is_prefill:
        assert torch.equal(md.extended_lens, g.extended_lens)
        assert md.qo_indptr[-1].item() == g.q.shape[0]
    # CSR flattening agrees with the page-table rows the reference reads
    for i in range(g.req_pool_indices.shape[0]):
        a, b = int(md.kv_indptr[i]), int(md.kv_indptr[i + 1])
        row = g.req_to_token[int(g.req_pool_indices[i]), : int(g.seq_lens[i])]
        assert torch.equal(md.kv_indices[a:b], row)


def test_kv_store_casts_round_to_nearest():
    torch.manual_seed(0)
    k_slab = torch.zeros(9, 2, 8, dtype=torch.bfloat16)
    v_slab = torch.zeros(9, 2, 8, dtype=torch.bfloat16)
    k = torch.randn(3, 2, 8)
    v = torch.randn(3, 2, 8)
    loc = torch.tensor([5, 1, 8], dtype=torch.int32)
    O.kv_store(k_slab, v_slab, loc, k, v)
    assert torch.equal(k_slab[loc.long()], k.to(torch.bfloat16))
    assert torch.equal(v_slab[loc.long()], v.to(torch.bfloat16))
    assert k_slab[0].abs().sum() == 0


def test_split_plan_partitions_every_tile_once():
    kv_lens = [1, 16, 17, 4096, 33, 0, 250]
    for n_ctas in (1, 3, 148):
        for nhg in (1, 2):
            tile_indptr, begins, first = O.split_plan(kv_lens, nhg, 16, n_ctas)
            NT = tile_indptr[-1]
            assert begins[0] == 0 and begins[-1] == NT
            assert all(b1 >= b0 for b0, b1 in zip(begins, begins[1:]))
            for c in range(n_ctas):
                if begins[c] < begins[c + 1]:
                    j = first[c]
                    assert tile_indptr[j] <= begins[c] < tile_indptr[j + 1]


def test_merge_partials_equals_unsplit():
    torch.manual_seed(1)
    H, D, N = 3, 16, 50
    s = torch.randn(H, N) * 3
    v = torch.randn(N, D)
    full = torch.softmax(s, -1) @ v
    parts_o, parts_l = [], []
    for a, b in ((0, 7), (7, 30), (30, 50)):
        ss = s[:, a:b]
        parts_o.append(torch.softmax(ss, -1) @ v[a:b])
        parts_l.append(torch.logsumexp(ss, -1) / torch.log(torch.tensor(2.0)))
    o, lse2 = O.merge_partials(torch.stack(parts_o), torch.stack(parts_l))
    assert (o - full).abs().max() < 1e-5
    assert (lse2 - torch.logsumexp(s, -1) / torch.log(torch.tensor(2.0))).abs().max() < 1e-5


@pytest.mark.parametrize("name", rope_golden_names())
def test_rope_matches_reference_bit_for_bit(name):
    """oracle.rope / rope_kv_store vs the reference's RotaryEmbedding run in the same 16-bit dtype."""
    g = load_rope_golden(name)
    T = g.qkv.shape[0]
    q, k, v = g.qkv.split([g.Hq * g.D, g.Hkv * g.D, g.Hkv * g.D], dim=-1)
    q_rot = O.rope(g.positions, q.reshape(T, g.Hq, g.D), g.inv_freq)
    k_rot = O.rope(g.positions, k.reshape(T, g.Hkv, g.D), g.inv_freq)
    assert torch.equal(q_rot.view(torch.int16), g.q_rot.view(torch.int16))
    assert torch.equal(k_rot.view(torch.int16), g.k_rot.view(torch.int16))
    # fused form: rotated k and plain v land in the slabs at loc, q comes back rotated
    S = T + 5
    k_slab = torch.zeros(S + 1, g.Hkv, g.D, dtype=g.dtype)
    v_slab = torch.zeros(S + 1, g.Hkv, g.D, dtype=g.dtype)
    loc = (torch.randperm(S, generator=torch.Generator().manual_seed(1))[:T] + 1).to(torch.int32)
    q2 = O.rope_kv_store(g.qkv, g.positions, g.inv_freq, g.Hq, g.Hkv, g.D, k_slab, v_slab, loc)
    assert torch.equal(q2.view(torch.int16), g.q_rot.view(torch.int16))
    assert torch.equal(k_slab[loc.long()].view(torch.int16), g.k_rot.view(torch.int16))
    assert torch.equal(v_slab[loc.long()], v.reshape(T, g.Hkv, g.D))
    assert k_slab[0].abs().sum() == 0


def test_greedy_sample_matches_reference_scheduler():
    """oracle.greedy_sample vs the ids the reference's Scheduler._sample_next_ids returned."""
    from tests.helpers import load_sample_golden
    for name, c in load_sample_golden().items():
        got = O.greedy_sample(c.logits, c.extended_lens if c.mode == "prefill" else None, c.prev_ids, c.penalties)
        assert torch.equal(got, c.next_ids), name
