"""Pin the oracle against the LIVE reference whenever a checkout is reachable ($BABYSGL_REF, /root/reference,
or baseline/_ref staged by scripts/stage_reference.py): the reference's own `SimpleAttentionBackend.forward`
(attn_backend.py:8-149) on its own pools (memory_pool.py) and batch factories (forward_batch.py), fp32 on CPU,
against `oracle.paged_attention_oracle.backend_forward` on the same seeded inputs.  The committed golden fixtures
(tests/test_oracle_vs_golden.py) hold the same comparison frozen; this one re-runs it on fresh, larger cases."""
import pytest
import torch

from baseline import ref_loader
from oracle import paged_attention_oracle as O
from tests.helpers import make_paged_case

pytestmark = pytest.mark.skipif(ref_loader.find_reference() is None, reason="no reference checkout on this machine")
FP32_TOL = 2e-5

CASES = [
    # (Hq, Hkv, D, seq_lens, prefix_lens or None, alias)
    (12, 2, 128, [1, 17, 200, 64], None, None),
    (28, 4, 128, [300, 5], None, None),
    (5, 1, 128, [90], None, None),
    (28, 4, 128, [120, 64, 33, 20], [100, 0, 33, 19], None),
    (10, 2, 128, [96, 100, 70], [64, 64, 40], 0),
    (14, 2, 64, [77, 200], [70, 100], None),
]


@pytest.mark.parametrize("Hq,Hkv,D,seq_lens,prefix_lens,alias", CASES)
def test_oracle_matches_live_reference(Hq, Hkv, D, seq_lens, prefix_lens, alias):
    ref = ref_loader.load_stack(names=("memory_pool", "forward_batch", "attn_backend"))
    c = make_paged_case(B=len(seq_lens), Hq=Hq, Hkv=Hkv, D=D, seq_lens=seq_lens, prefix_lens=prefix_lens,
                        seed=Hq * 31 + len(seq_lens), dtype=torch.float32, alias_prefix_of=alias)
    kv = ref.memory_pool.MHATokenToKVPool(size=c.S, dtype=torch.float32, head_num=Hkv, head_dim=D, layer_num=1,
                                          device="cpu")
    kv.k_buffer[0].copy_(c.k_slab)
    kv.v_buffer[0].copy_(c.v_slab)
    r2t = ref.memory_pool.ReqToTokenPool(size=c.req_to_token.shape[0], max_context_len=c.C, device="cpu")
    r2t.req_to_token.copy_(c.req_to_token)
    common = dict(input_ids=torch.zeros(c.T, dtype=torch.int64), req_pool_indices=c.req_pool_indices,
                  seq_lens=c.seq_lens, out_cache_loc=c.out_cache_loc, req_to_token_pool=r2t, token_to_kv_pool=kv)
    if c.is_prefill:
        fb = ref.forward_batch.SimplifiedForwardBatch.create_prefill_batch(prefix_lens=c.prefix_lens, **common)
    else:
        fb = ref.forward_batch.SimplifiedForwardBatch.create_decode_batch(**common)

    class Layer:
        layer_id, head_dim, scaling = 0, D, D ** -0.5
        tp_q_head_num, tp_k_head_num, tp_v_head_num = Hq, Hkv, Hkv

    want = ref.attn_backend.SimpleAttentionBackend.forward(c.q, c.k, c.v, Layer(), fb, True)
    ks, vs = c.k_slab.clone(), c.v_slab.clone()
    got = O.backend_forward(c.q, c.k, c.v, ks, vs, c.req_to_token, c.req_pool_indices, c.seq_lens, c.out_cache_loc,
                            c.is_prefill, c.prefix_lens, c.extended_lens)
    assert got.shape == want.shape
    assert (got - want).abs().max().item() <= FP32_TOL
    assert torch.equal(ks, kv.k_buffer[0]) and torch.equal(vs, kv.v_buffer[0])          # kv_store bit-exact
    md = O.build_metadata(c.req_to_token, c.req_pool_indices, c.seq_lens, c.prefix_lens)
    assert torch.equal(md.positions, fb.positions)
    if c.is_prefill:
        assert torch.equal(md.extended_lens, fb.extended_lens)
