"""BASELINE.json configs[0]: Qwen2.5-1.5B geometry (12 q / 2 kv heads, D 128), batch 8, 512-token
contexts -- the one configuration the reference itself can run (its torch attn_backend on CPU, fp32).

The reference's own `SimpleAttentionBackend.forward` on its own pools is executed LIVE wherever a checkout is
reachable (/root/reference in the build container, the staged baseline/_ref on the GPU box) on seeded inputs of
exactly that size: the prefill of the 8 x 512 prompts and the first decode step over the 512-token contexts.
  * not gpu: the oracle restatement against it (fp32, 2e-5; stored K/V rows bit-exact)
  * gpu:     the CUDA path through the drop-in modules (bf16) against it (2e-2, the north-star tolerance;
             stored K/V rows bit-exact against the reference's bf16-rounded rows)
"""
import sys

import pytest
import torch

from baseline import ref_loader
from oracle import paged_attention_oracle as O
from tests.helpers import make_paged_case

pytestmark = pytest.mark.skipif(ref_loader.find_reference() is None, reason="no reference checkout on this machine")
HQ, HKV, D, B, CTX = 12, 2, 128, 8, 512
PHASES = {"prefill": ([CTX] * B, [0] * B), "decode": ([CTX + 1] * B, None)}


def _reference_forward(c):
    """The unmodified reference modules on CPU fp32; returns (output, K slab, V slab) after the call."""
    ref = ref_loader.load_stack(names=("memory_pool", "forward_batch", "attn_backend"))
    kv = ref.memory_pool.MHATokenToKVPool(size=c.S, dtype=torch.float32, head_num=HKV, head_dim=D, layer_num=1,
                                          device="cpu")
    kv.k_buffer[0].copy_(c.k_slab.float())
    kv.v_buffer[0].copy_(c.v_slab.float())
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
        tp_q_head_num, tp_k_head_num, tp_v_head_num = HQ, HKV, HKV

    out = ref.attn_backend.SimpleAttentionBackend.forward(c.q.float(), c.k.float(), c.v.float(), Layer(), fb, True)
    return out, kv.k_buffer[0], kv.v_buffer[0]


def _case(phase):
    seq, pre = PHASES[phase]
    return make_paged_case(B=B, Hq=HQ, Hkv=HKV, D=D, seq_lens=seq, prefix_lens=pre, seed=20250 + len(phase))


@pytest.mark.parametrize("phase", sorted(PHASES))
def test_c1_oracle_matches_live_reference(phase):
    c = _case(phase)
    want, k_ref, v_ref = _reference_forward(c)
    ks, vs = c.k_slab.float().clone(), c.v_slab.float().clone()
    got = O.backend_forward(c.q.float(), c.k.float(), c.v.float(), ks, vs, c.req_to_token, c.req_pool_indices,
                            c.seq_lens, c.out_cache_loc, c.is_prefill, c.prefix_lens, c.extended_lens)
    assert (got - want).abs().max().item() <= 2e-5
    assert torch.equal(ks, k_ref) and torch.equal(vs, v_ref)


@pytest.mark.gpu
@pytest.mark.parametrize("phase", sorted(PHASES))
def test_c1_cuda_path_matches_live_reference(phase):
    from baby_sglang_b200 import DROPIN_DIR
    sys.path.insert(0, DROPIN_DIR)
    import attn_backend
    import forward_batch
    import memory_pool

    c = _case(phase)
    want, k_ref, v_ref = _reference_forward(c)
    dev = "cuda"
    kv = memory_pool.MHATokenToKVPool(size=c.S, dtype=torch.bfloat16, head_num=HKV, head_dim=D, layer_num=1, device=dev)
    kv.k_buffer[0].copy_(c.k_slab)
    kv.v_buffer[0].copy_(c.v_slab)
    r2t = memory_pool.ReqToTokenPool(size=c.req_to_token.shape[0], max_context_len=c.C, device=dev)
    r2t.req_to_token.copy_(c.req_to_token)
    qkv = c.qkv.to(dev)
    T = qkv.shape[0]
    q, k, v = qkv.split([HQ * D, HKV * D, HKV * D], dim=-1)
    q, k, v = q.contiguous().view(T, HQ, D), k.contiguous().view(T, HKV, D), v.view(T, HKV, D)   # v: strided view
    common = dict(input_ids=torch.zeros(T, dtype=torch.int64, device=dev), req_pool_indices=c.req_pool_indices.to(dev),
                  seq_lens=c.seq_lens.to(dev), out_cache_loc=c.out_cache_loc.to(dev), req_to_token_pool=r2t,
                  token_to_kv_pool=kv)
    if c.is_prefill:
        fb = forward_batch.SimplifiedForwardBatch.create_prefill_batch(prefix_lens=c.prefix_lens.to(dev), **common)
    else:
        fb = forward_batch.SimplifiedForwardBatch.create_decode_batch(**common)

    class Layer:
        layer_id, head_dim, scaling = 0, D, D ** -0.5

    out = attn_backend.SimpleAttentionBackend.forward(q, k, v, Layer(), fb, True)
    torch.cuda.synchronize()
    assert out.shape == want.shape and out.dtype == torch.bfloat16
    err = (out.float().cpu() - want).abs().max().item()
    assert err <= 2e-2, f"C1 {phase}: max abs err {err} vs the reference's fp32 CPU output"
    assert torch.equal(kv.k_buffer[0].cpu(), k_ref.to(torch.bfloat16))
    assert torch.equal(kv.v_buffer[0].cpu(), v_ref.to(torch.bfloat16))
