"""TEST INFRASTRUCTURE ONLY — the same three-module surface as the drop-ins, computed by the oracle with
plain torch ops on whatever device the tensors live on (CPU here, or the GPU for same-dtype comparisons).

Pools are the drop-in pools themselves (their allocation logic is host code and must be exercised); only the
batch factories and the backend are replaced, because the drop-in ones launch CUDA kernels."""
from __future__ import annotations

import os
import sys
from dataclasses import dataclass
from typing import Optional

import torch

from baby_sglang_b200 import DROPIN_DIR
from oracle import paged_attention_oracle as O

if DROPIN_DIR not in sys.path:
    sys.path.insert(0, DROPIN_DIR)
import memory_pool as dropin_memory_pool  # noqa: E402

from tests.harness_model import Stack  # noqa: E402


@dataclass
class OracleForwardBatch:
    batch_size: int
    input_ids: torch.Tensor
    req_pool_indices: torch.Tensor
    seq_lens: torch.Tensor
    out_cache_loc: torch.Tensor
    positions: torch.Tensor
    req_to_token_pool: object
    token_to_kv_pool: object
    is_prefill: bool = True
    prefix_lens: Optional[torch.Tensor] = None
    extended_lens: Optional[torch.Tensor] = None

    @classmethod
    def create_prefill_batch(cls, input_ids, req_pool_indices, seq_lens, out_cache_loc, req_to_token_pool,
                             token_to_kv_pool, prefix_lens):
        md = O.build_metadata(req_to_token_pool.req_to_token.cpu(), req_pool_indices.cpu(), seq_lens.cpu(),
                              prefix_lens.cpu())
        dev = input_ids.device
        return cls(len(req_pool_indices), input_ids, req_pool_indices, seq_lens, out_cache_loc,
                   md.positions.to(dev), req_to_token_pool, token_to_kv_pool, True, prefix_lens,
                   md.extended_lens.to(dev))

    @classmethod
    def create_decode_batch(cls, input_ids, req_pool_indices, seq_lens, out_cache_loc, req_to_token_pool,
                            token_to_kv_pool):
        return cls(len(req_pool_indices), input_ids, req_pool_indices, seq_lens, out_cache_loc, seq_lens - 1,
                   req_to_token_pool, token_to_kv_pool, False)


def oracle_backend_forward(q, k, v, layer, fb, save_kv_cache=True):
    """fp32 accumulation, output in q's dtype — what SDPA does for half inputs."""
    pool = fb.token_to_kv_pool
    ks, vs = pool.k_buffer[layer.layer_id], pool.v_buffer[layer.layer_id]
    if q.dim() == 2:
        q = q.view(q.shape[0], -1, layer.head_dim)
    return O.backend_forward(q, k, v, ks, vs, fb.req_to_token_pool.req_to_token, fb.req_pool_indices, fb.seq_lens,
                             fb.out_cache_loc, fb.is_prefill, fb.prefix_lens, fb.extended_lens,
                             save_kv_cache=save_kv_cache, out_dtype=q.dtype)


def oracle_stack() -> Stack:
    return Stack(dropin_memory_pool.ReqToTokenPool, dropin_memory_pool.MHATokenToKVPool, OracleForwardBatch,
                 oracle_backend_forward)


def dropin_stack() -> Stack:
    """The product: drop-in pools, batch factories and backend (CUDA only)."""
    import attn_backend
    import forward_batch
    return Stack(dropin_memory_pool.ReqToTokenPool, dropin_memory_pool.MHATokenToKVPool,
                 forward_batch.SimplifiedForwardBatch, attn_backend.SimpleAttentionBackend.forward)
