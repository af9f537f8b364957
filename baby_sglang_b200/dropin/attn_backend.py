"""Drop-in for the reference's `attn_backend` module (reference attn_backend.py:6-149).

`SimpleAttentionBackend.forward(q, k, v, attention_layer, forward_batch, save_kv_cache=True)` is
the one call the unchanged model makes (qwen2.py:315).  It keeps the reference's order — store
this step's K/V first, then attend, reading the new tokens back from the pool
(attn_backend.py:20-32) — but every per-request Python loop, `.item()` sync, gather,
`repeat_interleave` and SDPA call is replaced by sm_100a kernels:

    forward_decode  -> K1 kv_store, K2 paged split-KV GQA attention, K2c LSE merge
    forward_extend  -> K1 kv_store, K3 ragged extend attention
    init_forward_metadata -> K4 (normally already run by the forward_batch factories)

The three SGLang-style names above are additions; `forward` dispatches to them.
"""
from __future__ import annotations

import torch

from baby_sglang_b200 import ops
from forward_batch import SimplifiedForwardBatch


class SimpleAttentionBackend:

    @staticmethod
    def init_forward_metadata(forward_batch: SimplifiedForwardBatch) -> ops.BatchMetadata:
        """Per-batch device metadata, built once and cached on the batch.  The factories in
        forward_batch.py already attach it; a batch built through the bare dataclass constructor
        gets it here on first use."""
        md = forward_batch.attn_metadata
        if md is None:
            pool = forward_batch.token_to_kv_pool
            r2t = forward_batch.req_to_token_pool.req_to_token
            if forward_batch.is_prefill:
                n_tokens = int(forward_batch.out_cache_loc.shape[0])
                md = ops.build_extend_metadata(r2t, forward_batch.req_pool_indices, forward_batch.seq_lens,
                                               forward_batch.prefix_lens, n_tokens, pool.head_num)
            else:
                md = ops.build_decode_metadata(r2t, forward_batch.req_pool_indices, forward_batch.seq_lens,
                                               pool.head_num)
            forward_batch.attn_metadata = md
            forward_batch.qo_indptr, forward_batch.kv_indptr = md.qo_indptr, md.kv_indptr
        md.use_on_current_stream()     # overlap mode: built on the scheduler's stream, consumed on the worker's
        return md

    @staticmethod
    def forward(q, k, v, attention_layer, forward_batch: SimplifiedForwardBatch, save_kv_cache: bool = True):
        if forward_batch.is_prefill:
            return SimpleAttentionBackend.forward_extend(q, k, v, attention_layer, forward_batch, save_kv_cache)
        return SimpleAttentionBackend.forward_decode(q, k, v, attention_layer, forward_batch, save_kv_cache)

    # ------------------------------------------------------------------------------------
    @staticmethod
    def _prepare(q, k, v, layer, forward_batch, save_kv_cache):
        pool = forward_batch.token_to_kv_pool
        if save_kv_cache and k is not None and v is not None:
            pool.set_kv_buffer(layer.layer_id, forward_batch.out_cache_loc, k, v)
        head_dim = int(getattr(layer, "head_dim", pool.head_dim))
        if q.dim() == 2:
            q = q.view(q.shape[0], -1, head_dim)
        k_slab, v_slab = pool.get_kv_buffer(layer.layer_id)
        # The reference accepts `layer.scaling` but SDPA's default 1/sqrt(D) is what it applies
        # (attn_backend.py:18,94-101,141-143); qwen2.py:162 sets the same value.
        return q, k_slab, v_slab, head_dim ** -0.5

    @staticmethod
    def forward_decode(q, k, v, layer, forward_batch: SimplifiedForwardBatch, save_kv_cache: bool = True):
        """One new token per request over its paged KV (reference `_decode_attention`)."""
        q, k_slab, v_slab, scale = SimpleAttentionBackend._prepare(q, k, v, layer, forward_batch, save_kv_cache)
        if q.shape[0] == 0:
            return q.new_zeros((0, q.shape[1] * q.shape[2]))
        md = SimpleAttentionBackend.init_forward_metadata(forward_batch)
        return ops.paged_attention(q, k_slab, v_slab, forward_batch.req_to_token_pool.req_to_token, md.plan, scale)

    @staticmethod
    def forward_extend(q, k, v, layer, forward_batch: SimplifiedForwardBatch, save_kv_cache: bool = True):
        """Radix-cache prefix plus new tokens, ragged (reference `_prefill_attention`): request i's
        row r sees KV columns 0..prefix_i+r."""
        q, k_slab, v_slab, scale = SimpleAttentionBackend._prepare(q, k, v, layer, forward_batch, save_kv_cache)
        if q.shape[0] == 0:
            return q.new_zeros((0, q.shape[1] * q.shape[2]))   # attn_backend.py:108-109
        md = SimpleAttentionBackend.init_forward_metadata(forward_batch)
        return ops.extend_attention(q, k_slab, v_slab, forward_batch, md, scale)
