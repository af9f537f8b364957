"""Drop-in for the reference's `forward_batch` module (reference forward_batch.py:8-84).

Same dataclass name, fields, dtypes and factory keywords, so scheduler.py:444-452,500-507 and
overlap_worker.py:72-130 run unchanged.  What changes is WHERE the per-batch bookkeeping happens:
the reference loops over requests on the host with two `.item()` syncs each
(forward_batch.py:40-45); here one K4 launch builds positions, extended_lens, qo_indptr,
kv_indptr and the split-KV plan on the device, once per batch, and the attention backend reuses
them for every layer.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

from baby_sglang_b200 import ops
from memory_pool import MHATokenToKVPool, ReqToTokenPool


@dataclass
class SimplifiedForwardBatch:
    batch_size: int
    input_ids: torch.Tensor          # int64 [T]
    req_pool_indices: torch.Tensor   # int64 [B]
    seq_lens: torch.Tensor           # int64 [B]  (decode: already counts the token being decoded)
    out_cache_loc: torch.Tensor      # int32 [T]
    positions: torch.Tensor          # int64 [T]
    req_to_token_pool: ReqToTokenPool
    token_to_kv_pool: MHATokenToKVPool
    is_prefill: bool = True

    # prefix-cache metadata (prefill only)
    prefix_lens: Optional[torch.Tensor] = None
    extended_lens: Optional[torch.Tensor] = None

    # device-built additions (None until init_forward_metadata ran)
    qo_indptr: Optional[torch.Tensor] = None     # int32 [B+1]
    kv_indptr: Optional[torch.Tensor] = None     # int32 [B+1]
    kv_indices: Optional[torch.Tensor] = None    # int32 [sum(seq)], only built on request
    attn_metadata: Optional[ops.BatchMetadata] = None

    @classmethod
    def create_prefill_batch(cls, input_ids: torch.Tensor, req_pool_indices: torch.Tensor,
                             seq_lens: torch.Tensor, out_cache_loc: torch.Tensor,
                             req_to_token_pool: ReqToTokenPool, token_to_kv_pool: MHATokenToKVPool,
                             prefix_lens: torch.Tensor) -> "SimplifiedForwardBatch":
        md = ops.build_extend_metadata(
            req_to_token_pool.req_to_token, req_pool_indices, seq_lens, prefix_lens,
            n_tokens=int(input_ids.shape[0]), n_kv_heads=token_to_kv_pool.head_num)
        return cls(
            batch_size=len(req_pool_indices), input_ids=input_ids, req_pool_indices=req_pool_indices,
            seq_lens=seq_lens, out_cache_loc=out_cache_loc, positions=md.positions,
            req_to_token_pool=req_to_token_pool, token_to_kv_pool=token_to_kv_pool, is_prefill=True,
            prefix_lens=prefix_lens, extended_lens=md.extended_lens,
            qo_indptr=md.qo_indptr, kv_indptr=md.kv_indptr, attn_metadata=md)

    @classmethod
    def create_decode_batch(cls, input_ids: torch.Tensor, req_pool_indices: torch.Tensor,
                            seq_lens: torch.Tensor, out_cache_loc: torch.Tensor,
                            req_to_token_pool: ReqToTokenPool,
                            token_to_kv_pool: MHATokenToKVPool) -> "SimplifiedForwardBatch":
        md = ops.build_decode_metadata(
            req_to_token_pool.req_to_token, req_pool_indices, seq_lens,
            n_kv_heads=token_to_kv_pool.head_num)
        return cls(
            batch_size=len(req_pool_indices), input_ids=input_ids, req_pool_indices=req_pool_indices,
            seq_lens=seq_lens, out_cache_loc=out_cache_loc, positions=md.positions,
            req_to_token_pool=req_to_token_pool, token_to_kv_pool=token_to_kv_pool, is_prefill=False,
            qo_indptr=md.qo_indptr, kv_indptr=md.kv_indptr, attn_metadata=md)

    def build_kv_indices(self) -> torch.Tensor:
        """Flattened CSR copy of the page-table rows (`kv_indices`), built lazily: the kernels read
        the page table in place, so this exists for callers that want the SGLang-style pair."""
        if self.kv_indices is None:
            total = int(self.seq_lens.sum().item())
            self.kv_indices = ops.gather_kv_indices(self.req_to_token_pool.req_to_token, self.req_pool_indices,
                                                    self.kv_indptr, total)
        return self.kv_indices
