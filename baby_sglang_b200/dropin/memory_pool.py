"""Drop-in for the reference's `memory_pool` module (same public classes, ctor keywords and
attributes; reference memory_pool.py:7-171), backed by B200 kernels where the reference runs
torch index-puts.

    ReqToTokenPool      page table `req_to_token` int32 [max_reqs, max_context_len] + FIFO of rows
    BaseTokenToKVPool   FIFO of KV slots 1..size (slot 0 is the pad slot); a device ring on CUDA pools (K7)
    MHATokenToKVPool    per-layer K / V slabs [size+1, H, D]; set_kv_buffer -> K1 (bsgl_kv_store)

Allocation ORDER is part of the contract (KV-slot indexing must be bit-exact with the reference):
both allocators hand out from the head and take returns at the tail, verbatim, with no sorting or
de-duplication (memory_pool.py:29-41, 68-84).
"""
from __future__ import annotations

import os
from collections import deque
from typing import Deque, List, Optional, Tuple, Union

import torch

from baby_sglang_b200 import ops

_FP8 = tuple(getattr(torch, n) for n in ("float8_e5m2", "float8_e4m3fn") if hasattr(torch, n))


class ReqToTokenPool:
    """Request row -> token slot table.  `req_to_token` is read and written directly by the
    scheduler and the radix cache (scheduler.py:333-335,486-488; radix_cache.py:158-199)."""

    def __init__(self, size: int, max_context_len: int, device: str):
        self.size = size
        self.max_context_len = max_context_len
        self.device = device
        self.req_to_token = torch.zeros((size, max_context_len), dtype=torch.int32, device=device)
        self.free_slots: Deque[int] = deque(range(size))

    def write(self, indices, values) -> None:
        self.req_to_token[indices] = values

    def available_size(self) -> int:
        return len(self.free_slots)

    def alloc(self, need_size: int) -> Optional[List[int]]:
        if need_size > len(self.free_slots):
            return None
        take = self.free_slots.popleft
        return [take() for _ in range(need_size)]

    def free(self, free_index: Union[int, List[int]]) -> None:
        if isinstance(free_index, int):
            self.free_slots.append(free_index)
        else:
            self.free_slots.extend(free_index)

    def clear(self) -> None:
        self.free_slots = deque(range(self.size))


class BaseTokenToKVPool:
    """FIFO of free KV slots, order identical to the reference's (head out, tail in; memory_pool.py:68-84).

    On a CUDA pool the queue is a ring in device memory (K7): `alloc` is one stream-ordered pop kernel instead of
    a CPU slice + H2D copy, `free` one push kernel that reads the freed slots where they are (usually a slice of
    a page-table row) instead of the reference's blocking `.cpu()` (memory_pool.py:77).  Head and count stay on
    the host: every size is known there.  On a CPU pool (host-logic tests) the ring is a CPU tensor.
    `free_slots` materialises the queue in order as a CPU int32 tensor, like the reference attribute."""

    def __init__(self, size: int, dtype: torch.dtype, device: str):
        self.size = size
        self.dtype = dtype
        self.store_dtype = torch.uint8 if dtype in _FP8 else dtype
        self.device = device
        self._dev_ring = torch.device(device).type == "cuda" and os.environ.get("BSGL_DEVICE_SLOT_RING", "1") != "0"
        self._ring: torch.Tensor = None
        self._head = 0
        self._count = 0
        self._mutations = 0          # bumped by every alloc / free / clear: invalidates the `free_slots` snapshot
        self._snapshot = None
        self.clear()

    # -- queue plumbing ------------------------------------------------------------------
    def _ordered(self) -> torch.Tensor:
        """The queue contents in FIFO order, on the ring's device."""
        if self._dev_ring:
            return ops.slot_ring_pop(self._ring, self._head, self._count)
        cap = int(self._ring.numel())
        end = self._head + self._count
        if end <= cap:
            return self._ring[self._head:end]
        return torch.cat((self._ring[self._head:], self._ring[: end - cap]))

    def _grow(self, need: int) -> None:
        cap = int(self._ring.numel())
        new_cap = max(2 * cap, self._count + need)
        fresh = torch.empty(new_cap, dtype=torch.int32, device=self._ring.device)
        fresh[: self._count] = self._ordered()
        self._ring, self._head = fresh, 0

    @property
    def free_slots(self) -> torch.Tensor:
        """The queue in FIFO order as a CPU int32 tensor (the reference's attribute; nothing in the reference's
        scheduler or radix cache reads it -- they use `available_size`, `alloc`, `free`).  Materialising it costs a
        kernel and a blocking D2H copy, so the copy is kept until the queue changes: a caller that polls it pays once
        per change, not once per access."""
        key = (self._head, self._count, id(self._ring), self._mutations)
        if self._snapshot is None or self._snapshot[0] != key:
            self._snapshot = (key, self._ordered().cpu())
        return self._snapshot[1]

    # -- reference API -------------------------------------------------------------------
    def available_size(self) -> int:
        return self._count

    def alloc(self, need_size: int) -> Optional[torch.Tensor]:
        if need_size > self._count:
            return None
        cap = int(self._ring.numel())
        end = self._head + need_size
        if self._dev_ring:
            picked = ops.slot_ring_pop(self._ring, self._head, need_size)
        elif end <= cap:
            picked = self._ring[self._head:end].clone()
        else:
            picked = torch.cat((self._ring[self._head:], self._ring[: end - cap]))
        self._head = end % cap
        self._count -= need_size
        self._mutations += 1
        return picked if self._dev_ring else picked.to(self.device, non_blocking=True)

    def free(self, free_index: torch.Tensor) -> None:
        n = int(free_index.numel())
        if n == 0:
            return
        if self._count + n > int(self._ring.numel()):
            self._grow(n)
        cap = int(self._ring.numel())
        tail = (self._head + self._count) % cap
        if self._dev_ring:
            vals = free_index.detach()
            if not vals.is_cuda:
                vals = vals.to(self._ring.device, non_blocking=True)
            if vals.dtype not in (torch.int32, torch.int64):
                vals = vals.to(torch.int64)
            ops.slot_ring_push(self._ring, tail, vals.reshape(-1) if vals.dim() != 1 else vals)
        else:
            vals = free_index.detach().to("cpu", torch.int32).reshape(-1)
            first = min(n, cap - tail)
            self._ring[tail:tail + first] = vals[:first]
            if first < n:
                self._ring[: n - first] = vals[first:]
        self._count += n
        self._mutations += 1

    def clear(self) -> None:
        # slots 1..size; slot 0 is reserved for padded tokens' dummy writes
        self._ring = torch.arange(1, self.size + 1, dtype=torch.int32, device=self.device if self._dev_ring else "cpu")
        self._head = 0
        self._count = self.size
        self._mutations += 1

    def get_key_buffer(self, layer_id: int) -> torch.Tensor:
        raise NotImplementedError()

    def get_value_buffer(self, layer_id: int) -> torch.Tensor:
        raise NotImplementedError()

    def get_kv_buffer(self, layer_id: int) -> Tuple[torch.Tensor, torch.Tensor]:
        raise NotImplementedError()

    def set_kv_buffer(self, layer_id: int, loc: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor) -> None:
        raise NotImplementedError()


class MHATokenToKVPool(BaseTokenToKVPool):
    """K and V slabs, token-major (`[slot, head, dim]`, page size 1) — the layout the paged
    attention kernel gathers from: one token's HPB adjacent heads are one contiguous burst.
    All layers live in one allocation per K / V; `k_buffer[l]` / `v_buffer[l]` are the per-layer
    views the reference exposes."""

    def __init__(self, size: int, dtype: torch.dtype, head_num: int, head_dim: int, layer_num: int, device: str):
        super().__init__(size, dtype, device)
        self.layer_num = layer_num
        self.head_num = head_num
        self.head_dim = head_dim
        self._create_buffers()

    def _create_buffers(self) -> None:
        shape = (self.layer_num, self.size + 1, self.head_num, self.head_dim)
        self._k_all = torch.empty(shape, dtype=self.store_dtype, device=self.device)
        self._v_all = torch.empty(shape, dtype=self.store_dtype, device=self.device)
        # the pad slot is never attended to, but keep it finite
        self._k_all[:, 0].zero_()
        self._v_all[:, 0].zero_()
        self.k_buffer = [self._k_all[i] for i in range(self.layer_num)]
        self.v_buffer = [self._v_all[i] for i in range(self.layer_num)]

    def _no_fp8(self, what: str) -> None:
        if self.store_dtype != self.dtype:
            raise NotImplementedError(f"float8 KV cache {what} is not supported (nor by the reference, memory_pool.py:141-168)")

    def get_key_buffer(self, layer_id: int) -> torch.Tensor:
        self._no_fp8("read")
        return self.k_buffer[layer_id]

    def get_value_buffer(self, layer_id: int) -> torch.Tensor:
        self._no_fp8("read")
        return self.v_buffer[layer_id]

    def get_kv_buffer(self, layer_id: int) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def set_kv_buffer(self, layer_id: int, loc: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor) -> None:
        self._no_fp8("write")
        ops.kv_store(self.k_buffer[layer_id], self.v_buffer[layer_id], loc, cache_k, cache_v)
