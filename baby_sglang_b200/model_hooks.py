"""Model-side hooks for the steps immediately before and after the attention path.

The reference model (qwen2.py) stays unchanged on disk; these functions mutate a loaded model
*instance*, the same mechanism `tp_runtime.shard_attention_heads` uses for KV-head TP.

`install_fused_rope_kv_store(model)`
    Replaces, per `Qwen2Attention.forward` (qwen2.py:179-197): the q/k/v split of the fused
    projection (:185, QKVLinear.forward :285-290), `rotary_emb` (:191, RotaryEmbedding.forward
    :249-268: cos/sin, four products, two `torch.cat`) and `set_kv_buffer` (memory_pool.py:152-171)
    by ONE launch of K5 (`bsgl_rope_kv_store`): rotated q comes back contiguous, rotated K and V go
    straight into the paged slabs.  Attention then runs with `save_kv_cache=False` on the drop-in
    backend, reading the rows K5 just stored (store-before-attend order, attn_backend.py:20-32).

There is no fallback: on CPU tensors the kernels raise.
"""
from __future__ import annotations

import types
from typing import List

import torch

from . import ops


def _attention_modules(model: torch.nn.Module):
    """Modules shaped like the reference's Qwen2Attention (qwen2.py:143-177)."""
    for m in model.modules():
        inner = getattr(getattr(m, "qkv_proj", None), "qkv_proj", None)
        if isinstance(inner, torch.nn.Linear) and isinstance(getattr(m, "o_proj", None), torch.nn.Linear) \
                and hasattr(m, "attn") and hasattr(m, "rotary_emb") and hasattr(m, "num_kv_heads"):
            yield m


def _inv_freq_f32(rotary: torch.nn.Module) -> torch.Tensor:
    """`inv_freq` as the fp32 values the reference multiplies with.  The loader casts the whole model, buffers
    included, to the 16-bit dtype (model_loader.py:77), and `torch.outer(positions.float(), inv_freq)`
    (qwen2.py:251) promotes it back: the VALUES are the rounded ones.  Cached per buffer."""
    inv = rotary.inv_freq
    if inv.dtype == torch.float32:
        return inv
    key = (inv.data_ptr(), inv.dtype, inv.device, inv._version)
    cached = rotary.__dict__.get("_bsgl_inv_freq_f32")
    if cached is None or cached[0] != key:
        cached = (key, inv.to(torch.float32))
        rotary.__dict__["_bsgl_inv_freq_f32"] = cached
    return cached[1]


def _module_global(obj, method: str, name: str, module: str):
    """The object `name` as the module that defines `type(obj).method` sees it (e.g. the `SimpleAttentionBackend`
    qwen2.py imported at :7, the `SimplifiedForwardBatch` scheduler.py imported at :8) -- the exact class the
    unchanged caller uses, whatever `sys.modules` holds by now.  Falls back to importing `module` by name."""
    fn = getattr(type(obj), method, None)
    found = getattr(fn, "__globals__", {}).get(name) if fn is not None else None
    if found is None:
        import importlib
        found = getattr(importlib.import_module(module), name)
    return found


def _fused_attention_forward(self, positions: torch.Tensor, hidden_states: torch.Tensor, forward_batch):
    """Drop-in for Qwen2Attention.forward (qwen2.py:179-197) with K5 in place of split + RoPE + store."""
    backend = self.__dict__.get("_bsgl_backend")
    if backend is None:
        backend = _module_global(self.attn, "forward", "SimpleAttentionBackend", "attn_backend")
        if not hasattr(backend, "forward_decode"):
            raise RuntimeError("install_fused_rope_kv_store needs the drop-in attn_backend (the model was built on "
                               "the reference's own backend)")
        self.__dict__["_bsgl_backend"] = backend
    qkv = self.qkv_proj.qkv_proj(hidden_states)                      # the fused Linear (qwen2.py:283)
    pool = forward_batch.token_to_kv_pool
    layer_id = self.attn.layer_id
    k_slab, v_slab = pool.get_kv_buffer(layer_id)
    if qkv.dtype != k_slab.dtype:
        # the reference casts k/v to the pool dtype inside set_kv_buffer (memory_pool.py:163-165);
        # the fused store keeps one dtype end to end
        raise RuntimeError(f"fused RoPE store needs activation dtype {qkv.dtype} == KV pool dtype {k_slab.dtype}")
    q = ops.rope_kv_store(qkv, positions, _inv_freq_f32(self.rotary_emb), self.num_heads, self.num_kv_heads,
                          self.head_dim, k_slab, v_slab, forward_batch.out_cache_loc)
    attn_output = backend.forward(q, None, None, self.attn, forward_batch, False)
    return self.o_proj(attn_output)


def install_fused_rope_kv_store(model: torch.nn.Module) -> int:
    """Route every attention module of `model` through K5.  Returns the number of modules patched.
    Works on a model already sharded by `tp_runtime.shard_attention_heads` (it reads the patched
    `num_heads` / `num_kv_heads`, and the all-reduce hook on `o_proj` still fires)."""
    mods: List[torch.nn.Module] = list(_attention_modules(model))
    if not mods:
        raise ValueError("model has no attention modules of the expected shape")
    for m in mods:
        if m.head_dim % 16 != 0:
            raise ValueError(f"head_dim {m.head_dim} is not a multiple of 16")
        if not torch.is_tensor(getattr(m.rotary_emb, "inv_freq", None)):
            raise ValueError("rotary_emb has no inv_freq buffer (qwen2.py:244-245)")
        m.forward = types.MethodType(_fused_attention_forward, m)
    return len(mods)


def uninstall_fused_rope_kv_store(model: torch.nn.Module) -> None:
    """Restore the class-level forward of every patched attention module."""
    for m in _attention_modules(model):
        m.__dict__.pop("forward", None)


# ------------------------------------------------------------------------------------------------------
# after the path: batched greedy sampling
# ------------------------------------------------------------------------------------------------------
def _batched_sample_next_ids(self, logits: torch.Tensor, batch_requests, mode: str) -> torch.Tensor:
    """Drop-in for Scheduler._sample_next_ids (scheduler.py:509-546).  Batches where every request is
    greedy and unconstrained go through K6 in one launch with no `.item()`; any other batch is handed
    to the reference's own method untouched (sampling / constrained decoding are outside this path)."""
    if any(req.do_sample or req.constraint_state for req in batch_requests):
        return type(self)._sample_next_ids(self, logits, batch_requests, mode)
    B = len(batch_requests)
    dev = logits.device
    qo = None
    if mode == "prefill":      # host lists the scheduler already holds (scheduler.py:514-516): no device sync
        ends, acc = [0], 0
        for req in batch_requests:
            acc += len(req.input_ids) - len(req.prefix_indices)
            ends.append(acc)
        qo = torch.tensor(ends, dtype=torch.int32).to(dev, non_blocking=True)
    prev = indptr = pens = None
    if any((req.repetition_penalty or 0) > 1.0 and len(req.output_ids) > 0 for req in batch_requests):
        flat, ptr = [], [0]
        for req in batch_requests:
            flat.extend(int(t) for t in req.output_ids)
            ptr.append(len(flat))
        prev = torch.tensor(flat, dtype=torch.int64).to(dev, non_blocking=True)
        indptr = torch.tensor(ptr, dtype=torch.int32).to(dev, non_blocking=True)
        pens = torch.tensor([float(req.repetition_penalty or 1.0) for req in batch_requests],
                            dtype=torch.float32).to(dev, non_blocking=True)
    if logits.dim() != 2:
        logits = logits.reshape(-1, logits.shape[-1])
    ids = ops.greedy_sample(logits, qo, prev, indptr, pens, batch=B)
    # `_process_results` reads every id with `.item()` (scheduler.py:553): hand it a host tensor so that is ONE
    # device->host copy per step instead of B blocking ones
    return ids.cpu()


def install_batched_greedy_sampler(scheduler) -> None:
    """Route `scheduler._sample_next_ids` (an instance of the reference's unchanged Scheduler) through K6."""
    if not hasattr(type(scheduler), "_sample_next_ids"):
        raise ValueError("object has no _sample_next_ids method (scheduler.py:509)")
    scheduler._sample_next_ids = types.MethodType(_batched_sample_next_ids, scheduler)


# ------------------------------------------------------------------------------------------------------
# around the path: device-side page-table update
# ------------------------------------------------------------------------------------------------------
def _device_prepare_prefill_batch(self, batch_requests):
    """Drop-in for Scheduler._prepare_prefill_batch (scheduler.py:381-452).  Same allocations in the same
    order, same errors, same request bookkeeping and the same page-table rows ([reused prefix slots |
    new slots | 0 ...] over all C cells), but: one H2D copy of a packed [5, B] descriptor instead of three
    `torch.tensor` uploads, one K7 launch instead of `zeros([B, C])` + two slice writes per request + a
    whole-row index-put, and no `.item()` (the reference syncs once per request at :439)."""
    fb_cls = _module_global(self, "_prepare_prefill_batch", "SimplifiedForwardBatch", "forward_batch")

    B = len(batch_requests)
    rows = self.req_to_token_pool.alloc(B)
    if rows is None:
        raise RuntimeError("Failed to allocate request pool indices")
    seq_lens = [len(req.input_ids) for req in batch_requests]
    prefix_lens = [int(req.num_cached_tokens) for req in batch_requests]
    concat_input_ids = torch.cat([req.input_ids[req.num_cached_tokens:] for req in batch_requests], dim=0)
    total_tokens = int(concat_input_ids.shape[0])

    out_cache_loc = self.token_to_kv_pool.alloc(total_tokens)
    if out_cache_loc is None:
        # need eviction from the radix cache (scheduler.py:410-421)
        self.tree_cache.evict(total_tokens, self.token_to_kv_pool.free)
        out_cache_loc = self.token_to_kv_pool.alloc(total_tokens)
        if out_cache_loc is None:
            avail = self.token_to_kv_pool.available_size() + self.tree_cache.evictable_size()
            raise RuntimeError(f"Prefill OOM: need {total_tokens}, available {avail}")

    new_offs, prefix_offs, n_new, n_pre = [], [], 0, 0
    for s, p in zip(seq_lens, prefix_lens):
        new_offs.append(n_new)
        prefix_offs.append(n_pre)
        n_new += s - p
        n_pre += p
    desc = torch.tensor([list(rows), prefix_lens, seq_lens, new_offs, prefix_offs], dtype=torch.int64)
    desc = desc.to(self.device, non_blocking=True)
    pool = self.req_to_token_pool
    prefix_flat = None
    if n_pre:
        parts = []
        for req, p in zip(batch_requests, prefix_lens):
            if p:
                if int(req.prefix_indices.shape[0]) != p:   # the reference's slice assignment would raise (:434)
                    raise RuntimeError(f"prefix_indices has {int(req.prefix_indices.shape[0])} slots, expected {p}")
                parts.append(req.prefix_indices)
        prefix_flat = torch.cat(parts).to(device=pool.req_to_token.device, dtype=torch.int32).contiguous()
    ops.page_table_write_rows(pool.req_to_token, desc, prefix_flat, out_cache_loc)

    for req, row, s in zip(batch_requests, rows, seq_lens):
        req.req_pool_idx = int(row)
        req.seq_len = s

    return fb_cls.create_prefill_batch(
        input_ids=concat_input_ids, req_pool_indices=desc[0], seq_lens=desc[2], out_cache_loc=out_cache_loc,
        req_to_token_pool=pool, token_to_kv_pool=self.token_to_kv_pool, prefix_lens=desc[1])


def _device_prepare_decode_batch(self, batch_requests):
    """Drop-in for Scheduler._prepare_decode_batch (scheduler.py:454-507): the B scalar device writes
    `req_to_token[r, seq_len] = slot` (:486-488) become one K7 launch and the three uploads one."""
    fb_cls = _module_global(self, "_prepare_decode_batch", "SimplifiedForwardBatch", "forward_batch")

    B = len(batch_requests)
    new_kv_slots = self.token_to_kv_pool.alloc(B)
    if new_kv_slots is None:
        raise RuntimeError("Failed to allocate KV slots for decode")
    rows, ids, lens = [], [], []
    for i, req in enumerate(batch_requests):
        if req.req_pool_idx is None:
            raise RuntimeError(f"Request {i} missing req_pool_idx for decode")
        if req.is_retracted:
            raise RuntimeError(f"Request {i} was retracted but is scheduled for decode")
        rows.append(req.req_pool_idx)
        # the last generated token is the input (a negative placeholder in overlap mode, overlap_worker.py:74-81)
        if req.overlap_next_input_id is not None:
            ids.append(req.overlap_next_input_id)
        else:
            if not req.output_ids:
                raise RuntimeError(f"Request {i} has no output tokens for decode")
            ids.append(req.output_ids[-1])
        lens.append(req.seq_len + 1)
    for req in batch_requests:          # only after every check passed, like the reference's single loop would
        req.seq_len += 1
    packed = torch.tensor([rows, lens, ids], dtype=torch.int64).to(self.device, non_blocking=True)
    pool = self.req_to_token_pool
    ops.page_table_append(pool.req_to_token, packed[0], packed[1], new_kv_slots)
    return fb_cls.create_decode_batch(
        input_ids=packed[2], req_pool_indices=packed[0], seq_lens=packed[1], out_cache_loc=new_kv_slots,
        req_to_token_pool=pool, token_to_kv_pool=self.token_to_kv_pool)


def install_device_batch_prep(scheduler) -> None:
    """Route `_prepare_prefill_batch` / `_prepare_decode_batch` of an instance of the reference's unchanged
    Scheduler through K7.  `_prepare_forward` (scheduler.py:371-379) dispatches to the instance attributes."""
    for name in ("_prepare_prefill_batch", "_prepare_decode_batch"):
        if not hasattr(type(scheduler), name):
            raise ValueError(f"object has no {name} method (scheduler.py:381,454)")
    scheduler._prepare_prefill_batch = types.MethodType(_device_prepare_prefill_batch, scheduler)
    scheduler._prepare_decode_batch = types.MethodType(_device_prepare_decode_batch, scheduler)


def uninstall_device_batch_prep(scheduler) -> None:
    for name in ("_prepare_prefill_batch", "_prepare_decode_batch"):
        scheduler.__dict__.pop(name, None)


# ------------------------------------------------------------------------------------------------------
# the whole decode step as one CUDA graph (small batches are launch-bound: ~37 us of host time per layer)
# ------------------------------------------------------------------------------------------------------
class _GraphedDecode:
    """One captured decode forward per batch size: K4 metadata + every layer (projections, RoPE / K5, K1, K2,
    K2c, MLP) + the LM head, replayed with in-place updates of five small input buffers.  Nothing in the step
    takes a size from the host (the page table and the KV slabs are read in place, K4 builds the plan on the
    device), so a replay after the batch moved on computes exactly what eager launches would."""

    def __init__(self, model: torch.nn.Module, max_graphs: int = 8):
        self.model = model
        self.eager_forward = model.forward            # the bound class-level forward
        self.max_graphs = max_graphs
        self.graphs = {}                              # (B, pools) -> captured state
        self.fb_cls = None
        self.replays = 0
        self.captures = 0

    def _capture(self, B: int, fb):
        dev = fb.input_ids.device
        i64 = dict(dtype=torch.int64, device=dev)
        st = {"ids": torch.zeros(B, **i64), "rpi": torch.zeros(B, **i64), "seq": torch.ones(B, **i64),
              "loc": torch.zeros(B, dtype=torch.int32, device=dev)}
        self._load(st, fb)
        fb_cls = type(fb)

        def step():
            sfb = fb_cls.create_decode_batch(input_ids=st["ids"], req_pool_indices=st["rpi"], seq_lens=st["seq"],
                                             out_cache_loc=st["loc"], req_to_token_pool=fb.req_to_token_pool,
                                             token_to_kv_pool=fb.token_to_kv_pool)
            return self.eager_forward(sfb.input_ids, sfb.positions, sfb), sfb

        # eager warm-up on a side stream (lazy library initialisation must not happen under capture); it also
        # performs this step's KV stores, which the captured replay below then repeats with identical values
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            step()
        torch.cuda.current_stream(dev).wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            logits, sfb = step()
        st["graph"], st["logits"], st["fb"] = g, logits, sfb
        self.captures += 1
        return st

    @staticmethod
    def _load(st, fb) -> None:
        st["ids"].copy_(fb.input_ids, non_blocking=True)
        st["rpi"].copy_(fb.req_pool_indices, non_blocking=True)
        st["seq"].copy_(fb.seq_lens, non_blocking=True)
        st["loc"].copy_(fb.out_cache_loc, non_blocking=True)

    def __call__(self, input_ids, positions, forward_batch, input_embeds=None):
        if input_embeds is not None or getattr(forward_batch, "is_prefill", True) or \
                not hasattr(type(forward_batch), "create_decode_batch") or not input_ids.is_cuda or \
                torch.cuda.is_current_stream_capturing():
            return self.eager_forward(input_ids, positions, forward_batch, input_embeds)
        B = int(input_ids.shape[0])
        key = (B, id(forward_batch.req_to_token_pool), id(forward_batch.token_to_kv_pool))
        st = self.graphs.get(key)
        if st is None:
            if len(self.graphs) >= self.max_graphs:
                self.graphs.pop(next(iter(self.graphs)))
            st = self.graphs[key] = self._capture(B, forward_batch)
        else:
            self._load(st, forward_batch)
        st["graph"].replay()
        self.replays += 1
        return st["logits"]


def install_cuda_graph_decode(model: torch.nn.Module, max_graphs: int = 8) -> "_GraphedDecode":
    """Replay the decode forward of `model` (the reference's unchanged BabyQwen2ForCausalLM on the drop-in
    backend) as one CUDA graph per batch size, captured on first use; prefill batches run eagerly.  The returned
    logits are a static buffer: valid until the next decode forward of the same batch size (the reference's
    scheduler and overlap worker sample from them at once, scheduler.py:154-160, overlap_worker.py:123-150).
    Install after the other hooks (K5 is captured inside the graph)."""
    hook = _GraphedDecode(model, max_graphs)
    model.forward = hook
    model.__dict__["_bsgl_graphed_decode"] = hook
    return hook


def uninstall_cuda_graph_decode(model: torch.nn.Module) -> None:
    model.__dict__.pop("forward", None)
    model.__dict__.pop("_bsgl_graphed_decode", None)
