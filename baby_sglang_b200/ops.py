"""Tensor-level wrappers over the C ABI (include/bsgl_b200.h).

PyTorch is plumbing here: it owns device memory and the current stream; every function below
validates its tensors, allocates outputs/workspaces from torch's stream-ordered caching
allocator and hands raw pointers plus `torch.cuda.current_stream()` to the library.  Nothing
here computes: without a CUDA device and the built library these functions raise.
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from . import _lib

_DTYPE_CODE = {torch.bfloat16: _lib.BF16, torch.float16: _lib.F16, torch.float32: _lib.F32}
_SM_COUNT = {}
#: counts launches issued through this module (bench.py reports it as `gpu_launches`)
LAUNCH_COUNT = 0


def _count(n: int = 1) -> None:
    global LAUNCH_COUNT
    LAUNCH_COUNT += n


def _require_cuda(*tensors: torch.Tensor) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise _lib.BsglError(
                "baby_sglang_b200 runs on CUDA tensors only (got a %s tensor); there is no CPU path" % t.device)
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise _lib.BsglError(f"tensors on different devices: {dev} vs {t.device}")
    return dev


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DTYPE_CODE[dt]
    except KeyError:
        raise _lib.BsglError(f"unsupported dtype {dt}") from None


def sm_count(device: Optional[torch.device] = None) -> int:
    idx = torch.cuda.current_device() if device is None or device.index is None else device.index
    if idx not in _SM_COUNT:
        _SM_COUNT[idx] = torch.cuda.get_device_properties(idx).multi_processor_count
    return _SM_COUNT[idx]


def heads_per_group(n_kv_heads: int) -> int:
    return _lib.load().bsgl_heads_per_group(int(n_kv_heads))


# --------------------------------------------------------------------------------------- K1
def kv_store(k_slab: torch.Tensor, v_slab: torch.Tensor, loc: torch.Tensor,
             cache_k: torch.Tensor, cache_v: torch.Tensor) -> None:
    """slab[loc[t]] = (cache_k[t], cache_v[t]); replaces memory_pool.py:152-171.

    `loc` is the int32 `out_cache_loc`; `cache_v` may be a strided view of the fused QKV output
    (qwen2.py:285-290).  Sources in another float dtype are cast round-to-nearest in the kernel.
    """
    _require_cuda(k_slab, v_slab, loc, cache_k, cache_v)
    T = int(loc.shape[0])
    if T == 0:
        return
    S1, H, D = k_slab.shape
    if not (k_slab.is_contiguous() and v_slab.is_contiguous()):
        raise _lib.BsglError("KV slabs must be contiguous [S+1, H, D]")
    if loc.dtype != torch.int32:
        loc = loc.to(torch.int32)
    if not loc.is_contiguous():
        loc = loc.contiguous()
    if cache_v.dtype != cache_k.dtype:
        cache_v = cache_v.to(cache_k.dtype)
    cache_k = cache_k.reshape(T, H, D) if cache_k.dim() != 3 else cache_k
    cache_v = cache_v.reshape(T, H, D) if cache_v.dim() != 3 else cache_v
    esz = cache_k.element_size()
    align = 16 // esz

    def _ok(t):
        return (t.stride(2) == 1 and t.stride(0) % align == 0 and t.stride(1) % align == 0
                and t.data_ptr() % 16 == 0)

    if not _ok(cache_k):
        cache_k = cache_k.contiguous()
    if not _ok(cache_v):
        cache_v = cache_v.contiguous()
    lib = _lib.load()
    rc = lib.bsgl_kv_store(k_slab.data_ptr(), v_slab.data_ptr(), S1, loc.data_ptr(), cache_k.data_ptr(),
                           cache_v.data_ptr(), T, H, D, cache_k.stride(0), cache_k.stride(1),
                           cache_v.stride(0), cache_v.stride(1), dtype_code(cache_k.dtype),
                           dtype_code(k_slab.dtype), _stream())
    _lib.check(rc, "bsgl_kv_store")
    _count()


# --------------------------------------------------------------------------------------- K5
def rope_kv_store(qkv: torch.Tensor, positions: torch.Tensor, inv_freq: torch.Tensor, n_q_heads: int,
                  n_kv_heads: int, head_dim: int, k_slab: Optional[torch.Tensor] = None,
                  v_slab: Optional[torch.Tensor] = None, loc: Optional[torch.Tensor] = None) -> torch.Tensor:
    """RoPE + q/k/v split + KV store in one launch (reference qwen2.py:249-268, 285-290 and
    memory_pool.py:152-171).  `qkv` is the fused projection output `[T, (Hq + 2 Hkv) D]`; returns the
    rotated q `[T, Hq, D]`.  With slabs and `loc` the rotated K rows and the V rows are written to
    `slab[loc[t]]`; without them only q is rotated."""
    _require_cuda(qkv, positions, inv_freq, k_slab, v_slab, loc)
    Hq, Hkv, D = int(n_q_heads), int(n_kv_heads), int(head_dim)
    if qkv.dim() != 2 or qkv.shape[1] != (Hq + 2 * Hkv) * D or qkv.stride(1) != 1:
        raise _lib.BsglError(f"qkv must be [T, (Hq + 2 Hkv) * D] with unit inner stride, got {tuple(qkv.shape)}")
    T = int(qkv.shape[0])
    store = k_slab is not None
    if store:
        if v_slab is None or loc is None:
            raise _lib.BsglError("k_slab, v_slab and loc go together")
        if k_slab.dtype != qkv.dtype or v_slab.dtype != qkv.dtype:
            raise _lib.BsglError("the fused RoPE store needs activation dtype == KV pool dtype")
        if tuple(k_slab.shape[1:]) != (Hkv, D) or not (k_slab.is_contiguous() and v_slab.is_contiguous()):
            raise _lib.BsglError("KV slabs must be contiguous [S+1, Hkv, D]")
        if loc.dtype != torch.int32:
            loc = loc.to(torch.int32)
        loc = loc.contiguous()
        if int(loc.shape[0]) != T:
            raise _lib.BsglError("loc must have one slot per token")
    if positions.dtype != torch.int64:
        positions = positions.to(torch.int64)
    positions = positions.contiguous()
    if int(positions.shape[0]) != T:
        raise _lib.BsglError("positions must have one entry per token")
    if inv_freq.dtype != torch.float32 or int(inv_freq.numel()) != D // 2:
        raise _lib.BsglError("inv_freq must be float32 [D/2]")
    inv_freq = inv_freq.contiguous()
    if qkv.stride(0) % 8 != 0 or qkv.data_ptr() % 16 != 0:
        qkv = qkv.contiguous()
    q_out = torch.empty((T, Hq, D), dtype=qkv.dtype, device=qkv.device)
    if T == 0:
        return q_out
    null = ctypes.c_void_p(0)
    rc = _lib.load().bsgl_rope_kv_store(
        qkv.data_ptr(), qkv.stride(0), positions.data_ptr(), inv_freq.data_ptr(),
        loc.data_ptr() if store else null, k_slab.data_ptr() if store else null,
        v_slab.data_ptr() if store else null, int(k_slab.shape[0]) if store else 0, q_out.data_ptr(), T, Hq, Hkv, D, dtype_code(qkv.dtype), _stream())
    _lib.check(rc, "bsgl_rope_kv_store")
    _count()
    return q_out


# --------------------------------------------------------------------------------------- K4
@dataclass
class AttnPlan:
    """Device-resident description of one batch for the paged-attention kernel (built by K4)."""
    n_rows: int
    n_kv_heads: int
    n_head_groups: int
    n_ctas: int
    row_idx_off: torch.Tensor      # int64 [rows]
    row_kv_len: torch.Tensor       # int32 [rows]
    plan_hdr: torch.Tensor         # int32 [4]  {n_tiles, n_ctas_eff, n_seqs, 0}
    tile_indptr: torch.Tensor      # int32 [rows*nhg + 1]
    cta_first_seq: torch.Tensor    # int32 [n_ctas + 1]


@dataclass
class BatchMetadata:
    positions: torch.Tensor                    # int64 [T]
    kv_indptr: torch.Tensor                    # int32 [B+1]
    qo_indptr: torch.Tensor                    # int32 [B+1]
    extended_lens: Optional[torch.Tensor]      # int64 [B] (extend only)
    plan: Optional[AttnPlan]
    kv_indices: Optional[torch.Tensor] = None  # int32 [sum(seq)] (built on request)
    heavy_order: Optional[torch.Tensor] = None  # int32 [B] requests by descending seq_len (extend only)
    work_counter: Optional[torch.Tensor] = None  # int32 [2], zero: K3's work-item dealer (returned to zero by every launch)
    _streams_seen: Optional[set] = None         # streams the buffers are already recorded on

    def use_on_current_stream(self) -> None:
        """The reference's overlap worker builds the batch on the scheduler thread's stream and runs the forward on
        another one (overlap_worker.py:83-106).  torch's caching allocator only knows the allocating stream, so a
        buffer released right after the launches were enqueued could be handed out again while the kernels still
        read it: record every buffer on the consuming stream, once per stream."""
        cur = torch.cuda.current_stream()
        key = cur.cuda_stream
        if self._streams_seen is None:
            self._streams_seen = set()
        if key in self._streams_seen:
            return
        self._streams_seen.add(key)
        tensors = [self.positions, self.kv_indptr, self.qo_indptr, self.extended_lens, self.kv_indices, self.heavy_order,
                   self.work_counter]
        if self.plan is not None:
            tensors += [self.plan.row_idx_off, self.plan.row_kv_len, self.plan.plan_hdr, self.plan.tile_indptr,
                        self.plan.cta_first_seq]
        for t in tensors:
            if t is not None and t.is_cuda:
                t.record_stream(cur)


def _plan_buffers(dev, n_rows: int, nhg: int, n_ctas: int):
    i32 = dict(dtype=torch.int32, device=dev)
    return (torch.empty(n_rows, dtype=torch.int64, device=dev), torch.empty(n_rows, **i32),
            torch.empty(_lib.PLAN_HDR_INTS, **i32), torch.empty(n_rows * nhg + 1, **i32),
            torch.empty(n_ctas + 1, **i32))


def _as_i64(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.int64:
        t = t.to(torch.int64)
    return t if t.is_contiguous() else t.contiguous()


def build_decode_metadata(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                          seq_lens: torch.Tensor, n_kv_heads: int,
                          n_ctas: Optional[int] = None) -> BatchMetadata:
    """Device-side replacement of create_decode_batch's host work (forward_batch.py:61-84) and of
    the per-request `.item()` chain in `_decode_attention` (attn_backend.py:117-125)."""
    dev = _require_cuda(req_to_token, req_pool_indices, seq_lens)
    B = int(req_pool_indices.shape[0])
    nhg = int(n_kv_heads) // heads_per_group(n_kv_heads)
    n_ctas = n_ctas or sm_count(dev)
    rpi, sl = _as_i64(req_pool_indices), _as_i64(seq_lens)
    positions = torch.empty(B, dtype=torch.int64, device=dev)
    kv_indptr = torch.empty(B + 1, dtype=torch.int32, device=dev)
    qo_indptr = torch.empty(B + 1, dtype=torch.int32, device=dev)
    row_idx_off, row_kv_len, plan_hdr, tile_indptr, cta_first_seq = _plan_buffers(dev, B, nhg, n_ctas)
    rc = _lib.load().bsgl_build_decode_metadata(
        req_to_token.stride(0), rpi.data_ptr(), sl.data_ptr(), B, nhg, n_ctas, positions.data_ptr(),
        kv_indptr.data_ptr(), qo_indptr.data_ptr(), row_idx_off.data_ptr(), row_kv_len.data_ptr(),
        plan_hdr.data_ptr(), tile_indptr.data_ptr(), cta_first_seq.data_ptr(), _stream())
    _lib.check(rc, "bsgl_build_decode_metadata")
    _count()
    plan = AttnPlan(B, int(n_kv_heads), nhg, n_ctas, row_idx_off, row_kv_len, plan_hdr, tile_indptr, cta_first_seq)
    return BatchMetadata(positions, kv_indptr, qo_indptr, None, plan)


def build_extend_metadata(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                          seq_lens: torch.Tensor, prefix_lens: torch.Tensor, n_tokens: int,
                          n_kv_heads: int, n_ctas: Optional[int] = None,
                          row_plan: Optional[bool] = None) -> BatchMetadata:
    """Device-side replacement of create_prefill_batch's host loop (forward_batch.py:36-45) and of
    the per-request `.item()` chain in `_prefill_attention` (attn_backend.py:43-57).  `n_tokens`
    (= q rows) is known on the host: it is `input_ids.shape[0]`."""
    dev = _require_cuda(req_to_token, req_pool_indices, seq_lens, prefix_lens)
    B = int(req_pool_indices.shape[0])
    T = int(n_tokens)
    if row_plan is None:   # the per-token split-KV plan is only needed by the row-wise path
        row_plan = extend_mode(T, B) == "rowwise"
    nhg = int(n_kv_heads) // heads_per_group(n_kv_heads)
    n_ctas = n_ctas or sm_count(dev)
    rpi, sl, pl = _as_i64(req_pool_indices), _as_i64(seq_lens), _as_i64(prefix_lens)
    extended = torch.empty(B, dtype=torch.int64, device=dev)
    positions = torch.empty(T, dtype=torch.int64, device=dev)
    kv_indptr = torch.empty(B + 1, dtype=torch.int32, device=dev)
    qo_indptr = torch.empty(B + 1, dtype=torch.int32, device=dev)
    row_idx_off, row_kv_len, plan_hdr, tile_indptr, cta_first_seq = _plan_buffers(dev, T, nhg, n_ctas)
    heavy_order = torch.empty(B, dtype=torch.int32, device=dev)
    null = ctypes.c_void_p(0)
    rc = _lib.load().bsgl_build_extend_metadata(
        req_to_token.stride(0), rpi.data_ptr(), sl.data_ptr(), pl.data_ptr(), B, T, nhg, n_ctas,
        extended.data_ptr(), positions.data_ptr(), qo_indptr.data_ptr(), kv_indptr.data_ptr(),
        row_idx_off.data_ptr(), row_kv_len.data_ptr(),
        plan_hdr.data_ptr() if row_plan else null, tile_indptr.data_ptr() if row_plan else null,
        cta_first_seq.data_ptr() if row_plan else null, heavy_order.data_ptr(), _stream())
    _lib.check(rc, "bsgl_build_extend_metadata")
    _count()
    plan = AttnPlan(T, int(n_kv_heads), nhg, n_ctas, row_idx_off, row_kv_len, plan_hdr, tile_indptr,
                    cta_first_seq) if row_plan else None
    return BatchMetadata(positions, kv_indptr, qo_indptr, extended, plan, heavy_order=heavy_order,
                         work_counter=torch.zeros(2, dtype=torch.int32, device=dev))


def gather_kv_indices(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                      kv_indptr: torch.Tensor, total_kv: int) -> torch.Tensor:
    """CSR flattening of `req_to_token[r, :seq]` (the slices at attn_backend.py:55-57,123-125)."""
    dev = _require_cuda(req_to_token, req_pool_indices, kv_indptr)
    out = torch.empty(int(total_kv), dtype=torch.int32, device=dev)
    rpi = _as_i64(req_pool_indices)
    rc = _lib.load().bsgl_gather_kv_indices(req_to_token.data_ptr(), req_to_token.stride(0), rpi.data_ptr(),
                                            kv_indptr.data_ptr(), int(rpi.shape[0]), out.data_ptr(), _stream())
    _lib.check(rc, "bsgl_gather_kv_indices")
    _count()
    return out


# --------------------------------------------------------------------------------------- K2
def partial_buffers(dev, n_ctas: int, n_q_heads: int, n_kv_heads: int, head_dim: int) -> Tuple[torch.Tensor, torch.Tensor]:
    o = ctypes.c_int64(0)
    l = ctypes.c_int64(0)
    rc = _lib.load().bsgl_partial_sizes(n_ctas, n_q_heads, n_kv_heads, head_dim, ctypes.byref(o), ctypes.byref(l))
    _lib.check(rc, "bsgl_partial_sizes")
    return (torch.empty(o.value, dtype=torch.float32, device=dev),
            torch.empty(l.value, dtype=torch.float32, device=dev))


def paged_attention(q: torch.Tensor, k_slab: torch.Tensor, v_slab: torch.Tensor, idx_base: torch.Tensor,
                    plan: AttnPlan, sm_scale: float, out: Optional[torch.Tensor] = None,
                    workspace: Optional[Tuple[torch.Tensor, torch.Tensor]] = None,
                    merge: bool = True) -> torch.Tensor:
    """One q row per plan row over the paged KV; returns [rows, Hq*D] in q's dtype.

    Replaces `_decode_attention` (attn_backend.py:112-149): K2 split-KV partial pass + K2c merge.
    `idx_base` is the int32 page table (`req_to_token`) or a flattened `kv_indices`.
    """
    dev = _require_cuda(q, k_slab, v_slab, idx_base)
    if q.dim() != 3:
        raise _lib.BsglError(f"q must be [rows, Hq, D], got {tuple(q.shape)}")
    rows, Hq, D = q.shape
    if rows != plan.n_rows:
        raise _lib.BsglError(f"q has {rows} rows but the plan was built for {plan.n_rows}")
    Hkv = int(k_slab.shape[1])
    if Hkv != plan.n_kv_heads or int(k_slab.shape[2]) != D:
        raise _lib.BsglError("KV slab geometry does not match q / plan")
    if q.dtype != k_slab.dtype or q.dtype != v_slab.dtype:
        raise _lib.BsglError(f"q dtype {q.dtype} != KV slab dtype {k_slab.dtype}")
    if idx_base.dtype != torch.int32:
        raise _lib.BsglError("page table must be int32")
    if not q.is_contiguous():
        q = q.contiguous()
    if out is None:
        out = torch.empty((rows, Hq * D), dtype=q.dtype, device=dev)
    if rows == 0:
        return out
    part_o, part_lse = workspace if workspace is not None else partial_buffers(dev, plan.n_ctas, Hq, Hkv, D)
    lib = _lib.load()
    code = dtype_code(q.dtype)
    st = _stream()
    rc = lib.bsgl_paged_attention(
        q.data_ptr(), k_slab.data_ptr(), v_slab.data_ptr(), idx_base.data_ptr(), plan.row_idx_off.data_ptr(),
        plan.row_kv_len.data_ptr(), plan.plan_hdr.data_ptr(), plan.tile_indptr.data_ptr(),
        plan.cta_first_seq.data_ptr(), rows, Hq, Hkv, D, float(sm_scale), code, out.data_ptr(),
        part_o.data_ptr(), part_lse.data_ptr(), plan.n_ctas, st)
    _lib.check(rc, "bsgl_paged_attention")
    _count()
    if merge:
        rc = lib.bsgl_merge_partials(part_o.data_ptr(), part_lse.data_ptr(), plan.plan_hdr.data_ptr(),
                                     plan.tile_indptr.data_ptr(), plan.cta_first_seq.data_ptr(), plan.n_ctas,
                                     rows, Hq, Hkv, D, code, out.data_ptr(), st)
        _lib.check(rc, "bsgl_merge_partials")
        _count()
    return out


# --------------------------------------------------------------------------------------- K3
#: extends with at most this many new tokens per request on average run row-wise on the split-KV
#: kernel (a 256-row tensor-core chunk would be mostly padding); everything else runs on tcgen05.
ROWWISE_MAX_AVG_EXTEND = 4


def extend_mode(n_tokens: int, batch: int) -> str:
    forced = os.environ.get("BSGL_EXTEND_MODE", "")
    if forced in ("tcgen05", "rowwise"):
        return forced
    return "rowwise" if n_tokens <= ROWWISE_MAX_AVG_EXTEND * max(batch, 1) else "tcgen05"


def extend_attention_tc(q: torch.Tensor, k_slab: torch.Tensor, v_slab: torch.Tensor, req_to_token: torch.Tensor,
                        req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, prefix_lens: torch.Tensor,
                        qo_indptr: torch.Tensor, sm_scale: float, out: Optional[torch.Tensor] = None,
                        req_order: Optional[torch.Tensor] = None,
                        work_counter: Optional[torch.Tensor] = None) -> torch.Tensor:
    """K3: ragged extend attention on tcgen05/TMEM (reference `_prefill_attention`,
    attn_backend.py:34-110).  q rows are packed in request order (`qo_indptr`); request i's row r
    sees KV columns 0..prefix_i + r of its page-table row.  `req_order` (K4's `heavy_order`) only
    changes the order in which requests are dispatched to the SMs, never the result.  `work_counter` (int32 [2],
    zero; K4's metadata owns one) lets the persistent CTAs draw their work items greedily from a device counter --
    each launch returns it to zero; without it the items are dealt statically.  Launches that share a counter must be
    stream-ordered (the layers of one forward are)."""
    dev = _require_cuda(q, k_slab, v_slab, req_to_token, req_pool_indices, seq_lens, prefix_lens, qo_indptr)
    if q.dim() != 3:
        raise _lib.BsglError(f"q must be [tokens, Hq, D], got {tuple(q.shape)}")
    T, Hq, D = q.shape
    Hkv = int(k_slab.shape[1])
    if int(k_slab.shape[2]) != D or Hq % Hkv != 0:
        raise _lib.BsglError("KV slab geometry does not match q")
    if q.dtype != k_slab.dtype or q.dtype != v_slab.dtype:
        raise _lib.BsglError(f"q dtype {q.dtype} != KV slab dtype {k_slab.dtype}")
    if req_to_token.dtype != torch.int32 or qo_indptr.dtype != torch.int32:
        raise _lib.BsglError("page table and qo_indptr must be int32")
    if not q.is_contiguous():
        q = q.contiguous()
    if out is None:
        out = torch.empty((T, Hq * D), dtype=q.dtype, device=dev)
    B = int(req_pool_indices.shape[0])
    if T == 0 or B == 0:
        return out
    rpi, sl, pl = _as_i64(req_pool_indices), _as_i64(seq_lens), _as_i64(prefix_lens)
    if req_order is not None and (req_order.dtype != torch.int32 or int(req_order.shape[0]) != B or
                                  not req_order.is_cuda):
        raise _lib.BsglError("req_order must be a CUDA int32 tensor of length batch")
    if work_counter is not None and (work_counter.dtype != torch.int32 or int(work_counter.numel()) < 2 or
                                     not work_counter.is_cuda or not work_counter.is_contiguous()):
        raise _lib.BsglError("work_counter must be a contiguous CUDA int32 tensor of at least 2 elements")
    rc = _lib.load().bsgl_extend_attention(
        q.data_ptr(), k_slab.data_ptr(), v_slab.data_ptr(), req_to_token.data_ptr(), req_to_token.stride(0),
        rpi.data_ptr(), sl.data_ptr(), pl.data_ptr(), qo_indptr.data_ptr(),
        req_order.data_ptr() if req_order is not None else ctypes.c_void_p(0), B, T, Hq, Hkv, D, float(sm_scale),
        dtype_code(q.dtype), out.data_ptr(),
        work_counter.data_ptr() if work_counter is not None else ctypes.c_void_p(0), _stream())
    _lib.check(rc, "bsgl_extend_attention")
    _count()
    return out


def extend_attention(q: torch.Tensor, k_slab: torch.Tensor, v_slab: torch.Tensor, forward_batch,
                     md: BatchMetadata, sm_scale: float, mode: Optional[str] = None) -> torch.Tensor:
    """Ragged extend (reference `_prefill_attention`, attn_backend.py:34-110).

    "tcgen05": K3, the tensor-core kernel.  "rowwise" (tiny extends): q token t of request i is a
    sequence of prefix_i + r + 1 KV tokens over request i's page-table row, which is exactly the
    mask `col <= prefix+row` (attn_backend.py:83-85); it runs on the split-KV kernel (K2) with the
    per-row plan K4 builds."""
    r2t = forward_batch.req_to_token_pool.req_to_token
    B = int(forward_batch.req_pool_indices.shape[0])
    mode = mode or extend_mode(int(q.shape[0]), B)
    if mode == "tcgen05":
        return extend_attention_tc(q, k_slab, v_slab, r2t, forward_batch.req_pool_indices, forward_batch.seq_lens,
                                   forward_batch.prefix_lens, md.qo_indptr, sm_scale, req_order=md.heavy_order,
                                   work_counter=md.work_counter)
    if md.plan is None:   # metadata was built for the tensor-core path: add the row plan now
        full = build_extend_metadata(r2t, forward_batch.req_pool_indices, forward_batch.seq_lens,
                                     forward_batch.prefix_lens, int(q.shape[0]), int(k_slab.shape[1]),
                                     row_plan=True)
        md.plan = full.plan
    return paged_attention(q, k_slab, v_slab, r2t, md.plan, sm_scale)


# --------------------------------------------------------------------------------------- K6
def greedy_sample(logits: torch.Tensor, qo_indptr: Optional[torch.Tensor] = None,
                  prev_ids: Optional[torch.Tensor] = None, prev_indptr: Optional[torch.Tensor] = None,
                  penalties: Optional[torch.Tensor] = None, batch: Optional[int] = None) -> torch.Tensor:
    """Batched greedy next-token ids, int64 [B], on the device (reference scheduler.py:509-546 for
    `do_sample=False` requests: last-token gather, sample.py:19-36 repetition penalty, argmax).
    `qo_indptr` (K4) selects each request's last row of a prefill's `[T, V]` logits; without it row i
    belongs to request i.  `prev_ids`/`prev_indptr`/`penalties` describe the per-request histories."""
    dev = _require_cuda(logits, qo_indptr, prev_ids, prev_indptr, penalties)
    if logits.dim() != 2 or logits.stride(1) != 1:
        raise _lib.BsglError(f"logits must be [rows, vocab] with unit inner stride, got {tuple(logits.shape)}")
    V = int(logits.shape[1])
    if qo_indptr is not None:
        if qo_indptr.dtype != torch.int32:
            raise _lib.BsglError("qo_indptr must be int32")
        B = int(qo_indptr.shape[0]) - 1
    else:
        B = int(logits.shape[0]) if batch is None else int(batch)
    have_prev = prev_ids is not None
    if have_prev:
        if prev_indptr is None or penalties is None:
            raise _lib.BsglError("prev_ids, prev_indptr and penalties go together")
        if prev_ids.dtype != torch.int64 or prev_indptr.dtype != torch.int32 or penalties.dtype != torch.float32:
            raise _lib.BsglError("prev_ids int64, prev_indptr int32, penalties float32")
        if int(prev_indptr.shape[0]) != B + 1 or int(penalties.shape[0]) != B:
            raise _lib.BsglError("prev_indptr / penalties do not match the batch")
        prev_ids, prev_indptr, penalties = prev_ids.contiguous(), prev_indptr.contiguous(), penalties.contiguous()
    out = torch.empty(B, dtype=torch.int64, device=dev)
    if B == 0:
        return out
    null = ctypes.c_void_p(0)
    rc = _lib.load().bsgl_greedy_sample(
        logits.data_ptr(), logits.stride(0), V, dtype_code(logits.dtype),
        qo_indptr.data_ptr() if qo_indptr is not None else null, B,
        prev_ids.data_ptr() if have_prev else null, prev_indptr.data_ptr() if have_prev else null,
        penalties.data_ptr() if have_prev else null, out.data_ptr(), _stream())
    _lib.check(rc, "bsgl_greedy_sample")
    _count()
    return out


# --------------------------------------------------------------------------------------- K7
def _check_page_table(req_to_token: torch.Tensor) -> None:
    if req_to_token.dtype != torch.int32 or req_to_token.dim() != 2 or req_to_token.stride(1) != 1:
        raise _lib.BsglError("req_to_token must be an int32 [rows, C] table with unit inner stride")


def page_table_write_rows(req_to_token: torch.Tensor, desc: torch.Tensor, prefix_flat: Optional[torch.Tensor],
                          new_slots: Optional[torch.Tensor]) -> None:
    """Prefill rows of one batch in one launch (reference scheduler.py:423-442).  `desc` is int64
    [5, B]: req_pool_idx, prefix_len, seq_len, new_off, prefix_off; row r becomes
    [prefix_flat[prefix_off:+prefix] | new_slots[new_off:+seq-prefix] | zeros] over all C cells."""
    _require_cuda(req_to_token, desc, prefix_flat, new_slots)
    _check_page_table(req_to_token)
    if desc.dtype != torch.int64 or desc.dim() != 2 or desc.shape[0] != 5 or not desc.is_contiguous():
        raise _lib.BsglError("desc must be a contiguous int64 [5, B] tensor")
    for name, t in (("prefix_flat", prefix_flat), ("new_slots", new_slots)):
        if t is not None and (t.dtype != torch.int32 or not t.is_contiguous()):
            raise _lib.BsglError(f"{name} must be contiguous int32")
    B = int(desc.shape[1])
    if B == 0:
        return
    null = ctypes.c_void_p(0)
    rc = _lib.load().bsgl_page_table_write_rows(
        req_to_token.data_ptr(), req_to_token.stride(0), int(req_to_token.shape[0]), int(req_to_token.shape[1]),
        desc.data_ptr(), B, prefix_flat.data_ptr() if prefix_flat is not None and prefix_flat.numel() else null,
        new_slots.data_ptr() if new_slots is not None and new_slots.numel() else null, _stream())
    _lib.check(rc, "bsgl_page_table_write_rows")
    _count()


def page_table_append(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
                      slots: torch.Tensor) -> None:
    """`req_to_token[req_pool_indices[i], seq_lens[i] - 1] = slots[i]` for the whole decode batch in one
    launch (reference scheduler.py:486-488; `seq_lens` already counts the new token, :489-490)."""
    _require_cuda(req_to_token, req_pool_indices, seq_lens, slots)
    _check_page_table(req_to_token)
    if req_pool_indices.dtype != torch.int64 or seq_lens.dtype != torch.int64 or slots.dtype != torch.int32:
        raise _lib.BsglError("req_pool_indices / seq_lens must be int64 and slots int32")
    B = int(slots.shape[0])
    if int(req_pool_indices.shape[0]) != B or int(seq_lens.shape[0]) != B:
        raise _lib.BsglError("req_pool_indices / seq_lens / slots lengths differ")
    if B == 0:
        return
    rc = _lib.load().bsgl_page_table_append(
        req_to_token.data_ptr(), req_to_token.stride(0), int(req_to_token.shape[0]), int(req_to_token.shape[1]),
        req_pool_indices.contiguous().data_ptr(), seq_lens.contiguous().data_ptr(), slots.contiguous().data_ptr(), B,
        _stream())
    _lib.check(rc, "bsgl_page_table_append")
    _count()


def slot_ring_pop(ring: torch.Tensor, head: int, n: int) -> torch.Tensor:
    """n slots from the head of the device FIFO `ring` (int32 [capacity]) in FIFO order (memory_pool.py:68-74)."""
    dev = _require_cuda(ring)
    out = torch.empty(n, dtype=torch.int32, device=dev)
    if n:
        rc = _lib.load().bsgl_slot_ring_pop(ring.data_ptr(), int(ring.numel()), int(head), int(n), out.data_ptr(),
                                            _stream())
        _lib.check(rc, "bsgl_slot_ring_pop")
        _count()
    return out


def slot_ring_push(ring: torch.Tensor, tail: int, vals: torch.Tensor) -> None:
    """Append `vals` (int32/int64 device tensor, any 1-D stride, e.g. a slice of a page-table row) at the
    tail of the device FIFO, in order, without a device->host copy (memory_pool.py:76-80)."""
    _require_cuda(ring, vals)
    if vals.dtype not in (torch.int32, torch.int64):
        raise _lib.BsglError(f"freed slots must be int32 or int64, got {vals.dtype}")
    if vals.dim() != 1:
        vals = vals.reshape(-1)
    n = int(vals.numel())
    if n == 0:
        return
    rc = _lib.load().bsgl_slot_ring_push(ring.data_ptr(), int(ring.numel()), int(tail), n, vals.data_ptr(),
                                         int(vals.stride(0)), 1 if vals.dtype == torch.int64 else 0, _stream())
    _lib.check(rc, "bsgl_slot_ring_push")
    _count()


# --------------------------------------------------------------------------------------- K8
class CommBuffer:
    """A device buffer other processes on the node can map (cudaMalloc + CUDA IPC through the C ABI).
    `ptr` is the local address, `handle` the 64 bytes a peer passes to `CommBuffer.open`."""

    def __init__(self, ptr: int, nbytes: int, handle: Optional[bytes], owned: bool):
        self.ptr, self.nbytes, self.handle, self._owned = ptr, nbytes, handle, owned

    @classmethod
    def alloc(cls, nbytes: int) -> "CommBuffer":
        if not torch.cuda.is_available():
            raise _lib.BsglError("communicator buffers need a CUDA device; there is no CPU path")
        p = ctypes.c_void_p(0)
        h = ctypes.create_string_buffer(64)
        _lib.check(_lib.load().bsgl_comm_alloc(int(nbytes), ctypes.byref(p), h), "bsgl_comm_alloc")
        return cls(int(p.value), int(nbytes), bytes(h.raw), True)

    @classmethod
    def open(cls, handle: bytes, nbytes: int) -> "CommBuffer":
        p = ctypes.c_void_p(0)
        _lib.check(_lib.load().bsgl_comm_open(ctypes.c_char_p(handle), ctypes.byref(p)), "bsgl_comm_open")
        return cls(int(p.value), int(nbytes), None, False)

    def release(self) -> None:
        if self.ptr:
            lib = _lib.load()
            _lib.check(lib.bsgl_comm_free(self.ptr) if self._owned else lib.bsgl_comm_close(self.ptr), "bsgl_comm_free")
            self.ptr = 0

    def as_tensor(self, offset_bytes: int, shape, dtype: torch.dtype, device) -> torch.Tensor:
        """Zero-copy torch view of a region of the buffer (via __cuda_array_interface__)."""
        typestr = {torch.bfloat16: "<u2", torch.float16: "<f2", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
        holder = type("_CAI", (), {})()
        holder.__cuda_array_interface__ = {"shape": tuple(int(x) for x in shape), "typestr": typestr,
                                           "data": (self.ptr + int(offset_bytes), False), "version": 3}
        t = torch.as_tensor(holder, device=device)
        return t.view(dtype) if t.dtype != dtype else t


def oproj_pointer_tables(staging_ptrs, out_ptrs, flag_ptrs):
    """The three per-rank pointer tables of a communicator as ctypes arrays (build once, pass to every call)."""
    arr = ctypes.c_void_p * len(staging_ptrs)
    return arr(*staging_ptrs), arr(*out_ptrs), arr(*flag_ptrs)


def oproj_allreduce(a: torch.Tensor, w: torch.Tensor, rank: int, world: int, staging_ptrs, out_ptrs, flag_ptrs,
                    max_slots: int, max_tiles: int, out_bytes: int, mc_staging: int = 0, mc_out: int = 0,
                    timeout_s: float = 30.0) -> None:
    """K8 launch: this rank's o_proj partial GEMM with the SUM all-reduce fused in (reference qwen2.py:195 +
    tp_runtime.py:28-33).  `a` [T, K_r] and `w` [hidden, K_r] are this rank's shard; the result lands in every
    rank's output buffer.  `mc_staging` / `mc_out` are the NVLS multicast views of the staging / output regions
    (0 = exchange by peer stores).  All ranks must make the same sequence of calls (the call counter lives in
    device memory, so the launch can be captured into a CUDA graph)."""
    _require_cuda(a, w)
    if a.dtype != w.dtype or a.dim() != 2 or w.dim() != 2 or a.shape[1] != w.shape[1]:
        raise _lib.BsglError(f"oproj_allreduce: a {tuple(a.shape)} {a.dtype} vs w {tuple(w.shape)} {w.dtype}")
    if a.stride(1) != 1 or w.stride(1) != 1:
        raise _lib.BsglError("oproj_allreduce: operands need unit inner stride")
    if not isinstance(staging_ptrs, ctypes.Array):
        staging_ptrs, out_ptrs, flag_ptrs = oproj_pointer_tables(staging_ptrs, out_ptrs, flag_ptrs)
    rc = _lib.load().bsgl_oproj_allreduce(
        a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), int(a.shape[0]), int(w.shape[0]), int(a.shape[1]),
        int(rank), int(world), staging_ptrs, out_ptrs, flag_ptrs, ctypes.c_void_p(int(mc_staging) or None),
        ctypes.c_void_p(int(mc_out) or None), int(max_slots), int(max_tiles), int(out_bytes), float(timeout_s),
        dtype_code(a.dtype), _stream())
    _lib.check(rc, "bsgl_oproj_allreduce")
    _count()
