"""Drop-in for the reference's `tp_runtime` module (reference tp_runtime.py:10-51) plus the wiring the
reference leaves as a roadmap item (it defines the four helpers but never calls them, README.md:50).

Same four public names and semantics:

    init_tp() -> (rank, world_size, tp_group)      env RANK / WORLD_SIZE / LOCAL_RANK          tp_runtime.py:10-25
    tp_all_reduce_sum(x, tp_group)                 in-place SUM all-reduce, returns x           :28-33
    tp_all_gather(x, tp_group)                     all-gather + cat(dim=0)                      :36-43
    tp_broadcast(x, src, tp_group)                 in-place broadcast                           :46-51

The attention path shards only along its natural, exchange-free axes:

  * KV-head tensor parallelism when `num_kv_heads % world == 0`: rank r owns KV heads
    [r*Hkv/t, (r+1)*Hkv/t) and their G q heads each.  Page table, slot allocator, radix cache and
    scheduler are replicated and identical on every rank (the same slot numbers index different head
    slices), so paged attention itself needs no collective; ONE collective per layer follows it: the SUM
    all-reduce of the bias-free `o_proj` partial `[T, hidden]` (qwen2.py:166,195) over NCCL / NVLink.
    `shard_attention_heads(model, ...)` wires this into the reference's UNCHANGED model by slicing the
    fused QKV / o_proj weights of each attention module in place and hooking `o_proj`.
  * request-batch data parallelism otherwise: `dp_partition` splits the request list; replicas are
    independent (no collective in the data path).
"""
from __future__ import annotations

import copy
import ctypes
import os
from dataclasses import dataclass
from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist

__all__ = ["init_tp", "tp_all_reduce_sum", "tp_all_gather", "tp_broadcast", "plan_parallelism", "ParallelPlan",
           "shard_attention_heads", "kv_pool_config", "dp_partition", "dp_gather_outputs", "FusedOProjAllReduce"]


# ------------------------------------------------------------------------------------------------------
# the reference's four helpers
# ------------------------------------------------------------------------------------------------------
def init_tp() -> Tuple[int, int, object]:
    rank = int(os.getenv("RANK", "0"))
    world_size = int(os.getenv("WORLD_SIZE", "1"))
    local_rank = int(os.getenv("LOCAL_RANK", "0"))
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(local_rank)
    if world_size > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        # NCCL over NVLink/NVSwitch on GPUs (the reference's backend); gloo only exists so the host logic can
        # be exercised on a machine without GPUs.
        if cuda:
            dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
        else:
            dist.init_process_group(backend="gloo")
    tp_group = dist.group.WORLD
    if dist.is_initialized():
        rank = dist.get_rank(group=tp_group)
        world_size = dist.get_world_size(group=tp_group)
    return rank, world_size, tp_group


def _world(tp_group) -> int:
    return dist.get_world_size(group=tp_group) if dist.is_initialized() else 1


def tp_all_reduce_sum(x: torch.Tensor, tp_group) -> torch.Tensor:
    if _world(tp_group) == 1:
        return x
    dist.all_reduce(x, op=dist.ReduceOp.SUM, group=tp_group)
    return x


def tp_all_gather(x: torch.Tensor, tp_group) -> torch.Tensor:
    world = _world(tp_group)
    if world == 1:
        return x
    out = torch.empty((world * x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x.contiguous(), group=tp_group)   # == cat(gather_list, dim=0)
    return out


def tp_broadcast(x: torch.Tensor, src: int, tp_group) -> torch.Tensor:
    if _world(tp_group) == 1:
        return x
    dist.broadcast(x, src=src, group=tp_group)
    return x


# ------------------------------------------------------------------------------------------------------
# which shard applies
# ------------------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class ParallelPlan:
    mode: str            # "single" | "kv_head_tp" | "request_dp"
    world_size: int
    kv_heads_per_rank: int
    q_heads_per_rank: int


def plan_parallelism(num_q_heads: int, num_kv_heads: int, world_size: int) -> ParallelPlan:
    """KV-head TP where `num_kv_heads >= world` and divisible; request-batch DP otherwise."""
    if world_size <= 1:
        return ParallelPlan("single", 1, num_kv_heads, num_q_heads)
    if num_kv_heads >= world_size and num_kv_heads % world_size == 0:
        return ParallelPlan("kv_head_tp", world_size, num_kv_heads // world_size, num_q_heads // world_size)
    return ParallelPlan("request_dp", world_size, num_kv_heads, num_q_heads)


# ------------------------------------------------------------------------------------------------------
# KV-head tensor parallelism on an unchanged model
# ------------------------------------------------------------------------------------------------------
def _attention_modules(model: torch.nn.Module):
    """Every module shaped like the reference's Qwen2Attention (qwen2.py:143-177): a fused `qkv_proj.qkv_proj`
    linear, a bias-free `o_proj`, an `attn` shim and the head bookkeeping attributes."""
    for m in model.modules():
        inner = getattr(getattr(m, "qkv_proj", None), "qkv_proj", None)
        if isinstance(inner, torch.nn.Linear) and isinstance(getattr(m, "o_proj", None), torch.nn.Linear) \
                and hasattr(m, "attn") and hasattr(m, "num_kv_heads"):
            yield m


class _RawView:
    """Zero-copy torch view of raw device memory (via __cuda_array_interface__)."""

    @staticmethod
    def make(ptr: int, shape, dtype: torch.dtype, device) -> torch.Tensor:
        typestr = {torch.bfloat16: "<u2", torch.float16: "<f2", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
        holder = type("_CAI", (), {})()
        holder.__cuda_array_interface__ = {"shape": tuple(int(x) for x in shape), "typestr": typestr,
                                           "data": (int(ptr), False), "version": 3}
        t = torch.as_tensor(holder, device=device)
        return t.view(dtype) if t.dtype != dtype else t


class FusedOProjAllReduce:
    """`o_proj` + SUM all-reduce as ONE kernel per rank over NVLink (K8, `bsgl_oproj_allreduce`), in place of
    `F.linear` + `dist.all_reduce` (qwen2.py:195 + tp_runtime.py:28-33).

    One communicator serves every layer: a staging area for partial tiles, an output buffer and the flags.
    The memory comes from torch's symmetric-memory allocator when it is available (plumbing only: allocation,
    peer mapping and the NVSwitch MULTICAST mapping; `exchange = "nvls"`): partial tiles then stay in each rank's
    own staging and the tile's owner reduces them inside the switch (`multimem.ld_reduce`) and broadcasts the sum
    (`multimem.st`).  Without multicast support the buffers are cudaMalloc + CUDA IPC through the C ABI and
    partial tiles travel by peer stores (`exchange = "peer-stores"`).  The choice is made PER CALL.  Measured on 4 and
    8 B200s (profiles/r2_k8_trace_w4*.jsonl, _w8*.jsonl) neither exchange is bandwidth-bound at these sizes -- both
    are chains of flag round trips -- and the in-switch chain (flag -> multimem.ld_reduce -> multimem.st -> flag) is
    the longer one for large reductions (TP8, 10 MiB: 69-71 vs 60 us; TP4: 72 vs 53-57 us) but the shorter one for
    small reductions at 8 ranks (TP8, 2.6 MiB: 32-35 vs 37-38 us).  Default: peer stores, except in-switch below
    BSGL_K8_NVLS_MAX_BYTES (default 4 MiB) at world >= 8.  `BSGL_K8_EXCHANGE` = `nvls` (always in-switch), `peer` (IPC
    buffers, peer stores only) or `symm-peer` (symmetric memory, peer stores only) overrides it.

    Every rank must make the same sequence of calls, which the replicated scheduler guarantees; the call counter
    lives in device memory, so the call can sit inside a captured CUDA graph.  The returned tensor is a view of
    this rank's output buffer: valid until this communicator's next call (the reference's layer consumes it at
    once: residual add + RMSNorm create new tensors, qwen2.py:113-121)."""

    TILE = 128

    def __init__(self, rank: int, world_size: int, tp_group, hidden: int, max_tokens: int,
                 dtype: torch.dtype = torch.bfloat16, device=None, timeout_s: float = None):
        from baby_sglang_b200 import ops
        if hidden % self.TILE:
            raise ValueError(f"hidden {hidden} is not a multiple of {self.TILE}")
        if not 1 <= world_size <= 8:
            raise ValueError("fused o_proj all-reduce supports 1..8 ranks")
        self._ops, self.rank, self.world, self.group = ops, rank, world_size, tp_group
        self.hidden, self.max_tokens, self.dtype = hidden, max_tokens, dtype
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.timeout_s = float(os.environ.get("BSGL_K8_TIMEOUT_S", "30")) if timeout_s is None else float(timeout_s)
        self.max_tiles = -(-max_tokens // self.TILE) * (hidden // self.TILE)
        self.max_slots = -(-self.max_tiles // world_size)
        up = lambda n: -(-n // 1024) * 1024
        self.out_bytes = up(max_tokens * hidden * 2)
        staging_bytes = world_size * self.max_slots * self.TILE * self.TILE * 2
        self._off_out = up(staging_bytes)
        self._off_flags = self._off_out + self.out_bytes
        self._n_flag_words = self.max_slots * 8 + self.max_tiles + 2
        self._off_trace = self._off_flags + up(self._n_flag_words * 4)              # 8-byte aligned
        self.nbytes = self._off_trace + up(256 * 4 * 8)
        need = ctypes.c_int64(0)
        ops._lib.check(ops._lib.load().bsgl_oproj_flag_bytes(self.max_slots, self.max_tiles, ctypes.byref(need)),
                       "bsgl_oproj_flag_bytes")
        if self.nbytes - self._off_flags < need.value:
            raise RuntimeError("flag block smaller than the kernel's layout (bsgl_oproj_flag_bytes)")
        self._mine = None
        self._peers = []
        self._symm = None
        mc = 0
        base = None
        mode = os.environ.get("BSGL_K8_EXCHANGE", "")
        if world_size > 1 and mode != "peer":
            base, mc = self._alloc_symmetric()
        if base is None:
            base = self._alloc_ipc()
        if mode == "symm-peer":
            mc = 0
        self._nvls_always = mode == "nvls"
        self._nvls_small = mode in ("", "auto", "nvls-small")
        self._nvls_max_bytes = int(os.environ.get("BSGL_K8_NVLS_MAX_BYTES", str(4 << 20)))
        self.exchange = ("peer-stores (two-shot over NVLink peer memory)" +
                         ("; nvls (multimem.ld_reduce / multimem.st over the NVSwitch multicast mapping) for small "
                          "reductions at 8 ranks or on request" if mc else "")) if world_size > 1 else "local (world 1)"
        self.last_exchange = None
        self._staging = base
        self._out = [p + self._off_out for p in base]
        self._flags = [p + self._off_flags for p in base]
        self._mc_staging = mc if mc else 0
        self._mc_out = mc + self._off_out if mc else 0
        # built once: the pointer tables the C ABI takes and the output view (as_tensor costs tens of us)
        self._ptr_tables = ops.oproj_pointer_tables(self._staging, self._out, self._flags)
        self._out_view = _RawView.make(base[rank] + self._off_out, (max_tokens, hidden), dtype, self.device)
        self.calls = 0
        # handshake: one tiny call absorbs the lazy peer mapping / module load on every rank before timing matters
        if world_size > 1:
            a0 = torch.zeros(1, 64, dtype=dtype, device=self.device)
            w0 = torch.zeros(hidden, 64, dtype=dtype, device=self.device)
            self(a0, w0)
            torch.cuda.synchronize()
            dist.barrier(group=tp_group)

    # -- memory ------------------------------------------------------------------------------------------
    def _alloc_symmetric(self):
        """torch symmetric memory: same-size buffer on every rank, peer-mapped, plus the multicast mapping.
        Every rank leaves this function with the SAME answer: each computes a local verdict, then ONE all-reduce
        (outside the try, so that failing and succeeding ranks issue the same collective) decides for the group
        whether symmetric memory is used at all and whether the multicast mapping is.  (`rendezvous` is itself a
        collective: a rank that fails before reaching it -- an import error -- fails on every rank alike.)"""
        ok, base, mc, t, hdl = 1, None, 0, None, None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=self.device)
            t.zero_()
            torch.cuda.synchronize()
            hdl = symm_mem.rendezvous(t, self.group)
            base = [int(p) for p in hdl.buffer_ptrs]
            mc = int(hdl.multicast_ptr) if getattr(hdl, "has_multicast_support", False) else 0
            if len(base) != self.world or base[self.rank] != t.data_ptr():
                raise RuntimeError("unexpected symmetric-memory pointer table")
        except Exception as e:   # older torch / no fabric: the group falls back to cudaMalloc + IPC
            self._symm_error = repr(e)
            ok, mc = 0, 0
        verdict = torch.tensor([ok, 1 if mc else 0], device=self.device)
        dist.all_reduce(verdict, op=dist.ReduceOp.MIN, group=self.group)     # all ranks or none
        all_ok, all_mc = (int(x) for x in verdict.tolist())
        if not all_ok:
            return None, 0
        self._symm = (t, hdl)
        return base, (mc if all_mc else 0)

    def _alloc_ipc(self):
        ops = self._ops
        self._mine = ops.CommBuffer.alloc(self.nbytes)
        self._peers = [None] * self.world
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, self._mine.handle, group=self.group)
            for r, h in enumerate(handles):
                self._peers[r] = self._mine if r == self.rank else ops.CommBuffer.open(h, self.nbytes)
            dist.barrier(group=self.group)
        else:
            self._peers[0] = self._mine
        return [b.ptr for b in self._peers]

    # -- the call ----------------------------------------------------------------------------------------
    def __call__(self, attn_out: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
        """attn_out [T, K_r] (this rank's heads), weight [hidden, K_r] (this rank's o_proj columns) ->
        the all-reduced [T, hidden], bit-identical on every rank."""
        T = int(attn_out.shape[0])
        if T > self.max_tokens:
            raise ValueError(f"{T} tokens exceed the communicator's max_tokens {self.max_tokens}")
        if attn_out.dtype != self.dtype or weight.dtype != self.dtype or int(weight.shape[0]) != self.hidden:
            raise ValueError("dtype / hidden size differ from the communicator's")
        self.calls += 1
        # same decision on every rank: it depends on the call's shape only
        nvls = bool(self._mc_staging) and (self._nvls_always or
                                           (self._nvls_small and self.world >= 8 and
                                            T * self.hidden * 2 < self._nvls_max_bytes))
        self.last_exchange = "nvls" if nvls else ("peer-stores" if self.world > 1 else "local")
        self._ops.oproj_allreduce(attn_out, weight, self.rank, self.world, *self._ptr_tables, self.max_slots,
                                  self.max_tiles, self.out_bytes, self._mc_staging if nvls else 0,
                                  self._mc_out if nvls else 0, self.timeout_s)
        return self._out_view[:T]

    def trace(self):
        """Phase timestamps of the LAST call, per CTA, in microseconds after the first CTA's entry:
        rows {entry, last partial tile published, last owned tile delivered, exit} (0 where a CTA had no such work).
        Diagnostics only (one device->host copy)."""
        torch.cuda.synchronize()
        raw = _RawView.make(self._staging[self.rank] + self._off_trace, (256 * 4 * 2,), torch.int32, self.device)
        t = raw.cpu().view(torch.int64).view(256, 4)
        t = t[t[:, 0] > 0]
        if t.numel():
            t = t[t[:, 0] >= int(t[:, 0].max()) - 1_000_000]      # CTAs of the last launch only
        t0 = int(t[:, 0].min()) if t.numel() else 0
        rel = (t - t0).clamp(min=0).double() / 1e3
        rel[t == 0] = 0.0
        return rel

    def close(self) -> None:
        if self.world > 1 and dist.is_initialized():
            torch.cuda.synchronize()
            dist.barrier(group=self.group)
        for r, b in enumerate(self._peers):
            if b is not None and r != self.rank:
                b.release()
        if self._mine is not None:
            self._mine.release()
        self._peers, self._mine, self._symm = [], None, None


def shard_attention_heads(model: torch.nn.Module, rank: int, world_size: int, tp_group,
                          fused_allreduce: "FusedOProjAllReduce" = None) -> ParallelPlan:
    """Mutate `model` in place so this rank computes its KV heads only.

    Per attention module: keep rows [q heads of this rank | k heads | v heads] of the fused QKV weight and
    bias (layout `[Hq*D | Hkv*D | Hkv*D]`, qwen2.py:279-289), keep the matching input columns of `o_proj`,
    patch `num_heads / num_kv_heads / q_size / kv_size` (qwen2.py:157-161,279-281) and the backend-facing
    `attn.tp_*_head_num` (qwen2.py:302-304), and all-reduce `o_proj`'s output with a forward hook.  `head_dim`
    is untouched (the reference derives it from `hidden_size // num_heads` at construction, qwen2.py:159).
    """
    mods = list(_attention_modules(model))
    if not mods:
        raise ValueError("model has no attention modules of the expected shape")
    plan = plan_parallelism(mods[0].num_heads, mods[0].num_kv_heads, world_size)
    if plan.mode != "kv_head_tp":
        return plan
    for m in mods:
        Hq, Hkv, D = m.num_heads, m.num_kv_heads, m.head_dim
        G = Hq // Hkv
        kv_lo, kv_hi = rank * plan.kv_heads_per_rank, (rank + 1) * plan.kv_heads_per_rank
        q_lo, q_hi = kv_lo * G, kv_hi * G
        lin = m.qkv_proj.qkv_proj
        rows = torch.cat([torch.arange(q_lo * D, q_hi * D),
                          Hq * D + torch.arange(kv_lo * D, kv_hi * D),
                          (Hq + Hkv) * D + torch.arange(kv_lo * D, kv_hi * D)]).to(lin.weight.device)
        with torch.no_grad():
            lin.weight = torch.nn.Parameter(lin.weight.index_select(0, rows).contiguous(), requires_grad=False)
            if lin.bias is not None:
                lin.bias = torch.nn.Parameter(lin.bias.index_select(0, rows).contiguous(), requires_grad=False)
            lin.out_features = int(rows.numel())
            o = m.o_proj
            o.weight = torch.nn.Parameter(o.weight[:, q_lo * D:q_hi * D].contiguous(), requires_grad=False)
            o.in_features = (q_hi - q_lo) * D
        m.num_heads, m.num_kv_heads = q_hi - q_lo, kv_hi - kv_lo
        m.q_size, m.kv_size = m.num_heads * D, m.num_kv_heads * D
        for holder in (m.qkv_proj,):
            if hasattr(holder, "num_heads"):
                holder.num_heads, holder.num_kv_heads = m.num_heads, m.num_kv_heads
                holder.q_size, holder.kv_size = m.q_size, m.kv_size
                holder.total_size = m.q_size + 2 * m.kv_size
        m.attn.tp_q_head_num, m.attn.tp_k_head_num, m.attn.tp_v_head_num = m.num_heads, m.num_kv_heads, m.num_kv_heads
        if fused_allreduce is None:
            m.o_proj.register_forward_hook(lambda _mod, _inp, out, g=tp_group: tp_all_reduce_sum(out, g))
        else:
            m.o_proj.forward = _fused_oproj_forward(m.o_proj, fused_allreduce, tp_group)
    return plan


def _mlp_modules(model: torch.nn.Module):
    """Every module shaped like the reference's Qwen2MLP (qwen2.py:198-211): a fused `gate_up_proj.gate_up_proj`
    linear of 2*I rows ([gate | up], split by `chunk(2)` at qwen2.py:221) and a bias-free `down_proj`."""
    for m in model.modules():
        inner = getattr(getattr(m, "gate_up_proj", None), "gate_up_proj", None)
        if isinstance(inner, torch.nn.Linear) and isinstance(getattr(m, "down_proj", None), torch.nn.Linear):
            yield m


def shard_mlp(model: torch.nn.Module, rank: int, world_size: int, tp_group,
              fused_allreduce: "FusedOProjAllReduce" = None) -> int:
    """Tensor-parallel SwiGLU MLP (second half), on the unchanged model: rank r keeps
    intermediate columns [r*I/t, (r+1)*I/t) -- rows `cat(gate_shard, up_shard)` of the fused gate/up weight, so the
    reference's `chunk(2, dim=-1)` (qwen2.py:221) still separates gate from up -- and the matching input columns of
    `down_proj` (bias-free, qwen2.py:206), whose partial output is SUM all-reduced: by K8 when a communicator is
    given (`down_proj` is the same `[T, K_r] . [hidden, K_r]^T` + SUM as `o_proj`, so the layer's second collective
    rides the same fused kernel; the view it returns is consumed by the next layer's residual add before the next
    K8 call, qwen2.py:113-121), else by a forward hook on NCCL.  Returns the
    per-rank intermediate size (0 if nothing was sharded: world 1, no MLP modules, or I % world != 0)."""
    if world_size <= 1:
        return 0
    per_rank = 0
    for m in _mlp_modules(model):
        lin, down = m.gate_up_proj.gate_up_proj, m.down_proj
        inter = down.in_features
        if inter % world_size or lin.out_features != 2 * inter:
            return 0
        per_rank = inter // world_size
        lo, hi = rank * per_rank, (rank + 1) * per_rank
        rows = torch.cat([torch.arange(lo, hi), inter + torch.arange(lo, hi)]).to(lin.weight.device)
        with torch.no_grad():
            lin.weight = torch.nn.Parameter(lin.weight.index_select(0, rows).contiguous(), requires_grad=False)
            if lin.bias is not None:
                lin.bias = torch.nn.Parameter(lin.bias.index_select(0, rows).contiguous(), requires_grad=False)
            lin.out_features = 2 * per_rank
            down.weight = torch.nn.Parameter(down.weight[:, lo:hi].contiguous(), requires_grad=False)
            down.in_features = per_rank
        if hasattr(m.gate_up_proj, "intermediate_size"):
            m.gate_up_proj.intermediate_size = per_rank
        if fused_allreduce is not None and per_rank % 64 == 0 and down.out_features == fused_allreduce.hidden \
                and down.bias is None:
            down.forward = _fused_oproj_forward(down, fused_allreduce, tp_group)
        else:
            down.register_forward_hook(lambda _mod, _inp, out, g=tp_group: tp_all_reduce_sum(out, g))
    return per_rank


def _fused_oproj_forward(lin: torch.nn.Linear, comm: "FusedOProjAllReduce", tp_group):
    """`forward` replacement for a bias-free row-parallel linear (`o_proj`, `down_proj`): K8 in place of
    `F.linear` + `tp_all_reduce_sum`.  Batches larger than the communicator was sized for (a 16 K-token prefill
    against a decode-sized `max_tokens`) run as consecutive `max_tokens`-row blocks, each copied out of the
    communicator's single output buffer before the next block's call reuses it; every rank cuts the same blocks
    (the token count is the same on all ranks), so the call sequence stays identical across the group."""
    def forward(x: torch.Tensor) -> torch.Tensor:
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        T, cap = int(x2.shape[0]), comm.max_tokens
        if T <= cap:
            return comm(x2, lin.weight).view(*lead, comm.hidden)
        out = torch.empty((T, comm.hidden), dtype=x2.dtype, device=x2.device)
        for lo in range(0, T, cap):
            hi = min(T, lo + cap)
            out[lo:hi].copy_(comm(x2[lo:hi], lin.weight))
        return out.view(*lead, comm.hidden)
    return forward


def kv_pool_config(config, plan: ParallelPlan):
    """Config copy for the (unchanged) scheduler's pool construction: it sizes the KV pool with
    `num_key_value_heads` and `hidden_size // num_attention_heads` (scheduler.py:50-57), so only the KV-head count
    shrinks under TP; everything else — and therefore head_dim — stays."""
    cfg = copy.copy(config)
    if plan.mode == "kv_head_tp":
        cfg.num_key_value_heads = plan.kv_heads_per_rank
    return cfg


# ------------------------------------------------------------------------------------------------------
# request-batch data parallelism
# ------------------------------------------------------------------------------------------------------
def dp_partition(n_requests: int, rank: int, world_size: int) -> List[int]:
    """Round-robin share of request indices for this replica."""
    return list(range(rank, n_requests, max(world_size, 1)))


def dp_gather_outputs(local_outputs: Sequence[Sequence[int]], n_requests: int, rank: int, world_size: int,
                      tp_group) -> List[List[int]]:
    """Reassemble per-request token lists from all replicas in the original request order."""
    if world_size <= 1 or not dist.is_initialized():
        return [list(o) for o in local_outputs]
    gathered: List[object] = [None] * world_size
    dist.all_gather_object(gathered, [list(o) for o in local_outputs], group=tp_group)
    out: List[List[int]] = [[] for _ in range(n_requests)]
    for r, part in enumerate(gathered):
        for j, idx in enumerate(dp_partition(n_requests, r, world_size)):
            out[idx] = part[j]
    return out
