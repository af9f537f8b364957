#!/usr/bin/env python
"""bench.py — headline benchmark of the paged-attention hot path (contract in the task brief).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload ns|c2|c4|c5]

Workload `ns` (default; BASELINE.json north star): Qwen2.5-7B attention geometry (28 q / 4 kv heads,
D=128, 28 layers), paged decode, batch 256, context 4096, page size 1, bf16, fragmented page table.
One STEP = one decode step of the hot path: the K4 metadata build (once) and, for each of the 28
layers, K1 kv_store of the new token + K2 paged split-KV attention + K2c merge.

`value`  algorithmic GB/s (SURVEY.md §8d formula) with q/k/v already resident in HBM.
`e2e`    the same step driven through the reference-facing plugin call
         (`SimpleAttentionBackend.forward`, drop-in module) from pinned HOST buffers: per layer
         q/k/v are copied host->device and the attention output device->host inside the timed region.
`roofline`  the dominant kernel (paged_attention_kernel) timed alone with CUDA events.
`cpu_baseline`  the oracle (CPU port of the reference path) on a bounded sample of the same workload.

N > 1: request-batch data parallelism (independent requests per rank, no collective in the data
path) -> "scaling": "weak".  Workload c5 (Qwen2.5-32B KV-head TP with the o_proj
all-reduce) is the sharded-with-collective configuration.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

GEOMETRY = {  # name: (Hq, Hkv, D, layers, hidden)
    "qwen2.5-1.5b": (12, 2, 128, 28, 1536),
    "qwen2.5-7b": (28, 4, 128, 28, 3584),
    "qwen2.5-32b": (40, 8, 128, 64, 5120),
}
WORKLOADS = {
    # name: (model, batch, ctx_lo, ctx_hi, layers_resident)
    "ns": ("qwen2.5-7b", 256, 4096, 4096, 28),
    "c2": ("qwen2.5-1.5b", 256, 2048, 2048, 28),
    "c4": ("qwen2.5-7b", 512, 4096, 32768, 1),
    "c5": ("qwen2.5-32b", 1024, 512, 8192, 1),
}
METRIC = "paged decode attn HBM GB/s, Qwen2.5-7B bs=256 ctx=4K"


def peaks():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def traffic_from_profiles(key):
    """dram bytes per launch of the dominant kernel, from the committed ncu --set full capture (or None)."""
    path = os.path.join(REPO, "profiles", "r1_traffic.json")
    try:
        with open(path) as f:
            return int(json.load(f)[key]["dram_bytes"])
    except Exception:
        return None


def tensor_peak():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            return float(json.load(f)["bf16_tflops"]), "measured burst (MEASURED_PEAKS.json bf16_tflops)"
    return 1590.0, "fallback (B200_PROFILING.md 1.59 PFLOP/s)"


def extend_flops(Hq, D, prefix, ext):
    """QK^T and PV, unmasked MACs only."""
    return 4.0 * Hq * D * sum(e * (p + (e + 1) / 2.0) for p, e in zip(prefix, ext))


def decode_bytes(sum_seq, B, Hq, Hkv, D, e=2):
    """K+V rows once per KV head + int32 page-table entries + q read + out write."""
    return sum_seq * Hkv * D * 2 * e + 4 * sum_seq + 2 * B * Hq * D * e


def kv_store_bytes(T, Hkv, D, e=2):
    return 2 * (2 * T * Hkv * D * e) + 4 * T


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# workload construction
# ------------------------------------------------------------------------------------------------
class _Layer:
    """What the backend reads off the model's attention layer (qwen2.py:302-307)."""

    def __init__(self, layer_id, Hq, Hkv, D):
        self.layer_id = layer_id
        self.scaling = D ** -0.5
        self.head_dim = D
        self.tp_q_head_num = Hq
        self.tp_k_head_num = Hkv
        self.tp_v_head_num = Hkv


def build_workload(name, device, seed, Hkv_override=None, Hq_override=None):
    """Pools (drop-in classes), fragmented page table and per-layer q/k/v for one decode step."""
    sys.path.insert(0, os.path.join(REPO, "baby_sglang_b200", "dropin"))
    import memory_pool  # drop-in

    model, B, lo, hi, L = WORKLOADS[name]
    Hq, Hkv, D, _, hidden = GEOMETRY[model]
    if Hkv_override:
        Hq, Hkv = Hq_override, Hkv_override
    g = torch.Generator(device=device).manual_seed(seed)
    if lo == hi:
        seq = torch.full((B,), lo, dtype=torch.int64, device=device)
    else:
        seq = torch.randint(lo, hi + 1, (B,), generator=g, device=device)
    C = int(seq.max().item()) + 64
    S = int(seq.sum().item()) + 64 * B
    kv_pool = memory_pool.MHATokenToKVPool(size=S, dtype=torch.bfloat16, head_num=Hkv, head_dim=D, layer_num=L,
                                           device=str(device))
    for l in range(L):           # finite synthetic KV (torch.empty is garbage)
        kv_pool.k_buffer[l].normal_()
        kv_pool.v_buffer[l].normal_()
    r2t_pool = memory_pool.ReqToTokenPool(size=B, max_context_len=C, device=str(device))
    perm = (torch.randperm(S, generator=g, device=device) + 1).to(torch.int32)   # fragmented slots
    ends = torch.cumsum(seq, 0)
    starts = ends - seq
    col = torch.arange(C, device=device)[None, :]
    src = (starts[:, None] + col).clamp_(max=S - 1)
    table = perm[src]
    table[col >= seq[:, None]] = 0
    r2t_pool.req_to_token.copy_(table)
    rpi = torch.arange(B, dtype=torch.int64, device=device)
    out_cache_loc = r2t_pool.req_to_token[rpi, seq - 1].contiguous()           # slot of the token being decoded
    w = dict(name=name, model=model, B=B, Hq=Hq, Hkv=Hkv, D=D, L=L, hidden=hidden, seq=seq, C=C, S=S,
             kv_pool=kv_pool, r2t_pool=r2t_pool, rpi=rpi, out_cache_loc=out_cache_loc,
             sum_seq=int(seq.sum().item()), layers=[_Layer(l, Hq, Hkv, D) for l in range(L)])
    return w


def measure_extend_c3(device, reps=20):
    """Config C3: Qwen2.5-7B geometry, B=16, prefix_i ~ U[0,8192] (first two forced to 0 and
    8192), 512 new tokens each, bf16, fragmented page table; K3 (tcgen05) timed alone with CUDA events.
    The 151 MB KV working set is re-read across launches from HBM/L2 mix; the kernel is tensor-bound."""
    from baby_sglang_b200 import ops
    Hq, Hkv, D, _, _ = GEOMETRY["qwen2.5-7b"]
    B, ext_len, max_prefix = 16, 512, 8192
    g = torch.Generator().manual_seed(0)
    prefix = torch.randint(0, max_prefix + 1, (B,), generator=g)
    prefix[0], prefix[1] = 0, max_prefix
    ext = torch.full((B,), ext_len, dtype=torch.int64)
    seq = prefix + ext
    C = int(seq.max()) + 64
    S = int(seq.sum()) + 1024
    perm = (torch.randperm(S, generator=g) + 1).to(torch.int32)
    r2t = torch.zeros(B, C, dtype=torch.int32)
    off = 0
    for i, n in enumerate(seq.tolist()):
        r2t[i, :n] = perm[off:off + n]
        off += n
    T = int(ext.sum())
    dt = torch.bfloat16
    k_slab = torch.randn(S + 1, Hkv, D, device=device, dtype=dt)
    v_slab = torch.randn(S + 1, Hkv, D, device=device, dtype=dt)
    q = torch.randn(T, Hq, D, device=device, dtype=dt)
    r2t, rpi = r2t.to(device), torch.arange(B, device=device)
    seq_d, pre_d = seq.to(device), prefix.to(device)
    md = ops.build_extend_metadata(r2t, rpi, seq_d, pre_d, T, Hkv, row_plan=False)
    out = torch.empty(T, Hq * D, device=device, dtype=dt)
    for _ in range(3):
        ops.extend_attention_tc(q, k_slab, v_slab, r2t, rpi, seq_d, pre_d, md.qo_indptr, D ** -0.5, out=out,
                                req_order=md.heavy_order)
    torch.cuda.synchronize()
    times = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ops.extend_attention_tc(q, k_slab, v_slab, r2t, rpi, seq_d, pre_d, md.qo_indptr, D ** -0.5, out=out,
                                req_order=md.heavy_order)
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    ms = statistics.mean(times)
    fl = extend_flops(Hq, D, prefix.tolist(), ext.tolist())
    peak, src = tensor_peak()
    ach = fl / ms / 1e9
    return {"bound": "tensor", "kernel": "extend_attention_kernel (tcgen05/TMEM)", "achieved": round(ach, 1),
            "peak": peak, "unit": "TFLOP/s", "frac": round(ach / peak, 4), "peak_source": src,
            "flops_per_launch": fl, "avg_launch_ms": round(ms, 5), "traffic": traffic_from_profiles("extend_c3"),
            "workload": "qwen2.5-7b extend (config C3): B=16, prefix U[0,8192], 512 new tokens each, bf16"}


def run_b200(args, rank, world, device):
    from baby_sglang_b200 import ops
    model_name = WORKLOADS[args.workload][0]
    Hq_full, Hkv_full = GEOMETRY[model_name][0], GEOMETRY[model_name][1]
    # c5: KV-head tensor parallelism (rank r owns Hkv/t KV heads of EVERY request; identical page table on all
    # ranks; one SUM all-reduce of the o_proj partial per layer).  Everything else: request-batch DP.
    tp = args.workload == "c5" and world > 1 and Hkv_full % world == 0
    # K3 is timed alone and first, like the burst peak it is compared with (after the 28-layer decode loop the
    # GPU sits at its power cap and the same kernel runs ~10 % slower).
    extend = measure_extend_c3(device) if (world == 1 and rank == 0) else None
    if tp:
        w = build_workload(args.workload, device, seed=1234, Hkv_override=Hkv_full // world,
                           Hq_override=Hq_full // world)
    else:
        w = build_workload(args.workload, device, seed=1234 + rank)
    import attn_backend   # drop-in modules (sys.path set by build_workload)
    import forward_batch
    import tp_runtime
    B, Hq, Hkv, D, L = w["B"], w["Hq"], w["Hkv"], w["D"], w["L"]
    dt = torch.bfloat16
    with_oproj = args.workload == "c5"
    tp_group = torch.distributed.group.WORLD if world > 1 else None
    if with_oproj:   # this rank's column shard of o_proj (bias-free, qwen2.py:166): partial sums are exact
        w_o = torch.randn(w["hidden"], Hq * D, device=device, dtype=dt) * 0.02

    fused_comm = None
    if tp and args.oproj == "fused":   # K8: one kernel per rank instead of GEMM + NCCL call
        fused_comm = tp_runtime.FusedOProjAllReduce(rank, world, tp_group, hidden=w["hidden"], max_tokens=B)

    def after_attention(out):
        if not with_oproj:
            return out
        if fused_comm is not None:
            return fused_comm(out, w_o)
        partial = torch.nn.functional.linear(out, w_o)
        return tp_runtime.tp_all_reduce_sum(partial, tp_group) if tp else partial
    # per-layer activations of one step, device-resident (for `value`) and pinned host copies (for `e2e`)
    qkv_dev = [torch.randn(B, (Hq + 2 * Hkv) * D, device=device, dtype=dt) for _ in range(L)]
    qkv_host = [t.cpu().pin_memory() for t in qkv_dev]
    out_host = [torch.empty(B, w["hidden"] if args.workload == "c5" else Hq * D, dtype=dt).pin_memory()
                for _ in range(L)]
    ids_host = torch.zeros(B, dtype=torch.int64).pin_memory()
    rpi_host = w["rpi"].cpu().pin_memory()
    seq_host = w["seq"].cpu().pin_memory()
    loc_host = w["out_cache_loc"].cpu().pin_memory()
    ids_dev = torch.zeros(B, dtype=torch.int64, device=device)

    def split(qkv):
        q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        # q is made contiguous by RoPE in the real model (qwen2.py:268); k likewise; v stays a strided view
        return q.contiguous().view(B, Hq, D), k.contiguous().view(B, Hkv, D), v.view(B, Hkv, D)

    qkv_split_dev = [split(t) for t in qkv_dev]

    def step_resident():
        fb = forward_batch.SimplifiedForwardBatch.create_decode_batch(
            input_ids=ids_dev, req_pool_indices=w["rpi"], seq_lens=w["seq"], out_cache_loc=w["out_cache_loc"],
            req_to_token_pool=w["r2t_pool"], token_to_kv_pool=w["kv_pool"])
        out = None
        for l in range(L):
            q, k, v = qkv_split_dev[l]
            out = after_attention(attn_backend.SimpleAttentionBackend.forward(q, k, v, w["layers"][l], fb, True))
        return out

    from baby_sglang_b200.host_staging import HostStagedAttention
    staged = HostStagedAttention(Hq, Hkv, D, device)

    def step_e2e():
        rpi = rpi_host.to(device, non_blocking=True)
        seq = seq_host.to(device, non_blocking=True)
        loc = loc_host.to(device, non_blocking=True)
        ids = ids_host.to(device, non_blocking=True)
        fb = forward_batch.SimplifiedForwardBatch.create_decode_batch(
            input_ids=ids, req_pool_indices=rpi, seq_lens=seq, out_cache_loc=loc,
            req_to_token_pool=w["r2t_pool"], token_to_kv_pool=w["kv_pool"])
        # uploads of layer l+1 and downloads of layer l overlap layer l's kernels (host_staging.py)
        staged.forward_layers(qkv_host, out_host, w["layers"], fb, after=after_attention if with_oproj else None)
        return out_host[L - 1]

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = ops.LAUNCH_COUNT
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=device)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t.item())
        return ms / steps, ops.LAUNCH_COUNT - n0

    step_bytes = L * (decode_bytes(w["sum_seq"], B, Hq, Hkv, D) + kv_store_bytes(B, Hkv, D))
    sampler = ClockSampler(device.index or 0)
    if rank == 0:
        sampler.start()
    ms_step, launches = timed(step_resident, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e, _ = timed(step_e2e, args.steps, args.warmup)

    # ---- roofline of the dominant kernel: paged_attention_kernel alone, CUDA events per launch
    md = ops.build_decode_metadata(w["r2t_pool"].req_to_token, w["rpi"], w["seq"], Hkv)
    ws = ops.partial_buffers(device, md.plan.n_ctas, Hq, Hkv, D)
    r2t = w["r2t_pool"].req_to_token
    scale = D ** -0.5
    outbuf = torch.empty(B, Hq * D, device=device, dtype=dt)
    reps = max(3, args.steps // 2)
    for l in range(L):
        ops.paged_attention(qkv_split_dev[l][0], *w["kv_pool"].get_kv_buffer(l), r2t, md.plan, scale, out=outbuf,
                            workspace=ws, merge=False)
    torch.cuda.synchronize()
    evs = []
    for _ in range(reps):
        for l in range(L):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            ops.paged_attention(qkv_split_dev[l][0], *w["kv_pool"].get_kv_buffer(l), r2t, md.plan, scale, out=outbuf,
                                workspace=ws, merge=False)
            b.record()
            evs.append((a, b))
    torch.cuda.synchronize()
    k_ms = statistics.mean(a.elapsed_time(b) for a, b in evs)
    k_bytes = decode_bytes(w["sum_seq"], B, Hq, Hkv, D)
    peak, peak_src = peaks()
    achieved = k_bytes / k_ms / 1e6

    allreduce = None
    if tp:
        buf = torch.randn(B, w["hidden"], device=device, dtype=dt)
        for _ in range(5):
            tp_runtime.tp_all_reduce_sum(buf, tp_group)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_ar = 20
        a.record()
        for _ in range(n_ar):
            tp_runtime.tp_all_reduce_sum(buf, tp_group)
        b.record()
        barrier()
        t = torch.tensor([a.elapsed_time(b) / n_ar], device=device)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ar_ms = float(t.item())
        nbytes = buf.numel() * buf.element_size()
        allreduce = {"bytes": nbytes, "avg_ms": round(ar_ms, 5),
                     "busbw_GBps": round(2 * (world - 1) / world * nbytes / ar_ms / 1e6, 1),
                     "backend": "NCCL over NVLink/NVSwitch (torch.distributed all_reduce SUM)"}
    used_fused = fused_comm is not None
    if used_fused:
        fused_comm.close()      # collective (barrier): every rank, before the others leave
    if rank != 0:
        return
    cpu = cpu_baseline(args.workload, seconds=12.0)
    value = world * step_bytes / ms_step / 1e6
    e2e_v = world * step_bytes / ms_e2e / 1e6
    h2d = sum(t.numel() * t.element_size() for t in qkv_host) + 8 * B * 3 + 4 * B
    d2h = sum(t.numel() * t.element_size() for t in out_host)
    line = {
        "metric": METRIC if args.workload == "ns" else f"paged decode attn HBM GB/s, workload {args.workload}",
        "value": round(value, 1), "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "strong" if tp else "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"{w['model']} paged GQA decode ({Hq}q/{Hkv}kv{' per rank' if tp else ''}, D={D}), "
                               f"batch {B}{'' if tp else '/GPU'}, "
                               f"ctx {WORKLOADS[args.workload][2]}..{WORKLOADS[args.workload][3]}, page_size 1, "
                               f"{L} layers/step, fragmented page table",
                   "parallelism": (f"KV-head TP x{world}: o_proj partial + "
                                   + ("all-reduce fused in one kernel over NVLink peer memory (K8)" if used_fused
                                      else "NCCL all-reduce per layer") if tp else
                                   "request-batch DP (no collective)" if world > 1 else "single GPU"),
                   "l2": f"inputs > L2: {k_bytes / 1e9:.2f} GB of KV per layer, {L} distinct layers per step",
                   "step": "K4 metadata + per layer (K1 kv_store, K2 paged attention, K2c merge"
                           + (", o_proj shard GEMM" + (", all-reduce" if tp else "") if with_oproj else "") + ")"},
        "allreduce": allreduce,
        "decode_tok_per_s_attention_path": round(world * B / (ms_step / 1e3), 1),
        "e2e": {"value": round(e2e_v, 1), "unit": "GB/s", "ms_per_step": round(ms_e2e, 4),
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "HostStagedAttention.forward_layers (dropin attn_backend.SimpleAttentionBackend.forward per layer) from pinned host q/k/v, copies on side streams"},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "paged_attention_kernel", "achieved": round(achieved, 1),
                     "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4), "peak_source": peak_src,
                     "bytes_per_launch": k_bytes, "avg_launch_ms": round(k_ms, 5),
                     "traffic": args.traffic if args.traffic is not None else
                     (traffic_from_profiles(args.workload) if world == 1 else None)},
        "extend_roofline": extend,
        "cpu_baseline": cpu,
    }
    if used_fused:
        line["oproj"] = {"impl": "K8 oproj_allreduce_kernel (tcgen05 GEMM + owner-reduce over NVLink peer memory)",
                         "replaces": "F.linear + NCCL all-reduce (bench.py --oproj nccl)"}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (port of the reference's torch path) on the box's host cores
# ------------------------------------------------------------------------------------------------
def cpu_baseline(workload, seconds=12.0, sample_batch=8):
    from oracle import paged_attention_oracle as O
    model, B, lo, hi, L = WORKLOADS[workload]
    Hq, Hkv, D, _, _ = GEOMETRY[model]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    g = torch.Generator().manual_seed(0)
    Bs = min(sample_batch, B)
    seq = torch.full((Bs,), lo, dtype=torch.int64) if lo == hi else torch.randint(lo, hi + 1, (Bs,), generator=g)
    S = int(seq.sum()) + 16
    C = int(seq.max()) + 8
    # fp32 on CPU: the reference's own CPU dtype policy (model_loader.py:56-64)
    k_slab = torch.randn(S + 1, Hkv, D, generator=g)
    v_slab = torch.randn(S + 1, Hkv, D, generator=g)
    perm = (torch.randperm(S, generator=g) + 1).to(torch.int32)
    r2t = torch.zeros(Bs, C, dtype=torch.int32)
    off = 0
    for i, n in enumerate(seq.tolist()):
        r2t[i, :n] = perm[off:off + n]
        off += n
    rpi = torch.arange(Bs)
    loc = r2t[rpi, seq - 1].contiguous()
    q = torch.randn(Bs, Hq, D, generator=g)
    k = torch.randn(Bs, Hkv, D, generator=g)
    v = torch.randn(Bs, Hkv, D, generator=g)
    nbytes = decode_bytes(int(seq.sum()), Bs, Hq, Hkv, D) + kv_store_bytes(Bs, Hkv, D)
    O.backend_forward(q, k, v, k_slab, v_slab, r2t, rpi, seq, loc, False)      # warm-up
    t0 = time.perf_counter()
    n = 0
    while True:
        O.backend_forward(q, k, v, k_slab, v_slab, r2t, rpi, seq, loc, False)
        n += 1
        dt = time.perf_counter() - t0
        if dt >= seconds or n >= 200:
            break
    return {"value": round(n * nbytes / dt / 1e9, 4), "unit": "GB/s", "cores": cores, "kind": "port",
            "ms_per_layer_call": round(dt / n * 1e3, 3),
            "sample": f"{n} calls of one layer (kv_store + decode attention) on {Bs} of the {B} requests, "
                      f"ctx {lo}..{hi}, fp32 (reference CPU dtype), oracle/paged_attention_oracle.py"}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path = the oracle port (the
    Python reference cannot travel to the GPU box), all host threads, bounded sample per step."""
    if rank != 0:
        return
    per = max(2.0, min(20.0, 90.0 / max(1, args.steps + args.warmup)))
    vals = []
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(args.workload, seconds=per)
        if i >= args.warmup:
            vals.append(r)
    v = statistics.mean(x["value"] for x in vals)
    model, B, lo, hi, L = WORKLOADS[args.workload]
    Hq, Hkv, D, _, _ = GEOMETRY[model]
    line = {
        "impl": "reference",
        "metric": METRIC if args.workload == "ns" else f"paged decode attn HBM GB/s, workload {args.workload}",
        "value": round(v, 4), "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(statistics.mean(x["ms_per_layer_call"] for x in vals), 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{model} paged GQA decode ({Hq}q/{Hkv}kv, D={D}), ctx {lo}..{hi}, page_size 1; "
                               f"CPU port of the reference path on a bounded sample"},
        "cpu_baseline": {**vals[-1], "value": round(v, 4)},
        "e2e": {"value": round(v, 4), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="ns", choices=sorted(WORKLOADS))
    ap.add_argument("--oproj", default="fused", choices=["fused", "nccl"],
                    help="c5 under TP: K8 (o_proj GEMM with the all-reduce fused over NVLink peer memory) or "
                         "F.linear + NCCL all-reduce")
    ap.add_argument("--traffic", type=float, default=None,
                    help="dram bytes per launch of the dominant kernel from an ncu --set full capture (profiles/)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.distributed.init_process_group("nccl", device_id=device)
    try:
        with torch.no_grad():
            run_b200(args, rank, world, device)
    finally:
        if world > 1:
            torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
