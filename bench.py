#!/usr/bin/env python
"""bench.py — headline benchmark of the paged-attention hot path (contract in the task brief).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload ns|c2|c4|c5]

Workload `ns` (default; BASELINE.json north star): Qwen2.5-7B attention geometry (28 q / 4 kv heads,
D=128, 28 layers), paged decode, batch 256, context 4096, page size 1, bf16, fragmented page table.
One STEP = one decode step of the hot path: the K4 metadata build (once) and, for each of the 28
layers, K1 kv_store of the new token + K2 paged split-KV attention + K2c merge (under KV-head TP also the
rank's o_proj shard with the SUM all-reduce, K8).

`value`        algorithmic GB/s (SURVEY.md §8d formula) with q/k/v already resident in HBM.
`e2e`          the same step driven through the reference-facing plugin call (`SimpleAttentionBackend.forward`,
               drop-in module) from pinned HOST buffers: per layer q/k/v are copied host->device and the attention
               output device->host inside the timed region.
`roofline`     the dominant kernel (paged_attention_kernel) timed alone with CUDA events.
`parity_check` sampled rows of the timed configuration against the oracle (tolerance 2e-2), per workload.
`decode_tok_per_s`  the other half of the metric: the reference's UNCHANGED scheduler + Qwen2.5-7B-geometry model
               (random init) on the drop-in path, B=256, ctx 4096, 64 greedy decode steps; tok/s = 256 / median step.
`sharded`      config C5 (Qwen2.5-32B, B=1024 ragged, KV-head TP = N with the o_proj all-reduce) in the same line.
`small_kernels` K1 / K5 / K6 rooflines.
`cpu_baseline` the reference's own modules (when a checkout is reachable, kind "reference") or the oracle port on
               the box's host cores, bounded sample.

N > 1: the sharding is `tp_runtime.plan_parallelism`'s: KV-head tensor parallelism where Hkv % N == 0 (7B: N = 2, 4;
32B: N <= 8) with the o_proj partial + SUM all-reduce inside the timed step ("scaling": "strong"); request-batch data
parallelism otherwise (7B at N = 8; independent requests per rank, no collective, "scaling": "weak").
"""
from __future__ import annotations

import argparse
import gc
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

GEOMETRY = {  # name: (Hq, Hkv, D, layers, hidden, intermediate, vocab)
    "qwen2.5-1.5b": (12, 2, 128, 28, 1536, 8960, 151936),
    "qwen2.5-7b": (28, 4, 128, 28, 3584, 18944, 152064),
    "qwen2.5-32b": (40, 8, 128, 64, 5120, 27648, 152064),
}
WORKLOADS = {
    # name: (model, batch, ctx_lo, ctx_hi, layers_resident)
    "ns": ("qwen2.5-7b", 256, 4096, 4096, 28),
    "c2": ("qwen2.5-1.5b", 256, 2048, 2048, 28),
    "c4": ("qwen2.5-7b", 512, 4096, 32768, 1),
    "c5": ("qwen2.5-32b", 1024, 512, 8192, 1),
}
METRIC = "paged decode attn HBM GB/s & decode tok/s, Qwen2.5-7B bs=256 ctx=4K"
TOL = 2e-2


def _peaks_file():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            return json.load(f)
    return None


def peaks():
    p = _peaks_file()
    if p:
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def tensor_peak():
    p = _peaks_file()
    if p:
        return float(p["bf16_tflops"]), "measured burst (MEASURED_PEAKS.json bf16_tflops)"
    return 1590.0, "fallback (B200_PROFILING.md 1.59 PFLOP/s)"


def traffic_from_profiles(key, with_source=False):
    """dram bytes per launch of the dominant kernel, from the committed ncu --set full capture (or None).  The figure
    is a CAPTURE, not something this run measured: `with_source` also returns where it came from (file, the capture
    it summarises, the commit that last touched the file) so that a stale figure is visible in the line."""
    for name in ("r3_traffic.json", "r2_traffic.json", "r1_traffic.json"):
        try:
            with open(os.path.join(REPO, "profiles", name)) as f:
                entry = json.load(f)[key]
            val = int(entry["dram_bytes"])
        except Exception:
            continue
        if not with_source:
            return val
        try:
            commit = subprocess.run(["git", "log", "-1", "--format=%h", "--", f"profiles/{name}"], cwd=REPO,
                                    capture_output=True, text=True, timeout=10).stdout.strip() or None
        except Exception:
            commit = None
        return val, {"file": f"profiles/{name}", "capture": entry.get("source"), "file_commit": commit,
                     "note": "committed ncu --set full capture, not measured by this run"}
    return (None, None) if with_source else None


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


def _free():
    gc.collect()
    torch.cuda.empty_cache()


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


def _dropin():
    d = os.path.join(REPO, "baby_sglang_b200", "dropin")
    if d not in sys.path:
        sys.path.insert(0, d)
    import attn_backend
    import forward_batch
    import memory_pool
    import tp_runtime
    return attn_backend, forward_batch, memory_pool, tp_runtime


def build_workload(name, device, seed, Hkv_override=None, Hq_override=None):
    """Pools (drop-in classes), fragmented page table and per-layer q/k/v for one decode step."""
    _, _, memory_pool, _ = _dropin()
    model, B, lo, hi, L = WORKLOADS[name]
    Hq, Hkv, D, _, hidden = GEOMETRY[model][:5]
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
    return dict(name=name, model=model, B=B, Hq=Hq, Hkv=Hkv, D=D, L=L, hidden=hidden, seq=seq, C=C, S=S,
                kv_pool=kv_pool, r2t_pool=r2t_pool, rpi=rpi, out_cache_loc=out_cache_loc,
                sum_seq=int(seq.sum().item()), layers=[_Layer(l, Hq, Hkv, D) for l in range(L)])


def measure_extend_c3(device, reps=20):
    """Config C3: Qwen2.5-7B geometry, B=16, prefix_i ~ U[0,8192] (first two forced to 0 and
    8192), 512 new tokens each, bf16, fragmented page table; K3 (tcgen05) timed alone with CUDA events, then a
    sample of its rows checked against the oracle.  The 151 MB KV working set is re-read across launches from an
    HBM/L2 mix; the kernel is tensor-bound."""
    from baby_sglang_b200 import ops
    from oracle import sampled_check as SC
    Hq, Hkv, D = GEOMETRY["qwen2.5-7b"][:3]
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
                                req_order=md.heavy_order, work_counter=md.work_counter)
    torch.cuda.synchronize()
    times = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ops.extend_attention_tc(q, k_slab, v_slab, r2t, rpi, seq_d, pre_d, md.qo_indptr, D ** -0.5, out=out,
                                req_order=md.heavy_order, work_counter=md.work_counter)
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    ms = statistics.mean(times)
    fl = extend_flops(Hq, D, prefix.tolist(), ext.tolist())
    peak, src = tensor_peak()
    ach = fl / ms / 1e9
    par = SC.extend_rows(q, k_slab, v_slab, r2t, rpi, seq_d, pre_d, out, SC.pick_requests(seq, 4, seed=2))
    par.update(tol=TOL, ok=par["max_abs_err"] <= TOL)
    return {"bound": "tensor", "kernel": "extend_attention_kernel (tcgen05/TMEM)", "achieved": round(ach, 1),
            "peak": peak, "unit": "TFLOP/s", "frac": round(ach / peak, 4), "peak_source": src,
            "flops_per_launch": fl, "avg_launch_ms": round(ms, 5), "traffic": traffic_from_profiles("extend_c3"),
            "workload": "qwen2.5-7b extend (config C3): B=16, prefix U[0,8192], 512 new tokens each, bf16",
            "parity_check": par}


def measure_decode_workload(workload, rank, world, device, steps, warmup, oproj="fused", sample_clocks=False,
                            with_e2e=True, roof_reps=None, force_oproj=False):
    """One decode workload under the sharding `tp_runtime.plan_parallelism` picks for `world`.  Returns a dict of
    measurements (times are the max over ranks)."""
    from baby_sglang_b200 import ops
    from baby_sglang_b200.host_staging import HostStagedAttention
    from oracle import sampled_check as SC
    attn_backend, forward_batch, _, tp_runtime = _dropin()
    model_name = WORKLOADS[workload][0]
    Hq_full, Hkv_full = GEOMETRY[model_name][0], GEOMETRY[model_name][1]
    plan = tp_runtime.plan_parallelism(Hq_full, Hkv_full, world)
    tp = plan.mode == "kv_head_tp"
    if tp:      # rank r owns Hkv/t KV heads of EVERY request; identical page table on all ranks
        w = build_workload(workload, device, seed=1234, Hkv_override=plan.kv_heads_per_rank,
                           Hq_override=plan.q_heads_per_rank)
    else:       # single GPU, or request-batch DP: independent requests per rank
        w = build_workload(workload, device, seed=1234 + rank)
    B, Hq, Hkv, D, L = w["B"], w["Hq"], w["Hkv"], w["D"], w["L"]
    dt = torch.bfloat16
    # the o_proj shard belongs to the step wherever a collective follows it; `force_oproj` adds the full o_proj (K8's
    # world-1 form) to a single-GPU step so that it does the same work as the TP step it is compared with
    with_oproj = tp or workload == "c5" or force_oproj
    tp_group = torch.distributed.group.WORLD if world > 1 else None
    w_o = torch.randn(w["hidden"], Hq * D, device=device, dtype=dt) * 0.02 if with_oproj else None
    fused_comm = None
    if (tp or (force_oproj and world == 1)) and oproj == "fused":   # K8: one kernel per rank instead of GEMM + NCCL call
        fused_comm = tp_runtime.FusedOProjAllReduce(rank, world, tp_group, hidden=w["hidden"], max_tokens=B)

    def nccl_pair(out):
        partial = torch.nn.functional.linear(out, w_o)
        return tp_runtime.tp_all_reduce_sum(partial, tp_group) if tp else partial

    def after_attention(out):
        if not with_oproj:
            return out
        return fused_comm(out, w_o) if fused_comm is not None else nccl_pair(out)

    qkv_dev = [torch.randn(B, (Hq + 2 * Hkv) * D, device=device, dtype=dt) for _ in range(L)]
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

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(fn, n, nw):
        for _ in range(nw):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = ops.LAUNCH_COUNT
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=device)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t.item())
        return ms / n, ops.LAUNCH_COUNT - n0

    res = dict(workload=workload, model=w["model"], B=B, Hq=Hq, Hkv=Hkv, D=D, L=L, tp=tp, mode=plan.mode,
               with_oproj=with_oproj, hidden=w["hidden"], fused=fused_comm is not None)
    res["step_bytes"] = L * (decode_bytes(w["sum_seq"], B, Hq, Hkv, D) + kv_store_bytes(B, Hkv, D))
    sampler = ClockSampler(device.index or 0) if (sample_clocks and rank == 0) else None
    if sampler:
        sampler.start()
    res["ms_step"], res["launches"] = timed(step_resident, steps, warmup)
    res["clocks"] = sampler.stop() if sampler else None

    # ---- e2e: the same step through the plugin call from pinned HOST buffers
    if with_e2e:
        qkv_host = [t.cpu().pin_memory() for t in qkv_dev]
        # under TP the all-reduced result is replicated on every rank: each rank downloads its 1/N of the rows
        rows = (rank * B // world, (rank + 1) * B // world) if tp else None
        n_rows_host = rows[1] - rows[0] if tp else B
        out_host = [torch.empty(n_rows_host, w["hidden"] if with_oproj else Hq * D, dtype=dt).pin_memory()
                    for _ in range(L)]
        ids_host = torch.zeros(B, dtype=torch.int64).pin_memory()
        rpi_host, seq_host = w["rpi"].cpu().pin_memory(), w["seq"].cpu().pin_memory()
        loc_host = w["out_cache_loc"].cpu().pin_memory()
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
            staged.forward_layers(qkv_host, out_host, w["layers"], fb, after=after_attention if with_oproj else None,
                                  out_rows=rows)
            return out_host[L - 1]

        res["ms_e2e"], _ = timed(step_e2e, steps, warmup)
        res["h2d"] = (sum(t.numel() * t.element_size() for t in qkv_host) + 8 * B * 3 + 4 * B) * world   # whole job
        res["d2h"] = sum(t.numel() * t.element_size() for t in out_host) * world   # whole job
        del qkv_host, out_host, staged

    # ---- roofline of the dominant kernel: paged_attention_kernel alone, CUDA events per launch
    md = ops.build_decode_metadata(w["r2t_pool"].req_to_token, w["rpi"], w["seq"], Hkv)
    ws = ops.partial_buffers(device, md.plan.n_ctas, Hq, Hkv, D)
    r2t = w["r2t_pool"].req_to_token
    scale = D ** -0.5
    outbuf = torch.empty(B, Hq * D, device=device, dtype=dt)
    reps = roof_reps or max(3, steps // 2)
    for l in range(L):
        ops.paged_attention(qkv_split_dev[l][0], *w["kv_pool"].get_kv_buffer(l), r2t, md.plan, scale, out=outbuf,
                            workspace=ws, merge=False)
    torch.cuda.synchronize()
    evs = []
    for _ in range(reps * (28 // L if L < 28 else 1)):
        for l in range(L):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            ops.paged_attention(qkv_split_dev[l][0], *w["kv_pool"].get_kv_buffer(l), r2t, md.plan, scale, out=outbuf,
                                workspace=ws, merge=False)
            b.record()
            evs.append((a, b))
    torch.cuda.synchronize()
    res["k_ms"] = statistics.mean(a.elapsed_time(b) for a, b in evs)
    res["k_bytes"] = decode_bytes(w["sum_seq"], B, Hq, Hkv, D)

    # ---- parity at the timed size: layer 0 of the step (store, attend, merge) vs the oracle on sampled requests
    fb = forward_batch.SimplifiedForwardBatch.create_decode_batch(
        input_ids=ids_dev, req_pool_indices=w["rpi"], seq_lens=w["seq"], out_cache_loc=w["out_cache_loc"],
        req_to_token_pool=w["r2t_pool"], token_to_kv_pool=w["kv_pool"])
    q0, k0, v0 = qkv_split_dev[0]
    attn0 = attn_backend.SimpleAttentionBackend.forward(q0, k0, v0, w["layers"][0], fb, True)
    torch.cuda.synchronize()
    ks, vs = w["kv_pool"].get_kv_buffer(0)
    par = SC.decode_rows(q0, ks, vs, r2t, w["rpi"], w["seq"], attn0, SC.pick_requests(w["seq"], 16, seed=3))
    loc = w["out_cache_loc"].long()
    par["kv_store_bit_exact"] = bool(torch.equal(ks[loc], k0) and torch.equal(vs[loc], v0))
    if tp:   # the collective: fused K8 (or the NCCL pair) against fp32 sums of every rank's partial
        got = after_attention(attn0).float()
        ref32 = torch.nn.functional.linear(attn0.float(), w_o.float())
        torch.distributed.all_reduce(ref32)
        par["oproj_allreduce_max_abs_err"] = float((got - ref32).abs().max())
        par["oproj_allreduce_ref_absmax"] = float(ref32.abs().max())
        same = got.clone()
        torch.distributed.broadcast(same, src=0)
        par["oproj_allreduce_identical_on_all_ranks"] = bool(torch.equal(same, got))
        t = torch.tensor([par["max_abs_err"], par["oproj_allreduce_max_abs_err"],
                          0.0 if par["kv_store_bit_exact"] else 1.0,
                          0.0 if par["oproj_allreduce_identical_on_all_ranks"] else 1.0], device=device)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        par["max_abs_err"], par["oproj_allreduce_max_abs_err"] = float(t[0]), float(t[1])
        par["kv_store_bit_exact"] = t[2].item() == 0.0
        par["oproj_allreduce_identical_on_all_ranks"] = t[3].item() == 0.0
        par["oproj_tol"] = "one bf16 ulp of the fp32 sum"
    par.update(tol=TOL, ok=par["max_abs_err"] <= TOL and par["kv_store_bit_exact"])
    res["parity_check"] = par

    # ---- the collective alone (NCCL) and K8 alone, for the sharded report
    res["allreduce"] = None
    if tp:
        buf = torch.randn(B, w["hidden"], device=device, dtype=dt)

        def ar():
            tp_runtime.tp_all_reduce_sum(buf, tp_group)
        ar_ms, _ = timed(ar, 20, 5)
        nbytes = buf.numel() * buf.element_size()
        res["allreduce"] = {"bytes": nbytes, "nccl_avg_us": round(ar_ms * 1e3, 2),
                            "nccl_busbw_GBps": round(2 * (world - 1) / world * nbytes / ar_ms / 1e6, 1),
                            "backend": "NCCL over NVLink/NVSwitch (torch.distributed all_reduce SUM)"}
        x = torch.randn(B, Hq * D, device=device, dtype=dt)
        pair_ms, _ = timed(lambda: nccl_pair(x), 20, 5)
        res["allreduce"]["linear_plus_nccl_us"] = round(pair_ms * 1e3, 2)
        if fused_comm is not None:
            k8_ms, _ = timed(lambda: fused_comm(x, w_o), 20, 5)
            res["allreduce"]["k8_fused_us"] = round(k8_ms * 1e3, 2)
            res["allreduce"]["k8_busbw_GBps"] = round(2 * (world - 1) / world * nbytes / k8_ms / 1e6, 1)
            res["allreduce"]["k8_exchange"] = fused_comm.last_exchange
    if fused_comm is not None:
        fused_comm.close()      # collective (barrier): every rank, before the others leave
    del w, qkv_dev, qkv_split_dev, fb, attn0
    _free()
    return res


# ------------------------------------------------------------------------------------------------
# the other half of the metric: decode tok/s through the unchanged scheduler + model on the drop-in path
# ------------------------------------------------------------------------------------------------
def measure_decode_tok_s(device, steps=64, batch=256, ctx=4096):
    from baby_sglang_b200 import DROPIN_DIR, model_hooks
    from baseline import ref_loader, ref_stack
    if ref_loader.find_reference() is None:
        return {"unavailable": "no reference checkout (baseline/_ref not staged: scripts/stage_reference.py)"}
    Hq, Hkv, D, L, hidden, inter, vocab = GEOMETRY["qwen2.5-7b"]
    cfg = dict(hidden_size=hidden, num_attention_heads=Hq, num_key_value_heads=Hkv, intermediate_size=inter,
               num_hidden_layers=L, vocab_size=vocab)
    pool = dict(max_total_tokens=batch * (ctx + steps), max_requests=batch)     # SURVEY §8d NS row: 256 * 4160
    out = {"workload": f"qwen2.5-7b geometry, random init, bf16, B={batch}, ctx {ctx}, {steps} greedy decode steps "
                       f"through the reference's unchanged Scheduler loop body (scheduler.py:137-163) and "
                       f"BabyQwen2ForCausalLM on the drop-in modules; MAX_REQUESTS={batch}, "
                       f"MAX_TOTAL_TOKENS={pool['max_total_tokens']}",
           "state": "decode phase seeded directly (baseline/ref_stack.seed_decode_state): the reference's own "
                    "prefill would need 319 GB of logits at this size",
           "unit": "tok/s", "variants": {}}
    # The PURE reference stack first (its own torch attention path: 28 x 256 SDPA calls and ~14 k host syncs per
    # step), same GPU, same seeded state, N_FORCED free-running steps with every sampled token's
    # top-1 / top-2 logits recorded: context for the tok/s numbers, and the trajectory + margins the drop-in
    # stack's ids are classified against below.
    N_FORCED = 4
    pure = None
    try:
        ref = ref_stack.build(None, cfg, str(device), torch.bfloat16, **pool, listen=True, init="device")
        ref_stack.seed_decode_state(ref, batch, ctx, 8, seed=0)
        r = ref_stack.timed_decode_steps(ref, N_FORCED - 1, warmup=1)
        med = statistics.median(r["host_ms"])
        vals = ref.log.values()             # one [top1, top2] per sample_next_ids call: step-major, request-minor
        assert len(vals) == N_FORCED * batch, (len(vals), N_FORCED, batch)
        pure = {"ids": r["ids"], "margins": [[vals[s * batch + i] for s in range(N_FORCED)] for i in range(batch)]}
        out["pure_reference_on_this_gpu"] = {
            "ms_per_step_median": round(med, 1), "decode_tok_per_s": round(batch / (med / 1e3), 1),
            "steps": N_FORCED - 1,
            "note": "the reference's unmodified attn_backend/memory_pool/forward_batch (torch gathers + SDPA per "
                    "request) under the same scheduler and model; context, not the baseline arm"}
        del ref
    except Exception as e:
        out["pure_reference_on_this_gpu"] = {"error": repr(e)[:300]}
    _free()
    run = ref_stack.build(DROPIN_DIR, cfg, str(device), torch.bfloat16, **pool, listen=False, init="device")
    model = run.model
    if pure is not None:
        # teacher-forced over N_FORCED steps of all `batch` requests: the drop-in stack is fed the reference's tokens
        # and its OWN argmax at every step is compared with them; a difference is a tie if the reference's own
        # top-1 / top-2 margin there is within 2 bf16 ulps, else a mismatch (which fails the run)
        out["id_parity_vs_pure_reference"] = {}
        for name, hooks in (("dropin", ()), ("dropin+K5K6K7", ("rope", "sampler", "prep"))):
            sched = run.stack.scheduler.Scheduler(model, tokenizer=None)
            run.scheduler = sched
            model_hooks.uninstall_fused_rope_kv_store(model)
            if "rope" in hooks:
                model_hooks.install_fused_rope_kv_store(model)
            if "sampler" in hooks:
                model_hooks.install_batched_greedy_sampler(sched)
            if "prep" in hooks:
                model_hooks.install_device_batch_prep(sched)
            ref_stack.seed_decode_state(run, batch, ctx, 8, seed=0)
            reqs = list(sched.decoding_batch)
            ref_ids = [[pure["ids"][s][i] for s in range(N_FORCED)] for i in range(batch)]
            forcer = ref_stack.Forcer(sched)
            # Forcer indexes `want` by len(output_ids): the seeded first token sits at position 0
            forcer.start_run([q.input_ids for q in reqs], [[int(q.output_ids[0])] + ids for q, ids in zip(reqs, ref_ids)])
            ref_stack.timed_decode_steps(run, N_FORCED - 1, warmup=1)
            res = ref_stack.classify_forced(ref_ids, [list(x) for x in forcer.own], pure["margins"])
            req_bad = {m["request"] for m in res["mismatches"]}
            req_tied = {t["request"] for t in res["ties"]} - req_bad
            out["id_parity_vs_pure_reference"][name] = {
                "requests": batch, "steps_per_request": N_FORCED, "steps_total": res["steps_total"],
                "steps_equal": res["steps_equal"], "equal": batch - len(req_bad) - len(req_tied),
                "tied": len(req_tied),
                "ties": [{"request": t["request"], "step": t["step"], "ref_margin": round(t["ref_margin"], 5),
                          "tie_tolerance": round(t["tie_tolerance"], 5)} for t in res["ties"]][:32],
                "mismatches": len(res["mismatches"]), "mismatch_detail": res["mismatches"][:4],
                "rule": "teacher-forced; tie = pure reference's own top-1/top-2 margin <= 2 bf16 ulps at that step"}
            model_hooks.uninstall_fused_rope_kv_store(model)
            del sched, forcer
            run.scheduler = None
            _free()
    ids_ref = None
    for name, hooks in (("dropin", ()), ("dropin+K5K6K7", ("rope", "sampler", "prep")),
                        ("dropin+K5K6K7+cuda_graph", ("rope", "sampler", "prep", "graph"))):
        sched = run.stack.scheduler.Scheduler(model, tokenizer=None)          # fresh pools for every variant
        run.scheduler = sched
        model_hooks.uninstall_cuda_graph_decode(model)
        model_hooks.uninstall_fused_rope_kv_store(model)
        if "rope" in hooks:
            model_hooks.install_fused_rope_kv_store(model)
        if "sampler" in hooks:
            model_hooks.install_batched_greedy_sampler(sched)
        if "prep" in hooks:
            model_hooks.install_device_batch_prep(sched)
        if "graph" in hooks:
            model_hooks.install_cuda_graph_decode(model)
        ref_stack.seed_decode_state(run, batch, ctx, steps + 4, seed=0)
        r = ref_stack.timed_decode_steps(run, steps - 2, warmup=2)
        med = statistics.median(r["host_ms"])
        v = {"ms_per_step_median": round(med, 3), "ms_per_step_mean": round(statistics.mean(r["host_ms"]), 3),
             "device_ms_per_step_median": round(statistics.median(r["dev_ms"]), 3),
             "decode_tok_per_s": round(batch / (med / 1e3), 1)}
        if ids_ref is None:
            ids_ref = r["ids"]
        else:   # same seeded state, same weights: the hooks must not change a sampled token
            v["first_step_ids_identical_to_dropin"] = r["ids"][0] == ids_ref[0]
            v["steps_with_identical_ids"] = sum(a == b for a, b in zip(r["ids"], ids_ref))
        out["variants"][name] = v
        model_hooks.uninstall_cuda_graph_decode(model)
        model_hooks.uninstall_fused_rope_kv_store(model)
        del sched
        run.scheduler = None
        _free()
    best = max(out["variants"].values(), key=lambda x: x["decode_tok_per_s"])
    out["value"] = best["decode_tok_per_s"]
    out["best_variant"] = [k for k, x in out["variants"].items() if x is best][0]
    del run, model
    _free()
    _free()
    return out


# ------------------------------------------------------------------------------------------------
# BASELINE.json configs[0] (C1): the one configuration the reference runs itself -- Qwen2.5-1.5B random
# init, 8 x 512-token prompts, 64 greedy tokens through Scheduler.run_batch on CPU in fp32 -- timed beside the same
# run_batch on the drop-in modules on the GPU, token ids compared
# ------------------------------------------------------------------------------------------------
def _timed_run_batch(run, prompts, sp, device_events, time_attention=True):
    """One `Scheduler.run_batch` (scheduler.py:99-163) with a timestamp at every `_select_batch_to_run` (= step
    boundary) and the time spent inside `SimpleAttentionBackend.forward` (attn_backend.py:8-32: KV store + attention,
    every layer) accumulated per step: host clock on CPU, CUDA events on the GPU."""
    sched, cls = run.scheduler, run.stack.qwen2.SimpleAttentionBackend
    inner_fwd, inner_sel = cls.forward, sched._select_batch_to_run
    marks, attn, modes, pending = [], [0.0], [], []

    def fwd(*a, **kw):
        if device_events:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = inner_fwd(*a, **kw)
            e1.record()
            pending.append((len(attn) - 1, e0, e1))
            return out
        t0 = time.perf_counter()
        out = inner_fwd(*a, **kw)
        attn[-1] += time.perf_counter() - t0
        return out

    def sel():
        marks.append(time.perf_counter())
        attn.append(0.0)
        r = inner_sel()
        modes.append(r[1])
        return r

    if time_attention:      # not under the CUDA-graph hook: timing events cannot be recorded inside a capture
        cls.forward = staticmethod(fwd)
    sched._select_batch_to_run = sel
    try:
        ids = sched.run_batch([p.clone() for p in prompts], sp)
        if device_events:
            torch.cuda.synchronize()
        marks.append(time.perf_counter())
    finally:
        cls.forward = staticmethod(inner_fwd)
        sched._select_batch_to_run = inner_sel
    for i, e0, e1 in pending:
        attn[i] += e0.elapsed_time(e1) / 1e3
    steps = [b - a for a, b in zip(marks, marks[1:])]
    attn = attn[1:]
    dec = [i for i, m in enumerate(modes) if m == "decode"]
    pre = [i for i, m in enumerate(modes) if m == "prefill"]
    n_tok = sum(len(x) for x in ids)
    return [[int(t) for t in r] for r in ids], {
        "total_s": round(marks[-1] - marks[0], 3), "tok_per_s": round(n_tok / (marks[-1] - marks[0]), 1),
        "prefill_s": round(sum(steps[i] for i in pre), 3), "prefill_attention_s": round(sum(attn[i] for i in pre), 4),
        "decode_steps": len(dec),
        "decode_ms_per_step_median": round(statistics.median(steps[i] for i in dec) * 1e3, 3),
        "decode_attention_ms_per_step_median":
            round(statistics.median(attn[i] for i in dec) * 1e3, 3) if time_attention else None,
        "decode_attention_share":
            round(sum(attn[i] for i in dec) / max(sum(steps[i] for i in dec), 1e-9), 3) if time_attention else None}


def measure_c1(device, batch=8, ctx=512, new=64):
    from baby_sglang_b200 import DROPIN_DIR
    from baseline import ref_loader, ref_stack
    if ref_loader.find_reference() is None:
        return {"unavailable": "no reference checkout (baseline/_ref not staged: scripts/stage_reference.py)"}
    Hq, Hkv, D, L, hidden, inter, vocab = GEOMETRY["qwen2.5-1.5b"]
    cfg = dict(hidden_size=hidden, num_attention_heads=Hq, num_key_value_heads=Hkv, intermediate_size=inter,
               num_hidden_layers=L, vocab_size=vocab, tie_word_embeddings=True)
    pool = dict(max_total_tokens=2 * batch * (ctx + new), max_requests=batch)
    g = torch.Generator().manual_seed(2025)
    prompts = [torch.randint(0, vocab, (ctx,), generator=g) for _ in range(batch)]
    out = {"workload": f"BASELINE.json configs[0]: qwen2.5-1.5b geometry ({Hq}q/{Hkv}kv, D={D}, {L} layers, vocab "
                       f"{vocab}), random init, {batch} x {ctx}-token prompts, {new} greedy tokens each through the "
                       f"reference's unchanged Scheduler.run_batch (scheduler.py:99-163)"}
    # ---- the reference itself: CPU, fp32 (model_loader.py:56-64 picks fp32 on CPU), all host threads
    ref = ref_stack.build(None, cfg, "cpu", torch.float32, **pool, listen=True, init="device")
    sp = ref.stack.sample.SamplingParams(max_new_tokens=new, do_sample=False, eos_id=-1)
    ref.tracker.start_run()
    cpu_ids, cpu_t = _timed_run_batch(ref, prompts, sp, device_events=False)
    cpu_margins = ref.tracker.margins(prompts)
    out["reference_cpu_fp32"] = {**cpu_t, "cores": torch.get_num_threads(),
                                 "impl": "unmodified reference stack (staged checkout), torch CPU kernels"}
    weights = {k: v.to(torch.bfloat16) if v.is_floating_point() else v for k, v in ref.model.state_dict().items()}
    del ref
    gc.collect()

    def gpu_stack(shadow, hooks=(), listen=False):
        r = ref_stack.build(shadow, cfg, str(device), torch.bfloat16, **pool, listen=listen, init="device", hooks=hooks)
        r.model.load_state_dict(weights, strict=False)       # the SAME random weights as the CPU run, rounded to bf16
        return r

    # ---- the pure reference on this GPU in bf16 (its own torch attention path): trajectory + margins to classify against
    gref = gpu_stack(None, listen=True)
    gref.tracker.start_run()
    gref_ids, gref_t = _timed_run_batch(gref, prompts, sp, device_events=True)
    gref_margins = gref.tracker.margins(prompts)
    out["pure_reference_gpu_bf16"] = gref_t
    del gref
    _free()
    # ---- the drop-in modules under the same scheduler and model
    for name, hooks in (("dropin", ()), ("dropin+K5K6K7+cuda_graph", ("rope", "sampler", "prep", "graph"))):
        got = gpu_stack(DROPIN_DIR, hooks=hooks)
        ta = not hooks
        _timed_run_batch(got, prompts, sp, True, time_attention=ta)           # warm-up (module loads, graph capture)
        ids, t = _timed_run_batch(got, prompts, sp, True, time_attention=ta)  # radix-hit rerun of the same prompts
        free = ref_stack.classify_divergence(gref_ids, ids, gref_margins)
        forcer = ref_stack.Forcer(got.scheduler)
        forcer.start_run(prompts, gref_ids)
        got.scheduler.run_batch([p.clone() for p in prompts], sp)
        forced = ref_stack.classify_forced(gref_ids, [list(x) for x in forcer.own], gref_margins)
        vs_cpu = ref_stack.classify_forced(cpu_ids, [list(x) for x in forcer.own], cpu_margins) \
            if cpu_ids == gref_ids else None
        t["ids_vs_pure_reference_gpu_bf16"] = {
            "requests": batch, "steps_total": forced["steps_total"],
            "free_running_identical_requests": sum(a == b for a, b in zip(gref_ids, ids)),
            "free_running_tie_departures": len(free["ties"]), "free_running_mismatches": len(free["mismatches"]),
            "teacher_forced_steps_equal": forced["steps_equal"], "teacher_forced_ties": len(forced["ties"]),
            "teacher_forced_mismatches": len(forced["mismatches"]), "mismatch_detail": forced["mismatches"][:3],
            "rule": "tie = the pure reference's own top-1/top-2 margin <= 2 bf16 ulps at that step"}
        t["note"] = "timed on the second run_batch call (radix hit on the whole prompt, like any repeated prompt)"
        if vs_cpu is not None:
            t["ids_vs_reference_cpu_fp32_teacher_forced"] = {"steps_equal": vs_cpu["steps_equal"],
                                                             "steps_total": vs_cpu["steps_total"]}
        out[name] = t
        del got, forcer
        _free()
    same = sum(a == b for a, b in zip(cpu_ids, gref_ids))
    out["reference_cpu_fp32_vs_reference_gpu_bf16"] = {
        "identical_requests_free_running": same, "requests": batch,
        "note": "the reference against itself across dtype/device: fp32 CPU vs bf16 GPU trajectories part wherever "
                "28 layers of bf16 rounding outweigh a margin; the drop-ins are therefore classified against the "
                "reference run in their own dtype on the same GPU"}
    return out


# ------------------------------------------------------------------------------------------------
# K1 / K5 / K6: HBM-bound byte kernels either side of the path
# ------------------------------------------------------------------------------------------------
def measure_small_kernels(device):
    from baby_sglang_b200 import ops
    peak, _ = peaks()
    Hq, Hkv, D = GEOMETRY["qwen2.5-7b"][:3]
    vocab = GEOMETRY["qwen2.5-7b"][6]
    dt = torch.bfloat16
    out = {}

    def timeit(fns, trains=7, per_train=24):
        """Median over `trains` replays of ONE CUDA graph holding `per_train` back-to-back launches, CUDA events
        around the replay, per launch (a graph so that the host's ~10 us per Python call is not what is timed).
        The launches rotate over `fns`: distinct input sets whose total footprint exceeds 2 x L2, so every launch
        reads its inputs from HBM and no foreign dirty lines are written back on its clock (zeroing a flush buffer
        between launches, as round 1 did, leaves 126 MB of dirty L2 lines for the timed kernel to evict)."""
        for fn in fns:
            fn()
        torch.cuda.synchronize()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(graph, stream=side):
                keep = [fns[i % len(fns)]() for i in range(per_train)]
        torch.cuda.synchronize()
        graph.replay()
        torch.cuda.synchronize()
        ts = []
        for _ in range(trains):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            graph.replay()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) / per_train)
        del graph, keep
        return statistics.median(ts)

    NSET = 4
    for T in (256, 16384):
        S = 1 << 20
        ks = torch.empty(S + 1, Hkv, D, device=device, dtype=dt)
        vs = torch.empty(S + 1, Hkv, D, device=device, dtype=dt)
        inv = 1.0 / (1e6 ** (torch.arange(0, D, 2, device=device).float() / D))
        k1, k5 = [], []
        for i in range(NSET):
            loc = (torch.randperm(S, device=device)[:T] + 1).to(torch.int32)
            qkv = torch.randn(T, (Hq + 2 * Hkv) * D, device=device, dtype=dt)
            q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
            k, v = k.contiguous().view(T, Hkv, D), v.view(T, Hkv, D)
            pos = torch.randint(0, 4096, (T,), device=device)
            k1.append(lambda loc=loc, k=k, v=v: ops.kv_store(ks, vs, loc, k, v))
            k5.append(lambda qkv=qkv, pos=pos, loc=loc: ops.rope_kv_store(qkv, pos, inv, Hq, Hkv, D, ks, vs, loc))
        ms = timeit(k1)
        nb = kv_store_bytes(T, Hkv, D)
        out[f"K1_kv_store_T{T}"] = {"bytes": nb, "us": round(ms * 1e3, 2), "GBps": round(nb / ms / 1e6, 1),
                                    "frac_of_hbm_peak": round(nb / ms / 1e6 / peak, 4)}
        ms = timeit(k5)
        nb = 2 * T * (Hq + 2 * Hkv) * D * 2 + 12 * T
        out[f"K5_rope_kv_store_T{T}"] = {"bytes": nb, "us": round(ms * 1e3, 2), "GBps": round(nb / ms / 1e6, 1),
                                         "frac_of_hbm_peak": round(nb / ms / 1e6 / peak, 4)}
        del ks, vs, k1, k5, qkv, q, k, v
        _free()
    logits = [torch.randn(256, vocab, device=device, dtype=dt) for _ in range(NSET)]
    ms = timeit([lambda x=x: ops.greedy_sample(x) for x in logits])
    nb = logits[0].numel() * 2 + 8 * 256
    out["K6_greedy_sample_B256"] = {"bytes": nb, "us": round(ms * 1e3, 2), "GBps": round(nb / ms / 1e6, 1),
                                    "frac_of_hbm_peak": round(nb / ms / 1e6 / peak, 4)}
    out["note"] = ("CUDA events around replays of one CUDA graph of 24 back-to-back launches rotating over 4 distinct "
                   "input sets (total footprint > 2 x L2, no flush writes, no host time), median of 7 replays, per launch; bytes = SURVEY §8d formulas "
                   "(K1: 2*(2*T*Hkv*D*e)+4T; K5: read + write of the fused row; K6: B*V*e).  At T=256 these are "
                   "launch-latency bound (a few us of work).")
    del logits
    _free()
    return out


def measure_small_batch(device, ctx=4096, iters=30):
    """Small decode batches are launch-bound when launched eagerly (~37 us of host time per layer); the step makes
    no host sync and takes every size from device buffers, so it replays as ONE CUDA graph.  Attention path of the
    metric's geometry (28 layers: K4 + per layer K1, K2, K2c), B = 1, 4, 16."""
    from baby_sglang_b200 import ops
    Hq, Hkv, D, L = GEOMETRY["qwen2.5-7b"][:4]
    out = {"workload": f"qwen2.5-7b attention path, ctx {ctx}, {L} layers per step, bf16", "cases": []}
    g = torch.Generator(device=device).manual_seed(0)
    for B in (1, 4, 16):
        seq = torch.full((B,), ctx, device=device, dtype=torch.int64)
        S = B * ctx + 64
        perm = torch.randperm(S, generator=g, device=device).to(torch.int32) + 1
        r2t = torch.zeros(B, ctx + 8, dtype=torch.int32, device=device)
        r2t[:, :ctx] = perm[:B * ctx].view(B, ctx)
        rpi = torch.arange(B, device=device)
        loc = r2t[rpi, seq - 1].contiguous()
        slabs = [(torch.randn(S + 1, Hkv, D, device=device, dtype=torch.bfloat16),
                  torch.randn(S + 1, Hkv, D, device=device, dtype=torch.bfloat16)) for _ in range(L)]
        q = torch.randn(B, Hq, D, device=device, dtype=torch.bfloat16)
        k = torch.randn(B, Hkv, D, device=device, dtype=torch.bfloat16)
        v = torch.randn(B, Hkv, D, device=device, dtype=torch.bfloat16)
        outs = [torch.empty(B, Hq * D, device=device, dtype=torch.bfloat16) for _ in range(L)]

        def step():
            md = ops.build_decode_metadata(r2t, rpi, seq, Hkv)
            for l in range(L):
                ops.kv_store(slabs[l][0], slabs[l][1], loc, k, v)
                ops.paged_attention(q, slabs[l][0], slabs[l][1], r2t, md.plan, D ** -0.5, out=outs[l])

        def timed(fn):
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / iters

        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            step()
        side.synchronize()
        eager = timed(step)
        ref = [o.clone() for o in outs]
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            step()
        for o in outs:
            o.zero_()
        replay = timed(graph.replay)
        same = all(torch.equal(a, b) for a, b in zip(ref, outs))
        nb = L * (decode_bytes(B * ctx, B, Hq, Hkv, D) + kv_store_bytes(B, Hkv, D))
        out["cases"].append({"batch": B, "eager_ms_per_step": round(eager, 4), "graph_ms_per_step": round(replay, 4),
                             "graph_us_per_layer": round(replay / L * 1e3, 1), "graph_GBps": round(nb / replay / 1e6, 1),
                             "replay_bit_identical_to_eager": same})
        del graph, slabs, outs
        _free()
    return out


# ------------------------------------------------------------------------------------------------
# the GPU arm
# ------------------------------------------------------------------------------------------------
def run_b200(args, rank, world, device):
    extend = None
    if world == 1 and rank == 0 and not args.skip_extend:
        # K3 is timed alone and first, like the burst peak it is compared with (after the 28-layer decode loop the
        # GPU sits at its power cap and the same kernel runs ~10 % slower).
        extend = measure_extend_c3(device)
        _free()
    main = measure_decode_workload(args.workload, rank, world, device, args.steps, args.warmup, oproj=args.oproj,
                                   sample_clocks=True)
    sharded = None
    if args.workload == "ns" and not args.skip_sharded:
        # config C5 in the same line, so the driver's SCALE run records the sharded-with-collective configuration
        s = measure_decode_workload("c5", rank, world, device, max(5, min(args.steps, 20)), 3, oproj=args.oproj,
                                    with_e2e=False, roof_reps=1)
        peak, _ = peaks()
        sharded = {
            "workload": f"{s['model']} decode (config C5): B={s['B']}, ctx U[512,8192], {s['Hq']}q/{s['Hkv']}kv per "
                        f"rank, one layer per step (K1, K2, K2c, o_proj shard"
                        + (" + all-reduce" if s["tp"] else "") + ")",
            "parallelism": f"KV-head TP x{world}" if s["tp"] else "single GPU",
            "oproj": ("K8 fused GEMM + all-reduce" if s["fused"] else "F.linear + NCCL all-reduce") if s["tp"]
            else "F.linear (full o_proj)",
            "ms_per_step": round(s["ms_step"], 4),
            "aggregate_GBps": round(world * s["step_bytes"] / s["ms_step"] / 1e6, 1),
            "decode_tok_per_s_attention_path": round(s["B"] / (s["ms_step"] / 1e3), 1),
            "k2_ms_per_rank": round(s["k_ms"], 5),
            "k2_frac_of_hbm_peak": round(s["k_bytes"] / s["k_ms"] / 1e6 / peak, 4),
            "allreduce": s["allreduce"], "parity_check": s["parity_check"]}
    # The same work on ONE GPU, measured inside this job (rank 0 alone, the others wait): full heads + the full o_proj,
    # so that the TP step is compared with a single-GPU step that also pays for the GEMM
    tp1 = None
    if world > 1 and main["tp"] and not args.skip_tp1:
        if rank == 0:
          try:       # a failure here must not leave the other ranks waiting at the barrier below
            _free()
            a = measure_decode_workload(args.workload, 0, 1, device, max(5, args.steps // 2), 3, oproj=args.oproj,
                                        with_e2e=False, roof_reps=1, force_oproj=True)
            tp1 = {"ms_per_step": round(a["ms_step"], 4), "oproj": "K8 world-1 form (full o_proj, no exchange)",
                   "speedup_of_this_run": round(a["ms_step"] / main["ms_step"], 3),
                   "efficiency_of_this_run": round(a["ms_step"] / main["ms_step"] / world, 3),
                   "note": "rank 0 alone, same job, same step as the TP ranks run (K4 + per layer K1, K2, K2c, o_proj) "
                           "on all heads of all requests"}
            if sharded is not None and sharded.get("parallelism", "").startswith("KV-head TP"):
                b = measure_decode_workload("c5", 0, 1, device, 5, 3, oproj=args.oproj, with_e2e=False, roof_reps=1,
                                            force_oproj=True)
                sharded["tp1_ms_per_step_same_job"] = round(b["ms_step"], 4)
                sharded["speedup_vs_tp1"] = round(b["ms_step"] / sharded["ms_per_step"], 3)
                sharded["efficiency"] = round(b["ms_step"] / sharded["ms_per_step"] / world, 3)
            _free()
          except Exception as e:
            tp1 = {"error": repr(e)[:400]}
            _free()
        torch.distributed.barrier()
    tp_parity = measure_tp_parity(rank, world, device) if (world > 1 and not args.skip_parity) else None
    tok = small = configs = None
    if world == 1 and rank == 0:
        if args.workload == "ns" and not args.skip_model:
            try:
                tok = measure_decode_tok_s(device)
            except Exception as e:   # the headline line must survive a failure of the model-level leg
                tok = {"error": repr(e)[:500]}
                _free()
        if not args.skip_small:
            small = measure_small_kernels(device)
            small["small_batch_decode"] = measure_small_batch(device)
        if args.workload == "ns" and not args.skip_configs:
            # BASELINE.json configs[1] (C2) and configs[3] (C4) in the same driver-run line: K2 alone (CUDA events
            # per launch) against the measured HBM peak, the whole step, and a sampled parity check vs the oracle
            peak, _ = peaks()
            configs = {}
            for wl, n_steps in (("c2", 10), ("c4", 5)):
                _free()
                try:
                    c = measure_decode_workload(wl, rank, world, device, n_steps, 3, with_e2e=False, roof_reps=3)
                    model_name, cB, clo, chi, cL = WORKLOADS[wl]
                    configs[wl] = {
                        "workload": f"{model_name} paged GQA decode ({c['Hq']}q/{c['Hkv']}kv, D={c['D']}), batch {cB}, "
                                    f"ctx {clo}..{chi}, {cL} layer(s) resident per step",
                        "k2_ms": round(c["k_ms"], 5), "k2_bytes_per_launch": c["k_bytes"],
                        "k2_GBps": round(c["k_bytes"] / c["k_ms"] / 1e6, 1),
                        "frac": round(c["k_bytes"] / c["k_ms"] / 1e6 / peak, 4),
                        "ms_per_step": round(c["ms_step"], 4),
                        "step_GBps": round(c["step_bytes"] / c["ms_step"] / 1e6, 1),
                        "traffic": traffic_from_profiles(wl), "parity_check": c["parity_check"]}
                except Exception as e:
                    configs[wl] = {"error": repr(e)[:400]}
            _free()
            try:
                configs["c1"] = measure_c1(device)
            except Exception as e:
                configs["c1"] = {"error": repr(e)[:600]}
            _free()
    if rank != 0:
        return
    cpu = cpu_baseline(args.workload, seconds=15.0)
    m = main
    tp, L, B = m["tp"], m["L"], m["B"]
    peak, peak_src = peaks()
    achieved = m["k_bytes"] / m["k_ms"] / 1e6
    total = (1 if tp else world) * m["step_bytes"] * (world if tp else 1)      # whole-job bytes either way
    value = total / m["ms_step"] / 1e6
    lo, hi = WORKLOADS[args.workload][2], WORKLOADS[args.workload][3]
    if tp:
        par = (f"KV-head TP x{world} (tp_runtime.plan_parallelism): o_proj partial + " +
               ("SUM all-reduce fused in one kernel over NVLink (K8)" if m["fused"] else "NCCL all-reduce") +
               " per layer, inside the timed step")
    elif world > 1:
        par = f"request-batch DP x{world} (tp_runtime.plan_parallelism: Hkv {m['Hkv']} < {world}; no collective)"
    else:
        par = "single GPU"
    line = {
        "metric": METRIC if args.workload == "ns" else f"paged decode attn HBM GB/s, workload {args.workload}",
        "value": round(value, 1), "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(m["ms_step"], 4), "higher_is_better": True, "scaling": "strong" if tp else "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"{m['model']} paged GQA decode ({m['Hq']}q/{m['Hkv']}kv{' per rank' if tp else ''}, "
                               f"D={m['D']}), batch {B}{'' if tp or world == 1 else '/GPU'}, ctx {lo}..{hi}, "
                               f"page_size 1, {L} layers/step, fragmented page table",
                   "parallelism": par,
                   "l2": f"inputs > L2: {m['k_bytes'] / 1e9:.2f} GB of KV per layer per rank, {L} distinct layers "
                         f"per step",
                   "step": "K4 metadata + per layer (K1 kv_store, K2 paged attention, K2c merge"
                           + (", o_proj shard GEMM" + (" + all-reduce" if tp else "") if m["with_oproj"] else "")
                           + ")"},
        "decode_tok_per_s": tok,
        "decode_tok_per_s_attention_path": round((1 if tp else world) * B / (m["ms_step"] / 1e3), 1),
        "e2e": {"value": round(total / m["ms_e2e"] / 1e6, 1), "unit": "GB/s", "ms_per_step": round(m["ms_e2e"], 4),
                "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"],
                "api": "HostStagedAttention.forward_layers (dropin attn_backend.SimpleAttentionBackend.forward per "
                       "layer) from pinned host q/k/v, copies on side streams"
                       + ("; each rank uploads its own heads' q/k/v and downloads 1/N of the rows of the replicated "
                          "all-reduced result" if tp else "")
                       + ("; byte counts are sums over the ranks" if world > 1 else "")},
        "gpu_launches": m["launches"],
        "clocks": m["clocks"],
        "roofline": {"bound": "hbm", "kernel": "paged_attention_kernel", "achieved": round(achieved, 1),
                     "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4), "peak_source": peak_src,
                     "bytes_per_launch": m["k_bytes"], "avg_launch_ms": round(m["k_ms"], 5),
                     "traffic": args.traffic if args.traffic is not None else
                     (traffic_from_profiles(args.workload) if world == 1 else None),
                     "traffic_source": ({"file": "--traffic flag"} if args.traffic is not None else
                                        (traffic_from_profiles(args.workload, with_source=True)[1] if world == 1
                                         else None))},
        "parity_check": m["parity_check"],
        "allreduce": m["allreduce"],
        "sharded": sharded,
        "tp1_same_work": tp1,
        "tp_parity": tp_parity,
        "extend_roofline": extend,
        "small_kernels": small,
        "configs": configs,
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    c1 = (configs or {}).get("c1") or {}
    c1_bad = [k for k, v in c1.items() if isinstance(v, dict) and
              (v.get("ids_vs_pure_reference_gpu_bf16") or {}).get("teacher_forced_mismatches")]
    if c1_bad:
        raise SystemExit("C1: greedy ids differ from the pure reference away from a tie: " + json.dumps(c1_bad))
    idp = (tok or {}).get("id_parity_vs_pure_reference") or {}
    if any(v["mismatches"] for v in idp.values()):
        raise SystemExit("greedy ids differ from the pure reference away from a tie: " + json.dumps(idp)[:2000])
    if tp_parity is not None and not tp_parity["ok"]:
        raise SystemExit("tp_parity failed: " + json.dumps(tp_parity["failed"])[:2000])


def measure_tp_parity(rank, world, device):
    """The multi-GPU parity scripts (tests/multi_gpu_oproj.py, tests/multi_gpu_tp.py) run inside this torchrun job,
    after every timed leg: K8 under both exchanges incl. graph replay and odd row counts, host-staged attention + K8,
    K8 in row blocks, and the reference's unchanged scheduler + model sharded by tp_runtime (NCCL / K8 / K8 + TP MLP
    / row blocks) against the unsharded drop-in stack.  Their per-case JSON lines are collected, not printed."""
    import contextlib
    import io
    from baseline import ref_loader
    from tests import multi_gpu_oproj, multi_gpu_tp
    group = torch.distributed.group.WORLD
    saved = os.environ.get("BSGL_K8_EXCHANGE")
    buf, ok, t0 = io.StringIO(), True, time.perf_counter()
    skipped = []
    try:
        with contextlib.redirect_stdout(buf):
            ok &= bool(multi_gpu_oproj.run("", rank, world, group))
            ok &= bool(multi_gpu_oproj.run("nvls", rank, world, group))
            ok &= bool(multi_gpu_oproj.host_staged(rank, world, group))
            ok &= bool(multi_gpu_oproj.chunked(rank, world, group))
            os.environ["BSGL_K8_EXCHANGE"] = ""
            if ref_loader.find_reference() is None:
                skipped.append("tp-sharded reference stack: no staged reference checkout")
            elif multi_gpu_tp.CFG["num_key_value_heads"] % world:
                skipped.append(f"tp-sharded reference stack: 8 KV heads do not shard over {world} ranks")
            else:
                ok &= bool(multi_gpu_tp.check(rank, world, group))
    finally:
        if saved is None:
            os.environ.pop("BSGL_K8_EXCHANGE", None)
        else:
            os.environ["BSGL_K8_EXCHANGE"] = saved
    flag = torch.tensor([1 if ok else 0], device=device)
    torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN)
    cases = []
    for ln in buf.getvalue().splitlines():
        try:
            cases.append(json.loads(ln))
        except ValueError:
            pass
    failed = [c for c in cases if not c.get("ok", True)]
    _free()
    return {"world": world, "cases": len(cases), "ok": bool(flag.item()) and not failed, "failed": failed[:4],
            "skipped": skipped, "seconds": round(time.perf_counter() - t0, 1),
            "what": "tests/multi_gpu_oproj.py (K8 both exchanges, 3 eager calls + 3 graph replays per shape, "
                    "host-staged + K8, row blocks) and tests/multi_gpu_tp.py (unchanged scheduler + model, TP-sharded, "
                    "4 variants x cold / radix-hit x 32 teacher-forced steps) inside this job, all ranks"}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own modules (when reachable) or the oracle port, on the box's host cores
# ------------------------------------------------------------------------------------------------
def _cpu_case(workload, Bs, seed=0):
    model, B, lo, hi, L = WORKLOADS[workload]
    Hq, Hkv, D = GEOMETRY[model][:3]
    g = torch.Generator().manual_seed(seed)
    Bs = min(Bs, B)
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
    return dict(B=B, Bs=Bs, Hq=Hq, Hkv=Hkv, D=D, L=L, lo=lo, hi=hi, seq=seq, S=S, C=C, k_slab=k_slab, v_slab=v_slab,
                r2t=r2t, rpi=rpi, loc=loc, q=q, k=k, v=v, nbytes=nbytes)


def _reference_layer_call(c):
    """The reference's own SimpleAttentionBackend.forward on its own pools (attn_backend.py:8-149), or None."""
    from baseline import ref_loader
    if ref_loader.find_reference() is None:
        return None
    ref = ref_loader.load_stack(names=("memory_pool", "forward_batch", "attn_backend"))
    kv = ref.memory_pool.MHATokenToKVPool(size=c["S"], dtype=torch.float32, head_num=c["Hkv"], head_dim=c["D"],
                                          layer_num=1, device="cpu")
    kv.k_buffer[0].copy_(c["k_slab"])
    kv.v_buffer[0].copy_(c["v_slab"])
    r2t = ref.memory_pool.ReqToTokenPool(size=c["Bs"], max_context_len=c["C"], device="cpu")
    r2t.req_to_token.copy_(c["r2t"])
    fb = ref.forward_batch.SimplifiedForwardBatch.create_decode_batch(
        input_ids=torch.zeros(c["Bs"], dtype=torch.int64), req_pool_indices=c["rpi"], seq_lens=c["seq"],
        out_cache_loc=c["loc"], req_to_token_pool=r2t, token_to_kv_pool=kv)
    layer = _Layer(0, c["Hq"], c["Hkv"], c["D"])
    fwd = ref.attn_backend.SimpleAttentionBackend.forward
    return lambda: fwd(c["q"], c["k"], c["v"], layer, fb, True)


def cpu_baseline(workload, seconds=15.0, sample_batch=64):
    """One layer call (kv_store + decode attention) of the workload on the host cores, fp32, all threads, on
    `sample_batch` of the requests, repeated for about `seconds`.  kind "reference" = the reference's own modules
    from the staged checkout; else the oracle port."""
    from oracle import paged_attention_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    c = _cpu_case(workload, sample_batch)
    call = _reference_layer_call(c)
    kind = "reference" if call is not None else "port"
    if call is None:
        def call():
            return O.backend_forward(c["q"], c["k"], c["v"], c["k_slab"], c["v_slab"], c["r2t"], c["rpi"], c["seq"],
                                     c["loc"], False)
    call()      # warm-up
    t0 = time.perf_counter()
    n = 0
    while True:
        call()
        n += 1
        dt = time.perf_counter() - t0
        if dt >= seconds or n >= 200:
            break
    ms_call = dt / n * 1e3
    # a STEP of the workload is L layers x B requests; the sample is one layer x Bs requests
    ms_step_est = ms_call * c["L"] * c["B"] / c["Bs"]
    out = {"value": round(n * c["nbytes"] / dt / 1e9, 4), "unit": "GB/s", "cores": cores, "kind": kind,
           "ms_per_layer_call_on_sample": round(ms_call, 3), "ms_per_step_extrapolated": round(ms_step_est, 1),
           "decode_tok_per_s_attention_path": round(c["B"] / (ms_step_est / 1e3), 2),
           "sample": f"{n} calls of one layer (kv_store + decode attention) on {c['Bs']} of the {c['B']} requests, "
                     f"ctx {c['lo']}..{c['hi']}, fp32 (the reference's CPU dtype, model_loader.py:56-64), "
                     + ("the reference's own attn_backend/memory_pool/forward_batch from the staged checkout"
                        if kind == "reference" else "oracle/paged_attention_oracle.py")}
    if kind == "reference":     # the port beside it, same sample, for continuity with round 1
        t0 = time.perf_counter()
        m = 0
        while m < 3 or time.perf_counter() - t0 < min(4.0, seconds / 3):
            O.backend_forward(c["q"], c["k"], c["v"], c["k_slab"], c["v_slab"], c["r2t"], c["rpi"], c["seq"],
                              c["loc"], False)
            m += 1
        out["oracle_port_GBps_same_sample"] = round(m * c["nbytes"] / (time.perf_counter() - t0) / 1e9, 4)
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (its own modules from the staged checkout,
    else the oracle port), all host threads, a bounded sample per step."""
    if rank != 0:
        return
    per = max(3.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    vals = []
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(args.workload, seconds=per)
        if i >= args.warmup:
            vals.append(r)
    v = statistics.mean(x["value"] for x in vals)
    model, B, lo, hi, L = WORKLOADS[args.workload]
    Hq, Hkv, D = GEOMETRY[model][:3]
    line = {
        "impl": "reference",
        "metric": METRIC if args.workload == "ns" else f"paged decode attn HBM GB/s, workload {args.workload}",
        "value": round(v, 4), "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(statistics.mean(x["ms_per_step_extrapolated"] for x in vals), 1),
        "ms_per_step_note": "one step = 28 layers x 256 requests, extrapolated per byte from the timed sample "
                            "(one layer call on 64 requests); the sample's own time is in cpu_baseline",
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{model} paged GQA decode ({Hq}q/{Hkv}kv, D={D}), batch {B}, ctx {lo}..{hi}, "
                               f"page_size 1, {L} layers/step; the reference's CPU path on a bounded sample"},
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
                    help="under KV-head TP: K8 (o_proj GEMM with the all-reduce fused over NVLink) or "
                         "F.linear + NCCL all-reduce")
    ap.add_argument("--traffic", type=float, default=None,
                    help="dram bytes per launch of the dominant kernel from an ncu --set full capture (profiles/)")
    for flag in ("extend", "sharded", "model", "small", "parity", "configs", "tp1"):
        ap.add_argument(f"--skip-{flag}", action="store_true", help=f"leave the `{flag}` leg out (profiling runs)")
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
