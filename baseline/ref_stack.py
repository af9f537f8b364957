"""Harness code (not product): build and run the reference's UNCHANGED scheduler / radix cache / overlap worker /
qwen2 model, either pure (every module the reference's own) or with the four drop-in modules shadowing
`attn_backend`, `forward_batch`, `memory_pool`, `tp_runtime` (the plug mechanism, qwen2.py:5-7, scheduler.py:8-9).

Used by tests/test_reference_stack.py (pure stack vs drop-in stack on the same GPU) and by bench.py (the
unchanged scheduler driving the drop-in path for the end-to-end decode tok/s number).

The recorder below only LISTENS: it wraps `sample_next_ids` where the reference's scheduler / overlap worker look
it up (scheduler.py:11, overlap_worker.py:10) and notes the top-2 logits of every sampled token, so that a
differing greedy token can be classified as a numerical tie or a real mismatch.  It never changes a result.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch

from . import ref_loader


def deterministic_weights(model: torch.nn.Module, scale: float = 1.0) -> None:
    """Same recipe as tests/harness_model.py: every parameter from a generator seeded by crc32(name)."""
    import zlib
    with torch.no_grad():
        for name, p in model.named_parameters():
            g = torch.Generator().manual_seed(zlib.crc32(name.encode()))
            r = torch.randn(p.shape, generator=g, dtype=torch.float32)
            if name.endswith("norm.weight") or "layernorm" in name:
                val = 1.0 + 0.1 * r
            elif name.endswith(".bias"):
                val = 0.1 * r
            elif "embed_tokens" in name:
                val = r
            else:
                val = r * (scale / (p.shape[-1] ** 0.5))
            p.copy_(val.to(p.dtype))


def peak_lm_head(model: torch.nn.Module, peak: float, stride: int = 389) -> torch.Tensor:
    """Widen the top-1 / top-2 margins of a random-init model without taking the context out of the decision.

    A random LM head gives logits ~ N(0, 1) over the vocabulary: the top two are a few bf16 ulps apart every few
    dozen steps, and rounding noise then decides the argmax in ANY implementation.  Here row j of `lm_head` gets, on
    top of its random part, a multiple of the embedding of token perm[j] (perm[j] = (stride * j + 1) mod V, stride
    odd): the token whose embedding dominates the residual stream -- normally the current input token t -- lifts the
    logit of its "successor" inv_perm[t] to about `peak` (in units of the random logits' standard deviation), while
    attention and MLP outputs still move every logit.  With `peak` around 5.5 most steps follow the successor chain
    with a wide margin, a few per cent are decided by the context-dependent random part, and near-ties become rare
    enough that a whole 64-step batch without one exists.  Returns inv_perm (successor of each token)."""
    emb = model.model.embed_tokens.weight
    V, H = emb.shape
    perm = (torch.arange(V) * stride + 1) % V
    with torch.no_grad():
        e = emb.float()
        # h = x / rms(x) with x = embedding + four block outputs of comparable size: h . e is about H / sqrt(5)
        model.lm_head.weight.add_((peak * (5.0 ** 0.5) / H) * e[perm].to(model.lm_head.weight.dtype))
    inv = torch.empty(V, dtype=torch.long)
    inv[perm] = torch.arange(V)
    return inv


def bf16_ulp(x: float) -> float:
    """Spacing of bf16 numbers at |x| (8 significant bits)."""
    return 2.0 ** (math.floor(math.log2(max(abs(x), 2.0 ** -126))) - 7)


@dataclass
class SampleLog:
    """top-1 / top-2 logit of every sampled token, in call order."""
    top2: List[torch.Tensor] = field(default_factory=list)      # each [2] float32 on the device

    def values(self) -> List[List[float]]:
        if not self.top2:
            return []
        return torch.stack(self.top2).cpu().tolist()


def _listen(stack, log: SampleLog) -> None:
    for mod in (stack.scheduler, stack.overlap_worker):
        if mod is None or not hasattr(mod, "sample_next_ids"):
            continue
        inner = mod.sample_next_ids

        def recorder(logits, *a, _inner=inner, **kw):
            log.top2.append(torch.topk(logits.reshape(-1).float(), 2).values)
            return _inner(logits, *a, **kw)

        mod.sample_next_ids = recorder


class StepTracker:
    """Maps the recorder's call-order log onto (request, step).  Both `_process_results` (scheduler.py:548-589) and
    `_process_results_overlap` (:700-761) commit batches in the order they were sampled, one `sample_next_ids`
    call per request of the batch; a request's output list grows by one exactly when its sample was kept (overlap
    mode drops the extra step of an already finished request, :741-747)."""

    def __init__(self, scheduler, log: SampleLog):
        self.log, self.kept = log, []            # kept: (record index, prompt key) per committed token
        self._cursor = 0
        cls = type(scheduler)

        def wrap(name, get_reqs):
            inner = getattr(cls, name)

            def method(this, *a, **kw):
                reqs = get_reqs(*a, **kw)
                before = [len(r.output_ids) for r in reqs]
                out = inner(this, *a, **kw)
                for i, (r, n0) in enumerate(zip(reqs, before)):
                    if len(r.output_ids) > n0:
                        self.kept.append((self._cursor + i, tuple(r.input_ids.tolist())))
                self._cursor += len(reqs)
                return out
            setattr(scheduler, name, method.__get__(scheduler, cls))

        wrap("_process_results", lambda next_ids, batch_requests, mode: batch_requests)
        wrap("_process_results_overlap", lambda batch, next_ids_cpu: batch.batch_requests)

    def start_run(self) -> None:
        self.kept, self._cursor, self.log.top2 = [], 0, []

    def margins(self, prompts: Sequence[torch.Tensor]) -> List[List[List[float]]]:
        """[request][step] = [top1, top2] of the logits the kept token was sampled from."""
        vals = self.log.values()
        index = {tuple(p.tolist()): i for i, p in enumerate(prompts)}
        out: List[List[List[float]]] = [[] for _ in prompts]
        for rec, key in self.kept:
            out[index[key]].append(vals[rec])
        return out


class Forcer:
    """Teacher forcing for `Scheduler.run_batch`: the stack's own `_sample_next_ids` (the reference's method, or the
    K6 hook) runs and its ids are RECORDED, but the ids handed back to the scheduler are the given ones, so the run
    follows another run's trajectory exactly.  With it the index state (page table, slot FIFO, radix tree sizes)
    of two stacks must agree bit for bit even where a bf16 tie would let their free-running tokens part, and every
    one of the 64 steps of every request gets compared, not only those before a first departure."""

    def __init__(self, scheduler):
        self.scheduler = scheduler
        self.inner = scheduler._sample_next_ids          # bound: class method or an installed hook
        self.want: Optional[List[List[int]]] = None
        self.index: Dict[tuple, int] = {}
        self.own: List[List[int]] = []
        scheduler._sample_next_ids = self._call

    def start_run(self, prompts: Sequence[torch.Tensor], want_ids: List[List[int]]) -> None:
        self.want = want_ids
        self.index = {tuple(p.tolist()): i for i, p in enumerate(prompts)}
        self.own = [[] for _ in prompts]

    def _call(self, logits, batch_requests, mode):
        own = self.inner(logits, batch_requests, mode)
        if self.want is None:
            return own
        own_list = [int(x) for x in own.tolist()]
        forced = []
        for req, o in zip(batch_requests, own_list):
            r = self.index[tuple(req.input_ids.tolist())]
            self.own[r].append(o)
            forced.append(self.want[r][len(req.output_ids)])
        return torch.tensor(forced, dtype=own.dtype, device=own.device)


def classify_forced(ref_ids: List[List[int]], own_ids: List[List[int]], ref_margins: List[List[List[float]]],
                    ulps: float = 2.0) -> dict:
    """Teacher-forced comparison: EVERY step of every request.  A differing own argmax is a TIE if the reference
    side's top-1/top-2 margin at that step is within `ulps` bf16 ulps, else a MISMATCH."""
    ties, bad, equal = [], [], 0
    for r, (a, b) in enumerate(zip(ref_ids, own_ids)):
        if len(a) != len(b):
            bad.append({"request": r, "step": -1, "why": f"lengths {len(a)} vs {len(b)}"})
            continue
        for s, (x, y) in enumerate(zip(a, b)):
            if x == y:
                equal += 1
                continue
            t1, t2 = ref_margins[r][s]
            tol = ulps * bf16_ulp(max(abs(t1), 1.0))
            rec = {"request": r, "step": s, "ref_id": x, "got_id": y, "ref_top1": t1, "ref_margin": t1 - t2,
                   "tie_tolerance": tol}
            (ties if (t1 - t2) <= tol else bad).append(rec)
    return {"steps_total": sum(len(a) for a in ref_ids), "steps_equal": equal, "ties": ties, "mismatches": bad}


@dataclass
class StackRun:
    stack: object
    scheduler: object
    model: object
    log: SampleLog
    tracker: Optional[StepTracker] = None
    comm: object = None            # FusedOProjAllReduce of a TP-sharded stack (close() it collectively)

    def state(self) -> Dict[str, object]:
        """The index state the north star requires to be bit-exact."""
        s = self.scheduler
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        return {"req_to_token": s.req_to_token_pool.req_to_token.clone().cpu(),
                "free_slots": s.token_to_kv_pool.free_slots.clone().cpu().to(torch.int32),
                "free_rows": list(s.req_to_token_pool.free_slots),
                "evictable_size": int(s.tree_cache.evictable_size_),
                "kv_available": int(s.token_to_kv_pool.available_size())}


def build(shadow_dir: Optional[str], cfg_kwargs: dict, device: str, dtype: torch.dtype, max_total_tokens: int,
          max_requests: int, hooks: Sequence[str] = (), listen: bool = True, init: str = "deterministic",
          tp: Optional[dict] = None, lm_head_peak: Optional[float] = None) -> StackRun:
    """Reference scheduler + model on `device`; `shadow_dir=None` is the pure reference, `DROPIN_DIR` the product.
    `hooks` ⊆ {"rope", "sampler", "prep", "graph"} installs the row-(f) kernels (baby_sglang_b200.model_hooks) on
    the instances.  `tp = dict(rank, world, group, fused: bool, mlp: bool)` shards the unchanged model with the
    drop-in `tp_runtime` (KV-head TP; the scheduler's pool shrinks through `kv_pool_config`) before the scheduler
    is built."""
    st = ref_loader.load_stack(shadow_dir=shadow_dir)
    os.environ["BABYSGL_MAX_TOTAL_TOKENS"] = str(max_total_tokens)     # scheduler.py:33-36
    os.environ["BABYSGL_MAX_REQUESTS"] = str(max_requests)
    cfg = st.model_loader.BabyQwenConfig(**cfg_kwargs)
    if init == "deterministic":
        model = st.qwen2.BabyQwen2ForCausalLM(cfg).float().eval()
        deterministic_weights(model)
        if lm_head_peak is not None:
            peak_lm_head(model, lm_head_peak)
        model = model.to(device=device, dtype=dtype)
    else:   # big geometries: construct on the device, default torch init, seeded
        torch.manual_seed(0)
        prev = torch.get_default_dtype()
        torch.set_default_dtype(dtype)
        try:
            with torch.device(device):
                model = st.qwen2.BabyQwen2ForCausalLM(cfg).eval()
        finally:
            torch.set_default_dtype(prev)
        for m in model.modules():       # rotary buffers are built with explicit CPU/float math in the reference
            for name, buf in list(m.named_buffers(recurse=False)):
                setattr(m, name, buf.to(device))
    comm = None
    if tp is not None:
        tpr = st.tp_runtime
        if tp.get("fused"):
            comm = tpr.FusedOProjAllReduce(tp["rank"], tp["world"], tp["group"], hidden=cfg.hidden_size,
                                           max_tokens=tp.get("max_tokens", 1024), dtype=dtype)
        plan = tpr.shard_attention_heads(model, tp["rank"], tp["world"], tp["group"], fused_allreduce=comm)
        if plan.mode != "kv_head_tp":
            raise ValueError(f"geometry does not shard over {tp['world']} ranks: {plan}")
        if tp.get("mlp"):
            tpr.shard_mlp(model, tp["rank"], tp["world"], tp["group"], fused_allreduce=comm)
        model.config = tpr.kv_pool_config(model.config, plan)      # scheduler.py:50-57 reads num_key_value_heads
    sched = st.scheduler.Scheduler(model, tokenizer=None)
    log = SampleLog()
    tracker = None
    if listen:
        _listen(st, log)
        tracker = StepTracker(sched, log)
    if hooks:
        from baby_sglang_b200 import model_hooks
        if "rope" in hooks:
            model_hooks.install_fused_rope_kv_store(model)
        if "sampler" in hooks:
            model_hooks.install_batched_greedy_sampler(sched)
        if "prep" in hooks:
            model_hooks.install_device_batch_prep(sched)
        if "graph" in hooks:
            model_hooks.install_cuda_graph_decode(model)
    run = StackRun(st, sched, model, log, tracker)
    run.comm = comm
    return run


def classify_divergence(ref_ids: List[List[int]], got_ids: List[List[int]], ref_margins: List[List[List[float]]],
                        ulps: float = 2.0) -> dict:
    """Free-running comparison of two greedy runs.  For every request: identical up to (excluding) the first
    differing step; that step is a TIE if the reference side's own top-1/top-2 margin there is within `ulps` bf16
    ulps of the top-1 logit (the argmax is then decided by rounding noise in either stack),
    otherwise a MISMATCH.  `ref_margins[r][s] = [top1, top2]`."""
    ties, bad, compared = [], [], 0
    for r, (a, b) in enumerate(zip(ref_ids, got_ids)):
        if len(a) != len(b):
            bad.append({"request": r, "step": -1, "why": f"lengths {len(a)} vs {len(b)}"})
            continue
        for s, (x, y) in enumerate(zip(a, b)):
            if x == y:
                compared += 1
                continue
            t1, t2 = ref_margins[r][s]
            tol = ulps * bf16_ulp(max(abs(t1), 1.0))
            rec = {"request": r, "step": s, "ref_id": x, "got_id": y, "ref_top1": t1, "ref_margin": t1 - t2,
                   "tie_tolerance": tol}
            (ties if (t1 - t2) <= tol else bad).append(rec)
            break
    total = sum(len(a) for a in ref_ids)
    return {"steps_total": total, "steps_compared_equal": compared, "ties": ties, "mismatches": bad}


# ------------------------------------------------------------------------------------------------------
# end-to-end decode step timing through the unchanged scheduler (bench.py: the "& decode tok/s" half of the metric)
# ------------------------------------------------------------------------------------------------------
def seed_decode_state(run: StackRun, batch: int, ctx: int, max_new_tokens: int, seed: int = 0,
                      chunk: int = 32) -> None:
    """Put `batch` requests with `ctx`-token contexts straight into the scheduler's decode phase.

    The reference cannot prefill this state itself at BASELINE size: its prefill computes LM-head logits for EVERY
    prompt token (qwen2.py:31-36) -- 256 x 4096 tokens x 152 064 vocab = 319 GB of bf16.  So the page-table rows and
    KV slots are allocated by the scheduler's own `_prepare_prefill_batch` (scheduler.py:381-452: row FIFO, slot
    FIFO, page-table rows written exactly as a real prefill would), the model forward of the prefill is skipped and
    the KV slabs are filled with N(0,1) instead; every request then holds one (arbitrary) first output token and
    sits in `decoding_batch`, as after `_process_results(..., "prefill")` (:566-579) without the radix insertion.
    From here on the scheduler's normal loop body runs unchanged."""
    sched, st = run.scheduler, run.stack
    dev = sched.device
    g = torch.Generator().manual_seed(seed)
    vocab = run.model.config.vocab_size
    sched.waiting_queue.clear()
    sched.decoding_batch.clear()
    sched.finished_requests.clear()
    sched.est_new_token_ratio = sched.r_init
    reqs = []
    for _ in range(batch):
        ids = torch.randint(0, vocab, (ctx,), generator=g).to(dev)
        reqs.append(st.request.Request(
            input_ids=ids, output_ids=[], max_new_tokens=max_new_tokens + 1, temperature=1.0, top_k=0, top_p=1.0,
            do_sample=False, repetition_penalty=1.0, eos_id=-1,
            prefix_indices=torch.tensor([], dtype=torch.int32), last_node=None, num_cached_tokens=0,
            constraint_state=None))
    for i in range(0, batch, chunk):
        sched._prepare_prefill_batch(reqs[i:i + chunk])          # rows + slots + page-table rows; batch discarded
    pool = sched.token_to_kv_pool
    gdev = torch.Generator(device=dev).manual_seed(seed)            # same KV contents for every stack / variant
    for l in range(pool.layer_num):
        pool.k_buffer[l].normal_(generator=gdev)
        pool.v_buffer[l].normal_(generator=gdev)
    first = torch.randint(0, vocab, (batch,), generator=g).tolist()
    for r, t in zip(reqs, first):
        r.output_ids.append(int(t))
        sched.decoding_batch.append(r)
    torch.cuda.synchronize()


def timed_decode_steps(run: StackRun, steps: int, warmup: int = 2) -> dict:
    """`steps` iterations of the body of `Scheduler.run_batch`'s loop (scheduler.py:137-163, steps 2-6), each
    timed on the host across a device synchronise (the step ends in `.item()` reads anyway) and on the device with
    CUDA events.  Returns per-step times and the ids sampled at every step."""
    import time
    sched, model = run.scheduler, run.model
    host_ms, dev_ms, ids = [], [], []
    for it in range(warmup + steps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        running_batch, mode = sched._select_batch_to_run()
        if mode != "decode":
            raise RuntimeError(f"expected a decode step, scheduler chose {mode!r}")
        fb = sched._prepare_forward(running_batch, mode)
        with torch.no_grad():
            logits = model(fb.input_ids, fb.positions, fb)
        next_ids = sched._sample_next_ids(logits, running_batch, mode)
        sched._process_results(next_ids, running_batch, mode)
        e1.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        if it >= warmup:
            host_ms.append((t1 - t0) * 1e3)
            dev_ms.append(e0.elapsed_time(e1))
        ids.append([int(r.output_ids[-1]) for r in running_batch])
    return {"host_ms": host_ms, "dev_ms": dev_ms, "ids": ids, "batch": len(sched.decoding_batch)}
