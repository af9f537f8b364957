"""The north-star gate itself: the reference's UNCHANGED scheduler.py / radix_cache.py / overlap_worker.py /
qwen2.py run on top of the drop-in `attn_backend`, `forward_batch`, `memory_pool` (module-name shadowing,
qwen2.py:5-7, scheduler.py:8-9), compared with the PURE reference stack on the same GPU, dtype, weights, prompts.

Needs a reference checkout: $BABYSGL_REF, /root/reference (build container) or baseline/_ref (staged by
scripts/stage_reference.py; git-ignored, travels to the GPU box with the snapshot).  Without one the tests skip.

What is asserted, per geometry (Qwen2.5-1.5B heads, Qwen2.5-7B heads) and per mode (`run_batch`,
`run_batch_overlap`), for a cold run and a radix-hit rerun of the same prompts, 64 greedy tokens each, pools sized
so retraction never fires:
  * index state bit-exact after every run: `req_to_token`, KV `free_slots` (FIFO order), free page-table rows,
    `evictable_size_`, `available_size()`                                            
  * token IDs: identical for all 64 steps, free running.  bf16 logits of a random-init model tie now and then;
    a request may leave the reference's trajectory ONLY at a step where the pure reference's own top-1/top-2
    margin is within 2 bf16 ulps (rounding noise decides the argmax in either stack).  Every such
    tie is written, with its margin, to gpurun_out/reference_stack_flips.json (committed copy:
    profiles/r2_reference_stack_flips.json); anything else fails the test.
  * the same again with K5 + K6 + K7 installed on the instances (`model_hooks.install_*`), and once more with the
    decode forward replayed as a CUDA graph (`install_cuda_graph_decode`; in overlap mode the graph is captured and
    replayed on the worker thread's side stream).
"""
import json
import os

import pytest
import torch

from baby_sglang_b200 import DROPIN_DIR, _lib
from baseline import ref_loader, ref_stack

needs_ref = pytest.mark.skipif(ref_loader.find_reference() is None, reason="no reference checkout on this machine")
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FLIP_LOG = os.path.join(REPO, "gpurun_out", "reference_stack_flips.json")

TINY = dict(hidden_size=256, num_attention_heads=2, num_key_value_heads=1, intermediate_size=512,
            num_hidden_layers=2, vocab_size=512)
GEOMETRIES = {
    # real head counts of the BASELINE configs (D=128), short MLP / vocab so the pure reference finishes in seconds
    "qwen1.5b_heads": (dict(hidden_size=12 * 128, num_attention_heads=12, num_key_value_heads=2,
                            intermediate_size=1024, num_hidden_layers=2, vocab_size=1024), [300, 17, 129, 64, 1, 255]),
    "qwen7b_heads": (dict(hidden_size=28 * 128, num_attention_heads=28, num_key_value_heads=4,
                          intermediate_size=1024, num_hidden_layers=2, vocab_size=1024), [257, 40, 700, 5]),
}
N_NEW = 64
POOL = dict(max_total_tokens=16384, max_requests=8)       # C = 2048 per request; nothing ever retracts


def _prompts(vocab, lens, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.randint(0, vocab, (n,), generator=g) for n in lens]


def _run(run: "ref_stack.StackRun", prompts, mode, n_new=N_NEW, forced=None):
    """Cold run + radix-hit rerun.  `forced` = [ids of the cold run, ids of the rerun] of another stack: this
    stack's own argmax is recorded but the other stack's tokens are fed forward (normal mode only)."""
    sp = run.stack.sample.SamplingParams(max_new_tokens=n_new, do_sample=False, eos_id=-1)
    fn = run.scheduler.run_batch if mode == "normal" else run.scheduler.run_batch_overlap
    forcer = None
    if forced is not None:
        assert mode == "normal"
        forcer = getattr(run, "forcer", None)
        if forcer is None:
            forcer = run.forcer = ref_stack.Forcer(run.scheduler)
    out = []
    for phase in range(2):
        if run.tracker is not None:
            run.tracker.start_run()
        if forcer is not None:
            forcer.start_run(prompts, forced[phase])
        ids = fn([p.clone() for p in prompts], sp)
        ids = [[int(t) for t in r] for r in ids]
        margins = run.tracker.margins(prompts) if run.tracker is not None else None
        own = [list(x) for x in forcer.own] if forcer is not None else None
        out.append((ids, margins, run.state(), own))
    return out


def _append_flip_log(entry):
    os.makedirs(os.path.dirname(FLIP_LOG), exist_ok=True)
    data = []
    if os.path.isfile(FLIP_LOG):
        with open(FLIP_LOG) as f:
            data = json.load(f)
    data = [d for d in data if d["case"] != entry["case"]] + [entry]
    with open(FLIP_LOG, "w") as f:
        json.dump(data, f, indent=1)


def _same_state(case, phase, rstate, gstate):
    for key in ("req_to_token", "free_slots"):
        assert torch.equal(rstate[key], gstate[key]), f"{case} {phase}: {key} differs"
    for key in ("free_rows", "evictable_size", "kv_available"):
        assert rstate[key] == gstate[key], f"{case} {phase}: {key} {rstate[key]} vs {gstate[key]}"


def _compare_free(case, ref_runs, got_runs):
    """Free-running: each request identical for all 64 steps, or parting from the reference at a tie.  Once a
    request has parted, its later tokens -- and through them the radix tree's node sizes -- legitimately differ,
    so the index state is compared only while no request has parted (the forced comparison pins it always)."""
    parted = False
    for phase, (r, g) in zip(("cold", "radix-hit"), zip(ref_runs, got_runs)):
        (rids, rmarg, rstate, _), (gids, _, gstate, _) = r, g
        assert all(len(x) == N_NEW for x in rids) and all(len(x) == N_NEW for x in gids)
        assert all(len(m) == N_NEW for m in rmarg), "recorder lost a step"
        res = ref_stack.classify_divergence(rids, gids, rmarg)
        _append_flip_log({"case": f"{case}/free/{phase}", "steps_total": res["steps_total"],
                          "identical_requests": sum(a == b for a, b in zip(rids, gids)), "requests": len(rids),
                          "ties": res["ties"], "mismatches": res["mismatches"]})
        assert not res["mismatches"], f"{case} {phase}: greedy IDs differ away from a tie: {res['mismatches'][:3]}"
        parted = parted or bool(res["ties"])
        if not parted:
            _same_state(case, phase, rstate, gstate)
        print(f"[real-stack] {case} free {phase}: {sum(a == b for a, b in zip(rids, gids))}/{len(rids)} requests "
              f"identical over {N_NEW} steps; tie departures (request, step, margin): "
              f"{[(t['request'], t['step'], round(t['ref_margin'], 5)) for t in res['ties']]}")


def _compare_forced(case, ref_runs, got_runs):
    """Teacher-forced: index state bit-exact after every run, and the stack's own argmax equal to the reference's
    token at EVERY step of every request, ties (reference margin <= 2 bf16 ulps) excepted and logged."""
    for phase, (r, g) in zip(("cold", "radix-hit"), zip(ref_runs, got_runs)):
        (rids, rmarg, rstate, _), (gids, _, gstate, own) = r, g
        assert gids == rids, "forcing did not keep the trajectory"
        _same_state(case, phase, rstate, gstate)
        res = ref_stack.classify_forced(rids, own, rmarg)
        _append_flip_log({"case": f"{case}/forced/{phase}", "steps_total": res["steps_total"],
                          "steps_equal": res["steps_equal"], "ties": res["ties"], "mismatches": res["mismatches"]})
        assert not res["mismatches"], f"{case} {phase}: argmax differs away from a tie: {res['mismatches'][:3]}"
        print(f"[real-stack] {case} forced {phase}: {res['steps_equal']}/{res['steps_total']} steps identical, "
              f"{len(res['ties'])} ties (margins {[round(t['ref_margin'], 5) for t in res['ties']]}); index state "
              f"bit-exact")


# ----------------------------------------------------------------------------------------------- CPU
@needs_ref
def test_unchanged_scheduler_builds_on_dropin_modules_cpu():
    st = ref_loader.load_stack(shadow_dir=DROPIN_DIR)
    assert st.memory_pool.__file__.startswith(DROPIN_DIR)
    assert st.attn_backend.__file__.startswith(DROPIN_DIR) and st.forward_batch.__file__.startswith(DROPIN_DIR)
    assert not st.scheduler.__file__.startswith(DROPIN_DIR) and not st.qwen2.__file__.startswith(DROPIN_DIR)
    run = ref_stack.build(DROPIN_DIR, TINY, "cpu", torch.float32, 4096, 8, listen=False)
    sched = run.scheduler
    assert type(sched.token_to_kv_pool).__module__ == "memory_pool"
    assert sched.tree_cache.token_to_kv_pool is sched.token_to_kv_pool          # radix cache shares the drop-in pool
    sp = run.stack.sample.SamplingParams(max_new_tokens=4, do_sample=False, eos_id=-1)
    with pytest.raises(_lib.BsglError, match="no CPU path"):                     # the product never computes on CPU
        sched.run_batch([torch.arange(5)], sp)


@needs_ref
def test_recorder_and_tracker_on_the_pure_reference_cpu():
    """The listening harness itself, on the pure reference stack (fp32, CPU): it must not change a token, must
    attribute every kept sample to its (request, step), and two identical runs must classify as identical."""
    prompts = _prompts(TINY["vocab_size"], [37, 5, 64], seed=3)
    plain = ref_stack.build(None, TINY, "cpu", torch.float32, 4096, 8, listen=False)
    sp = plain.stack.sample.SamplingParams(max_new_tokens=12, do_sample=False, eos_id=-1)
    want = plain.scheduler.run_batch([p.clone() for p in prompts], sp)
    run = ref_stack.build(None, TINY, "cpu", torch.float32, 4096, 8)
    run.tracker.start_run()
    got = run.scheduler.run_batch([p.clone() for p in prompts], sp)
    assert got == want
    marg = run.tracker.margins(prompts)
    assert [len(m) for m in marg] == [12, 12, 12]
    assert all(t1 >= t2 for m in marg for t1, t2 in m)
    res = ref_stack.classify_divergence(want, got, marg)
    assert not res["ties"] and not res["mismatches"] and res["steps_compared_equal"] == 36
    # a forged departure is a mismatch unless the margin is a tie
    forged = [list(x) for x in got]
    forged[1][4] += 1
    res = ref_stack.classify_divergence(want, forged, marg)
    assert len(res["ties"]) + len(res["mismatches"]) == 1
    marg[1][4] = [1.0, 1.0]
    assert len(ref_stack.classify_divergence(want, forged, marg)["ties"]) == 1
    # teacher forcing: a second pure-reference stack fed a DIFFERENT trajectory follows it, records its own argmax,
    # and ends in the same index state as a stack that produced that trajectory itself would have sizes for
    other = ref_stack.build(None, TINY, "cpu", torch.float32, 4096, 8, listen=False)
    forcer = ref_stack.Forcer(other.scheduler)
    forcer.start_run(prompts, forged)
    assert other.scheduler.run_batch([p.clone() for p in prompts], sp) == forged
    assert forcer.own[0] == want[0] and forcer.own[2] == want[2]        # untouched requests predict the same
    assert forcer.own[1][:5] == want[1][:5]                             # ... and so does request 1 up to the forgery
    res = ref_stack.classify_forced(want, forcer.own, marg)
    assert not res["mismatches"] or all(m["request"] == 1 and m["step"] > 4 for m in res["mismatches"])


# ----------------------------------------------------------------------------------------------- GPU
@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["normal", "overlap"])
@pytest.mark.parametrize("name", sorted(GEOMETRIES))
def test_unchanged_stack_on_dropins_matches_pure_reference(name, mode):
    cfg, lens = GEOMETRIES[name]
    prompts = _prompts(cfg["vocab_size"], lens, seed=len(name))
    ref = ref_stack.build(None, cfg, "cuda", torch.bfloat16, **POOL)
    ref_runs = _run(ref, prompts, mode)
    ref_ids = [r[0] for r in ref_runs]
    for label, hooks in (("dropin", ()), ("dropin+K5K6K7", ("rope", "sampler", "prep")),
                         ("dropin+K5K6K7+graph", ("rope", "sampler", "prep", "graph"))):
        got = ref_stack.build(DROPIN_DIR, cfg, "cuda", torch.bfloat16, **POOL, listen=False, hooks=hooks)
        assert type(got.scheduler.token_to_kv_pool).__module__ == "memory_pool"
        assert got.stack.attn_backend.__file__.startswith(DROPIN_DIR)
        _compare_free(f"{name}/{mode}/{label}", ref_runs, _run(got, prompts, mode))
        del got
        if mode == "normal":     # the overlap worker samples on its own thread: free-running only
            forced = ref_stack.build(DROPIN_DIR, cfg, "cuda", torch.bfloat16, **POOL, listen=False, hooks=hooks)
            _compare_forced(f"{name}/{mode}/{label}", ref_runs, _run(forced, prompts, mode, forced=ref_ids))
            del forced


# Tie-free cases: the same geometries and prompt lengths, with the LM head "peaked" (ref_stack.peak_lm_head) so that
# the PURE REFERENCE's own top-1 / top-2 margin stays far above rounding noise at every one of the 64 steps of every
# request.  (peak, prompt seed) were chosen once from a probe of the pure reference alone (scripts/probe_margins.py,
# profiles/r3_probe_margins_bf16.jsonl: smallest margin 15 / 22 bf16 ulps at peak 4.0, where the 1.5B-heads case
# still has steps decided by the context-dependent part of the logits; >= 33 ulps at peak 5.0).  fp16 is the dtype the
# reference itself picks on CUDA (model_loader.py:56-64).
TIE_FREE = {
    ("qwen1.5b_heads", "bf16"): (4.0, 1003), ("qwen7b_heads", "bf16"): (4.0, 1000),
    ("qwen1.5b_heads", "fp16"): (5.0, 1000), ("qwen7b_heads", "fp16"): (5.0, 1000),
}
MIN_REF_MARGIN_ULPS = 4.0       # asserted about the reference BEFORE anything is compared (2 ulps is the tie rule)


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["normal", "overlap"])
@pytest.mark.parametrize("name,dt", sorted(TIE_FREE))
def test_tie_free_case_is_identical_for_all_64_steps(name, dt, mode):
    """north_star, literally: greedy token IDs identical for 64 steps, free running, EVERY request, cold and
    radix-hit, normal and overlap mode, on the drop-ins alone and with K5/K6/K7 and the CUDA graph installed --
    no tie rule involved: `ties` must be empty because the reference's margins leave no room for one."""
    cfg, lens = GEOMETRIES[name]
    peak, seed = TIE_FREE[(name, dt)]
    dtype = torch.bfloat16 if dt == "bf16" else torch.float16
    prompts = _prompts(cfg["vocab_size"], lens, seed=seed)
    ref = ref_stack.build(None, cfg, "cuda", dtype, **POOL, lm_head_peak=peak)
    ref_runs = _run(ref, prompts, mode)
    del ref
    worst = min((t1 - t2) / ref_stack.bf16_ulp(max(abs(t1), 1.0))
                for _, margins, _, _ in ref_runs for m in margins for t1, t2 in m)
    assert all(len(m) == N_NEW for _, margins, _, _ in ref_runs for m in margins), "recorder lost a step"
    assert worst > MIN_REF_MARGIN_ULPS, f"harness case is not tie-free: reference margin {worst} bf16 ulps"
    for label, hooks in (("dropin", ()), ("dropin+K5K6K7", ("rope", "sampler", "prep")),
                         ("dropin+K5K6K7+graph", ("rope", "sampler", "prep", "graph"))):
        got = ref_stack.build(DROPIN_DIR, cfg, "cuda", dtype, **POOL, listen=False, hooks=hooks, lm_head_peak=peak)
        got_runs = _run(got, prompts, mode)
        del got
        for phase, (r, g) in zip(("cold", "radix-hit"), zip(ref_runs, got_runs)):
            (rids, rmarg, rstate, _), (gids, _, gstate, _) = r, g
            same = sum(a == b for a, b in zip(rids, gids))
            res = ref_stack.classify_divergence(rids, gids, rmarg)
            _append_flip_log({"case": f"{name}/{dt}/{mode}/{label}/tie-free/{phase}", "steps_total": res["steps_total"],
                              "identical_requests": same, "requests": len(rids), "ties": res["ties"],
                              "mismatches": res["mismatches"], "reference_min_margin_bf16_ulps": round(worst, 2)})
            assert all(len(x) == N_NEW for x in gids)
            assert gids == rids, f"{name} {dt} {mode} {label} {phase}: {same}/{len(rids)} requests identical"
            assert not res["ties"] and not res["mismatches"]
            _same_state(f"{name}/{dt}/{mode}/{label}", phase, rstate, gstate)
        print(f"[real-stack] {name} {dt} {mode} {label}: {len(prompts)}/{len(prompts)} requests identical for "
              f"{N_NEW} free-running steps, cold and radix-hit; reference min margin {worst:.1f} bf16 ulps")


@needs_ref
@pytest.mark.gpu
def test_unchanged_stack_shared_prefixes_and_eviction():
    """Prompts that share prefixes (shared radix nodes -> aliased page-table cells, `cache_unfinished_req` cell
    rewrites, duplicate-slot frees, radix_cache.py:174-205) in a pool small enough that the radix cache must EVICT
    between waves (scheduler.py:409-421), but never retract.  Teacher-forced, so the radix tree holds the same
    token sequences in both stacks and the index state must stay bit-exact through every wave."""
    cfg, _ = GEOMETRIES["qwen7b_heads"]
    g = torch.Generator().manual_seed(7)
    base = torch.randint(0, cfg["vocab_size"], (400,), generator=g)
    prompts = [torch.cat([base[:n], torch.randint(0, cfg["vocab_size"], (m,), generator=g)])
               for n, m in ((400, 30), (400, 70), (250, 100), (250, 9), (100, 300), (0, 120))]
    # C = 640; probed on the pure reference (fp32, CPU): evicts from wave 2 on, never retracts
    pool = dict(max_total_tokens=2560, max_requests=4)
    ref = ref_stack.build(None, cfg, "cuda", torch.bfloat16, **pool)
    got = ref_stack.build(DROPIN_DIR, cfg, "cuda", torch.bfloat16, **pool, listen=False, hooks=("prep",))
    evictions = []
    inner_evict = got.scheduler.tree_cache.evict
    got.scheduler.tree_cache.evict = lambda n, cb: (evictions.append(n), inner_evict(n, cb))[1]
    for wave in range(4):      # later waves hit the tree, then outgrow it and force evictions
        g2 = torch.Generator().manual_seed(100 + wave)
        extra = [torch.cat([p[: 200 + 60 * wave], torch.randint(0, cfg["vocab_size"], (90,), generator=g2)])
                 for p in prompts[:3]]
        batch = prompts + extra if wave else prompts
        r = _run(ref, batch, "normal")
        d = _run(got, batch, "normal", forced=[x[0] for x in r])
        _compare_forced(f"shared-prefix/wave{wave}", r, d)
    assert evictions, "the scenario was meant to evict"
