"""KV-head tensor parallelism on GPUs (one process per GPU under torchrun, NCCL): the reference's UNCHANGED scheduler +
qwen2 model, sharded by the drop-in `tp_runtime` (`shard_attention_heads`, `kv_pool_config`, optionally
`FusedOProjAllReduce` = K8 and `shard_mlp`), against the UNSHARDED drop-in stack on the same GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node {2,4,8} --master-addr 127.0.0.1 \
        --master-port 29521 tests/multi_gpu_tp.py

Per variant (o_proj all-reduce by NCCL / by K8 / K8 + TP MLP with down_proj on K8 too / the same with a communicator
smaller than the prefill, i.e. K8 in row blocks), cold run + radix-hit rerun, 32 greedy tokens,
teacher-forced with the unsharded run's tokens:
  * index state (page table, KV free-slot FIFO, free rows, radix sizes) bit-exact with the unsharded run's;
  * every rank's own argmax at every step equal to the unsharded token, except at ties: the unsharded side's
    top-1/top-2 margin within 4 bf16 ulps (each rank rounds its o_proj partial to bf16 before the sum, so the TP
    logits carry a little more rounding noise than the unsharded ones);
  * every rank predicts the SAME ids (the all-reduced activations are bit-identical on all ranks, which is what keeps
    the replicated schedulers in lock-step).
Needs the staged reference checkout (scripts/stage_reference.py).  One JSON line per variant on rank 0; exit code 1
on failure.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from baby_sglang_b200 import DROPIN_DIR  # noqa: E402
from baseline import ref_loader, ref_stack  # noqa: E402

N_NEW = 32
# Qwen2.5-32B head geometry (40 q / 8 kv, D = 128: shards over 2, 4 and 8 ranks; per-rank Hkv 4 / 2 / 1 exercises all
# three head-group widths of K2 and the 256-byte-row producer path), short MLP / vocab
CFG = dict(hidden_size=40 * 128, num_attention_heads=40, num_key_value_heads=8, intermediate_size=2048,
           num_hidden_layers=2, vocab_size=1024)
POOL = dict(max_total_tokens=16384, max_requests=8)
LENS = [300, 17, 129, 64, 1, 700]


def _run(run, prompts, forced=None):
    sp = run.stack.sample.SamplingParams(max_new_tokens=N_NEW, do_sample=False, eos_id=-1)
    forcer = ref_stack.Forcer(run.scheduler) if forced is not None else None
    out = []
    for phase in range(2):
        if run.tracker is not None:
            run.tracker.start_run()
        if forcer is not None:
            forcer.start_run(prompts, forced[phase])
        ids = [[int(t) for t in r] for r in run.scheduler.run_batch([p.clone() for p in prompts], sp)]
        out.append(dict(ids=ids, margins=run.tracker.margins(prompts) if run.tracker else None, state=run.state(),
                        own=[list(x) for x in forcer.own] if forcer else None))
    return out


def check(rank, world, group):
    """All variants on an initialised process group (also called by bench.py's `tp_parity` leg).  True if all pass."""
    g = torch.Generator().manual_seed(11)
    prompts = [torch.randint(0, CFG["vocab_size"], (n,), generator=g) for n in LENS]
    base = ref_stack.build(DROPIN_DIR, CFG, "cuda", torch.bfloat16, **POOL)
    ref_runs = _run(base, prompts)
    del base
    ok = True
    # the last variant sizes the communicator BELOW the 1 211-token prefill: o_proj and down_proj then run K8 in
    # 500-row blocks (500 + 500 + 211) instead of leaving the fused path
    for name, fused, mlp, cap in (("nccl", False, False, 2048), ("k8", True, False, 2048),
                                  ("k8+tp_mlp", True, True, 2048), ("k8+tp_mlp, 500-row blocks", True, True, 500)):
        run = ref_stack.build(DROPIN_DIR, CFG, "cuda", torch.bfloat16, **POOL, listen=False,
                              tp=dict(rank=rank, world=world, group=group, fused=fused, mlp=mlp, max_tokens=cap))
        if fused:       # no library GEMM + NCCL pair left on the TP layer path
            fwd = [m.o_proj.forward for m in run.scheduler.model.modules() if hasattr(m, "o_proj")]
            assert fwd and all(getattr(f, "__name__", "") == "forward" and not hasattr(f, "__self__") for f in fwd)
            if mlp:
                fwd = [m.down_proj.forward for m in run.scheduler.model.modules() if hasattr(m, "down_proj")]
                assert fwd and all(not hasattr(f, "__self__") for f in fwd), "down_proj is not on K8"
        assert run.scheduler.token_to_kv_pool.head_num == CFG["num_key_value_heads"] // world
        runs = _run(run, prompts, forced=[r["ids"] for r in ref_runs])
        line = {"variant": name, "world": world, "exchange": getattr(run.comm, "last_exchange", "nccl all_reduce")}
        good = True
        for phase, (r, d) in zip(("cold", "radix-hit"), zip(ref_runs, runs)):
            same_state = all(torch.equal(r["state"][k], d["state"][k]) for k in ("req_to_token", "free_slots")) and \
                all(r["state"][k] == d["state"][k] for k in ("free_rows", "evictable_size", "kv_available"))
            res = ref_stack.classify_forced(r["ids"], d["own"], r["margins"], ulps=4.0)
            gathered = [None] * world
            dist.all_gather_object(gathered, d["own"], group=group)
            lockstep = all(x == gathered[0] for x in gathered)
            line[phase] = {"index_state_bit_exact": same_state, "steps": res["steps_total"],
                           "steps_equal": res["steps_equal"], "ties": len(res["ties"]),
                           "tie_margins": [round(t["ref_margin"], 5) for t in res["ties"]][:8],
                           "mismatches": res["mismatches"][:3], "ranks_predict_identical_ids": lockstep}
            good &= same_state and not res["mismatches"] and lockstep
        flag = torch.tensor([1 if good else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        line["ok"] = bool(flag.item())
        ok &= line["ok"]
        if rank == 0:
            print(json.dumps(line), flush=True)
        if run.comm is not None:
            run.comm.close()
        del run
        torch.cuda.empty_cache()
    return ok


def main():
    if ref_loader.find_reference() is None:
        print(json.dumps({"skipped": "no reference checkout (scripts/stage_reference.py)"}))
        return 0
    sys.path.insert(0, DROPIN_DIR)
    import tp_runtime
    rank, world, group = tp_runtime.init_tp()
    ok = check(rank, world, group)
    dist.barrier(group=group)
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
