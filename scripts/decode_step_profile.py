"""Development aid: where one end-to-end decode step goes (7B geometry, B=256, ctx 4096, the reference's unchanged
scheduler + model on the drop-ins with K5/K6/K7 installed, eager launches): per-kernel device time from
torch.profiler over a few steps.  bench.py is the contract; this only explains its decode tok/s number."""
import collections
import json
import os
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from baby_sglang_b200 import DROPIN_DIR, model_hooks  # noqa: E402
from baseline import ref_stack  # noqa: E402

Hq, Hkv, D, L, hidden, inter, vocab = 28, 4, 128, 28, 3584, 18944, 152064
batch, ctx, steps = 256, 4096, 3
cfg = dict(hidden_size=hidden, num_attention_heads=Hq, num_key_value_heads=Hkv, intermediate_size=inter,
           num_hidden_layers=L, vocab_size=vocab)
run = ref_stack.build(DROPIN_DIR, cfg, "cuda:0", torch.bfloat16, max_total_tokens=batch * (ctx + 16),
                      max_requests=batch, listen=False, init="device")
sched = run.stack.scheduler.Scheduler(run.model, tokenizer=None)
run.scheduler = sched
model_hooks.install_fused_rope_kv_store(run.model)
model_hooks.install_batched_greedy_sampler(sched)
model_hooks.install_device_batch_prep(sched)
ref_stack.seed_decode_state(run, batch, ctx, steps + 8, seed=0)
ref_stack.timed_decode_steps(run, 1, warmup=2)
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    r = ref_stack.timed_decode_steps(run, steps, warmup=0)
agg = collections.defaultdict(lambda: [0, 0.0])
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        n = ev.name.split("<")[0].replace("void ", "")[:60]
        agg[n][0] += 1
        agg[n][1] += ev.device_time_total if hasattr(ev, "device_time_total") else ev.cuda_time_total
tot = sum(v[1] for v in agg.values())
print(json.dumps({"dev_ms_per_step": [round(x, 2) for x in r["dev_ms"]], "kernel_us_per_step": round(tot / steps, 1)}))
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:25]:
    print(f"{t / steps:10.1f} us/step  {c // steps:5d} launches/step  {n}")
