"""Probe (pure reference only, on the GPU): for each geometry / LM-head peak / prompt seed, the smallest top-1 / top-2
margin over all 64 steps of all requests (cold run + radix-hit rerun), in bf16 ulps of the top-1 logit, and how many
steps were NOT the successor-chain token (decided by the context-dependent part of the logits).  Used once to choose
the tie-free cases of tests/test_reference_stack.py; output kept under profiles/."""
import json
import os
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from baseline import ref_stack  # noqa: E402
from tests.test_reference_stack import GEOMETRIES, N_NEW, POOL, _prompts, _run  # noqa: E402


def main():
    dtype = torch.float16 if "--fp16" in sys.argv else torch.bfloat16
    peaks = [float(x) for x in os.environ.get("PEAKS", "5,5.75,6.5").split(",")]
    seeds = range(int(os.environ.get("SEEDS", "6")))
    for name in sorted(GEOMETRIES):
        cfg, lens = GEOMETRIES[name]
        for peak in peaks:
            ref = ref_stack.build(None, cfg, "cuda", dtype, **POOL, lm_head_peak=peak)
            probe = ref_stack.build(None, cfg, "cpu", torch.float32, **POOL, listen=False, lm_head_peak=peak)
            inv = ref_stack.peak_lm_head(probe.model, 0.0)
            del probe
            for seed in seeds:
                prompts = _prompts(cfg["vocab_size"], lens, seed=1000 + seed)
                runs = _run(ref, prompts, "normal")
                worst, off_chain, top1s = 1e9, 0, []
                for ids, margins, _, _ in runs:
                    for r, (row, m) in enumerate(zip(ids, margins)):
                        prev = int(prompts[r][-1])
                        for tok, (t1, t2) in zip(row, m):
                            worst = min(worst, (t1 - t2) / ref_stack.bf16_ulp(max(abs(t1), 1.0)))
                            off_chain += int(tok != int(inv[prev]))
                            top1s.append(t1)
                            prev = tok
                print(json.dumps({"geometry": name, "dtype": str(dtype), "peak": peak, "prompt_seed": 1000 + seed,
                                  "min_margin_bf16_ulps": round(worst, 2), "steps": len(top1s),
                                  "steps_off_successor_chain": off_chain,
                                  "mean_top1": round(sum(top1s) / len(top1s), 2),
                                  "rerun_ids_equal_cold": runs[0][0] == runs[1][0]}), flush=True)
            del ref
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
