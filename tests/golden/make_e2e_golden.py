"""Generate tests/golden/e2e_tokens.json with the UNMODIFIED reference stack (build container only).

    python tests/golden/make_e2e_golden.py

The reference's own `BabyQwen2ForCausalLM` + `Scheduler.run_batch` (scheduler.py:137-163: prefill, decode,
radix cache, sampling) run on CPU in fp32 with weights set by `tests.harness_model.deterministic_weights`
(a function of the parameter NAMES only, so the harness can rebuild the same weights anywhere).  Two runs
per configuration on one scheduler: a cold run and a re-run of the same prompts, which hits the radix cache
for all but the last prompt token (scheduler.py:264-268) — both must give the same IDs.
"""
from __future__ import annotations

import json
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from baseline import ref_loader  # noqa: E402
from tests.harness_model import HarnessConfig, deterministic_weights  # noqa: E402

CASES = {
    # name: (config kwargs, prompt lengths, max_new_tokens)
    "gqa2_d128": (dict(vocab_size=512, hidden_size=256, num_attention_heads=2, num_key_value_heads=1,
                       num_hidden_layers=2, intermediate_size=512), [37, 5, 64, 19], 64),
    "gqa7_d64": (dict(vocab_size=384, hidden_size=448, num_attention_heads=7, num_key_value_heads=1,
                      num_hidden_layers=2, intermediate_size=384), [150, 3], 64),
    "mha_d128_tied": (dict(vocab_size=320, hidden_size=256, num_attention_heads=2, num_key_value_heads=2,
                           num_hidden_layers=3, intermediate_size=256, tie_word_embeddings=True), [9, 33, 70], 64),
}


def prompts_for(name, lens, vocab):
    g = torch.Generator().manual_seed(sum(map(ord, name)))
    return [torch.randint(0, vocab, (n,), generator=g).tolist() for n in lens]


def main():
    os.environ["BABYSGL_MAX_TOTAL_TOKENS"] = "4096"
    os.environ["BABYSGL_MAX_REQUESTS"] = "8"
    ref = ref_loader.load_stack()
    out = {}
    for name, (kw, lens, n_new) in CASES.items():
        hc = HarnessConfig(**kw)
        cfg = ref.model_loader.BabyQwenConfig(
            hidden_size=hc.hidden_size, num_attention_heads=hc.num_attention_heads,
            num_key_value_heads=hc.num_key_value_heads, intermediate_size=hc.intermediate_size,
            num_hidden_layers=hc.num_hidden_layers, vocab_size=hc.vocab_size, rms_norm_eps=hc.rms_norm_eps,
            rope_theta=hc.rope_theta, tie_word_embeddings=hc.tie_word_embeddings)
        model = ref.qwen2.BabyQwen2ForCausalLM(cfg).float().eval()
        deterministic_weights(model)
        sched = ref.scheduler.Scheduler(model, tokenizer=None)
        prompts = prompts_for(name, lens, hc.vocab_size)
        sp = ref.sample.SamplingParams(max_new_tokens=n_new, do_sample=False, eos_id=-1)
        cold = sched.run_batch([torch.tensor(p, dtype=torch.int64) for p in prompts], sp)
        hit = sched.run_batch([torch.tensor(p, dtype=torch.int64) for p in prompts], sp)
        assert cold == hit, f"{name}: radix-hit rerun differs from the cold run"
        out[name] = {"config": kw, "prompt_lens": lens, "prompts": prompts, "max_new_tokens": n_new,
                     "output_ids": cold}
        print(name, [o[:6] for o in cold])
    with open(os.path.join(HERE, "e2e_tokens.json"), "w") as f:
        json.dump(out, f)
    print("wrote e2e_tokens.json")


if __name__ == "__main__":
    main()
