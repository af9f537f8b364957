"""End-to-end greedy decoding (north star: "greedy-decoded token IDs identical for 64 steps").

Chain of evidence:
  1. CPU, fp32:  harness model + oracle backend  ==  tests/golden/e2e_tokens.json, the IDs the reference's own
     `BabyQwen2ForCausalLM` + `Scheduler.run_batch` produced (tests/golden/make_e2e_golden.py) — this pins the
     harness (model wiring, RoPE, page-table handling, greedy loop) to the reference.
  2. GPU, bf16:  harness model + CUDA drop-in backend  ==  harness model + oracle backend on the same device,
     dtype, weights and prompts: only the three shadowed modules differ.
Both legs also re-run the prompts with a reused prefix (what a radix-cache hit produces, scheduler.py:423-442):
the extend path with prefix > 0 must give the same IDs as the cold run.
"""
import json
import os

import pytest
import torch

from tests.harness_model import HarnessConfig, HarnessQwen2, deterministic_weights, greedy_generate

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "e2e_tokens.json")
with open(GOLDEN) as _f:
    CASES = json.load(_f)


def _model(case, backend, device, dtype):
    cfg = HarnessConfig(**case["config"])
    m = HarnessQwen2(cfg, backend).float().eval()
    deterministic_weights(m)
    return m.to(device=device, dtype=dtype)


@pytest.mark.parametrize("name", sorted(CASES))
def test_harness_reproduces_reference_scheduler_ids(name):
    from tests.oracle_stack import oracle_stack
    case = CASES[name]
    st = oracle_stack()
    m = _model(case, st.backend_forward, "cpu", torch.float32)
    n_new = case["max_new_tokens"]
    ids, pools, _ = greedy_generate(m, st, case["prompts"], n_new, torch.device("cpu"), torch.float32)
    assert ids == case["output_ids"]
    # prefix reuse (all but the last prompt token, like the reference's radix hit): same IDs
    reuse = [len(p) - 1 for p in case["prompts"]]
    ids2, _, _ = greedy_generate(m, st, case["prompts"], n_new, torch.device("cpu"), torch.float32,
                                 reuse_prefix=reuse, pools=pools)
    assert ids2 == case["output_ids"]


GPU_CASES = {
    # bigger than the CPU goldens: real head counts of the BASELINE configs, short hidden size
    "qwen1.5b_heads": (dict(vocab_size=1024, hidden_size=12 * 128, num_attention_heads=12, num_key_value_heads=2,
                            num_hidden_layers=2, intermediate_size=1024), [300, 17, 129, 64, 1, 255]),
    "qwen7b_heads": (dict(vocab_size=1024, hidden_size=28 * 128, num_attention_heads=28, num_key_value_heads=4,
                          num_hidden_layers=2, intermediate_size=1024), [257, 40, 700]),
}


def _gpu_compare(cfg_kw, prompts, n_new, dtype, mode_env=None):
    from tests.oracle_stack import dropin_stack, oracle_stack
    dev = torch.device("cuda")
    case = {"config": cfg_kw}
    so, sd = oracle_stack(), dropin_stack()
    mo = _model(case, so.backend_forward, dev, dtype)
    md = _model(case, sd.backend_forward, dev, dtype)
    ref_ids, ref_pools, margins = greedy_generate(mo, so, prompts, n_new, dev, dtype)
    got_ids, got_pools, _ = greedy_generate(md, sd, prompts, n_new, dev, dtype)
    torch.cuda.synchronize()
    assert got_ids == ref_ids, f"token IDs diverge (min top-1/top-2 margin on the oracle side {min(margins):.3e})"
    # page table and slot allocator state are bit-exact between the stacks
    assert torch.equal(ref_pools[0].req_to_token, got_pools[0].req_to_token)
    assert torch.equal(ref_pools[1].free_slots, got_pools[1].free_slots)
    # prefix reuse: keep an uneven share of every prompt, extend the rest (K3 with prefix > 0)
    reuse = [max(0, (2 * len(p)) // 3) for p in prompts]
    got2, _, _ = greedy_generate(md, sd, prompts, n_new, dev, dtype, reuse_prefix=reuse, pools=got_pools)
    ref2, _, _ = greedy_generate(mo, so, prompts, n_new, dev, dtype, reuse_prefix=reuse, pools=ref_pools)
    assert got2 == ref2
    return min(margins)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_greedy_ids_match_oracle_backend_golden_configs(name):
    case = CASES[name]
    _gpu_compare(case["config"], case["prompts"], case["max_new_tokens"], torch.bfloat16)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(GPU_CASES))
def test_gpu_greedy_ids_match_oracle_backend(name):
    kw, lens = GPU_CASES[name]
    g = torch.Generator().manual_seed(len(name))
    prompts = [torch.randint(0, kw["vocab_size"], (n,), generator=g).tolist() for n in lens]
    _gpu_compare(kw, prompts, 64, torch.bfloat16)
