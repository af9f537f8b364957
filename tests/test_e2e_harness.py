"""End-to-end greedy decoding (north star: "greedy-decoded token IDs identical for 64 steps").

Chain of evidence:
  1. CPU, fp32:  harness model + oracle backend  ==  tests/golden/e2e_tokens.json, the IDs the reference's own
     `BabyQwen2ForCausalLM` + `Scheduler.run_batch` produced (tests/golden/make_e2e_golden.py) — this pins the
     harness (model wiring, RoPE, page-table handling, greedy loop) to the reference.
  2. GPU, bf16:  harness model + CUDA drop-in backend  ==  harness model + oracle backend on the same device,
     dtype, weights and prompts: only the three shadowed modules differ.
  (bf16 logits tie often on a random-init model; the GPU leg feeds the oracle-side tokens to the CUDA stack and
  excuses a differing argmax ONLY where the oracle side's own top-1/top-2 margin is within 2 bf16 ulps.)
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
    tr = greedy_generate(m, st, case["prompts"], n_new, torch.device("cpu"), torch.float32)
    assert tr.ids == case["output_ids"]
    # prefix reuse (all but the last prompt token, like the reference's radix hit): same IDs
    reuse = [len(p) - 1 for p in case["prompts"]]
    tr2 = greedy_generate(m, st, case["prompts"], n_new, torch.device("cpu"), torch.float32,
                          reuse_prefix=reuse, pools=tr.pools)
    assert tr2.ids == case["output_ids"]


GPU_CASES = {
    # bigger than the CPU goldens: real head counts of the BASELINE configs, short hidden size
    "qwen1.5b_heads": (dict(vocab_size=1024, hidden_size=12 * 128, num_attention_heads=12, num_key_value_heads=2,
                            num_hidden_layers=2, intermediate_size=1024), [300, 17, 129, 64, 1, 255]),
    "qwen7b_heads": (dict(vocab_size=1024, hidden_size=28 * 128, num_attention_heads=28, num_key_value_heads=4,
                          num_hidden_layers=2, intermediate_size=1024), [257, 40, 700]),
}


import math


def bf16_ulp(x: float) -> float:
    """Spacing of bf16 numbers at |x| (8 significant bits): the reference computes logits in the model dtype."""
    return 2.0 ** (math.floor(math.log2(max(abs(x), 2.0 ** -126))) - 7)


def _check_against(ref, got, what):
    """`ref` ran free (oracle backend), `got` was fed ref's tokens (CUDA backend).  Every step must predict the
    same token, except where the oracle side itself had a numerical tie: top-1/top-2 logits within 2 bf16 ulps of
    each other — there the argmax is decided by rounding noise in either stack.  The
    margin of every excused step is reported in the assertion message; nothing else is excused."""
    steps = sum(len(x) for x in ref.ids)
    excused, bad = [], []
    for i, (a, b) in enumerate(zip(ref.ids, got.pred)):
        for s_, (x, y) in enumerate(zip(a, b)):
            if x != y:
                tol = 2 * bf16_ulp(max(abs(ref.top1[i][s_]), 1.0))
                (excused if ref.margin[i][s_] <= tol else bad).append((i, s_, ref.margin[i][s_], tol))
    assert not bad, f"{what}: token IDs differ away from ties (request, step, margin, tie tolerance): {bad[:5]}"
    assert len(excused) <= max(2, steps // 20), f"{what}: too many tie flips {len(excused)}/{steps}: {excused[:5]}"
    return len(excused), steps


def _gpu_compare(cfg_kw, prompts, n_new, dtype):
    from tests.oracle_stack import dropin_stack, oracle_stack
    dev = torch.device("cuda")
    case = {"config": cfg_kw}
    so, sd = oracle_stack(), dropin_stack()
    mo = _model(case, so.backend_forward, dev, dtype)
    md = _model(case, sd.backend_forward, dev, dtype)
    ref = greedy_generate(mo, so, prompts, n_new, dev, dtype)
    got = greedy_generate(md, sd, prompts, n_new, dev, dtype, forced_ids=ref.ids)
    torch.cuda.synchronize()
    flips, steps = _check_against(ref, got, "cold run")
    # page table and slot allocator state are bit-exact between the stacks
    assert torch.equal(ref.pools[0].req_to_token, got.pools[0].req_to_token)
    assert torch.equal(ref.pools[1].free_slots, got.pools[1].free_slots)
    # prefix reuse: keep an uneven share of every prompt, extend the rest (K3 with prefix > 0)
    reuse = [max(0, (2 * len(p)) // 3) for p in prompts]
    ref2 = greedy_generate(mo, so, prompts, n_new, dev, dtype, reuse_prefix=reuse, pools=ref.pools)
    got2 = greedy_generate(md, sd, prompts, n_new, dev, dtype, reuse_prefix=reuse, pools=got.pools,
                           forced_ids=ref2.ids)
    flips2, _ = _check_against(ref2, got2, "prefix-reuse run")
    free_running = greedy_generate(md, sd, prompts, n_new, dev, dtype)
    identical = free_running.ids == ref.ids
    print(f"[e2e] {steps} steps: tie flips cold={flips} reuse={flips2}; free-running IDs identical: {identical}; "
          f"min oracle margin {min(min(m) for m in ref.margin):.3e}")
    return identical


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


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(GPU_CASES))
def test_gpu_fused_rope_kv_store_hook_is_bit_identical(name):
    """`install_fused_rope_kv_store` (K5: split + RoPE + KV store in one launch, attention
    with save_kv_cache=False) must not change a single bit: same greedy IDs, same logit margins, same KV slabs
    and page table as the unfused drop-in path (which stores through K1 after torch RoPE)."""
    from baby_sglang_b200.model_hooks import install_fused_rope_kv_store, uninstall_fused_rope_kv_store
    from tests.oracle_stack import dropin_stack
    kw, lens = GPU_CASES[name]
    g = torch.Generator().manual_seed(len(name))
    prompts = [torch.randint(0, kw["vocab_size"], (n,), generator=g).tolist() for n in lens]
    dev, dtype = torch.device("cuda"), torch.bfloat16
    sd = dropin_stack()
    m = _model({"config": kw}, sd.backend_forward, dev, dtype)
    plain = greedy_generate(m, sd, prompts, 24, dev, dtype)
    assert install_fused_rope_kv_store(m) == kw["num_hidden_layers"]
    fused = greedy_generate(m, sd, prompts, 24, dev, dtype)
    assert fused.ids == plain.ids
    assert fused.top1 == plain.top1 and fused.margin == plain.margin
    assert torch.equal(plain.pools[0].req_to_token, fused.pools[0].req_to_token)
    for l in range(kw["num_hidden_layers"]):
        used = plain.pools[0].req_to_token.flatten().long().unique()
        assert torch.equal(plain.pools[1].k_buffer[l][used], fused.pools[1].k_buffer[l][used])
        assert torch.equal(plain.pools[1].v_buffer[l][used], fused.pools[1].v_buffer[l][used])
    # prefix reuse (extend with prefix > 0) through the fused path
    reuse = [max(0, (2 * len(p)) // 3) for p in prompts]
    fused2 = greedy_generate(m, sd, prompts, 24, dev, dtype, reuse_prefix=reuse, pools=fused.pools)
    uninstall_fused_rope_kv_store(m)
    plain2 = greedy_generate(m, sd, prompts, 24, dev, dtype, reuse_prefix=reuse, pools=plain.pools)
    assert fused2.ids == plain2.ids and fused2.margin == plain2.margin
