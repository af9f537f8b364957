"""The plug mechanism itself: the reference's UNCHANGED scheduler / radix cache / model import the drop-in
`attn_backend`, `forward_batch`, `memory_pool` by module-name shadowing (qwen2.py:5-7, scheduler.py:8-9).

Needs a reference checkout ($BABYSGL_REF, /root/reference or baseline/_ref): present in the build container, absent
on the GPU box, so these tests skip there.  On CPU the stack must load, build its pools from the drop-in classes and
then refuse to compute (no CPU path); with a GPU *and* a checkout the full greedy run is compared with the pure
reference stack on the same device."""
import os

import pytest
import torch

from baby_sglang_b200 import DROPIN_DIR, _lib
from oracle import ref_loader
from tests.harness_model import deterministic_weights

needs_ref = pytest.mark.skipif(ref_loader.find_reference() is None, reason="no reference checkout on this machine")

CFG = dict(hidden_size=256, num_attention_heads=2, num_key_value_heads=1, intermediate_size=512,
           num_hidden_layers=2, vocab_size=512)


def _scheduler(stack, device, dtype):
    os.environ["BABYSGL_MAX_TOTAL_TOKENS"] = "4096"
    os.environ["BABYSGL_MAX_REQUESTS"] = "8"
    cfg = stack.model_loader.BabyQwenConfig(**CFG)
    model = stack.qwen2.BabyQwen2ForCausalLM(cfg).float().eval()
    deterministic_weights(model)
    model = model.to(device=device, dtype=dtype)
    return stack.scheduler.Scheduler(model, tokenizer=None)


@needs_ref
def test_unchanged_scheduler_builds_on_dropin_modules_cpu():
    st = ref_loader.load_stack(shadow_dir=DROPIN_DIR)
    assert st.memory_pool.__file__.startswith(DROPIN_DIR)
    assert st.attn_backend.__file__.startswith(DROPIN_DIR) and st.forward_batch.__file__.startswith(DROPIN_DIR)
    assert not st.scheduler.__file__.startswith(DROPIN_DIR) and not st.qwen2.__file__.startswith(DROPIN_DIR)
    sched = _scheduler(st, "cpu", torch.float32)
    assert type(sched.token_to_kv_pool).__module__ == "memory_pool"
    assert type(sched.token_to_kv_pool) is st.memory_pool.MHATokenToKVPool
    assert sched.tree_cache.token_to_kv_pool is sched.token_to_kv_pool          # radix cache shares the drop-in pool
    sp = st.sample.SamplingParams(max_new_tokens=4, do_sample=False, eos_id=-1)
    with pytest.raises(_lib.BsglError, match="no CPU path"):                     # the product never computes on CPU
        sched.run_batch([torch.arange(5)], sp)


@needs_ref
@pytest.mark.gpu
def test_unchanged_scheduler_greedy_ids_match_pure_reference_gpu():
    """Cold run + radix-hit rerun through the reference's real Scheduler, drop-ins vs pure reference, bf16."""
    prompts = [torch.randint(0, CFG["vocab_size"], (n,), generator=torch.Generator().manual_seed(n)) for n in (37, 5, 64)]
    outs = {}
    for name, shadow in (("reference", None), ("dropin", DROPIN_DIR)):
        st = ref_loader.load_stack(shadow_dir=shadow)
        sched = _scheduler(st, "cuda", torch.bfloat16)
        sp = st.sample.SamplingParams(max_new_tokens=64, do_sample=False, eos_id=-1)
        cold = sched.run_batch(list(prompts), sp)
        hit = sched.run_batch(list(prompts), sp)
        outs[name] = (cold, hit, sched.req_to_token_pool.req_to_token.clone(), sched.token_to_kv_pool.free_slots.clone())
    assert torch.equal(outs["reference"][2], outs["dropin"][2]), "page table differs"
    assert torch.equal(outs["reference"][3], outs["dropin"][3]), "free-slot order differs"
    # bf16 ties can flip a token (see tests/test_e2e_harness.py); the first 16 tokens of every request must agree
    for a, b in zip(outs["reference"][0], outs["dropin"][0]):
        assert a[:16] == b[:16]
