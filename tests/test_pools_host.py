"""Host-side allocators of the drop-in `memory_pool` replayed against a trace recorded on the reference's own
pools (tests/golden/pool_trace.json, made by tests/golden/make_pool_trace.py; memory_pool.py:29-41,68-84).
Bit-exact: every alloc result, every available_size, and the final free-list order."""
import json
import os
import sys

import pytest
import torch

from baby_sglang_b200 import DROPIN_DIR

if DROPIN_DIR not in sys.path:
    sys.path.insert(0, DROPIN_DIR)
import memory_pool  # noqa: E402  (the drop-in)

with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pool_trace.json")) as _f:
    TRACES = json.load(_f)


@pytest.mark.parametrize("name", sorted(TRACES))
def test_pool_trace_matches_reference(name):
    tr = TRACES[name]
    kv = memory_pool.MHATokenToKVPool(size=tr["size"], dtype=torch.float32, head_num=1, head_dim=8, layer_num=1,
                                      device="cpu")
    rq = memory_pool.ReqToTokenPool(size=16, max_context_len=8, device="cpu")
    for kind, arg, want, avail in tr["ops"]:
        if kind == "alloc":
            got = kv.alloc(arg)
            assert (None if got is None else got.tolist()) == want
            if got is not None:
                assert got.dtype == torch.int32
            assert kv.available_size() == avail
        elif kind == "free":
            kv.free(torch.tensor(arg, dtype=torch.int32))
            assert kv.available_size() == avail
        elif kind == "ralloc":
            got = rq.alloc(arg)
            assert (None if got is None else list(got)) == want
            assert rq.available_size() == avail
        elif kind == "rfree":
            rq.free(arg)
            assert rq.available_size() == avail
    assert kv.free_slots.tolist() == tr["final_free_slots"]
    assert list(rq.free_slots) == tr["final_rows"]


def test_pool_surface_matches_reference_names():
    kv = memory_pool.MHATokenToKVPool(size=10, dtype=torch.bfloat16, head_num=2, head_dim=64, layer_num=3,
                                      device="cpu")
    assert len(kv.k_buffer) == 3 and kv.k_buffer[0].shape == (11, 2, 64) and kv.v_buffer[2].dtype == torch.bfloat16
    k, v = kv.get_kv_buffer(1)
    assert k.data_ptr() == kv.k_buffer[1].data_ptr() and v.data_ptr() == kv.v_buffer[1].data_ptr()
    assert kv.free_slots.tolist() == list(range(1, 11)) and kv.free_slots.dtype == torch.int32   # slot 0 is the pad
    kv.clear()
    assert kv.available_size() == 10
    rq = memory_pool.ReqToTokenPool(size=4, max_context_len=16, device="cpu")
    assert rq.req_to_token.shape == (4, 16) and rq.req_to_token.dtype == torch.int32
    rq.write((torch.tensor([1, 2]),), torch.ones(2, 16, dtype=torch.int32))   # scheduler.py:442 style
    rq.write((3, slice(2, 5)), torch.tensor([7, 8, 9], dtype=torch.int32))    # radix_cache.py:196-199 style
    assert rq.req_to_token[3, 2:5].tolist() == [7, 8, 9] and int(rq.req_to_token[1].sum()) == 16
    if hasattr(torch, "float8_e4m3fn"):
        f8 = memory_pool.MHATokenToKVPool(size=4, dtype=torch.float8_e4m3fn, head_num=1, head_dim=16, layer_num=1,
                                          device="cpu")
        with pytest.raises(NotImplementedError):       # memory_pool.py:141-168
            f8.get_kv_buffer(0)
