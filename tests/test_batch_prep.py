"""K7: page-table writes and the free-slot FIFO on the device, bit-exact against a trace the
reference's own `Scheduler._prepare_prefill_batch` / `_prepare_decode_batch` produced
(tests/golden/batch_prep_trace.json, generator tests/golden/make_batch_prep_golden.py; scheduler.py:381-507,
memory_pool.py:68-84).

CPU: the oracle restatement replays the trace.  GPU: the scheduler hooks (`install_device_batch_prep`) replay it on
the drop-in pools with the device ring; every batch tensor, every live page-table row and the FIFO order must match.
"""
import json
import os
import sys
from types import SimpleNamespace

import pytest
import torch

from baby_sglang_b200 import DROPIN_DIR
from oracle import paged_attention_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "batch_prep_trace.json")) as _f:
    TRACE = json.load(_f)
with open(os.path.join(HERE, "golden", "pool_trace.json")) as _f:
    POOL_TRACES = json.load(_f)


def test_oracle_replays_reference_batch_prep_trace():
    C, R = TRACE["C"], TRACE["max_requests"]
    table = torch.zeros(R, C, dtype=torch.int32)
    fifo = O.SlotFifo(TRACE["total_tokens"])
    free_rows = list(range(R))
    live = {}     # rid -> dict(row, seq_len, prefix)
    for op in TRACE["ops"]:
        kind = op["op"]
        if kind == "alloc_prefix":
            assert fifo.alloc(op["n"]) == op["got"]
            continue
        if kind == "prefill":
            reqs = op["requests"]
            rows, free_rows = free_rows[:len(reqs)], free_rows[len(reqs):]
            seq = [len(r["input_ids"]) for r in reqs]
            n_new = sum(s - len(r["prefix_indices"]) for s, r in zip(seq, reqs))
            new = fifo.alloc(n_new)
            O.prefill_page_table_rows(table, rows, [r["prefix_indices"] for r in reqs], new, seq)
            for rid, row, s in zip(op["rids"], rows, seq):
                live[rid] = dict(row=row, seq_len=s)
            assert new == op["batch"]["out_cache_loc"] and rows == op["batch"]["req_pool_indices"]
            md = O.build_metadata(table, torch.tensor(rows), torch.tensor(seq),
                                  torch.tensor([len(r["prefix_indices"]) for r in reqs]))
            assert md.positions.tolist() == op["batch"]["positions"]
            assert md.extended_lens.tolist() == op["batch"]["extended_lens"]
        elif kind == "decode":
            rows = [live[r]["row"] for r in op["rids"]]
            before = [live[r]["seq_len"] for r in op["rids"]]
            new = fifo.alloc(len(rows))
            O.decode_page_table_append(table, rows, before, new)
            for r in op["rids"]:
                live[r]["seq_len"] += 1
            assert new == op["batch"]["out_cache_loc"]
            assert [live[r]["seq_len"] for r in op["rids"]] == op["batch"]["seq_lens"] == op["seq_len"]
            assert [s - 1 for s in op["batch"]["seq_lens"]] == op["batch"]["positions"]     # forward_batch.py:72
        elif kind == "finish":
            st = live.pop(op["rid"])
            fifo.free(table[st["row"], op["keep_prefix"]:st["seq_len"]].tolist())
            free_rows.append(st["row"])
        after = op["after"]
        assert fifo.slots == after["free_slots"] and free_rows == after["free_rows"]
        for rid, want in after["rows"].items():
            assert table[live[int(rid)]["row"]].tolist() == want, (kind, rid)


def _fake_scheduler(device):
    if DROPIN_DIR not in sys.path:
        sys.path.insert(0, DROPIN_DIR)
    import memory_pool
    C, R = TRACE["C"], TRACE["max_requests"]
    kv = memory_pool.MHATokenToKVPool(size=TRACE["total_tokens"], dtype=torch.bfloat16, head_num=1, head_dim=64,
                                      layer_num=1, device=device)
    rq = memory_pool.ReqToTokenPool(size=R, max_context_len=C, device=device)
    tree = SimpleNamespace(evict=lambda n, cb: None, evictable_size=lambda: 0)
    return SimpleNamespace(req_to_token_pool=rq, token_to_kv_pool=kv, tree_cache=tree, device=device)


@pytest.mark.gpu
def test_device_batch_prep_hooks_replay_reference_trace():
    from baby_sglang_b200 import model_hooks, ops
    dev = "cuda"
    sched = _fake_scheduler(dev)
    assert sched.token_to_kv_pool._dev_ring
    live = {}
    launches0 = ops.LAUNCH_COUNT
    for op in TRACE["ops"]:
        kind = op["op"]
        if kind == "alloc_prefix":
            got = sched.token_to_kv_pool.alloc(op["n"])
            assert got.is_cuda and got.dtype == torch.int32 and got.tolist() == op["got"]
            continue
        if kind == "prefill":
            reqs = []
            for rid, r in zip(op["rids"], op["requests"]):
                q = SimpleNamespace(input_ids=torch.tensor(r["input_ids"]), output_ids=[],
                                    prefix_indices=torch.tensor(r["prefix_indices"], dtype=torch.int32, device=dev),
                                    num_cached_tokens=len(r["prefix_indices"]), req_pool_idx=None, seq_len=0,
                                    overlap_next_input_id=None, is_retracted=False)
                live[rid] = q
                reqs.append(q)
            fb = model_hooks._device_prepare_prefill_batch(sched, reqs)
            assert [q.req_pool_idx for q in reqs] == op["req_pool_idx"] and [q.seq_len for q in reqs] == op["seq_len"]
            assert all(isinstance(q.req_pool_idx, int) for q in reqs)
        elif kind == "decode":
            reqs = [live[r] for r in op["rids"]]
            for q, t in zip(reqs, op["last_ids"]):
                q.output_ids.append(t)
            fb = model_hooks._device_prepare_decode_batch(sched, reqs)
            assert [q.seq_len for q in reqs] == op["seq_len"]
        elif kind == "finish":
            q = live.pop(op["rid"])
            view = sched.req_to_token_pool.req_to_token[q.req_pool_idx, op["keep_prefix"]:q.seq_len]
            sched.token_to_kv_pool.free(view)                      # a device view: no .cpu() round trip
            sched.req_to_token_pool.free(q.req_pool_idx)
            fb = None
        if fb is not None:
            want = op["batch"]
            assert fb.is_prefill == (kind == "prefill") and fb.batch_size == len(want["req_pool_indices"])
            for name in ("input_ids", "req_pool_indices", "seq_lens", "out_cache_loc", "positions", "prefix_lens",
                         "extended_lens"):
                got = getattr(fb, name)
                assert (None if got is None else got.tolist()) == want[name], (kind, name)
            assert fb.out_cache_loc.dtype == torch.int32 and fb.seq_lens.dtype == torch.int64
            assert fb.req_pool_indices.dtype == torch.int64 and fb.positions.dtype == torch.int64
        after = op["after"]
        assert sched.token_to_kv_pool.free_slots.tolist() == after["free_slots"], kind
        assert list(sched.req_to_token_pool.free_slots) == after["free_rows"]
        table = sched.req_to_token_pool.req_to_token.cpu()
        for rid, row in after["rows"].items():
            assert table[live[int(rid)].req_pool_idx].tolist() == row, (kind, rid)
    assert ops.LAUNCH_COUNT > launches0     # the CUDA path ran (K7 + K4 launches)


@pytest.mark.gpu
def test_decode_hook_rejects_bad_requests_like_the_reference():
    from baby_sglang_b200 import model_hooks
    sched = _fake_scheduler("cuda")
    ok = dict(output_ids=[3], req_pool_idx=0, seq_len=4, overlap_next_input_id=None, is_retracted=False)
    for patch, msg in ((dict(req_pool_idx=None), "missing req_pool_idx"), (dict(is_retracted=True), "was retracted"),
                       (dict(output_ids=[]), "no output tokens")):
        with pytest.raises(RuntimeError, match=msg):                 # scheduler.py:466-482
            model_hooks._device_prepare_decode_batch(sched, [SimpleNamespace(**{**ok, **patch})])
    small = _fake_scheduler("cuda")
    small.token_to_kv_pool.alloc(TRACE["total_tokens"])
    with pytest.raises(RuntimeError, match="Failed to allocate KV slots"):   # scheduler.py:460-461
        model_hooks._device_prepare_decode_batch(small, [SimpleNamespace(**ok)])
    # overlap mode feeds a negative placeholder id through unchanged (overlap_worker.py:74-81)
    q = SimpleNamespace(**{**ok, "overlap_next_input_id": -7})
    fb = model_hooks._device_prepare_decode_batch(sched, [q])
    assert fb.input_ids.tolist() == [-7] and q.seq_len == 5
    assert int(sched.req_to_token_pool.req_to_token[0, 4]) == int(fb.out_cache_loc[0])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(POOL_TRACES))
def test_device_slot_ring_matches_reference_pool_trace(name):
    """The alloc/free trace recorded on the reference's CPU FIFO (tests/golden/pool_trace.json), replayed on the
    device ring: same slots out, same order left."""
    if DROPIN_DIR not in sys.path:
        sys.path.insert(0, DROPIN_DIR)
    import memory_pool
    tr = POOL_TRACES[name]
    kv = memory_pool.MHATokenToKVPool(size=tr["size"], dtype=torch.bfloat16, head_num=1, head_dim=16, layer_num=1,
                                      device="cuda")
    for i, (kind, arg, want, avail) in enumerate(tr["ops"]):
        if kind == "alloc":
            got = kv.alloc(arg)
            assert (None if got is None else got.tolist()) == want
            assert got is None or (got.is_cuda and got.dtype == torch.int32)
            assert kv.available_size() == avail
        elif kind == "free":
            dt = torch.int64 if i % 3 == 0 else torch.int32        # radix-cache frees can be either width
            vals = torch.tensor(arg, dtype=dt)
            kv.free(vals.cuda() if i % 2 else vals)                 # device views and host tensors both land in order
            assert kv.available_size() == avail
    assert kv.free_slots.tolist() == tr["final_free_slots"]
    assert kv.free_slots.dtype == torch.int32 and not kv.free_slots.is_cuda


@pytest.mark.gpu
def test_device_slot_ring_grows_and_wraps():
    if DROPIN_DIR not in sys.path:
        sys.path.insert(0, DROPIN_DIR)
    import memory_pool
    kv = memory_pool.MHATokenToKVPool(size=10, dtype=torch.bfloat16, head_num=1, head_dim=16, layer_num=1, device="cuda")
    fifo = O.SlotFifo(10)
    a, b = kv.alloc(7), fifo.alloc(7)
    kv.free(a[:5]); fifo.free(b[:5])                 # wraps the 10-cell ring
    assert kv.alloc(6).tolist() == fifo.alloc(6)
    kv.free(torch.arange(1, 21, dtype=torch.int32, device="cuda")); fifo.free(range(1, 21))   # over capacity: the ring grows
    assert kv.available_size() == fifo.available_size() == 22
    assert kv.free_slots.tolist() == fifo.slots
    strided = torch.arange(100, 140, dtype=torch.int32, device="cuda")[::4]
    kv.free(strided); fifo.free(strided.tolist())
    assert kv.alloc(30).tolist() == fifo.alloc(30) and kv.free_slots.tolist() == fifo.slots
