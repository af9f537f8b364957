"""Overlap mode (reference overlap_worker.py:46-160): the model forward of step t runs on a daemon thread under a
side stream while the main thread prepares step t+1 on its own stream -- it appends the next page-table cell
(scheduler.py:486-488), takes slots from the pool and builds the next forward batch (our K4 runs there).  The
reference's worker cannot run on the GPU box (the reference does not travel), so this test re-creates its threading and
stream pattern around the drop-in modules and checks every step's attention output bit for bit against the same steps
executed one after the other on one thread.  What it pins: the library launches on the caller's current stream, keeps
no process-global device scratch, drops the GIL during calls and reads only `req_to_token[r, :seq_len]`."""
import queue
import sys
import threading

import pytest
import torch

from baby_sglang_b200 import DROPIN_DIR

pytestmark = pytest.mark.gpu


class _Layer:
    def __init__(self, layer_id, Hq, Hkv, D):
        self.layer_id, self.scaling, self.head_dim = layer_id, D ** -0.5, D
        self.tp_q_head_num, self.tp_k_head_num, self.tp_v_head_num = Hq, Hkv, Hkv


def _modules():
    if DROPIN_DIR not in sys.path:
        sys.path.insert(0, DROPIN_DIR)
    import attn_backend
    import forward_batch
    import memory_pool
    return attn_backend, forward_batch, memory_pool


def _run(overlap: bool, steps: int = 6):
    attn_backend, forward_batch, memory_pool = _modules()
    dev = torch.device("cuda")
    B, Hq, Hkv, D, L = 48, 28, 4, 128, 3
    g = torch.Generator().manual_seed(3)
    seq0 = torch.randint(200, 1200, (B,), generator=g)
    C = int(seq0.max()) + steps + 8
    S = int(seq0.sum()) + B * (steps + 2)
    kv_pool = memory_pool.MHATokenToKVPool(size=S, dtype=torch.bfloat16, head_num=Hkv, head_dim=D, layer_num=L,
                                           device="cuda")
    for l in range(L):
        kv_pool.k_buffer[l].copy_(torch.randn(kv_pool.k_buffer[l].shape, generator=g).to(torch.bfloat16))
        kv_pool.v_buffer[l].copy_(torch.randn(kv_pool.v_buffer[l].shape, generator=g).to(torch.bfloat16))
    r2t_pool = memory_pool.ReqToTokenPool(size=B, max_context_len=C, device="cuda")
    rows = r2t_pool.alloc(B)
    rpi = torch.tensor(rows, dtype=torch.int64, device=dev)
    for i, n in enumerate(seq0.tolist()):                      # prefill history: slots in pool FIFO order
        r2t_pool.req_to_token[rows[i], :n] = kv_pool.alloc(n)
    qkv = [[torch.randn(B, (Hq + 2 * Hkv) * D, generator=g).to(torch.bfloat16).to(dev) for _ in range(L)]
           for _ in range(steps)]
    layers = [_Layer(l, Hq, Hkv, D) for l in range(L)]
    results = [None] * steps
    torch.cuda.synchronize()

    @torch.no_grad()
    def forward(t, fb):
        outs = []
        for l in range(L):
            q, k, v = qkv[t][l].split([Hq * D, Hkv * D, Hkv * D], dim=-1)
            outs.append(attn_backend.SimpleAttentionBackend.forward(
                q.contiguous().view(B, Hq, D), k.contiguous().view(B, Hkv, D), v.view(B, Hkv, D), layers[l], fb, True))
        results[t] = torch.stack(outs)

    def prepare(t):
        """The scheduler's decode preparation (scheduler.py:476-507) on the main thread's stream."""
        seq = (seq0 + t + 1).to(dev)                           # seq_lens include the token being decoded (:489-490)
        loc = kv_pool.alloc(B)
        r2t_pool.req_to_token[rpi, seq - 1] = loc              # the cell the NEXT forward reads, beyond the current one
        return forward_batch.SimplifiedForwardBatch.create_decode_batch(
            input_ids=torch.zeros(B, dtype=torch.int64, device=dev), req_pool_indices=rpi, seq_lens=seq,
            out_cache_loc=loc, req_to_token_pool=r2t_pool, token_to_kv_pool=kv_pool)

    if not overlap:
        for t in range(steps):
            forward(t, prepare(t))
        torch.cuda.synchronize()
        return results, r2t_pool.req_to_token.clone()

    work: "queue.Queue" = queue.Queue()
    side = torch.cuda.Stream()
    errors = []

    def worker():                                              # overlap_worker.py:104-106: thread + side stream
        try:
            with torch.cuda.stream(side):
                while True:
                    item = work.get()
                    if item is None:
                        return
                    forward(*item)
                    # the reference's worker samples with one `.item()` per request here (overlap_worker.py:129-141),
                    # i.e. it drains the forward stream before it lets go of the batch
                    float(results[item[0]][0, 0, 0].item())
        except Exception as e:                                 # surfaced by the main thread below
            errors.append(e)

    th = threading.Thread(target=worker, daemon=True)
    th.start()
    for t in range(steps):
        fb = prepare(t)                                        # runs while the worker is still inside step t-1
        torch.cuda.current_stream().synchronize()              # overlap_worker.py:83-84
        work.put((t, fb))
    work.put(None)
    th.join(timeout=120)
    assert not th.is_alive() and not errors, errors
    side.synchronize()
    torch.cuda.synchronize()
    return results, r2t_pool.req_to_token.clone()


def test_overlap_mode_matches_sequential_steps():
    seq_out, seq_table = _run(overlap=False)
    ovl_out, ovl_table = _run(overlap=True)
    assert torch.equal(seq_table, ovl_table), "page table differs between modes"
    for t, (a, b) in enumerate(zip(seq_out, ovl_out)):
        assert a is not None and b is not None
        assert torch.equal(a, b), f"step {t}: overlap-mode attention output differs from the sequential run"
