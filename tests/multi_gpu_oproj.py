"""K8 under torchrun (one process per GPU, NCCL for the plumbing only): fused o_proj + all-reduce vs the pair it
replaces (`F.linear` on the rank's shard + `dist.all_reduce`, qwen2.py:195 + tp_runtime.py:28-33).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/multi_gpu_oproj.py [--bench]

Checks per case: (1) the result is bit-identical on every rank; (2) it is within bf16 rounding of the fp32 sum of
the ranks' exact partial products (the NCCL pair rounds every partial to bf16 too, and then adds in bf16);
(3) it is within tolerance of the NCCL pair.  Prints one JSON line per case on rank 0 and exits non-zero on failure.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from baby_sglang_b200 import DROPIN_DIR  # noqa: E402

sys.path.insert(0, DROPIN_DIR)
import tp_runtime  # noqa: E402


def run(mode, rank, world, group):
    """All cases under one exchange: "" = the communicator's default (peer stores), "nvls" = in-switch reduction."""
    bench = "--bench" in sys.argv
    os.environ["BSGL_K8_EXCHANGE"] = mode
    dev = torch.device("cuda", torch.cuda.current_device())
    # (tokens, hidden, full K = Hq * D): 32B decode C5, a ragged token count, 7B geometry, a tiny batch
    cases = [(1024, 5120, 5120), (333, 5120, 5120), (256, 3584, 3584), (7, 1536, 1536)]
    comm = tp_runtime.FusedOProjAllReduce(rank, world, group, hidden=5120, max_tokens=1024)
    ok = True
    for T, hidden, Kfull in cases:
        if Kfull % (world * 128) or hidden > comm.hidden:
            continue
        if hidden != comm.hidden:
            comm.close()
            comm = tp_runtime.FusedOProjAllReduce(rank, world, group, hidden=hidden, max_tokens=1024)
        K = Kfull // world
        g = torch.Generator(device="cpu").manual_seed(1000 * T + rank)
        a = torch.randn(T, K, generator=g).to(dev, torch.bfloat16)
        w = (torch.randn(hidden, K, generator=g) * 0.05).to(dev, torch.bfloat16)
        for rep in range(3):                      # consecutive calls exercise the epoch hand-over
            out = comm(a, w).clone()
        # the same call captured into a CUDA graph and replayed: the call counter lives in device memory, so every
        # replay is a new epoch on every rank (a host-side epoch would have been baked into the graph)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            comm(a, w)
        torch.cuda.current_stream().wait_stream(side)
        with torch.cuda.graph(graph):
            g_out = comm(a, w).clone()
        replay_same = True
        for rep in range(3):
            g_out.zero_()
            graph.replay()
            torch.cuda.synchronize()
            replay_same &= bool(torch.equal(g_out, out))
        del graph
        # the pair it replaces
        ref = torch.nn.functional.linear(a, w)
        dist.all_reduce(ref, group=group)
        # fp32 truth: sum over ranks of the exact partials
        exact = a.float() @ w.float().t()
        dist.all_reduce(exact, group=group)
        gathered = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(gathered, out, group=group)
        same = all(torch.equal(gathered[0], t) for t in gathered)
        err_exact = (out.float() - exact).abs().max().item()
        err_nccl = (out.float() - ref.float()).abs().max().item()
        scale = exact.abs().max().item()
        tol = 2.0 ** -7 * scale * (1 + world / 4)        # each partial is rounded to bf16 once, the sum once more
        good = same and replay_same and err_exact <= tol and err_nccl <= 2 * tol
        ok &= good
        line = {"case": f"T={T} hidden={hidden} K/rank={K}", "world": world, "bit_identical_across_ranks": same,
                "cuda_graph_replays_identical": replay_same, "exchange": comm.last_exchange,
                "max_abs_err_vs_fp32": err_exact, "max_abs_err_vs_nccl_pair": err_nccl, "tol": tol, "ok": good}
        if bench and T == 1024:
            def timed(fn, n=50):
                for _ in range(5):
                    fn()
                torch.cuda.synchronize()
                dist.barrier(group=group)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(n):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
                return float(t)

            def pair():
                o = torch.nn.functional.linear(a, w)
                dist.all_reduce(o, group=group)
                return o

            line["fused_us"] = round(timed(lambda: comm(a, w)) * 1e3, 1)
            line["nccl_pair_us"] = round(timed(pair) * 1e3, 1)
            line["gemm_only_us"] = round(timed(lambda: torch.nn.functional.linear(a, w)) * 1e3, 1)
            line["allreduce_bytes"] = T * hidden * 2
        if rank == 0:
            print(json.dumps(line), flush=True)
    comm.close()
    dist.barrier(group=group)
    return ok


def host_staged(rank, world, group):
    """HostStagedAttention with `after` = K8 (bench.py's TP e2e leg): every layer's downloaded result must equal the
    resident path's.  K8 has ONE output buffer, so layer l+1's call may only start once layer l's download has left
    it; the download (10 MiB per layer over PCIe) is far slower than a layer of this tiny attention, so a missing
    dependency shows up as layer l+1's tiles inside out_host[l]."""
    import attn_backend
    import forward_batch
    import memory_pool
    from baby_sglang_b200.host_staging import HostStagedAttention
    os.environ["BSGL_K8_EXCHANGE"] = ""
    dev = torch.device("cuda", torch.cuda.current_device())
    B, Hq, Hkv, D, L, hidden, ctx = 1024, 8, 2, 128, 6, 5120, 24
    dt = torch.bfloat16
    g = torch.Generator(device="cpu").manual_seed(77 + rank)
    S = B * ctx
    kv = memory_pool.MHATokenToKVPool(size=S, dtype=dt, head_num=Hkv, head_dim=D, layer_num=L, device=dev)
    for l in range(L):
        kv.k_buffer[l].copy_(torch.randn(S + 1, Hkv, D, generator=g).to(dt))
        kv.v_buffer[l].copy_(torch.randn(S + 1, Hkv, D, generator=g).to(dt))
    r2t = memory_pool.ReqToTokenPool(size=B, max_context_len=ctx + 4, device=dev)
    r2t.req_to_token[:, :ctx] = (torch.arange(S, dtype=torch.int32).view(B, ctx) + 1).to(dev)
    rpi = torch.arange(B, dtype=torch.int64, device=dev)
    seq = torch.full((B,), ctx, dtype=torch.int64, device=dev)
    loc = r2t.req_to_token[:, ctx - 1].contiguous()
    fb = forward_batch.SimplifiedForwardBatch.create_decode_batch(
        input_ids=torch.zeros(B, dtype=torch.int64, device=dev), req_pool_indices=rpi, seq_lens=seq,
        out_cache_loc=loc, req_to_token_pool=r2t, token_to_kv_pool=kv)
    layers = [type("Layer", (), dict(layer_id=l, head_dim=D, scaling=D ** -0.5))() for l in range(L)]
    w_o = [(torch.randn(hidden, Hq * D, generator=g) * 0.05).to(dev, dt) for _ in range(L)]
    qkv_host = [torch.randn(B, (Hq + 2 * Hkv) * D, generator=g).to(dt).pin_memory() for _ in range(L)]
    comm = tp_runtime.FusedOProjAllReduce(rank, world, group, hidden=hidden, max_tokens=B)
    staged = HostStagedAttention(Hq, Hkv, D, dev)
    # resident path: same backend call, K8 output cloned before the next call reuses the buffer
    want = []
    for l in range(L):
        q, k, v = staged._split(qkv_host[l].to(dev))
        want.append(comm(attn_backend.SimpleAttentionBackend.forward(q, k, v, layers[l], fb, True), w_o[l]).clone())
    torch.cuda.synchronize()
    calls = [0]

    def after(out):
        calls[0] += 1
        return comm(out, w_o[(calls[0] - 1) % L])

    good = True
    bad_layers = []
    for rep in range(3):
        out_host = [torch.zeros(B, hidden, dtype=dt).pin_memory() for _ in range(L)]
        staged.forward_layers(qkv_host, out_host, layers, fb, after=after)
        for l in range(L):
            if not torch.equal(out_host[l], want[l].cpu()):
                good = False
                bad_layers.append((rep, l))
    flag = torch.tensor([1 if good else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    if rank == 0:
        print(json.dumps({"case": f"host-staged attention + K8, {L} layers x 3 steps, T={B} hidden={hidden}",
                          "world": world, "every_layer_download_equals_resident": bool(flag.item()),
                          "bad_layers_rank0": bad_layers[:6], "ok": bool(flag.item())}), flush=True)
    comm.close()
    dist.barrier(group=group)
    return bool(flag.item())


def chunked(rank, world, group):
    """Batches larger than the communicator: `_fused_oproj_forward` cuts them into max_tokens-row blocks (one row
    over, a ragged tail, a 16 K-token prefill) instead of leaving K8 for F.linear + NCCL."""
    os.environ["BSGL_K8_EXCHANGE"] = ""
    dev = torch.device("cuda", torch.cuda.current_device())
    hidden, K, cap = 3584, 3584 // world // 128 * 128, 1024
    comm = tp_runtime.FusedOProjAllReduce(rank, world, group, hidden=hidden, max_tokens=cap)
    g = torch.Generator(device="cpu").manual_seed(4242 + rank)
    lin = torch.nn.Linear(K, hidden, bias=False, device=dev, dtype=torch.bfloat16)
    with torch.no_grad():
        lin.weight.copy_((torch.randn(hidden, K, generator=g) * 0.05).to(dev, torch.bfloat16))
    fwd = tp_runtime._fused_oproj_forward(lin, comm, group)
    ok = True
    for T in (cap, cap + 1, 2 * cap + 333, 16384):
        a = torch.randn(T, K, generator=g).to(dev, torch.bfloat16)
        calls0 = comm.calls
        with torch.no_grad():
            out = fwd(a.view(1, T, K))
            # the same rows through single calls of at most `cap` rows: blocks must reproduce them bit for bit
            want = torch.cat([comm(a[lo:lo + cap], lin.weight).clone() for lo in range(0, T, cap)])
            ref = torch.nn.functional.linear(a, lin.weight)
        n_calls = (comm.calls - calls0) // 2
        dist.all_reduce(ref, group=group)
        gathered = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(gathered, out.contiguous(), group=group)
        same = all(torch.equal(gathered[0], t) for t in gathered)
        exact = bool(torch.equal(out.view(T, hidden), want))
        err = (out.view(T, hidden).float() - ref.float()).abs().max().item()
        tol = 2.0 ** -6 * ref.float().abs().max().item() * (1 + world / 4)
        good = same and exact and tuple(out.shape) == (1, T, hidden) and n_calls == -(-T // cap) and err <= tol
        ok &= good
        if rank == 0:
            print(json.dumps({"case": f"row blocks: T={T} max_tokens={cap} hidden={hidden} K/rank={K}", "world": world,
                              "k8_calls": n_calls, "bit_identical_across_ranks": same,
                              "equal_to_blockwise_single_calls": exact, "max_abs_err_vs_nccl_pair": err, "tol": tol,
                              "ok": good}), flush=True)
    comm.close()
    dist.barrier(group=group)
    return ok


def main():
    rank, world, group = tp_runtime.init_tp()
    ok = run("", rank, world, group)
    ok &= host_staged(rank, world, group)
    ok &= chunked(rank, world, group)
    if world > 1:
        ok &= run("nvls", rank, world, group)     # falls back to peer stores where multicast is unavailable
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
