"""`tp_runtime` drop-in on CPU: world_size-2 gloo processes (the N>1 path without GPUs).

  * the four reference helpers behave like tp_runtime.py:10-51 (in-place SUM, gather + cat(dim=0), broadcast);
  * KV-head tensor parallelism wired into a model by `shard_attention_heads` (weight slicing + o_proj all-reduce
    hook) reproduces the unsharded model's greedy token IDs and logits, with per-rank KV pools of Hkv/t heads;
  * request-batch data parallelism: partition + gather restores the request order.
"""
import os
import socket
import sys
import traceback

import pytest
import torch
import torch.multiprocessing as mp

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, fn_name, q):
    try:
        sys.path.insert(0, REPO)
        os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                          MASTER_PORT=str(port))
        torch.set_num_threads(2)
        from baby_sglang_b200 import DROPIN_DIR
        sys.path.insert(0, DROPIN_DIR)
        import tp_runtime
        r, w, group = tp_runtime.init_tp()
        assert (r, w) == (rank, world)
        globals()[fn_name](tp_runtime, r, w, group)
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
        q.put((rank, "ok"))
    except Exception:
        q.put((rank, traceback.format_exc()))


def _spawn(fn_name, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, fn_name, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank} failed:\n{msg}"


# ---------------------------------------------------------------------------------------- rank bodies
def _body_collectives(tp, rank, world, group):
    x = torch.full((3, 4), float(rank + 1))
    y = tp.tp_all_reduce_sum(x, group)
    assert y is x and torch.equal(x, torch.full((3, 4), float(sum(range(1, world + 1)))))
    g = tp.tp_all_gather(torch.full((2, 3), float(rank)), group)
    assert g.shape == (2 * world, 3)
    for r in range(world):
        assert torch.equal(g[2 * r:2 * r + 2], torch.full((2, 3), float(r)))
    b = torch.arange(5.0) if rank == 1 else torch.zeros(5)
    assert torch.equal(tp.tp_broadcast(b, 1, group), torch.arange(5.0))
    # request-batch DP: round-robin partition, gather restores order
    n = 7
    mine = tp.dp_partition(n, rank, world)
    outs = tp.dp_gather_outputs([[i, i * i] for i in mine], n, rank, world, group)
    assert outs == [[i, i * i] for i in range(n)]


def _body_kv_head_tp(tp, rank, world, group):
    from tests.harness_model import HarnessConfig, HarnessQwen2, deterministic_weights, greedy_generate
    from tests.oracle_stack import oracle_stack
    cfg = HarnessConfig(vocab_size=300, hidden_size=256, num_attention_heads=4, num_key_value_heads=2,
                        num_hidden_layers=2, intermediate_size=256)
    st = oracle_stack()
    g = torch.Generator().manual_seed(5)
    prompts = [torch.randint(0, cfg.vocab_size, (n,), generator=g).tolist() for n in (21, 4, 40)]
    full = HarnessQwen2(cfg, st.backend_forward).float().eval()
    deterministic_weights(full)
    want = greedy_generate(full, st, prompts, 12, torch.device("cpu"), torch.float32).ids

    sharded = HarnessQwen2(cfg, st.backend_forward).float().eval()
    deterministic_weights(sharded)
    plan = tp.shard_attention_heads(sharded, rank, world, group)
    assert plan.mode == "kv_head_tp" and plan.kv_heads_per_rank == 1 and plan.q_heads_per_rank == 2
    at = sharded.model.layers[0].self_attn
    assert (at.num_heads, at.num_kv_heads, at.q_size, at.kv_size) == (2, 1, 128, 64)
    assert at.qkv_proj.qkv_proj.weight.shape == (256, 256) and at.o_proj.weight.shape == (256, 128)
    assert at.attn.tp_q_head_num == 2 and at.attn.tp_k_head_num == 1
    pool_cfg = tp.kv_pool_config(cfg, plan)
    assert pool_cfg.num_key_value_heads == 1 and cfg.num_key_value_heads == 2
    tr = greedy_generate(sharded, st, prompts, 12, torch.device("cpu"), torch.float32)
    assert tr.pools[1].head_num == 1         # per-rank KV pool holds Hkv / t heads
    assert tr.ids == want, f"rank {rank}: TP token ids differ"
    # ... and with the MLP sharded too (gate/up rows + down_proj columns, all-reduce after down_proj)
    assert tp.shard_mlp(sharded, rank, world, group) == cfg.intermediate_size // world
    mlp = sharded.model.layers[0].mlp
    assert mlp.gate_up_proj.gate_up_proj.weight.shape == (cfg.intermediate_size, cfg.hidden_size)   # 2 * I / 2 rows
    assert mlp.down_proj.weight.shape == (cfg.hidden_size, cfg.intermediate_size // world)
    ref_mlp = full.model.layers[0].mlp
    lo = rank * (cfg.intermediate_size // world)
    assert torch.equal(mlp.gate_up_proj.gate_up_proj.weight[0], ref_mlp.gate_up_proj.gate_up_proj.weight[lo])
    assert torch.equal(mlp.gate_up_proj.gate_up_proj.weight[cfg.intermediate_size // world],
                       ref_mlp.gate_up_proj.gate_up_proj.weight[cfg.intermediate_size + lo])
    tr2 = greedy_generate(sharded, st, prompts, 12, torch.device("cpu"), torch.float32)
    assert tr2.ids == want, f"rank {rank}: TP (attention + MLP) token ids differ"
    assert tp.shard_mlp(full, rank, 1, None) == 0


def _body_dp_plan(tp, rank, world, group):
    assert tp.plan_parallelism(12, 2, 2).mode == "kv_head_tp"
    assert tp.plan_parallelism(12, 2, 4).mode == "request_dp"       # Hkv < world
    assert tp.plan_parallelism(28, 4, 8).mode == "request_dp"
    p = tp.plan_parallelism(40, 8, 8)
    assert (p.mode, p.kv_heads_per_rank, p.q_heads_per_rank) == ("kv_head_tp", 1, 5)
    assert tp.plan_parallelism(40, 8, 1).mode == "single"


# ---------------------------------------------------------------------------------------- tests
def test_collectives_world2_gloo():
    _spawn("_body_collectives")


def test_kv_head_tp_matches_unsharded_world2_gloo():
    _spawn("_body_kv_head_tp")


def test_plans_world2_gloo():
    _spawn("_body_dp_plan")


def test_single_process_short_circuits():
    sys.path.insert(0, os.path.join(REPO, "baby_sglang_b200", "dropin"))
    import tp_runtime
    x = torch.ones(2, 2)
    assert tp_runtime.tp_all_reduce_sum(x, None) is x
    assert tp_runtime.tp_all_gather(x, None) is x
    assert tp_runtime.tp_broadcast(x, 0, None) is x
    assert tp_runtime.dp_partition(5, 0, 1) == [0, 1, 2, 3, 4]
