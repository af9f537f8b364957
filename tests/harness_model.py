"""TEST INFRASTRUCTURE ONLY — a small Qwen2-style decoder and a static-batch greedy loop that drive the
attention backend the way the reference's unchanged model and scheduler do.

Why it exists: the reference's `qwen2.py` / `scheduler.py` cannot travel to the GPU box (and must not be
copied), yet the north-star parity gate is "greedy token IDs identical for 64 steps".  This harness
restates, compactly and in its own words, what surrounds the backend call:

    embed -> [ RMSNorm(+residual) -> fused QKV linear (bias) -> neox RoPE -> BACKEND.forward -> o_proj
               -> RMSNorm(+residual) -> SwiGLU MLP ] x L -> RMSNorm -> lm_head          (qwen2.py:57-75,105-123,
                                                                                         179-197,206-222,249-268)
    prefill:  rows of the page table = [reused prefix slots | newly allocated slots]     (scheduler.py:423-452)
    decode:   one slot per request appended at req_to_token[r, seq_len]                  (scheduler.py:486-507)
    greedy:   argmax of the last-token logits                                            (sample.py:56-57)

Parameter names equal the reference's state-dict names, so `deterministic_weights()` can be applied to the
reference's own model too; `tests/golden/make_e2e_golden.py` did exactly that to produce
`tests/golden/e2e_tokens.json` with the reference's real `Scheduler.run_batch`, and
`tests/test_e2e_harness.py` checks this harness (fp32, oracle backend, CPU) against those IDs.  On the GPU the
same harness runs in bf16 with the CUDA drop-in backend and with the oracle backend, and the IDs must agree.

The attention layer object handed to the backend exposes what `BabyRadixAttention` does (qwen2.py:302-307) and the
module tree uses the reference's attribute names (`self_attn.qkv_proj.qkv_proj`, `o_proj`, `num_heads`, ...) because
`tp_runtime.shard_attention_heads` mutates exactly those attributes on the reference's unchanged model.
"""
from __future__ import annotations

import zlib
from dataclasses import dataclass
from typing import Callable, List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class HarnessConfig:
    vocab_size: int = 512
    hidden_size: int = 256
    num_attention_heads: int = 2          # head_dim = hidden_size // num_attention_heads (qwen2.py:159)
    num_key_value_heads: int = 1
    num_hidden_layers: int = 2
    intermediate_size: int = 512
    rms_norm_eps: float = 1e-6
    rope_theta: float = 1e6
    tie_word_embeddings: bool = False


def _rms(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    xf = x.float()
    return w * (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype)


class _Norm(nn.Module):
    def __init__(self, n, eps):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(n))
        self.eps = eps


class _Fused(nn.Module):
    """A linear layer wrapped under an attribute of the given name (the reference nests its fused projections:
    `qkv_proj.qkv_proj`, `gate_up_proj.gate_up_proj`)."""

    def __init__(self, name, n_in, n_out, bias):
        super().__init__()
        setattr(self, name, nn.Linear(n_in, n_out, bias=bias))


class _AttnLayer(nn.Module):
    """What the backend reads off the attention layer (qwen2.py:302-307)."""

    def __init__(self, n_heads, head_dim, n_kv_heads, layer_id):
        super().__init__()
        self.tp_q_head_num, self.tp_k_head_num, self.tp_v_head_num = n_heads, n_kv_heads, n_kv_heads
        self.head_dim, self.scaling, self.layer_id = head_dim, head_dim ** -0.5, layer_id


class _Rotary(nn.Module):
    """Same surface as the reference's RotaryEmbedding (qwen2.py:227-268): an `inv_freq` buffer built on the CPU
    at construction (and therefore cast along with the model, model_loader.py:77) and forward(positions, q, k)."""

    def __init__(self, head_dim: int, theta: float):
        super().__init__()
        inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2).float() / head_dim))
        self.register_buffer("inv_freq", inv, persistent=False)

    def forward(self, positions, q, k):
        return _rope(q, positions, self.inv_freq), _rope(k, positions, self.inv_freq)


class _SelfAttn(nn.Module):
    """Same attributes and forward(positions, hidden_states, forward_batch) as the reference's Qwen2Attention
    (qwen2.py:143-197); `backend` is the drop-in boundary (SimpleAttentionBackend.forward)."""
    backend: Callable = None

    def forward(self, positions, hidden_states, forward_batch):
        qkv = self.qkv_proj.qkv_proj(hidden_states)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        q = q.view(-1, self.num_heads, self.head_dim)
        k = k.view(-1, self.num_kv_heads, self.head_dim)
        v = v.view(-1, self.num_kv_heads, self.head_dim)        # strided view of the fused output
        q, k = self.rotary_emb(positions, q, k)
        return self.o_proj(self.backend(q, k, v, self.attn, forward_batch, True))

    def __init__(self, cfg: HarnessConfig, layer_id: int):
        super().__init__()
        self.hidden_size = cfg.hidden_size
        self.num_heads, self.num_kv_heads = cfg.num_attention_heads, cfg.num_key_value_heads
        self.head_dim = cfg.hidden_size // cfg.num_attention_heads
        self.q_size, self.kv_size = self.num_heads * self.head_dim, self.num_kv_heads * self.head_dim
        self.qkv_proj = _Fused("qkv_proj", cfg.hidden_size, self.q_size + 2 * self.kv_size, True)
        self.o_proj = nn.Linear(self.q_size, cfg.hidden_size, bias=False)
        self.attn = _AttnLayer(self.num_heads, self.head_dim, self.num_kv_heads, layer_id)
        self.rope_theta = cfg.rope_theta
        self.rotary_emb = _Rotary(self.head_dim, cfg.rope_theta)


class _Mlp(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.gate_up_proj = _Fused("gate_up_proj", cfg.hidden_size, 2 * cfg.intermediate_size, False)
        self.down_proj = nn.Linear(cfg.intermediate_size, cfg.hidden_size, bias=False)


class _Layer(nn.Module):
    def __init__(self, cfg, i):
        super().__init__()
        self.self_attn = _SelfAttn(cfg, i)
        self.mlp = _Mlp(cfg)
        self.input_layernorm = _Norm(cfg.hidden_size, cfg.rms_norm_eps)
        self.post_attention_layernorm = _Norm(cfg.hidden_size, cfg.rms_norm_eps)


class _Body(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        self.layers = nn.ModuleList([_Layer(cfg, i) for i in range(cfg.num_hidden_layers)])
        self.norm = _Norm(cfg.hidden_size, cfg.rms_norm_eps)


def _rope(x: torch.Tensor, positions: torch.Tensor, inv_freq: torch.Tensor) -> torch.Tensor:
    """neox-style rotation of [T, H, D]: halves (x1, x2) -> (x1 cos - x2 sin, x1 sin + x2 cos); cos/sin are
    computed in fp32 and cast to x's dtype before the multiply (qwen2.py:249-268)."""
    D = x.shape[-1]
    ang = torch.outer(positions.float(), inv_freq)
    c, s = torch.cos(ang)[:, None, :].to(x.dtype), torch.sin(ang)[:, None, :].to(x.dtype)
    a, b = x[..., : D // 2], x[..., D // 2:]
    return torch.cat((a * c - b * s, a * s + b * c), dim=-1)


class HarnessQwen2(nn.Module):
    """`backend(q, k, v, attn_layer, forward_batch, save_kv_cache)` is the drop-in boundary
    (`SimpleAttentionBackend.forward`, qwen2.py:315)."""

    def __init__(self, cfg: HarnessConfig, backend: Callable):
        super().__init__()
        self.config = cfg
        self.model = _Body(cfg)
        self.lm_head = None if cfg.tie_word_embeddings else nn.Linear(cfg.hidden_size, cfg.vocab_size, bias=False)
        self.backend = backend
        for layer in self.model.layers:
            layer.self_attn.__dict__["backend"] = backend      # plain attribute, not a submodule

    @torch.no_grad()
    def forward(self, input_ids, positions, forward_batch):
        eps = self.config.rms_norm_eps
        h = self.model.embed_tokens(input_ids)
        res = None
        for layer in self.model.layers:
            if res is None:
                res = h
            else:
                res = h + res
            x = _rms(res, layer.input_layernorm.weight, eps)
            h = layer.self_attn(positions, x, forward_batch)
            res = h + res
            x = _rms(res, layer.post_attention_layernorm.weight, eps)
            gate, up = layer.mlp.gate_up_proj.gate_up_proj(x).chunk(2, dim=-1)
            h = layer.mlp.down_proj(F.silu(gate) * up)
        h = _rms(h + res, self.model.norm.weight, eps)
        w = self.model.embed_tokens.weight if self.lm_head is None else self.lm_head.weight
        return F.linear(h, w)


def deterministic_weights(model: nn.Module, scale: float = 1.0) -> None:
    """Fill every parameter from a generator seeded by crc32(parameter name): the same recipe gives the same
    weights on the reference's model (same state-dict names) and on this harness, on any machine."""
    with torch.no_grad():
        for name, p in model.named_parameters():
            g = torch.Generator().manual_seed(zlib.crc32(name.encode()))
            r = torch.randn(p.shape, generator=g, dtype=torch.float32)
            if name.endswith("norm.weight") or "layernorm" in name:
                val = 1.0 + 0.1 * r
            elif name.endswith(".bias"):
                val = 0.1 * r
            elif "embed_tokens" in name:
                val = r
            else:
                val = r * (scale / (p.shape[-1] ** 0.5))
            p.copy_(val.to(p.dtype))


# ----------------------------------------------------------------------------------------------------
# static-batch greedy loop over the pools / batch factories of a given module set
# ----------------------------------------------------------------------------------------------------
@dataclass
class Stack:
    """The three shadowed modules' public classes plus the backend entry point."""
    ReqToTokenPool: type
    MHATokenToKVPool: type
    SimplifiedForwardBatch: type
    backend_forward: Callable


@dataclass
class GreedyTrace:
    ids: List[List[int]]          # tokens fed forward (argmax, or the forced ones)
    pred: List[List[int]]         # this stack's own argmax at every step
    top1: List[List[float]]       # top-1 logit value per step
    margin: List[List[float]]     # top-1 minus top-2 logit per step
    pools: tuple


def greedy_generate(model: HarnessQwen2, stack: Stack, prompts: List[List[int]], max_new_tokens: int, device,
                    dtype, reuse_prefix: Optional[List[int]] = None, pools=None, max_total_tokens: int = 0,
                    forced_ids: Optional[List[List[int]]] = None) -> GreedyTrace:
    """Prefill all prompts in one extend batch, then `max_new_tokens - 1` decode steps.  With `reuse_prefix`
    (tokens per request) and the `pools` of an earlier call on the same prompts, the first `reuse_prefix[i]`
    page-table cells of request i are kept (a radix-cache hit, scheduler.py:423-442) and only the rest of the
    prompt is extended.  `forced_ids` feeds the given tokens forward instead of this stack's own argmax
    (teacher forcing), while `pred` still records the argmax."""
    cfg = model.config
    B = len(prompts)
    head_dim = cfg.hidden_size // cfg.num_attention_heads
    kv_heads = model.model.layers[0].self_attn.num_kv_heads        # per-rank under TP
    lens = [len(p) for p in prompts]
    C = max(lens) + max_new_tokens + 8
    total = max_total_tokens or (sum(lens) + B * max_new_tokens + 16) * 2
    if pools is None:
        r2t = stack.ReqToTokenPool(size=B, max_context_len=C, device=str(device))
        kvp = stack.MHATokenToKVPool(size=total, dtype=dtype, head_num=kv_heads, head_dim=head_dim,
                                     layer_num=cfg.num_hidden_layers, device=str(device))
        rows = r2t.alloc(B)
        prefix = [0] * B
    else:
        r2t, kvp, rows = pools
        prefix = list(reuse_prefix)
    ext = [n - p for n, p in zip(lens, prefix)]
    new_slots = kvp.alloc(sum(ext))
    assert new_slots is not None
    off = 0
    for i in range(B):
        r2t.req_to_token[rows[i], prefix[i]:lens[i]] = new_slots[off:off + ext[i]].to(torch.int32)
        off += ext[i]
    ids = torch.tensor([t for p, pre in zip(prompts, prefix) for t in p[pre:]], dtype=torch.int64, device=device)
    rpi = torch.tensor(rows, dtype=torch.int64, device=device)
    seq = torch.tensor(lens, dtype=torch.int64, device=device)
    fb = stack.SimplifiedForwardBatch.create_prefill_batch(
        input_ids=ids, req_pool_indices=rpi, seq_lens=seq, out_cache_loc=new_slots.to(torch.int32),
        req_to_token_pool=r2t, token_to_kv_pool=kvp,
        prefix_lens=torch.tensor(prefix, dtype=torch.int64, device=device))
    logits = model(fb.input_ids, fb.positions, fb)
    last = torch.cumsum(torch.tensor(ext, device=device), 0) - 1
    logits = logits[last].float()
    tr = GreedyTrace([[] for _ in range(B)], [[] for _ in range(B)], [[] for _ in range(B)],
                     [[] for _ in range(B)], (r2t, kvp, rows))

    def record(lg, step):
        top2 = lg.topk(2, dim=-1)
        pred = top2.indices[:, 0]
        nxt = pred if forced_ids is None else torch.tensor([f[step] for f in forced_ids], device=device)
        for i in range(B):
            tr.pred[i].append(int(pred[i]))
            tr.ids[i].append(int(nxt[i]))
            tr.top1[i].append(float(top2.values[i, 0]))
            tr.margin[i].append(float(top2.values[i, 0] - top2.values[i, 1]))
        return nxt

    nxt = record(logits, 0)
    for step in range(1, max_new_tokens):
        slots = kvp.alloc(B)
        r2t.req_to_token[rpi, seq] = slots.to(torch.int32)
        seq = seq + 1
        fb = stack.SimplifiedForwardBatch.create_decode_batch(
            input_ids=nxt, req_pool_indices=rpi, seq_lens=seq, out_cache_loc=slots.to(torch.int32),
            req_to_token_pool=r2t, token_to_kv_pool=kvp)
        nxt = record(model(fb.input_ids, fb.positions, fb).float(), step)
    return tr
