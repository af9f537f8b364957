"""GPU parity tests proper: the CUDA path (through the C ABI) against the golden fixtures the
reference produced and against the oracle on seeded inputs.

Tolerances (BASELINE.json north_star): bf16/fp16 outputs max-abs-err <= 2e-2 versus the fp32
reference; KV-slot / index work bit-exact.
"""
import sys

import pytest
import torch

from baby_sglang_b200 import _lib
from oracle import paged_attention_oracle as O
from tests.helpers import golden_names, load_golden, make_paged_case, slab_checksum

pytestmark = pytest.mark.gpu

TOL = 2e-2          # bf16 output vs fp32 oracle (north_star)
SPLIT_TOL = 1e-3    # same inputs, different split counts -- on fp32-normalised outputs


@pytest.fixture(scope="module")
def ops():
    from baby_sglang_b200 import ops as _ops
    return _ops


def _dev(t):
    return None if t is None else t.cuda()


def _run_backend(ops, c, dtype=torch.bfloat16, n_ctas=None, store=True, extend_mode="tcgen05"):
    """kv_store + attention through ops on the case `c` (CPU tensors).  Returns (out, k_slab, v_slab, md).
    Extend cases run on K3 (`extend_mode="tcgen05"`) or row-wise on K2 (`"rowwise"`)."""
    k_slab = c.k_slab.to(dtype).cuda().contiguous()
    v_slab = c.v_slab.to(dtype).cuda().contiguous()
    r2t = c.req_to_token.cuda()
    qkv = c.qkv.to(dtype).cuda()
    T = qkv.shape[0]
    q, k, v = qkv.split([c.Hq * c.D, c.Hkv * c.D, c.Hkv * c.D], dim=-1)
    q = q.contiguous().view(T, c.Hq, c.D)
    k = k.contiguous().view(T, c.Hkv, c.D)
    v = v.view(T, c.Hkv, c.D)                    # strided view
    if store:
        ops.kv_store(k_slab, v_slab, c.out_cache_loc.cuda(), k, v)
    if c.is_prefill:
        md = ops.build_extend_metadata(r2t, c.req_pool_indices.cuda(), c.seq_lens.cuda(), c.prefix_lens.cuda(),
                                       T, c.Hkv, n_ctas=n_ctas, row_plan=(extend_mode == "rowwise"))
        if extend_mode == "rowwise":
            out = ops.paged_attention(q, k_slab, v_slab, r2t, md.plan, c.D ** -0.5)
        else:
            out = ops.extend_attention_tc(q, k_slab, v_slab, r2t, c.req_pool_indices.cuda(), c.seq_lens.cuda(),
                                          c.prefix_lens.cuda(), md.qo_indptr, c.D ** -0.5, req_order=md.heavy_order)
            # the dispatch order is bit-exact (stable, longest KV first) and never changes the result
            want = torch.sort(c.seq_lens, descending=True, stable=True).indices.to(torch.int32)
            assert torch.equal(md.heavy_order.cpu(), want)
            plain = ops.extend_attention_tc(q, k_slab, v_slab, r2t, c.req_pool_indices.cuda(), c.seq_lens.cuda(),
                                            c.prefix_lens.cuda(), md.qo_indptr, c.D ** -0.5)
            assert torch.equal(out, plain)
            # neither does the dealing of the work items to the persistent CTAs: greedy draws from the metadata's
            # device counter (twice: every launch must hand the counter back as zeros) vs the static dealing above
            for _ in range(2):
                drawn = ops.extend_attention_tc(q, k_slab, v_slab, r2t, c.req_pool_indices.cuda(), c.seq_lens.cuda(),
                                                c.prefix_lens.cuda(), md.qo_indptr, c.D ** -0.5,
                                                req_order=md.heavy_order, work_counter=md.work_counter)
                assert torch.equal(out, drawn)
                assert int(md.work_counter.abs().sum()) == 0
    else:
        md = ops.build_decode_metadata(r2t, c.req_pool_indices.cuda(), c.seq_lens.cuda(), c.Hkv, n_ctas=n_ctas)
        out = ops.paged_attention(q, k_slab, v_slab, r2t, md.plan, c.D ** -0.5)
    torch.cuda.synchronize()
    return out, k_slab, v_slab, md


def _golden_as_case(g):
    g.k_slab, g.v_slab = g.k_slab0, g.v_slab0
    return g


@pytest.mark.parametrize("extend_mode", ["tcgen05", "rowwise"])
@pytest.mark.parametrize("name", golden_names())
def test_golden_fixture(ops, name, extend_mode):
    g = _golden_as_case(load_golden(name))
    if not g.is_prefill and extend_mode == "rowwise":
        pytest.skip("decode fixtures have one path")
    out, k_slab, v_slab, md = _run_backend(ops, g, extend_mode=extend_mode)
    # K1: bit-exact against the rows the reference wrote, nothing else touched
    loc = g.out_cache_loc.long()
    assert torch.equal(k_slab.cpu()[loc].view(torch.int16), g.k_rows1.view(torch.int16))
    assert torch.equal(v_slab.cpu()[loc].view(torch.int16), g.v_rows1.view(torch.int16))
    assert slab_checksum(k_slab) == g.k_slab1_sum
    assert slab_checksum(v_slab) == g.v_slab1_sum
    # K4: positions / extended_lens bit-exact with the reference's factories
    assert torch.equal(md.positions.cpu(), g.positions)
    if g.is_prefill:
        assert torch.equal(md.extended_lens.cpu(), g.extended_lens)
    # attention: bf16 out vs the reference's fp32 out
    assert out.shape == g.out.shape and out.dtype == torch.bfloat16
    err = (out.float().cpu() - g.out).abs().max().item()
    assert err <= TOL, err


DECODE_CASES = [
    # (Hq, Hkv, D, seq_lens)
    (12, 2, 128, [1, 2, 15, 16, 17, 31, 32, 33, 255, 256, 257]),
    (28, 4, 128, [4096, 3, 777, 2048]),
    (40, 8, 128, [513, 1200, 64]),
    (14, 2, 64, [100, 1000, 5]),
    (5, 1, 128, [300, 17]),              # one KV head (32B under TP8)
    (8, 8, 128, [90, 91]),               # MHA
    (16, 2, 128, [1025, 130]),           # G = 8
]


@pytest.mark.parametrize("Hq,Hkv,D,seq_lens", DECODE_CASES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_decode_vs_oracle(ops, Hq, Hkv, D, seq_lens, dtype):
    c = make_paged_case(B=len(seq_lens), Hq=Hq, Hkv=Hkv, D=D, seq_lens=seq_lens, seed=len(seq_lens) * 7 + Hq,
                        dtype=dtype)
    out, k_slab, v_slab, _ = _run_backend(ops, c, dtype=dtype)
    ks, vs = c.k_slab.float().clone(), c.v_slab.float().clone()
    ref = O.backend_forward(c.q.float(), c.k.float(), c.v.float(), ks, vs, c.req_to_token, c.req_pool_indices,
                            c.seq_lens, c.out_cache_loc, False)
    assert torch.equal(k_slab.cpu(), ks.to(dtype)) and torch.equal(v_slab.cpu(), vs.to(dtype))
    err = (out.float().cpu() - ref).abs().max().item()
    assert err <= TOL, err


EXTEND_CASES = [
    # (Hq, Hkv, D, seq_lens, prefix_lens, alias_prefix_of)
    (12, 2, 128, [40, 9, 130], [0, 0, 0], None),
    (28, 4, 128, [300, 64, 33, 20], [256, 0, 33, 19], None),
    (14, 2, 64, [77, 200], [70, 100], None),
    (40, 8, 128, [96, 100, 70], [64, 64, 40], 0),        # aliased (shared) prefix slots
    (28, 4, 128, [1500, 700, 129], [1000, 0, 128], None),  # many KV tiles, several q chunks
    (12, 2, 128, [513, 600], [1, 0], None),              # G = 6, 15 q chunks
    (8, 8, 128, [300, 10], [0, 9], None),                # MHA: one q head per KV head
    (16, 2, 128, [260, 255], [200, 0], None),            # G = 8
    (4, 2, 64, [700], [100], None),                      # G = 2, D = 64
    (5, 1, 128, [400, 3], [399, 0], None),               # one-token extend after a long prefix
]


@pytest.mark.parametrize("extend_mode", ["tcgen05", "rowwise"])
@pytest.mark.parametrize("Hq,Hkv,D,seq_lens,prefix_lens,alias", EXTEND_CASES)
def test_extend_vs_oracle(ops, Hq, Hkv, D, seq_lens, prefix_lens, alias, extend_mode):
    c = make_paged_case(B=len(seq_lens), Hq=Hq, Hkv=Hkv, D=D, seq_lens=seq_lens, prefix_lens=prefix_lens,
                        seed=Hq + D, alias_prefix_of=alias)
    out, k_slab, v_slab, md = _run_backend(ops, c, extend_mode=extend_mode)
    ks, vs = c.k_slab.float().clone(), c.v_slab.float().clone()
    ref = O.backend_forward(c.q.float(), c.k.float(), c.v.float(), ks, vs, c.req_to_token, c.req_pool_indices,
                            c.seq_lens, c.out_cache_loc, True, c.prefix_lens, c.extended_lens)
    assert torch.equal(k_slab.cpu(), ks.to(torch.bfloat16))
    assert out.shape == ref.shape
    err = (out.float().cpu() - ref).abs().max().item()
    assert err <= TOL, err
    omd = O.build_metadata(c.req_to_token, c.req_pool_indices, c.seq_lens, c.prefix_lens)
    assert torch.equal(md.positions.cpu(), omd.positions)
    assert torch.equal(md.extended_lens.cpu(), omd.extended_lens)
    assert torch.equal(md.qo_indptr.cpu(), omd.qo_indptr)
    assert torch.equal(md.kv_indptr.cpu(), omd.kv_indptr)


def test_extend_fp16_and_modes_agree(ops):
    """fp16 through K3, and K3 against the row-wise K2 path on the same inputs."""
    c = make_paged_case(B=3, Hq=28, Hkv=4, D=128, seq_lens=[640, 200, 77], prefix_lens=[512, 0, 70], seed=21,
                        dtype=torch.float16)
    a, *_ = _run_backend(ops, c, dtype=torch.float16, extend_mode="tcgen05")
    b, *_ = _run_backend(ops, c, dtype=torch.float16, extend_mode="rowwise")
    ks, vs = c.k_slab.float().clone(), c.v_slab.float().clone()
    ref = O.backend_forward(c.q.float(), c.k.float(), c.v.float(), ks, vs, c.req_to_token, c.req_pool_indices,
                            c.seq_lens, c.out_cache_loc, True, c.prefix_lens, c.extended_lens)
    assert (a.float().cpu() - ref).abs().max().item() <= TOL
    assert (a.float() - b.float()).abs().max().item() <= 2 ** -6


def test_extend_many_requests(ops):
    """More requests than one scan round of the work-item search (B > 32), ragged, some empty extends."""
    g = torch.Generator().manual_seed(77)
    B = 70
    pre = torch.randint(0, 300, (B,), generator=g).tolist()
    ext = torch.randint(0, 90, (B,), generator=g).tolist()
    ext[3] = 0
    ext[40] = 0
    seq = [p + e for p, e in zip(pre, ext)]
    seq = [max(s, 1) for s in seq]
    pre = [min(p, s) for p, s in zip(pre, seq)]
    c = make_paged_case(B=B, Hq=12, Hkv=2, D=128, seq_lens=seq, prefix_lens=pre, seed=9)
    out, *_ = _run_backend(ops, c, extend_mode="tcgen05")
    ks, vs = c.k_slab.float().clone(), c.v_slab.float().clone()
    ref = O.backend_forward(c.q.float(), c.k.float(), c.v.float(), ks, vs, c.req_to_token, c.req_pool_indices,
                            c.seq_lens, c.out_cache_loc, True, c.prefix_lens, c.extended_lens)
    assert out.shape == ref.shape
    assert (out.float().cpu() - ref).abs().max().item() <= TOL


def test_split_count_invariance(ops):
    """Same batch under 1, 2, 5, 37 and 148 CTAs: the LSE merge must not change the answer."""
    c = make_paged_case(B=3, Hq=28, Hkv=4, D=128, seq_lens=[3000, 500, 1234], seed=5)
    outs = []
    for n in (1, 2, 5, 37, 148):
        out, *_ = _run_backend(ops, c, n_ctas=n)
        outs.append(out.float().cpu())
    ref = O.decode_attention(c.q.float(), *_stored(c), c.req_to_token, c.req_pool_indices, c.seq_lens)
    for o in outs:
        assert (o - ref).abs().max().item() <= TOL
    for o in outs[1:]:
        # outputs are rounded to bf16 (ulp 2^-8 relative), so compare at bf16 resolution of |x|<=4
        assert (o - outs[0]).abs().max().item() <= 2 ** -6


@pytest.mark.parametrize("Hq,Hkv,D", [(5, 1, 128), (4, 2, 64), (8, 1, 64)])
def test_narrow_rows_many_items_per_cta(ops, Hq, Hkv, D):
    """Rows of <= 256 B take the warp-per-tile producers (K2, Cfg::WPT).  200 ragged sequences on 1, 2 and 148 CTAs:
    with 1-2 CTAs a CTA walks more than kMaxItems (64) sequences, so the tile counter carries across passes and the
    tile -> warp assignment starts at every residue."""
    g = torch.Generator().manual_seed(Hq * 100 + D)
    seq = torch.randint(1, 200, (200,), generator=g).tolist()
    seq[7], seq[100] = 1500, 16
    c = make_paged_case(B=200, Hq=Hq, Hkv=Hkv, D=D, seq_lens=seq, seed=11)
    ref = O.decode_attention(c.q.float(), *_stored(c), c.req_to_token, c.req_pool_indices, c.seq_lens)
    for n in (1, 2, 148):
        out, *_ = _run_backend(ops, c, n_ctas=n)
        assert (out.float().cpu() - ref).abs().max().item() <= TOL, n


@pytest.mark.parametrize("n_ctas", [1, 3, 148])
def test_rows_without_kv_tokens_are_zero(ops, n_ctas):
    """A row whose KV length is 0 owns no tile, so no attention CTA writes it: the merge launch zero-fills it
    (K2c's zero-fill CTAs) and leaves every other row alone.  (The reference never decodes an empty request; the
    row-wise extend plan and padded batches can produce such rows.)"""
    c = make_paged_case(B=6, Hq=12, Hkv=2, D=128, seq_lens=[300, 1, 77, 513, 40, 9], seed=4)
    ref = O.decode_attention(c.q.float(), *_stored(c), c.req_to_token, c.req_pool_indices, c.seq_lens)
    k_slab, v_slab = (t.to(torch.bfloat16).cuda() for t in _stored(c))
    seq = c.seq_lens.clone()
    seq[1] = 0
    seq[4] = 0
    md = ops.build_decode_metadata(c.req_to_token.cuda(), c.req_pool_indices.cuda(), seq.cuda(), c.Hkv, n_ctas=n_ctas)
    out = torch.full((6, 12 * 128), 7.0, dtype=torch.bfloat16, device="cuda")
    ops.paged_attention(c.q.to(torch.bfloat16).cuda(), k_slab, v_slab, c.req_to_token.cuda(), md.plan, 128 ** -0.5,
                        out=out)
    torch.cuda.synchronize()
    out = out.float().cpu()
    assert torch.count_nonzero(out[1]) == 0 and torch.count_nonzero(out[4]) == 0
    keep = [0, 2, 3, 5]
    assert (out[keep] - ref[keep]).abs().max().item() <= TOL


def test_decode_step_is_cuda_graph_capturable(ops):
    """K4 + per layer (K1, K2, K2c) make no host sync and take all sizes from device buffers, so a whole decode step
    -- programmatic dependent launches included -- captures into one CUDA graph.  Replaying it after the batch moved
    on (new seq_lens / slots / q, k, v written into the same buffers) must give what eager launches give."""
    L = 2
    c = make_paged_case(B=5, Hq=28, Hkv=4, D=128, seq_lens=[700, 33, 1500, 16, 255], seed=8)
    dev = torch.device("cuda")
    r2t = c.req_to_token.cuda()
    rpi = c.req_pool_indices.cuda()
    seq = (c.seq_lens - 1).cuda()                        # static buffers, updated in place between replays
    loc = c.out_cache_loc.cuda().clone()
    slabs = [(c.k_slab.to(torch.bfloat16).cuda().clone(), c.v_slab.to(torch.bfloat16).cuda().clone()) for _ in range(L)]
    T = c.qkv.shape[0]
    qkv = c.qkv.to(torch.bfloat16).cuda().clone()
    outs = [torch.empty(T, 28 * 128, dtype=torch.bfloat16, device=dev) for _ in range(L)]

    def step():
        q, k, v = qkv.split([28 * 128, 4 * 128, 4 * 128], dim=-1)
        q = q.contiguous().view(T, 28, 128)
        k = k.contiguous().view(T, 4, 128)
        v = v.view(T, 4, 128)
        md = ops.build_decode_metadata(r2t, rpi, seq, 4)
        for l in range(L):
            ops.kv_store(slabs[l][0], slabs[l][1], loc, k, v)
            ops.paged_attention(q, slabs[l][0], slabs[l][1], r2t, md.plan, 128 ** -0.5, out=outs[l])

    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        step()                                             # warm-up outside capture (function attributes, allocator)
    side.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        step()
    # the batch moves on: one more token per request, written into the SAME buffers
    seq.add_(1)
    qkv.copy_(torch.randn(qkv.shape, generator=torch.Generator().manual_seed(99)).to(torch.bfloat16))
    graph.replay()
    torch.cuda.synchronize()
    got = [o.clone() for o in outs]
    for o in outs:
        o.zero_()
    step()                                                 # eager, same inputs (the stores are idempotent)
    torch.cuda.synchronize()
    for l in range(L):
        assert torch.equal(got[l], outs[l]), f"layer {l}: graph replay differs from eager launches"
    # and it is the right answer for the moved-on batch
    q, k, v = qkv.cpu().float().split([28 * 128, 4 * 128, 4 * 128], dim=-1)
    ks, vs = c.k_slab.float().clone(), c.v_slab.float().clone()
    O.kv_store(ks, vs, c.out_cache_loc, k.reshape(T, 4, 128).to(torch.bfloat16).float(),
               v.reshape(T, 4, 128).to(torch.bfloat16).float())
    ref = O.decode_attention(q.reshape(T, 28, 128), ks.to(torch.bfloat16).float(), vs.to(torch.bfloat16).float(),
                             c.req_to_token, c.req_pool_indices, c.seq_lens)
    assert (got[0].float().cpu() - ref).abs().max().item() <= TOL


def _stored(c):
    ks, vs = c.k_slab.float().clone(), c.v_slab.float().clone()
    O.kv_store(ks, vs, c.out_cache_loc, c.k.float(), c.v.float())
    return ks, vs


@pytest.mark.parametrize("nhg_heads", [1, 2, 4, 8])
@pytest.mark.parametrize("n_ctas", [1, 7, 148])
def test_decode_metadata_bit_exact(ops, nhg_heads, n_ctas):
    g = torch.Generator().manual_seed(n_ctas + nhg_heads)
    B = 300
    seq = torch.randint(1, 5000, (B,), generator=g)
    C = 5000
    r2t = torch.zeros(B + 5, C, dtype=torch.int32)
    rpi = torch.randperm(B + 5, generator=g)[:B]
    md = ops.build_decode_metadata(r2t.cuda(), rpi.cuda(), seq.cuda(), nhg_heads, n_ctas=n_ctas)
    torch.cuda.synchronize()
    omd = O.build_metadata(r2t, rpi, seq, None)
    assert torch.equal(md.positions.cpu(), omd.positions)
    assert torch.equal(md.kv_indptr.cpu(), omd.kv_indptr)
    assert torch.equal(md.qo_indptr.cpu(), omd.qo_indptr)
    assert torch.equal(md.plan.row_kv_len.cpu(), seq.to(torch.int32))
    assert torch.equal(md.plan.row_idx_off.cpu(), rpi * C)
    nhg = nhg_heads // ops.heads_per_group(nhg_heads)
    hdr = md.plan.plan_hdr.cpu().tolist()
    n_eff = hdr[1]
    tile_indptr, begins, first = O.split_plan(seq.tolist(), nhg, 16, n_eff)
    assert hdr[0] == tile_indptr[-1] and hdr[2] == B * nhg
    assert n_eff == max(1, min(n_ctas, tile_indptr[-1] // _lib.load().bsgl_min_tiles_per_cta()))
    assert md.plan.tile_indptr.cpu().tolist() == tile_indptr
    assert md.plan.cta_first_seq.cpu().tolist()[: n_eff + 1] == first


def test_kv_indices_gather(ops):
    c = make_paged_case(B=5, Hq=4, Hkv=2, D=64, seq_lens=[5, 1, 300, 17, 64], seed=3)
    md = ops.build_decode_metadata(c.req_to_token.cuda(), c.req_pool_indices.cuda(), c.seq_lens.cuda(), 2)
    got = ops.gather_kv_indices(c.req_to_token.cuda(), c.req_pool_indices.cuda(), md.kv_indptr, int(c.seq_lens.sum()))
    want = O.build_metadata(c.req_to_token, c.req_pool_indices, c.seq_lens, None).kv_indices
    assert torch.equal(got.cpu(), want)


def test_kv_store_casts_and_strides(ops):
    torch.manual_seed(0)
    S, H, D, T = 50, 2, 128, 9
    for src_dt, dst_dt in [(torch.float32, torch.bfloat16), (torch.float16, torch.bfloat16),
                           (torch.bfloat16, torch.float16), (torch.float32, torch.float16)]:
        k_slab = torch.zeros(S + 1, H, D, dtype=dst_dt, device="cuda")
        v_slab = torch.zeros(S + 1, H, D, dtype=dst_dt, device="cuda")
        fused = torch.randn(T, 3 * H * D, device="cuda").to(src_dt)
        k = fused[:, H * D:2 * H * D].view(T, H, D)        # both strided
        v = fused[:, 2 * H * D:].view(T, H, D)
        loc = (torch.randperm(S)[:T] + 1).to(torch.int32).cuda()
        ops.kv_store(k_slab, v_slab, loc, k, v)
        torch.cuda.synchronize()
        assert torch.equal(k_slab[loc.long()], k.to(dst_dt))
        assert torch.equal(v_slab[loc.long()], v.to(dst_dt))
        assert k_slab[0].abs().sum().item() == 0


def test_cpu_tensors_are_refused(ops):
    from baby_sglang_b200._lib import BsglError
    with pytest.raises(BsglError):
        ops.kv_store(torch.zeros(3, 1, 64, dtype=torch.bfloat16), torch.zeros(3, 1, 64, dtype=torch.bfloat16),
                     torch.tensor([1], dtype=torch.int32), torch.zeros(1, 1, 64, dtype=torch.bfloat16),
                     torch.zeros(1, 1, 64, dtype=torch.bfloat16))


def test_host_staged_layers_match_direct_calls(ops):
    """HostStagedAttention (pinned host q/k/v in, host outputs out, copies on side streams) returns
    bit-for-bit what per-layer calls of the drop-in backend return, for decode and extend."""
    from baby_sglang_b200 import DROPIN_DIR
    from baby_sglang_b200.host_staging import HostStagedAttention
    if DROPIN_DIR not in sys.path:
        sys.path.insert(0, DROPIN_DIR)
    import attn_backend
    import forward_batch
    import memory_pool

    Hq, Hkv, D, L = 12, 2, 128, 3
    for prefill in (False, True):
        seq = [300, 17, 129, 64]
        pre = [260, 0, 100, 1] if prefill else None
        c = make_paged_case(B=4, Hq=Hq, Hkv=Hkv, D=D, seq_lens=seq, prefix_lens=pre, seed=5)
        T = c.qkv.shape[0]

        class Layer:
            def __init__(self, i):
                self.layer_id, self.head_dim, self.scaling = i, D, D ** -0.5
                self.tp_q_head_num, self.tp_k_head_num, self.tp_v_head_num = Hq, Hkv, Hkv

        layers = [Layer(i) for i in range(L)]
        g = torch.Generator().manual_seed(3)
        qkv_host = [torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).to(torch.bfloat16).pin_memory() for _ in range(L)]

        def fresh():
            kv = memory_pool.MHATokenToKVPool(size=c.S, dtype=torch.bfloat16, head_num=Hkv, head_dim=D, layer_num=L,
                                              device="cuda")
            for l in range(L):
                kv.k_buffer[l].copy_(c.k_slab)
                kv.v_buffer[l].copy_(c.v_slab)
            r2t = memory_pool.ReqToTokenPool(size=c.req_to_token.shape[0], max_context_len=c.C, device="cuda")
            r2t.req_to_token.copy_(c.req_to_token)
            common = dict(input_ids=torch.zeros(T, dtype=torch.int64, device="cuda"),
                          req_pool_indices=c.req_pool_indices.cuda(), seq_lens=c.seq_lens.cuda(),
                          out_cache_loc=c.out_cache_loc.cuda(), req_to_token_pool=r2t, token_to_kv_pool=kv)
            if prefill:
                return forward_batch.SimplifiedForwardBatch.create_prefill_batch(prefix_lens=c.prefix_lens.cuda(), **common)
            return forward_batch.SimplifiedForwardBatch.create_decode_batch(**common)

        fb = fresh()
        direct = []
        for l in range(L):
            x = qkv_host[l].cuda()
            q, k, v = x.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
            direct.append(attn_backend.SimpleAttentionBackend.forward(
                q.contiguous().view(T, Hq, D), k.contiguous().view(T, Hkv, D), v.view(T, Hkv, D), layers[l], fb, True).cpu())
        fb2 = fresh()
        out_host = [torch.empty(T, Hq * D, dtype=torch.bfloat16).pin_memory() for _ in range(L)]
        HostStagedAttention(Hq, Hkv, D, torch.device("cuda")).forward_layers(qkv_host, out_host, layers, fb2)
        for l in range(L):
            assert torch.equal(out_host[l], direct[l]), f"layer {l} prefill={prefill}"
            assert torch.equal(fb.token_to_kv_pool.k_buffer[l], fb2.token_to_kv_pool.k_buffer[l])


# ---------------------------------------------------------------------------- K5 RoPE + split + store
def _ulp_mismatch(a: torch.Tensor, b: torch.Tensor):
    """(#elements that differ, max difference in units of the 16-bit pattern)."""
    ia, ib = a.view(torch.int16).int(), b.view(torch.int16).int()
    d = (ia - ib).abs()
    return int((d != 0).sum()), int(d.max()) if d.numel() else 0


@pytest.mark.parametrize("name", __import__("tests.helpers", fromlist=["x"]).rope_golden_names())
def test_rope_kv_store_vs_reference_fixture(ops, name):
    """K5 against RotaryEmbedding outputs the reference produced (CPU libm cos/sin) and, bit for bit,
    against the oracle evaluated on the device (CUDA cosf/sinf, what the reference calls on a GPU)."""
    from tests.helpers import load_rope_golden
    g = load_rope_golden(name)
    T = g.qkv.shape[0]
    S = T + 9
    gen = torch.Generator().manual_seed(2)
    loc = (torch.randperm(S, generator=gen)[:T] + 1).to(torch.int32)
    k_slab = torch.randn(S + 1, g.Hkv, g.D, generator=gen).to(g.dtype).cuda()
    v_slab = torch.randn(S + 1, g.Hkv, g.D, generator=gen).to(g.dtype).cuda()
    k0, v0 = k_slab.clone(), v_slab.clone()
    qkv, pos, inv = g.qkv.cuda(), g.positions.cuda(), g.inv_freq.cuda()
    q_rot = ops.rope_kv_store(qkv, pos, inv, g.Hq, g.Hkv, g.D, k_slab, v_slab, loc.cuda())
    torch.cuda.synchronize()
    # (1) device-evaluated oracle: bit-exact, including every untouched slab row
    ko, vo = k0.clone(), v0.clone()
    q_ref = O.rope_kv_store(qkv, pos, inv, g.Hq, g.Hkv, g.D, ko, vo, loc.cuda())
    assert torch.equal(q_rot.view(torch.int16), q_ref.view(torch.int16))
    assert torch.equal(k_slab.view(torch.int16), ko.view(torch.int16))
    assert torch.equal(v_slab.view(torch.int16), vo.view(torch.int16))
    # (2) the reference's own CPU output: identical up to the last bit of cos/sin on a few elements
    nq, dq = _ulp_mismatch(q_rot.cpu(), g.q_rot)
    nk, dk = _ulp_mismatch(k_slab.cpu()[loc.long()], g.k_rot)
    assert dq <= 1 and dk <= 1, (dq, dk)
    assert nq <= 0.01 * g.q_rot.numel() and nk <= 0.01 * g.k_rot.numel(), (nq, nk)
    # (3) rotate-only form leaves the slabs alone
    k1 = k_slab.clone()
    q_only = ops.rope_kv_store(qkv, pos, inv, g.Hq, g.Hkv, g.D)
    assert torch.equal(q_only, q_rot) and torch.equal(k1, k_slab)


def test_rope_kv_store_large_and_strided(ops):
    """Full-size decode/prefill shapes (7B geometry), a row-strided qkv view, positions up to 32K."""
    Hq, Hkv, D = 28, 4, 128
    for T in (256, 4099):
        gen = torch.Generator(device="cuda").manual_seed(T)
        wide = torch.randn(T, (Hq + 2 * Hkv) * D + 64, device="cuda", generator=gen).to(torch.bfloat16)
        qkv = wide[:, : (Hq + 2 * Hkv) * D]                     # row stride != row length
        pos = torch.randint(0, 32768, (T,), device="cuda", generator=gen)
        inv = (1.0 / (1e6 ** (torch.arange(0, D, 2).float() / D))).cuda()
        S = T + 3
        loc = (torch.randperm(S, device="cuda", generator=gen)[:T] + 1).to(torch.int32)
        k_slab = torch.zeros(S + 1, Hkv, D, device="cuda", dtype=torch.bfloat16)
        v_slab = torch.zeros_like(k_slab)
        q_rot = ops.rope_kv_store(qkv, pos, inv, Hq, Hkv, D, k_slab, v_slab, loc)
        ko, vo = torch.zeros_like(k_slab), torch.zeros_like(v_slab)
        q_ref = O.rope_kv_store(qkv, pos, inv, Hq, Hkv, D, ko, vo, loc)
        assert torch.equal(q_rot, q_ref) and torch.equal(k_slab, ko) and torch.equal(v_slab, vo)


# ---------------------------------------------------------------------------- K6 batched greedy sampling
def test_greedy_sample_vs_reference_fixture(ops):
    from tests.helpers import load_sample_golden, prev_ids_csr
    for name, c in load_sample_golden().items():
        prev, indptr = prev_ids_csr(c.prev_ids, "cuda")
        pens = torch.tensor(c.penalties, dtype=torch.float32, device="cuda")
        qo = None
        if c.mode == "prefill":
            qo = torch.tensor([0] + list(torch.tensor(c.extended_lens).cumsum(0)), dtype=torch.int32, device="cuda")
        got = ops.greedy_sample(c.logits.cuda(), qo, prev, indptr, pens)
        assert got.dtype == torch.int64 and torch.equal(got.cpu(), c.next_ids), name
        if all(p <= 1.0 or not h for p, h in zip(c.penalties, c.prev_ids)):
            assert torch.equal(ops.greedy_sample(c.logits.cuda(), qo).cpu(), c.next_ids)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_greedy_sample_full_vocab_vs_oracle(ops, dtype):
    """Qwen2.5 vocabulary (152 064), B=64: ties, NaN rows, histories with repeats, prefill gather, odd row stride."""
    from tests.helpers import prev_ids_csr
    V, B = 152064, 64
    g = torch.Generator().manual_seed(9)
    ext = torch.randint(1, 4, (B,), generator=g).tolist()
    T = sum(ext)
    wide = (torch.randn(T, V + 8, generator=g) * 4).to(dtype)
    logits = wide[:, :V]
    ends = torch.tensor(ext).cumsum(0) - 1
    for i in range(B):
        r = int(ends[i])
        j = int(torch.argmax(logits[r].float()))
        if i % 3 == 0:
            logits[r, (j * 7 + 1) % V] = logits[r, j]                # exact tie: lowest index wins
        if i == 5:
            logits[r, 4242] = float("nan")
    prev = [[] if i % 4 == 0 else torch.topk(logits[int(ends[i])].float(), 3).indices.tolist() * 2 +
            torch.randint(0, V, (50,), generator=g).tolist() for i in range(B)]
    pens = [1.0 + (i % 5) * 0.35 for i in range(B)]
    want = O.greedy_sample(logits, ext, prev, pens)
    p, ip = prev_ids_csr(prev, "cuda")
    qo = torch.tensor([0] + torch.tensor(ext).cumsum(0).tolist(), dtype=torch.int32, device="cuda")
    got = ops.greedy_sample(wide.cuda()[:, :V], qo, p, ip, torch.tensor(pens, dtype=torch.float32, device="cuda"))
    assert torch.equal(got.cpu(), want)
    # decode form (row i = request i), no penalty: plain argmax with first-index ties
    dec = logits[ends]
    got2 = ops.greedy_sample(dec.cuda().contiguous())
    assert torch.equal(got2.cpu(), O.greedy_sample(dec, None, [[]] * B, [1.0] * B))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("B,V", [(1, 152064), (7, 32000), (300, 151936), (3, 1000), (5, 50001)])
def test_greedy_sample_register_path_edges(ops, dtype, B, V):
    """The no-penalty 16-bit path (slice in registers, cluster-wide max then lowest index) against torch.argmax
    semantics: exact ties, -0.0 vs +0.0, NaN of either sign (first NaN wins), rows of -inf, every cluster split,
    and vocabularies that are not a multiple of 8 (which take the exact path)."""
    g = torch.Generator().manual_seed(B * 1000 + V)
    logits = (torch.randn(B, V, generator=g) * 3).to(dtype)
    for i in range(B):
        j = int(torch.argmax(logits[i].float()))
        kind = i % 6
        if kind == 0:                                  # tie at a lower and a higher index
            logits[i, (j * 5 + 3) % V] = logits[i, j]
        elif kind == 1:                                # NaN with the sign bit set, then a positive one later
            logits[i, V // 3] = torch.tensor(float("nan"), dtype=dtype).view(torch.int16).bitwise_or(-32768).view(dtype)
            logits[i, V // 2] = float("nan")
        elif kind == 2:                                # the whole row -inf: index 0
            logits[i] = float("-inf")
        elif kind == 3:                                # maximum is zero: -0.0 first, +0.0 later -> first index
            logits[i] = -logits[i].abs() - 1
            logits[i, 11] = -0.0
            logits[i, V - 1] = 0.0
        elif kind == 4:                                # maximum in the very last element
            logits[i, V - 1] = 100.0
    want = O.greedy_sample(logits, None, [[]] * B, [1.0] * B)
    got = ops.greedy_sample(logits.cuda())
    assert torch.equal(got.cpu(), want)
    for i in range(B):                                 # and the oracle itself against torch on the edge rows
        if i % 6 in (2, 3):
            assert int(want[i]) == int(torch.argmax(logits[i].float()))


def test_batched_greedy_sampler_hook(ops):
    """`install_batched_greedy_sampler` on an object with the reference scheduler's method: greedy batches go
    through K6 (ids equal to the fixture the reference produced), other batches reach the original method."""
    from types import SimpleNamespace
    from baby_sglang_b200.model_hooks import install_batched_greedy_sampler
    from tests.helpers import load_sample_golden

    class FakeScheduler:
        calls = 0

        def _sample_next_ids(self, logits, batch_requests, mode):     # stands in for scheduler.py:509-546
            FakeScheduler.calls += 1
            return torch.full((len(batch_requests),), -7, dtype=torch.long, device=logits.device)

    sched = FakeScheduler()
    install_batched_greedy_sampler(sched)
    for name, c in load_sample_golden().items():
        reqs = []
        for i, e in enumerate(c.extended_lens):
            reqs.append(SimpleNamespace(input_ids=list(range(e + 2)), prefix_indices=[0, 1], constraint_state=None,
                                        do_sample=False, output_ids=c.prev_ids[i], repetition_penalty=c.penalties[i]))
        got = sched._sample_next_ids(c.logits.cuda(), reqs, c.mode)
        assert torch.equal(got.cpu(), c.next_ids), name
        assert FakeScheduler.calls == 0
    reqs[0].do_sample = True
    assert int(sched._sample_next_ids(c.logits.cuda(), reqs, c.mode)[0]) == -7 and FakeScheduler.calls == 1
