"""bsgl_paged_attention_static: decode attention for hosts that cannot allocate device memory.

Some users of the C ABI have no way to provide the split-KV partial buffers that
bsgl_paged_attention borrows: a serving worker whose HBM was fully divided between weights and
the KV pool at start-up (a later cudaMalloc fails, or stalls both of the overlap worker's
streams because it synchronises the device), or a pass-through cgo / JNI / C binding with no
device allocator of its own.  For them the library carries an 8 MiB partials array in its own
device image and this entry point uses it.

What is checked here, through the C ABI only:
  * two calls in a row (different geometries, both with sequences split over many CTAs, so the
    second call finds the first call's leftovers in the shared array) each give exactly the
    bits that bsgl_paged_attention + bsgl_merge_partials give with caller-provided buffers --
    the same kernels run, only the scratch address differs, so the comparison is exact, which
    is inside the 2e-2 tolerance the decode tests use against the fp32 oracle;
  * repeating both calls changes neither the result nor the amount of free device memory;
  * a request whose partials exceed the built-in capacity is refused with a message that names
    both sizes, and nothing is launched.

This file sorts last on purpose: under `pytest -x` a failure here cannot hide other tests.
It passes alone in a fresh process (the first call resolves the array's address) and after the
rest of the suite.
"""
import ctypes

import pytest
import torch

from baby_sglang_b200 import _lib
from tests.helpers import make_paged_case

pytestmark = pytest.mark.gpu

TOL = 2e-2   # the decode tests' bound vs the fp32 oracle; the comparisons below are exact (0 <= TOL)


def _prepare(ops, Hq, Hkv, D, seq_lens, dtype, seed):
    c = make_paged_case(B=len(seq_lens), Hq=Hq, Hkv=Hkv, D=D, seq_lens=seq_lens, seed=seed, dtype=dtype,
                        device="cuda")
    md = ops.build_decode_metadata(c.req_to_token, c.req_pool_indices, c.seq_lens, Hkv)
    return c, md.plan


def _common_args(c, plan, out):
    return (c.q.data_ptr(), c.k_slab.data_ptr(), c.v_slab.data_ptr(), c.req_to_token.data_ptr(),
            plan.row_idx_off.data_ptr(), plan.row_kv_len.data_ptr(), plan.plan_hdr.data_ptr(),
            plan.tile_indptr.data_ptr(), plan.cta_first_seq.data_ptr(), c.q.shape[0], c.Hq, c.Hkv, c.D,
            float(c.D ** -0.5), _lib.BF16 if c.q.dtype == torch.bfloat16 else _lib.F16, out.data_ptr())


def _with_caller_buffers(lib, ops, c, plan):
    """The existing pair of entry points, partial buffers allocated by the caller (torch)."""
    out = torch.full((c.q.shape[0], c.Hq * c.D), float("nan"), dtype=c.q.dtype, device="cuda")
    part_o, part_lse = ops.partial_buffers(out.device, plan.n_ctas, c.Hq, c.Hkv, c.D)
    st = torch.cuda.current_stream().cuda_stream
    a = _common_args(c, plan, out)
    _lib.check(lib.bsgl_paged_attention(*a, part_o.data_ptr(), part_lse.data_ptr(), plan.n_ctas, st),
               "bsgl_paged_attention")
    _lib.check(lib.bsgl_merge_partials(part_o.data_ptr(), part_lse.data_ptr(), plan.plan_hdr.data_ptr(),
                                       plan.tile_indptr.data_ptr(), plan.cta_first_seq.data_ptr(), plan.n_ctas,
                                       c.q.shape[0], c.Hq, c.Hkv, c.D, a[14], out.data_ptr(), st),
               "bsgl_merge_partials")
    torch.cuda.synchronize()
    return out


def _with_builtin_scratch(lib, c, plan, out):
    st = torch.cuda.current_stream().cuda_stream
    out.fill_(float("nan"))
    _lib.check(lib.bsgl_paged_attention_static(*_common_args(c, plan, out), plan.n_ctas, st),
               "bsgl_paged_attention_static")
    torch.cuda.synchronize()
    return out


def test_static_scratch_matches_caller_buffers_twice_in_a_row():
    from baby_sglang_b200 import ops
    lib = _lib.load()
    assert lib.bsgl_static_scratch_bytes() == 8 << 20
    # long sequences on 148 CTAs: most of them are split, so the partials really go through the array
    cases = [_prepare(ops, 28, 4, 128, [4096, 3, 777, 2048, 1], torch.bfloat16, seed=11),
             _prepare(ops, 12, 2, 128, [1500, 2600, 17, 900], torch.float16, seed=12)]
    want = [_with_caller_buffers(lib, ops, c, plan) for c, plan in cases]
    outs = [torch.empty_like(w) for w in want]

    # two calls in a row; the second one starts with the first one's partials still in the array
    for (c, plan), w, o in zip(cases, want, outs):
        got = _with_builtin_scratch(lib, c, plan, o)
        assert not torch.isnan(got.float()).any()
        err = (got.float() - w.float()).abs().max().item()
        assert err <= TOL, err
        assert torch.equal(got.view(torch.int16), w.view(torch.int16)), "same kernels, same bits expected"

    # again, in the other order: same bits, and the device's free memory does not move (nothing allocated)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for (c, plan), w, o in reversed(list(zip(cases, want, outs))):
        _with_builtin_scratch(lib, c, plan, o)
    assert torch.cuda.mem_get_info()[0] == free_before
    for w, o in zip(want, outs):
        assert torch.equal(o.view(torch.int16), w.view(torch.int16))


def test_static_scratch_rejects_oversized_requests():
    from baby_sglang_b200 import ops
    lib = _lib.load()
    c, plan = _prepare(ops, 28, 4, 128, [300, 40], torch.bfloat16, seed=13)
    out = torch.zeros((c.q.shape[0], c.Hq * c.D), dtype=c.q.dtype, device="cuda")
    o, l = ctypes.c_int64(), ctypes.c_int64()
    n_ctas = 1024                                  # 1024 CTAs x 2 slots x 28 heads x 128 floats = 28 MiB of part_o
    assert lib.bsgl_partial_sizes(n_ctas, c.Hq, c.Hkv, c.D, ctypes.byref(o), ctypes.byref(l)) == 0
    assert (o.value + l.value) * 4 > lib.bsgl_static_scratch_bytes()
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.bsgl_paged_attention_static(*_common_args(c, plan, out), n_ctas, st)
    assert rc == 1
    msg = lib.bsgl_last_error().decode()
    assert "built-in scratch" in msg and str(8 << 20) in msg, msg
    torch.cuda.synchronize()
    assert int(out.view(torch.int16).abs().sum()) == 0      # nothing ran
