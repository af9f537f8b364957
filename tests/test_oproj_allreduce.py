"""K8: o_proj GEMM with the fused NVLink all-reduce epilogue.

Single GPU (`world = 1`): the whole kernel runs (TMA -> tcgen05 -> staging -> owner reduce -> output), only the
peers are missing; checked against an fp32 matmul of the same bf16 operands.  Two or more GPUs: tests/multi_gpu_oproj.py
under torchrun (bit-identical across ranks, vs fp32 truth, vs `F.linear` + NCCL all-reduce); skipped with one GPU.
"""
import os
import subprocess
import sys

import pytest
import torch

from baby_sglang_b200 import DROPIN_DIR

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tp():
    if DROPIN_DIR not in sys.path:
        sys.path.insert(0, DROPIN_DIR)
    import tp_runtime
    return tp_runtime


@pytest.mark.gpu
@pytest.mark.parametrize("gemm", ["auto", "single", "pair"])
@pytest.mark.parametrize("T,hidden,K,dtype", [(1024, 5120, 640, torch.bfloat16), (333, 1536, 768, torch.bfloat16),
                                              (1, 3584, 3584, torch.bfloat16), (130, 256, 64, torch.float16),
                                              (1000, 5120, 2560, torch.bfloat16)])
def test_fused_oproj_world1_matches_fp32_matmul(T, hidden, K, dtype, gemm, monkeypatch):
    """`gemm`: the single-CTA 128 x 128 tiles, the cta_group::2 256 x 256 pair blocks (forced: also odd numbers of tile
    rows, where the peer CTA of the last row pair has no rows), or the library's own choice."""
    if gemm != "auto":
        monkeypatch.setenv("BSGL_K8_GEMM", gemm)
    else:
        monkeypatch.delenv("BSGL_K8_GEMM", raising=False)
    tp = _tp()
    comm = tp.FusedOProjAllReduce(0, 1, None, hidden=hidden, max_tokens=1024, dtype=dtype)
    try:
        g = torch.Generator().manual_seed(T + K)
        a = torch.randn(T, K, generator=g).to("cuda", dtype)
        w = (torch.randn(hidden, K, generator=g) * 0.05).to("cuda", dtype)
        ref = a.float() @ w.float().t()
        outs = [comm(a, w).clone() for _ in range(3)]          # consecutive epochs
        torch.cuda.synchronize()
        tol = 2.0 ** (-8 if dtype == torch.bfloat16 else -11) * ref.abs().max().item() * 1.01   # one rounding of the result
        for o in outs:
            assert o.shape == (T, hidden) and o.dtype == dtype
            assert (o.float() - ref).abs().max().item() <= tol
        assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])
        # a strided activation view (rows further apart than K) goes through the tensor map's row pitch
        wide = torch.randn(T, K + 64, generator=g).to("cuda", dtype)
        o2 = comm(wide[:, :K], w).clone()
        assert (o2.float() - wide[:, :K].float() @ w.float().t()).abs().max().item() <= tol * 1.5
    finally:
        comm.close()


@pytest.mark.gpu
def test_fused_oproj_rejects_bad_shapes():
    tp = _tp()
    from baby_sglang_b200 import _lib
    comm = tp.FusedOProjAllReduce(0, 1, None, hidden=256, max_tokens=128)
    try:
        a = torch.zeros(200, 64, device="cuda", dtype=torch.bfloat16)
        w = torch.zeros(256, 64, device="cuda", dtype=torch.bfloat16)
        with pytest.raises(ValueError, match="max_tokens"):
            comm(a, w)
        with pytest.raises(_lib.BsglError, match="multiple of 64"):
            comm(a[:100, :32].contiguous(), w[:, :32].contiguous())
        with pytest.raises(_lib.BsglError, match="no CPU path"):
            comm(a[:100].cpu(), w)
    finally:
        comm.close()


@pytest.mark.gpu
def test_fused_oproj_is_cuda_graph_capturable():
    """The call counter (epoch) lives in device memory and the last CTA of each launch advances it, so a captured
    call replays as a fresh epoch each time (the host-side epoch of round 1 would have been baked into
    the graph and every later replay would have passed its flag waits at once).  Multi-rank: tests/test_tp_gpu.py."""
    tp = _tp()
    T, hidden, K = 300, 1536, 256
    comm = tp.FusedOProjAllReduce(0, 1, None, hidden=hidden, max_tokens=512)
    try:
        g = torch.Generator().manual_seed(1)
        a = torch.randn(T, K, generator=g).to("cuda", torch.bfloat16)
        w = (torch.randn(hidden, K, generator=g) * 0.05).to("cuda", torch.bfloat16)
        want = comm(a, w).clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            comm(a, w)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = comm(a, w).clone()
        for _ in range(4):
            out.zero_()
            a2 = a.clone()                       # eager calls between replays keep numbering consistent too
            graph.replay()
            torch.cuda.synchronize()
            assert torch.equal(out, want)
            assert torch.equal(comm(a2, w), want)
    finally:
        comm.close()
