"""Generate tests/golden/rope_*.npz by running the UNMODIFIED reference's RotaryEmbedding
(qwen2.py:227-268, imported through baseline/ref_loader.py) on seeded inputs, in the 16-bit
activation dtypes the CUDA path supports.  Build container only; the .npz files are committed.

    python tests/golden/make_rope_golden.py
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from baseline import ref_loader  # noqa: E402


def bits16(x: torch.Tensor) -> np.ndarray:
    return x.contiguous().view(torch.int16).numpy().astype(np.uint16)


def make(ref, name, *, dtype, Hq, Hkv, D, T, max_pos, theta, seed):
    gen = torch.Generator().manual_seed(seed)
    emb = ref.qwen2.RotaryEmbedding(D, 32768, theta)
    positions = torch.randint(0, max_pos, (T,), generator=gen, dtype=torch.int64)
    positions[0] = 0
    positions[-1] = max_pos - 1
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=gen).to(dtype)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)      # qwen2.py:285-290
    q_rot, k_rot = emb(positions, q.view(T, Hq, D), k.view(T, Hkv, D))   # qwen2.py:185-191
    assert q_rot.dtype == dtype
    np.savez_compressed(os.path.join(HERE, name + ".npz"), Hq=Hq, Hkv=Hkv, D=D, is_bf16=int(dtype == torch.bfloat16),
                        positions=positions.numpy(), inv_freq=emb.inv_freq.numpy(), qkv=bits16(qkv),
                        q_rot=bits16(q_rot), k_rot=bits16(k_rot))
    print(name, tuple(q_rot.shape), tuple(k_rot.shape))


def main():
    ref = ref_loader.load_stack(names=("memory_pool", "forward_batch", "attn_backend", "qwen2"))
    make(ref, "rope_bf16_d128_g7", dtype=torch.bfloat16, Hq=28, Hkv=4, D=128, T=67, max_pos=32768, theta=1e6, seed=301)
    make(ref, "rope_bf16_d64", dtype=torch.bfloat16, Hq=14, Hkv=2, D=64, T=33, max_pos=4096, theta=1e4, seed=302)
    make(ref, "rope_f16_d128_g6", dtype=torch.float16, Hq=12, Hkv=2, D=128, T=40, max_pos=32768, theta=1e6, seed=303)


if __name__ == "__main__":
    main()
