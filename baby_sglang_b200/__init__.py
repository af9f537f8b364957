"""baby_sglang_b200 — B200-native (sm_100a) paged-attention path for baby-sglang.

Scope: KV store, paged GQA decode (split-KV + LSE merge), ragged extend and the
device-side batch metadata, behind the reference's own `attn_backend` / `forward_batch` /
`memory_pool` / `tp_runtime` module interface (`baby_sglang_b200/dropin/`).  Host code is
Python/PyTorch (device memory, streams, torch.distributed); the compute is hand-written CUDA in
`csrc/`, reached through the C ABI declared in `include/bsgl_b200.h`.

There is no CPU path: importing `ops` without the built library raises.
"""
import os

PACKAGE_DIR = os.path.dirname(os.path.abspath(__file__))
DROPIN_DIR = os.path.join(PACKAGE_DIR, "dropin")
LIB_PATH = os.path.join(PACKAGE_DIR, "lib", "libbsgl_b200.so")

__all__ = ["PACKAGE_DIR", "DROPIN_DIR", "LIB_PATH"]
