"""CPU-side checks: the C-ABI library loads and exports every symbol include/bsgl_b200.h declares,
host-only entry points behave, and the product path refuses to run without CUDA."""
import os
import re

import pytest
import torch

from baby_sglang_b200 import LIB_PATH, _lib

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.isfile(LIB_PATH):
        from baby_sglang_b200 import build
        build.build()
    return _lib.load()


def _declared_symbols():
    text = open(os.path.join(REPO, "include", "bsgl_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bsgl_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(lib):
    names = _declared_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/bsgl_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes table and header disagree"


def test_version_and_host_queries(lib):
    assert lib.bsgl_version() == 100
    assert [lib.bsgl_heads_per_group(h) for h in (1, 2, 3, 4, 6, 8, 12)] == [1, 2, 1, 4, 2, 4, 4]
    import ctypes
    env = os.environ.get("BSGL_MIN_TILES_PER_CTA")
    assert lib.bsgl_min_tiles_per_cta() == (max(1, int(env)) if env and int(env) >= 1 else _lib.MIN_TILES_PER_CTA)
    n = ctypes.c_int64()
    assert lib.bsgl_oproj_flag_bytes(10, 40, ctypes.byref(n)) == 0 and n.value == 1024 + 256 * 32
    o, l = ctypes.c_int64(), ctypes.c_int64()
    assert lib.bsgl_partial_sizes(148, 28, 4, 128, ctypes.byref(o), ctypes.byref(l)) == 0
    assert o.value == 2 * 148 * 28 * 128 and l.value == 2 * 148 * 28


def test_invalid_arguments_report_errors(lib):
    # bad geometry is rejected before any CUDA call, with a message in the thread-local slot
    rc = lib.bsgl_partial_sizes(148, 28, 5, 128, None, None)
    assert rc == 1
    assert b"n_q_heads" in lib.bsgl_last_error()
    rc = lib.bsgl_kv_store(None, None, 9, None, None, None, 4, 2, 128, 256, 128, 256, 128, 0, 0, None)
    assert rc == 1 and b"null" in lib.bsgl_last_error()
    rc = lib.bsgl_kv_store(None, None, 9, None, None, None, 0, 2, 128, 256, 128, 256, 128, 0, 0, None)
    assert rc == 0   # empty store is a no-op


def test_no_cpu_fallback():
    from baby_sglang_b200 import ops
    z = torch.zeros(3, 1, 64, dtype=torch.bfloat16)
    with pytest.raises(_lib.BsglError, match="no CPU path"):
        ops.kv_store(z, z.clone(), torch.tensor([1], dtype=torch.int32), z[:1], z[:1])
    with pytest.raises(_lib.BsglError, match="no CPU path"):
        ops.build_decode_metadata(torch.zeros(2, 8, dtype=torch.int32), torch.tensor([0]), torch.tensor([3]), 1)


def test_product_never_imports_oracle():
    pkg = os.path.join(REPO, "baby_sglang_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(root, f)).read()
                assert "oracle" not in src.replace("paged_attention_oracle", "oracle") or "import oracle" not in src
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f"{f} imports the oracle"
