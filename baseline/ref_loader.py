"""Load the UNMODIFIED reference (baby-llm/baby-sglang) in isolation — harness code, not product: the package
`baby_sglang_b200/` never imports it.  Consumers: tests/ (reference stack vs drop-in stack), tests/golden/make_*.py
(fixture generation), bench.py (the reference's own scheduler + model DRIVING the drop-in path for the end-to-end
decode tok/s number, and the reference's CPU path as the timed baseline).

The reference is a flat directory of modules that import each other by bare top-level
name (`from attn_backend import ...`, qwen2.py:5-7, scheduler.py:8-9), so "which
attn_backend/forward_batch/memory_pool/tp_runtime is first on sys.path" IS its plug
mechanism.  `load_stack()` imports the reference's modules with an optional shadow
directory in front (our drop-in modules), then removes them from `sys.modules` again so
two stacks (reference-only, and reference + drop-in) can live in one process.

Reference location: $BABYSGL_REF, else /root/reference, else <repo>/baseline/_ref.  The last one is
staged by `scripts/stage_reference.py` (git-ignored, not gpurun-ignored, so it travels to the GPU box with
the snapshot, like the built .so); callers must skip when `find_reference()` returns None.
"""
from __future__ import annotations

import importlib
import os
import sys
from types import SimpleNamespace
from typing import Optional

_REF_MODULES = (
    "memory_pool", "forward_batch", "attn_backend", "tp_runtime", "radix_tree", "constraints",
    "request", "sample", "radix_cache", "qwen2", "model_loader", "overlap_worker", "scheduler",
)
_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def find_reference() -> Optional[str]:
    for cand in (os.environ.get("BABYSGL_REF"), "/root/reference", os.path.join(_REPO, "baseline", "_ref")):
        if cand and os.path.isfile(os.path.join(cand, "attn_backend.py")) and \
                os.path.isfile(os.path.join(cand, "scheduler.py")):
            return cand
    return None


def _transformers_shim() -> None:
    """scheduler.py:16 imports `transformers.tokenization_utils.PreTrainedTokenizerBase`,
    which transformers>=5 no longer exposes there (the reference pins <5.0.0,
    requirements.txt:2).  One attribute alias; nothing in the reference is touched."""
    try:
        import transformers
        import transformers.tokenization_utils as tu
        if not hasattr(tu, "PreTrainedTokenizerBase"):
            tu.PreTrainedTokenizerBase = transformers.PreTrainedTokenizerBase
    except Exception:  # pragma: no cover - transformers missing altogether
        pass


def load_stack(shadow_dir: Optional[str] = None, names=_REF_MODULES, optional=("constraints", "scheduler", "radix_cache", "request", "overlap_worker")) -> SimpleNamespace:
    """Import the reference modules (with `shadow_dir` taking precedence for same-named
    modules) and return them as attributes of a namespace.  sys.modules / sys.path are
    restored afterwards."""
    ref = find_reference()
    if ref is None:
        raise FileNotFoundError("no reference checkout found ($BABYSGL_REF, /root/reference, baseline/_ref)")
    _transformers_shim()
    saved_modules = {n: sys.modules.pop(n) for n in list(sys.modules) if n in _REF_MODULES}
    saved_path = list(sys.path)
    front = ([shadow_dir] if shadow_dir else []) + [ref]
    sys.path[:] = front + [p for p in saved_path if p not in front]
    out = {}
    try:
        for n in names:
            try:
                out[n] = importlib.import_module(n)
            except Exception as e:  # e.g. lm-format-enforcer missing
                if n in optional:
                    out[n] = None
                    out[n + "_error"] = repr(e)
                else:
                    raise
    finally:
        for n in _REF_MODULES:
            sys.modules.pop(n, None)
        sys.modules.update(saved_modules)
        sys.path[:] = saved_path
    out["ref_dir"] = ref
    out["shadow_dir"] = shadow_dir
    return SimpleNamespace(**out)
