#!/usr/bin/env python
"""Stage the UNMODIFIED reference (baby-llm/baby-sglang) under baseline/_ref/ so it travels to the GPU box.

    python scripts/stage_reference.py [--src /root/reference] [--check]

Why: the north-star gate is "scheduler.py, radix_cache.py, overlap_worker.py and qwen2.py run unchanged on the
drop-in modules".  Those files are the reference's; they are never copied into the tracked tree, never imported
by the package, and `baseline/_ref/` is git-ignored (it is NOT gpurun-ignored, so a `gpurun` snapshot and the
driver's round-end snapshot carry it, like the built `.so`).  Test infrastructure only:
`baseline/ref_loader.find_reference()` is its one consumer (tests/, bench.py's reference arm).

How: the contract's install (`pip install --no-index --no-build-isolation --target baseline/_ref <src>`) is tried
first.  The reference is a flat directory of modules with neither setup.py nor pyproject.toml, so pip answers
"not installable"; the install of such a project IS its top-level modules on a path, so the `.py` files (plus
README/requirements for attribution) are placed there byte for byte.  A sha256 manifest is written next to them;
`--check` re-verifies an existing staging against the source (or, without the source, against the manifest).
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import shutil
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEST = os.path.join(REPO, "baseline", "_ref")
KEEP = (".py", ".md", ".txt")


def _sha(path: str) -> str:
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def _files(src: str):
    return sorted(n for n in os.listdir(src) if n.endswith(KEEP) and os.path.isfile(os.path.join(src, n)))


def stage(src: str) -> dict:
    os.makedirs(DEST, exist_ok=True)
    pip = subprocess.run(
        [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
         "/opt/wheelhouse", "--target", DEST, src], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    pip_tail = pip.stdout.strip().splitlines()[-1] if pip.stdout.strip() else ""
    manifest = {"source": src, "pip_install_rc": pip.returncode, "pip_install_says": pip_tail, "files": {}}
    if pip.returncode != 0:          # flat module directory: place the modules themselves
        for n in _files(src):
            shutil.copyfile(os.path.join(src, n), os.path.join(DEST, n))
            os.chmod(os.path.join(DEST, n), 0o444)
    for n in _files(DEST):
        manifest["files"][n] = _sha(os.path.join(DEST, n))
    with open(os.path.join(DEST, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1)
    return manifest


def check(src: str | None) -> bool:
    mpath = os.path.join(DEST, "MANIFEST.json")
    if not os.path.isfile(mpath):
        print("no staging at", DEST)
        return False
    with open(mpath) as f:
        manifest = json.load(f)
    ok = True
    for n, h in manifest["files"].items():
        here = _sha(os.path.join(DEST, n))
        if here != h:
            print("MODIFIED since staging:", n)
            ok = False
        if src and os.path.isfile(os.path.join(src, n)) and _sha(os.path.join(src, n)) != here:
            print("differs from source:", n)
            ok = False
    print(f"{len(manifest['files'])} files, {'all identical' if ok else 'MISMATCH'}")
    return ok


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--src", default="/root/reference")
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    have_src = os.path.isdir(a.src)
    if a.check:
        sys.exit(0 if check(a.src if have_src else None) else 1)
    if not have_src:
        sys.exit(f"{a.src} not found")
    m = stage(a.src)
    print(f"staged {len(m['files'])} files under {DEST} (pip: rc={m['pip_install_rc']}: {m['pip_install_says']})")
