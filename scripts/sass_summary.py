#!/usr/bin/env python
"""Opcode counts per kernel from `cuobjdump -sass` of the shipped library -> profiles/sass_summary.txt, so the
Blackwell-native claim (tcgen05 = UTCHMMA, TMEM ld/st = LDTM/STTM, tcgen05.commit = UTCBAR, TMA = UTMALDG,
cp.async = LDGSTS, NVLS multimem = LDGMC/REDG..., mma.sync = HMMA) is checkable from the repo.

    python scripts/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(REPO, "baby_sglang_b200", "lib", "libbsgl_b200.so")
WATCH = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UTCCP", "UTMALDG", "UTMASTG", "UBLKCP", "LDGSTS", "HMMA",
         "LDSM", "MOVM", "LDGMC", "STGMC", "REDGMC", "MUFU", "FFMA2", "SYNCS", "ACQBULK", "BAR", "LDG", "STG", "ATOMG",
         "MEMBAR", "NANOSLEEP"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    total = collections.Counter()
    print(f"# {os.path.relpath(LIB, REPO)}: SASS opcode counts per kernel (cuobjdump -sass, sm_100a)")
    print("# only opcodes that occur are listed; 'insts' = all instructions of the kernel\n")
    for (name, ctr), pretty in zip(kernels.items(), demangle):
        short = re.sub(r"\(.*", "", pretty)
        hits = [f"{op}={ctr[op]}" for op in WATCH if ctr[op]]
        print(f"{short}\n    insts={sum(ctr.values())}  " + "  ".join(hits))
        total.update(ctr)
    print("\n# library totals: " + "  ".join(f"{op}={total[op]}" for op in WATCH if total[op]))


if __name__ == "__main__":
    sys.exit(main())
