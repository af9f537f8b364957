"""Ad-hoc timing of the K3 extend kernel (development aid; bench.py is the contract).

Config C3: Qwen2.5-7B geometry, B=16, prefix_i ~ U[0, 8192] (element 0 forced to 0,
element 1 to 8192), ext_i = 512, bf16, fragmented page table.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from baby_sglang_b200 import ops  # noqa: E402

GEOM = {"1.5b": (12, 2, 128), "7b": (28, 4, 128), "32b": (40, 8, 128), "32b_tp8": (5, 1, 128)}


def extend_flops(Hq, D, prefix, ext):
    """QK^T and PV, unmasked MACs only."""
    return 4.0 * Hq * D * sum(e * (p + (e + 1) / 2.0) for p, e in zip(prefix, ext))


def build(geom, batch, max_prefix, ext_len, seed=0, fixed_prefix=None):
    Hq, Hkv, D = GEOM[geom]
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(seed)
    if fixed_prefix is not None:
        prefix = torch.full((batch,), fixed_prefix, dtype=torch.int64)
    else:
        prefix = torch.randint(0, max_prefix + 1, (batch,), generator=g)
        prefix[0] = 0
        if batch > 1:
            prefix[1] = max_prefix
    ext = torch.full((batch,), ext_len, dtype=torch.int64)
    seq = prefix + ext
    C = int(seq.max()) + 64
    S = int(seq.sum()) + 1024
    perm = (torch.randperm(S, generator=g) + 1).to(torch.int32)
    r2t = torch.zeros(batch, C, dtype=torch.int32)
    off = 0
    for i, n in enumerate(seq.tolist()):
        r2t[i, :n] = perm[off:off + n]
        off += n
    T = int(ext.sum())
    k_slab = torch.randn(S + 1, Hkv, D, device=dev, dtype=torch.bfloat16)
    v_slab = torch.randn(S + 1, Hkv, D, device=dev, dtype=torch.bfloat16)
    q = torch.randn(T, Hq, D, device=dev, dtype=torch.bfloat16)
    return dict(Hq=Hq, Hkv=Hkv, D=D, prefix=prefix, ext=ext, seq=seq, r2t=r2t.to(dev), k_slab=k_slab, v_slab=v_slab,
                q=q, T=T, rpi=torch.arange(batch, device=dev))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geom", default="7b")
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--max-prefix", type=int, default=8192)
    ap.add_argument("--fixed-prefix", type=int, default=None)
    ap.add_argument("--ext", type=int, default=512)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--mode", default="tcgen05", choices=["tcgen05", "rowwise"])
    ap.add_argument("--no-order", action="store_true", help="dispatch requests in index order (no heavy-first)")
    ap.add_argument("--static", action="store_true", help="deal the work items statically (no device work counter)")
    ap.add_argument("--sustain", type=float, default=0.0,
                    help="also launch back to back for this many seconds while nvidia-smi samples SM clock and power")
    ap.add_argument("--lib", default=None, help="load this build of the library instead of the in-tree one "
                                                "(e.g. one compiled with -DBSGL_EXT_EXPERIMENTS)")
    a = ap.parse_args()
    if a.lib:
        from baby_sglang_b200 import _lib
        _lib.LIB_PATH = os.path.abspath(a.lib)
    w = build(a.geom, a.batch, a.max_prefix, a.ext, fixed_prefix=a.fixed_prefix)
    dev = w["q"].device
    seq, pre = w["seq"].to(dev), w["prefix"].to(dev)
    md = ops.build_extend_metadata(w["r2t"], w["rpi"], seq, pre, w["T"], w["Hkv"], row_plan=(a.mode == "rowwise"))
    out = torch.empty(w["T"], w["Hq"] * w["D"], device=dev, dtype=torch.bfloat16)
    scale = w["D"] ** -0.5

    def run():
        if a.mode == "tcgen05":
            ops.extend_attention_tc(w["q"], w["k_slab"], w["v_slab"], w["r2t"], w["rpi"], seq, pre, md.qo_indptr,
                                    scale, out=out, req_order=None if a.no_order else md.heavy_order,
                                    **({} if a.static or a.lib else {"work_counter": md.work_counter}))
        else:
            ops.paged_attention(w["q"], w["k_slab"], w["v_slab"], w["r2t"], md.plan, scale, out=out)

    for _ in range(3):
        run()
    torch.cuda.synchronize()
    ms = []
    for _ in range(a.iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    med = ms[len(ms) // 2]
    fl = extend_flops(w["Hq"], w["D"], w["prefix"].tolist(), w["ext"].tolist())
    sustained = None
    if a.sustain > 0:
        import subprocess
        import time
        smi = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,power.draw,power.limit,"
                                "clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-lms", "100"],
                               stdout=subprocess.PIPE, text=True)
        time.sleep(0.3)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n, t0 = 0, time.time()
        e0.record()
        while time.time() - t0 < a.sustain:
            for _ in range(50):
                run()
            n += 50
            torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
        smi.terminate()
        rows = [[x.strip() for x in ln.split(",")] for ln in smi.stdout.read().strip().splitlines()]
        rows = [r for r in rows if len(r) == 4][3:]
        clk = sorted(float(r[0]) for r in rows)
        pw = sorted(float(r[1]) for r in rows)
        sustained = {"launches": n, "ms_avg": round(e0.elapsed_time(e1) / n, 4),
                     "TFLOPs": round(fl / (e0.elapsed_time(e1) / n) / 1e9, 1),
                     "sm_mhz_median": clk[len(clk) // 2] if clk else None, "power_w_median": pw[len(pw) // 2] if pw else None,
                     "power_limit_w": float(rows[0][2]) if rows else None,
                     "power_cap_active_share": round(sum(r[3].lower().startswith("active") for r in rows) / max(1, len(rows)), 2)}
    print(json.dumps({"geom": a.geom, "batch": a.batch, "ext": a.ext, "mode": a.mode,
                      "prefix": "fixed %d" % a.fixed_prefix if a.fixed_prefix is not None else "U[0,%d]" % a.max_prefix,
                      "ms_med": round(med, 4), "ms_min": round(ms[0], 4), "TFLOPs_med": round(fl / med / 1e9, 1),
                      "TFLOPs_best": round(fl / ms[0] / 1e9, 1), "frac_of_1665.9": round(fl / med / 1e9 / 1665.9, 4),
                      "flops": fl, "sustained": sustained, "expmode": os.environ.get("BSGL_EXT_EXPMODE")}))


if __name__ == "__main__":
    main()
