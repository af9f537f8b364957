"""Developer tool: run K3 (tcgen05 extend) on a few shapes and print WHERE it disagrees with the
oracle (per request / q head / token block / head-dim block), to localise layout bugs in one GPU call.

    python scripts/debug_extend.py [case ...]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from baby_sglang_b200 import ops  # noqa: E402
from oracle import paged_attention_oracle as O  # noqa: E402
from tests.helpers import make_paged_case  # noqa: E402

CASES = {
    # name: (Hq, Hkv, D, seq_lens, prefix_lens)
    "tiny": (1, 1, 128, [16], [0]),                  # one head, one KV tile, one q tile
    "tile": (1, 1, 128, [128], [0]),
    "twotile": (1, 1, 128, [256], [0]),
    "prefix": (1, 1, 128, [300], [290]),
    "long": (1, 1, 128, [1000], [900]),
    "gqa": (7, 1, 128, [200], [100]),
    "d64": (2, 1, 64, [200], [100]),
    "multi": (28, 4, 128, [700, 64, 300], [500, 0, 299]),
}


def run(name):
    Hq, Hkv, D, seq, pre = CASES[name]
    c = make_paged_case(B=len(seq), Hq=Hq, Hkv=Hkv, D=D, seq_lens=seq, prefix_lens=pre, seed=3)
    dt = torch.bfloat16
    k_slab, v_slab = c.k_slab.cuda(), c.v_slab.cuda()
    r2t = c.req_to_token.cuda()
    T = c.T
    q = c.q.cuda()
    ops.kv_store(k_slab, v_slab, c.out_cache_loc.cuda(), c.k.cuda(), c.v.cuda())
    md = ops.build_extend_metadata(r2t, c.req_pool_indices.cuda(), c.seq_lens.cuda(), c.prefix_lens.cuda(), T, Hkv,
                                   row_plan=False)
    out = ops.extend_attention_tc(q, k_slab, v_slab, r2t, c.req_pool_indices.cuda(), c.seq_lens.cuda(),
                                  c.prefix_lens.cuda(), md.qo_indptr, D ** -0.5)
    torch.cuda.synchronize()
    ks, vs = c.k_slab.float().clone(), c.v_slab.float().clone()
    ref = O.backend_forward(c.q.float(), c.k.float(), c.v.float(), ks, vs, c.req_to_token, c.req_pool_indices,
                            c.seq_lens, c.out_cache_loc, True, c.prefix_lens, c.extended_lens)
    o = out.float().cpu().view(T, Hq, D)
    r = ref.view(T, Hq, D)
    err = (o - r).abs()
    print(f"== {name}: Hq={Hq} Hkv={Hkv} D={D} seq={seq} prefix={pre}  max_err={err.max().item():.4e} "
          f"nan={int(torch.isnan(o).sum())} mean|ref|={r.abs().mean().item():.3f}")
    if err.max().item() > 2e-2 or torch.isnan(o).any():
        e = torch.nan_to_num(err, nan=99.0)
        print("   per head max:", [round(x, 3) for x in e.amax(dim=(0, 2)).tolist()])
        print("   per 16-wide d block:", [round(x, 3) for x in e.view(T, Hq, D // 16, 16).amax(dim=(0, 1, 3)).tolist()])
        tb = 8
        n = (T + tb - 1) // tb
        rows = e.amax(dim=(1, 2))
        print("   per 8-token block:", [round(rows[i * tb:(i + 1) * tb].max().item(), 3) for i in range(min(n, 48))])
        t_bad = int(rows.argmax())
        print(f"   worst token {t_bad}: out[:8]={o[t_bad, 0, :8].tolist()}\n                   ref[:8]={r[t_bad, 0, :8].tolist()}")
        # is the output a scaled / permuted version of the reference?
        ratio = (o[t_bad, 0] / r[t_bad, 0])
        print(f"   ratio out/ref at worst token (first 8): {ratio[:8].tolist()}")
    return err.max().item()


if __name__ == "__main__":
    names = sys.argv[1:] or list(CASES)
    bad = 0
    for n in names:
        try:
            if run(n) > 2e-2:
                bad += 1
        except Exception as ex:  # keep going: a trap poisons the context, so report and stop
            print(f"== {n}: EXCEPTION {type(ex).__name__}: {ex}")
            bad += 1
            break
    print("FAILED" if bad else "ALL OK")
    sys.exit(1 if bad else 0)
