"""Development aid: a few world-1 launches of K8 on 1024 x 5120 x 5120 for an ncu capture
(`BSGL_K8_GEMM=pair|single ncu -k regex:oproj ... python scripts/k8_world1_once.py`)."""
import os
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from baby_sglang_b200 import DROPIN_DIR  # noqa: E402

sys.path.insert(0, DROPIN_DIR)
import tp_runtime  # noqa: E402

torch.cuda.set_device(0)
comm = tp_runtime.FusedOProjAllReduce(0, 1, None, hidden=5120, max_tokens=1024)
a = torch.randn(1024, 5120, device="cuda", dtype=torch.bfloat16)
w = torch.randn(5120, 5120, device="cuda", dtype=torch.bfloat16) * 0.05
for _ in range(6):
    comm(a, w)
torch.cuda.synchronize()
comm.close()
print("done")
