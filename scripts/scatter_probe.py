"""Times clsm_emb_scatter_plan / _reduce alone on the cfg2 document-tower token stream (2.05 M tokens, D = 300)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as BN
from tests.gpu_util import P, ST, call, lib
from dnn_page_vectors_b200 import _lib

c = dict(BN.CFG2)
rng = np.random.RandomState(1)
tok = torch.as_tensor(np.stack([BN.zipf_tokens(rng, c["V"], c["B"], c["Ld"]) for _ in range(1 + c["J"])], 1).reshape(-1)).cuda()
n, D, Dpad, V1 = tok.numel(), 300, 320, c["V"] + 1
bf16 = (sys.argv[1] if len(sys.argv) > 1 else "bf16") == "bf16"
dX = torch.randn((n, Dpad), device="cuda").to(torch.bfloat16 if bf16 else torch.float32)
bits = torch.randint(-32768, 32767, (n, Dpad // 16), dtype=torch.int16, device="cuda")
wsb = lib.clsm_emb_scatter_workspace_bytes(n, D)
ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
uniq = torch.empty(n, dtype=torch.int32, device="cuda"); nu = torch.zeros(1, dtype=torch.int32, device="cuda")
rows = torch.empty((min(n, V1), Dpad), dtype=torch.float32, device="cuda"); roi = torch.empty(V1, dtype=torch.int32, device="cuda")
code = _lib.BF16 if bf16 else _lib.F32
def plan(): call("clsm_emb_scatter_plan", P(tok), n, Dpad, D, V1, P(uniq), P(nu), P(roi), P(ws), wsb, ST())
def red(): call("clsm_emb_scatter_reduce", n, P(dX), code, Dpad, D, 0.75, 5, P(bits), P(rows), P(ws), wsb, ST())
plan()
for _ in range(3): red()
torch.cuda.synchronize()
e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
e[0].record(); plan(); e[1].record()
for _ in range(5): red()
e[2].record(); torch.cuda.synchronize()
print("n_tok %d uniq %d | plan %.3f ms | reduce %.3f ms (%.0f GB/s of dX)" % (n, int(nu.item()), e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]) / 5,
      n * Dpad * (2 if bf16 else 4) / 1e9 / (e[1].elapsed_time(e[2]) / 5 * 1e-3)))
