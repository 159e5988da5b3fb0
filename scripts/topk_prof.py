import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dnn_page_vectors_b200 import _lib
from tests.gpu_util import P, ST, DEV
lib = _lib.lib
N, Q, H, k = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, int(sys.argv[2]) if len(sys.argv) > 2 else 128, 128, 100
g = torch.Generator(device="cuda").manual_seed(3)
c = torch.nn.functional.normalize(torch.randn((N, H), generator=g, device="cuda"), dim=1).to(torch.bfloat16)
q = torch.nn.functional.normalize(torch.randn((Q, H), generator=g, device="cuda"), dim=1).to(torch.bfloat16)
wsb = lib.clsm_cos_topk_bf16_workspace_bytes(Q, N, k)
ws = torch.empty(wsb, dtype=torch.uint8, device=DEV)
sc = torch.empty((Q, k), dtype=torch.float32, device=DEV); ids = torch.empty((Q, k), dtype=torch.int64, device=DEV)
for _ in range(3):
    _lib.call("clsm_cos_topk_bf16", P(q), Q, P(c), N, H, k, 0, P(sc), P(ids), P(ws), wsb, ST())
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
_lib.call("clsm_cos_topk_bf16", P(q), Q, P(c), N, H, k, 0, P(sc), P(ids), P(ws), wsb, ST())
e1.record(); torch.cuda.synchronize()
print("ms", e0.elapsed_time(e1))
