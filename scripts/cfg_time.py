"""Device time of one resident training step at the other BASELINE.json shapes (cfg3: 50 negatives of 1000 tokens,
cfg4: four window sizes) -- parity-test shapes, timed here once for DESIGN.md; bench.py stays on cfg2.
    python scripts/cfg_time.py cfg3|cfg4|ref [batch]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from dnn_page_vectors_b200 import CLSM, ModelParams

CFG = {"cfg3": dict(V=1_000_000, D=300, windows=(3,), F=300, H=128, J=50, Lq=20, Ld=1000, B=256, act="relu"),
       "ref": dict(V=50_000, D=100, windows=(3, 4), F=150, H=150, J=3, Lq=20, Ld=975, B=128, act="relu"),     # the reference script's literals
       "cfg4": dict(V=1_000_000, D=300, windows=(1, 2, 3, 5), F=300, H=128, J=4, Lq=20, Ld=100, B=8192, act="relu")}
name = sys.argv[1]
c = dict(CFG[name])
if len(sys.argv) > 2:
    c["B"] = int(sys.argv[2])
mp = ModelParams(embedding_dim=c["D"], batch_size=c["B"], filter_sizes=c["windows"], num_filters=c["F"], hidden_dims=c["H"],
                 J=c["J"], dropout_prob=(0.25, 0.5), activation=c["act"], precision="bf16")
rng = np.random.RandomState(1337)
table = rng.uniform(-0.25, 0.25, size=(c["V"] + 1, c["D"])).astype(np.float32)
net = CLSM(mp, c["Lq"], c["Ld"], embedding_weights=table, device="cuda:0").compile()
del table
x = bench.make_batches(c, 1, 1337)[0]
B = c["B"]
bq, bd = net.query_model.buffers(B, True), net.doc_model.buffers(B * (1 + c["J"]), True)
bq["tok"].copy_(torch.as_tensor(x[0]).cuda())
bd["tok"].copy_(torch.as_tensor(np.stack(x[1:], axis=1)).cuda().view(B * (1 + c["J"]), c["Ld"]))
for _ in range(3):
    h = net.step_resident(B, bq, bd)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
steps = 5
e0.record()
for _ in range(steps):
    h = net.step_resident(B, bq, bd)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
fl = 3 * bench.flops_pair_fwd(c) * B
if os.environ.get("PHASES"):
    ph, _ = bench.profile_phases(net, lambda i: net.step_resident(B, bq, bd), torch)
    print(json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in ph.items()}))
print(json.dumps({"config": name, "B": B, "ms_per_step": ms, "pairs_per_s": B / (ms * 1e-3), "model_tflops_3x": fl / (ms * 1e-3) / 1e12,
                  "loss": float(h["loss_acc"][0]), "mem_GB": torch.cuda.max_memory_allocated() / 1e9}))
