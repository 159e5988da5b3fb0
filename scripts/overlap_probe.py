"""Where the two streams of CLSM.step_resident sit on the device time line (cfg2, one GPU): start/end of the query
tower's and the document tower's post-loss work relative to the fork, and the host time to issue a step."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as BN
from dnn_page_vectors_b200 import CLSM, ModelParams

c = dict(BN.CFG2)
mp = ModelParams(embedding_dim=c["D"], batch_size=c["B"], filter_sizes=c["windows"], num_filters=c["F"], hidden_dims=c["H"], J=c["J"],
                 dropout_prob=(0.25, 0.5), activation=c["act"], precision="bf16")
rng = np.random.RandomState(1337)
table = rng.uniform(-0.25, 0.25, size=(c["V"] + 1, c["D"])).astype(np.float32)
net = CLSM(mp, c["Lq"], c["Ld"], embedding_weights=table).compile()
B = c["B"]
xs = BN.make_batches(c, 12, 1)
net.assume_optimizer_state()
bq, bd = net.query_model.buffers(B, True), net.doc_model.buffers(B * (1 + c["J"]), True)
dev_x = [(torch.as_tensor(x[0]).cuda(), torch.as_tensor(np.stack(x[1:], 1)).view(B * (1 + c["J"]), c["Ld"]).cuda()) for x in xs]
_n = [0]
def load():
    q, d = dev_x[_n[0] % len(dev_x)]; _n[0] += 1
    bq["tok"].copy_(q); bd["tok"].copy_(d)
for _ in range(5):
    load()
    net.step_resident(B, bq, bd)
torch.cuda.synchronize()
ev = lambda: torch.cuda.Event(enable_timing=True)
for it in range(4):
    net._probe = {k: ev() for k in ("q0", "q1", "d0", "d1", "fq", "fd", "hd", "dx", "sc", "tab", "dw")}
    load()
    s0, s1 = ev(), ev()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    s0.record()
    net.step_resident(B, bq, bd)
    s1.record()
    t_issue = time.perf_counter() - t0
    torch.cuda.synchronize()
    p = net._probe
    print("step %.2f ms | host issue %.2f ms | query work [%.2f, %.2f] | doc work [%.2f, %.2f] (ms after step start)" % (
        s0.elapsed_time(s1), t_issue * 1e3, s0.elapsed_time(p["q0"]), s0.elapsed_time(p["q1"]), s0.elapsed_time(p["d0"]), s0.elapsed_time(p["d1"])))
    if net.split_bwd:
        print("   ms after step start: query fwd done %.2f | doc fwd done %.2f | doc dense bwd done %.2f | dX done %.2f | dW done %.2f | scatter done %.2f | table optimizer done %.2f" % tuple(
            s0.elapsed_time(p[k]) for k in ("fq", "fd", "hd", "dx", "dw", "sc", "tab")))
net._probe = None
# host-only issue rate: same step with nothing to wait for
t0 = time.perf_counter()
for _ in range(10):
    load()
    net.step_resident(B, bq, bd)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("10 steps: host issue %.2f ms/step, wall %.2f ms/step" % ((t1 - t0) * 100, (t2 - t0) * 100))
