import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from dnn_page_vectors_b200 import CLSM, ModelParams
act = sys.argv[1] if len(sys.argv) > 1 else "tanh"
V = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
c = dict(bench.CFG2, V=V, act=act)
mp = ModelParams(embedding_dim=c["D"], batch_size=c["B"], filter_sizes=c["windows"], num_filters=c["F"],
                 hidden_dims=c["H"], J=c["J"], dropout_prob=(0.25, 0.5), activation=c["act"], precision="bf16")
rng = np.random.RandomState(1337)
table = rng.uniform(-0.25, 0.25, size=(c["V"] + 1, c["D"])).astype(np.float32)
net = CLSM(mp, c["Lq"], c["Ld"], embedding_weights=table, device="cuda:0").compile()
batches = bench.make_batches(c, 2, 1)
for i in range(8):
    l = net.train_on_batch(batches[i % 2])
    bq = net.query_model.buffers(c["B"], True); bd = net.doc_model.buffers(c["B"] * 5, True)
    h = net._head_bufs(c["B"])
    print(i, l, "semq|", float(bq["sem"].abs().mean()), "semd|", float(bd["sem"].abs().mean()),
          "pooled", float(bd["pooled"].mean()), "cos", float(h["cos"].mean()), float(h["cos"].max()),
          "gWd", float(net.doc_model.g_Wd.abs().max()), "n_uniq", int(bd["n_uniq"].item()),
          "dE", float(bd["dE_rows"][:10].abs().max()), flush=True)
