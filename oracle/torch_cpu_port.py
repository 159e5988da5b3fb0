"""CPU port of the CLSM training / encode / retrieval path for TIMING -- TEST INFRASTRUCTURE ONLY.

The reference's own stack (Keras 1.0.x on Theano 0.8 / TensorFlow 0.x, Python 2) cannot be
installed on this box, so the timed CPU baseline is this torch-CPU float32 restatement of exactly
the graph dssm_cnn_v2/cnn_dssm_tf.py:63-182 builds: two towers with separate embedding tables ->
Conv1D(valid)+act per window -> global max-pool -> concat -> Dense+act -> clipped cosine x(1+J) ->
gamma softmax -> -log p0 mean -> Keras-1 Adam applied densely to every weight including both
tables.  oneDNN conv/GEMM with all host threads.  Only bench.py (cpu_baseline / --impl reference)
and tests may import it.  PARITY UNPINNED for the same reason as clsm_oracle.py; it is checked
against clsm_oracle.py in tests/test_oracle_model.py.
"""
from __future__ import annotations

import math
import os
import time

import numpy as np
import torch
import torch.nn.functional as Fn

TINY32 = float(np.finfo(np.float32).tiny)


class CpuCLSM(object):
    def __init__(self, V1, D, windows, F, H, J, gamma=10.0, act="relu", seed=1337, threads=None):
        torch.set_num_threads(threads or os.cpu_count() or 1)
        self.threads = torch.get_num_threads()
        g = torch.Generator().manual_seed(seed)
        self.act, self.gamma, self.J = act, gamma, J
        self.params = []

        def tower():
            P = len(windows) * F
            lim = math.sqrt(6.0 / (P + H))
            t = {"E": (torch.rand((V1, D), generator=g) - 0.5) * 0.5,
                 "convW": [(torch.rand((F, D, k), generator=g) - 0.5) * 0.1 for k in windows],
                 "convb": [torch.zeros(F) for _ in windows],
                 "Wd": (torch.rand((P, H), generator=g) * 2 - 1) * lim, "bd": torch.zeros(H)}
            flat = [t["E"]] + t["convW"] + t["convb"] + [t["Wd"], t["bd"]]
            for p in flat:
                p.requires_grad_(True)
            self.params += flat
            return t
        self.q, self.d = tower(), tower()
        self.m = [torch.zeros_like(p) for p in self.params]
        self.v = [torch.zeros_like(p) for p in self.params]
        self.t = 0

    def _act(self, x):
        return torch.relu(x) if self.act == "relu" else torch.tanh(x) if self.act == "tanh" else x

    def tower_fwd(self, tok, t):
        X = Fn.embedding(tok, t["E"]).transpose(1, 2)            # (S, D, L)
        pooled = [self._act(Fn.conv1d(X, W, b)).amax(dim=2) for W, b in zip(t["convW"], t["convb"])]
        return self._act(torch.cat(pooled, 1) @ t["Wd"] + t["bd"])

    def loss(self, q_tok, d_tok):
        B, J1, Ld = d_tok.shape
        zq = self.tower_fwd(q_tok, self.q)
        zd = self.tower_fwd(d_tok.reshape(B * J1, Ld), self.d).reshape(B, J1, -1)

        def norm(x):
            return torch.sqrt(torch.clamp((x * x).sum(-1), min=TINY32))
        cos = torch.clamp((zq[:, None, :] * zd).sum(-1) / (norm(zq)[:, None] * norm(zd)), 0.0, 1.0)
        e = torch.exp(self.gamma * cos)
        p0 = e[:, 0] / e.sum(1)
        return (-torch.log(torch.clamp(p0, 1e-7, 1 - 1e-7))).mean()

    def train_step(self, q_tok, d_tok, lr=1e-3, b1=0.9, b2=0.999, eps=1e-8):
        for p in self.params:
            p.grad = None
        loss = self.loss(q_tok, d_tok)
        loss.backward()
        self.t += 1
        lr_t = lr * math.sqrt(1 - b2 ** self.t) / (1 - b1 ** self.t)
        with torch.no_grad():
            for p, m, v in zip(self.params, self.m, self.v):     # dense, every weight (Keras)
                g = p.grad if p.grad is not None else torch.zeros_like(p)
                m.mul_(b1).add_(g, alpha=1 - b1)
                v.mul_(b2).addcmul_(g, g, value=1 - b2)
                p.addcdiv_(m, v.sqrt().add_(eps), value=-lr_t)
        return float(loss.detach())

    @torch.no_grad()
    def encode(self, tok):
        z = self.tower_fwd(tok, self.d)
        return z / torch.sqrt(torch.clamp((z * z).sum(-1, keepdim=True), min=TINY32))

    @torch.no_grad()
    def topk(self, qn, cn, k):
        return torch.topk(qn @ cn.T, k, dim=1)


def time_train(cfg, B, steps, warmup, seed=1337, batches=None, budget_s=None):
    """Times `steps` training steps after `warmup` untimed ones.  batches: optional list of [q (B, Lq), pos (B, Ld),
    neg_0.. (B, Ld)] integer arrays (the reference's batch format, data_helpers.py:196-197) cycled through; default
    uniform random ids.  budget_s: stop early once the timed steps have used that many seconds (at least one step
    is always timed).  Returns (pairs_per_s, threads, seconds_per_step, steps_done)."""
    net = CpuCLSM(cfg["V"] + 1, cfg["D"], cfg["windows"], cfg["F"], cfg["H"], cfg["J"], act=cfg["act"], seed=seed)
    if batches is None:
        g = torch.Generator().manual_seed(seed)
        data = [(torch.randint(0, cfg["V"], (B, cfg["Lq"]), generator=g),
                 torch.randint(0, cfg["V"], (B, 1 + cfg["J"], cfg["Ld"]), generator=g))]
    else:
        data = [(torch.as_tensor(np.asarray(x[0])).long(), torch.as_tensor(np.stack(x[1:], axis=1)).long()) for x in batches]
        B = int(data[0][0].shape[0])
    for i in range(warmup):
        net.train_step(*data[i % len(data)])
    done, t0 = 0, time.perf_counter()
    for i in range(max(steps, 1)):
        net.train_step(*data[i % len(data)])
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = (time.perf_counter() - t0) / done
    return B / dt, net.threads, dt, done
