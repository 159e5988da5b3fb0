"""Independent derivation of the CLSM graph with torch CPU autograd (float64 or float32).
Test infrastructure: the NumPy oracle's hand-written backward is checked against this, and
its float32 run is the 'reference fp32' for the 1e-5 tolerance."""
import numpy as np
import torch
import torch.nn.functional as Fn

TINY32 = float(np.finfo(np.float32).tiny)


def act(x, name):
    if name == "relu":
        return torch.relu(x)
    if name == "tanh":
        return torch.tanh(x)
    return x


def tower(tok, p, actn, mask=None, keep=1.0):
    X = p["E"][tok]                                   # (S, L, D)
    if mask is not None:
        X = X * mask / keep
    pooled = []
    for W, b in zip(p["convW"], p["convb"]):
        # conv1d wants (S, D, L) and weight (F, D, k); cross-correlation like TF
        y = Fn.conv1d(X.transpose(1, 2), W.permute(2, 1, 0)) + b[None, :, None]
        y = act(y, actn)
        pooled.append(y.max(dim=2).values)           # torch max -> first index gradient
    P = torch.cat(pooled, dim=1)
    return act(P @ p["Wd"] + p["bd"], actn), P


def loss_fn(zq, zd, gamma):
    def norm(x):
        return torch.sqrt(torch.clamp((x * x).sum(-1), min=TINY32))
    raw = (zq[:, None, :] * zd).sum(-1) / (norm(zq)[:, None] * norm(zd))
    cos = torch.clamp(raw, 0.0, 1.0)
    e = torch.exp(gamma * cos)
    p0 = e[:, 0] / e.sum(1)
    ph = torch.clamp(p0, 1e-7, 1 - 1e-7)
    return (-torch.log(ph)).mean(), cos, p0


def to_torch(p, dtype):
    out = {}
    for k, v in p.items():
        if isinstance(v, list):
            out[k] = [torch.tensor(a, dtype=dtype, requires_grad=True) for a in v]
        else:
            out[k] = torch.tensor(v, dtype=dtype, requires_grad=True)
    return out


def loss_and_grads(q_tok, d_tok, pq, pd, actn="relu", gamma=10.0, dtype=torch.float64,
                   mask_q=None, mask_d=None, keep=1.0):
    tq, td = to_torch(pq, dtype), to_torch(pd, dtype)
    B, J1, Ld = d_tok.shape
    mq = None if mask_q is None else torch.tensor(mask_q, dtype=dtype)
    md = None if mask_d is None else torch.tensor(mask_d, dtype=dtype)
    zq, Pq = tower(torch.tensor(q_tok, dtype=torch.long), tq, actn, mq, keep)
    zd, Pd = tower(torch.tensor(d_tok.reshape(B * J1, Ld), dtype=torch.long), td, actn, md, keep)
    loss, cos, p0 = loss_fn(zq, zd.reshape(B, J1, -1), gamma)
    loss.backward()

    def grads(t):
        g = {}
        for k, v in t.items():
            g[k] = [a.grad.numpy() for a in v] if isinstance(v, list) else v.grad.numpy()
        return g
    return {"loss": float(loss.detach()), "cos": cos.detach().numpy(), "p0": p0.detach().numpy(),
            "zq": zq.detach().numpy(), "zd": zd.detach().numpy(),
            "Pq": Pq.detach().numpy(), "Pd": Pd.detach().numpy()}, grads(tq), grads(td)
