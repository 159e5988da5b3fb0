"""CPU oracle for the CLSM (convolutional DSSM) hot path -- TEST INFRASTRUCTURE ONLY.

This file is a NumPy restatement of the arithmetic the reference builds out of
Keras-1 layers in dssm_cnn_v2/cnn_dssm_tf.py (and its Theano twin
cnn_dssm_th.py).  It is the checker for the CUDA path; only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import it.  Nothing under dnn_page_vectors_b200/ imports it.

PARITY UNPINNED: the reference ships no tests, golden vectors or recorded
outputs for this path, and its Keras 1.0.x / Theano 0.8 / TF 0.x stack cannot
be installed here (Python 2 only, no network), so the Keras-internal
conventions restated below (first-index max-pool gradient, inclusive clip
gradient, BCE epsilon 1e-7, Adam bias-correction form) are written down from
the published Keras 1.0.x / backend sources, not checked against a run of the
reference.  The token-packing half of the path IS pinned: see
oracle/packing_oracle.py, which is validated against the reference's own
utils/data_utils.py (it runs under Python 3) through tests/golden/.

Reference map (all paths relative to /root/reference):
  embedding lookup ........ dssm_cnn_v2/cnn_dssm_tf.py:111-116
  dropout on embeddings ... dssm_cnn_v2/cnn_dssm_tf.py:117-122
  Convolution1D(valid) .... dssm_cnn_v2/cnn_dssm_tf.py:84-91
  MaxPooling1D + Flatten .. dssm_cnn_v2/cnn_dssm_tf.py:92-100
  concat of windows ....... dssm_cnn_v2/cnn_dssm_tf.py:102-107
  Dense + relu ............ dssm_cnn_v2/cnn_dssm_tf.py:124-126
  clipped cosine (RTF) .... dssm_cnn_v2/cnn_dssm_tf.py:63-77
  gamma / exp / p0 ........ dssm_cnn_v2/cnn_dssm_tf.py:164-176
  loss / optimizer ........ dssm_cnn_v2/cnn_dssm_tf.py:182
"""
from __future__ import annotations

import numpy as np

TINY32 = float(np.finfo(np.float32).tiny)  # cnn_dssm_tf.py:70-72
BCE_EPS = 1e-7                             # Keras-1 backend epsilon


# --------------------------------------------------------------------------
# activations
# --------------------------------------------------------------------------
def act_fwd(x, act):
    if act == "relu":
        return np.maximum(x, 0.0)
    if act == "tanh":
        return np.tanh(x)
    if act == "linear":
        return x
    raise ValueError(act)


def act_grad_from_output(y, act):
    """d act / d pre-activation, expressed through the activation OUTPUT y."""
    if act == "relu":
        return (y > 0).astype(y.dtype)
    if act == "tanh":
        return 1.0 - y * y
    if act == "linear":
        return np.ones_like(y)
    raise ValueError(act)


# --------------------------------------------------------------------------
# tower forward  (cnn_dssm_tf.py:81-127)
# --------------------------------------------------------------------------
def conv_pool_fwd(X, W, b, act="relu"):
    """Convolution1D(border_mode='valid') + global max over time.

    X (S, L, D), W (k, D, F) in TF cross-correlation order, b (F,).
    Returns pooled (S, F) and argmax (S, F) int32.  The arg-max is taken over
    the raw accumulator (before bias/activation), ties -> lowest t; for relu a
    column whose pooled activation is 0 reports position 0 (every position ties
    at 0 after the activation, cnn_dssm_tf.py:88,92).
    """
    S, L, D = X.shape
    k, D2, F = W.shape
    assert D == D2 and L >= k
    T = L - k + 1
    Y = np.zeros((S, T, F), dtype=X.dtype)
    for j in range(k):
        Y += X[:, j:j + T, :] @ W[j]
    arg = np.argmax(Y, axis=1).astype(np.int32)          # first index on ties
    mx = np.take_along_axis(Y, arg[:, None, :].astype(np.int64), axis=1)[:, 0, :]
    pooled = act_fwd(mx + b[None, :], act)
    if act == "relu":
        arg = np.where(pooled > 0, arg, 0).astype(np.int32)
    return pooled, arg


def tower_fwd(tok, p, act="relu", mask=None, keep=1.0, post_mask=None, post_keep=1.0):
    """tok (S, L) int -> dict with X, pooled, argmax, sem.

    p: {'E': (V1, D), 'convW': [(k, D, F)], 'convb': [(F,)], 'Wd': (P, H), 'bd': (H,)}
    mask: optional (S, L, D) 0/1 keep-mask of Dropout(1-keep); X = E[tok]*mask/keep.
    post_mask: optional (S, H) keep-mask of the v1 model's second Dropout, which sits between Dense and the
    activation (dssm_cnn/cnn_dssm.py:148-150: Dense -> Dropout(dropout_prob[1]) -> relu).  relu commutes with a
    non-negative scaling, so relu(pre * m / keep) = relu(pre) * m / keep, which is what is computed here.
    """
    X = p["E"][tok]
    if mask is not None:
        X = X * mask / keep
    pooled, args = [], []
    for W, b in zip(p["convW"], p["convb"]):
        pl, ag = conv_pool_fwd(X, W, b, act)
        pooled.append(pl)
        args.append(ag)
    P = np.concatenate(pooled, axis=1)
    A = np.concatenate(args, axis=1)
    Z = act_fwd(P @ p["Wd"] + p["bd"][None, :], act)
    if post_mask is not None:
        assert act in ("relu", "linear"), "the second dropout commutes only with relu / linear"
        Z = Z * post_mask / post_keep
    return {"X": X, "pooled": P, "argmax": A, "sem": Z}


# --------------------------------------------------------------------------
# similarity head + loss  (cnn_dssm_tf.py:63-77,159-182)
# --------------------------------------------------------------------------
def _norm(x):
    return np.sqrt(np.maximum(np.sum(x * x, axis=-1), TINY32))


def sim_loss_fwd(zq, zd, gamma=10.0):
    """zq (B, H), zd (B, 1+J, H) (positive first).  Returns dict."""
    nq = _norm(zq)                       # (B,)
    nd = _norm(zd)                       # (B, 1+J)
    dot = np.einsum("bh,bjh->bj", zq, zd)
    raw = dot / (nq[:, None] * nd)
    cos = np.clip(raw, 0.0, 1.0)
    e = np.exp(gamma * cos)              # no max-subtraction (cnn_dssm_tf.py:171)
    prob = e / np.sum(e, axis=1, keepdims=True)
    p0 = prob[:, 0]
    ph = np.clip(p0, BCE_EPS, 1.0 - BCE_EPS)
    loss_rows = -np.log(ph)
    return {
        "cos": cos, "raw": raw, "prob": prob, "p0": p0,
        "loss": float(np.mean(loss_rows)),
        "acc": float(np.mean((p0 > 0.5).astype(np.float64))),
        "nq": nq, "nd": nd,
    }


def sim_loss_bwd(zq, zd, fw, gamma=10.0, batch_total=None):
    """Gradient of mean_b(-log clip(p0)) wrt zq, zd.  batch_total overrides the
    1/B scale (data-parallel: the global batch)."""
    B = zq.shape[0]
    Bt = float(batch_total if batch_total is not None else B)
    prob, raw, cos, nq, nd = fw["prob"], fw["raw"], fw["cos"], fw["nq"], fw["nd"]
    p0 = prob[:, 0]
    live = ((p0 >= BCE_EPS) & (p0 <= 1.0 - BCE_EPS)).astype(zq.dtype)
    ds = prob.copy()
    ds[:, 0] -= 1.0
    ds *= (live / Bt)[:, None]           # dL/d(gamma*cos)
    dc = gamma * ds
    dc = dc * ((raw >= 0.0) & (raw <= 1.0))   # inclusive clip gradient
    sq = np.sum(zq * zq, axis=-1)
    sd = np.sum(zd * zd, axis=-1)
    fq = (sq > TINY32).astype(zq.dtype)       # floor active -> norm is constant
    fd = (sd > TINY32).astype(zq.dtype)
    inv = 1.0 / (nq[:, None] * nd)            # (B, 1+J)
    # d raw / d zq = zd/(nq nd) - raw * zq / nq^2
    dzq = np.einsum("bj,bjh->bh", dc * inv, zd) \
        - (np.sum(dc * raw, axis=1) * fq / (nq * nq))[:, None] * zq
    dzd = (dc * inv)[:, :, None] * zq[:, None, :] \
        - (dc * raw * fd / (nd * nd))[:, :, None] * zd
    return dzq, dzd


# --------------------------------------------------------------------------
# tower backward
# --------------------------------------------------------------------------
def tower_bwd(tok, p, fw, dZ, act="relu", mask=None, keep=1.0, post_mask=None, post_keep=1.0):
    """Returns grads dict {'E','convW','convb','Wd','bd'} for one tower call."""
    X, P, A, Z = fw["X"], fw["pooled"], fw["argmax"], fw["sem"]
    S, L, D = X.shape
    if post_mask is not None:                    # Z = act(pre) * m / keep: dropped units get no gradient
        dZ = dZ * post_mask / post_keep
    dpre = dZ * act_grad_from_output(Z, act)
    g = {"Wd": P.T @ dpre, "bd": dpre.sum(axis=0)}
    dP = dpre @ p["Wd"].T
    G = dP * act_grad_from_output(P, act)           # (S, P) routed to the arg-max
    dX = np.zeros_like(X)
    g["convW"], g["convb"] = [], []
    off = 0
    sidx = np.arange(S)
    for W in p["convW"]:
        k, _, F = W.shape
        Gi = G[:, off:off + F]
        Ai = A[:, off:off + F].astype(np.int64)
        dW = np.zeros_like(W)
        for j in range(k):
            rows = X[sidx[:, None], Ai + j, :]               # (S, F, D)
            dW[j] = np.einsum("sfd,sf->df", rows, Gi)
            np.add.at(dX, (np.repeat(sidx, F), (Ai + j).ravel()),
                      (Gi[:, :, None] * W[j].T[None, :, :]).reshape(S * F, D))
        g["convW"].append(dW)
        g["convb"].append(Gi.sum(axis=0))
        off += F
    if mask is not None:
        dX = dX * mask / keep
    dE = np.zeros_like(p["E"])
    np.add.at(dE, tok.ravel(), dX.reshape(S * L, D))
    g["E"] = dE
    g["dX"] = dX
    return g


# --------------------------------------------------------------------------
# whole model: loss + grads for one batch  (cnn_dssm_tf.py:130-182)
# --------------------------------------------------------------------------
def model_loss_and_grads(q_tok, d_tok, pq, pd, act="relu", gamma=10.0,
                         mask_q=None, mask_d=None, keep=1.0, batch_total=None):
    """q_tok (B, Lq); d_tok (B, 1+J, Ld), positive first.  pq / pd tower params."""
    B, J1, Ld = d_tok.shape
    fq = tower_fwd(q_tok, pq, act, mask_q, keep)
    fd = tower_fwd(d_tok.reshape(B * J1, Ld), pd, act, mask_d, keep)
    zd = fd["sem"].reshape(B, J1, -1)
    head = sim_loss_fwd(fq["sem"], zd, gamma)
    dzq, dzd = sim_loss_bwd(fq["sem"], zd, head, gamma, batch_total)
    gq = tower_bwd(q_tok, pq, fq, dzq, act, mask_q, keep)
    gd = tower_bwd(d_tok.reshape(B * J1, Ld), pd, fd, dzd.reshape(B * J1, -1), act, mask_d, keep)
    return head, gq, gd, fq, fd


def adam_step(p, m, v, g, t, lr=1e-3, b1=0.9, b2=0.999, eps=1e-8):
    """Keras-1.0.x Adam (optimizers.py): t = iterations+1 (1-based)."""
    lr_t = lr * np.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t)
    m[...] = b1 * m + (1.0 - b1) * g
    v[...] = b2 * v + (1.0 - b2) * g * g
    p[...] = p - lr_t * m / (np.sqrt(v) + eps)


def flat_params(p):
    """Trainable tensors of a tower in a fixed order (name, array)."""
    out = [("E", p["E"])]
    for i, (W, b) in enumerate(zip(p["convW"], p["convb"])):
        out += [("convW%d" % i, W), ("convb%d" % i, b)]
    out += [("Wd", p["Wd"]), ("bd", p["bd"])]
    return out


def flat_grads(g):
    out = [("E", g["E"])]
    for i, (W, b) in enumerate(zip(g["convW"], g["convb"])):
        out += [("convW%d" % i, W), ("convb%d" % i, b)]
    out += [("Wd", g["Wd"]), ("bd", g["bd"])]
    return out


# --------------------------------------------------------------------------
# retrieval (no counterpart script in the reference; composition of RTF)
# --------------------------------------------------------------------------
def normalise_rows(z):
    return z / _norm(z)[:, None]


def cos_topk(qn, cn, k, id_base=0):
    """qn (Q, H), cn (N, H) both L2-normalised.  Order: score desc, id asc."""
    scores = qn @ cn.T
    N = cn.shape[0]
    k = min(k, N)
    ids = np.empty((qn.shape[0], k), dtype=np.int64)
    sc = np.empty((qn.shape[0], k), dtype=scores.dtype)
    for i in range(qn.shape[0]):
        order = np.argsort(-scores[i], kind="stable")[:k]
        ids[i] = order + id_base
        sc[i] = scores[i][order]
    return sc, ids


def topk_merge(score_parts, id_parts, k):
    """Merge per-shard (Q, k_i) lists; order (score desc, id asc)."""
    sc = np.concatenate(score_parts, axis=1)
    ids = np.concatenate(id_parts, axis=1)
    Q = sc.shape[0]
    k = min(k, sc.shape[1])
    out_s = np.empty((Q, k), dtype=sc.dtype)
    out_i = np.empty((Q, k), dtype=np.int64)
    for i in range(Q):
        order = np.lexsort((ids[i], -sc[i]))[:k]
        out_s[i] = sc[i][order]
        out_i[i] = ids[i][order]
    return out_s, out_i


def model_v1_loss_and_grads(q_tok, d_tok, pq, pds, act="relu", gamma=10.0, masks=None, keep=1.0,
                            post_masks=None, post_keep=1.0):
    """The v1 model (dssm_cnn/cnn_dssm.py:153-185): FIVE independent towers -- the query, the positive document and
    every negative document each have their own embedding table and weights (:160-164) -- and a second dropout after
    the Dense layer (:149).  d_tok (B, 1+J, Ld); pds: list of 1+J parameter dicts; masks / post_masks: optional lists
    [query, doc_0 .. doc_J] of (B, L, D) / (B, H) keep masks.  Returns head, gq, [gd_j], fq, [fd_j]."""
    B, J1, Ld = d_tok.shape
    m = masks or [None] * (1 + J1)
    pm = post_masks or [None] * (1 + J1)
    fq = tower_fwd(q_tok, pq, act, m[0], keep, pm[0], post_keep)
    fds = [tower_fwd(d_tok[:, j], pds[j], act, m[1 + j], keep, pm[1 + j], post_keep) for j in range(J1)]
    zd = np.stack([f["sem"] for f in fds], axis=1)
    head = sim_loss_fwd(fq["sem"], zd, gamma)
    dzq, dzd = sim_loss_bwd(fq["sem"], zd, head, gamma)
    gq = tower_bwd(q_tok, pq, fq, dzq, act, m[0], keep, pm[0], post_keep)
    gds = [tower_bwd(d_tok[:, j], pds[j], fds[j], dzd[:, j], act, m[1 + j], keep, pm[1 + j], post_keep) for j in range(J1)]
    return head, gq, gds, fq, fds


# --------------------------------------------------------------------------
# parameter initialisation used by tests / bench (synthetic)
# --------------------------------------------------------------------------
def init_tower(rng, V1, D, windows, F, H, dtype=np.float64):
    """E ~ U(-0.25, 0.25) (data_helpers.py:62), conv W ~ U(+-0.05) (Keras-1
    'uniform'), dense glorot-uniform, biases 0.  Last table row zero
    (data_helpers.py:56)."""
    E = rng.uniform(-0.25, 0.25, size=(V1, D))
    E[-1] = 0.0
    P = len(windows) * F
    lim = np.sqrt(6.0 / (P + H))
    return {
        "E": E.astype(dtype),
        "convW": [rng.uniform(-0.05, 0.05, size=(k, D, F)).astype(dtype) for k in windows],
        "convb": [np.zeros(F, dtype=dtype) for _ in windows],
        "Wd": rng.uniform(-lim, lim, size=(P, H)).astype(dtype),
        "bd": np.zeros(H, dtype=dtype),
    }
