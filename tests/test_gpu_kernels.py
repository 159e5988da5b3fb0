"""Per-kernel parity on the GPU, through the C ABI, against the CPU oracle (-m gpu)."""
import ctypes as C
import numpy as np
import pytest
import torch

from oracle import clsm_oracle as O
from tests.gpu_util import DEV, P, ST, bf16_round, call, decode_keys, dev, lib, make_X, pack_w, rup
from dnn_page_vectors_b200 import _lib

pytestmark = pytest.mark.gpu
F32, BF16 = _lib.F32, _lib.BF16


# ------------------------------------------------------------------ gather / dropout
@pytest.mark.parametrize("bf16", [True, False])
@pytest.mark.parametrize("D", [100, 300, 64])
def test_gather_bit_exact(bf16, D):
    rng = np.random.RandomState(0)
    V1, n_tok = 1000, 777
    Dpad = rup(D, 64)
    E = rng.uniform(-0.25, 0.25, size=(V1, D)).astype(np.float32)
    tok = rng.randint(0, V1, size=n_tok).astype(np.int32)
    tok[:5] = [0, V1 - 1, 0, V1 - 1, 7]
    if bf16:
        table = torch.zeros((V1, Dpad), dtype=torch.bfloat16, device=DEV)
        E_d = dev(E)
        call("clsm_table_to_bf16", P(E_d), V1, D, P(table), Dpad, ST())
        ld, ref = Dpad, bf16_round(E)
    else:
        table, ld, ref = dev(E), D, E
    X = torch.full((n_tok, Dpad), 7.0, dtype=table.dtype, device=DEV)
    tok_d = dev(tok)
    call("clsm_emb_gather", P(tok_d), n_tok, P(table), V1, D, ld, BF16 if bf16 else F32, P(X), Dpad, 1.0, 0, ST())
    got = X.float().cpu().numpy()
    np.testing.assert_array_equal(got[:, :D], ref[tok])
    assert (got[:, D:] == 0).all()


def test_gather_host_ids_entry_point():
    rng = np.random.RandomState(1)
    V1, D, n_tok = 50, 64, 33
    E = rng.randn(V1, D).astype(np.float32)
    tok_h = torch.as_tensor(rng.randint(0, V1, size=n_tok).astype(np.int32)).pin_memory()
    tok_d = torch.empty(n_tok, dtype=torch.int32, device=DEV)
    X = torch.empty((n_tok, 64), dtype=torch.float32, device=DEV)
    E_d = dev(E)
    call("clsm_emb_gather_h", tok_h.data_ptr(), P(tok_d), n_tok, P(E_d), V1, D, D, F32, P(X), 64, 1.0, 0, ST())
    np.testing.assert_array_equal(X.cpu().numpy(), E[tok_h.numpy()])


@pytest.mark.parametrize("bf16", [True, False])
def test_gather_dropout_matches_exported_mask(bf16):
    rng = np.random.RandomState(2)
    V1, D, n_tok, keep, seed = 300, 100, 4096, 0.75, 123456789
    Dpad = rup(D, 64)
    E = rng.uniform(-0.25, 0.25, size=(V1, D)).astype(np.float32)
    tok = rng.randint(0, V1, size=n_tok).astype(np.int32)
    mask = torch.empty((n_tok, D), dtype=torch.uint8, device=DEV)
    call("clsm_dropout_mask", n_tok, D, keep, seed, P(mask), ST())
    m = mask.cpu().numpy().astype(np.float32)
    assert abs(m.mean() - keep) < 0.01
    assert abs(m[:, 0].mean() - keep) < 0.05 and abs(m[0].mean() - keep) < 0.15
    if bf16:
        table = torch.zeros((V1, Dpad), dtype=torch.bfloat16, device=DEV)
        E_d = dev(E)
        call("clsm_table_to_bf16", P(E_d), V1, D, P(table), Dpad, ST())
        src, ld = bf16_round(E), Dpad
    else:
        table, src, ld = dev(E), E, D
    X = torch.empty((n_tok, Dpad), dtype=table.dtype, device=DEV)
    tok_d = dev(tok)
    call("clsm_emb_gather", P(tok_d), n_tok, P(table), V1, D, ld, BF16 if bf16 else F32, P(X), Dpad, keep, seed, ST())
    want = src[tok] * m * np.float32(1.0 / keep)
    if bf16:
        want = bf16_round(want)
    np.testing.assert_array_equal(X.float().cpu().numpy()[:, :D], want)
    # a different seed gives a different mask
    mask2 = torch.empty_like(mask)
    call("clsm_dropout_mask", n_tok, D, keep, seed + 1, P(mask2), ST())
    assert (mask2 != mask).float().mean().item() > 0.2


# ------------------------------------------------------------------ conv + pool (CUDA-core path)
def conv_case(seed, S, L, D, k, F, exact):
    rng = np.random.RandomState(seed)
    if exact:       # multiples of 1/8 in a small range: every product and sum is exact in fp32
        X = rng.randint(-4, 5, size=(S, L, D)) / 8.0
        W = rng.randint(-4, 5, size=(k, D, F)) / 8.0
    else:
        X = rng.uniform(-0.25, 0.25, size=(S, L, D))
        W = rng.uniform(-0.05, 0.05, size=(k, D, F))
    return X, W


def run_conv(X, W, bf16, impl, col0=0, ldk=None):
    S, L, D = X.shape
    k, _, F = W.shape
    Dpad, Fpad = rup(D, 64), rup(F, 16)
    Xd, Wt = make_X(X, Dpad, bf16), pack_w(W, Dpad, Fpad, bf16)
    ldk = ldk or (col0 + F)
    keys = torch.full((S, ldk), -1, dtype=torch.int64, device=DEV)
    call("clsm_conv_pool_fwd", P(Xd), BF16 if bf16 else F32, S, L, Dpad, P(Wt), k, F, Fpad, P(keys), ldk, col0,
         _lib.IMPL[impl], ST())
    torch.cuda.synchronize()
    mx, am = decode_keys(keys[:, col0:col0 + F])
    return mx, am, keys


def oracle_conv(X, W):
    S, L, D = X.shape
    k = W.shape[0]
    T = L - k + 1
    Y = np.zeros((S, T, W.shape[2]))
    for j in range(k):
        Y += X[:, j:j + T, :] @ W[j]
    return Y.max(axis=1), Y.argmax(axis=1), Y


@pytest.mark.parametrize("bf16", [False, True])
@pytest.mark.parametrize("shape", [(3, 9, 10, 2, 5), (5, 20, 100, 3, 150), (2, 100, 300, 3, 300),
                                   (4, 7, 64, 7, 33), (1, 3, 16, 3, 1), (7, 33, 100, 1, 70), (3, 130, 100, 4, 150)])
def test_conv_pool_simt_exact_arithmetic(bf16, shape):
    S, L, D, k, F = shape
    X, W = conv_case(10, S, L, D, k, F, exact=True)
    mx, am, _ = run_conv(X, W, bf16, "simt")
    omx, oam, _ = oracle_conv(X, W)
    np.testing.assert_array_equal(mx, omx.astype(np.float32))     # bit-exact values
    np.testing.assert_array_equal(am, oam)                        # bit-exact arg-max, ties -> lowest t


@pytest.mark.parametrize("bf16", [False, True])
def test_conv_pool_simt_random_and_ties(bf16):
    S, L, D, k, F = 6, 100, 300, 3, 300
    X, W = conv_case(11, S, L, D, k, F, exact=False)
    X[2, 40:] = X[2, 40]                     # right-padded document: identical windows from t=40 on
    X[3, :] = X[3, 0]                        # all-PAD document
    if bf16:
        X, W = bf16_round(X).astype(np.float64), bf16_round(W).astype(np.float64)
    mx, am, _ = run_conv(X, W, bf16, "simt")
    omx, oam, Y = oracle_conv(X, W)
    np.testing.assert_allclose(mx, omx, rtol=1e-5, atol=1e-6)
    assert (am[3] == 0).all()                # ties -> lowest index
    # arg-max may differ from the f64 oracle only where the top two are closer than fp32 noise
    bad = np.argwhere(am != oam)
    for s, f in bad:
        assert abs(Y[s, am[s, f], f] - Y[s, oam[s, f], f]) < 1e-5


def test_conv_pool_column_offset_and_clearing():
    X, W = conv_case(12, 4, 12, 64, 2, 20, exact=True)
    mx, am, keys = run_conv(X, W, False, "simt", col0=7, ldk=40)
    omx, oam, _ = oracle_conv(X, W)
    np.testing.assert_array_equal(mx, omx.astype(np.float32))
    k = keys.cpu().numpy()
    assert (k[:, :7] == -1).all() and (k[:, 27:] == -1).all()     # other columns untouched


@pytest.mark.parametrize("act", ["relu", "tanh", "linear"])
def test_pool_finalize(act):
    rng = np.random.RandomState(3)
    X, W = conv_case(13, 5, 15, 64, 3, 24, exact=False)
    b = rng.uniform(-0.2, 0.2, size=24)
    _, _, keys = run_conv(X, W, False, "simt")
    pooled = torch.empty((5, 24), dtype=torch.float32, device=DEV)
    am = torch.empty((5, 24), dtype=torch.int32, device=DEV)
    b_d = dev(b, torch.float32)
    call("clsm_pool_finalize", P(keys), 5, 24, 24, P(b_d), _lib.ACT[act], P(pooled), P(am), 24, ST())
    op, oa = O.conv_pool_fwd(X, W, b, act)
    np.testing.assert_allclose(pooled.cpu().numpy(), op, rtol=1e-5, atol=1e-6)
    np.testing.assert_array_equal(am.cpu().numpy(), oa)


# ------------------------------------------------------------------ dense
@pytest.mark.parametrize("S,Pn,H", [(5, 7, 3), (130, 300, 128), (257, 600, 150), (1, 1, 1)])
@pytest.mark.parametrize("act", ["relu", "tanh"])
def test_dense_fwd_bwd(S, Pn, H, act):
    rng = np.random.RandomState(4)
    pooled = rng.uniform(0, 1, size=(S, Pn)) * (rng.rand(S, Pn) > 0.3)
    Wd = rng.uniform(-0.2, 0.2, size=(Pn, H))
    bd = rng.uniform(-0.1, 0.1, size=H)
    dsem = rng.randn(S, H)
    sem = torch.empty((S, H), dtype=torch.float32, device=DEV)
    pd_, wd_, bd_ = dev(pooled, torch.float32), dev(Wd, torch.float32), dev(bd, torch.float32)
    call("clsm_dense_fwd", P(pd_), S, Pn, Pn, P(wd_), P(bd_), H, _lib.ACT[act], P(sem), ST())
    Z = O.act_fwd(pooled @ Wd + bd, act)
    np.testing.assert_allclose(sem.cpu().numpy(), Z, rtol=2e-5, atol=2e-6)

    ws_bytes = lib.clsm_dense_bwd_workspace_bytes(S, Pn, H)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV)
    dWd = torch.empty((Pn, H), dtype=torch.float32, device=DEV)
    dbd = torch.empty(H, dtype=torch.float32, device=DEV)
    G = torch.empty((S, Pn), dtype=torch.float32, device=DEV)
    dsem_d = dev(dsem, torch.float32)
    call("clsm_dense_bwd", P(pd_), P(sem), P(dsem_d), S, Pn, Pn, P(wd_), H, _lib.ACT[act],
         P(dWd), P(dbd), P(G), P(ws), ws_bytes, ST())
    semn = sem.cpu().numpy().astype(np.float64)
    dpre = dsem * O.act_grad_from_output(semn, act)
    np.testing.assert_allclose(dWd.cpu().numpy(), pooled.T @ dpre, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(dbd.cpu().numpy(), dpre.sum(0), rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(G.cpu().numpy(), (dpre @ Wd.T) * O.act_grad_from_output(pooled, act), rtol=1e-4, atol=1e-5)
    with pytest.raises(_lib.ClsmError):
        call("clsm_dense_bwd", P(pd_), P(sem), P(sem), S, Pn, Pn, P(wd_), H, _lib.ACT[act], P(dWd), P(dbd), P(G), P(ws), 8, ST())


@pytest.mark.parametrize("S,Pn,H", [(5, 7, 3), (130, 300, 128), (257, 600, 150), (1000, 1200, 128), (1, 1, 1)])
@pytest.mark.parametrize("act", ["relu", "tanh"])
def test_dense_fwd_bwd_tf32(S, Pn, H, act):
    """act | CLSM_MATH_TF32: the same GEMMs on the tensor cores with TF32 operands.  Each operand is rounded to a
    10-bit mantissa (relative 2^-11), so every result is within 2 * 2^-11 (+ fp32 accumulation) of the sum of |terms|."""
    rng = np.random.RandomState(4)
    pooled = rng.uniform(0, 1, size=(S, Pn)) * (rng.rand(S, Pn) > 0.3)
    Wd = rng.uniform(-0.2, 0.2, size=(Pn, H))
    bd = rng.uniform(-0.1, 0.1, size=H)
    dsem = rng.randn(S, H)
    flag = _lib.ACT[act] | _lib.MATH_TF32
    tol = 1.2e-3
    sem = torch.empty((S, H), dtype=torch.float32, device=DEV)
    pd_, wd_, bd_ = dev(pooled, torch.float32), dev(Wd, torch.float32), dev(bd, torch.float32)
    call("clsm_dense_fwd", P(pd_), S, Pn, Pn, P(wd_), P(bd_), H, flag, P(sem), ST())
    pre = pooled @ Wd + bd
    mag = np.abs(pooled) @ np.abs(Wd) + np.abs(bd)
    got = sem.cpu().numpy()
    # the activations used here are 1-Lipschitz: an error of e before them is at most e after
    assert (np.abs(got - O.act_fwd(pre, act)) <= tol * mag + 1e-6).all()

    ws_bytes = lib.clsm_dense_bwd_workspace_bytes(S, Pn, H)
    ws = torch.full((ws_bytes,), 0xA5, dtype=torch.uint8, device=DEV)
    dWd = torch.empty((Pn, H), dtype=torch.float32, device=DEV)
    dbd = torch.empty(H, dtype=torch.float32, device=DEV)
    G = torch.empty((S, Pn), dtype=torch.float32, device=DEV)
    dsem_d = dev(dsem, torch.float32)
    call("clsm_dense_bwd", P(pd_), P(sem), P(dsem_d), S, Pn, Pn, P(wd_), H, flag, P(dWd), P(dbd), P(G), P(ws), ws_bytes, ST())
    semn = got.astype(np.float64)
    dpre = dsem * O.act_grad_from_output(semn, act)
    assert (np.abs(dWd.cpu().numpy() - pooled.T @ dpre) <= tol * (np.abs(pooled).T @ np.abs(dpre)) + 1e-5).all()
    np.testing.assert_allclose(dbd.cpu().numpy(), dpre.sum(0), rtol=1e-4, atol=1e-5)          # column sum stays fp32
    gate = O.act_grad_from_output(pooled, act)
    assert (np.abs(G.cpu().numpy() - (dpre @ Wd.T) * gate) <= tol * (np.abs(dpre) @ np.abs(Wd).T) * np.abs(gate) + 1e-5).all()


# ------------------------------------------------------------------ cosine / softmax / loss
@pytest.mark.parametrize("B,J,H", [(7, 3, 150), (64, 4, 128), (5, 50, 128), (3, 0, 8)])
def test_sim_loss_fwd_bwd(B, J, H):
    rng = np.random.RandomState(5)
    zq = np.maximum(rng.randn(B, H), 0)          # post-ReLU vectors
    zd = np.maximum(rng.randn(B, 1 + J, H), 0)
    zq[0] = 0.0                                   # dead query -> cosine exactly 0
    if J >= 1:
        zd[1, 1] = 0.0
        zd[2, 0] = zq[2] * 3.0                    # cosine exactly at the upper clip
    fw = O.sim_loss_fwd(zq, zd, 10.0)
    dq, dd = O.sim_loss_bwd(zq, zd, fw, 10.0, batch_total=2 * B)
    q_, d_ = dev(zq, torch.float32), dev(zd.reshape(B * (1 + J), H), torch.float32)
    cosv = torch.empty((B, 1 + J), dtype=torch.float32, device=DEV)
    prob = torch.empty_like(cosv)
    la = torch.empty(2, dtype=torch.float32, device=DEV)
    gq = torch.empty((B, H), dtype=torch.float32, device=DEV)
    gd = torch.empty((B * (1 + J), H), dtype=torch.float32, device=DEV)
    call("clsm_sim_loss", P(q_), P(d_), B, J, H, 10.0, 2 * B, P(cosv), P(prob), P(la), P(gq), P(gd), ST())
    c = cosv.cpu().numpy()
    assert (c[0] == 0.0).all() and np.isfinite(gq.cpu().numpy()).all()
    np.testing.assert_allclose(c, fw["cos"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(prob.cpu().numpy(), fw["prob"], rtol=1e-4, atol=1e-7)
    np.testing.assert_allclose(la.cpu().numpy(), [fw["loss"], fw["acc"]], rtol=1e-5, atol=2e-7)   # 1-1e-7 is not an fp32 number
    # rows whose cosine sits exactly on a clip edge may legitimately differ in the gate
    edge = (np.abs(fw["raw"] - 1.0) < 1e-6) | (np.abs(fw["raw"]) < 1e-9)
    keep_rows = ~edge.any(axis=1)
    np.testing.assert_allclose(gq.cpu().numpy()[keep_rows], dq[keep_rows], rtol=2e-4, atol=1e-7)
    np.testing.assert_allclose(gd.cpu().numpy().reshape(B, 1 + J, H)[keep_rows], dd[keep_rows], rtol=2e-4, atol=1e-7)


# ------------------------------------------------------------------ conv backward
@pytest.mark.parametrize("bf16", [False, True])
@pytest.mark.parametrize("shape", [(5, 20, 100, 3, 150), (9, 12, 64, 2, 20), (3, 100, 300, 3, 300), (4, 30, 100, 5, 130),
                                   (6, 9, 40, 1, 17), (2, 1100, 64, 3, 40), (3, 500, 100, 4, 150), (70, 100, 300, 3, 300),
                                   # token-chunk counts that do not divide over the dW token splits (48 chunks / 37 CTAs,
                                   # 14 / 49): the split count must shrink so that no CTA is left without work
                                   (128, 24, 100, 4, 150), (37, 24, 100, 3, 150), (11, 200, 100, 6, 64)])
def test_conv_pool_bwd(bf16, shape):
    S, L, D, k, F = shape
    rng = np.random.RandomState(6)
    X, W = conv_case(14, S, L, D, k, F, exact=False)
    if bf16:
        X, W = bf16_round(X).astype(np.float64), bf16_round(W).astype(np.float64)
    Dpad, Fpad = rup(D, 64), rup(F, 16)
    Pn, col0 = F + 9, 4
    G = np.zeros((S, Pn)); A = np.zeros((S, Pn), dtype=np.int32)
    G[:, col0:col0 + F] = rng.randn(S, F) * (rng.rand(S, F) > 0.4)      # ~40 % dead (ReLU)
    A[:, col0:col0 + F] = rng.randint(0, L - k + 1, size=(S, F))
    A[0, col0:col0 + F] = 0; A[1, col0:col0 + F] = L - k                  # edges
    Xd, Wt = make_X(X, Dpad, bf16), pack_w(W, Dpad, Fpad, bf16)
    ws_bytes = lib.clsm_conv_bwd_workspace_bytes(S, L, k, D, F)
    ws = torch.full((ws_bytes,), 0xA5, dtype=torch.uint8, device=DEV)       # poisoned workspace
    dW = torch.empty((k, D, F), dtype=torch.float32, device=DEV)
    db = torch.empty(F, dtype=torch.float32, device=DEV)
    dX = torch.full((S * L, Dpad), 3.0, dtype=torch.float32, device=DEV)
    G_d, A_d = dev(G, torch.float32), dev(A)        # keep the device buffers alive across the calls
    args = (P(Xd), BF16 if bf16 else F32, S, L, D, Dpad, P(Wt), k, F, Fpad, P(G_d), P(A_d), Pn, col0,
            P(dW), P(db), P(dX))
    call("clsm_conv_pool_bwd", *args, 1, P(ws), ws_bytes, ST())
    Gi, Ai = G[:, col0:col0 + F], A[:, col0:col0 + F].astype(np.int64)
    odW = np.zeros_like(W); odX = np.zeros_like(X)
    sidx = np.arange(S)
    for j in range(k):
        odW[j] = np.einsum("sfd,sf->df", X[sidx[:, None], Ai + j, :], Gi)
        np.add.at(odX, (np.repeat(sidx, F), (Ai + j).ravel()), (Gi[:, :, None] * W[j].T[None]).reshape(S * F, D))
    if bf16:     # tensor-core dW: G enters as bf16 -> error bounded by 2^-8 of the sum of |terms|
        magW = np.zeros_like(W)
        for j in range(k):
            magW[j] = np.einsum("sfd,sf->df", np.abs(X[sidx[:, None], Ai + j, :]), np.abs(Gi))
        assert (np.abs(dW.cpu().numpy() - odW) <= 8e-3 * magW + 1e-6).all()
    else:
        np.testing.assert_allclose(dW.cpu().numpy(), odW, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(db.cpu().numpy(), Gi.sum(0), rtol=1e-4, atol=1e-5)
    got = dX.cpu().numpy()
    # bf16 mode: dX comes from the tensor core with G rounded to bf16 (8-bit mantissa) -> 2e-2 class tolerance,
    # checked per element against the sum of |terms|; fp32 mode is the exact CUDA-core path
    if bf16:
        mag = np.zeros_like(odX)
        for j in range(k):
            np.add.at(mag, (np.repeat(sidx, F), (Ai + j).ravel()), np.abs(Gi[:, :, None] * W[j].T[None]).reshape(S * F, D))
        assert (np.abs(got[:, :D] - odX.reshape(S * L, D)) <= 8e-3 * mag.reshape(S * L, D) + 1e-6).all()
    else:
        np.testing.assert_allclose(got[:, :D], odX.reshape(S * L, D), rtol=1e-4, atol=1e-5)
    assert (got[:, D:] == 0).all()
    # second window accumulates into dX
    call("clsm_conv_pool_bwd", *args, 0, P(ws), ws_bytes, ST())
    got2 = dX.cpu().numpy()[:, :D]
    if bf16:
        assert (np.abs(got2 - 2 * odX.reshape(S * L, D)) <= 1.6e-2 * mag.reshape(S * L, D) + 2e-6).all()
    else:
        np.testing.assert_allclose(got2, 2 * odX.reshape(S * L, D), rtol=1e-4, atol=1e-5)


# ------------------------------------------------------------------ embedding scatter
def run_scatter(tok, dX, D, V1, keep=1.0, seed=0):
    n_tok, Dpad = dX.shape
    ws_bytes = lib.clsm_emb_scatter_workspace_bytes(n_tok, D)
    ws = torch.full((ws_bytes,), 0xA5, dtype=torch.uint8, device=DEV)       # poisoned workspace
    uniq = torch.full((n_tok,), -7, dtype=torch.int32, device=DEV)
    nu = torch.zeros(1, dtype=torch.int32, device=DEV)
    rows = torch.full((min(n_tok, V1), Dpad), 9.0, dtype=torch.float32, device=DEV)
    roi = torch.empty(V1, dtype=torch.int32, device=DEV)
    call("clsm_emb_scatter", P(tok), n_tok, P(dX), Dpad, D, keep, seed, V1, P(uniq), P(nu), P(rows), P(roi),
         P(ws), ws_bytes, ST())
    torch.cuda.synchronize()
    n = int(nu.item())
    return uniq[:n].cpu().numpy(), rows[:n].cpu().numpy(), roi.cpu().numpy()


@pytest.mark.parametrize("n_tok,V1,D,skew", [(1, 10, 100, 0.0), (63, 10, 100, 0.5), (64, 1000, 100, 0.0), (65, 5, 64, 0.9),
                                             (5000, 200, 100, 0.6), (70000, 50000, 300, 0.4), (4097, 3, 100, 0.0),
                                             (140000, 10, 64, 0.97)])
def test_emb_scatter_matches_add_at_and_is_deterministic(n_tok, V1, D, skew):
    rng = np.random.RandomState(7)
    Dpad = rup(D, 64)
    tok = rng.randint(0, V1, size=n_tok).astype(np.int32)
    tok[rng.rand(n_tok) < skew] = V1 // 2                 # PAD-like heavy id
    dXn = np.zeros((n_tok, Dpad), dtype=np.float32)
    dXn[:, :D] = rng.randn(n_tok, D)
    tok_d, dX_d = dev(tok), dev(dXn)
    u1, r1, roi = run_scatter(tok_d, dX_d, D, V1)
    want = np.zeros((V1, D)); np.add.at(want, tok, dXn[:, :D].astype(np.float64))
    ids = np.unique(tok)
    np.testing.assert_array_equal(u1, ids)                 # sorted unique ids
    np.testing.assert_allclose(r1[:, :D], want[ids], rtol=2e-5, atol=2e-4 if n_tok > 50000 else 2e-5)
    assert (r1[:, D:] == 0).all()
    exp_roi = np.full(V1, -1); exp_roi[ids] = np.arange(len(ids))
    np.testing.assert_array_equal(roi, exp_roi)
    u2, r2, _ = run_scatter(tok_d, dX_d, D, V1)
    assert np.array_equal(r1, r2) and np.array_equal(u1, u2)      # bit-identical rerun


def test_emb_scatter_reapplies_dropout_mask():
    rng = np.random.RandomState(8)
    n_tok, V1, D, keep, seed = 3000, 40, 100, 0.75, 99
    Dpad = rup(D, 64)
    tok = rng.randint(0, V1, size=n_tok).astype(np.int32)
    dXn = np.zeros((n_tok, Dpad), dtype=np.float32); dXn[:, :D] = rng.randn(n_tok, D)
    mask = torch.empty((n_tok, D), dtype=torch.uint8, device=DEV)
    call("clsm_dropout_mask", n_tok, D, keep, seed, P(mask), ST())
    m = mask.cpu().numpy().astype(np.float64)
    u, r, _ = run_scatter(dev(tok), dev(dXn), D, V1, keep, seed)
    want = np.zeros((V1, D)); np.add.at(want, tok, dXn[:, :D] * m / keep)
    np.testing.assert_allclose(r[:, :D], want[u], rtol=2e-5, atol=2e-5)


def test_mask_bits_bf16_dx_and_split_scatter_equal_the_one_call_path():
    """The training path's variants: clsm_emb_gather_mask stores the keep decisions as bits, clsm_emb_scatter_plan /
    _reduce split the scatter and read those bits (no Philox in the backward), dX may arrive as bf16.  With fp32 dX
    the split path must be BIT-IDENTICAL to clsm_emb_scatter (same tree, same order); with bf16 dX it must equal the
    one-call path fed the bf16-rounded dX."""
    rng = np.random.RandomState(8)
    n_tok, V1, D, keep, seed = 70001, 300, 100, 0.75, 4242
    Dpad = rup(D, 64)
    E = rng.uniform(-0.25, 0.25, size=(V1, D)).astype(np.float32)
    tok = rng.randint(0, V1, size=n_tok).astype(np.int32)
    tok[rng.rand(n_tok) < 0.4] = 7
    tok_d, E_d = dev(tok), dev(E)
    table = torch.zeros((V1, Dpad), dtype=torch.bfloat16, device=DEV)
    call("clsm_table_to_bf16", P(E_d), V1, D, P(table), Dpad, ST())
    X1 = torch.empty((n_tok, Dpad), dtype=torch.bfloat16, device=DEV)
    X2 = torch.empty_like(X1)
    bits = torch.zeros((n_tok, Dpad // 16), dtype=torch.int16, device=DEV)
    call("clsm_emb_gather", P(tok_d), n_tok, P(table), V1, D, Dpad, BF16, P(X1), Dpad, keep, seed, ST())
    call("clsm_emb_gather_mask", P(tok_d), n_tok, P(table), V1, D, Dpad, BF16, P(X2), Dpad, keep, seed, P(bits), ST())
    assert torch.equal(X1, X2)
    mask = torch.empty((n_tok, D), dtype=torch.uint8, device=DEV)
    call("clsm_dropout_mask", n_tok, D, keep, seed, P(mask), ST())
    b = bits.cpu().numpy().view(np.uint16)
    unpacked = ((b[:, :, None] >> np.arange(16, dtype=np.uint16)[None, None, :]) & 1).reshape(n_tok, Dpad)[:, :D]
    np.testing.assert_array_equal(unpacked, mask.cpu().numpy())
    assert abs(lib.clsm_dropout_scale(0.75) - 4.0 / 3.0) < 1e-7 and abs(lib.clsm_dropout_scale(0.9) - 256.0 / 230.0) < 1e-7
    # the same gather straight from the fp32 master table: ONE rounding, same mask, same stored bits
    X3 = torch.empty_like(X1)
    bits3 = torch.zeros_like(bits)
    call("clsm_emb_gather_master", P(tok_d), n_tok, P(E_d), V1, D, P(X3), Dpad, keep, seed, P(bits3), ST())
    assert torch.equal(bits3, bits)
    want3 = bf16_round(E[tok] * mask.cpu().numpy().astype(np.float32) * np.float32(4.0 / 3.0))
    np.testing.assert_array_equal(X3.float().cpu().numpy()[:, :D], want3)
    assert (X3.float().cpu().numpy()[:, D:] == 0).all()
    call("clsm_emb_gather_master", P(tok_d), n_tok, P(E_d), V1, D, P(X3), Dpad, 1.0, 0, P(None), ST())
    np.testing.assert_array_equal(X3.float().cpu().numpy()[:, :D], bf16_round(E[tok]))
    # scatter: one call (Philox regenerated) vs plan + reduce (stored bits)
    dXn = np.zeros((n_tok, Dpad), dtype=np.float32); dXn[:, :D] = rng.randn(n_tok, D)
    dX_d = dev(dXn)
    u1, r1, roi1 = run_scatter(tok_d, dX_d, D, V1, keep, seed)

    def split(dX_t, code):
        wsb = lib.clsm_emb_scatter_workspace_bytes(n_tok, D)
        ws = torch.full((wsb,), 0xA5, dtype=torch.uint8, device=DEV)
        uniq = torch.full((n_tok,), -7, dtype=torch.int32, device=DEV)
        nu = torch.zeros(1, dtype=torch.int32, device=DEV)
        rows = torch.full((min(n_tok, V1), Dpad), 9.0, dtype=torch.float32, device=DEV)
        roi = torch.empty(V1, dtype=torch.int32, device=DEV)
        call("clsm_emb_scatter_plan", P(tok_d), n_tok, Dpad, D, V1, P(uniq), P(nu), P(roi), P(ws), wsb, ST())
        call("clsm_emb_scatter_reduce", n_tok, P(dX_t), code, Dpad, D, keep, seed, P(bits), P(rows), P(ws), wsb, ST())
        n = int(nu.item())
        return uniq[:n].cpu().numpy(), rows[:n].cpu().numpy(), roi.cpu().numpy()
    u2, r2, roi2 = split(dX_d, F32)
    assert np.array_equal(u1, u2) and np.array_equal(roi1, roi2) and np.array_equal(r1, r2)
    dX16 = dX_d.to(torch.bfloat16)
    u3, r3, _ = split(dX16, BF16)
    _, r4, _ = run_scatter(tok_d, dX16.float(), D, V1, keep, seed)
    assert np.array_equal(u1, u3) and np.array_equal(r3, r4)
    # id validation
    bad = torch.zeros(1, dtype=torch.int32, device=DEV)
    call("clsm_check_token_ids", P(tok_d), n_tok, V1, P(bad), ST())
    assert int(bad.item()) == 0
    tok2 = tok.copy(); tok2[5] = V1; tok2[77] = -1; tok2[n_tok - 1] = 1 << 30
    call("clsm_check_token_ids", P(dev(tok2)), n_tok, V1, P(bad), ST())
    assert int(bad.item()) == 3


def test_gather_scales_by_the_realised_keep_probability():
    """keep is realised in steps of 1/256: 0.9 -> 230/256.  Kept values are scaled by 256/230, not 1/0.9."""
    rng = np.random.RandomState(9)
    V1, D, n_tok, keep, seed = 50, 64, 2000, 0.9, 5
    E = rng.uniform(-1, 1, size=(V1, D)).astype(np.float32)
    tok = rng.randint(0, V1, size=n_tok).astype(np.int32)
    X = torch.empty((n_tok, 64), dtype=torch.float32, device=DEV)
    mask = torch.empty((n_tok, D), dtype=torch.uint8, device=DEV)
    E_d, tok_d = dev(E), dev(tok)
    call("clsm_emb_gather", P(tok_d), n_tok, P(E_d), V1, D, D, F32, P(X), 64, keep, seed, ST())
    call("clsm_dropout_mask", n_tok, D, keep, seed, P(mask), ST())
    m = mask.cpu().numpy().astype(np.float32)
    assert abs(m.mean() - 230.0 / 256.0) < 0.01
    np.testing.assert_array_equal(X.cpu().numpy(), E[tok] * m * np.float32(256.0 / 230.0))


@pytest.mark.parametrize("shape", [(70, 100, 300, 3, 300), (5, 1000, 100, 4, 150), (9, 12, 64, 2, 20)])
def test_conv_bwd_bf16_dx_rows_are_the_rounded_fp32_rows(shape):
    """clsm_conv_pool_bwd_ex with dx_dtype = bf16: the same accumulators, rounded once on the way out."""
    S, L, D, k, F = shape
    assert lib.clsm_conv_bwd_dx_bf16_ok(S, L, k, D, F) == 1
    rng = np.random.RandomState(16)
    X, W = conv_case(17, S, L, D, k, F, exact=False)
    Dpad, Fpad = rup(D, 64), rup(F, 16)
    G = (rng.randn(S, F) * (rng.rand(S, F) > 0.4)).astype(np.float32)
    A = rng.randint(0, L - k + 1, size=(S, F)).astype(np.int32)
    Xd, Wt = make_X(X, Dpad, True), pack_w(W, Dpad, Fpad, True)
    ws_bytes = lib.clsm_conv_bwd_workspace_bytes(S, L, k, D, F)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV)
    G_d, A_d = dev(G), dev(A)
    outs = []
    for code, dt in ((F32, torch.float32), (BF16, torch.bfloat16)):
        dW = torch.empty((k, D, F), dtype=torch.float32, device=DEV)
        db = torch.empty(F, dtype=torch.float32, device=DEV)
        dX = torch.full((S * L, Dpad), 3.0, dtype=dt, device=DEV)
        call("clsm_conv_pool_bwd_ex", P(Xd), BF16, S, L, D, Dpad, P(Wt), k, F, Fpad, P(G_d), P(A_d), F, 0,
             P(dW), P(db), P(dX), code, 1, P(ws), ws_bytes, ST())
        outs.append((dW, dX))
    assert torch.equal(outs[0][0], outs[1][0])
    assert torch.equal(outs[0][1].to(torch.bfloat16), outs[1][1])
    with pytest.raises(_lib.ClsmError):                     # accumulating windows need fp32 rows
        call("clsm_conv_pool_bwd_ex", P(Xd), BF16, S, L, D, Dpad, P(Wt), k, F, Fpad, P(G_d), P(A_d), F, 0,
             P(outs[1][0]), P(db), P(outs[1][1]), BF16, 0, P(ws), ws_bytes, ST())


# ------------------------------------------------------------------ Adam
def test_adam_dense_and_table():
    rng = np.random.RandomState(9)
    n = 1001
    p0, g = rng.randn(n), rng.randn(n)
    p, m, v = p0.copy(), np.zeros(n), np.zeros(n)
    pd_, md_, vd_ = dev(p0, torch.float32), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    for t in (1, 2, 3):
        O.adam_step(p, m, v, g * t, t)
        g_d = dev(g * t, torch.float32)
        call("clsm_adam_dense", P(pd_), P(md_), P(vd_), P(g_d), n, t, 1e-3, 0.9, 0.999, 1e-8, ST())
        torch.cuda.synchronize()
    np.testing.assert_allclose(pd_.cpu().numpy(), p, rtol=1e-5, atol=1e-6)
    # (1 - 0.999) evaluated in fp32, as Keras does on the device, is 1.00005e-3: 5e-5 relative on v
    np.testing.assert_allclose(vd_.cpu().numpy(), v, rtol=1e-4, atol=1e-9)

    for D in (100, 30, 7):           # vectorised and scalar variants
        V1, Dpad = 50, rup(D, 64)
        E0 = rng.randn(V1, D)
        ids = np.array([3, 7, 8, 49], dtype=np.int32)
        rows = np.zeros((4, Dpad), dtype=np.float32); rows[:, :D] = rng.randn(4, D)
        roi = np.full(V1, -1, dtype=np.int32); roi[ids] = np.arange(4)
        Ed, Em, Ev = dev(E0, torch.float32), torch.zeros((V1, D), device=DEV), torch.zeros((V1, D), device=DEV)
        copy = torch.full((V1, Dpad), 5.0, dtype=torch.bfloat16, device=DEV) if D % 4 == 0 else None
        p, m, v = E0.copy(), np.zeros((V1, D)), np.zeros((V1, D))
        gfull = np.zeros((V1, D)); gfull[ids] = rows[:, :D]
        roi_d, rows_d, ids_d, cnt_d = dev(roi), dev(rows), dev(ids), dev(np.array([4], dtype=np.int32))
        for t in (1, 2):
            O.adam_step(p, m, v, gfull, t)
            call("clsm_adam_table", P(Ed), P(Em), P(Ev), V1, D, P(roi_d), P(rows_d), Dpad, t,
                 1e-3, 0.9, 0.999, 1e-8, P(copy), Dpad, ST())
        np.testing.assert_allclose(Ed.cpu().numpy(), p, rtol=1e-5, atol=1e-6)
        assert np.abs(Ed.cpu().numpy()[0] - E0[0].astype(np.float32)).max() == 0   # zero-grad rows with m=v=0 stay put
        if copy is not None:
            np.testing.assert_array_equal(copy.float().cpu().numpy()[:, :D], bf16_round(Ed.cpu().numpy()))
        # lazy variant touches only the listed rows
        E2, m2, v2 = dev(E0, torch.float32), torch.zeros((V1, D), device=DEV), torch.zeros((V1, D), device=DEV)
        call("clsm_adam_rows", P(E2), P(m2), P(v2), D, P(ids_d), P(cnt_d), 4,
             P(rows_d), Dpad, 1, 1e-3, 0.9, 0.999, 1e-8, P(None), Dpad, ST())
        pp, mm, vv = E0.copy(), np.zeros((V1, D)), np.zeros((V1, D))
        O.adam_step(pp, mm, vv, gfull, 1)
        np.testing.assert_allclose(E2.cpu().numpy(), pp, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("D,loaded", [(300, False), (300, True), (20, False), (401, True)])
def test_deferred_rows_adam_is_bit_identical_to_the_dense_pass(D, loaded):
    """clsm_adam_defer_*: rows without gradient take their zero-gradient steps later, in registers.  After every step
    the rows the NEXT batch reads, and after settle_all the whole table with m and v, must equal clsm_adam_table's
    bit for bit -- from zeroed state (rows marked -1) and from loaded state (rows marked with the current step)."""
    rng = np.random.RandomState(91)
    V1, ldg, steps, t0 = 3000, rup(D, 64), 45, (7 if loaded else 0)
    lr, b1, b2, eps = 1e-3, 0.9, 0.999, 1e-8
    E0 = (rng.randn(V1, D) * 0.05).astype(np.float32)
    m0 = (rng.randn(V1, D) * 1e-3).astype(np.float32) if loaded else np.zeros((V1, D), np.float32)
    v0 = (rng.rand(V1, D) * 1e-5).astype(np.float32) if loaded else np.zeros((V1, D), np.float32)
    A = [dev(E0), dev(m0), dev(v0)]                    # dense
    Bt = [dev(E0), dev(m0), dev(v0)]                   # deferred
    last = torch.full((V1,), t0 if loaded else -1, dtype=torch.int32, device=DEV)
    use_copy = D % 4 == 0
    copyA = torch.zeros((V1, ldg), dtype=torch.bfloat16, device=DEV) if use_copy else None
    copyB = torch.zeros((V1, ldg), dtype=torch.bfloat16, device=DEV) if use_copy else None
    if use_copy:
        for c in (copyA, copyB):
            call("clsm_table_to_bf16", P(A[0]), V1, D, P(c), ldg, ST())
    tab = np.zeros(t0 + steps + 2, dtype=np.float32)
    call("clsm_adam_lr_table", lr, b1, b2, 0, tab.size, C.c_void_p(tab.ctypes.data))
    lr_tab = dev(tab)
    n_tok = 700
    ws_bytes = lib.clsm_adam_defer_workspace_bytes(n_tok, V1)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV)
    for i in range(steps):
        step = t0 + i + 1
        tok = (rng.zipf(1.3, size=n_tok) % V1).astype(np.int32)
        if i % 7 == 3:
            tok[:50] = rng.randint(0, V1, size=50)                # rare rows: long gaps
        ids = np.unique(tok).astype(np.int32)
        rows = np.zeros((ids.size, ldg), dtype=np.float32); rows[:, :D] = rng.randn(ids.size, D) * 0.1
        roi = np.full(V1, -1, dtype=np.int32); roi[ids] = np.arange(ids.size)
        tok_d, ids_d, rows_d, roi_d, cnt_d = dev(tok), dev(ids), dev(rows), dev(roi), dev(np.array([ids.size], dtype=np.int32))
        # what the gather of this step reads: the batch's rows as the dense pass left them after step - 1
        call("clsm_adam_defer_settle_tokens", P(Bt[0]), P(Bt[1]), P(Bt[2]), V1, D, P(last), P(tok_d), n_tok, step - 1, P(lr_tab),
             b1, b2, eps, P(copyB), ldg, P(ws), ws_bytes, ST())
        np.testing.assert_array_equal(Bt[0][ids_d.long()].cpu().numpy(), A[0][ids_d.long()].cpu().numpy())
        if use_copy:
            np.testing.assert_array_equal(copyB[ids_d.long()].float().cpu().numpy(), copyA[ids_d.long()].float().cpu().numpy())
        call("clsm_adam_table", P(A[0]), P(A[1]), P(A[2]), V1, D, P(roi_d), P(rows_d), ldg, step, lr, b1, b2, eps, P(copyA), ldg, ST())
        call("clsm_adam_defer_rows", P(Bt[0]), P(Bt[1]), P(Bt[2]), D, P(last), P(ids_d), P(cnt_d), n_tok, P(rows_d), ldg, step,
             P(lr_tab), b1, b2, eps, P(copyB), ldg, ST())
    torch.cuda.synchronize()
    assert int((last.cpu().numpy() == t0 + steps).sum()) < V1          # some rows do owe steps: the test tests something
    if not loaded:
        assert (last.cpu().numpy() == -1).any()
    call("clsm_adam_defer_settle_all", P(Bt[0]), P(Bt[1]), P(Bt[2]), V1, D, P(last), t0 + steps, P(lr_tab), b1, b2, eps, P(copyB), ldg, ST())
    torch.cuda.synchronize()
    for a, b in zip(A, Bt):
        np.testing.assert_array_equal(b.cpu().numpy(), a.cpu().numpy())
    if use_copy:
        np.testing.assert_array_equal(copyB.float().cpu().numpy(), copyA.float().cpu().numpy())
    lv = last.cpu().numpy()
    assert ((lv == t0 + steps) | (lv == -1)).all()


# ------------------------------------------------------------------ top-k
@pytest.mark.parametrize("Q,N,H,k", [(3, 50, 8, 10), (9, 5000, 128, 100), (130, 2048, 128, 100), (2, 2049, 16, 128),
                                     (1, 7, 4, 10), (5, 300000, 32, 100)])
def test_cos_topk_matches_oracle(Q, N, H, k):
    rng = np.random.RandomState(10)
    c = O.normalise_rows(rng.randn(N, H)).astype(np.float32)
    q = O.normalise_rows(rng.randn(Q, H)).astype(np.float32)
    if N > 40:
        c[17] = c[5]; c[N - 1] = c[5]; q[0] = c[5]            # exact ties -> lowest id first
    cd, qd = dev(c), dev(q)
    ws_bytes = lib.clsm_cos_topk_workspace_bytes(Q, N, k)
    ws = torch.full((ws_bytes,), 0xA5, dtype=torch.uint8, device=DEV)       # poisoned workspace
    sc = torch.empty((Q, k), dtype=torch.float32, device=DEV)
    ids = torch.empty((Q, k), dtype=torch.int64, device=DEV)
    call("clsm_cos_topk", P(qd), Q, P(cd), N, H, k, 1000, P(sc), P(ids), P(ws), ws_bytes, ST())
    # oracle on the GPU's own fp32 scores is not available; use f64 scores and compare ids where the
    # margin to the next candidate exceeds fp32 noise, and scores everywhere
    osc, oid = O.cos_topk(q.astype(np.float64), c.astype(np.float64), k, id_base=1000)
    kk = min(k, N)
    got_s, got_i = sc.cpu().numpy(), ids.cpu().numpy()
    np.testing.assert_allclose(got_s[:, :kk], osc, rtol=1e-5, atol=1e-6)
    if kk < k:
        assert (got_i[:, kk:] == -1).all()
    mism = got_i[:, :kk] != oid
    assert mism.mean() < 0.02
    for r, j in np.argwhere(mism):
        assert abs(got_s[r, j] - osc[r, j]) < 1e-5
    if N > 40:
        assert list(got_i[0, :3]) == [1005, 1017, 1000 + N - 1]
    assert (np.diff(got_s[:, :kk], axis=1) <= 0).all()


def test_topk_merge_equals_single_pass():
    rng = np.random.RandomState(11)
    Q, N, H, k, parts = 6, 9000, 32, 100, 4
    c = O.normalise_rows(rng.randn(N, H)).astype(np.float32)
    q = O.normalise_rows(rng.randn(Q, H)).astype(np.float32)
    c[8000] = c[10]; q[1] = c[10]
    cd, qd = dev(c), dev(q)

    def topk(c_t, base):
        n = c_t.shape[0]
        wsb = lib.clsm_cos_topk_workspace_bytes(Q, n, k)
        ws = torch.empty(wsb, dtype=torch.uint8, device=DEV)
        sc = torch.empty((Q, k), dtype=torch.float32, device=DEV)
        ids = torch.empty((Q, k), dtype=torch.int64, device=DEV)
        call("clsm_cos_topk", P(qd), Q, P(c_t), n, H, k, base, P(sc), P(ids), P(ws), wsb, ST())
        return sc, ids
    full_s, full_i = topk(cd, 0)
    step = N // parts
    ps = [topk(cd[i * step:(i + 1) * step].contiguous(), i * step) for i in range(parts)]
    S_ = torch.stack([p[0] for p in ps]); I_ = torch.stack([p[1] for p in ps])
    wsb = lib.clsm_topk_merge_workspace_bytes(parts, Q, k)
    ws = torch.empty(wsb, dtype=torch.uint8, device=DEV)
    ms = torch.empty((Q, k), dtype=torch.float32, device=DEV)
    mi = torch.empty((Q, k), dtype=torch.int64, device=DEV)
    call("clsm_topk_merge", P(S_), P(I_), parts, Q, k, P(ms), P(mi), P(ws), wsb, ST())
    assert torch.equal(mi, full_i) and torch.equal(ms, full_s)     # independent of the sharding


def test_topk_merge_ids_beyond_32_bits_and_span_check():
    """Merged ids are int64: shards of a corpus whose ids start past 2^32 merge exactly (keys hold id - smallest id);
    a call whose ids span 2^32 - 1 or more cannot be merged exactly and says so in its outputs."""
    rng = np.random.RandomState(15)
    parts, Q, k = 3, 5, 20
    base = 6_000_000_000
    sc = np.sort(rng.rand(parts, Q, k).astype(np.float32), axis=2)[:, :, ::-1].copy()
    sc[1, 2, :4] = sc[0, 2, :4]                                   # cross-shard score ties -> lowest id first
    ids = np.stack([np.stack([np.sort(rng.choice(1000, size=k, replace=False)) for _ in range(Q)]) + base + p_ * 1000
                    for p_ in range(parts)])
    ids[2, 0, -3:] = -1; sc[2, 0, -3:] = -np.inf                  # short list
    S_, I_ = dev(sc), dev(ids.astype(np.int64))
    wsb = lib.clsm_topk_merge_workspace_bytes(parts, Q, k)
    ws = torch.full((wsb,), 0xA5, dtype=torch.uint8, device=DEV)
    ms = torch.empty((Q, k), dtype=torch.float32, device=DEV)
    mi = torch.empty((Q, k), dtype=torch.int64, device=DEV)
    call("clsm_topk_merge", P(S_), P(I_), parts, Q, k, P(ms), P(mi), P(ws), wsb, ST())
    for q in range(Q):
        valid = [(s, i) for s, i in zip(sc[:, q].ravel(), ids[:, q].ravel()) if i >= 0]
        want = sorted(valid, key=lambda t: (-t[0], t[1]))[:k]
        np.testing.assert_array_equal(mi[q].cpu().numpy(), [w[1] for w in want])
        np.testing.assert_array_equal(ms[q].cpu().numpy(), np.array([w[0] for w in want], dtype=np.float32))
    ids[0, 0, 0] = 5                                              # 6e9 below the others: the span is not representable
    I2 = dev(ids.astype(np.int64))
    call("clsm_topk_merge", P(S_), P(I2), parts, Q, k, P(ms), P(mi), P(ws), wsb, ST())
    assert (mi.cpu().numpy() == -2).all() and np.isnan(ms.cpu().numpy()).all()


def test_normalize_rows():
    rng = np.random.RandomState(12)
    z = np.maximum(rng.randn(100, 128), 0).astype(np.float32); z[3] = 0
    out = torch.empty((100, 128), dtype=torch.float32, device=DEV)
    z_d = dev(z)
    call("clsm_normalize_rows", P(z_d), 100, 128, P(out), ST())
    np.testing.assert_allclose(out.cpu().numpy(), O.normalise_rows(z.astype(np.float64)), rtol=1e-5, atol=1e-7)
    assert (out[3] == 0).all()


# ------------------------------------------------------------------ top-k, bf16 corpus on the tensor cores
@pytest.mark.parametrize("Q,N,H,k", [(3, 50, 64, 10), (9, 5000, 128, 100), (130, 70000, 128, 100), (200, 300000, 128, 100),
                                     (2, 65537, 64, 128), (1, 7, 64, 10)])
def test_cos_topk_bf16_matches_oracle(Q, N, H, k):
    rng = np.random.RandomState(13)
    c = bf16_round(O.normalise_rows(rng.randn(N, H)))
    q = bf16_round(O.normalise_rows(rng.randn(Q, H)))
    if N > 40:
        c[17] = c[5]; c[N - 1] = c[5]; q[0] = c[5]            # exact ties across segments -> lowest id first
    cd, qd = dev(c).to(torch.bfloat16), dev(q).to(torch.bfloat16)
    ws_bytes = lib.clsm_cos_topk_bf16_workspace_bytes(Q, N, k)
    ws = torch.full((ws_bytes,), 0xA5, dtype=torch.uint8, device=DEV)       # poisoned: nothing may be read before written
    sc = torch.empty((Q, k), dtype=torch.float32, device=DEV)
    ids = torch.empty((Q, k), dtype=torch.int64, device=DEV)
    call("clsm_cos_topk_bf16", P(qd), Q, P(cd), N, H, k, 1000, P(sc), P(ids), P(ws), ws_bytes, ST())
    osc, oid = O.cos_topk(q.astype(np.float64), c.astype(np.float64), k, id_base=1000)   # same bf16-rounded inputs
    kk = min(k, N)
    got_s, got_i = sc.cpu().numpy(), ids.cpu().numpy()
    np.testing.assert_allclose(got_s[:, :kk], osc, rtol=2e-5, atol=2e-6)       # exact bf16 products, fp32 accumulate
    if kk < k:
        assert (got_i[:, kk:] == -1).all()
    mism = got_i[:, :kk] != oid
    assert mism.mean() < 0.02
    for r, j in np.argwhere(mism):
        assert abs(got_s[r, j] - osc[r, j]) < 1e-5
    if N > 40:
        assert list(got_i[0, :3]) == [1005, 1017, 1000 + N - 1]
    assert (np.diff(got_s[:, :kk], axis=1) <= 0).all()
    # second call straight after (workspace reuse): identical answer
    sc2 = torch.empty_like(sc); ids2 = torch.empty_like(ids)
    call("clsm_cos_topk_bf16", P(qd), Q, P(cd), N, H, k, 1000, P(sc2), P(ids2), P(ws), ws_bytes, ST())
    assert torch.equal(ids2, ids) and torch.equal(sc2, sc)


def test_cos_topk_bf16_adversarial_order_and_many_queries():
    """Rows ordered so that every later row scores higher for query 0: the running threshold never prunes anything,
    every row of every later segment survives into the compact lists (their capacity is a whole segment) and the
    per-query fold has to walk them in many rounds.  Also more than one 128-query block."""
    rng = np.random.RandomState(14)
    N, H, Q, k = 200000, 64, 130, 100
    c = O.normalise_rows(rng.randn(N, H))
    u = O.normalise_rows(rng.randn(1, H))[0]
    order = np.argsort(c @ u, kind="stable")                 # ascending similarity to u
    c = bf16_round(c[order])
    q = bf16_round(O.normalise_rows(rng.randn(Q, H)))
    q[0] = bf16_round(u[None, :])[0]
    cd, qd = dev(c).to(torch.bfloat16), dev(q).to(torch.bfloat16)
    ws_bytes = lib.clsm_cos_topk_bf16_workspace_bytes(Q, N, k)
    ws = torch.full((ws_bytes,), 0xA5, dtype=torch.uint8, device=DEV)
    sc = torch.empty((Q, k), dtype=torch.float32, device=DEV)
    ids = torch.empty((Q, k), dtype=torch.int64, device=DEV)
    call("clsm_cos_topk_bf16", P(qd), Q, P(cd), N, H, k, 0, P(sc), P(ids), P(ws), ws_bytes, ST())
    osc, oid = O.cos_topk(q.astype(np.float64), c.astype(np.float64), k, id_base=0)
    got_s, got_i = sc.cpu().numpy(), ids.cpu().numpy()
    np.testing.assert_allclose(got_s, osc, rtol=2e-5, atol=2e-6)
    mism = got_i != oid
    assert mism.mean() < 0.02
    for r, j in np.argwhere(mism):
        assert abs(got_s[r, j] - osc[r, j]) < 1e-5
    assert (got_i[0] >= N - 2000).all()                      # query 0's neighbours are the last rows
