"""Parity at the sizes that are benchmarked (-m gpu).

The tcgen05 kernels are persistent: at BASELINE cfg2 one CTA walks ~62 token tiles and refills its weight ring
thousands of times, the retrieval scan walks segments of 65 k, 262 k, 1 M ... rows.  None of that is reached by the
small shapes of test_gpu_conv_tc.py / test_gpu_kernels.py, so these tests run the cfg2 document-tower shape
(S = 4096 sequences x 100 tokens, D = 300, window 3, 300 filters), a char-level L = 5000 shape
(dssm_cnn_v2/config.py:86-87) and a 2.5 M-row retrieval shard (cfg5's per-GPU shard) with EXACT-ARITHMETIC inputs:
every operand is a small multiple of 2^-3 / 2^-4, every product and every partial sum is representable in fp32, so
the result does not depend on the accumulation order and must equal the NumPy oracle bit for bit -- values,
arg-max positions and top-k ids (ties -> lowest index) compared with assert_array_equal.
"""
import numpy as np
import pytest
import torch

from tests.gpu_util import DEV, P, ST, call, decode_keys, dev, lib, make_X, pack_w, rup
from dnn_page_vectors_b200 import _lib

pytestmark = pytest.mark.gpu
BF16 = _lib.BF16


def exact_case(seed, S, L, D, k, F):
    rng = np.random.RandomState(seed)
    X = rng.randint(-4, 5, size=(S, L, D), dtype=np.int8).astype(np.float32) * np.float32(0.125)
    W = rng.randint(-4, 5, size=(k, D, F), dtype=np.int8).astype(np.float32) * np.float32(0.125)
    return X, W


def oracle_conv_chunked(X, W, chunk=128):
    """max over time and FIRST arg-max of the valid cross-correlation (cnn_dssm_tf.py:84-100), in sequence chunks.
    float32 is enough here: with exact-arithmetic inputs every partial sum is a multiple of 2^-6 below 2^8, so the
    matmul is exact in any order (the small shapes of test_gpu_conv_tc.py check the same thing against float64)."""
    S, L, D = X.shape
    k, _, F = W.shape
    T = L - k + 1
    mx = np.empty((S, F), dtype=np.float32)
    am = np.empty((S, F), dtype=np.int64)
    for s0 in range(0, S, chunk):
        x = X[s0:s0 + chunk]
        Y = np.zeros((x.shape[0], T, F), dtype=np.float32)
        for j in range(k):
            Y += x[:, j:j + T, :] @ W[j]
        am[s0:s0 + chunk] = Y.argmax(axis=1)
        mx[s0:s0 + chunk] = Y.max(axis=1)
    return mx, am


def run_fwd(X, W, impl):
    S, L, D = X.shape
    k, _, F = W.shape
    Dpad, Fpad = rup(D, 64), rup(F, 16)
    Xd, Wt = make_X(X, Dpad, True), pack_w(W, Dpad, Fpad, True)
    keys = torch.full((S, F), -1, dtype=torch.int64, device=DEV)
    call("clsm_conv_pool_fwd", P(Xd), BF16, S, L, Dpad, P(Wt), k, F, Fpad, P(keys), F, 0, _lib.IMPL[impl], ST())
    torch.cuda.synchronize()
    return decode_keys(keys)


@pytest.mark.parametrize("shape", [(4096, 100, 300, 3, 300),      # cfg2 document tower: 1829 tiles, 12 per persistent CTA ... x5 of bench
                                   (20480, 100, 300, 3, 300),     # the benchmarked launch itself (B*(1+J) = 20480 sequences, 9143 tiles)
                                   (3, 5000, 100, 3, 150),        # char-level documents (config.py:86-87): 23 tiles per sequence
                                   (5, 2000, 100, 4, 150)],       # n-gram level documents (config.py:90-91)
                         ids=lambda s: "S%d-L%d-D%d-k%d-F%d" % s)
def test_tcgen05_forward_bit_exact_at_bench_scale(shape):
    S, L, D, k, F = shape
    X, W = exact_case(31, S, L, D, k, F)
    mx, am = run_fwd(X, W, "tcgen05")
    omx, oam = oracle_conv_chunked(X, W)
    np.testing.assert_array_equal(mx, omx)
    np.testing.assert_array_equal(am, oam)
    if S <= 4096:                                   # the CUDA-core kernel agrees as well (it is the fp32 parity path)
        smx, sam = run_fwd(X, W, "simt")
        np.testing.assert_array_equal(smx, omx)
        np.testing.assert_array_equal(sam, oam)


def oracle_conv_bwd_chunked(X, W, G, A, chunk=128):
    """dW (k, D, F), db (F), dX (S, L, D) of conv + global max-pool given the gradient G (S, F) at the pooled
    maxima and their positions A (S, F): dY is one-hot per (s, f) column."""
    S, L, D = X.shape
    k, _, F = W.shape
    T = L - k + 1
    dW = np.zeros((k, D, F))
    dX = np.zeros((S, L, D), dtype=np.float32)
    fidx = np.arange(F)[None, :]
    for s0 in range(0, S, chunk):            # float32 chunks are exact (multiples of 2^-6, small); dW accumulates in float64
        x = X[s0:s0 + chunk]
        n = x.shape[0]
        dY = np.zeros((n, T, F), dtype=np.float32)
        dY[np.arange(n)[:, None], A[s0:s0 + n], fidx] = G[s0:s0 + n]
        for j in range(k):
            dW[j] += x[:, j:j + T, :].reshape(n * T, D).T @ dY.reshape(n * T, F)
            dX[s0:s0 + n, j:j + T, :] += dY @ W[j].T
    return dW, G.astype(np.float64).sum(axis=0), dX


@pytest.mark.parametrize("shape", [(40, 100, 300, 3, 300), (64, 20, 100, 4, 150), (37, 24, 100, 3, 150), (3, 1000, 64, 5, 256),
                                   (11, 200, 100, 6, 64), (2, 129, 64, 2, 48), (300, 7, 64, 1, 16),
                                   (4096, 100, 300, 3, 300),      # cfg2 document tower, S cut 5x
                                   (20480, 100, 300, 3, 300),     # the benchmarked launch
                                   (3, 5000, 100, 3, 150)],       # char-level documents
                         ids=lambda s: "S%d-L%d-D%d-k%d-F%d" % s)
def test_tcgen05_backward_bit_exact(shape):
    """dW, db and dX of the tensor-core backward (dw2_tc / dx2_tc, or the first tilings for k > 4) on exact-arithmetic
    inputs: G is a multiple of 1/8 (exact in bf16, the format it enters the MMA in), so every dW / dX element must
    equal the float64 oracle exactly, whatever the split-K / tile order."""
    S, L, D, k, F = shape
    rng = np.random.RandomState(32)
    X, W = exact_case(33, S, L, D, k, F)
    G = (rng.randint(-4, 5, size=(S, F)) / 8.0 * (rng.rand(S, F) > 0.4)).astype(np.float32)
    A = rng.randint(0, L - k + 1, size=(S, F)).astype(np.int32)
    A[0] = 0
    A[S - 1] = L - k
    Dpad, Fpad = rup(D, 64), rup(F, 16)
    Xd, Wt = make_X(X, Dpad, True), pack_w(W, Dpad, Fpad, True)
    ws_bytes = lib.clsm_conv_bwd_workspace_bytes(S, L, k, D, F)
    ws = torch.full((ws_bytes,), 0xA5, dtype=torch.uint8, device=DEV)
    dW = torch.empty((k, D, F), dtype=torch.float32, device=DEV)
    db = torch.empty(F, dtype=torch.float32, device=DEV)
    dX = torch.full((S * L, Dpad), 3.0, dtype=torch.float32, device=DEV)
    G_d, A_d = dev(G), dev(A)
    call("clsm_conv_pool_bwd", P(Xd), BF16, S, L, D, Dpad, P(Wt), k, F, Fpad, P(G_d), P(A_d), F, 0,
         P(dW), P(db), P(dX), 1, P(ws), ws_bytes, ST())
    torch.cuda.synchronize()
    odW, odb, odX = oracle_conv_bwd_chunked(X, W, G, A.astype(np.int64))
    np.testing.assert_array_equal(dW.cpu().numpy(), odW.astype(np.float32))
    np.testing.assert_array_equal(db.cpu().numpy(), odb.astype(np.float32))
    got = dX.cpu().numpy()
    np.testing.assert_array_equal(got[:, :D], odX.reshape(S * L, D).astype(np.float32))
    assert (got[:, D:] == 0).all()


@pytest.mark.parametrize("shape", [(40, 100, 300, (1, 2, 3, 5), 300),        # BASELINE cfg4's windows
                                   (64, 30, 100, (3, 4), 150),               # the reference's literal model (config.py)
                                   (5, 1000, 64, (2, 7, 3), 256),            # long rows, widest window, two M tiles
                                   (300, 7, 64, (1, 7), 16),                 # rows as short as the widest window
                                   (2048, 100, 300, (1, 2, 3, 5), 300)],     # ~915 tiles: every CTA walks several
                         ids=lambda s: "S%d-L%d-D%d-k%s-F%d" % (s[0], s[1], s[2], "_".join(map(str, s[3])), s[4]))
@pytest.mark.parametrize("dx16", [False, True], ids=["fp32rows", "bf16rows"])
def test_fused_multi_window_dx_bit_exact(shape, dx16):
    """clsm_conv_pool_bwd_dx_multi: the dX of all window sizes accumulated on chip and written once.  Exact-arithmetic
    inputs: the sum over windows equals the oracle's in any order; bf16 rows are that sum rounded once."""
    S, L, D, ks, F = shape
    rng = np.random.RandomState(35)
    Dpad, Fpad = rup(D, 64), rup(F, 16)
    nw = len(ks)
    X = np.zeros((S, L, D), dtype=np.float32)                    # dX does not depend on X
    G = (rng.randint(-4, 5, size=(S, nw * F)) / 8.0 * (rng.rand(S, nw * F) > 0.4)).astype(np.float32)
    A = np.zeros((S, nw * F), dtype=np.int32)
    want = np.zeros((S, L, D), dtype=np.float64)
    Wts = []
    for w, k in enumerate(ks):
        _, W = exact_case(36 + w, 1, L, D, k, F)
        Wts.append(pack_w(W, Dpad, Fpad, True))
        Aw = rng.randint(0, L - k + 1, size=(S, F)).astype(np.int32)
        Aw[0] = 0
        Aw[S - 1] = L - k
        A[:, w * F:(w + 1) * F] = Aw
        want += oracle_conv_bwd_chunked(X, W, G[:, w * F:(w + 1) * F], Aw.astype(np.int64))[2]
    import ctypes as C
    ks_c = (C.c_int * nw)(*ks)
    wt_c = (C.c_void_p * nw)(*[t.data_ptr() for t in Wts])
    assert lib.clsm_conv_bwd_dx_multi_ok(S, L, nw, ks_c, D, F) == 1
    ws_bytes = lib.clsm_conv_bwd_dx_multi_workspace_bytes(nw, ks_c, D, F)
    ws = torch.full((ws_bytes,), 0xA5, dtype=torch.uint8, device=DEV)
    dX = torch.full((S * L, Dpad), 3.0, dtype=torch.bfloat16 if dx16 else torch.float32, device=DEV)
    G_d, A_d = dev(G), dev(A)
    call("clsm_conv_pool_bwd_dx_multi", S, L, D, Dpad, wt_c, ks_c, nw, F, P(G_d), P(A_d), nw * F, 0, P(dX),
         BF16 if dx16 else _lib.F32, P(ws), ws_bytes, ST())
    torch.cuda.synchronize()
    got = dX.float().cpu().numpy()
    ref = want.reshape(S * L, D).astype(np.float32)
    if dx16:
        ref = torch.as_tensor(ref).to(torch.bfloat16).float().numpy()
    np.testing.assert_array_equal(got[:, :D], ref)
    assert (got[:, D:] == 0).all()


# ------------------------------------------------------------------ retrieval: exact ids
def exact_topk_oracle(q, c, k, id_base=0, qchunk=8):
    """Top-k by (score desc, id asc) of exact float64 scores, without a full sort per row."""
    Q, N = q.shape[0], c.shape[0]
    k = min(k, N)
    out_s = np.empty((Q, k)); out_i = np.empty((Q, k), dtype=np.int64)
    c64 = c.astype(np.float64)
    for q0 in range(0, Q, qchunk):
        sc = q[q0:q0 + qchunk].astype(np.float64) @ c64.T
        for r in range(sc.shape[0]):
            s = sc[r]
            kth = np.partition(s, N - k)[N - k]
            cand = np.nonzero(s >= kth)[0]                      # ascending ids
            order = cand[np.argsort(-s[cand], kind="stable")[:k]]
            out_s[q0 + r], out_i[q0 + r] = s[order], order + id_base
    return out_s, out_i


def exact_vectors(rng, n, H):
    """entries in {-4..4}/16: bf16-exact, dot products are multiples of 2^-8 below 2^3 -- exact in fp32 in any order"""
    return (rng.randint(-4, 5, size=(n, H)) / 16.0).astype(np.float32)


@pytest.mark.parametrize("Q,N,H,k,bf16", [(9, 5000, 128, 100, True), (130, 70000, 128, 100, True), (200, 300000, 64, 100, True),
                                         (130, 2500000, 128, 100, True),       # cfg5's per-GPU shard: every segment size, 1 GB workspace
                                         (9, 5000, 128, 100, False), (5, 300000, 32, 100, False), (3, 2049, 150, 128, False)])
def test_topk_ids_bit_exact(Q, N, H, k, bf16):
    rng = np.random.RandomState(34)
    c, q = exact_vectors(rng, N, H), exact_vectors(rng, Q, H)
    c[17] = c[5]; c[N - 1] = c[5]; q[0] = c[5]
    if bf16:
        cd, qd = dev(c).to(torch.bfloat16), dev(q).to(torch.bfloat16)
        ws_bytes = lib.clsm_cos_topk_bf16_workspace_bytes(Q, N, k)
        name = "clsm_cos_topk_bf16"
    else:
        cd, qd = dev(c), dev(q)
        ws_bytes = lib.clsm_cos_topk_workspace_bytes(Q, N, k)
        name = "clsm_cos_topk"
    ws = torch.full((ws_bytes,), 0xA5, dtype=torch.uint8, device=DEV)
    sc = torch.empty((Q, k), dtype=torch.float32, device=DEV)
    ids = torch.empty((Q, k), dtype=torch.int64, device=DEV)
    base = 7_000_000_000                                   # ids beyond 32 bits (20 M-page corpora keep them small; the ABI is int64)
    call(name, P(qd), Q, P(cd), N, H, k, base, P(sc), P(ids), P(ws), ws_bytes, ST())
    torch.cuda.synchronize()
    osc, oid = exact_topk_oracle(q, c, k, id_base=base)
    np.testing.assert_array_equal(ids.cpu().numpy(), oid)              # bit-exact ids, ties -> lowest id
    np.testing.assert_array_equal(sc.cpu().numpy(), osc.astype(np.float32))
