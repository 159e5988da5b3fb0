"""Oracle self-consistency (CPU): the NumPy restatement's hand-derived backward against torch
autograd float64, plus the edge cases the domain has (ties, dead ReLU, clip, tiny hand case)."""
import numpy as np
import pytest
import torch

from oracle import clsm_oracle as O
from tests import torch_ref as R


def make(seed, B=6, J=3, Lq=7, Ld=12, D=10, windows=(2, 3), F=8, H=5, V=50):
    rng = np.random.RandomState(seed)
    pq = O.init_tower(rng, V + 1, D, windows, F, H)
    pd = O.init_tower(rng, V + 1, D, windows, F, H)
    for p in (pq, pd):                      # non-zero biases so the bias path is exercised
        p["convb"] = [rng.uniform(-0.1, 0.1, size=F) for _ in windows]
        p["bd"] = rng.uniform(-0.1, 0.1, size=H)
        p["E"] *= 4.0
        p["convW"] = [w * 6.0 for w in p["convW"]]
    q = rng.randint(0, V, size=(B, Lq))
    d = rng.randint(0, V, size=(B, 1 + J, Ld))
    return q, d, pq, pd


@pytest.mark.parametrize("act", ["relu", "tanh"])
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_forward_backward_matches_autograd(act, seed):
    q, d, pq, pd = make(seed)
    head, gq, gd, fq, fd = O.model_loss_and_grads(q, d, pq, pd, act=act, gamma=10.0)
    ref, rq, rd = R.loss_and_grads(q, d, pq, pd, act, 10.0)
    assert abs(head["loss"] - ref["loss"]) < 1e-12
    np.testing.assert_allclose(head["cos"], ref["cos"], atol=1e-13)
    np.testing.assert_allclose(fq["sem"], ref["zq"], atol=1e-13)
    np.testing.assert_allclose(fd["pooled"], ref["Pd"], atol=1e-13)
    for g, r in ((gq, rq), (gd, rd)):
        np.testing.assert_allclose(g["E"], r["E"], atol=1e-12)
        np.testing.assert_allclose(g["Wd"], r["Wd"], atol=1e-12)
        np.testing.assert_allclose(g["bd"], r["bd"], atol=1e-12)
        for a, b in zip(g["convW"], r["convW"]):
            np.testing.assert_allclose(a, b, atol=1e-12)
        for a, b in zip(g["convb"], r["convb"]):
            np.testing.assert_allclose(a, b, atol=1e-12)


def test_dropout_mask_path():
    q, d, pq, pd = make(3)
    rng = np.random.RandomState(9)
    mq = (rng.rand(q.shape[0], q.shape[1], 10) < 0.75).astype(np.float64)
    md = (rng.rand(d.shape[0] * d.shape[1], d.shape[2], 10) < 0.75).astype(np.float64)
    head, gq, gd, _, _ = O.model_loss_and_grads(q, d, pq, pd, mask_q=mq, mask_d=md, keep=0.75)
    ref, rq, rd = R.loss_and_grads(q, d, pq, pd, mask_q=mq, mask_d=md, keep=0.75)
    assert abs(head["loss"] - ref["loss"]) < 1e-12
    np.testing.assert_allclose(gd["E"], rd["E"], atol=1e-12)
    np.testing.assert_allclose(gq["E"], rq["E"], atol=1e-12)


def test_ties_go_to_lowest_index():
    # an all-PAD document: every window is identical, so every position ties
    rng = np.random.RandomState(4)
    p = O.init_tower(rng, 6, 4, (2,), 3, 2)
    p["E"] = np.abs(p["E"]) + 0.1
    p["convW"] = [np.abs(p["convW"][0])]
    tok = np.full((2, 6), 3)
    fw = O.tower_fwd(tok, p, "relu")
    assert (fw["argmax"] == 0).all()
    assert (fw["pooled"] > 0).all()
    g = O.tower_bwd(tok, p, fw, np.ones((2, 2)), "relu")
    # gradient reaches rows 0 and 1 of each sequence only (window 2 at t=0)
    assert np.abs(g["dX"][:, 2:, :]).max() == 0.0
    assert np.abs(g["dX"][:, :2, :]).min() > 0.0


def test_dead_relu_gives_exact_zero_cosine_and_no_nan():
    zq = np.zeros((2, 4))
    zd = np.random.RandomState(0).rand(2, 3, 4)
    fw = O.sim_loss_fwd(zq, zd)
    assert (fw["cos"] == 0.0).all()
    dq, dd = O.sim_loss_bwd(zq, zd, fw)
    assert np.isfinite(dq).all() and np.isfinite(dd).all()
    assert abs(fw["loss"] - np.log(3.0)) < 1e-12


def test_negative_cosine_is_clipped_and_blocks_gradient():
    zq = np.array([[1.0, 0.0]])
    zd = np.array([[[1.0, 0.0], [-1.0, 0.2]]])
    fw = O.sim_loss_fwd(zq, zd)
    assert fw["cos"][0, 0] == 1.0 and fw["cos"][0, 1] == 0.0
    dq, dd = O.sim_loss_bwd(zq, zd, fw)
    assert np.abs(dd[0, 1]).max() == 0.0          # clipped branch: no gradient


def test_tiny_hand_case():
    # D=2, L=4, k=2, F=1, H=1, J=1: computed by hand below
    E = np.array([[1.0, 0.0], [0.0, 1.0], [1.0, 1.0], [0.0, 0.0]])
    p = {"E": E, "convW": [np.array([[[1.0], [2.0]], [[3.0], [4.0]]])], "convb": [np.array([0.5])],
         "Wd": np.array([[2.0]]), "bd": np.array([-1.0])}
    tok = np.array([[0, 1, 2, 0]])
    # windows: t0: x0=[1,0],x1=[0,1] -> 1*1+0*2 + 0*3+1*4 = 5 ; t1: [0,1],[1,1] -> 2 + 7 = 9 ;
    #          t2: [1,1],[1,0] -> 3 + 3 = 6  -> max 9 at t=1, +0.5 = 9.5 ; dense 2*9.5-1 = 18
    fw = O.tower_fwd(tok, p, "relu")
    assert fw["argmax"][0, 0] == 1
    assert fw["pooled"][0, 0] == 9.5
    assert fw["sem"][0, 0] == 18.0


def test_adam_matches_formula():
    p = np.array([1.0, -2.0]); m = np.zeros(2); v = np.zeros(2); g = np.array([0.5, -0.25])
    O.adam_step(p, m, v, g, 1)
    lr_t = 1e-3 * np.sqrt(1 - 0.999) / (1 - 0.9)
    np.testing.assert_allclose(m, 0.1 * g)
    np.testing.assert_allclose(v, 0.001 * g * g)
    np.testing.assert_allclose(p, np.array([1.0, -2.0]) - lr_t * (0.1 * g) / (np.sqrt(0.001 * g * g) + 1e-8))


def test_topk_order_and_merge():
    rng = np.random.RandomState(0)
    c = O.normalise_rows(rng.randn(300, 8))
    c[17] = c[5]; c[200] = c[5]                       # exact score ties
    qv = O.normalise_rows(rng.randn(4, 8)); qv[0] = c[5]
    sc, ids = O.cos_topk(qv, c, 10)
    assert list(ids[0][:3]) == [5, 17, 200]
    parts = [O.cos_topk(qv, c[s:s + 100], 10, id_base=s) for s in (0, 100, 200)]
    ms, mi = O.topk_merge([p[0] for p in parts], [p[1] for p in parts], 10)
    np.testing.assert_array_equal(mi, ids)
    np.testing.assert_allclose(ms, sc)
