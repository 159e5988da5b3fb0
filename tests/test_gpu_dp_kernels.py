"""The peer-memory data-parallel kernels (dp.cu) on ONE GPU (-m gpu): W ranks are simulated by W sets of buffers on
the same device, so the push / sharded Adam / broadcast arithmetic is covered without a multi-GPU box
(tests/test_gpu_dp.py + dp_check.py cover the real multi-process path).  The result must equal, bit for bit,
clsm_adam_table applied to the rank-ordered fp32 sum of the ranks' row-sparse gradients."""
import ctypes

import numpy as np
import pytest
import torch

from tests.gpu_util import DEV, P, ST, call, dev, lib

pytestmark = pytest.mark.gpu


def ptr_array(ptrs):
    return (ctypes.c_void_p * len(ptrs))(*[int(x) for x in ptrs])


def scatter(tok, dX, D, Dpad, V1):
    n = tok.numel()
    uniq = torch.empty(n, dtype=torch.int32, device=DEV)
    nu = torch.zeros(1, dtype=torch.int32, device=DEV)
    rows = torch.full((min(n, V1), Dpad), 7.0, dtype=torch.float32, device=DEV)
    roi = torch.empty(V1, dtype=torch.int32, device=DEV)
    wsb = lib.clsm_emb_scatter_workspace_bytes(n, D)
    ws = torch.empty(wsb, dtype=torch.uint8, device=DEV)
    call("clsm_emb_scatter", P(tok), n, P(dX), Dpad, D, 1.0, 0, V1, P(uniq), P(nu), P(rows), P(roi), P(ws), wsb, ST())
    return uniq, nu, rows, roi


@pytest.mark.parametrize("W,V1,D,n_tok", [(2, 1003, 100, 700), (3, 1003, 100, 2500), (8, 997, 300, 300), (4, 64, 64, 4000), (1, 50, 100, 90)])
def test_push_and_sharded_adam_equal_single_process_adam(W, V1, D, n_tok):
    rng = np.random.RandomState(21)
    Dpad = (D + 63) // 64 * 64
    R = (V1 + W - 1) // W
    cap = min(R, n_tok)
    E0 = rng.uniform(-0.25, 0.25, size=(V1, D)).astype(np.float32)
    m0 = (rng.randn(V1, D) * 1e-3).astype(np.float32)
    v0 = (rng.rand(V1, D) * 1e-6).astype(np.float32)
    E, m, v = dev(E0, torch.float32), dev(m0, torch.float32), dev(v0, torch.float32)
    Eref, mref, vref = E.clone(), m.clone(), v.clone()
    # W replicas of the bf16 compute table, padded to W*R rows, identical on entry
    Ec0 = torch.zeros((W * R, Dpad), dtype=torch.bfloat16, device=DEV)
    call("clsm_table_to_bf16", P(E), V1, D, P(Ec0), Dpad, ST())
    copies = [Ec0.clone() for _ in range(W)]
    Ec_ref = Ec0.clone()
    rows_in = [torch.full((W, cap, Dpad), float("nan"), dtype=torch.float32, device=DEV) for _ in range(W)]     # poisoned
    map_in = [torch.full((W, R), 12345, dtype=torch.int32, device=DEV) for _ in range(W)]
    seg = torch.zeros(W + 1, dtype=torch.int32, device=DEV)
    for step in (3, 4):
        g_sum = np.zeros((V1, D), dtype=np.float32)
        keep = []
        for s in range(W):                                  # "rank" s: its tokens, its scatter, its pushes
            tok = dev((rng.zipf(1.2, size=n_tok) % V1).astype(np.int32))
            dX = dev(rng.randn(n_tok, Dpad).astype(np.float32) * 1e-3, torch.float32)
            dX[:, D:] = 0
            uniq, nu, rows, roi = scatter(tok, dX, D, Dpad, V1)
            keep.append((tok, dX, uniq, nu, rows, roi))
            n_u = int(nu.item())
            ids = uniq[:n_u].cpu().numpy()
            g_sum[ids] = g_sum[ids] + rows[:n_u, :D].cpu().numpy()          # fp32, rank order: what the owner computes
            peer_rows = ptr_array([rows_in[o][s].data_ptr() for o in range(W)])
            peer_map = ptr_array([map_in[o][s].data_ptr() for o in range(W)])
            call("clsm_dp_push_rows", P(uniq), P(nu), rows.shape[0], P(rows), Dpad, D, P(roi), V1, R, W, peer_rows, peer_map, cap,
                 P(seg), ST())
        torch.cuda.synchronize()
        for o in range(W):                                  # owner o: sharded Adam + broadcast of its rows
            lo, hi = min(o * R, V1), min((o + 1) * R, V1)
            if hi <= lo:
                continue
            cps = ptr_array([copies[s].data_ptr() + o * R * Dpad * 2 for s in range(W)])
            call("clsm_adam_table_dp", P(E[lo:hi]), P(m[lo:hi]), P(v[lo:hi]), hi - lo, D, W, P(map_in[o]), R, P(rows_in[o]), cap, Dpad,
                 step, 1e-3, 0.9, 0.999, 1e-8, cps, o, Dpad, ST())
        # single-process reference: the dense sum as an all-rows "sparse" gradient through clsm_adam_table
        g_rows = torch.zeros((V1, Dpad), dtype=torch.float32, device=DEV)
        g_rows[:, :D] = dev(g_sum, torch.float32)
        ident = torch.arange(V1, dtype=torch.int32, device=DEV)
        call("clsm_adam_table", P(Eref), P(mref), P(vref), V1, D, P(ident), P(g_rows), Dpad, step, 1e-3, 0.9, 0.999, 1e-8,
             P(Ec_ref), Dpad, ST())
        torch.cuda.synchronize()
        assert torch.equal(E, Eref) and torch.equal(m, mref) and torch.equal(v, vref)
        for s in range(W):
            assert torch.equal(copies[s][:V1].view(torch.int16), Ec_ref[:V1].view(torch.int16)), "replica %d differs" % s
            assert (copies[s][V1:] == 0).all()


def test_dp_entry_points_reject_bad_arguments():
    from dnn_page_vectors_b200._lib import ClsmError
    t = torch.zeros(16, dtype=torch.int32, device=DEV)
    f = torch.zeros(64, dtype=torch.float32, device=DEV)
    pa = ptr_array([t.data_ptr()])
    with pytest.raises(ClsmError):       # world out of range
        call("clsm_dp_push_rows", P(t), P(t), 4, P(f), 4, 4, P(t), 16, 16, 17, pa, pa, 4, P(t), ST())
    with pytest.raises(ClsmError):       # owners do not cover the table
        call("clsm_dp_push_rows", P(t), P(t), 4, P(f), 4, 4, P(t), 16, 4, 1, pa, pa, 4, P(t), ST())
    with pytest.raises(ClsmError):       # D not a multiple of 4
        call("clsm_adam_table_dp", P(f), P(f), P(f), 4, 3, 1, P(t), 16, P(f), 4, 4, 1, 1e-3, 0.9, 0.999, 1e-8, pa, 0, 4, ST())
    with pytest.raises(ClsmError):       # self_rank outside the world
        call("clsm_adam_table_dp", P(f), P(f), P(f), 4, 4, 1, P(t), 16, P(f), 4, 4, 1, 1e-3, 0.9, 0.999, 1e-8, pa, 1, 4, ST())
