"""tcgen05/TMEM/TMA convolution + max-pool kernel against the oracle and the CUDA-core kernel
(-m gpu).  Exact-arithmetic inputs (multiples of 1/8) make every partial sum representable, so
values AND arg-max positions must match bit for bit whatever the accumulation order."""
import numpy as np
import pytest
import torch

from tests.gpu_util import bf16_round
from tests.test_gpu_kernels import conv_case, oracle_conv, run_conv

pytestmark = pytest.mark.gpu

SHAPES = [(3, 9, 10, 2, 5), (5, 20, 100, 3, 150), (2, 100, 300, 3, 300), (40, 100, 300, 3, 300),
          (64, 20, 100, 4, 150), (3, 1000, 64, 5, 256), (300, 7, 64, 1, 16), (11, 33, 128, 7, 500),
          (1, 128, 64, 3, 32), (2, 129, 64, 2, 48), (37, 975, 100, 3, 150)]


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "S%d-L%d-D%d-k%d-F%d" % s)
def test_tc_exact_arithmetic_bit_exact(shape):
    S, L, D, k, F = shape
    X, W = conv_case(20, S, L, D, k, F, exact=True)
    mx, am, _ = run_conv(X, W, True, "tcgen05")
    omx, oam, _ = oracle_conv(X, W)
    np.testing.assert_array_equal(mx, omx.astype(np.float32))
    np.testing.assert_array_equal(am, oam)


@pytest.mark.parametrize("shape", [(6, 100, 300, 3, 300), (33, 20, 300, 3, 300), (9, 60, 100, 4, 150)])
def test_tc_random_matches_simt_and_oracle(shape):
    S, L, D, k, F = shape
    X, W = conv_case(21, S, L, D, k, F, exact=False)
    X[2, L // 2:] = X[2, L // 2]                 # right padding -> exact ties
    X[3, :] = X[3, 0]
    X, W = bf16_round(X).astype(np.float64), bf16_round(W).astype(np.float64)
    mx, am, _ = run_conv(X, W, True, "tcgen05")
    smx, sam, _ = run_conv(X, W, True, "simt")
    omx, oam, Y = oracle_conv(X, W)
    np.testing.assert_allclose(mx, omx, rtol=1e-4, atol=1e-5)       # fp32 accumulate of exact bf16 products
    np.testing.assert_allclose(mx, smx, rtol=1e-4, atol=1e-5)
    assert (am[3] == 0).all()
    for s, f in np.argwhere(am != oam):
        assert abs(Y[s, am[s, f], f] - Y[s, oam[s, f], f]) < 1e-4


def test_tc_is_default_for_bf16_and_rejects_fp32():
    from dnn_page_vectors_b200 import _lib
    X, W = conv_case(22, 4, 30, 64, 3, 32, exact=True)
    a = run_conv(X, W, True, "auto")
    b = run_conv(X, W, True, "tcgen05")
    assert torch.equal(a[2], b[2])
    with pytest.raises(_lib.ClsmError):
        run_conv(X, W, False, "tcgen05")


@pytest.mark.parametrize("shape", [(40, 100, 300, 3, 300), (5, 1000, 100, 4, 150), (300, 7, 64, 1, 16), (3, 9, 10, 2, 5),
                                   (2, 129, 64, 2, 48), (11, 33, 128, 7, 500), (4096, 100, 300, 3, 300)],
                         ids=lambda s: "S%d-L%d-D%d-k%d-F%d" % s)
def test_token_gather_inside_the_conv_kernel_is_bit_identical(shape):
    """clsm_conv_pool_fwd_tok: the token tile comes from the embedding table by TMA gather4 (ids = row coordinates).
    Same keys, bit for bit, as clsm_emb_gather + clsm_conv_pool_fwd on the same ids -- Zipf-like ids with a heavy PAD
    row, the first and last table rows, and a token count that is not a multiple of the tile."""
    from tests.gpu_util import DEV, P, ST, call, dev, pack_w, rup
    from dnn_page_vectors_b200 import _lib
    S, L, D, k, F = shape
    rng = np.random.RandomState(23)
    V1 = 5000
    Dpad, Fpad = rup(D, 64), rup(F, 16)
    E = rng.uniform(-0.25, 0.25, size=(V1, D)).astype(np.float32)
    W = rng.uniform(-0.05, 0.05, size=(k, D, F))
    tok = (rng.zipf(1.2, size=(S, L)) % V1).astype(np.int32)
    tok[rng.rand(S, L) < 0.3] = V1 - 1
    tok[0, 0], tok[-1, -1] = 0, V1 - 1
    table = torch.zeros((V1, Dpad), dtype=torch.bfloat16, device=DEV)
    E_d = dev(E)
    call("clsm_table_to_bf16", P(E_d), V1, D, P(table), Dpad, ST())
    Wt = pack_w(W, Dpad, Fpad, True)
    tok_d = dev(tok.reshape(-1))
    X = torch.zeros((S * L + 8, Dpad), dtype=torch.bfloat16, device=DEV)
    call("clsm_emb_gather", P(tok_d), S * L, P(table), V1, D, Dpad, _lib.BF16, P(X), Dpad, 1.0, 0, ST())
    k1 = torch.full((S, F), -1, dtype=torch.int64, device=DEV)
    k2 = torch.full((S, F), -1, dtype=torch.int64, device=DEV)
    call("clsm_conv_pool_fwd", P(X), _lib.BF16, S, L, Dpad, P(Wt), k, F, Fpad, P(k1), F, 0, _lib.IMPL["tcgen05"], ST())
    call("clsm_conv_pool_fwd_tok", P(tok_d), P(table), V1, Dpad, S, L, D, Dpad, P(Wt), k, F, Fpad, P(k2), F, 0, ST())
    k3 = torch.full((S, F), -1, dtype=torch.int64, device=DEV)       # told D: padding K steps skipped, same bits
    call("clsm_conv_pool_fwd_d", P(X), _lib.BF16, S, L, D, Dpad, P(Wt), k, F, Fpad, P(k3), F, 0, _lib.IMPL["tcgen05"], ST())
    torch.cuda.synchronize()
    assert torch.equal(k1, k2) and torch.equal(k1, k3)
