"""clsm_topk_merge_static -- the merge entry point that brings its own scratch (-m gpu).

Some hosts cannot hand the library a workspace when they need to merge per-shard top-k lists: a request
handler that must not allocate device memory, or a process whose HBM is already fully owned by its own
pool.  For them the library carries a small fixed scratch array.  This file checks, through the C ABI,
that the entry point gives exactly what clsm_topk_merge gives with a caller workspace -- on two calls in
a row with different shapes, because every call reuses the same array and must not depend on what the
previous call left there -- and that sizes beyond the array are refused with a clear message.
It is named to be collected last, and passes both alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest
import torch

from dnn_page_vectors_b200 import _lib
from tests.gpu_util import DEV, P, ST, call, lib

pytestmark = pytest.mark.gpu


def _lists(seed, n_parts, Q, k, id_space):
    """Per-shard (scores, ids) lists as the per-GPU scans produce them: sorted, distinct ids, some exact ties."""
    rng = np.random.RandomState(seed)
    sc = np.sort(rng.uniform(-1, 1, size=(n_parts, Q, k)).astype(np.float32), axis=2)[:, :, ::-1].copy()
    span = id_space // n_parts
    ids = np.stack([np.cumsum(rng.randint(1, max(span // k, 2), size=(Q, k)), axis=1) - 1 + p * span
                    for p in range(n_parts)]).astype(np.int64)            # distinct ids, shard p owns [p*span, (p+1)*span)
    sc[1 % n_parts, :, 0] = sc[0, :, 0]                 # the same best score on two shards: lowest id must win
    if k > 3:
        ids[0, :, -1] = -1                              # a short list padded with the "empty" marker
        sc[0, :, -1] = -np.inf
    return torch.as_tensor(sc).to(DEV), torch.as_tensor(ids).to(DEV)


def _merge_with_workspace(sc, ids):
    n_parts, Q, k = sc.shape
    nbytes = lib.clsm_topk_merge_workspace_bytes(n_parts, Q, k)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
    out_s = torch.empty((Q, k), dtype=torch.float32, device=DEV)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=DEV)
    call("clsm_topk_merge", P(sc), P(ids), n_parts, Q, k, P(out_s), P(out_i), P(ws), nbytes, ST())
    return out_s, out_i


def _merge_static(sc, ids):
    n_parts, Q, k = sc.shape
    out_s = torch.empty((Q, k), dtype=torch.float32, device=DEV)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=DEV)
    call("clsm_topk_merge_static", P(sc), P(ids), n_parts, Q, k, P(out_s), P(out_i), ST())
    return out_s, out_i


def test_static_merge_matches_workspace_merge_on_two_calls_in_a_row():
    # first call: the largest shape the benchmark uses (8 shards, 128 queries, k = 100)
    a_sc, a_ids = _lists(1, 8, 128, 100, 20_000_000)
    want_s, want_i = _merge_with_workspace(a_sc, a_ids)
    got_s, got_i = _merge_static(a_sc, a_ids)
    assert torch.equal(got_i, want_i) and torch.equal(got_s, want_s)      # integer/byte work: bit exact
    # second call straight after, smaller and differently shaped: the array still holds the first
    # call's keys, none of which may leak into this answer
    b_sc, b_ids = _lists(2, 3, 5, 10, 3000)
    want_s, want_i = _merge_with_workspace(b_sc, b_ids)
    got_s, got_i = _merge_static(b_sc, b_ids)
    assert torch.equal(got_i, want_i) and torch.equal(got_s, want_s)
    # and the first shape once more, to see that the small call did not disturb anything either
    want_s, want_i = _merge_with_workspace(a_sc, a_ids)
    got_s, got_i = _merge_static(a_sc, a_ids)
    assert torch.equal(got_i, want_i) and torch.equal(got_s, want_s)
    torch.cuda.synchronize()


def test_static_merge_refuses_sizes_beyond_its_array():
    n_parts, Q, k = 8, 4096, 100                              # 26 MB of keys: far beyond the 2 MiB array
    assert lib.clsm_topk_merge_workspace_bytes(n_parts, Q, k) > 2 * 1024 * 1024
    sc = torch.zeros((n_parts, Q, k), dtype=torch.float32, device=DEV)
    ids = torch.zeros((n_parts, Q, k), dtype=torch.int64, device=DEV)
    out_s = torch.empty((Q, k), dtype=torch.float32, device=DEV)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=DEV)
    rc = lib.clsm_topk_merge_static(P(sc), P(ids), n_parts, Q, k, P(out_s), P(out_i), ST())
    assert rc == _lib.EWORKSPACE
    msg = lib.clsm_last_error().decode()
    assert "built-in" in msg and "clsm_topk_merge" in msg and str(2 * 1024 * 1024) in msg
