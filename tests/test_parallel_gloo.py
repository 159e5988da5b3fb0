"""Host-side multi-rank plumbing on CPU (gloo, world_size 2): shard ranges, the padded allgather used
for the row-sparse embedding-gradient exchange, the one-bucket gradient allreduce, and that a
sharded top-k merged in (score desc, id asc) order equals the single-shard answer (oracle merge)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dnn_page_vectors_b200.parallel import allgather_padded, allreduce_flat, shard_range
from oracle import clsm_oracle as O


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. uneven row-sparse exchange: rank r owns r+2 valid rows out of a capacity of 5
        rows = torch.arange(5 * 3, dtype=torch.float32).view(5, 3) + 100 * rank
        ids = torch.arange(5, dtype=torch.int32) + 10 * rank
        n = rank + 2
        all_rows, counts = allgather_padded(rows, n)
        all_ids, _ = allgather_padded(ids, n)
        assert counts == [2, 3]
        want_rows = torch.cat([torch.arange(15, dtype=torch.float32).view(5, 3)[:2],
                               (torch.arange(15, dtype=torch.float32).view(5, 3) + 100)[:3]])
        assert torch.equal(all_rows, want_rows)
        assert all_ids.tolist() == [0, 1, 10, 11, 12]
        # 2. one-bucket allreduce of two flat gradient buffers
        a = torch.full((7,), float(rank + 1))
        b = torch.full((3,), float(10 * (rank + 1)))
        allreduce_flat([a, b])
        assert torch.equal(a, torch.full((7,), 3.0)) and torch.equal(b, torch.full((3,), 30.0))
        # 3. sharded retrieval == single shard (oracle scoring + oracle merge stand in for the kernels)
        rng = np.random.RandomState(0)
        corpus = O.normalise_rows(rng.randn(101, 8))
        corpus[60] = corpus[7]                                    # cross-shard exact tie
        q = O.normalise_rows(rng.randn(3, 8)); q[0] = corpus[7]
        lo, hi = shard_range(101, rank, world)
        sc, idx = O.cos_topk(q, corpus[lo:hi], 5, id_base=lo)
        S = [torch.zeros(3, 5, dtype=torch.float64) for _ in range(world)]
        I = [torch.zeros(3, 5, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(S, torch.as_tensor(sc)); dist.all_gather(I, torch.as_tensor(idx))
        ms, mi = O.topk_merge([s.numpy() for s in S], [i.numpy() for i in I], 5)
        fs, fi = O.cos_topk(q, corpus, 5)
        assert np.array_equal(mi, fi) and np.allclose(ms, fs)
        assert list(mi[0][:2]) == [7, 60]
        open(os.path.join(out_dir, "ok%d" % rank), "w").close()
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions_exactly():
    for n, w in ((20_000_000, 8), (101, 2), (7, 8), (0, 4)):
        spans = [shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 1


def test_two_rank_gloo(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")
