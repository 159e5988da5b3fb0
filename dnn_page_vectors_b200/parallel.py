"""Multi-GPU plumbing for the CLSM path: one process per GPU, torch.distributed (NCCL over
NVLink 5 / NVSwitch on the GPU box, gloo in the CPU tests).

The reference has no data parallelism (it pins tower instances to four devices with tf.device,
dssm_cnn_v2/cnn_dssm_tf.py:139-156); SURVEY 8e replaces that with

  training  : batch split over ranks; ONE allreduce of the flat small-parameter gradient per tower,
              and a row-sparse exchange of the embedding gradient: allgather (unique ids, summed rows)
              of every rank, then every rank runs the same deterministic sort + segmented reduce over
              the gathered set, so replicas stay bit-identical without a (V, D) dense allreduce;
  encode    : pages split by contiguous row ranges, no communication;
  retrieval : queries replicated, each rank scans its shard, allgather of the (Q, k) lists and a
              k-way merge ordered (score desc, id asc), so the answer does not depend on the shard count.

Only collectives and copies happen here; all arithmetic stays in libclsm_b200.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n, rank, world):
    """Contiguous row range [lo, hi) of shard `rank` (first n % world shards get one extra row)."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allgather_padded(rows, count, group=None):
    """rows: (cap, ...) tensor whose first `count` rows are valid on this rank.
    Returns (rows of all ranks concatenated in rank order, per-rank counts list)."""
    world = dist.get_world_size(group)
    cnt = torch.tensor([int(count)], dtype=torch.int64, device=rows.device)
    counts = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(counts, cnt, group=group)
    counts = [int(c.item()) for c in counts]
    cap = max(max(counts), 1)
    send = rows[:cap] if rows.shape[0] >= cap else torch.cat(
        [rows, rows.new_zeros((cap - rows.shape[0],) + tuple(rows.shape[1:]))])
    out = rows.new_empty((world * cap,) + tuple(rows.shape[1:]))
    dist.all_gather_into_tensor(out, send.contiguous(), group=group)
    parts = [out[r * cap:r * cap + counts[r]] for r in range(world)]
    return torch.cat(parts, dim=0), counts


def allreduce_flat(tensors, group=None):
    """Sum-allreduce a list of flat gradient buffers in one bucket."""
    if len(tensors) == 1:
        dist.all_reduce(tensors[0], group=group)
        return
    flat = torch.cat([t.reshape(-1) for t in tensors])
    dist.all_reduce(flat, group=group)
    off = 0
    for t in tensors:
        n = t.numel()
        t.copy_(flat[off:off + n].view_as(t))
        off += n


class DataParallel(object):
    """Attach to a CLSM: net.dp = DataParallel(net).  step_resident() then calls exchange()."""

    def __init__(self, net, group=None):
        if not dist.is_initialized():
            raise RuntimeError("initialise torch.distributed first")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.net = net
        self._bufs = {}
        # identical replicas: rank 0's weights and optimizer state everywhere
        for tower in (net.query_model, net.doc_model):
            for t in tower.state().values():
                dist.broadcast(t, src=0, group=group)
            tower.refresh_compute_copies(table_too=True)
        net.dp = self

    def _global_sparse(self, tower, b, tag):
        """allgather this tower's (uniq ids, rows) and re-reduce them identically on every rank."""
        import ctypes as C
        from .model import _ptr, _stream
        from ._lib import call, lib
        n_local = int(b["n_uniq"].item())                 # one small D2H sync per tower per step
        ids_all, _ = allgather_padded(b["uniq"], n_local, self.group)
        rows_all, _ = allgather_padded(b["dE_rows"], n_local, self.group)
        n_all = int(ids_all.shape[0])
        g = self._bufs.get(tag)
        if g is None or g["cap"] < n_all:
            cap = int(n_all * 1.25) + 64
            dev = rows_all.device
            g = {"cap": cap,
                 "uniq": torch.empty(cap, dtype=torch.int32, device=dev),
                 "n_uniq": torch.zeros(1, dtype=torch.int32, device=dev),
                 "rows": torch.empty((min(cap, tower.V1), tower.Dpad), dtype=torch.float32, device=dev),
                 "row_of_id": torch.empty(tower.V1, dtype=torch.int32, device=dev)}
            g["ws_bytes"] = lib.clsm_emb_scatter_workspace_bytes(cap, tower.D)
            g["ws"] = torch.empty(g["ws_bytes"], dtype=torch.uint8, device=dev)
            self._bufs[tag] = g
        ids_all = ids_all.contiguous()
        rows_all = rows_all.contiguous()
        call("clsm_emb_scatter", _ptr(ids_all), n_all, _ptr(rows_all), tower.Dpad, tower.D, 1.0, 0, tower.V1,
             _ptr(g["uniq"]), _ptr(g["n_uniq"]), _ptr(g["rows"]), _ptr(g["row_of_id"]),
             _ptr(g["ws"]), g["ws_bytes"], _stream())
        g["keepalive"] = (ids_all, rows_all)
        return g["row_of_id"], g["rows"], g["uniq"], g["n_uniq"]

    def exchange(self, net, bq, bd):
        allreduce_flat([net.query_model.flat_g, net.doc_model.flat_g], self.group)
        gq = self._global_sparse(net.query_model, bq, "q")
        gd = self._global_sparse(net.doc_model, bd, "d")
        return gq, gd

    def reduce_metrics(self, loss_acc):
        """mean loss / accuracy over ranks (equal per-rank batch)."""
        t = loss_acc.clone()
        dist.all_reduce(t, group=self.group)
        return t / self.world

    def sharded_topk(self, query_vecs, local_corpus, k, id_base):
        """Per-rank scan of its shard + allgather + merge.  Returns merged (scores, ids)."""
        sc, ids = self.net.topk(query_vecs, local_corpus, k=k, id_base=id_base)
        all_s = sc.new_empty((self.world,) + tuple(sc.shape))
        all_i = ids.new_empty((self.world,) + tuple(ids.shape))
        dist.all_gather_into_tensor(all_s, sc.contiguous(), group=self.group)
        dist.all_gather_into_tensor(all_i, ids.contiguous(), group=self.group)
        return self.net.topk_merge(all_s, all_i)
