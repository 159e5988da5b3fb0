"""Multi-GPU plumbing for the CLSM path: one process per GPU, torch.distributed (NCCL over
NVLink 5 / NVSwitch on the GPU box, gloo in the CPU tests).

The reference has no data parallelism (it pins tower instances to four devices with tf.device,
dssm_cnn_v2/cnn_dssm_tf.py:139-156); SURVEY 8e replaces that with

  training  : batch split over ranks.  Small parameters: ONE allreduce of the two flat gradient buffers,
              replicated Adam.  Embedding tables: the optimizer is SHARDED by vocabulary rows -- rank r owns
              rows [r*R, (r+1)*R) of both tables.  Each rank's row-sparse gradient (ids sorted ascending) is
              cut at the ownership boundaries and sent to the owners with one all-to-all; every owner runs
              the deterministic sort + segmented reduce over what it received, applies the dense Keras Adam
              to ITS rows only (1/N of the 14 GB optimizer pass), and the refreshed bf16 compute rows are
              allgathered in place.  No (V, D) gradient is ever allreduced and replicas stay bit-identical.
              On NVLink (bf16 mode, the default) the all-to-all and the allgather are not NCCL calls at all: the
              receive buffers and compute tables of all ranks are mapped into every process (CUDA IPC) and two
              kernels of libclsm_b200 do the moving -- clsm_dp_push_rows stores each owner's piece of the
              gradient straight into the owner's memory, clsm_adam_table_dp runs the owner's Adam pass and
              stores every refreshed bf16 row into all ranks' tables from the same kernel, so the NVLink
              traffic overlaps the HBM-bound optimizer row by row and the host never waits for a count.
  encode    : pages split by contiguous row ranges, no communication;
  retrieval : queries replicated, each rank scans its shard, allgather of the (Q, k) lists and a
              k-way merge ordered (score desc, id asc), so the answer does not depend on the shard count.

Only collectives and copies happen here; all arithmetic stays in libclsm_b200.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


# ---- collectives that also work on CUDA tensors under the gloo backend --------------------------------------------
# NCCL is the backend on a multi-GPU box.  gloo carries CUDA tensors only for broadcast and all_reduce; the others are
# staged through the host here, so that the SAME DataParallel code (peer-memory kernels, CUDA IPC mappings, ordering
# collectives) can run as several processes on ONE GPU -- which is how the single-GPU test box exercises it
# (tests/test_gpu_dp.py).  Staging synchronises the stream, which only strengthens the ordering the callers rely on.
def _staged(t, group):
    return t.is_cuda and dist.get_backend(group) == "gloo"


def c_all_gather_into(out, inp, group=None):
    if _staged(inp, group):
        o = torch.empty(out.numel(), dtype=out.dtype)           # gloo wants flat (world * n) <- (n)
        dist.all_gather_into_tensor(o, inp.contiguous().cpu().reshape(-1), group=group)
        out.copy_(o.view(out.shape))
    else:
        dist.all_gather_into_tensor(out, inp, group=group)


def c_all_gather(outs, inp, group=None):
    if _staged(inp, group):
        o = [torch.empty(t.shape, dtype=t.dtype) for t in outs]
        dist.all_gather(o, inp.contiguous().cpu(), group=group)
        for a, b in zip(outs, o):
            a.copy_(b)
    else:
        dist.all_gather(outs, inp, group=group)


def c_all_to_all_single(out, inp, out_splits=None, in_splits=None, group=None):
    if _staged(inp, group):
        o = torch.empty(out.shape, dtype=out.dtype)
        dist.all_to_all_single(o, inp.contiguous().cpu(), out_splits, in_splits, group=group)
        out.copy_(o)
    else:
        dist.all_to_all_single(out, inp, out_splits, in_splits, group=group)


def shard_range(n, rank, world):
    """Contiguous row range [lo, hi) of shard `rank` (first n % world shards get one extra row)."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def owner_rows(n_rows, world):
    """Rows per owner for the sharded table optimizer: equal slices of ceil(n/world) rows."""
    return (int(n_rows) + world - 1) // world


def allgather_padded(rows, count, group=None):
    """rows: (cap, ...) tensor whose first `count` rows are valid on this rank.
    Returns (rows of all ranks concatenated in rank order, per-rank counts list)."""
    world = dist.get_world_size(group)
    cnt = torch.tensor([int(count)], dtype=torch.int64, device=rows.device)
    counts = [torch.zeros_like(cnt) for _ in range(world)]
    c_all_gather(counts, cnt, group=group)
    counts = [int(c.item()) for c in counts]
    cap = max(max(counts), 1)
    send = rows[:cap] if rows.shape[0] >= cap else torch.cat(
        [rows, rows.new_zeros((cap - rows.shape[0],) + tuple(rows.shape[1:]))])
    out = rows.new_empty((world * cap,) + tuple(rows.shape[1:]))
    c_all_gather_into(out, send.contiguous(), group=group)
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


def split_points(sorted_ids, rows_per_owner, world):
    """sorted_ids ascending (n,) -> (world,) counts of ids falling into each owner's row range."""
    bounds = torch.arange(1, world, device=sorted_ids.device, dtype=sorted_ids.dtype) * rows_per_owner
    cuts = torch.searchsorted(sorted_ids, bounds)
    edges = torch.cat([cuts.new_zeros(1), cuts, cuts.new_full((1,), sorted_ids.numel())])
    return edges[1:] - edges[:-1]


def share_across_ranks(t, group=None):
    """Map tensor `t` of every rank into this process (CUDA IPC through torch's own tensor-sharing plumbing).
    Returns one tensor per rank, all addressable from this rank's device; entry [rank] is `t` itself.
    Collective: every rank of the group must call it with a tensor of the same role."""
    import inspect
    from torch.multiprocessing import reductions
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    fn, args = reductions.reduce_tensor(t)
    objs = [None] * world
    dist.all_gather_object(objs, (fn, args), group=group)
    dev_arg = list(inspect.signature(reductions.rebuild_cuda_tensor).parameters).index("storage_device")
    out = []
    for r in range(world):
        if r == rank:
            out.append(t)
            continue
        f, a = objs[r]
        a = list(a)
        a[dev_arg] = torch.cuda.current_device()       # open the peer's allocation in THIS device's context
        out.append(f(*a))
    return out


def _ptr_array(ptrs):
    import ctypes
    return (ctypes.c_void_p * len(ptrs))(*[int(x) for x in ptrs])


class DataParallel(object):
    """Attach to a CLSM: DataParallel(net) sets net.dp; step_resident() then calls exchange()/after_adam()."""

    def __init__(self, net, group=None):
        if not dist.is_initialized():
            raise RuntimeError("initialise torch.distributed first")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.net = net
        self._bufs = {}
        for tower in (net.query_model, net.doc_model):
            # identical replicas: rank 0's weights and optimizer state everywhere
            for t in tower.state().values():
                dist.broadcast(t, src=0, group=group)
            tower.lazy_copy = False                      # the bf16 compute tables are what the ranks exchange: kept current
            R = owner_rows(tower.V1, self.world)
            tower.rows_per_owner = R
            tower.shard = (min(self.rank * R, tower.V1), min((self.rank + 1) * R, tower.V1))
            if tower.E_c is not None and tower.E_c.shape[0] != R * self.world:
                # compute copy padded to world * R rows so the in-place allgather has equal slices
                tower.E_c = torch.zeros((R * self.world, tower.Dpad), dtype=tower.E_c.dtype, device=tower.E_c.device)
            tower.refresh_compute_copies(table_too=True)
        # peer-memory path: bf16 compute tables present, rows are whole float4s, CUDA tensors, not switched off
        import os
        towers = (net.query_model, net.doc_model)
        self.p2p = (self.world > 1 and os.environ.get("CLSM_DP_P2P", "1") != "0"
                    and all(t.E_c is not None and t.E_c.is_cuda and t.D % 4 == 0 for t in towers)
                    and self.world <= 16 and not getattr(net, "lazy_table_adam", False))
        if self.p2p:
            for tower in towers:
                R = tower.rows_per_owner
                peers = share_across_ranks(tower.E_c, group)
                tower._peer_Ec = peers                      # keep the mappings alive
                off = self.rank * R * tower.Dpad * tower.E_c.element_size()
                tower._peer_Ec_rows = _ptr_array([t.data_ptr() + off for t in peers])
        net.dp = self

    # ------------------------------------------------------------------ training exchange
    def exchange(self, net, bq, bd):
        """Allreduce the small gradients; route each tower's row-sparse table gradient to the row owners and
        reduce it there.  Returns per tower (row_of_id, dE_rows, uniq, n_uniq) of the rows this rank owns."""
        from .model import _ptr, _stream
        from ._lib import call, lib
        if self.p2p:
            return self._exchange_p2p(net, bq, bd)
        allreduce_flat([net.query_model.flat_g, net.doc_model.flat_g], self.group)
        towers = ((net.query_model, bq, "q"), (net.doc_model, bd, "d"))
        W = self.world
        # one host sync for both towers' unique counts
        n_loc = torch.cat([bq["n_uniq"], bd["n_uniq"]]).tolist()
        send_cnt = torch.stack([split_points(b["uniq"][:n], t.rows_per_owner, W) for (t, b, _), n in zip(towers, n_loc)])
        send_cnt = send_cnt.to(torch.int64).contiguous()                       # (2, W)
        recv_cnt = torch.empty_like(send_cnt)
        # counts: row i of send_cnt goes to ... transpose so that the all-to-all splits by destination rank
        s_t = send_cnt.t().contiguous()                                        # (W, 2): what I send to each rank
        r_t = torch.empty_like(s_t)
        c_all_to_all_single(r_t, s_t, group=self.group)
        recv_cnt = r_t.t().contiguous()
        sc, rc = send_cnt.tolist(), recv_cnt.tolist()                          # second (and last) host sync
        out = []
        for i, (tower, b, tag) in enumerate(towers):
            n = n_loc[i]
            n_in = int(sum(rc[i]))
            g = self._bufs.get(tag)
            if g is None or g["cap"] < n_in:
                cap = int(n_in * 1.25) + 64
                dev = b["dE_rows"].device
                g = {"cap": cap,
                     "ids_in": torch.empty(cap, dtype=torch.int32, device=dev),
                     "rows_in": torch.empty((cap, tower.Dpad), dtype=torch.float32, device=dev),
                     "uniq": torch.empty(cap, dtype=torch.int32, device=dev),
                     "n_uniq": torch.zeros(1, dtype=torch.int32, device=dev),
                     "rows": torch.empty((min(cap, tower.V1), tower.Dpad), dtype=torch.float32, device=dev),
                     "row_of_id": torch.empty(tower.V1, dtype=torch.int32, device=dev)}
                g["ws_bytes"] = lib.clsm_emb_scatter_workspace_bytes(cap, tower.D)
                g["ws"] = torch.empty(g["ws_bytes"], dtype=torch.uint8, device=dev)
                self._bufs[tag] = g
            ids_in, rows_in = g["ids_in"][:n_in], g["rows_in"][:n_in]
            c_all_to_all_single(ids_in, b["uniq"][:n], rc[i], sc[i], group=self.group)
            c_all_to_all_single(rows_in, b["dE_rows"][:n], rc[i], sc[i], group=self.group)
            if n_in > 0:
                call("clsm_emb_scatter", _ptr(ids_in), n_in, _ptr(rows_in), tower.Dpad, tower.D, 1.0, 0, tower.V1,
                     _ptr(g["uniq"]), _ptr(g["n_uniq"]), _ptr(g["rows"]), _ptr(g["row_of_id"]),
                     _ptr(g["ws"]), g["ws_bytes"], _stream())
            else:
                g["row_of_id"].fill_(-1)
                g["n_uniq"].zero_()
            out.append((g["row_of_id"], g["rows"], g["uniq"], g["n_uniq"]))
        return out[0], out[1]

    def _exchange_p2p(self, net, bq, bd):
        """Peer-memory exchange: push each owner's piece of both towers' row-sparse gradients into the owner's
        receive buffers, then allreduce the small gradients -- the collective also orders the pushes before the
        owners' Adam kernels.  Returns per tower the dict Tower.adam() hands to clsm_adam_table_dp."""
        out = [self._push_tower(tower, b, tag) for tower, b, tag in ((net.query_model, bq, "q"), (net.doc_model, bd, "d"))]
        allreduce_flat([net.query_model.flat_g, net.doc_model.flat_g], self.group)
        return out[0], out[1]

    def exchange_tower(self, tower, b, tag):
        """One tower's half of _exchange_p2p, on the CURRENT stream: push its row-sparse table gradient to the owners,
        then allreduce its small gradients (which orders the pushes before the owners' Adam).  CLSM.step_resident runs
        the query tower's exchange and optimizer on one stream while the document tower's backward runs on another;
        every rank issues the two towers' collectives in the same order."""
        g = self._push_tower(tower, b, tag)
        dist.all_reduce(tower.flat_g, group=self.group)
        return g

    def push_tower(self, tower, b, tag):
        """the push alone (the caller follows it, on a stream that waits for this one, with allreduce_small)"""
        return self._push_tower(tower, b, tag)

    def allreduce_small(self, tower):
        dist.all_reduce(tower.flat_g, group=self.group)

    def _push_tower(self, tower, b, tag):
        from .model import _ptr, _stream
        from ._lib import call
        W = self.world
        g = self._bufs.get(tag)
        if g is None:
            # first step: size and map the receive buffers (collective)
            nmax = torch.tensor([int(b["n_tok"])], dtype=torch.int64, device=b["dE_rows"].device)
            dist.all_reduce(nmax, op=dist.ReduceOp.MAX, group=self.group)
            R = tower.rows_per_owner
            cap = int(min(R, int(nmax.item())))
            dev = b["dE_rows"].device
            rows_in = torch.empty((W, cap, tower.Dpad), dtype=torch.float32, device=dev)
            map_in = torch.full((W, R), -1, dtype=torch.int32, device=dev)
            peers_rows = share_across_ranks(rows_in, self.group)
            peers_map = share_across_ranks(map_in, self.group)
            g = {"p2p": True, "cap": cap, "tok_basis": int(nmax.item()), "rows_in": rows_in, "map_in": map_in,
                 "_keep": (peers_rows, peers_map),
                 "peer_rows": _ptr_array([t.data_ptr() + self.rank * cap * tower.Dpad * 4 for t in peers_rows]),
                 "peer_map": _ptr_array([t.data_ptr() + self.rank * R * 4 for t in peers_map]),
                 "seg": torch.zeros(W + 1, dtype=torch.int32, device=dev)}
            self._bufs[tag] = g
        if int(b["n_tok"]) > g["tok_basis"]:
            raise RuntimeError("data-parallel batch grew after the first step (%d > %d tokens per rank): the peer "
                               "receive buffers were sized from the first step" % (int(b["n_tok"]), g["tok_basis"]))
        call("clsm_dp_push_rows", _ptr(b["uniq"]), _ptr(b["n_uniq"]), b["dE_rows"].shape[0], _ptr(b["dE_rows"]), tower.Dpad,
             tower.D, _ptr(b["row_of_id"]), tower.V1, tower.rows_per_owner, W, g["peer_rows"], g["peer_map"], g["cap"],
             _ptr(g["seg"]), _stream())
        return g

    def after_adam(self, net):
        """Every owner has refreshed the bf16 compute rows of its slice: allgather them in place.  (Peer-memory
        path: the Adam kernels already stored them everywhere; one tiny collective orders those stores before
        the next step's gathers.)"""
        if self.p2p:
            if not hasattr(self, "_tick"):
                self._tick = torch.zeros(1, dtype=torch.float32, device=net.query_model.E_c.device)
            dist.all_reduce(self._tick, group=self.group)
            return
        for tower in (net.query_model, net.doc_model):
            if tower.E_c is None:
                self.gather_masters(tower, state=False)
                continue
            R = tower.rows_per_owner
            c_all_gather_into(tower.E_c, tower.E_c[self.rank * R:(self.rank + 1) * R], group=self.group)

    def gather_masters(self, tower, state=True):
        """Make the fp32 master table (and, with state=True, Adam m/v) complete on every rank
        (each rank only keeps its own rows current).  Used by get_weights()/save() and by the fp32 mode."""
        R = tower.rows_per_owner
        lo, hi = tower.shard
        for t in ([tower.E, tower.E_m, tower.E_v] if state else [tower.E]):
            full = t.new_zeros((R * self.world, t.shape[1]))
            mine = full[self.rank * R:(self.rank + 1) * R]
            mine[:hi - lo].copy_(t[lo:hi])
            c_all_gather_into(full, mine, group=self.group)
            t.copy_(full[:t.shape[0]])

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
        c_all_gather_into(all_s, sc.contiguous(), group=self.group)
        c_all_gather_into(all_i, ids.contiguous(), group=self.group)
        return self.net.topk_merge(all_s, all_i)
