"""CLSM two-tower model: build / train / predict entry points over libclsm_b200.

Mirrors the model-level surface of the reference scripts (dssm_cnn_v2/cnn_dssm_tf.py):

  model(sequence_length)            -> a tower (Embedding -> Dropout -> {Conv1D -> max-pool}_k ->
                                       concat -> Dense -> act), :81-127            -> ``CLSM.model``
  RTF(vects)                        -> clipped cosine, :63-77                       -> ``RTF``
  Model(inputs, prob).compile(...)  -> :178-182                                     -> ``CLSM.compile``
  fit_generator / save / save_weights / to_json -> :189-206                         -> same names
  (added, absent in the reference)  predict, encode_pages, topk

PyTorch is used for device memory, streams and (in parallel.py) NCCL only.  Every arithmetic
step is a call into the C ABI (include/clsm_b200.h) on raw device pointers; there is no torch
fallback for any of them.
"""
from __future__ import annotations

import ctypes as C
import json
import math
import time

import numpy as np
import torch

from . import _lib
from ._lib import ACT, BF16, F32, IMPL, MATH_TF32, call, lib
from .config import ModelParams


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _round_up(x, m):
    return (x + m - 1) // m * m


class Tower(object):
    """Device-resident parameters and optimizer state of one tower (query or document).

    Reference weight shapes are kept for the master copies (so they can be exchanged with a
    Keras checkpoint): table (V+1, D); conv kernel (k, D, F) [Keras (k,1,D,F)]; conv bias (F,);
    dense (|K|F, H); dense bias (H,).  All small tensors live in ONE flat fp32 buffer (p/m/v/g) so
    the optimizer and the data-parallel allreduce are one call each.
    """

    def __init__(self, net, sequence_length, table):
        p = net.params
        dev = net.device
        self.net = net
        self.L = int(sequence_length)
        self.V1, self.D = int(table.shape[0]), int(table.shape[1])
        assert self.D == p.embedding_dim
        self.Dpad = _round_up(self.D, 64)
        self.windows = tuple(int(k) for k in p.filter_sizes)
        self.F = int(p.num_filters)
        self.Fpad = _round_up(self.F, 16)
        self.P = len(self.windows) * self.F
        self.H = int(p.hidden_dims)
        if self.L < max(self.windows):
            raise ValueError("sequence_length %d shorter than the widest window %d" % (self.L, max(self.windows)))
        self.bf16 = p.precision == "bf16"
        self.dtype_code = BF16 if self.bf16 else F32
        self.cdtype = torch.bfloat16 if self.bf16 else torch.float32

        # --- embedding table: master fp32 (V1, D) + Adam state, bf16 compute copy (V1, Dpad)
        self._E = torch.as_tensor(np.ascontiguousarray(table, dtype=np.float32)).to(dev)
        self._E_m = torch.zeros_like(self._E)
        self._E_v = torch.zeros_like(self._E)
        # Deferred rows (adam.cu): the step of a row WITHOUT gradient (g = 0, pure momentum) is not taken when the batch
        # is stepped but replayed, bit-exactly, when the row is next read.  _last[row] = the last step the stored row
        # includes, -1 while its m and v are still zero.  E / E_m / E_v (properties) settle every row before they
        # hand the tensors out; the training step itself works on _E / _E_m / _E_v.
        self._last = torch.full((self.V1,), -1, dtype=torch.int32, device=dev)
        self._owes = False                 # some row may owe steps
        self._dense_to = None              # the dense pass ran last, through this step: _last is stale until re-stamped
        self.E_c = torch.zeros((self.V1, self.Dpad), dtype=torch.bfloat16, device=dev) if self.bf16 else None
        # Single-GPU bf16 training with the DENSE table pass (defer_table_adam=False) gathers straight from the fp32
        # master (clsm_emb_gather_master), so the optimizer does not rewrite the bf16 copy every step; the copy is
        # refreshed on demand, when inference reads it (compute_table()).  Data-parallel training keeps it current: it
        # is what the ranks send each other.
        import os
        # With deferred row steps (the default) the optimizer writes only the batch's rows, so keeping their bf16 images
        # current costs ~0.1 GB per step and the gather reads 640-byte instead of 1200-byte rows: the copy stays eager.
        self.lazy_copy = (self.bf16 and self.D % 4 == 0 and not net.lazy_table_adam and not net.defer_table_adam
                          and os.environ.get("CLSM_LAZY_COPY", "1") != "0")
        self._copy_dirty = False

        # --- flat buffer of the small tensors
        sizes = [k * self.D * self.F for k in self.windows] + [self.P, self.P * self.H, self.H]
        offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        self.n_small = int(offs[-1])
        self.flat = torch.zeros(self.n_small, dtype=torch.float32, device=dev)
        self.flat_m = torch.zeros_like(self.flat)
        self.flat_v = torch.zeros_like(self.flat)
        self.flat_g = torch.zeros_like(self.flat)
        nw = len(self.windows)

        def views(buf):
            cw = [buf[offs[i]:offs[i + 1]].view(k, self.D, self.F) for i, k in enumerate(self.windows)]
            cb = buf[offs[nw]:offs[nw + 1]]
            wd = buf[offs[nw + 1]:offs[nw + 2]].view(self.P, self.H)
            bd = buf[offs[nw + 2]:offs[nw + 3]]
            return cw, cb, wd, bd

        self.convW, self.convb, self.Wd, self.bd = views(self.flat)
        self.g_convW, self.g_convb, self.g_Wd, self.g_bd = views(self.flat_g)
        # compute-layout conv weights Wt (Fpad, k*Dpad)
        self.Wt = [torch.zeros((self.Fpad, k * self.Dpad), dtype=self.cdtype, device=dev) for k in self.windows]
        self._bufs = {}
        self.shard = (0, self.V1)          # rows whose optimizer state this rank owns
        self._init_small(net.rng)
        self.refresh_compute_copies(table_too=True)

    # ---- initialisation: Keras-1 defaults (Convolution1D 'uniform' +-0.05, Dense glorot_uniform)
    def _init_small(self, rng):
        for w in self.convW:
            w.copy_(torch.as_tensor(rng.uniform(-0.05, 0.05, size=tuple(w.shape)).astype(np.float32)))
        lim = math.sqrt(6.0 / (self.P + self.H))
        self.Wd.copy_(torch.as_tensor(rng.uniform(-lim, lim, size=(self.P, self.H)).astype(np.float32)))

    @property
    def E(self):
        self.settle()
        return self._E

    @property
    def E_m(self):
        self.settle()
        return self._E_m

    @property
    def E_v(self):
        self.settle()
        return self._E_v

    @property
    def defer(self):
        return self.net.defer_table_adam and self.net.dp is None

    def _defer_copy(self):
        """the bf16 table the deferred kernels keep current row by row: none when training reads the fp32 master"""
        return self.E_c if (self.E_c is not None and not self.lazy_copy) else None

    def settle(self):
        """Every row takes the zero-gradient steps it still owes (up to net.iterations).  Called wherever the table or
        its optimizer state is read as a whole: export, checkpoint, inference, set_weights, the data-parallel set-up."""
        self._restamp()
        if not self._owes:
            return
        p, net = self.net.params, self.net
        call("clsm_adam_defer_settle_all", _ptr(self._E), _ptr(self._E_m), _ptr(self._E_v), self.V1, self.D, _ptr(self._last),
             net.iterations, _ptr(net.lr_table(net.iterations)), p.beta_1, p.beta_2, p.epsilon,
             _ptr(self._defer_copy()), self.Dpad, _stream())
        self._owes = False

    def _restamp(self):
        """after steps taken by the dense pass (which moves every row and does not stamp them)"""
        if self._dense_to is not None:
            self._last.fill_(self._dense_to)
            self._dense_to = None

    def mark_settled(self):
        """optimizer state was loaded: every row is current and none is known to have zero m, v"""
        self._last.fill_(self.net.iterations)
        self._owes, self._dense_to = False, None

    def refresh_compute_copies(self, table_too=False):
        st = _stream()
        for w, wt, k in zip(self.convW, self.Wt, self.windows):
            call("clsm_pack_conv_w", _ptr(w), k, self.D, self.F, self.Dpad, self.Fpad, self.dtype_code, _ptr(wt), st)
        if table_too and self.bf16:
            call("clsm_table_to_bf16", _ptr(self.E), self.V1, self.D, _ptr(self.E_c), self.Dpad, st)
            self._copy_dirty = False

    def compute_table(self):
        """The bf16 compute copy of the embedding table, brought up to date if training steps have moved the master
        since it was last written."""
        self.settle()
        if self.bf16 and self._copy_dirty:
            call("clsm_table_to_bf16", _ptr(self._E), self.V1, self.D, _ptr(self.E_c), self.Dpad, _stream())
            self._copy_dirty = False
        return self.E_c

    # ---- numpy import/export in reference layouts
    def get_weights(self):
        if self.net.dp is not None:
            self.net.dp.gather_masters(self)
        out = {"E": self.E.cpu().numpy()}
        for i, w in enumerate(self.convW):
            out["convW%d" % i] = w.cpu().numpy().copy()
            out["convb%d" % i] = self.convb[i * self.F:(i + 1) * self.F].cpu().numpy().copy()
        out["Wd"] = self.Wd.cpu().numpy().copy()
        out["bd"] = self.bd.cpu().numpy().copy()
        return out

    def set_weights(self, w, theano_kernels=False):
        """theano_kernels: conv kernels trained under Keras-Theano are applied flipped along the
        window axis (filter_flip=True); flip them once on import."""
        self.settle()                             # owed steps belong to the OLD table
        self._E.copy_(torch.as_tensor(np.asarray(w["E"], dtype=np.float32)))
        for i, cw in enumerate(self.convW):
            a = np.asarray(w["convW%d" % i], dtype=np.float32)
            if a.ndim == 4:                       # Keras (k, 1, D, F)
                a = a[:, 0]
            if theano_kernels:
                a = a[::-1].copy()
            cw.copy_(torch.as_tensor(np.ascontiguousarray(a)))
            self.convb[i * self.F:(i + 1) * self.F].copy_(torch.as_tensor(np.asarray(w["convb%d" % i], dtype=np.float32)))
        self.Wd.copy_(torch.as_tensor(np.asarray(w["Wd"], dtype=np.float32)))
        self.bd.copy_(torch.as_tensor(np.asarray(w["bd"], dtype=np.float32)))
        self.refresh_compute_copies(table_too=True)

    def state(self, settle=True):
        if settle:
            self.settle()
        return {"E": self._E, "E_m": self._E_m, "E_v": self._E_v,
                "flat": self.flat, "flat_m": self.flat_m, "flat_v": self.flat_v}

    # ---- per-shape scratch (allocated once per (S, train) and reused)
    def buffers(self, S, train, slot=0):
        """slot: a second (third, ...) independent set of the same shape, for callers that keep two batches in flight"""
        key = (int(S), bool(train)) if slot == 0 else (int(S), bool(train), int(slot))
        b = self._bufs.get(key)
        if b is not None:
            return b
        dev = self.net.device
        n_tok = S * self.L
        b = {"S": S, "n_tok": n_tok}
        b["tok"] = torch.empty((S, self.L), dtype=torch.int32, device=dev)
        b["X"] = torch.zeros((n_tok + 8, self.Dpad), dtype=self.cdtype, device=dev)
        b["keys"] = torch.empty((S, self.P), dtype=torch.int64, device=dev)
        b["pooled"] = torch.empty((S, self.P), dtype=torch.float32, device=dev)
        b["argmax"] = torch.empty((S, self.P), dtype=torch.int32, device=dev)
        b["sem"] = torch.empty((S, self.H), dtype=torch.float32, device=dev)
        if train:
            b["dsem"] = torch.empty((S, self.H), dtype=torch.float32, device=dev)
            b["G"] = torch.empty((S, self.P), dtype=torch.float32, device=dev)
            # dX rows: bf16 when ONE window size writes them from the tensor-core kernel (the largest stream of the step,
            # halved); fp32 when several windows accumulate into them or the CUDA-core kernels run
            dx16 = (self.bf16 and len(self.windows) == 1 and self.net.conv_impl != "simt" and
                    lib.clsm_conv_bwd_dx_bf16_ok(S, self.L, self.windows[0], self.D, self.F) == 1)
            # several window sizes: ONE fused dX pass keeps the accumulators on chip across the windows and writes the
            # rows once, bf16 too (clsm_conv_pool_bwd_dx_multi), instead of adding into fp32 rows in HBM per window
            b["dx_multi"] = False
            if self.bf16 and len(self.windows) > 1 and self.net.conv_impl != "simt":
                ks_c = (C.c_int * len(self.windows))(*self.windows)
                if lib.clsm_conv_bwd_dx_multi_ok(S, self.L, len(self.windows), ks_c, self.D, self.F) == 1:
                    b["dx_multi"], dx16 = True, True
                    b["ks_c"] = ks_c
                    b["wt_c"] = (C.c_void_p * len(self.windows))(*[w.data_ptr() for w in self.Wt])
                    b["ws_dxm_bytes"] = lib.clsm_conv_bwd_dx_multi_workspace_bytes(len(self.windows), ks_c, self.D, self.F)
                    b["ws_dxm"] = torch.empty(b["ws_dxm_bytes"], dtype=torch.uint8, device=dev)
            b["dx_dtype"] = BF16 if dx16 else F32
            b["dX"] = torch.empty((n_tok, self.Dpad), dtype=torch.bfloat16 if dx16 else torch.float32, device=dev)
            b["mask"] = torch.empty((n_tok, self.Dpad // 16), dtype=torch.int16, device=dev)       # 1 keep bit per element
            max_u = min(n_tok, self.V1)
            b["uniq"] = torch.empty(n_tok, dtype=torch.int32, device=dev)
            b["n_uniq"] = torch.zeros(1, dtype=torch.int32, device=dev)
            b["dE_rows"] = torch.empty((max_u, self.Dpad), dtype=torch.float32, device=dev)
            b["row_of_id"] = torch.empty(self.V1, dtype=torch.int32, device=dev)
            ws = max(
                lib.clsm_dense_bwd_workspace_bytes(S, self.P, self.H),
                max(lib.clsm_conv_bwd_workspace_bytes(S, self.L, k, self.D, self.F) for k in self.windows))
            b["ws"] = torch.empty(ws, dtype=torch.uint8, device=dev)
            b["ws_bytes"] = ws
            # the scatter's plan (sort + segments) is made early in the step and must survive until its reduce
            b["ws_sc_bytes"] = lib.clsm_emb_scatter_workspace_bytes(n_tok, self.D)
            b["ws_sc"] = torch.empty(b["ws_sc_bytes"], dtype=torch.uint8, device=dev)
            b["planned"] = False
            b["ws_df_bytes"] = lib.clsm_adam_defer_workspace_bytes(n_tok, self.V1)
            b["ws_df"] = torch.empty(b["ws_df_bytes"], dtype=torch.uint8, device=dev)
        self._bufs[key] = b
        return b

    # ---- forward: tok already in b['tok']
    def forward(self, b, train, seed=0):
        self.forward_conv(b, train, seed)
        return self.forward_tail(b, train)

    def forward_conv(self, b, train, seed=0):
        """embedding lookup (+ dropout) and convolution + max-pool of every window: token ids -> pool keys"""
        net = self.net
        st = _stream()
        S, n_tok = b["S"], b["n_tok"]
        keep = 1.0 - net.params.dropout_prob[0] if train else 1.0
        master = train and self.lazy_copy                   # training reads the fp32 master, see __init__
        if train and self.defer:
            self._restamp()
        if train and self.defer and self._owes:
            # the rows this batch reads take the steps they owe (all but the coming one) before they are gathered
            p = net.params
            call("clsm_adam_defer_settle_tokens", _ptr(self._E), _ptr(self._E_m), _ptr(self._E_v), self.V1, self.D, _ptr(self._last),
                 _ptr(b["tok"]), n_tok, net.iterations, _ptr(net.lr_table(net.iterations + 1)), p.beta_1, p.beta_2, p.epsilon,
                 _ptr(self._defer_copy()), self.Dpad, _ptr(b["ws_df"]), b["ws_df_bytes"], st)
        elif not (train and self.defer):
            self.settle()                                   # inference, or training with the dense pass: whole table current
        if master or not self.bf16:
            table = self._E
        else:
            table = self.E_c if (train and self.defer) else self.compute_table()
        ld_table = self.Dpad if (self.bf16 and not master) else self.D
        if not train and self.bf16 and net.conv_impl != "simt" and b.get("fused_gather", True):
            # inference: the embedding lookup happens inside the conv kernel (TMA gather4 of table rows by token id);
            # the gathered matrix X is never written.  Shapes outside that kernel's plan fall back to gather + conv.
            try:
                for i, k in enumerate(self.windows):
                    call("clsm_conv_pool_fwd_tok", _ptr(b["tok"]), _ptr(table), self.V1, self.Dpad, S, self.L, self.D, self.Dpad,
                         _ptr(self.Wt[i]), k, self.F, self.Fpad, _ptr(b["keys"]), self.P, i * self.F, st)
                fused = True
            except _lib.ClsmError:
                b["fused_gather"] = fused = False
        else:
            fused = False
        if not fused:
            self._gather_and_conv(b, train, keep, seed, table, ld_table, st)
        b["keep"], b["seed"] = keep, seed

    def forward_tail(self, b, train):
        """pool keys -> pooled activations -> semantic vectors"""
        net = self.net
        st = _stream()
        S, seed = b["S"], b["seed"]
        call("clsm_pool_finalize", _ptr(b["keys"]), S, self.P, self.P, _ptr(self.convb), ACT[net.params.activation],
             _ptr(b["pooled"]), _ptr(b["argmax"]), self.P, st)
        call("clsm_dense_fwd", _ptr(b["pooled"]), S, self.P, self.P, _ptr(self.Wd), _ptr(self.bd), self.H,
             ACT[net.params.activation] | (MATH_TF32 if self.bf16 else 0), _ptr(b["sem"]), st)
        b["keep2"] = 1.0
        if train and net.post_dense_dropout:           # v1's second Dropout, dssm_cnn/cnn_dssm.py:149
            b["keep2"], b["seed2"] = 1.0 - net.params.dropout_prob[1], seed ^ 0x5DEECE66D
            call("clsm_dropout_rows", _ptr(b["sem"]), S, self.H, b["keep2"], b["seed2"], _ptr(b["sem"]), st)
        return b["sem"]

    def _gather_and_conv(self, b, train, keep, seed, table, ld_table, st):
        net = self.net
        S, n_tok = b["S"], b["n_tok"]
        mask = _ptr(b["mask"]) if (train and keep < 1.0) else C.c_void_p(0)
        if train and self.lazy_copy:
            call("clsm_emb_gather_master", _ptr(b["tok"]), n_tok, _ptr(self._E), self.V1, self.D, _ptr(b["X"]), self.Dpad,
                 keep, seed, mask, st)
        else:
            call("clsm_emb_gather_mask", _ptr(b["tok"]), n_tok, _ptr(table), self.V1, self.D, ld_table,
                 self.dtype_code, _ptr(b["X"]), self.Dpad, keep, seed, mask, st)
        for i, k in enumerate(self.windows):
            call("clsm_conv_pool_fwd_d", _ptr(b["X"]), self.dtype_code, S, self.L, self.D, self.Dpad, _ptr(self.Wt[i]),
                 k, self.F, self.Fpad, _ptr(b["keys"]), self.P, i * self.F, IMPL[net.conv_impl], st)

    # ---- backward: gradient of the loss wrt sem in b['dsem']; fills flat_g and the sparse table grad
    def backward(self, b):
        self.backward_head(b)
        self.backward_conv(b, 3)
        self.backward_scatter(b)

    def backward_head(self, b):
        net = self.net
        st = _stream()
        S = b["S"]
        act = ACT[net.params.activation]
        if b.get("keep2", 1.0) < 1.0:                  # dropped semantic units get no gradient, kept ones 1/keep of it
            call("clsm_dropout_rows", _ptr(b["dsem"]), S, self.H, b["keep2"], b["seed2"], _ptr(b["dsem"]), st)
        call("clsm_dense_bwd", _ptr(b["pooled"]), _ptr(b["sem"]), _ptr(b["dsem"]), S, self.P, self.P, _ptr(self.Wd),
             self.H, act | (MATH_TF32 if self.bf16 else 0), _ptr(self.g_Wd), _ptr(self.g_bd), _ptr(b["G"]),
             _ptr(b["ws"]), b["ws_bytes"], st)

    def backward_conv(self, b, parts):
        """parts: 1 = conv weight / bias gradients, 2 = the gradient of the gathered rows (dX), 3 = both.  The two read the
        same tensors and write different ones (disjoint workspace regions too), so step_resident issues them on two
        streams: dX first, then dW beside the embedding scatter and the table optimizer that consume dX."""
        st = _stream()
        S = b["S"]
        multi = b.get("dx_multi", False)
        if multi and parts & 2:
            call("clsm_conv_pool_bwd_dx_multi", S, self.L, self.D, self.Dpad, b["wt_c"], b["ks_c"], len(self.windows), self.F,
                 _ptr(b["G"]), _ptr(b["argmax"]), self.P, 0, _ptr(b["dX"]), b["dx_dtype"], _ptr(b["ws_dxm"]), b["ws_dxm_bytes"], st)
            parts &= 1
            if not parts:
                return
        for i, k in enumerate(self.windows):
            gb = self.g_convb[i * self.F:(i + 1) * self.F]
            call("clsm_conv_pool_bwd_parts", _ptr(b["X"]), self.dtype_code, S, self.L, self.D, self.Dpad, _ptr(self.Wt[i]),
                 k, self.F, self.Fpad, _ptr(b["G"]), _ptr(b["argmax"]), self.P, i * self.F,
                 _ptr(self.g_convW[i]), _ptr(gb), _ptr(b["dX"]), F32 if multi else b["dx_dtype"], 1 if (i == 0 or multi) else 0, parts,
                 _ptr(b["ws"]), b["ws_bytes"], st)

    def backward_scatter(self, b):
        st = _stream()
        n_tok = b["n_tok"]
        if not b["planned"]:
            self.plan_scatter(b)
        b["planned"] = False
        use_bits = b["keep"] < 1.0
        call("clsm_emb_scatter_reduce", n_tok, _ptr(b["dX"]), b["dx_dtype"], self.Dpad, self.D, b["keep"], b["seed"],
             _ptr(b["mask"]) if use_bits else C.c_void_p(0), _ptr(b["dE_rows"]), _ptr(b["ws_sc"]), b["ws_sc_bytes"], st)

    def plan_scatter(self, b):
        """The id-only half of the embedding-gradient scatter (sort, segments, unique ids): needs no gradient, so
        step_resident issues it at the start of the step on its own stream, underneath the forward pass."""
        call("clsm_emb_scatter_plan", _ptr(b["tok"]), b["n_tok"], self.Dpad, self.D, self.V1, _ptr(b["uniq"]),
             _ptr(b["n_uniq"]), _ptr(b["row_of_id"]), _ptr(b["ws_sc"]), b["ws_sc_bytes"], _stream())
        b["planned"] = True

    def adam(self, b, step, sparse_grad=None, parts=3):
        """sparse_grad: optional (row_of_id, dE_rows, uniq, n_uniq) replacing this rank's own
        (the data-parallel path passes the globally reduced gradient of the rows this rank owns).
        self.shard = (lo, hi): table rows this rank's optimizer owns (all rows when not data parallel).
        parts: 1 = the small tensors (+ their compute copies), 2 = the embedding table, 3 = both."""
        p = self.net.params
        st = _stream()
        if parts & 1:
            call("clsm_adam_dense", _ptr(self.flat), _ptr(self.flat_m), _ptr(self.flat_v), _ptr(self.flat_g),
                 self.n_small, step, p.lr, p.beta_1, p.beta_2, p.epsilon, st)
            self.refresh_compute_copies(table_too=False)
        if not parts & 2:
            return
        lo, hi = self.shard
        if isinstance(sparse_grad, dict):
            # peer-memory data parallel path (parallel.py): gradient pieces of all ranks sit in this rank's
            # receive buffers; the kernel also stores the refreshed bf16 rows into every rank's compute table
            self._dense_to = step
            if hi > lo:
                dp = self.net.dp
                call("clsm_adam_table_dp", _ptr(self._E[lo:hi]), _ptr(self._E_m[lo:hi]), _ptr(self._E_v[lo:hi]), hi - lo, self.D,
                     dp.world, _ptr(sparse_grad["map_in"]), self.rows_per_owner, _ptr(sparse_grad["rows_in"]),
                     sparse_grad["cap"], self.Dpad, step, p.lr, p.beta_1, p.beta_2, p.epsilon,
                     self._peer_Ec_rows, dp.rank, self.Dpad, st)
            return
        row_of_id, dE_rows, uniq, n_uniq = sparse_grad if sparse_grad is not None else (
            b["row_of_id"], b["dE_rows"], b["uniq"], b["n_uniq"])
        if self.defer and sparse_grad is None and not self.net.lazy_table_adam:
            # rows with a gradient take step `step`; every other row owes it (replayed when next read, see settle())
            self._restamp()
            call("clsm_adam_defer_rows", _ptr(self._E), _ptr(self._E_m), _ptr(self._E_v), self.D, _ptr(self._last), _ptr(uniq),
                 _ptr(n_uniq), dE_rows.shape[0], _ptr(dE_rows), self.Dpad, step, _ptr(self.net.lr_table(step)),
                 p.beta_1, p.beta_2, p.epsilon, _ptr(self._defer_copy()), self.Dpad, st)
            self._owes = True
            self._copy_dirty = self._copy_dirty or (self.E_c is not None and self.lazy_copy)
        elif self.net.lazy_table_adam:
            call("clsm_adam_rows", _ptr(self._E), _ptr(self._E_m), _ptr(self._E_v), self.D, _ptr(uniq), _ptr(n_uniq),
                 dE_rows.shape[0], _ptr(dE_rows), self.Dpad, step, p.lr, p.beta_1, p.beta_2, p.epsilon,
                 _ptr(self.E_c), self.Dpad, st)
        elif hi > lo:
            copy = self.E_c[lo:hi] if (self.E_c is not None and not self.lazy_copy) else None
            self._copy_dirty = self._copy_dirty or (self.E_c is not None and self.lazy_copy)
            self.settle()
            self._dense_to = step
            call("clsm_adam_table", _ptr(self._E[lo:hi]), _ptr(self._E_m[lo:hi]), _ptr(self._E_v[lo:hi]), hi - lo, self.D,
                 _ptr(row_of_id[lo:hi]), _ptr(dE_rows), self.Dpad, step, p.lr, p.beta_1, p.beta_2, p.epsilon,
                 _ptr(copy), self.Dpad, st)

    # ---- Keras-like predict on a tower: tokens (N, L) -> semantic vectors (N, H) numpy
    def predict(self, tokens, batch_size=4096):
        tokens = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.empty((tokens.shape[0], self.H), dtype=np.float32)
        for s in range(0, tokens.shape[0], batch_size):
            chunk = tokens[s:s + batch_size]
            b = self.buffers(chunk.shape[0], False)
            b["tok"].copy_(torch.as_tensor(chunk))
            out[s:s + chunk.shape[0]] = self.forward(b, False).cpu().numpy()
        return out


def RTF(vects):
    """Clipped cosine of two (B, H) arrays, dssm_cnn_v2/cnn_dssm_tf.py:63-77, on the GPU.
    Returns (B, 1) float32 like the reference's Lambda layer."""
    x, y = (np.ascontiguousarray(v, dtype=np.float32) for v in vects)
    B, H = x.shape
    dev = torch.device("cuda", torch.cuda.current_device())
    xq, yd = torch.as_tensor(x).to(dev), torch.as_tensor(y).to(dev)
    cosv = torch.empty((B, 1), dtype=torch.float32, device=dev)
    prob = torch.empty((B, 1), dtype=torch.float32, device=dev)
    la = torch.empty(2, dtype=torch.float32, device=dev)
    call("clsm_sim_loss", _ptr(xq), _ptr(yd), B, 0, H, 1.0, B, _ptr(cosv), _ptr(prob), _ptr(la),
         C.c_void_p(0), C.c_void_p(0), _stream())
    return cosv.cpu().numpy()


class History(object):
    def __init__(self):
        self.history = {"loss": [], "acc": [], "val_loss": [], "val_acc": []}


class CLSM(object):
    """The compiled two-tower model of cnn_dssm_tf.py:130-182.

    query_length / document_length / num_negative_examples come from a Configuration
    (config.py:76-91); embedding weights from DataHelpers.get_vocab_index_embedding_weights
    (one table per tower, each initialised from the same matrix, cnn_dssm_tf.py:111-116,135-137)
    or, for char/ngram levels, random uniform +-0.05 (cnn_dssm_th.py:121-126).
    """

    def __init__(self, params=None, query_length=20, document_length=975, embedding_weights=None,
                 vocab_size=None, device=None, conv_impl="auto", lazy_table_adam=False,
                 shared_doc_tower=True, post_dense_dropout=False, defer_table_adam=None):
        """shared_doc_tower=False, post_dense_dropout=True is the v1 model (dssm_cnn/cnn_dssm.py:105-167): the
        query, the positive document and every negative document each get their OWN tower (embedding table included,
        :160-164) and a second Dropout(dropout_prob[1]) sits between Dense and relu (:148-150).  The default is v2
        (dssm_cnn_v2): one document tower shared by the positive and all negatives, no second dropout."""
        if not torch.cuda.is_available():
            raise RuntimeError("dnn_page_vectors_b200 needs a CUDA device (sm_100a); there is no CPU path")
        self.params = params if params is not None else ModelParams()
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        self.conv_impl = conv_impl
        self.lazy_table_adam = lazy_table_adam
        import os
        # dense Keras Adam with the zero-gradient steps of untouched rows deferred and replayed bit-exactly (Tower.settle)
        self.defer_table_adam = (os.environ.get("CLSM_DEFER_ADAM", "1") != "0") if defer_table_adam is None else bool(defer_table_adam)
        self.iterations = 0
        self._lr_tab, self._lr_key, self._lr_old = None, None, []
        self.rng = np.random.RandomState(self.params.seed)
        self.query_length, self.document_length = int(query_length), int(document_length)
        self.J = int(self.params.J)
        if embedding_weights is None or len(embedding_weights) == 0:
            if vocab_size is None:
                raise ValueError("pass embedding_weights (word level) or vocab_size (char/ngram level)")
            embedding_weights = self.rng.uniform(-0.05, 0.05, size=(vocab_size, self.params.embedding_dim))
        self.embedding_weights = np.asarray(embedding_weights)
        self.shared_doc_tower = bool(shared_doc_tower)
        self.post_dense_dropout = bool(post_dense_dropout)
        if self.post_dense_dropout and self.params.activation not in ("relu", "linear"):
            raise ValueError("the second dropout sits before the activation (dssm_cnn/cnn_dssm.py:148-150); it is applied "
                             "after it here, which is the same thing only for relu / linear")
        self.query_model = self.model(self.query_length)
        if self.shared_doc_tower:
            self.doc_model = self.model(self.document_length)
            self.doc_models = [self.doc_model]
        else:
            self.doc_models = [self.model(self.document_length) for _ in range(1 + self.J)]
            self.doc_model = self.doc_models[0]          # the positive-document tower (encode_pages uses it)
        self.iterations = 0
        self.compiled = False
        self.dp = None                 # set by parallel.DataParallel
        import os
        self.overlap = os.environ.get("CLSM_OVERLAP", "1") != "0"      # two-stream step, see step_resident
        self._side = None
        self.dp_pipeline = os.environ.get("CLSM_DP_PIPELINE", "1") != "0"
        self._enc = None               # encode_batches: two streams and their events
        self._tail = None              # third stream: scatter + table optimizer beside the conv weight gradients
        self.split_bwd = os.environ.get("CLSM_SPLIT_BWD", "1") != "0"
        self._plan_stream = None
        self._copy_stream = None       # host -> device token copies (see _copy_begin)
        self._ev_tok_free = None       # the last training step has finished reading its token buffers
        self._loss_pin = None          # pinned landing place of (loss, accuracy): train_on_batch waits for it alone
        self._probe = None             # scripts/overlap_probe.py: events around the two towers' post-loss work
        self._head = {}
        self._pinned = {}

    def lr_table(self, upto):
        """Device array lr_tab[s] = Adam's lr_t of step s, valid for s <= upto (the deferred-row kernels replay old steps
        with the step size each of them had).  A change of lr / beta settles every row under the old table first."""
        p = self.params
        key = (float(p.lr), float(p.beta_1), float(p.beta_2), float(p.epsilon))
        if self._lr_tab is not None and self._lr_key == key and upto < self._lr_tab.shape[0]:
            return self._lr_tab
        if self._lr_tab is not None and self._lr_key != key:
            saved, self.params = self.params, self._lr_params          # the steps owed so far were taken with the old values
            for t in [self.query_model] + list(self.doc_models):
                t.settle()
            self.params = saved
        n = max(1 << 16, 2 * (int(upto) + 1))
        host = np.zeros(n, dtype=np.float32)
        call("clsm_adam_lr_table", key[0], key[1], key[2], 0, n, C.c_void_p(host.ctypes.data))
        if self._lr_tab is not None:
            self._lr_old.append(self._lr_tab)                          # kernels already queued may still read it
        self._lr_tab, self._lr_key = torch.as_tensor(host).to(self.device), key
        import copy
        self._lr_params = copy.copy(p)
        return self._lr_tab

    def settle(self):
        """All towers: every table row takes the zero-gradient steps it still owes (Tower.settle)."""
        for t in [self.query_model] + list(self.doc_models):
            t.settle()

    def assume_optimizer_state(self):
        """Treat every table row as carrying Adam state (as after load()): its zero-gradient steps are replayed even
        while m and v are still zero.  Same results; bench.py uses it so that no row step is free."""
        for t in [self.query_model] + list(self.doc_models):
            t.settle()
            t.mark_settled()

    # cnn_dssm_tf.py:81-127
    def model(self, sequence_length=None):
        return Tower(self, sequence_length, self.embedding_weights)

    # cnn_dssm_tf.py:182
    def compile(self, optimizer="adam", loss="binary_crossentropy", metrics=("accuracy",)):
        if optimizer != "adam" or loss != "binary_crossentropy":
            raise ValueError("only optimizer='adam', loss='binary_crossentropy' (the reference's choice) exist")
        self.compiled = True
        return self

    # ---- batch staging: [q, pos, neg_0..neg_{J-1}] (host, any int dtype) -> device int32
    # ---- host -> device staging.  The copies run on their own stream: they wait only until the previous step has
    # finished READING its token buffers (forward + scatter plan, event _ev_tok_free), not for its backward and
    # optimizer -- train_on_batch returns as soon as the loss is on the host, so the next batch's transfer overlaps the
    # rest of the current step.
    def _copy_begin(self):
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._ev_copied = torch.cuda.Event()
        if self._ev_tok_free is not None:
            self._copy_stream.wait_event(self._ev_tok_free)
        return self._copy_stream

    def _copy_end(self):
        self._ev_copied.record(self._copy_stream)
        torch.cuda.current_stream().wait_event(self._ev_copied)

    def _stage(self, x, train):
        if len(x) != 2 + self.J:
            raise ValueError("expected %d input arrays [query, pos_doc, %d neg_docs], got %d" % (2 + self.J, self.J, len(x)))
        if all(isinstance(a, torch.Tensor) for a in x):
            return self._stage_tensors(x, train)
        B = int(np.asarray(x[0]).shape[0])
        key = (B, train)
        pin = self._pinned.get(key)
        if pin is None:
            pin = (torch.empty((B, self.query_length), dtype=torch.int32).pin_memory(),
                   torch.empty((B, 1 + self.J, self.document_length), dtype=torch.int32).pin_memory())
            self._pinned[key] = pin
        pq, pdoc = pin
        pq.numpy()[...] = x[0]
        dn = pdoc.numpy()
        for j in range(1 + self.J):
            dn[:, j, :] = x[1 + j]
        bq = self.query_model.buffers(B, train)
        bd = self.doc_model.buffers(B * (1 + self.J), train)
        with torch.cuda.stream(self._copy_begin()):
            bq["tok"].copy_(pq, non_blocking=True)
            bd["tok"].view(B, 1 + self.J, self.document_length).copy_(pdoc, non_blocking=True)
        self._copy_end()
        return B, bq, bd

    def _stage_tensors(self, x, train):
        """Batches handed over as int32 torch tensors (pinned host memory, or already on the device): no host-side
        repacking.  Each host array goes to the device with one asynchronous contiguous copy (a strided host copy of
        400-byte rows was measured at 2 ms per step on one of the test hosts); a strided device-to-device copy
        (clsm_copy_rows_strided) then puts each document input into its slot of the document tower's interleaved
        (pair, document) token matrix."""
        B = int(x[0].shape[0])
        for a in x:
            if a.dtype != torch.int32 or not a.is_contiguous():
                raise ValueError("tensor batches must be contiguous int32 (pinned host or device memory)")
        bq = self.query_model.buffers(B, train)
        bd = self.doc_model.buffers(B * (1 + self.J), train)
        Ld, n = self.document_length, 1 + self.J
        stage = bd.get("tok_stage")
        if stage is None:
            stage = bd["tok_stage"] = torch.empty((n, B, Ld), dtype=torch.int32, device=self.device)
            bq["tok_stage"] = torch.empty((B, self.query_length), dtype=torch.int32, device=self.device)
        src = list(x)
        if any(not a.is_cuda for a in x):
            with torch.cuda.stream(self._copy_begin()):
                for i, a in enumerate(x):
                    if not a.is_cuda:
                        dst = bq["tok_stage"] if i == 0 else stage[i - 1]
                        dst.copy_(a, non_blocking=True)
                        src[i] = dst
            self._copy_end()
        st = _stream()
        bq["tok"].copy_(src[0], non_blocking=True)
        for j in range(n):
            call("clsm_copy_rows_strided", C.c_void_p(src[1 + j].data_ptr()), Ld * 4, C.c_void_p(bd["tok"].data_ptr() + j * Ld * 4),
                 n * Ld * 4, Ld * 4, B, st)
        return B, bq, bd

    def _head_bufs(self, B):
        h = self._head.get(B)
        if h is None:
            dev = self.device
            h = {"cos": torch.empty((B, 1 + self.J), dtype=torch.float32, device=dev),
                 "prob": torch.empty((B, 1 + self.J), dtype=torch.float32, device=dev),
                 "loss_acc": torch.empty(2, dtype=torch.float32, device=dev)}
            self._head[B] = h
        return h

    def _forward_head(self, B, bq, bd, train, batch_total=None):
        h = self._head_bufs(B)
        call("clsm_sim_loss", _ptr(bq["sem"]), _ptr(bd["sem"]), B, self.J, self.query_model.H,
             float(self.params.GAMMA), int(batch_total or B), _ptr(h["cos"]), _ptr(h["prob"]), _ptr(h["loss_acc"]),
             _ptr(bq["dsem"]) if train else C.c_void_p(0), _ptr(bd["dsem"]) if train else C.c_void_p(0), _stream())
        return h

    def _seeds(self):
        base = (self.params.seed << 20) + self.iterations * 2
        rank = self.dp.rank if self.dp is not None else 0
        return base + (rank << 44), base + 1 + (rank << 44)

    # ---- device-resident step (bench `value`): tokens already staged in the tower buffers
    def _mark(self, name):
        """scripts/overlap_probe.py: an event on the current stream at a named point of the step"""
        if self._probe is not None and name in self._probe:
            self._probe[name].record(torch.cuda.current_stream())

    def step_resident(self, B, bq, bd):
        """One training step.  After the loss kernel the two towers are independent: the query tower's backward,
        scatter and table Adam (HBM-bound: a 1.2 GB table and its Adam state) and the document tower's tensor-bound
        conv gradients run on two streams, and the streams join before the step returns.  Every kernel is
        deterministic and the towers share no buffer, so the weights are bit-identical to the one-stream order
        (CLSM_OVERLAP=0 selects it; tests/test_gpu_model.py compares the two)."""
        sq, sd = self._seeds()
        if self.overlap:
            # the scatters' plans (sort of the token ids, segments) on their own stream, under the forward pass
            main = torch.cuda.current_stream()
            if self._plan_stream is None:
                self._plan_stream = torch.cuda.Stream(device=self.device)
                self._ev_tok, self._ev_plan = torch.cuda.Event(), torch.cuda.Event()
            self._ev_tok.record(main)
            self._plan_stream.wait_event(self._ev_tok)
            with torch.cuda.stream(self._plan_stream):
                self.doc_model.plan_scatter(bd)
                self.query_model.plan_scatter(bq)
                self._ev_plan.record(self._plan_stream)
        self.query_model.forward(bq, True, sq)
        self._mark("fq")
        self.doc_model.forward(bd, True, sd)
        self._mark("fd")
        bt = B * (self.dp.world if self.dp is not None else 1)
        h = self._forward_head(B, bq, bd, True, bt)
        self.iterations += 1
        if self.overlap:
            main.wait_event(self._ev_plan)
        if self._loss_pin is None:
            self._loss_pin = torch.empty(2, dtype=torch.float32).pin_memory()
            self._ev_loss = torch.cuda.Event()
            self._ev_tok_free = torch.cuda.Event()
        self._loss_pin.copy_(h["loss_acc"], non_blocking=True)      # the step's result leaves right after the loss kernel
        self._ev_loss.record(torch.cuda.current_stream())
        if self.overlap:                                            # forward and scatter plans are done with the token buffers
            self._ev_tok_free.record(torch.cuda.current_stream())
        if self.dp is not None and self.dp.p2p and self.overlap and self.dp_pipeline:
            # Data parallel, peer-memory path.  The towers are independent after the loss, so the query tower's whole
            # tail -- backward, row push, small-gradient allreduce, sharded table optimizer with its NVLink broadcast --
            # runs on this stream while the document tower's tensor-bound backward runs on the high-priority one.
            # Collectives are issued in the same order on every rank (query tower's first).
            main = torch.cuda.current_stream()
            if self._side is None:
                self._side = torch.cuda.Stream(device=self.device, priority=-1)
                self._ev_fork, self._ev_join = torch.cuda.Event(), torch.cuda.Event()
            self._ev_fork.record(main)
            self._side.wait_event(self._ev_fork)
            self.query_model.backward(bq)
            gq = self.dp.exchange_tower(self.query_model, bq, "q")
            self.query_model.adam(bq, self.iterations, gq)
            dm = self.doc_model
            if self._tail is None:
                self._tail = torch.cuda.Stream(device=self.device)
                self._ev_dx, self._ev_tab = torch.cuda.Event(), torch.cuda.Event()
            with torch.cuda.stream(self._side):
                if self.split_bwd:
                    # dX first: its scatter and the row push to the owners (HBM / NVLink) then run beside the conv
                    # weight gradients (tensor cores), and the allreduce that orders the pushes follows both
                    dm.backward_head(bd)
                    dm.backward_conv(bd, 2)
                    self._ev_dx.record(self._side)
                    self._tail.wait_event(self._ev_dx)
                    with torch.cuda.stream(self._tail):
                        dm.backward_scatter(bd)
                        gd = self.dp.push_tower(dm, bd, "d")
                        self._ev_tab.record(self._tail)
                    dm.backward_conv(bd, 1)
                    self._side.wait_event(self._ev_tab)
                    self.dp.allreduce_small(dm)
                else:
                    dm.backward(bd)
                    gd = self.dp.exchange_tower(dm, bd, "d")
                dm.adam(bd, self.iterations, gd)
                self._ev_join.record(self._side)
            main.wait_event(self._ev_join)
            self.dp.after_adam(self)
        elif self.dp is not None:
            self.query_model.backward(bq)
            self.doc_model.backward(bd)
            gq, gd = self.dp.exchange(self, bq, bd)
            self.query_model.adam(bq, self.iterations, gq)
            self.doc_model.adam(bd, self.iterations, gd)
            self.dp.after_adam(self)
        elif self.overlap:
            # The document tower's persistent tensor-core kernels go to a HIGH-priority stream: their 148 CTAs become
            # resident as soon as they are launched and the query tower's short HBM-bound CTAs (current stream) fill
            # the thread slots that are left, instead of the first-launched grid holding every SM until it drains.
            main = torch.cuda.current_stream()
            if self._side is None:
                self._side = torch.cuda.Stream(device=self.device, priority=-1)
                self._ev_fork, self._ev_join = torch.cuda.Event(), torch.cuda.Event()
            self._ev_fork.record(main)
            self._side.wait_event(self._ev_fork)
            if self._probe is not None:
                self._probe["q0"].record(main)
            self.query_model.backward(bq)             # issued first: the host needs ~1 ms to issue a tower's backward
            self.query_model.adam(bq, self.iterations)
            if self._probe is not None:
                self._probe["q1"].record(main)
            dm = self.doc_model
            if self._tail is None:
                self._tail = torch.cuda.Stream(device=self.device)
                self._ev_dx, self._ev_tab = torch.cuda.Event(), torch.cuda.Event()
            with torch.cuda.stream(self._side):
                if self._probe is not None:
                    self._probe["d0"].record(self._side)
                dm.backward_head(bd)
                if self.split_bwd:
                    # dX first; then the conv weight gradients (tensor-bound, persistent CTAs, this high-priority stream)
                    # run beside the scatter of dX and the table optimizer (HBM-bound, short CTAs, a third stream)
                    self._mark("hd")
                    dm.backward_conv(bd, 2)
                    self._mark("dx")
                    self._ev_dx.record(self._side)
                    self._tail.wait_event(self._ev_dx)
                    with torch.cuda.stream(self._tail):
                        dm.backward_scatter(bd)
                        self._mark("sc")
                        dm.adam(bd, self.iterations, parts=2)
                        self._mark("tab")
                        self._ev_tab.record(self._tail)
                    dm.backward_conv(bd, 1)
                    self._mark("dw")
                    dm.adam(bd, self.iterations, parts=1)
                    self._side.wait_event(self._ev_tab)
                else:
                    dm.backward_conv(bd, 3)
                    dm.backward_scatter(bd)
                    dm.adam(bd, self.iterations)
                if self._probe is not None:
                    self._probe["d1"].record(self._side)
                self._ev_join.record(self._side)
            main.wait_event(self._ev_join)
        else:
            self.query_model.backward(bq)
            self.doc_model.backward(bd)
            self.query_model.adam(bq, self.iterations)
            self.doc_model.adam(bd, self.iterations)
        if not self.overlap:                                        # one-stream order: the plans read the tokens in backward()
            self._ev_tok_free.record(torch.cuda.current_stream())
        return h

    # ---- Keras-like batch API (host numpy in, python floats out)
    def train_on_batch(self, x, y=None):
        if not self.shared_doc_tower:
            h = self._step_independent(x, True)
            if h is None:
                return [0.0, 0.0]
            la = h["loss_acc"].cpu().numpy()
            return [float(la[0]), float(la[1])]
        B, bq, bd = self._stage(x, True)
        if B == 0:
            return [0.0, 0.0]
        self.step_resident(B, bq, bd)
        self._ev_loss.synchronize()             # loss and accuracy of THIS step; its backward / optimizer run on behind
        return [float(self._loss_pin[0]), float(self._loss_pin[1])]

    def test_on_batch(self, x, y=None):
        if not self.shared_doc_tower:
            la = self._step_independent(x, False)["loss_acc"].cpu().numpy()
            return [float(la[0]), float(la[1])]
        B, bq, bd = self._stage(x, False)
        self.query_model.forward(bq, False)
        self.doc_model.forward(bd, False)
        la = self._forward_head(B, bq, bd, False)["loss_acc"].cpu().numpy()
        return [float(la[0]), float(la[1])]

    def predict_on_batch(self, x):
        if not self.shared_doc_tower:
            return self._step_independent(x, False)["prob"][:, :1].cpu().numpy()
        B, bq, bd = self._stage(x, False)
        self.query_model.forward(bq, False)
        self.doc_model.forward(bd, False)
        return self._forward_head(B, bq, bd, False)["prob"][:, :1].cpu().numpy()

    # ---- the v1 wiring: 2 + J independent towers (dssm_cnn/cnn_dssm.py:153-185), same kernels
    def _step_independent(self, x, train):
        """One step (train) or forward (eval) of the model with one tower per input.  Each document tower sees its own
        (B, Ld) token matrix; their semantic vectors are interleaved into the (B, 1+J, H) layout the loss kernel reads,
        and its gradient is dealt back tower by tower.  Towers are stepped one after the other on one stream."""
        if len(x) != 2 + self.J:
            raise ValueError("expected %d input arrays [query, pos_doc, %d neg_docs], got %d" % (2 + self.J, self.J, len(x)))
        if self.dp is not None:
            raise RuntimeError("data parallelism is wired for the shared-document-tower model only")
        B = int(np.asarray(x[0]).shape[0])
        if B == 0:
            return None
        towers = [self.query_model] + self.doc_models
        bufs = [t.buffers(B, train) for t in towers]
        for b, a in zip(bufs, x):
            b["tok"].copy_(torch.as_tensor(np.ascontiguousarray(a, dtype=np.int32)), non_blocking=False)
        H = self.query_model.H
        key = ("indep", B, train)
        hb = self._head.get(key)
        if hb is None:
            hb = {"sem": torch.empty((B, 1 + self.J, H), dtype=torch.float32, device=self.device),
                  "dsem": torch.empty((B, 1 + self.J, H), dtype=torch.float32, device=self.device)}
            self._head[key] = hb
        base = (self.params.seed << 20) + self.iterations * (2 + self.J)
        for i, (t, b) in enumerate(zip(towers, bufs)):
            t.forward(b, train, base + i)
            if i > 0:
                hb["sem"][:, i - 1, :].copy_(b["sem"])
        h = self._head_bufs(B)
        bq = bufs[0]
        call("clsm_sim_loss", _ptr(bq["sem"]), _ptr(hb["sem"]), B, self.J, H, float(self.params.GAMMA), B,
             _ptr(h["cos"]), _ptr(h["prob"]), _ptr(h["loss_acc"]),
             _ptr(bq["dsem"]) if train else C.c_void_p(0), _ptr(hb["dsem"]) if train else C.c_void_p(0), _stream())
        if train:
            self.iterations += 1
            for i, (t, b) in enumerate(zip(towers, bufs)):
                if i > 0:
                    b["dsem"].copy_(hb["dsem"][:, i - 1, :])
                t.backward(b)
                t.adam(b, self.iterations)
        return h

    predict = predict_on_batch

    # cnn_dssm_tf.py:194
    def fit_generator(self, generator, samples_per_epoch, nb_epoch, validation_data=None,
                      nb_val_samples=None, callbacks=None, verbose=1):
        if not self.compiled:
            raise RuntimeError("compile() the model first")
        hist = History()
        callbacks = list(callbacks or [])
        for epoch in range(nb_epoch):
            seen, lsum, asum, t0 = 0, 0.0, 0.0, time.time()
            while seen < samples_per_epoch:
                x, y = next(generator)
                n = len(y)
                if n == 0:
                    continue
                l, a = self.train_on_batch(x, y)
                lsum += l * n
                asum += a * n
                seen += n
            logs = {"loss": lsum / max(seen, 1), "acc": asum / max(seen, 1)}
            if validation_data is not None:
                vseen, vl, va = 0, 0.0, 0.0
                while vseen < (nb_val_samples or 0):
                    x, y = next(validation_data)
                    n = len(y)
                    if n == 0:
                        continue
                    l, a = self.test_on_batch(x, y)
                    vl += l * n
                    va += a * n
                    vseen += n
                logs["val_loss"] = vl / max(vseen, 1)
                logs["val_acc"] = va / max(vseen, 1)
            for k, v in logs.items():
                hist.history[k].append(v)
            if verbose:
                print("Epoch %d/%d - %.1fs - %s" % (epoch + 1, nb_epoch, time.time() - t0,
                      " - ".join("%s: %.4f" % kv for kv in sorted(logs.items()))))
            for cb in callbacks:
                cb(self, epoch, logs)
        return hist

    # ---- page vectors and retrieval (absent in the reference; composition of model() and RTF)
    def encode_pages(self, tokens, batch_size=8192, out=None, normalise=True, dtype=torch.float32):
        """tokens (N, Ld) host ints -> device tensor (N, H) of L2-normalised page vectors.
        dtype=torch.bfloat16 gives the compact corpus format the tensor-core retrieval path scans."""
        tokens = np.ascontiguousarray(tokens, dtype=np.int32)
        return self._encode(self.doc_model, tokens, batch_size, out, normalise, dtype)

    def encode_queries(self, tokens, batch_size=8192, normalise=True, dtype=torch.float32):
        tokens = np.ascontiguousarray(tokens, dtype=np.int32)
        return self._encode(self.query_model, tokens, batch_size, None, normalise, dtype)

    def _encode(self, tower, tokens, batch_size, out, normalise, dtype=torch.float32):
        N = tokens.shape[0]
        if out is None:
            out = torch.empty((N, tower.H), dtype=dtype, device=self.device)
        starts = list(range(0, N, batch_size))

        def fill(i, b):
            b["tok"].copy_(torch.as_tensor(tokens[starts[i]:starts[i] + batch_size]), non_blocking=False)
        self.encode_batches(tower, [min(batch_size, N - s0) for s0 in starts], fill,
                            lambda i: out[starts[i]:starts[i] + batch_size], normalise)
        return out

    def encode_batches(self, tower, sizes, fill, out_of, normalise=True):
        """Encode a run of batches with two of them in flight: batch i's convolution (tensor cores, every SM) on one
        stream, batch i-1's tail -- pool finalize, dense projection, normalisation, ~0.2 of a batch's time as short
        launches -- on another, two buffer sets alternating.  sizes[i] = rows of batch i; fill(i, b) writes its token ids
        into b["tok"] (called under the convolution stream); out_of(i) = its (rows, H) destination."""
        if not sizes:
            return
        tower.settle()
        main = torch.cuda.current_stream()
        if self._enc is None:
            self._enc = {"conv": torch.cuda.Stream(device=self.device, priority=-1), "tail": torch.cuda.Stream(device=self.device),
                         "start": torch.cuda.Event(), "conv_done": [torch.cuda.Event(), torch.cuda.Event()],
                         "tail_done": [torch.cuda.Event(), torch.cuda.Event()], "end": torch.cuda.Event()}
        e = self._enc
        e["start"].record(main)
        e["conv"].wait_event(e["start"])
        e["tail"].wait_event(e["start"])
        used = [False, False]
        for i, S in enumerate(sizes):
            slot = i & 1
            b = tower.buffers(S, False, slot=slot)
            with torch.cuda.stream(e["conv"]):
                if used[slot]:
                    e["conv"].wait_event(e["tail_done"][slot])       # the tail of batch i-2 has read these buffers
                fill(i, b)
                tower.forward_conv(b, False)
                e["conv_done"][slot].record(e["conv"])
            with torch.cuda.stream(e["tail"]):
                e["tail"].wait_event(e["conv_done"][slot])
                self._encode_tail(tower, b, out_of(i), normalise)
                e["tail_done"][slot].record(e["tail"])
            used[slot] = True
        e["end"].record(e["tail"])
        main.wait_event(e["end"])
        main.wait_event(e["conv_done"][(len(sizes) - 1) & 1])

    def _encode_tail(self, tower, b, out, normalise):
        sem = tower.forward_tail(b, False)
        if not normalise:
            out.copy_(sem)
        elif out.dtype == torch.bfloat16:
            call("clsm_normalize_rows_bf16", _ptr(sem), b["S"], tower.H, _ptr(out), _stream())
        else:
            call("clsm_normalize_rows", _ptr(sem), b["S"], tower.H, _ptr(out), _stream())

    def encode_resident(self, tower, b, out, normalise=True):
        """one batch whose token ids are already in b["tok"], on the current stream"""
        tower.forward_conv(b, False)
        self._encode_tail(tower, b, out, normalise)
        return out

    def topk(self, query_vecs, corpus_vecs, k=100, id_base=0):
        """query_vecs (Q, H), corpus_vecs (N, H): normalised fp32 DEVICE tensors.
        Returns (scores (Q, k) fp32, ids (Q, k) int64) device tensors, order (score desc, id asc)."""
        Q, H = query_vecs.shape
        N = corpus_vecs.shape[0]
        k = int(k)
        if corpus_vecs.dtype == torch.bfloat16 and H in (64, 128):
            # tensor-core scan of a bf16 corpus (queries are rounded to bf16 too)
            qv = query_vecs if query_vecs.dtype == torch.bfloat16 else query_vecs.to(torch.bfloat16)
            qv = qv.contiguous()
            ws_bytes = lib.clsm_cos_topk_bf16_workspace_bytes(Q, N, k)
            ws = self._scratch("topk_tc", ws_bytes)
            sc = torch.empty((Q, k), dtype=torch.float32, device=self.device)
            ids = torch.empty((Q, k), dtype=torch.int64, device=self.device)
            call("clsm_cos_topk_bf16", _ptr(qv), Q, _ptr(corpus_vecs), N, H, k, int(id_base), _ptr(sc), _ptr(ids),
                 _ptr(ws), ws_bytes, _stream())
            return sc, ids
        if corpus_vecs.dtype == torch.bfloat16:
            # a width the tensor-core scan does not take (it wants H = 64 or 128; the reference's own hidden_dims is
            # 150, cnn_dssm_tf.py:38): the exact fp32 scan over fp32 views of the SAME bf16 vectors, one bounded slice
            # of the corpus at a time, merged in (score desc, id asc) order -- same answer, CUDA-core speed.
            qf = query_vecs.to(torch.bfloat16).float().contiguous()
            step = 1 << 20
            if N <= step:
                return self.topk(qf, corpus_vecs.float(), k=k, id_base=id_base)
            parts = [self.topk(qf, corpus_vecs[s:s + step].float(), k=k, id_base=id_base + s) for s in range(0, N, step)]
            return self.topk_merge(torch.stack([p_[0] for p_ in parts]), torch.stack([p_[1] for p_ in parts]))
        query_vecs = query_vecs.float().contiguous()
        ws_bytes = lib.clsm_cos_topk_workspace_bytes(Q, N, k)
        ws = self._scratch("topk", ws_bytes)
        sc = torch.empty((Q, k), dtype=torch.float32, device=self.device)
        ids = torch.empty((Q, k), dtype=torch.int64, device=self.device)
        call("clsm_cos_topk", _ptr(query_vecs), Q, _ptr(corpus_vecs), N, H, k, int(id_base), _ptr(sc), _ptr(ids),
             _ptr(ws), ws_bytes, _stream())
        return sc, ids

    def topk_merge(self, scores, ids):
        """scores/ids (n_parts, Q, k) device tensors -> merged (Q, k)."""
        n_parts, Q, k = scores.shape
        ws_bytes = lib.clsm_topk_merge_workspace_bytes(n_parts, Q, k)
        ws = self._scratch("merge", ws_bytes)
        sc = torch.empty((Q, k), dtype=torch.float32, device=self.device)
        out = torch.empty((Q, k), dtype=torch.int64, device=self.device)
        call("clsm_topk_merge", _ptr(scores.contiguous()), _ptr(ids.contiguous()), n_parts, Q, k, _ptr(sc), _ptr(out),
             _ptr(ws), ws_bytes, _stream())
        return sc, out

    def _scratch(self, name, nbytes):
        t = self._head.get(("ws", name))
        if t is None or t.numel() < nbytes:
            t = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=self.device)
            self._head[("ws", name)] = t
        return t

    # ---- persistence (cnn_dssm_tf.py:193,199-206)
    def to_json(self):
        return json.dumps({"class_name": "CLSM", "config": dict(self.params.to_dict(),
                           query_length=self.query_length, document_length=self.document_length,
                           vocab_rows=int(self.embedding_weights.shape[0]))})

    def get_weights(self):
        w = {"query": self.query_model.get_weights(), "doc": self.doc_model.get_weights()}
        if not self.shared_doc_tower:
            w["docs"] = [t.get_weights() for t in self.doc_models]       # 'doc' is docs[0], the positive-document tower
        return w

    def set_weights(self, w, theano_kernels=False):
        self.query_model.set_weights(w["query"], theano_kernels)
        if self.shared_doc_tower:
            self.doc_model.set_weights(w["doc"], theano_kernels)
        else:
            for t, wt in zip(self.doc_models, w["docs"]):
                t.set_weights(wt, theano_kernels)

    def save_weights(self, path):
        """model.save_weights (cnn_dssm_tf.py:206).  The container follows the extension: '.npz' = the dependency-free
        interchange with Keras-1 layer / weight names and kernel layouts (keras_io.py), '.h5' / '.hdf5' = the reference's
        own HDF5 layout (needs h5py), anything else = a torch file of plain tensors."""
        from . import keras_io
        kind = keras_io.container_of(path)
        if not self.shared_doc_tower:
            if kind != "torch":
                raise ValueError("the Keras-named containers describe the two-tower v2 model; save the v1 model as a torch file")
            w = self.get_weights()
            torch.save({"weights": {"query": {k: torch.as_tensor(v) for k, v in w["query"].items()},
                                    "docs": [{k: torch.as_tensor(v) for k, v in d.items()} for d in w["docs"]]}}, path)
            return
        if kind == "torch":
            w = self.get_weights()
            torch.save({"weights": {t: {k: torch.as_tensor(v) for k, v in w[t].items()} for t in w}}, path)
            return
        layers = keras_io.export_layers(self.get_weights(), len(self.query_model.windows))
        (keras_io.save_npz if kind == "npz" else keras_io.save_h5)(path, layers)

    def load_weights(self, path, theano_kernels=False):
        """Loads what save_weights wrote -- or, for '.npz' / '.h5', a checkpoint trained by the REFERENCE
        (ModelCheckpoint / save_weights / model.save of cnn_dssm_tf.py:193,199-206; theano_kernels=True for the
        Theano backend, whose kernels are applied flipped)."""
        from . import keras_io
        kind = keras_io.container_of(path)
        if kind == "torch":
            w = torch.load(path, weights_only=True)["weights"]        # tensors only: nothing in the file is executed
            if "docs" in w:
                self.set_weights({"query": {k: v.numpy() for k, v in w["query"].items()},
                                  "docs": [{k: v.numpy() for k, v in d.items()} for d in w["docs"]]}, theano_kernels)
            else:
                self.set_weights({t: {k: v.numpy() for k, v in w[t].items()} for t in w}, theano_kernels)
            return
        layers = keras_io.load_npz(path) if kind == "npz" else keras_io.load_h5(path)
        self.set_weights(keras_io.import_layers(layers, theano_kernels))

    def save(self, path):
        """Whole model: architecture JSON, weights, optimizer state and step (resume-able, which the reference's
        save is not).  A torch file of plain tensors / strings / ints (loadable with weights_only=True)."""
        if self.dp is not None:
            for tower in (self.query_model, self.doc_model):
                self.dp.gather_masters(tower)
        st = {"json": self.to_json(), "iterations": self.iterations,
              "query": {k: v.cpu() for k, v in self.query_model.state().items()},
              "doc": {k: v.cpu() for k, v in self.doc_model.state().items()}}
        if not self.shared_doc_tower:
            st["docs"] = [{k: v.cpu() for k, v in t.state().items()} for t in self.doc_models]
        torch.save(st, path)

    def load(self, path):
        st = torch.load(path, weights_only=True)
        self.iterations = int(st["iterations"])
        pairs = [(self.query_model, st["query"])]
        pairs += [(self.doc_model, st["doc"])] if self.shared_doc_tower else list(zip(self.doc_models, st["docs"]))
        for tower, saved in pairs:
            for name, t in tower.state(settle=False).items():          # whatever the old rows owed is overwritten
                t.copy_(saved[name])
            tower.mark_settled()
            tower.refresh_compute_copies(table_too=True)
        return self


class ModelCheckpoint(object):
    """callbacks.ModelCheckpoint(filepath) of cnn_dssm_tf.py:193: saves weights each epoch."""

    def __init__(self, filepath, verbose=0):
        self.filepath, self.verbose = filepath, verbose

    def __call__(self, model, epoch, logs):
        path = self.filepath.format(epoch=epoch, **logs)
        model.save_weights(path)
        if self.verbose:
            print("Epoch %05d: saving model to %s" % (epoch, path))
