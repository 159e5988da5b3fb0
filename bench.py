#!/usr/bin/env python
"""bench.py -- CLSM training throughput (train pairs/s) on BASELINE.json configs[1]:
bf16 CLSM, D=300 w2v, window 3, 300 filters, 128-d semantic, J=4, 20-token query / 100-token page,
batch 4096 per GPU, 1M-word embedding table, dense Keras Adam.  One pair = 1 query + 1 positive +
J negatives through forward, loss, backward and the optimizer step.

  python bench.py [--gpus N --steps K --warmup W]          our arm (one process per GPU under torchrun)
  python bench.py --impl reference [...]                   CPU arm: torch-CPU port of the reference graph,
                                                           SAME batch, shapes and token stream as our arm

Prints ONE JSON line (rank 0).  Extra keys beyond the contract:
  step_ms          per-step device times of the timed region (median, p90, min, max)
  kernels          device time of every C-ABI call class in an instrumented one-stream step (median of 3)
  roofline(_all)   per-kernel achieved / peak, each launch timed directly with CUDA events (clsm_conv_pool_bwd is
                   split into its dW and dX kernels by two events recorded INSIDE the call); `traffic` comes from
                   the ncu capture summarised in profiles/r02_ncu_traffic.json
  other_configs    BASELINE configs[2] (J=50, 1000-token pages) and configs[3] (windows 1/2/3/5, batch 8192) on one GPU
  sharded          configs[4] at this N: page encoding over a row shard per GPU, per-GPU top-100 + allgather + merge
  cpu_baseline     the CPU port on the host cores (N = 1 only)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# activation: the reference's own literal is ReLU (cnn_dssm_tf.py:88,126).  north_star says "tanh" (the CLSM paper),
# but tanh + the reference's cosine clip to [0, 1] (cnn_dssm_tf.py:75-76) dies after one Adam step on random
# init: every cosine goes negative, is clipped, and all gradients are zero from then on (measured).
# A dead backward would be skipped work, so the benchmark trains the reference's ReLU model.
CFG1 = dict(V=50_000, D=300, windows=(3,), F=300, H=128, J=4, Lq=20, Ld=100, B=128, act="relu")
CFG2 = dict(V=1_000_000, D=300, windows=(3,), F=300, H=128, J=4, Lq=20, Ld=100, B=4096, act="relu")
CFG3 = dict(V=1_000_000, D=300, windows=(3,), F=300, H=128, J=50, Lq=20, Ld=1000, B=256, act="relu")
CFG4 = dict(V=1_000_000, D=300, windows=(1, 2, 3, 5), F=300, H=128, J=4, Lq=20, Ld=100, B=8192, act="relu")
WORKLOAD = "cfg2: CLSM bf16 train step, B=4096/GPU, J=4, Lq=20, Ld=100, D=300, window 3, F=300, H=128, V=1M, dense Adam"
SHARD_PAGES = 2_500_000          # configs[4]: 20 M pages over 8 GPUs
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm": p["hbm_gbs"], "tf_burst": p["bf16_tflops"], "tf_sust": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "src": "measured (MEASURED_PEAKS.json)"}
    except Exception:
        return {"hbm": 6650.0, "tf_burst": 1590.0, "tf_sust": 1400.0, "src": "fallback (B200_PROFILING.md)"}


def conv_flops_seq(c, L):
    return sum(2 * (L - k + 1) * k * c["D"] * c["F"] for k in c["windows"])


def flops_pair_fwd(c):
    def seq(L):
        return conv_flops_seq(c, L) + 2 * len(c["windows"]) * c["F"] * c["H"]
    return seq(c["Lq"]) + (1 + c["J"]) * seq(c["Ld"])


_ZIPF = {}


def zipf_tokens(rng, V, n, L, s=1.07):
    """Zipf(s) ids over [0, V), true length ~ U[L/4, L], right-padded with PAD id V-1.
    The frequency rank -> id map is a fixed random permutation: vocabulary ids are assigned in sorted-word
    order (data_helpers.py:109-118 enumerates a set; our DataHelpers sorts), never in frequency order, so the hot
    rows are spread over the whole table rather than packed at its head."""
    if (V, s) not in _ZIPF:
        ranks = np.arange(1, V + 1, dtype=np.float64)
        cdf = np.cumsum(ranks ** (-s))
        cdf /= cdf[-1]
        perm = np.random.RandomState(20261106).permutation(V).astype(np.int32)
        _ZIPF[(V, s)] = (cdf, perm)
    cdf, perm = _ZIPF[(V, s)]
    tok = perm[np.searchsorted(cdf, rng.random_sample((n, L)))]
    lens = rng.randint(max(L // 4, 1), L + 1, size=n)
    tok[np.arange(L)[None, :] >= lens[:, None]] = V - 1
    return tok


def make_batches(c, n_batches, seed):
    rng = np.random.RandomState(seed)
    out = []
    for _ in range(n_batches):
        q = zipf_tokens(rng, c["V"], c["B"], c["Lq"])
        docs = [zipf_tokens(rng, c["V"], c["B"], c["Ld"]) for _ in range(1 + c["J"])]
        out.append([q] + docs)
    return out


class ClockSampler(object):
    """nvidia-smi clocks + throttle reasons every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------ reference arm
def run_reference(args, rank, world):
    """The reference graph (dssm_cnn_v2/cnn_dssm_tf.py:63-182) on the host cores: oracle/torch_cpu_port.py, all
    threads, on the SAME workload as our arm -- cfg2 shapes, batch --batch (4096) per step, the same Zipf token
    stream, --steps / --warmup honoured.  pairs/s is not batch invariant (every step carries a dense Adam over two
    1M x 300 tables), which is why the batch must match.  A wall-clock budget (--ref-budget-s) bounds the run: when it
    cuts the timed steps short the line says so in `steps` / `requested_steps`."""
    if rank != 0:
        return 0
    from oracle import torch_cpu_port as CP
    c = dict(CFG2, B=args.batch, V=args.vocab)
    steps, warmup = max(args.steps, 1), max(args.warmup, 0)
    batches = make_batches(c, 4, 1337)
    pps, threads, dt, done = CP.time_train(c, c["B"], steps, warmup, batches=batches, budget_s=args.ref_budget_s)
    sample = "torch-CPU fp32 port of the reference graph, cfg2 shapes, B=%d per step, Zipf token stream of the GPU arm, %d timed steps" % (c["B"], done)
    line = {
        "impl": "reference", "metric": "train_pairs_per_s", "value": pps, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": done, "requested_steps": steps, "warmup": warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": c["B"], "activation": c["act"],
                   "tokens": "Zipf(1.07) ids (frequency rank -> id by a fixed random permutation), true length U[L/4,L], right-padded with one PAD id",
                   "sample": sample},
        "cpu_baseline": {"value": pps, "unit": "pairs/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": pps, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference stack (Keras 1.0.x / Theano 0.8 / TF 0.x, Python 2) not installable here; oracle port timed",
    }
    if done < steps:
        line["note"] += "; timed steps cut from %d to %d by the %.0f s budget" % (steps, done, args.ref_budget_s)
    emit(line)
    return 0


# ------------------------------------------------------------------------------ our arm
def build_net(c, local_rank, args, torch):
    from dnn_page_vectors_b200 import CLSM, ModelParams
    mp = ModelParams(embedding_dim=c["D"], batch_size=c["B"], filter_sizes=c["windows"], num_filters=c["F"],
                     hidden_dims=c["H"], J=c["J"], dropout_prob=(0.25, 0.5), activation=c["act"], precision="bf16")
    rng = np.random.RandomState(1337)
    table = rng.uniform(-0.25, 0.25, size=(c["V"] + 1, c["D"])).astype(np.float32)   # data_helpers.py:62
    table[-1] = 0.0
    net = CLSM(mp, c["Lq"], c["Ld"], embedding_weights=table, device="cuda:%d" % local_rank,
               conv_impl=args.conv_impl, lazy_table_adam=args.lazy_adam).compile()
    return net


def stage_resident(net, c, batches, torch):
    B = c["B"]
    bq = net.query_model.buffers(B, True)
    bd = net.doc_model.buffers(B * (1 + c["J"]), True)
    dev_batches = []
    for x in batches:
        q = torch.as_tensor(x[0]).cuda()
        d = torch.as_tensor(np.stack(x[1:], axis=1)).cuda().view(B * (1 + c["J"]), c["Ld"])
        dev_batches.append((q, d))

    def resident_step(i):
        q, d = dev_batches[i % len(dev_batches)]
        bq["tok"].copy_(q)
        bd["tok"].copy_(d)
        return net.step_resident(B, bq, bd)
    return bq, bd, resident_step


def pct(xs, q):
    xs = sorted(xs)
    return float(xs[min(len(xs) - 1, int(round(q * (len(xs) - 1))))])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=CFG2["B"])
    ap.add_argument("--vocab", type=int, default=CFG2["V"])
    ap.add_argument("--conv-impl", default="auto")
    ap.add_argument("--lazy-adam", action="store_true", help="row-sparse table Adam (NOT the reference's semantics)")
    ap.add_argument("--no-extras", action="store_true", help="skip kernels / roofline / other_configs / sharded / cpu_baseline")
    ap.add_argument("--dp-profile", action="store_true", help="N>1: device time of exchange / Adam / allgather per step, to stderr")
    ap.add_argument("--ref-budget-s", type=float, default=420.0, help="--impl reference: wall-clock budget of the timed steps")
    ap.add_argument("--shard-pages", type=int, default=SHARD_PAGES)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    from dnn_page_vectors_b200._lib import lib

    assert torch.cuda.is_available(), "bench.py needs a CUDA device"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    warmup = max(args.warmup, 3)
    steps = max(args.steps, 1)
    c = dict(CFG2, B=args.batch, V=args.vocab)
    net = build_net(c, local_rank, args, torch)
    if world > 1:
        from dnn_page_vectors_b200.parallel import DataParallel
        DataParallel(net)
    # a DISTINCT batch for every warm-up and timed step (up to 32): which table rows a step touches, and how many steps
    # the others owe, is then that of a fresh token stream, not of a 4-batch cycle
    batches = make_batches(c, min(max(warmup + steps, 4), 32), 1337 + rank)
    B = c["B"]
    deferred = world == 1 and net.defer_table_adam and not args.lazy_adam
    if deferred:
        # every table row counts as carrying optimizer state (what a loaded checkpoint, or a long run, looks like):
        # each of its zero-gradient steps is then really executed -- in a later step or in the settle that closes the
        # timed region -- instead of being skipped as a no-op on zero m, v
        net.assume_optimizer_state()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: tokens resident in HBM before the timed region
    bq, bd, resident_step = stage_resident(net, c, batches, torch)
    for i in range(warmup):
        resident_step(i)
    net.settle()                     # the warm-up's owed row steps are paid before the clock starts
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = lib.clsm_launch_count()
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    marks[0].record()
    for i in range(steps):
        resident_step(warmup + i)
        marks[i + 1].record()
    net.settle()                     # INSIDE the timed region: every row step the K steps deferred is taken here at the latest
    m_end = torch.cuda.Event(enable_timing=True)
    m_end.record()
    barrier()
    launches = int(lib.clsm_launch_count() - l0)
    ms = marks[0].elapsed_time(m_end)
    settle_ms = marks[steps].elapsed_time(m_end)
    per_step = [marks[i].elapsed_time(marks[i + 1]) for i in range(steps)]
    # ---- e2e: host batches in PINNED memory through the public API (train_on_batch): ids H2D and loss D2H
    # inside every step
    pinned = [[torch.as_tensor(a).pin_memory() for a in x] for x in batches]
    for i in range(2):
        net.train_on_batch(pinned[i % len(pinned)])
    net.settle()
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    last = None
    for i in range(steps):
        last = net.train_on_batch(pinned[i % len(pinned)])
    net.settle()
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms, ms_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    pairs = B * world * steps
    value = pairs / (ms * 1e-3)
    e2e_value = pairs / (ms_e2e * 1e-3)
    tok_bytes = 4 * B * (c["Lq"] + (1 + c["J"]) * c["Ld"])

    line = {
        "metric": "train_pairs_per_s", "value": value, "unit": "pairs/s", "n_gpus": world, "steps": steps,
        "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": B * world, "activation": c["act"], "dropout": 0.25,
                   "table_adam": ("lazy-rows" if args.lazy_adam else
                                  "dense (Keras semantics); zero-gradient row steps deferred and replayed bit-exactly, every row treated as "
                                  "carrying optimizer state, the settle of all rows INSIDE the timed region (settle_ms)" if deferred else
                                  "dense (Keras semantics)"),
                   "distinct_batches": len(batches),
                   "conv_impl": args.conv_impl, "parallelism": "dp%d" % world,
                   "streams": ("three after the loss (query tower | document dX then conv weight gradients | document scatter + table optimizer beside those)"
                               if (world == 1 and net.overlap and net.split_bwd) else
                               "two (query tower's backward/scatter/Adam beside the document tower's)" if (world == 1 and net.overlap) else "one"),
                   "l2": ("inputs larger than L2: a step writes and re-reads the 1.3 GB gathered matrix and its 1.3 GB gradient and touches "
                          "~0.2 M table rows with their Adam state; a different batch every step" if deferred else
                          "inputs larger than L2: each step streams both 1M x 300 tables and their Adam state (14 GB)"),
                   "tokens": "Zipf(1.07) ids (frequency rank -> id by a fixed random permutation), true length U[L/4,L], right-padded with one PAD id"},
        "step_ms": {"median": pct(per_step, 0.5), "p90": pct(per_step, 0.9), "min": min(per_step), "max": max(per_step),
                    "note": "rank 0's device time per step of the timed region (events between steps)"},
        "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": tok_bytes, "d2h_bytes_per_step": 8,
                "ms_per_step": ms_e2e / steps, "loss": last[0] if last else None},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if deferred:
        line["settle_ms"] = settle_ms
        line["step_ms"]["note"] += "; ms_per_step = (the K steps + the closing settle of every table row, settle_ms) / K"
        if not args.no_extras:
            # the same steps with the DENSE table pass (every row of both tables streamed every step, round 1's and
            # early round 2's optimizer; fp32-master gather, no bf16 copy written): what deferring the row steps buys
            net.settle()
            towers = (net.query_model, net.doc_model)
            was = [(t.lazy_copy, ) for t in towers]
            net.defer_table_adam = False
            for t in towers:
                t.lazy_copy = t.bf16 and t.D % 4 == 0
            for i in range(3):
                resident_step(i)
            torch.cuda.synchronize()
            d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n_dense = min(steps, 10)
            d0.record()
            for i in range(n_dense):
                resident_step(3 + i)
            d1.record()
            torch.cuda.synchronize()
            dms = d0.elapsed_time(d1) / n_dense
            line["dense_table_pass"] = {"ms_per_step": dms, "pairs_per_s": B / (dms * 1e-3), "steps": n_dense,
                                        "note": "same model and batches, defer_table_adam=False: clsm_adam_table over all rows of both tables each step"}
            net.defer_table_adam = True
            for t, (lc,) in zip(towers, was):
                t.lazy_copy = lc
                t.refresh_compute_copies(table_too=True)

    if world > 1 and args.dp_profile:
        dp_profile(net, resident_step, torch, rank)

    if not args.no_extras:
        pk = peaks()
        # ---- per-call device times of instrumented ONE-STREAM steps.  Every rank runs them (the steps contain
        # collectives); rank 0 reports its own.
        phases, launch = profile_phases(net, resident_step, torch)
        if rank == 0:
            line["kernels"] = phases
            from dnn_page_vectors_b200._lib import BF16 as _BF16
            fill_roofline(line, c, launch, clocks, pk, steps, ms, dx_bytes=2 if bd.get("dx_dtype") == _BF16 else 4,
                          n_uniq_doc=int(bd["n_uniq"].item()))
        # ---- configs[4] at this N: sharded encode and retrieval
        try:
            line_sh = bench_sharded(net, c, torch, dist, pk, rank, world, args.shard_pages)
            if rank == 0:
                line["sharded"] = line_sh
        except Exception as e:          # keep the headline line even if the extra leg fails
            if rank == 0:
                line["sharded"] = {"error": repr(e)}
        if world == 1:
            del bq, bd, resident_step
            net = None
            torch.cuda.empty_cache()
            line["other_configs"] = {}
            for name, cc in (("cfg3 (J=50, 1000-token pages, B=256/GPU)", CFG3), ("cfg4 (windows 1/2/3/5, 4x300 filters, B=8192)", CFG4)):
                try:
                    line["other_configs"][name] = bench_other(cc, local_rank, args, torch, pk)
                except Exception as e:
                    line["other_configs"][name] = {"error": repr(e)}
                torch.cuda.empty_cache()
            line["cpu_baseline"] = cpu_baseline(c, batches)
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def fill_roofline(line, c, launch, clocks, pk, steps, ms, dx_bytes=4, n_uniq_doc=0):
    """launch: {kernel: directly measured device ms of ONE launch of the document tower (or of one table)}.
    dx_bytes: element size of the dX rows the scatter reads (2 when the tensor-core backward writes them as bf16)."""
    B = c["B"]
    n_doc = B * (1 + c["J"]) * c["Ld"]
    Dpad = (c["D"] + 63) // 64 * 64
    conv_doc_tf = conv_flops_seq(c, c["Ld"]) * B * (1 + c["J"]) / 1e12
    dense_doc_tf = 2 * len(c["windows"]) * c["F"] * c["H"] * B * (1 + c["J"]) / 1e12
    n_table = (c["V"] + 1) * c["D"]
    try:
        with open(TRAFFIC_FILE) as f:
            traffic = json.load(f)
    except Exception:
        traffic = {}
    # the bf16 denominator: the burst figure unless the clock record of the timed region shows the power cap (then the
    # sustained one); both fractions are printed
    capped = bool(clocks and "sw_power_cap" in (clocks.get("reasons") or []))
    cand = {
        # kernel: (bound, algorithmic work per launch [TFLOP or GB], what the work is)
        "conv_pool_tc": ("tensor", conv_doc_tf, "2 T k D F per sequence, document tower"),
        "dw2_tc": ("tensor", conv_doc_tf, "dense-equivalent GEMM of the conv weight gradient (= forward FLOPs)"),
        "dx2_tc": ("tensor", conv_doc_tf, "dense-equivalent GEMM of the conv input gradient (= forward FLOPs)"),
        "dense_fwd": ("tensor", dense_doc_tf, "2 |K|F H per sequence"),
        "dense_bwd": ("tensor", 2 * dense_doc_tf, "dWd and G GEMMs"),
        "adam_table": ("hbm", 24.0 * n_table / 1e9, "24 B per parameter of one 1M x 300 table (p, m, v read + write)"),
        "adam_defer_rows": ("hbm", 28.0 * n_uniq_doc * c["D"] / 1e9, "28 B per parameter of the %d document-table rows of the batch (p, m, v read + write, gradient row read)" % n_uniq_doc),
        # SURVEY 8(d) figures.  gather: ids + table rows read; "the write is not counted when fused" -- it is NOT fused
        # here (the kernel must write X, Dpad bf16 per token), so both accountings are printed (frac_with_write below).
        "emb_gather": ("hbm", n_doc * (4 + c["D"] * 2) / 1e9, "SURVEY 8(d): ids + table rows read (the X write, which this standalone kernel also does, is not counted)"),
        # scatter: SURVEY 8(d) counts dX as D*4 bytes per token (+ 8 B key/index); this build's single-window backward
        # hands the rows over as bf16 (dx_bytes = 2), i.e. it moves LESS than the algorithmic figure
        "emb_scatter": ("hbm", n_doc * (c["D"] * 4 + 8) / 1e9, "SURVEY 8(d): tokens x (D x 4 + 8) B; rows actually read with %d B per element (clsm_emb_scatter_reduce; the id-only plan runs under the forward)" % dx_bytes),
    }
    out = {}
    for k, (bound, work, what) in cand.items():
        tms = launch.get(k)
        if not tms:
            continue
        ach = work / (tms * 1e-3)
        if bound == "tensor":
            e = {"bound": bound, "achieved": ach, "unit": "TFLOP/s", "peak": pk["tf_sust"] if capped else pk["tf_burst"],
                 "frac_of_burst_peak": ach / pk["tf_burst"], "frac_of_sustained_peak": ach / pk["tf_sust"]}
        else:
            e = {"bound": bound, "achieved": ach, "unit": "GB/s", "peak": pk["hbm"]}
        e["frac"] = ach / e["peak"]
        e["launch_ms"] = tms
        e["algorithmic_work_per_launch"] = work
        e["work"] = what
        e["traffic"] = traffic.get(k)
        if k == "emb_gather":
            e["frac_with_write"] = (n_doc * (4 + c["D"] * 2 + Dpad * 2) / 1e9) / (tms * 1e-3) / pk["hbm"]
        if k == "emb_scatter":
            e["frac_of_bytes_actually_read"] = (n_doc * (c["D"] * dx_bytes + 8) / 1e9) / (tms * 1e-3) / pk["hbm"]
        out[k] = e
    line["roofline_all"] = out
    if out:
        top = max(out, key=lambda k: out[k]["launch_ms"])
        r = dict(out[top])
        r["kernel"] = top
        r["peak_source"] = pk["src"]
        r["traffic_unit"] = ("GB per launch (dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full, %s)"
                             % os.path.relpath(TRAFFIC_FILE, ROOT)) if r.get("traffic") else None
        how = "launch_ms = CUDA events around this kernel's launch in a one-stream instrumented step, median of 3; "
        if r["bound"] == "tensor":
            how += ("bf16 peak = sustained figure (sw_power_cap seen in the timed region)" if capped
                    else "bf16 peak = burst figure (no power-cap sample in the timed region)")
        else:
            how += "peak = measured HBM copy bandwidth"
        r["note"] = how
        line["roofline"] = r
    line["train_flops_per_pair_3x"] = 3 * flops_pair_fwd(c)
    line["model_tflops_3x"] = 3 * flops_pair_fwd(c) * B * (steps / (ms * 1e-3)) / 1e12


def cpu_baseline(c, batches):
    """Rank 0, N = 1: the CPU port on the host cores.  cfg2 at the GPU arm's own batch (a bounded sample: 1 warm-up +
    2 timed steps), and BASELINE.json configs[0] (cfg1: B=128, V=50k -- the reference's own CPU-runnable case)."""
    try:
        from oracle import torch_cpu_port as CP
        pps, threads, dt, done = CP.time_train(dict(c), c["B"], 2, 1, batches=batches, budget_s=60.0)
        out = {"value": pps, "unit": "pairs/s", "cores": threads, "kind": "port",
               "sample": "torch-CPU fp32 port of the reference graph, cfg2 shapes, B=%d per step (same batch and Zipf tokens as the "
                         "GPU arm), 1 warm-up + %d timed steps, %.1f s per step" % (c["B"], done, dt)}
        try:
            c1 = dict(CFG1)
            p1, _, dt1, n1 = CP.time_train(c1, c1["B"], 10, 2, batches=make_batches(c1, 2, 1337), budget_s=20.0)
            out["cfg1"] = {"value": p1, "unit": "pairs/s", "sample": "BASELINE configs[0]: B=128, V=50k, %d timed steps, %.3f s per step" % (n1, dt1)}
        except Exception as e:
            out["cfg1"] = {"error": repr(e)}
        return out
    except Exception as e:
        return {"value": None, "unit": "pairs/s", "cores": os.cpu_count(), "kind": "port", "sample": "failed: %r" % (e,)}


def dp_profile(net, resident_step, torch, rank):
    """Device time of the data-parallel phases of a step (5 steps, median): everything before the exchange
    (forward + backward), the gradient exchange, the sharded Adam, the compute-row allgather."""
    marks = []

    def wrap(obj, name, tag):
        orig = getattr(obj, name)

        def f(*a, **k):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            r = orig(*a, **k)
            e.record()
            marks.append((tag, s, e))
            return r
        setattr(obj, name, f)
        return orig
    import dnn_page_vectors_b200.parallel as PAR
    was_pipe, net.dp_pipeline = net.dp_pipeline, False          # phases are timed on the one-stream order
    o0 = wrap(PAR, "allreduce_flat", "allreduce_small (inside exchange)")
    o1 = wrap(net.dp, "exchange", "exchange")
    o2 = wrap(net.dp, "after_adam", "allgather")
    o3 = wrap(net.query_model, "adam", "adam")
    o4 = wrap(net.doc_model, "adam", "adam")
    runs = []
    for i in range(5):
        marks.clear()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        resident_step(i)
        s1.record()
        torch.cuda.synchronize()
        agg = {"step": s0.elapsed_time(s1)}
        for tag, s, e in marks:
            agg[tag] = agg.get(tag, 0.0) + s.elapsed_time(e)
        agg["fwd_bwd"] = min(s0.elapsed_time(m[1]) for m in marks)
        runs.append(agg)
    net.dp.exchange, net.dp.after_adam, net.query_model.adam, net.doc_model.adam = o1, o2, o3, o4
    PAR.allreduce_flat = o0
    net.dp_pipeline = was_pipe
    med = {k: round(float(np.median([r[k] for r in runs])), 3) for k in runs[0]}
    sys.stderr.write("[dp profile, rank %d, ms] %s\n" % (rank, json.dumps(med)))


def profile_phases(net, resident_step, torch):
    """Device time of each C-ABI call over 3 instrumented steps (median), with the two-stream overlap switched off so
    that every launch is timed alone on the stream.  Returns (per-call-class totals per step, per-launch times of the
    document tower: for a class called once per tower the LARGER launch is the document tower's).  clsm_conv_pool_bwd
    is split into dW kernel / reductions / dX kernel by two events the library records inside the call."""
    import ctypes as C
    from dnn_page_vectors_b200 import _lib
    import dnn_page_vectors_b200.model as M
    orig = _lib.call
    rec = []

    def new_event():
        e = torch.cuda.Event(enable_timing=True)
        e.record()                          # creates the underlying cudaEvent_t
        return e

    def timed(name, *a):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        mids = None
        if name in ("clsm_conv_pool_bwd", "clsm_conv_pool_bwd_ex", "clsm_conv_pool_bwd_parts"):
            mids = (new_event(), new_event())
            _lib.lib.clsm_debug_mark_conv_bwd(C.c_void_p(mids[0].cuda_event), C.c_void_p(mids[1].cuda_event))
        s.record()
        r = orig(name, *a)
        e.record()
        rec.append((name, s, e, mids, a))
        return r
    was = net.overlap
    net.overlap = False
    M.call = timed
    totals, launches = [], []
    try:
        for i in range(3):
            rec.clear()
            resident_step(i)
            torch.cuda.synchronize()
            agg, per = {}, {}
            for name, s, e, mids, a in rec:
                key = name.replace("clsm_", "")
                d = s.elapsed_time(e)
                agg[key] = agg.get(key, 0.0) + d
                per.setdefault(key, []).append(d)
                if mids is not None and (name != "clsm_conv_pool_bwd_parts" or a[-4] == 3):
                    per.setdefault("dw2_tc", []).append(s.elapsed_time(mids[0]))
                    per.setdefault("conv_bwd_reductions", []).append(mids[0].elapsed_time(mids[1]))
                    per.setdefault("dx2_tc", []).append(mids[1].elapsed_time(e))
            totals.append(agg)
            launches.append({k: max(v) for k, v in per.items()})
    finally:
        M.call = orig
        net.overlap = was
    keys = sorted(set().union(*totals), key=lambda k: min(list(r).index(k) if k in r else 99 for r in totals))
    out = {k + "_ms": float(np.median([r.get(k, 0.0) for r in totals])) for k in keys}
    launch = {k: float(np.median([r.get(k, 0.0) for r in launches])) for k in set().union(*launches)}
    rename = {"conv_pool_fwd": "conv_pool_tc", "conv_pool_fwd_d": "conv_pool_tc", "emb_gather_mask": "emb_gather", "emb_gather_master": "emb_gather", "conv_pool_bwd_ex": "conv_pool_bwd", "conv_pool_bwd_parts": "conv_pool_bwd",
              "emb_scatter_reduce": "emb_scatter"}
    launch = {rename.get(k, k): v for k, v in launch.items()}
    out["doc_tower_launch_ms"] = {k: round(v, 4) for k, v in launch.items()}
    return out, launch


def bench_other(cc, local_rank, args, torch, pk):
    """One of the other BASELINE training shapes on one GPU: ms/step, pairs/s, and the fraction of the roofline-implied
    floor of SURVEY 8(d) (3x forward FLOPs at the sustained bf16 peak -- these steps run for seconds)."""
    net = build_net(cc, local_rank, args, torch)
    n = 5
    batches = make_batches(cc, 3 + n, 4242)          # the headline's protocol: a batch per step, every row carries
    if net.defer_table_adam and not args.lazy_adam:  # optimizer state, the settle of all rows inside the timed region
        net.assume_optimizer_state()
    bq, bd, step = stage_resident(net, cc, batches, torch)
    for i in range(3):
        step(i)
    net.settle()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        step(3 + i)
    net.settle()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    flops3 = 3 * flops_pair_fwd(cc) * cc["B"]
    floor_ms = flops3 / (pk["tf_sust"] * 1e12) * 1e3
    phases, launch = profile_phases(net, step, torch)
    out = {"ms_per_step": ms, "pairs_per_s": cc["B"] / (ms * 1e-3), "batch": cc["B"], "model_tflops_3x": flops3 / (ms * 1e-3) / 1e12,
           "floor_ms_3x_flops_at_sustained_bf16_peak": floor_ms, "frac_of_floor": floor_ms / ms,
           "kernels_ms": {k: round(v, 3) for k, v in phases.items() if isinstance(v, float)}}
    del net, bq, bd, step
    return out


def bench_sharded(net, c, torch, dist, pk, rank, world, shard_pages):
    """BASELINE configs[4] at this N.  Encode: each rank encodes its contiguous row shard (shard_pages pages of Ld tokens,
    no communication) into bf16 page vectors.  Retrieval: each rank scans its shard (tensor-core scan, k = 100), the
    (Q, k) lists are allgathered and merged in (score desc, id asc) order.  Rank 0 also scans the whole corpus alone and
    checks that the merged ids are IDENTICAL.  Rates are whole-job (all ranks), times the max over ranks."""
    H, Ld, k = c["H"], c["Ld"], 100
    dev = torch.device("cuda", torch.cuda.current_device())

    def maxr(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])
    # ---- encode: 8 distinct resident batches of 16384 pages, cycled until the shard is done
    NB = 16384
    rng = np.random.RandomState(7 + rank)
    pool = [torch.as_tensor(zipf_tokens(rng, c["V"], NB, Ld)).to(dev) for _ in range(8)]
    outs = [torch.empty((NB, H), dtype=torch.bfloat16, device=dev) for _ in range(2)]
    n_calls = max(shard_pages // NB, 1)

    def fill(i, b):
        b["tok"].copy_(pool[i % len(pool)])
    # two batches in flight (CLSM.encode_batches): batch i's convolution beside batch i-1's finalize / dense / normalise
    net.encode_batches(net.doc_model, [NB] * 3, fill, lambda i: outs[i & 1])
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    net.encode_batches(net.doc_model, [NB] * n_calls, fill, lambda i: outs[i & 1])
    e1.record()
    torch.cuda.synchronize()
    enc_ms = maxr(e0.elapsed_time(e1))
    pages = n_calls * NB
    page_flops = conv_flops_seq(c, Ld) + 2 * len(c["windows"]) * c["F"] * H
    enc = {"value": pages * world / (enc_ms * 1e-3), "unit": "pages/s", "pages_per_gpu": pages, "pages_per_call": NB,
           "ms_per_call": enc_ms / n_calls, "per_gpu_pages_per_s": pages / (enc_ms * 1e-3), "out_dtype": "bf16",
           "frac_of_tensor_floor_sustained": (pages * page_flops / (enc_ms * 1e-3) / 1e12) / pk["tf_sust"]}
    del pool

    # ---- retrieval: shard r = unit vectors from seed 1000 + r
    def shard(r):
        g = torch.Generator(device=dev).manual_seed(1000 + r)
        return torch.nn.functional.normalize(torch.randn((shard_pages, H), generator=g, device=dev), dim=1).to(torch.bfloat16)
    g = torch.Generator(device=dev).manual_seed(3)
    qall = torch.nn.functional.normalize(torch.randn((1024, H), generator=g, device=dev), dim=1).to(torch.bfloat16)
    corpus = shard(rank)
    res = {}
    merged_ids = {}
    for Q in (128, 1024):
        qv = qall[:Q].contiguous()

        def one():
            sc, ids = net.topk(qv, corpus, k=k, id_base=rank * shard_pages)
            m0 = torch.cuda.Event(enable_timing=True)
            m0.record()
            if world > 1:
                all_s = sc.new_empty((world,) + tuple(sc.shape))
                all_i = ids.new_empty((world,) + tuple(ids.shape))
                dist.all_gather_into_tensor(all_s, sc.contiguous())
                dist.all_gather_into_tensor(all_i, ids.contiguous())
                sc, ids = net.topk_merge(all_s, all_i)
            return sc, ids, m0
        for _ in range(2):
            one()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        reps = 5
        scan_ms, merge_ms = [], []
        for _ in range(reps):
            a = torch.cuda.Event(enable_timing=True)
            a.record()
            sc, ids, m0 = one()
            z = torch.cuda.Event(enable_timing=True)
            z.record()
            torch.cuda.synchronize()
            scan_ms.append(a.elapsed_time(m0))
            merge_ms.append(m0.elapsed_time(z))
        tot = maxr(float(np.sum(scan_ms) + np.sum(merge_ms))) / reps
        scan = maxr(float(np.median(scan_ms)))
        blocks = (Q + 127) // 128
        res["q%d" % Q] = {"value": Q / (tot * 1e-3), "unit": "queries/s", "queries_per_call": Q, "k": k, "corpus_rows_per_gpu": shard_pages,
                          "corpus_rows_total": shard_pages * world, "corpus_dtype": "bf16", "ms_per_call": tot, "scan_ms": scan,
                          "allgather_merge_ms": maxr(float(np.median(merge_ms))),
                          "hbm_frac_of_one_scan_per_query_block": (blocks * shard_pages * H * 2 / 1e9) / (scan * 1e-3) / pk["hbm"]}
        merged_ids[Q] = ids
    # ---- the merged answer against one GPU scanning the whole corpus
    if world > 1:
        check = None
        del corpus
        if rank == 0:
            whole = torch.cat([shard(r) for r in range(world)], dim=0)
            _, ids1 = net.topk(qall[:128].contiguous(), whole, k=k, id_base=0)
            check = bool(torch.equal(ids1, merged_ids[128]))
            del whole
        res["merged_ids_identical_to_single_gpu_scan_of_whole_corpus"] = check
        dist.barrier()
    return {"encode": enc, "topk": res}


_RESULT_FD = None


def emit(line):
    """The ONE JSON line of the contract, written to the process's original stdout."""
    text = json.dumps(line) + "\n"
    if _RESULT_FD is None:
        sys.stdout.write(text)
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, text.encode())


if __name__ == "__main__":
    # Libraries write banners to fd 1 (NCCL prints its version there): keep stdout for the result line only.
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    sys.exit(main())
