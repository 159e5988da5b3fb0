"""Per-window device times of the conv backward (dW kernel | reductions | dX kernel) and forward for one BASELINE shape:
python scripts/bwd_split.py cfg4"""
import ctypes as C, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as BN
from dnn_page_vectors_b200 import _lib
import dnn_page_vectors_b200.model as M

cfg = {"cfg2": BN.CFG2, "cfg3": BN.CFG3, "cfg4": BN.CFG4}[sys.argv[1] if len(sys.argv) > 1 else "cfg4"]
class A: conv_impl = "auto"; lazy_adam = False
net = BN.build_net(cfg, 0, A, torch)
bq, bd, step = BN.stage_resident(net, cfg, BN.make_batches(cfg, 2, 7), torch)
for i in range(3):
    step(i)
torch.cuda.synchronize()
net.overlap = False
orig, rec = _lib.call, []
def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e
def timed(name, *a):
    mids = None
    if name == "clsm_conv_pool_bwd_parts":
        mids = (ev(), ev())
        _lib.lib.clsm_debug_mark_conv_bwd(C.c_void_p(mids[0].cuda_event), C.c_void_p(mids[1].cuda_event))
    s = ev(); r = orig(name, *a); e = ev()
    rec.append((name, s, e, mids, a))
    return r
M.call = timed
step(0)
torch.cuda.synchronize()
M.call = orig
for name, s, e, mids, a in rec:
    if name == "clsm_conv_pool_bwd_parts" and a[-4] == 3:
        print("bwd  S=%d k=%d: dW %.3f | reductions %.3f | dX %.3f ms" % (a[2], a[7], s.elapsed_time(mids[0]), mids[0].elapsed_time(mids[1]), mids[1].elapsed_time(e)))
    elif name == "clsm_conv_pool_bwd_parts":
        print("bwd  S=%d k=%d: dW + reductions %.3f ms" % (a[2], a[7], s.elapsed_time(e)))
    elif name == "clsm_conv_pool_bwd_dx_multi":
        print("bwd  S=%d all windows, fused dX %.3f ms" % (a[0], s.elapsed_time(e)))
    elif name in ("clsm_conv_pool_fwd_d", "clsm_emb_scatter_reduce", "clsm_emb_gather_mask", "clsm_adam_defer_rows", "clsm_adam_defer_settle_tokens", "clsm_dense_bwd"):
        print("%-28s %.3f ms" % (name[5:] + (" k=%d" % a[7] if name == "clsm_conv_pool_fwd_d" else ""), s.elapsed_time(e)))
