"""ctypes binding of libclsm_b200.so (the C ABI declared in include/clsm_b200.h).

There is deliberately no fallback: if the shared library is missing or a symbol does not
resolve, importing this module raises.  Build it with ``python -c "import __graft_entry__ as g;
g.build()"`` or ``make -C dnn_page_vectors_b200/csrc``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libclsm_b200.so")

OK, EINVAL, ECUDA, EWORKSPACE, EUNSUPPORTED = 0, -1, -2, -3, -4
F32, BF16 = 0, 1
ACT = {"linear": 0, "relu": 1, "tanh": 2}
MATH_TF32 = 0x100          # OR into the act argument of clsm_dense_fwd / clsm_dense_bwd (include/clsm_b200.h)
IMPL = {"auto": 0, "simt": 1, "tcgen05": 2}

i32, i64, u64, f32, vp, sz = C.c_int, C.c_int64, C.c_uint64, C.c_float, C.c_void_p, C.c_size_t

# name -> (restype, argtypes); every symbol include/clsm_b200.h declares
SIGNATURES = {
    "clsm_version": (i32, []),
    "clsm_last_error": (C.c_char_p, []),
    "clsm_num_sms": (i32, []),
    "clsm_launch_count": (C.c_ulonglong, []),
    "clsm_debug_mark_conv_bwd": (None, [vp, vp]),
    "clsm_emb_gather": (i32, [vp, i64, vp, i64, i32, i32, i32, vp, i32, f32, u64, vp]),
    "clsm_emb_gather_h": (i32, [vp, vp, i64, vp, i64, i32, i32, i32, vp, i32, f32, u64, vp]),
    "clsm_dropout_mask": (i32, [i64, i32, f32, u64, vp, vp]),
    "clsm_emb_gather_mask": (i32, [vp, i64, vp, i64, i32, i32, i32, vp, i32, f32, u64, vp, vp]),
    "clsm_dropout_scale": (f32, [f32]),
    "clsm_emb_gather_master": (i32, [vp, i64, vp, i64, i32, vp, i32, f32, u64, vp, vp]),
    "clsm_copy_rows_strided": (i32, [vp, sz, vp, sz, sz, i64, vp]),
    "clsm_dropout_rows": (i32, [vp, i64, i32, f32, u64, vp, vp]),
    "clsm_check_token_ids": (i32, [vp, i64, i64, vp, vp]),
    "clsm_conv_bwd_dx_bf16_ok": (i32, [i64, i32, i32, i32, i32]),
    "clsm_conv_pool_bwd_ex": (i32, [vp, i32, i64, i32, i32, i32, vp, i32, i32, i32, vp, vp, i32, i32,
                                    vp, vp, vp, i32, i32, vp, sz, vp]),
    "clsm_conv_pool_bwd_parts": (i32, [vp, i32, i64, i32, i32, i32, vp, i32, i32, i32, vp, vp, i32, i32,
                                       vp, vp, vp, i32, i32, i32, vp, sz, vp]),
    "clsm_conv_bwd_dx_multi_ok": (i32, [i64, i32, i32, vp, i32, i32]),
    "clsm_conv_bwd_dx_multi_workspace_bytes": (sz, [i32, vp, i32, i32]),
    "clsm_conv_pool_bwd_dx_multi": (i32, [i64, i32, i32, i32, vp, vp, i32, i32, vp, vp, i32, i32, vp, i32, vp, sz, vp]),
    "clsm_emb_scatter_plan": (i32, [vp, i64, i32, i32, i64, vp, vp, vp, vp, sz, vp]),
    "clsm_emb_scatter_reduce": (i32, [i64, vp, i32, i32, i32, f32, u64, vp, vp, vp, sz, vp]),
    "clsm_pack_conv_w": (i32, [vp, i32, i32, i32, i32, i32, i32, vp, vp]),
    "clsm_conv_pool_fwd": (i32, [vp, i32, i64, i32, i32, vp, i32, i32, i32, vp, i32, i32, i32, vp]),
    "clsm_conv_pool_fwd_tok": (i32, [vp, vp, i64, i32, i64, i32, i32, i32, vp, i32, i32, i32, vp, i32, i32, vp]),
    "clsm_conv_pool_fwd_d": (i32, [vp, i32, i64, i32, i32, i32, vp, i32, i32, i32, vp, i32, i32, i32, vp]),
    "clsm_pool_finalize": (i32, [vp, i64, i32, i32, vp, i32, vp, vp, i32, vp]),
    "clsm_dense_fwd": (i32, [vp, i64, i32, i32, vp, vp, i32, i32, vp, vp]),
    "clsm_dense_bwd_workspace_bytes": (sz, [i64, i32, i32]),
    "clsm_dense_bwd": (i32, [vp, vp, vp, i64, i32, i32, vp, i32, i32, vp, vp, vp, vp, sz, vp]),
    "clsm_sim_loss": (i32, [vp, vp, i64, i32, i32, f32, i64, vp, vp, vp, vp, vp, vp]),
    "clsm_conv_bwd_workspace_bytes": (sz, [i64, i32, i32, i32, i32]),
    "clsm_conv_pool_bwd": (i32, [vp, i32, i64, i32, i32, i32, vp, i32, i32, i32, vp, vp, i32, i32,
                                 vp, vp, vp, i32, vp, sz, vp]),
    "clsm_emb_scatter_workspace_bytes": (sz, [i64, i32]),
    "clsm_emb_scatter": (i32, [vp, i64, vp, i32, i32, f32, u64, i64, vp, vp, vp, vp, vp, sz, vp]),
    "clsm_adam_dense": (i32, [vp, vp, vp, vp, i64, i32, f32, f32, f32, f32, vp]),
    "clsm_adam_table": (i32, [vp, vp, vp, i64, i32, vp, vp, i32, i32, f32, f32, f32, f32, vp, i32, vp]),
    "clsm_adam_lr_table": (i32, [f32, f32, f32, i32, i32, vp]),
    "clsm_adam_defer_workspace_bytes": (C.c_size_t, [i64, i64]),
    "clsm_adam_defer_settle_tokens": (i32, [vp, vp, vp, i64, i32, vp, vp, i64, i32, vp, f32, f32, f32, vp, i32, vp, C.c_size_t, vp]),
    "clsm_adam_defer_rows": (i32, [vp, vp, vp, i32, vp, vp, vp, i64, vp, i32, i32, vp, f32, f32, f32, vp, i32, vp]),
    "clsm_adam_defer_settle_all": (i32, [vp, vp, vp, i64, i32, vp, i32, vp, f32, f32, f32, vp, i32, vp]),
    "clsm_adam_rows": (i32, [vp, vp, vp, i32, vp, vp, i64, vp, i32, i32, f32, f32, f32, f32, vp, i32, vp]),
    "clsm_table_to_bf16": (i32, [vp, i64, i32, vp, i32, vp]),
    "clsm_dp_push_rows": (i32, [vp, vp, i64, vp, i32, i32, vp, i64, i64, i32, vp, vp, i64, vp, vp]),
    "clsm_adam_table_dp": (i32, [vp, vp, vp, i64, i32, i32, vp, i64, vp, i64, i32, i32, f32, f32, f32, f32, vp, i32, i32, vp]),
    "clsm_normalize_rows": (i32, [vp, i64, i32, vp, vp]),
    "clsm_cos_topk_workspace_bytes": (sz, [i64, i64, i32]),
    "clsm_cos_topk": (i32, [vp, i64, vp, i64, i32, i32, i64, vp, vp, vp, sz, vp]),
    "clsm_normalize_rows_bf16": (i32, [vp, i64, i32, vp, vp]),
    "clsm_cos_topk_bf16_workspace_bytes": (sz, [i64, i64, i32]),
    "clsm_cos_topk_bf16": (i32, [vp, i64, vp, i64, i32, i32, i64, vp, vp, vp, sz, vp]),
    "clsm_topk_merge_workspace_bytes": (sz, [i32, i64, i32]),
    "clsm_topk_merge": (i32, [vp, vp, i32, i64, i32, vp, vp, vp, sz, vp]),
    "clsm_topk_merge_static": (i32, [vp, vp, i32, i64, i32, vp, vp, vp]),
}


class ClsmError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libclsm_b200.so not found at %s -- build it first (python -c 'import __graft_entry__ as g; "
            "g.build()' or make -C dnn_page_vectors_b200/csrc).  There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc, what=""):
    if rc != OK:
        msg = lib.clsm_last_error()
        raise ClsmError("%s failed (%d): %s" % (what or "libclsm_b200 call", rc,
                                               msg.decode("utf-8", "replace") if msg else ""))
    return rc


def call(name, *args):
    """Invoke an int-returning entry point and raise ClsmError on a non-zero code."""
    return check(getattr(lib, name)(*args), name)
