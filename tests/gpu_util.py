"""Helpers for the -m gpu tests: raw-pointer calls into the C ABI on torch-owned memory."""
import ctypes as C

import numpy as np
import torch

from dnn_page_vectors_b200 import _lib
from dnn_page_vectors_b200._lib import call, lib  # noqa: F401

DEV = "cuda:0"


def P(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def ST():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def dev(a, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(DEV)


def bf16_round(a):
    """float array -> the float32 value of its bf16 rounding (round-to-nearest-even)."""
    return torch.as_tensor(np.asarray(a, dtype=np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


def rup(x, m):
    return (x + m - 1) // m * m


def decode_keys(keys):
    """(S, P) int64 tensor of pool keys -> (max fp32, argmax int)"""
    k = keys.cpu().numpy().view(np.uint64)
    hi = (k >> np.uint64(32)).astype(np.uint32)
    lo = (k & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    neg = (hi & np.uint32(0x80000000)) == 0
    u = np.where(neg, ~hi, hi & np.uint32(0x7FFFFFFF)).astype(np.uint32)
    return u.view(np.float32), (np.uint32(0xFFFFFFFF) - lo).astype(np.int64)


def pack_w(W, Dpad, Fpad, bf16):
    """(k, D, F) numpy -> device Wt (Fpad, k*Dpad) through clsm_pack_conv_w"""
    k, D, F = W.shape
    w = dev(W, torch.float32)
    wt = torch.zeros((Fpad, k * Dpad), dtype=torch.bfloat16 if bf16 else torch.float32, device=DEV)
    call("clsm_pack_conv_w", P(w), k, D, F, Dpad, Fpad, _lib.BF16 if bf16 else _lib.F32, P(wt), ST())
    return wt


def make_X(Xnp, Dpad, bf16):
    """(S, L, D) numpy -> device (S*L+8, Dpad) zero padded"""
    S, L, D = Xnp.shape
    X = torch.zeros((S * L + 8, Dpad), dtype=torch.bfloat16 if bf16 else torch.float32, device=DEV)
    X[:S * L, :D] = dev(Xnp.reshape(S * L, D)).to(X.dtype)
    return X
