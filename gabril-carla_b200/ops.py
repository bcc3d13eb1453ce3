"""Thin typed wrappers over the dense-contraction entry points of libgabril_b200.so.

Layouts (see include/gabril_b200.h): activations NHWC bf16, weights [Cout][R*S][Cin] bf16,
weight gradients fp32 in the same order.  Every wrapper validates shapes/dtypes on the host and
raises on a non-zero status; none of them falls back to PyTorch math.
"""
import os

import torch

from . import _lib
from ._lib import c_int, ptr, stream

_NAIVE = os.environ.get("GABRIL_NAIVE_CONV", "0") == "1"


def _fn(name):
    lib = _lib.init()
    return getattr(lib, name + ("_naive" if _NAIVE else ""))


def _chk_bf16(t, name):
    if t is None:
        return
    if t.dtype != torch.bfloat16 or not t.is_cuda or not t.is_contiguous():
        raise ValueError(f"{name}: expected contiguous CUDA bf16 tensor, got {t.dtype} {t.device}")


def conv_out_hw(H, W, R, S, stride, pad):
    return (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1


def conv_fwd(x, w, bias, y, R, S, stride, pad, relu=False, resid=None, rowmask=None,
             out_f32_atomic=False, splits=1):
    """x [N,H,W,Cin] bf16, w [Cout,R*S,Cin] bf16, bias fp32[Cout] or None -> y [N,OH,OW,Cout]."""
    _chk_bf16(x, "x"); _chk_bf16(w, "w"); _chk_bf16(resid, "resid")
    N, H, W, Cin = x.shape
    Cout = w.shape[0]
    assert w.shape[1] == R * S and w.shape[2] == Cin, (w.shape, R, S, Cin)
    OH, OW = conv_out_hw(H, W, R, S, stride, pad)
    assert y.numel() == N * OH * OW * Cout, (y.shape, N, OH, OW, Cout)
    assert y.dtype == (torch.float32 if out_f32_atomic else torch.bfloat16)
    st = _fn("gabril_conv_fwd")(ptr(x), ptr(w), ptr(bias), ptr(resid), ptr(rowmask), ptr(y),
                                c_int(N), c_int(H), c_int(W), c_int(Cin), c_int(Cout), c_int(R),
                                c_int(S), c_int(stride), c_int(pad), c_int(int(relu)),
                                c_int(int(out_f32_atomic)), c_int(splits), stream())
    _lib.check(st, "gabril_conv_fwd")
    return y


def conv_dgrad(dy, w, dx, R, S, stride, pad, addend=None, gate=None, rowmask=None):
    """dy [N,OH,OW,Cout] bf16 -> dx [N,H,W,Cin] bf16 (shape taken from dx)."""
    _chk_bf16(dy, "dy"); _chk_bf16(w, "w"); _chk_bf16(dx, "dx")
    _chk_bf16(addend, "addend"); _chk_bf16(gate, "gate")
    N, H, W, Cin = dx.shape
    Cout = w.shape[0]
    OH, OW = conv_out_hw(H, W, R, S, stride, pad)
    assert dy.numel() == N * OH * OW * Cout, (dy.shape, N, OH, OW, Cout)
    st = _fn("gabril_conv_dgrad")(ptr(dy), ptr(w), ptr(addend), ptr(gate), ptr(rowmask), ptr(dx),
                                  c_int(N), c_int(H), c_int(W), c_int(Cin), c_int(Cout), c_int(R),
                                  c_int(S), c_int(stride), c_int(pad), stream())
    _lib.check(st, "gabril_conv_dgrad")
    return dx


def conv_wgrad(dy, x, dw, R, S, stride, pad, direct_store=False):
    """dw [Cout,R*S,Cin] fp32 += dy^T * im2col(x).  direct_store overwrites (no split)."""
    _chk_bf16(dy, "dy"); _chk_bf16(x, "x")
    N, H, W, Cin = x.shape
    Cout = dw.shape[0]
    assert dw.dtype == torch.float32 and dw.numel() == Cout * R * S * Cin
    OH, OW = conv_out_hw(H, W, R, S, stride, pad)
    assert dy.numel() == N * OH * OW * Cout, (dy.shape, N, OH, OW, Cout)
    st = _fn("gabril_conv_wgrad")(ptr(dy), ptr(x), ptr(dw), c_int(N), c_int(H), c_int(W),
                                  c_int(Cin), c_int(Cout), c_int(R), c_int(S), c_int(stride),
                                  c_int(pad), c_int(int(direct_store)), stream())
    _lib.check(st, "gabril_conv_wgrad")
    return dw


# ---- layout helpers (host-side plumbing, torch ops) ------------------------------------------
def w_oihw_to_internal(w_oihw: torch.Tensor) -> torch.Tensor:
    """[Cout,Cin,R,S] -> [Cout,R*S,Cin] (same dtype)."""
    Co, Ci, R, S = w_oihw.shape
    return w_oihw.permute(0, 2, 3, 1).reshape(Co, R * S, Ci).contiguous()


def w_internal_to_oihw(w: torch.Tensor, R: int, S: int) -> torch.Tensor:
    Co, T, Ci = w.shape
    return w.reshape(Co, R, S, Ci).permute(0, 3, 1, 2).contiguous()
