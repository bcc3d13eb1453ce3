"""Thin typed wrappers over the dense-contraction entry points of libgabril_b200.so.

Layouts (see include/gabril_b200.h): activations NHWC bf16, weights [Cout][R*S][Cin] bf16,
weight gradients fp32 in the same order.  Every wrapper validates shapes/dtypes on the host and
raises on a non-zero status; none of them falls back to PyTorch math.
"""
import os

import torch

from . import _lib
from ._lib import c_int, ptr, stream

def _fn(name):
    return getattr(_lib.init(), name)


# ---- optional per-launch tracing (bench.py's roofline leg): CUDA events around every call ----
_TRACE = None


class trace:
    """with ops.trace() as t: ...; t.rows() -> [(kernel, label, work, unit, ms)] after a sync."""

    def __init__(self):
        self.items = []

    def __enter__(self):
        global _TRACE
        _TRACE = self
        return self

    def __exit__(self, *exc):
        global _TRACE
        _TRACE = None

    def rows(self):
        torch.cuda.synchronize()
        return [(k, lab, work, unit, e0.elapsed_time(e1)) for k, lab, work, unit, e0, e1 in self.items]


class _span:
    def __init__(self, kernel, label, work, unit):
        self.args = (kernel, label, work, unit)

    def __enter__(self):
        if _TRACE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if _TRACE is not None:
            self.e1.record()
            _TRACE.items.append(self.args + (self.e0, self.e1))


def _chk_bf16(t, name):
    if t is None:
        return
    if t.dtype != torch.bfloat16 or not t.is_cuda or not t.is_contiguous():
        raise ValueError(f"{name}: expected contiguous CUDA bf16 tensor, got {t.dtype} {t.device}")


def _epilogue_bits(gate_bits, bits_out, out):
    """Attach packed ReLU sign words (int32 [out.numel() / 32]) to the next conv launch:
    gate_bits is read instead of a `gate` tensor, bits_out receives (stored value > 0)."""
    if gate_bits is None and bits_out is None:
        return
    for t, name in ((gate_bits, "gate_bits"), (bits_out, "bits_out")):
        if t is None:
            continue
        if t.dtype != torch.int32 or not t.is_cuda or not t.is_contiguous() or t.numel() * 32 != out.numel():
            raise ValueError(f"{name}: expected contiguous CUDA int32 tensor of {out.numel() // 32} words")
    _lib.check(_lib.init().gabril_conv_epilogue_bits(ptr(gate_bits), ptr(bits_out)), "gabril_conv_epilogue_bits")


def _wgrad_dbias(dbias, cout, mod=0):
    """Fuse the bias gradient into the next weight-gradient launch: dbias[c % mod] += column sums of dy."""
    if dbias is None:
        return
    _chk(dbias, torch.float32, "dbias")
    if dbias.numel() != (mod if mod else cout):
        raise ValueError(f"dbias: expected {mod if mod else cout} elements, got {dbias.numel()}")
    _lib.check(_lib.init().gabril_conv_wgrad_dbias_next(ptr(dbias), c_int(mod)), "gabril_conv_wgrad_dbias_next")


def sign_bits_like(t):
    """Zeroed sign-word buffer for a bf16 activation tensor (1 bit per element)."""
    assert t.numel() % 32 == 0
    return torch.zeros(t.numel() // 32, device=t.device, dtype=torch.int32)


def conv_out_hw(H, W, R, S, stride, pad):
    return (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1


def conv_fwd_planes(Cin, R, S, splits):
    """Split-K planes conv_fwd(out_f32_atomic=2) writes for a requested split count."""
    return int(_lib.load().gabril_conv_fwd_splits(c_int(Cin), c_int(R), c_int(S), c_int(splits)))


def conv_fwd(x, w, bias, y, R, S, stride, pad, relu=False, resid=None, rowmask=None,
             out_f32_atomic=False, splits=1, bits_out=None):
    """x [N,H,W,Cin] bf16, w [Cout,R*S,Cin] bf16, bias fp32[Cout] or None -> y [N,OH,OW,Cout]."""
    _chk_bf16(x, "x"); _chk_bf16(w, "w"); _chk_bf16(resid, "resid")
    N, H, W, Cin = x.shape
    Cout = w.shape[0]
    assert w.shape[1] == R * S and w.shape[2] == Cin, (w.shape, R, S, Cin)
    OH, OW = conv_out_hw(H, W, R, S, stride, pad)
    planes = conv_fwd_planes(Cin, R, S, splits) if int(out_f32_atomic) == 2 else 1
    assert y.numel() == planes * N * OH * OW * Cout, (y.shape, planes, N, OH, OW, Cout)
    assert y.dtype == (torch.float32 if out_f32_atomic else torch.bfloat16)
    with _span("tc_gemm_kernel", f"fwd {Cin}->{Cout} k{R}s{stride}", 2.0 * N * OH * OW * Cout * R * S * Cin, "flop"):
        _epilogue_bits(None, bits_out, y)
        st = _fn("gabril_conv_fwd")(ptr(x), ptr(w), ptr(bias), ptr(resid), ptr(rowmask), ptr(y),
                                    c_int(N), c_int(H), c_int(W), c_int(Cin), c_int(Cout), c_int(R),
                                    c_int(S), c_int(stride), c_int(pad), c_int(int(relu)),
                                    c_int(int(out_f32_atomic)), c_int(splits), stream())  # 0 / 1 atomic / 2 planes
    _lib.check(st, "gabril_conv_fwd")
    return y


def conv_pack_wpairs(w, w_pair, R, S):
    """w [Cout,R*S,Cin] bf16 -> w_pair [2*Cout, R*(S+1), Cin] bf16: the operand of conv_fwd_pairs."""
    _chk_bf16(w, "w"); _chk_bf16(w_pair, "w_pair")
    Cout, T, Cin = w.shape
    assert T == R * S and tuple(w_pair.shape) == (2 * Cout, R * (S + 1), Cin), (w.shape, w_pair.shape)
    with _span("gabril_conv_pack_wpairs", "", 2.0 * (w.numel() + w_pair.numel()), "byte"):
        st = _fn("gabril_conv_pack_wpairs")(ptr(w), ptr(w_pair), c_int(Cout), c_int(R), c_int(S), c_int(Cin), stream())
    _lib.check(st, "gabril_conv_pack_wpairs")
    return w_pair


def conv_fwd_pairs(x, w_pair, bias, y, R, S, pad, relu=False, rowmask=None):
    """Narrow stride-1 same convolution with two adjacent output pixels stacked along N (bit-identical to conv_fwd):
    x [N,H,W,Cin] bf16, w_pair from conv_pack_wpairs, bias fp32 [Cout] or None -> y [N,H,W,Cout] bf16."""
    _chk_bf16(x, "x"); _chk_bf16(w_pair, "w_pair"); _chk_bf16(y, "y")
    N, H, W, Cin = x.shape
    Cout = w_pair.shape[0] // 2
    assert w_pair.shape[1] == R * (S + 1) and w_pair.shape[2] == Cin and y.numel() == N * H * W * Cout
    with _span("tc_gemm_kernel", f"fwd {Cin}->{Cout} k{R}s1", 2.0 * N * H * W * Cout * R * S * Cin, "flop"):
        assert rowmask is None or (rowmask.dtype == torch.float32 and rowmask.numel() == N * H * W)
        st = _fn("gabril_conv_fwd_pairs_masked")(ptr(x), ptr(w_pair), ptr(bias), ptr(rowmask), ptr(y), c_int(N),
                                                 c_int(H), c_int(W), c_int(Cin), c_int(Cout), c_int(R), c_int(S),
                                                 c_int(pad), c_int(int(relu)), stream())
    _lib.check(st, "gabril_conv_fwd_pairs_masked")
    return y


def conv_dgrad(dy, w, dx, R, S, stride, pad, addend=None, gate=None, rowmask=None, gate_bits=None):
    """dy [N,OH,OW,Cout] bf16 -> dx [N,H,W,Cin] bf16 (shape taken from dx)."""
    _chk_bf16(dy, "dy"); _chk_bf16(w, "w"); _chk_bf16(dx, "dx")
    _chk_bf16(addend, "addend"); _chk_bf16(gate, "gate")
    N, H, W, Cin = dx.shape
    Cout = w.shape[0]
    OH, OW = conv_out_hw(H, W, R, S, stride, pad)
    assert dy.numel() == N * OH * OW * Cout, (dy.shape, N, OH, OW, Cout)
    with _span("tc_gemm_kernel", f"dgrad {Cin}<-{Cout} k{R}s{stride}", 2.0 * N * OH * OW * Cout * R * S * Cin, "flop"):
        _epilogue_bits(gate_bits, None, dx)
        st = _fn("gabril_conv_dgrad")(ptr(dy), ptr(w), ptr(addend), ptr(gate), ptr(rowmask), ptr(dx),
                                      c_int(N), c_int(H), c_int(W), c_int(Cin), c_int(Cout), c_int(R),
                                      c_int(S), c_int(stride), c_int(pad), stream())
    _lib.check(st, "gabril_conv_dgrad")
    return dx


def res_block_fwd(x, w3, w1, h, bits_h, out, bits_out):
    """One Residual block forward in one launch: h = relu(conv3x3(x, w3)), out = relu(x + conv1x1(h, w1)), with the
    sign words of both (x [N,H,W,128] bf16 rectified, w3 [32,9,128], w1 [128,1,32])."""
    for t, n in ((x, "x"), (w3, "w3"), (w1, "w1"), (h, "h"), (out, "out")):
        _chk_bf16(t, n)
    N, H, W, C = x.shape
    Ch = w3.shape[0]
    assert w3.shape == (Ch, 9, C) and w1.shape == (C, 1, Ch) and h.numel() == N * H * W * Ch and out.numel() == x.numel()
    assert bits_h.dtype == torch.int32 and bits_h.numel() * 32 == h.numel()
    assert bits_out.dtype == torch.int32 and bits_out.numel() * 32 == out.numel()
    with _span("res_block_fwd_kernel", f"res-block {C}->{Ch}->{C}", 2.0 * N * H * W * (9 * C * Ch + Ch * C), "flop"):
        st = _fn("gabril_res_block_fwd")(ptr(x), ptr(w3), ptr(w1), ptr(h), ptr(bits_h), ptr(out), ptr(bits_out),
                                          c_int(N), c_int(H), c_int(W), c_int(C), c_int(Ch), stream())
    _lib.check(st, "gabril_res_block_fwd")
    return out


def conv_wgrad(dy, x, dw, R, S, stride, pad, direct_store=False, dbias=None):
    """dw [Cout,R*S,Cin] fp32 += dy^T * im2col(x).  direct_store overwrites (no split).
    dbias fp32 [Cout] (optional) += column sums of dy, inside the same launch."""
    _chk_bf16(dy, "dy"); _chk_bf16(x, "x")
    N, H, W, Cin = x.shape
    Cout = dw.shape[0]
    assert dw.dtype == torch.float32 and dw.numel() == Cout * R * S * Cin
    OH, OW = conv_out_hw(H, W, R, S, stride, pad)
    assert dy.numel() == N * OH * OW * Cout, (dy.shape, N, OH, OW, Cout)
    with _span("tc_gemm_kernel", f"wgrad {Cin}x{Cout} k{R}s{stride}", 2.0 * N * OH * OW * Cout * R * S * Cin, "flop"):
        _wgrad_dbias(dbias, Cout)
        st = _fn("gabril_conv_wgrad")(ptr(dy), ptr(x), ptr(dw), c_int(N), c_int(H), c_int(W),
                                      c_int(Cin), c_int(Cout), c_int(R), c_int(S), c_int(stride),
                                      c_int(pad), c_int(int(direct_store)), stream())
    _lib.check(st, "gabril_conv_wgrad")
    return dw



# ---- flat-slab convolutions over zero-bordered tensors (csrc/tc_plan.cu) ----------------------
def conv_flat(x, w, y, R, S, *, bias=None, resid=None, gate=None, valid=None, out_geom=None, relu=False,
              transpose_w=False, flop_pixels=None, rowmask=None, gate_bits=None, bits_out=None):
    """x [N,Hg,Wg,Cx] bf16 zero-bordered; out(q) = sum_{r,s} x_flat[q + r*Wg + s] . W(r,s) for every
    flat pixel q.  valid=(vh, vw), out_geom=(o_rows, o_pitch, off_h, off_w): pixel (n,i,j) with
    i<vh, j<vw is stored at y row (n*o_rows + i+off_h)*o_pitch + j+off_w; valid=None stores every
    pixel at its own index.  transpose_w: w is [Cx,R*S,Cy] and taps are flipped (dgrad)."""
    _chk_bf16(x, "x"); _chk_bf16(w, "w"); _chk_bf16(y, "y"); _chk_bf16(resid, "resid"); _chk_bf16(gate, "gate")
    N, Hg, Wg, Cx = x.shape
    Cy = w.shape[2] if transpose_w else w.shape[0]
    assert w.shape[1] == R * S and (w.shape[0] if transpose_w else w.shape[2]) == Cx, (w.shape, R, S, Cx)
    vh, vw = valid if valid is not None else (0, 0)
    o_rows, o_pitch, off_h, off_w = out_geom if out_geom is not None else (Hg, Wg, 0, 0)
    if valid is None:
        assert y.numel() == N * Hg * Wg * Cy, (y.shape, N, Hg, Wg, Cy)
    else:
        assert y.numel() == N * o_rows * o_pitch * Cy, (y.shape, N, o_rows, o_pitch, Cy)
        assert vh + off_h <= o_rows and vw + off_w <= o_pitch
    with _span("tc_gemm_kernel", f"{'dgrad' if transpose_w else 'fwd'}-flat {Cx}->{Cy} k{R}x{S}",
               2.0 * (flop_pixels if flop_pixels is not None else N * (vh or Hg) * (vw or Wg)) * Cy * R * S * Cx,
               "flop"):
        _epilogue_bits(gate_bits, bits_out, y)
        st = _fn("gabril_conv_flat_masked")(ptr(x), ptr(w), ptr(bias), ptr(resid), ptr(gate), ptr(rowmask),
                                     ptr(y), c_int(N), c_int(Hg), c_int(Wg), c_int(Cx), c_int(Cy), c_int(R), c_int(S),
                                     c_int(vh), c_int(vw), c_int(o_rows), c_int(o_pitch), c_int(off_h),
                                     c_int(off_w), c_int(int(relu)), c_int(int(transpose_w)), stream())
    _lib.check(st, "gabril_conv_flat")
    return y


def conv_wgrad_flat(dy, x, dw, R, S, dy_row_off, flop_pixels=None, dbias=None):
    """dw [Cy,R*S,Cx] fp32 += sum_q dy_flat[q + dy_row_off]^T x_flat[q + r*Wg + s]; dy [N,Hg,Wg,Cy] and
    x [N,Hg,Wg,Cx] share one zero-bordered grid."""
    _chk_bf16(dy, "dy"); _chk_bf16(x, "x")
    N, Hg, Wg, Cx = x.shape
    Cy = dy.shape[-1]
    assert dy.numel() == N * Hg * Wg * Cy and dw.dtype == torch.float32 and dw.numel() == Cy * R * S * Cx
    px = flop_pixels if flop_pixels is not None else N * Hg * Wg
    with _span("tc_gemm_kernel", f"wgrad-flat {Cx}x{Cy} k{R}x{S}", 2.0 * px * Cy * R * S * Cx, "flop"):
        _wgrad_dbias(dbias, Cy)
        st = _fn("gabril_conv_wgrad_flat")(ptr(dy), ptr(x), ptr(dw), c_int(N), c_int(Hg), c_int(Wg), c_int(Cx),
                                           c_int(Cy), c_int(R), c_int(S), c_int(dy_row_off), stream())
    _lib.check(st, "gabril_conv_wgrad_flat")
    return dw


def conv1_s2d_fwd(patches, w, bias, y, N, SH, SW, img_h, img_w, relu=True, rowmask=None, bits_out=None):
    """patches bf16 [N*SH*SW*4, Kp] (space-to-depth row order), w [Cout,1,Kp] -> y [N,SH,SW,4*Cout];
    rows of padding pixels are stored as exact zeros."""
    _chk_bf16(patches, "patches"); _chk_bf16(w, "w"); _chk_bf16(y, "y")
    Kp = patches.shape[1]
    Cout = w.shape[0]
    assert patches.shape[0] == N * SH * SW * 4 and y.numel() == N * SH * SW * 4 * Cout
    with _span("tc_gemm_kernel", f"fwd {Kp}->{Cout} k1s1", 2.0 * N * img_h * img_w * Cout * Kp, "flop"):
        _epilogue_bits(None, bits_out, y)
        st = _fn("gabril_conv1_s2d_fwd_masked")(ptr(patches), ptr(w), ptr(bias), ptr(rowmask), ptr(y), c_int(N),
                                         c_int(SH), c_int(SW),
                                         c_int(img_h), c_int(img_w), c_int(Kp), c_int(Cout), c_int(int(relu)),
                                         stream())
    _lib.check(st, "gabril_conv1_s2d_fwd")
    return y


def obs_patches_s2d(frames, S, center, grayscale, SH, SW, out):
    """uint8 [B,L,H,W,3] -> bf16 patch rows in space-to-depth order (out must start zeroed: rows of
    padding pixels are never written)."""
    _chk(frames, torch.uint8, "frames")
    B, L, H, W, C = frames.shape
    Kp = conv1_kp(S, grayscale)
    assert out.shape == (B * SH * SW * 4, Kp) and out.dtype == torch.bfloat16
    _call("gabril_obs_patches_s2d", ptr(frames), ptr(out), c_int(B), c_int(L), c_int(S), c_int(center),
          c_int(H), c_int(W), c_int(int(grayscale)), c_int(Kp), c_int(SH), c_int(SW), stream(),
          work=float(B * S * H * W * 3 + B * (H // 2) * (W // 2) * Kp * 2))
    return out


# ---- conv_1 over the 4x4 space-to-depth frame image (csrc/conv1_s4d.cu) ------------------------
def obs_s4d(frames, S, center, grayscale, SH, SW, out):
    """uint8 [B,L,H,W,3] -> bf16 cells [B*SH*SW (+1), Cp] (gather + /255 + gray + space-to-depth)."""
    _chk(frames, torch.uint8, "frames")
    B, L, H, W, C = frames.shape
    Cp = out.shape[1]
    assert out.shape[0] >= B * SH * SW + 1 and out.dtype == torch.bfloat16 and C == 3
    _call("gabril_obs_s4d", ptr(frames), ptr(out), c_int(B), c_int(L), c_int(S), c_int(center), c_int(H), c_int(W),
          c_int(int(grayscale)), c_int(Cp), c_int(SH), c_int(SW), stream(),
          work=float(B * S * H * W * 3 + B * SH * SW * Cp * 2))
    return out


def obs_s4d_store(store_frames, widx, grayscale, SH, SW, out):
    """obs_s4d reading frame widx[b,s] of the uint8 store [N,H,W,3] (no batch tensor)."""
    _chk(store_frames, torch.uint8, "store_frames"); _chk(widx, torch.int64, "widx")
    N, H, W, C = store_frames.shape
    B, S = widx.shape
    Cp = out.shape[1]
    assert out.shape[0] >= B * SH * SW + 1 and out.dtype == torch.bfloat16
    _call("gabril_obs_s4d_store", ptr(store_frames), ptr(widx), ptr(out), c_int(B), c_int(S), c_int(H), c_int(W),
          c_int(int(grayscale)), c_int(Cp), c_int(SH), c_int(SW), stream(),
          work=float(B * S * H * W * 3 + B * SH * SW * Cp * 2))
    return out


def obs_s4d_gaze(frames, widx, heat, sel, mode, S, center, SH, SW, out):
    """obs_s4d with the gaze heat maps folded in (gray stacks): mode 1 = obs * (ivg*heat + 1 - ivg)
    (gaze.method Mask / AGIL's second encoder), mode 2 = channels (obs, ivg*heat) (ViSaRL).  frames: the uint8
    batch [B,L,H,W,3] (widx None) or the frame store [N,H,W,3] read through widx int64 [B,S]."""
    _chk(frames, torch.uint8, "frames"); _chk(heat, torch.float32, "heat"); _chk(sel, torch.float32, "sel")
    _chk(widx, torch.int64, "widx")
    if widx is None:
        B, L, H, W, _ = frames.shape
    else:
        (_, H, W, _), (B, L) = frames.shape, widx.shape
    assert heat.shape == (B, S, H, W) and out.shape[0] >= B * SH * SW + 1 and out.dtype == torch.bfloat16
    _call("gabril_obs_s4d_gaze", ptr(frames), ptr(widx), ptr(heat), ptr(sel), c_int(mode), ptr(out), c_int(B), c_int(L),
          c_int(S), c_int(center), c_int(H), c_int(W), c_int(1), c_int(out.shape[1]), c_int(SH), c_int(SW), stream(),
          work=float(B * S * H * W * 7 + B * SH * SW * out.shape[1] * 2))
    return out


def conv1_s4d_expand(w1, w4, C0):
    """fp32 master conv_1 weight [C1,1,K1p] -> bf16 [4*C1, 2, 2*Cp] weights of the 2x2 convolution."""
    _chk(w1, torch.float32, "w1"); _chk_bf16(w4, "w4")
    C1, K1p = w1.shape[0], w1.shape[-1]
    Cp = w4.shape[2] // 2
    assert w4.shape == (4 * C1, 2, 2 * Cp)
    _call("gabril_conv1_s4d_expand", ptr(w1), ptr(w4), c_int(C1), c_int(C0), c_int(K1p), c_int(Cp), stream())
    return w4


def conv1_s4d_fwd(x4, w4, bias, y, N, SH, SW, img_h, img_w, relu=True, rowmask=None, bits_out=None):
    """x4 bf16 cells, w4 [4*C1,2,2*Cp] -> y [N,SH,SW,4*C1]: conv_1 + ReLU in space-to-depth storage."""
    _chk_bf16(x4, "x4"); _chk_bf16(w4, "w4"); _chk_bf16(y, "y")
    Cp, C1 = x4.shape[1], w4.shape[0] // 4
    assert w4.shape[2] == 2 * Cp and y.numel() == N * SH * SW * 4 * C1 and x4.shape[0] >= N * SH * SW + 1
    with _span("tc_gemm_kernel", f"fwd-s4d {16 * Cp // 16}->{C1} k4s2", 2.0 * N * img_h * img_w * C1 * Cp, "flop"):
        _epilogue_bits(None, bits_out, y)
        st = _lib.init().gabril_conv1_s4d_fwd(ptr(x4), ptr(w4), ptr(bias), ptr(rowmask), ptr(y), c_int(N), c_int(SH),
                                              c_int(SW), c_int(Cp), c_int(C1), c_int(img_h), c_int(img_w),
                                              c_int(int(relu)), stream())
    _lib.check(st, "gabril_conv1_s4d_fwd")
    return y


def conv1_s4d_wgrad(dy, x4, dw4, N, SH, SW, img_h, img_w, dbias=None):
    """dw4 fp32 [4*C1,2,2*Cp] += dy [N,SH,SW,4*C1]^T . x4 windows; dbias fp32 [C1] (optional) += column sums of
    dy folded over its four quadrants."""
    _chk_bf16(dy, "dy"); _chk_bf16(x4, "x4"); _chk(dw4, torch.float32, "dw4")
    Cp, C1 = x4.shape[1], dw4.shape[0] // 4
    assert dw4.shape == (4 * C1, 2, 2 * Cp) and dy.numel() == N * SH * SW * 4 * C1
    with _span("tc_gemm_kernel", f"wgrad-s4d {Cp}x{C1} k4s2", 2.0 * N * img_h * img_w * C1 * Cp, "flop"):
        _wgrad_dbias(dbias, 4 * C1, C1)
        st = _lib.init().gabril_conv1_s4d_wgrad(ptr(dy), ptr(x4), ptr(dw4), c_int(N), c_int(SH), c_int(SW), c_int(Cp),
                                                c_int(C1), stream())
    _lib.check(st, "gabril_conv1_s4d_wgrad")
    return dw4


def conv1_s4d_fold(dw4, dw1, C0):
    """dw1 fp32 [C1,1,K1p] += the four quadrants of dw4."""
    _chk(dw4, torch.float32, "dw4"); _chk(dw1, torch.float32, "dw1")
    C1, K1p = dw1.shape[0], dw1.shape[-1]
    Cp = dw4.shape[2] // 2
    _call("gabril_conv1_s4d_fold", ptr(dw4), ptr(dw1), c_int(C1), c_int(C0), c_int(K1p), c_int(Cp), stream())
    return dw1


def s2d_weight_taps():
    """Permutation of the 16 taps of a k4/s2 filter into (r', s', qh, qw) order: the 2x2 stride-1
    filter over the space-to-depth input has tap (r', s') = [W(2r'+qh, 2s'+qw) for qh, qw]."""
    return [(2 * r + qh) * 4 + (2 * s + qw) for r in range(2) for s in range(2) for qh in range(2) for qw in range(2)]


# ---- nn.ConvTranspose2d on the convolution kernels + the gaze-predictor head --------------------
def convt_fwd(x, w, bias, y, R, S, stride, pad, relu=False):
    """x [N,OH,OW,Cin_t] bf16, w [Cin_t,R*S,Cout_t] bf16 -> y [N,H,W,Cout_t] (H, W taken from y)."""
    _chk_bf16(x, "x"); _chk_bf16(w, "w"); _chk_bf16(y, "y")
    N, H, W, Cout_t = y.shape
    Cin_t = w.shape[0]
    OH, OW = conv_out_hw(H, W, R, S, stride, pad)
    assert w.shape[1] == R * S and w.shape[2] == Cout_t and x.numel() == N * OH * OW * Cin_t, (x.shape, w.shape, y.shape)
    with _span("tc_gemm_kernel", f"convT fwd {Cin_t}->{Cout_t} k{R}s{stride}", 2.0 * N * OH * OW * Cin_t * R * S * Cout_t, "flop"):
        st = _fn("gabril_convt_fwd")(ptr(x), ptr(w), ptr(bias), ptr(y), c_int(N), c_int(H), c_int(W), c_int(Cout_t),
                                     c_int(Cin_t), c_int(R), c_int(S), c_int(stride), c_int(pad), c_int(int(relu)),
                                     stream())
    _lib.check(st, "gabril_convt_fwd")
    return y


def convt_dgrad(dy, w, dx, R, S, stride, pad, gate=None):
    """dy [N,H,W,Cout_t] -> dx [N,OH,OW,Cin_t] (gated by `gate`, same shape as dx)."""
    _chk_bf16(dy, "dy"); _chk_bf16(w, "w"); _chk_bf16(dx, "dx"); _chk_bf16(gate, "gate")
    N, H, W, Cout_t = dy.shape
    Cin_t = w.shape[0]
    OH, OW = conv_out_hw(H, W, R, S, stride, pad)
    assert w.shape[1] == R * S and w.shape[2] == Cout_t and dx.numel() == N * OH * OW * Cin_t
    with _span("tc_gemm_kernel", f"convT dgrad {Cin_t}<-{Cout_t} k{R}s{stride}", 2.0 * N * OH * OW * Cin_t * R * S * Cout_t, "flop"):
        st = _fn("gabril_convt_dgrad")(ptr(dy), ptr(w), ptr(gate), ptr(dx), c_int(N), c_int(H), c_int(W),
                                       c_int(Cout_t), c_int(Cin_t), c_int(R), c_int(S), c_int(stride), c_int(pad),
                                       stream())
    _lib.check(st, "gabril_convt_dgrad")
    return dx


def gp_head_fwd(x, wp, bias, target_base, target_bstride, dout, loss_part, dbias, gscale, out=None):
    """x bf16 [B,h,w,C]; wp bf16 [C,1,Kp]; target_base: fp32 tensor whose data_ptr is sample 0's plane."""
    _chk_bf16(x, "x"); _chk_bf16(wp, "wp"); _chk(dout, torch.float32, "dout")
    B, h, w, C = x.shape
    Kp = wp.shape[-1]
    _call("gabril_gp_head_fwd", ptr(x), ptr(wp), c_int(Kp), ptr(bias), ptr(target_base), c_ll(target_bstride),
          ptr(out), ptr(dout), ptr(loss_part), ptr(dbias), c_int(B), c_int(h), c_int(w), c_int(C), c_float(gscale),
          stream(), work=float(x.numel() * 2 + B * 4 * h * w * 12))


def gp_dout_patches(dout, patches):
    _chk(dout, torch.float32, "dout"); _chk_bf16(patches, "patches")
    B, H, W = dout.shape
    Kp = patches.shape[1]
    assert patches.shape[0] == B * (H // 2) * (W // 2)
    _call("gabril_gp_dout_patches", ptr(dout), ptr(patches), c_int(Kp), c_int(B), c_int(H // 2), c_int(W // 2),
          stream(), work=float(dout.numel() * 4 + patches.numel() * 2))

# ---- layout helpers (host-side plumbing, torch ops) ------------------------------------------
def w_oihw_to_internal(w_oihw: torch.Tensor) -> torch.Tensor:
    """[Cout,Cin,R,S] -> [Cout,R*S,Cin] (same dtype)."""
    Co, Ci, R, S = w_oihw.shape
    return w_oihw.permute(0, 2, 3, 1).reshape(Co, R * S, Ci).contiguous()


def w_internal_to_oihw(w: torch.Tensor, R: int, S: int) -> torch.Tensor:
    Co, T, Ci = w.shape
    return w.reshape(Co, R, S, Ci).permute(0, 3, 1, 2).contiguous()


# ---- memory-bound kernels (gaze_ops.cu / train_ops.cu) ---------------------------------------
from ._lib import c_float, c_ll  # noqa: E402


def _chk(t, dtype, name):
    if t is None:
        return
    if t.dtype != dtype or not t.is_cuda or not t.is_contiguous():
        raise ValueError(f"{name}: expected contiguous CUDA {dtype} tensor, got {t.dtype} on {t.device}"
                         f" contiguous={t.is_contiguous()}")


def _call(name, *args, work=0.0):
    lib = _lib.init()
    with _span(name, "", work, "byte"):
        st = getattr(lib, name)(*args)
    _lib.check(st, name)


def obs_format(frames, S, center, grayscale, out=None):
    """uint8 [B,L,H,W,3] -> fp32 [B,S*C',H,W] (data_loader_robomimic.py:144-157,179-202)."""
    _chk(frames, torch.uint8, "frames")
    B, L, H, W, C = frames.shape
    assert C == 3
    Cp = 1 if grayscale else 3
    if out is None:
        out = torch.empty(B, S * Cp, H, W, device=frames.device, dtype=torch.float32)
    _call("gabril_obs_format", ptr(frames), ptr(out), c_int(B), c_int(L), c_int(S), c_int(center),
          c_int(H), c_int(W), c_int(int(grayscale)), stream())
    return out


def conv1_kp(S, grayscale):
    k = 16 * S * (1 if grayscale else 3)
    # the patch GEMM's K is the N side of its wgrad: 32, 64 or a multiple of 128
    return 32 if k <= 32 else 64 if k <= 64 else (k + 127) // 128 * 128


def obs_patches(frames, S, center, grayscale, out=None):
    """uint8 [B,L,H,W,3] -> bf16 [B*OH*OW, Kp] conv_1 patch matrix (format + im2col fused)."""
    _chk(frames, torch.uint8, "frames")
    B, L, H, W, C = frames.shape
    OH, OW = conv_out_hw(H, W, 4, 4, 2, 1)
    Kp = conv1_kp(S, grayscale)
    if out is None:
        out = torch.empty(B * OH * OW, Kp, device=frames.device, dtype=torch.bfloat16)
    _call("gabril_obs_patches", ptr(frames), ptr(out), c_int(B), c_int(L), c_int(S), c_int(center),
          c_int(H), c_int(W), c_int(int(grayscale)), c_int(Kp), stream(),
          work=float(B * S * H * W * 3 + out.numel() * 2))
    return out


_splat_ws = {}


def gaze_splat(gaze, out, ktab, Lk_host, agg_coeff, step_coeff, normalize_base, aggregate, S, center,
               P, H, W, ws=None):
    """gaze fp32 [B,L,P*2] -> out fp32 [B,S,H,W]; ktab fp32 [S,stride] (device), Lk_host: ints."""
    import ctypes
    _chk(gaze, torch.float32, "gaze"); _chk(out, torch.float32, "out"); _chk(ktab, torch.float32, "ktab")
    B, L = gaze.shape[0], gaze.shape[1]
    Lk = (ctypes.c_int * S)(*[int(v) for v in Lk_host])
    agg = (ctypes.c_float * (S * S))(*[float(v) for v in agg_coeff]) if agg_coeff is not None else None
    stc = (ctypes.c_float * S)(*[float(v) for v in step_coeff]) if step_coeff is not None else None
    if ws is None:  # per-(device, B) extrema workspace, reused across calls (stream ordered)
        key = (gaze.device, B)
        ws = _splat_ws.get(key)
        if ws is None:
            ws = _splat_ws[key] = torch.empty(B, 32, device=gaze.device, dtype=torch.float32)
    _call("gabril_gaze_splat", ptr(gaze), ptr(out), ptr(ws), ptr(ktab), Lk, c_int(ktab.shape[1]), agg, stc,
          c_int(int(normalize_base)), c_int(int(aggregate)), c_int(B), c_int(L), c_int(S),
          c_int(center), c_int(P), c_int(H), c_int(W), stream(),
          work=float(B * S * P * 8 + out.numel() * 4))
    return out


MASK_LAYOUTS = {"plain": 0, "s2d": 1, "bordered": 2}


def rng_advance(rng_state):
    """Bump the step counter of the device-side Philox state (int64 [2] = seed, step)."""
    _chk(rng_state, torch.int64, "rng_state")
    _call("gabril_rng_advance", ptr(rng_state), stream())


def gmd_mask(g, A, h, w, drop_p=0.7, test_mode=False, rowmask=None, keep_out=None, ws=None, sel=None,
             layout="plain", grid=(0, 0), rng_state=None, draw_id=0):
    """g fp32 [B,S,H,W], A fp32 [B,h*w] -> rowmask fp32 [B,h*w] (gaze_utils.py:33-52); sel fp32 [B]
    (optional): samples with sel <= 0 see a zero heat map (train_bc.py:258).  layout / grid=(SH, SW):
    storage order of the mask (include/gabril_b200.h, gabril_gmd_mask_layout); rowmask must then be
    given, sized for that storage."""
    _chk(g, torch.float32, "g"); _chk(A, torch.float32, "A"); _chk(rng_state, torch.int64, "rng_state")
    B, S, H, W = g.shape
    if rowmask is None:
        rowmask = torch.empty(B, h * w, device=g.device, dtype=torch.float32)
    if ws is None:
        ws = (torch.empty(B, h * w, device=g.device, dtype=torch.float32),
              torch.empty(2, device=g.device, dtype=torch.int32))
    assert layout == "plain" or rowmask is not None
    _call("gabril_gmd_mask_rng", ptr(g), ptr(sel), ptr(A), ptr(rng_state), c_int(draw_id), ptr(ws[0]), ptr(ws[1]),
          ptr(rowmask), ptr(keep_out),
          c_int(B), c_int(S), c_int(H), c_int(W), c_int(h), c_int(w), c_float(drop_p),
          c_int(int(test_mode)), c_int(MASK_LAYOUTS[layout]), c_int(grid[0]), c_int(grid[1]), stream(), work=float(B * h * w * (16 * S * 4 + 16)))
    return rowmask


def rowmask_apply_nchw(z, rowmask, out=None):
    _chk(z, torch.float32, "z"); _chk(rowmask, torch.float32, "rowmask")
    B, C = z.shape[0], z.shape[1]
    n = z.numel() // (B * C)
    if out is None:
        out = torch.empty_like(z)
    _call("gabril_rowmask_apply_nchw", ptr(z), ptr(rowmask), ptr(out), c_int(B), c_int(C), c_int(n),
          stream())
    return out


def _z_layout(z):
    if z.dtype == torch.bfloat16:
        return 1
    if z.dtype == torch.float32:
        return 0
    raise ValueError("z must be bf16 NHWC [B,h,w,C] or fp32 NCHW [B,C,h,w]")


def _z_dims(z):
    if z.dtype == torch.bfloat16:
        B, h, w, C = z.shape
    else:
        B, C, h, w = z.shape
    return B, C, h, w


def gaze_mask_fwd(z, H, W, beta, out=None, stats=None):
    assert z.is_cuda and z.is_contiguous()
    B, C, h, w = _z_dims(z)
    if out is None:
        out = torch.empty(B, 1, H, W, device=z.device, dtype=torch.float32)
    _call("gabril_gaze_mask_fwd", ptr(z), c_int(_z_layout(z)), ptr(out), ptr(stats), c_int(B), c_int(C),
          c_int(h), c_int(w), c_int(H), c_int(W), c_float(beta), stream())
    return out


def gaze_mask_bwd(z, grad_mask, H, W, beta, dz=None):
    assert z.is_cuda and z.is_contiguous()
    _chk(grad_mask, torch.float32, "grad_mask")
    B, C, h, w = _z_dims(z)
    if dz is None:
        dz = torch.empty_like(z)
    _call("gabril_gaze_mask_bwd", ptr(z), c_int(_z_layout(z)), ptr(grad_mask), ptr(dz), c_int(B),
          c_int(C), c_int(h), c_int(w), c_int(H), c_int(W), c_float(beta), stream())
    return dz


def reg_mse(z, g1_base, g1_bstride, H, W, beta, grad_scale, dz, loss_part, sel=None, stats=None):
    """Fused Reg-MSE forward+backward; g1_base: fp32 tensor whose data_ptr is sample 0's plane."""
    assert z.is_cuda and z.is_contiguous()
    B, C, h, w = _z_dims(z)
    _call("gabril_reg_mse", ptr(z), c_int(_z_layout(z)), ptr(g1_base), c_ll(g1_bstride), ptr(sel),
          ptr(dz), ptr(loss_part), ptr(stats), c_int(B), c_int(C), c_int(h), c_int(w), c_int(H),
          c_int(W), c_float(beta), c_float(grad_scale), stream(),
          work=float(B * (2 * z.element_size() * C * h * w + 4 * H * W)))


DIST_KINDS = {"MSE": 0, "KL": 1, "TV": 2, "JS": 3}


def reg_dist(z, g1_base, g1_bstride, H, W, beta, grad_scale, dz, loss_part, kind="MSE", sel=None,
             scale_dev=None, stats=None):
    """Fused regulariser forward+backward for gaze.prob_dist_type in {MSE, KL, TV, JS}
    (train_bc.py:152-169); scale_dev: optional device scalar multiplied into grad_scale."""
    assert z.is_cuda and z.is_contiguous()
    if kind not in DIST_KINDS:
        raise ValueError(f"Invalid prob_dist_type: {kind}")
    B, C, h, w = _z_dims(z)
    _call("gabril_reg_dist", ptr(z), c_int(_z_layout(z)), ptr(g1_base), c_ll(g1_bstride), ptr(sel),
          ptr(scale_dev), c_int(DIST_KINDS[kind]), ptr(dz), ptr(loss_part), ptr(stats), c_int(B),
          c_int(C), c_int(h), c_int(w), c_int(H), c_int(W), c_float(beta), c_float(grad_scale), stream(),
          work=float(B * (2 * z.element_size() * C * h * w + 4 * H * W)))


def ivg_select(frames, S, center, grayscale, ratio, sel=None, sel_stats=None):
    """gaze.ratio < 1 sample selection (train_bc.py:230-240) from the uint8 frames [B,L,H,W,3].
    Returns (sel fp32 [B], sel_stats fp32 [2] = (n_selected, B / n_selected))."""
    _chk(frames, torch.uint8, "frames")
    B, L, H, W, _ = frames.shape
    if sel is None:
        sel = torch.empty(B, device=frames.device, dtype=torch.float32)
    if sel_stats is None:
        sel_stats = torch.empty(2, device=frames.device, dtype=torch.float32)
    _call("gabril_ivg_select", ptr(frames), ptr(sel), ptr(sel_stats), c_int(B), c_int(L), c_int(S),
          c_int(center), c_int(H), c_int(W), c_int(int(grayscale)), c_float(ratio), stream(),
          work=float(B * S * H * W * 3))
    return sel, sel_stats


# ---- device-resident demonstration store ----------------------------------------
def store_windows(step_idx, demo_first, S, out=None):
    """[B] sampled step indices -> int64 [B,S] frame indices (robomimic frame_stack=S, seq_length=1:
    steps idx-S+1..idx, clamped to the first step of the demo; dataset.py:609-628)."""
    _chk(step_idx, torch.int64, "step_idx"); _chk(demo_first, torch.int64, "demo_first")
    B = step_idx.numel()
    if out is None:
        out = torch.empty(B, S, device=step_idx.device, dtype=torch.int64)
    _call("gabril_store_windows", ptr(step_idx), ptr(demo_first), ptr(out), c_int(B), c_int(S),
          c_ll(demo_first.numel()), stream(), work=float(B * S * 24))
    return out


def store_gather_rows(src, widx, out=None):
    """src fp32 [N,R], widx int64 [B,S] -> fp32 [B,S,R]."""
    _chk(src, torch.float32, "src"); _chk(widx, torch.int64, "widx")
    R = src.shape[1]
    if out is None:
        out = torch.empty(*widx.shape, R, device=src.device, dtype=torch.float32)
    _call("gabril_store_gather_rows", ptr(src), ptr(widx), ptr(out), c_ll(widx.numel()), c_int(R),
          stream(), work=float(widx.numel() * (8 + 8 * R)))
    return out


def obs_patches_s2d_store(store_frames, widx, grayscale, SH, SW, out):
    """obs_patches_s2d reading frame widx[b,s] of the uint8 store [N,H,W,3] (no batch tensor)."""
    _chk(store_frames, torch.uint8, "store_frames"); _chk(widx, torch.int64, "widx")
    N, H, W, C = store_frames.shape
    B, S = widx.shape
    Kp = conv1_kp(S, grayscale)
    assert out.shape == (B * SH * SW * 4, Kp) and out.dtype == torch.bfloat16
    _call("gabril_obs_patches_s2d_store", ptr(store_frames), ptr(widx), ptr(out), c_int(B), c_int(S),
          c_int(H), c_int(W), c_int(int(grayscale)), c_int(Kp), c_int(SH), c_int(SW), stream(),
          work=float(B * S * H * W * 3 + B * (H // 2) * (W // 2) * Kp * 2))
    return out


def ivg_select_store(store_frames, widx, grayscale, ratio, sel=None, sel_stats=None):
    _chk(store_frames, torch.uint8, "store_frames"); _chk(widx, torch.int64, "widx")
    N, H, W, _ = store_frames.shape
    B, S = widx.shape
    if sel is None:
        sel = torch.empty(B, device=widx.device, dtype=torch.float32)
    if sel_stats is None:
        sel_stats = torch.empty(2, device=widx.device, dtype=torch.float32)
    _call("gabril_ivg_select_store", ptr(store_frames), ptr(widx), ptr(sel), ptr(sel_stats), c_int(B),
          c_int(S), c_int(H), c_int(W), c_int(int(grayscale)), c_float(ratio), stream(),
          work=float(B * S * H * W * 3))
    return sel, sel_stats


def colsum_bf16(dy, db):
    _chk(dy, torch.bfloat16, "dy"); _chk(db, torch.float32, "db")
    C = db.numel()
    M = dy.numel() // C
    _call("gabril_colsum_bf16", ptr(dy), ptr(db), c_ll(M), c_int(C), stream(), work=float(M * C * 2))


def bias_act_cast(x, bias, y_bf16, relu=False, y_f32=None):
    """y = act(sum_planes x + bias) as bf16; x fp32 [M, C] or split-K planes [planes, M, C] (summed in order)."""
    _chk(x, torch.float32, "x"); _chk(y_bf16, torch.bfloat16, "y")
    planes = x.shape[0] if x.dim() == 3 else 1
    M, C = x.shape[-2], x.shape[-1]
    _call("gabril_bias_act_cast_planes", ptr(x), c_int(planes), ptr(bias), ptr(y_bf16), ptr(y_f32), c_ll(M), c_int(C),
          c_int(int(relu)), stream())


def mse_head(raw, bias, actions_view, A, part, gscale, dlogits=None, dbias=None, logits_out=None):
    """raw fp32 [B, ld]; actions_view: fp32 view whose row b starts at data_ptr + b*stride(0)."""
    B, ld = raw.shape
    ld_d = dlogits.shape[1] if dlogits is not None else 0
    _call("gabril_mse_head", ptr(raw), c_int(ld), ptr(bias), ptr(actions_view),
          c_ll(actions_view.stride(0)), c_int(A), ptr(logits_out), ptr(dlogits), c_int(ld_d),
          ptr(dbias), ptr(part), c_int(B), c_float(gscale), stream())


def gril_head(raw, bias, target_view, P2, sel, dpred, dbias, reg_part, gscale):
    """GRIL coordinate loss (train_bc.py:182-192); target_view: fp32 view whose row b starts at data_ptr + b*stride(0)."""
    B, ld = raw.shape
    _call("gabril_gril_head", ptr(raw), c_int(ld), ptr(bias), ptr(target_view), c_ll(target_view.stride(0)), ptr(sel),
          c_int(P2), ptr(dpred), c_int(dpred.shape[1]), ptr(dbias), ptr(reg_part), c_int(B), c_float(gscale), stream())


def agil_combine(z, z_agil, sel, rowmask=None):
    """z = where(sel > 0, 0.5 (z + z_agil), z) [* rowmask]; bf16 [B,h,w,C] in place (train_bc.py:268-272)."""
    _chk_bf16(z, "z"); _chk_bf16(z_agil, "z_agil")
    B, h, w, C = z.shape
    _call("gabril_agil_combine", ptr(z), ptr(z_agil), ptr(sel), ptr(rowmask), c_int(B), c_int(h * w), c_int(C), stream(),
          work=float(z.numel() * 6))


def agil_split(dz, dz_agil, sel):
    _chk_bf16(dz, "dz"); _chk_bf16(dz_agil, "dz_agil")
    B, h, w, C = dz.shape
    _call("gabril_agil_split", ptr(dz), ptr(dz_agil), ptr(sel), c_int(B), c_int(h * w), c_int(C), stream(),
          work=float(dz.numel() * 6))


def finalize_losses(actor_part, reg_part, B, actor_scale, reg_scale, lam, losses, reg_scale_dev=None):
    _call("gabril_finalize_losses_sel", ptr(actor_part), ptr(reg_part), c_int(B), c_float(actor_scale),
          c_float(reg_scale), ptr(reg_scale_dev), c_float(lam), ptr(losses), stream())


def adam(p, g, m, v, shadow, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0,
         decoupled=False, grad_scale=1.0):
    n = p.numel()
    _call("gabril_adam", ptr(p), ptr(g), ptr(m), ptr(v), ptr(shadow), c_ll(n), c_float(lr),
          c_float(beta1), c_float(beta2), c_float(eps), c_float(weight_decay), c_int(int(decoupled)),
          c_int(step), c_float(grad_scale), stream(), work=float(n * (28 + (2 if shadow is not None else 0))))


def cast_f32_bf16(x, y):
    _call("gabril_cast_f32_bf16", ptr(x), ptr(y), c_ll(x.numel()), stream())


def rowscale_bf16(t, scale):
    _chk(t, torch.bfloat16, "t"); _chk(scale, torch.float32, "scale")
    rows = scale.numel()
    C = t.numel() // rows
    _call("gabril_rowscale_bf16", ptr(t), ptr(scale), c_ll(rows), c_int(C), stream(),
          work=float(rows * (4 * C + 4)))
