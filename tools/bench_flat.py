"""Flat-slab vs im2col timing of the stride-1 layers and of conv_2 in space-to-depth form.
Usage: python tools/bench_flat.py [B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import gabril_carla_b200  # noqa: F401
from gabril_carla_b200 import _lib, ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = "cuda:0"
_lib.init()
bf = torch.bfloat16


def timeit(fn, reps=7):
    fn(); fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


# name, H, W, Cin, Cout, R, pad, out_pad (border of the output tensor)
LAYERS = [("conv4", 22, 40, 128, 128, 3, 0, 1), ("res3x3", 20, 38, 128, 32, 3, 1, 1),
          ("conv5", 20, 38, 128, 128, 5, 2, 2), ("conv6", 20, 38, 128, 64, 5, 2, 0)]
for name, H, W, Cin, Cout, R, pad, opad in LAYERS:
    Hg, Wg = H + 2 * pad, W + 2 * pad
    OH, OW = Hg - R + 1, Wg - R + 1
    flop = 2.0 * B * OH * OW * Cout * R * R * Cin
    x = torch.zeros(B, Hg, Wg, Cin, device=dev, dtype=bf)
    x[:, pad:pad + H, pad:pad + W] = (torch.randn(B, H, W, Cin, device=dev) * 0.5).to(bf)
    w = (torch.randn(Cout, R * R, Cin, device=dev) * 0.05).to(bf)
    bias = torch.randn(Cout, device=dev)
    y = torch.zeros(B, OH + 2 * opad, OW + 2 * opad, Cout, device=dev, dtype=bf)
    t_f = timeit(lambda: ops.conv_flat(x, w, y, R, R, bias=bias, valid=(OH, OW),
                                       out_geom=(OH + 2 * opad, OW + 2 * opad, opad, opad), relu=True))
    line = f"{name:7s} flat fwd {t_f:7.3f} ms {flop / t_f / 1e9:7.1f} TF/s"
    if Cout % 64 == 0:
        dyp = torch.zeros(B, Hg + R - 1, Wg + R - 1, Cout, device=dev, dtype=bf)
        dyp[:, R - 1:R - 1 + OH, R - 1:R - 1 + OW] = (torch.randn(B, OH, OW, Cout, device=dev) * 0.1).to(bf)
        dx = torch.empty(B, Hg, Wg, Cin, device=dev, dtype=bf)
        t_d = timeit(lambda: ops.conv_flat(dyp, w, dx, R, R, gate=x, valid=(Hg, Wg), out_geom=(Hg, Wg, 0, 0),
                                           transpose_w=True))
        line += f" | dgrad {t_d:7.3f} ms {flop / t_d / 1e9:7.1f}"
        del dyp, dx
    dyg = torch.zeros(B, Hg, Wg, Cout, device=dev, dtype=bf)
    dyg[:, :OH, :OW] = (torch.randn(B, OH, OW, Cout, device=dev) * 0.1).to(bf)
    dw = torch.zeros(Cout, R * R, Cin, device=dev)
    t_w = timeit(lambda: ops.conv_wgrad_flat(dyg, x, dw, R, R, 0))
    line += f" | wgrad {t_w:7.3f} ms {flop / t_w / 1e9:7.1f}"
    print(line, flush=True)
    del x, y, dyg

# conv_2 (k4 s2 p1, 32 -> 64) as a 2x2 stride-1 conv over the space-to-depth conv_1 output
SH, SW = 47, 82
flop = 2.0 * B * 45 * 80 * 64 * 16 * 32
a1s = (torch.randn(B, SH, SW, 128, device=dev) * 0.5).to(bf)
w2 = (torch.randn(64, 4, 128, device=dev) * 0.05).to(bf)
b2 = torch.randn(64, device=dev)
a2p = torch.zeros(B, SH, SW, 64, device=dev, dtype=bf)
t_f = timeit(lambda: ops.conv_flat(a1s, w2, a2p, 2, 2, bias=b2, valid=(45, 80), out_geom=(SH, SW, 1, 1), relu=True))
da2p = torch.zeros(B, SH, SW, 64, device=dev, dtype=bf)
da2p[:, 1:46, 1:81] = (torch.randn(B, 45, 80, 64, device=dev) * 0.1).to(bf)
da1s = torch.empty(B, SH, SW, 128, device=dev, dtype=bf)
t_d = timeit(lambda: ops.conv_flat(da2p, w2, da1s, 2, 2, gate=a1s, transpose_w=True))
dw2 = torch.zeros(64, 4, 128, device=dev)
t_w = timeit(lambda: ops.conv_wgrad_flat(da2p, a1s, dw2, 2, 2, SW + 1, flop_pixels=B * 45 * 80))
print(f"conv2   s2d  fwd {t_f:7.3f} ms {flop / t_f / 1e9:7.1f} TF/s | dgrad {t_d:7.3f} ms {flop / t_d / 1e9:7.1f}"
      f" | wgrad {t_w:7.3f} ms {flop / t_w / 1e9:7.1f}", flush=True)
_lib.poll_error()
