"""Per-layer micro-benchmark of the tcgen05 implicit-GEMM kernel: every conv / linear shape of the
encoder at batch B, fwd / dgrad / wgrad, CUDA-event timed (median of reps, L2 flushed by the
working set itself at B=1024).  Usage: python tools/bench_layers.py [B] [filter-substring]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import gabril_carla_b200  # noqa: F401
from gabril_carla_b200 import _lib, ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
flt = sys.argv[2] if len(sys.argv) > 2 else ""
dev = "cuda:0"
_lib.init()
bf = torch.bfloat16
# name, H, W, Cin, Cout, R, stride, pad
LAYERS = [("conv1", 14400 // 1, 1, 32, 32, 1, 1, 0),   # patch GEMM: rows = 90*160 per sample
          ("conv2", 90, 160, 32, 64, 4, 2, 1), ("conv3", 45, 80, 64, 128, 4, 2, 1),
          ("conv4", 22, 40, 128, 128, 3, 1, 0), ("res3x3", 20, 38, 128, 32, 3, 1, 1),
          ("res1x1", 20, 38, 32, 128, 1, 1, 0), ("conv5", 20, 38, 128, 128, 5, 1, 2),
          ("conv6", 20, 38, 128, 64, 5, 1, 2)]


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


print(f"B={B} occ={os.environ.get('GABRIL_TC_OCC', 'default')}")
tot = 0.0
for name, H, W, Cin, Cout, R, stride, pad in LAYERS:
    if flt and flt not in name:
        continue
    N = B
    if name == "conv1":
        N, H, W = B * 14400, 1, 1
    OH, OW = ops.conv_out_hw(H, W, R, R, stride, pad)
    x = (torch.randn(N, H, W, Cin, device=dev) * 0.5).to(bf)
    w = (torch.randn(Cout, R * R, Cin, device=dev) * 0.05).to(bf)
    bias = torch.randn(Cout, device=dev)
    y = torch.empty(N, OH, OW, Cout, device=dev, dtype=bf)
    dy = (torch.randn(N, OH, OW, Cout, device=dev) * 0.1).to(bf)
    dx = torch.empty_like(x)
    dw = torch.zeros(Cout, R * R, Cin, device=dev)
    flop = 2.0 * N * OH * OW * Cout * R * R * Cin
    gbits = torch.randint(-2 ** 31, 2 ** 31 - 1, (x.numel() // 32,), device=dev, dtype=torch.int32)
    addend = torch.randn_like(x) if name == "res3x3" else None
    t_f = timeit(lambda: ops.conv_fwd(x, w, bias, y, R, R, stride, pad, relu=True))
    line = f"{name:7s} fwd {t_f:7.3f} ms {flop / t_f / 1e9:7.1f} TF/s"
    if name != "conv1":
        t_d = timeit(lambda: ops.conv_dgrad(dy, w, dx, R, R, stride, pad, gate_bits=gbits, addend=addend))
        line += f" | dgrad {t_d:7.3f} ms {flop / t_d / 1e9:7.1f}"
    else:
        t_d = 0.0
    t_w = timeit(lambda: ops.conv_wgrad(dy, x, dw, R, R, stride, pad))
    line += f" | wgrad {t_w:7.3f} ms {flop / t_w / 1e9:7.1f}"
    tot += t_f + t_d + t_w
    print(line, flush=True)
    del x, y, dy, dx
print(f"sum {tot:.3f} ms")
_lib.poll_error()
