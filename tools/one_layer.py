"""Run one conv layer's fwd/dgrad/wgrad once (for compute-sanitizer / ncu): python tools/one_layer.py conv2 wgrad 8"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import _lib, ops
L = {"conv1": (1, 1, 32, 32, 1, 1, 0), "conv2": (90, 160, 32, 64, 4, 2, 1), "conv3": (45, 80, 64, 128, 4, 2, 1),
     "conv4": (22, 40, 128, 128, 3, 1, 0), "res3x3": (20, 38, 128, 32, 3, 1, 1), "res1x1": (20, 38, 32, 128, 1, 1, 0),
     "conv5": (20, 38, 128, 128, 5, 1, 2), "conv6": (20, 38, 128, 64, 5, 1, 2)}
name, what, N = sys.argv[1], sys.argv[2], int(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 1
H, W, Cin, Cout, R, stride, pad = L[name]
if name == "conv1":
    N = N * 14400
_lib.init()
dev, bf = "cuda:0", torch.bfloat16
OH, OW = ops.conv_out_hw(H, W, R, R, stride, pad)
x = (torch.randn(N, H, W, Cin, device=dev) * 0.5).to(bf)
w = (torch.randn(Cout, R * R, Cin, device=dev) * 0.05).to(bf)
y = torch.empty(N, OH, OW, Cout, device=dev, dtype=bf)
dy = (torch.randn(N, OH, OW, Cout, device=dev) * 0.1).to(bf)
dx = torch.empty_like(x)
dw = torch.zeros(Cout, R * R, Cin, device=dev)
bias = torch.randn(Cout, device=dev)
# the engine's data-gradient launches: ReLU gate as packed sign words, skip gradient added in the Residual 3x3
gbits = torch.randint(-2 ** 31, 2 ** 31 - 1, (x.numel() // 32,), device=dev, dtype=torch.int32)
addend = (torch.randn_like(x) if name == "res3x3" else None)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    if what == "fwd":
        ops.conv_fwd(x, w, bias, y, R, R, stride, pad, relu=True)
    elif what == "dgrad":
        ops.conv_dgrad(dy, w, dx, R, R, stride, pad, gate_bits=gbits, addend=addend)
    else:
        ops.conv_wgrad(dy, x, dw, R, R, stride, pad)
e1.record()
torch.cuda.synchronize()
print("ok", name, what, N, f"{e0.elapsed_time(e1) / reps:.3f} ms/launch (first launch included)", "TC_DEBUG", os.environ.get("GABRIL_TC_DEBUG"),
      "err flag", _lib.load().gabril_poll_error())
import ctypes
buf = (ctypes.c_ulonglong * 16)()
if hasattr(_lib.load(), "gabril_tc_prof"):
    _lib.load().gabril_tc_prof(buf, 1)
    n = max(1, buf[4])
    m = max(1, buf[12])
    print(f"epilogue warp 1 per tile: pre {buf[8]/m:.0f} wait_tfull {buf[9]/m:.0f} tmem_ld {buf[10]/m:.0f} math {buf[13]/m:.0f} stores {buf[14]/m:.0f} total {buf[11]/m:.0f} clk; tiles {buf[12]}")
    print(f"MMA thread per k-block: wait_full {buf[0]/n:.0f} issue {buf[1]/n:.0f} commit {buf[2]/n:.0f} clk; per-tile pre {buf[3]/max(1,n)*1.0:.1f}/kb; kblocks {buf[4]}")
