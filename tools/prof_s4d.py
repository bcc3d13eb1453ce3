"""conv_1 over the 4x4 space-to-depth image, alone (for ncu / timing): python tools/prof_s4d.py [B] [reps] [what]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import _lib, ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
what = sys.argv[3] if len(sys.argv) > 3 else "fwd"
_lib.init()
dev, bf = "cuda:0", torch.bfloat16
SH, SW, Cp, C1, C0 = 47, 82, 32, 32, 2
x4 = (torch.rand(B * SH * SW + 1, Cp, device=dev)).to(bf)
w1 = torch.randn(C1, 1, Cp, device=dev) * 0.1
w4 = torch.empty(4 * C1, 2, 2 * Cp, device=dev, dtype=bf)
ops.conv1_s4d_expand(w1, w4, C0)
bias = torch.randn(C1, device=dev) * 0.1
y = torch.empty(B, SH, SW, 4 * C1, device=dev, dtype=bf)
sb = torch.zeros(y.numel() // 32, device=dev, dtype=torch.int32)
dy = (torch.randn(B, SH, SW, 4 * C1, device=dev) * 0.1).to(bf)
dw4 = torch.zeros(4 * C1, 2, 2 * Cp, device=dev)
a2p = torch.zeros(B, SH, SW, 64, device=dev, dtype=bf)
w2 = (torch.randn(64, 4, 128, device=dev) * 0.05).to(bf)
b2 = torch.randn(64, device=dev)
da1s = torch.empty(B, SH, SW, 128, device=dev, dtype=bf)
sb2 = torch.zeros(a2p.numel() // 32, device=dev, dtype=torch.int32)


def run():
    if what == "fwd":
        ops.conv1_s4d_fwd(x4, w4, bias, y, B, SH, SW, 90, 160, relu=True, bits_out=sb)
    elif what == "wgrad":
        ops.conv1_s4d_wgrad(dy, x4, dw4, B, SH, SW, 90, 160)
    elif what == "conv2fwd":
        ops.conv_flat(y, w2, a2p, 2, 2, bias=b2, valid=(45, 80), out_geom=(SH, SW, 1, 1), relu=True, bits_out=sb2)
    elif what == "conv2dgrad":
        ops.conv_flat(a2p, w2, da1s, 2, 2, gate_bits=sb, transpose_w=True)


run(); run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    run()
e1.record(); torch.cuda.synchronize()
print(f"{what} B={B}: {e0.elapsed_time(e1) / reps:.3f} ms", "err", _lib.load().gabril_poll_error())
import ctypes
buf = (ctypes.c_ulonglong * 16)()
_lib.load().gabril_tc_prof(buf, 1)
m = max(1, buf[12])
if buf[12]:
    print(f"epilogue warp 4 per tile: pre {buf[8]/m:.0f} wait_tfull {buf[9]/m:.0f} tmem_ld {buf[10]/m:.0f} math {buf[13]/m:.0f} stores {buf[14]/m:.0f} total {buf[11]/m:.0f} clk; tiles {buf[12]}")
