"""Profiling target: conv_2 (space-to-depth, flat-slab) fwd / dgrad / wgrad, a few launches each."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import _lib, ops
which = sys.argv[1] if len(sys.argv) > 1 else "fwd"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
dev, bf = "cuda:0", torch.bfloat16
_lib.init()
SH, SW = 47, 82
a1s = (torch.randn(B, SH, SW, 128, device=dev) * 0.5).to(bf)
w2 = (torch.randn(64, 4, 128, device=dev) * 0.05).to(bf)
b2 = torch.randn(64, device=dev)
a2p = torch.zeros(B, SH, SW, 64, device=dev, dtype=bf)
da2p = torch.zeros(B, SH, SW, 64, device=dev, dtype=bf)
da2p[:, 1:46, 1:81] = (torch.randn(B, 45, 80, 64, device=dev) * 0.1).to(bf)
da1s = torch.empty(B, SH, SW, 128, device=dev, dtype=bf)
dw2 = torch.zeros(64, 4, 128, device=dev)
for _ in range(3):
    if which == "fwd":
        ops.conv_flat(a1s, w2, a2p, 2, 2, bias=b2, valid=(45, 80), out_geom=(SH, SW, 1, 1), relu=True)
    elif which == "dgrad":
        ops.conv_flat(da2p, w2, da1s, 2, 2, gate=a1s, transpose_w=True)
    else:
        ops.conv_wgrad_flat(da2p, a1s, dw2, 2, 2, SW + 1)
torch.cuda.synchronize()
_lib.poll_error()
print("ok")
