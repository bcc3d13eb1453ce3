"""Fused Residual block forward alone (timing / ncu): python tools/prof_res.py [B] [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import _lib, ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
_lib.init()
dev, bf = "cuda:0", torch.bfloat16
x = torch.relu(torch.randn(B, 20, 38, 128, device=dev) * 0.5).to(bf)
w3 = (torch.randn(32, 9, 128, device=dev) * 0.03).to(bf)
w1 = (torch.randn(128, 1, 32, device=dev) * 0.1).to(bf)
h = torch.empty(B, 20, 38, 32, device=dev, dtype=bf)
o = torch.empty(B, 20, 38, 128, device=dev, dtype=bf)
sbh, sbo = ops.sign_bits_like(h), ops.sign_bits_like(o)
for _ in range(3):
    ops.res_block_fwd(x, w3, w1, h, sbh, o, sbo)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    ops.res_block_fwd(x, w3, w1, h, sbh, o, sbo)
e1.record(); torch.cuda.synchronize()
print(f"res_block_fwd B={B} debug={os.environ.get('GABRIL_RB_DEBUG', '0')}: {e0.elapsed_time(e1) / reps:.4f} ms", "err", _lib.load().gabril_poll_error())
