"""Profiling target: a few launches of ONE memory-bound kernel (argv[1]) at batch argv[2]."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import ops
from gabril_carla_b200.gaze import GazePreprocessor
from tests.synth import make_batch
which = sys.argv[1]
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
dev = "cuda:0"
base = make_batch(min(B, 64), seed=3)
frames, gaze, actions = [t.repeat((B // t.shape[0],) + (1,) * (t.dim() - 1)).to(dev) for t in base]
pre = GazePreprocessor(device=dev)
heat = torch.empty(B, 2, 180, 320, device=dev)
pre.splat_into(gaze, 2, 1, heat)
z = torch.randn(B, 20, 38, 64, device=dev).to(torch.bfloat16)
dz = torch.empty_like(z)
part = torch.zeros(B, device=dev)
stats = torch.empty(B, 4, device=dev)
for _ in range(4):
    if which == "reg":
        ops.reg_mse(z, heat[:, 1], heat.stride(0), 180, 320, 50.0, 1e-6, dz, part, stats=stats)
    elif which == "splat":
        pre.splat_into(gaze, 2, 1, heat)
    elif which == "obs":
        SH, SW = 47, 82
        if "x4" not in globals():
            x4 = torch.zeros(B * SH * SW + 1, ops.conv1_kp(2, True), device=dev, dtype=torch.bfloat16)
        ops.obs_s4d(frames, 2, 1, True, SH, SW, x4)
torch.cuda.synchronize()
print("ok")
