"""Isolated CUDA-event timings of the memory-bound kernels at a given batch (default 1024)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import ops
from gabril_carla_b200.gaze import GazePreprocessor
from tests.synth import make_batch

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = "cuda:0"
base = make_batch(min(B, 64), seed=3)
frames, gaze, actions = [t.repeat((B // t.shape[0],) + (1,) * (t.dim() - 1)).to(dev) for t in base]
pre = GazePreprocessor(device=dev)
heat = torch.empty(B, 2, 180, 320, device=dev)
z = torch.randn(B, 20, 38, 64, device=dev).to(torch.bfloat16)
dz = torch.empty_like(z)
part = torch.zeros(B, device=dev)
stats = torch.empty(B, 4, device=dev)
patches = torch.empty(B * 90 * 160, 32, device=dev, dtype=torch.bfloat16)
A = torch.rand(B, 760, device=dev)
big = torch.randn(B * 14400, 32, device=dev).to(torch.bfloat16)
db = torch.zeros(32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(name, fn, nbytes, reps=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    print(f"{name:24s} {ms:8.4f} ms  {nbytes / ms / 1e6:8.1f} GB/s", flush=True)


timeit("gaze_splat", lambda: pre.splat_into(gaze, 2, 1, heat), B * (80 + 460800))
timeit("reg_mse", lambda: ops.reg_mse(z, heat[:, 1], heat.stride(0), 180, 320, 50.0, 1e-6, dz, part, stats=stats), B * 424960)
x4 = torch.zeros(B * 47 * 82 + 1, 32, device=dev, dtype=torch.bfloat16)
timeit("obs_s4d", lambda: ops.obs_s4d(frames, 2, 1, True, 47, 82, x4), B * (345600 + 47 * 82 * 64))
timeit("gmd_mask", lambda: ops.gmd_mask(heat, A, 20, 38), B * (460800 + 6080))
timeit("colsum da1", lambda: ops.colsum_bf16(big, db), big.numel() * 2)
