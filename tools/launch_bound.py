"""Is the step launch-bound at small per-GPU batches?  For each batch size: host time to ENQUEUE one
step (perf_counter, queue drained before, no sync inside) vs device time per step (CUDA events over
20 back-to-back steps).  python tools/launch_bound.py [B ...]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gabril_carla_b200  # noqa: E402,F401
from gabril_carla_b200.engine import StepEngine  # noqa: E402
from gabril_carla_b200.gaze import GazePreprocessor  # noqa: E402
from gabril_carla_b200.model import ParamArena, PolicySpec  # noqa: E402
from tests.synth import make_batch  # noqa: E402

dev = "cuda:0"
spec = PolicySpec(2)
arena = ParamArena(spec, dev)
arena.init_reference_style(seed=42)
for B in [int(a) for a in sys.argv[1:]] or [128, 256, 1024]:
    eng = StepEngine(spec, arena, GazePreprocessor(device=dev), gaze_method="Reg", dp_method="None")
    base = make_batch(min(B, 64), seed=1)
    reps = (B + 63) // 64
    batch = tuple(t.repeat((reps,) + (1,) * (t.dim() - 1))[:B].contiguous().to(dev) for t in base)
    n = [0]

    def step():
        n[0] += 1
        eng.zero_grad(); eng.forward(*batch); eng.backward(); eng.adam_step(n[0], 5e-4)

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    host = []
    for _ in range(10):
        torch.cuda.synchronize()
        t0 = time.perf_counter(); step(); host.append(time.perf_counter() - t0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        step()
    e1.record(); torch.cuda.synchronize()
    dev_ms = e0.elapsed_time(e1) / 20
    print(f"B={B}: host enqueue {1e3 * sorted(host)[len(host) // 2]:.3f} ms/step, device {dev_ms:.3f} ms/step, "
          f"{B / dev_ms:.1f}k samples/s", flush=True)
