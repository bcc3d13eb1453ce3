"""Upper bounds for fusion work: the graph-replayed training step at batch argv[1] with one kernel family at a
time replaced by a no-op (results are WRONG in those runs; only the step time is read).  Says how much of the
step each family really costs once the side-stream overlap and launch gaps are accounted for.

    python tools/exp_skip.py 1024
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import gabril_carla_b200  # noqa: E402,F401
from bench import synth_batch  # noqa: E402
from gabril_carla_b200 import ops  # noqa: E402
from gabril_carla_b200.engine import StepEngine  # noqa: E402
from gabril_carla_b200.gaze import GazePreprocessor  # noqa: E402
from gabril_carla_b200.model import ParamArena, PolicySpec  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = torch.device("cuda:0")
base = synth_batch(min(B, 256), seed=300)
reps = (B + base[0].shape[0] - 1) // base[0].shape[0]
batch = tuple(t.repeat((reps,) + (1,) * (t.dim() - 1))[:B].contiguous().to(dev) for t in base)


def run(tag, patch):
    saved = {k: getattr(ops, k) for k in patch}
    for k, v in patch.items():
        setattr(ops, k, v)
    try:
        spec = PolicySpec(2)
        arena = ParamArena(spec, dev)
        arena.init_reference_style(seed=42)
        eng = StepEngine(spec, arena, GazePreprocessor(device=str(dev)), gaze_method="Reg", dp_method="None",
                         grayscale=True, seed=0)
        eng.use_graph = True
        n = [0]

        def step():
            n[0] += 1
            eng.train_step(*batch)
            eng.adam_step(n[0], 5e-4)
        for _ in range(4):
            step()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                step()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 20)
        print(f"{tag:28s} {best:8.3f} ms/step  {B / best:8.1f} k samples/s", flush=True)
        del eng, arena
        torch.cuda.empty_cache()
    finally:
        for k, v in saved.items():
            setattr(ops, k, v)


def nop(*a, **k):
    return None


run("full step", {})
run("no colsum", {"colsum_bf16": nop})
run("no reg_dist", {"reg_dist": nop, "reg_mse": nop})
run("no colsum, no reg_dist", {"colsum_bf16": nop, "reg_dist": nop, "reg_mse": nop})
run("no gaze_splat", {"gaze_splat": nop})
