"""Inference-path numbers: batch-1 control-step latency of BCPolicy (host frame in,
action vector out) eager vs CUDA-graph replay, and batched predict throughput.
python tools/bench_agent.py [--dp GMD]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gabril_carla_b200  # noqa: E402,F401
from gabril_carla_b200.agent import BCPolicy  # noqa: E402
from gabril_carla_b200.gaze import GazePreprocessor  # noqa: E402
from gabril_carla_b200.gaze_predictor import GazePredictorEngine  # noqa: E402
from gabril_carla_b200.model import GazePredictorSpec, ParamArena, PolicySpec  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--dp", default="None")
ap.add_argument("--steps", type=int, default=200)
args = ap.parse_args()
dev = "cuda:0"


def make(use_graph):
    spec = PolicySpec(2)
    arena = ParamArena(spec, dev)
    arena.init_reference_style(seed=42)
    gp = None
    if args.dp in ("GMD", "IGMD"):
        gs = GazePredictorSpec(2)
        ga = ParamArena(gs, dev)
        ga.init_reference_style(seed=43)
        gp = GazePredictorEngine(gs, ga, GazePreprocessor(device=dev), grayscale=True, frame_stack=2)
    return BCPolicy(spec, arena, dp_method=args.dp, grayscale=True, frame_stack=2, gaze_predictor=gp,
                    use_graph=use_graph)


rng = np.random.default_rng(0)
frames = [rng.integers(0, 256, (180, 320, 3), dtype=np.uint8) for _ in range(8)]
res = {"dp_method": args.dp}
for name, use_graph in (("eager", False), ("graph", True)):
    pol = make(use_graph)
    pol.reset()
    for i in range(10):
        pol.run_step(frames[i % 8])
    torch.cuda.synchronize()
    lat = []
    for i in range(args.steps):
        t0 = time.perf_counter()
        pol.run_step(frames[i % 8])
        lat.append(time.perf_counter() - t0)
    lat = np.sort(np.asarray(lat)) * 1e3
    res[name] = {"p50_ms": float(lat[len(lat) // 2]), "p95_ms": float(lat[int(len(lat) * 0.95)])}
# batched forward-only throughput
pol = make(False)
for B in (64, 256, 1024):
    x = torch.randint(0, 256, (B, 2, 180, 320, 3), device=dev, dtype=torch.uint8)
    for _ in range(3):
        pol.predict(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        pol.predict(x)
    e1.record(); torch.cuda.synchronize()
    res[f"predict_B{B}_samples_per_s"] = B * 10 / (e0.elapsed_time(e1) * 1e-3)
print(json.dumps(res))
