"""Gaze-predictor pre-training throughput (BASELINE config 4) at batch B on one B200: CUDA-event timed
steps (forward + backward + Adam), inputs resident in HBM.  Usage: python tools/bench_gp.py [B] [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import json
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import _lib
from gabril_carla_b200.gaze import GazePreprocessor
from gabril_carla_b200.gaze_predictor import GazePredictorEngine
from gabril_carla_b200.model import GazePredictorSpec, ParamArena
from tests.synth import make_batch

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
dev = "cuda:0"
spec = GazePredictorSpec(2)
arena = ParamArena(spec, dev)
arena.init_reference_style(seed=1)
eng = GazePredictorEngine(spec, arena, GazePreprocessor(device=dev))
base = make_batch(min(B, 64), seed=7)
img, gaze, _ = [t.repeat((B // t.shape[0],) + (1,) * (t.dim() - 1)).to(dev) for t in base]


def step(i):
    eng.zero_grad(); eng.forward(img, gaze); eng.backward(); eng.adam_step(i, 5e-4)


for i in range(1, 4):
    step(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(4, 4 + steps):
    step(i)
e1.record()
torch.cuda.synchronize()
_lib.poll_error()
ms = e0.elapsed_time(e1) / steps
print(json.dumps({"workload": "train_gaze_predictor.py AutoEncoder step (gray stack 2, 180x320, MSE vs heat map, Adam)",
                  "per_gpu_batch": B, "ms_per_step": ms, "samples_per_s": B / ms * 1e3,
                  "algorithmic_tflops": 8.21e9 * B / ms / 1e9, "loss": eng.loss_host()}))
