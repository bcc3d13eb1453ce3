import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import _lib
from tests.test_step_gpu import build
from tests.synth import make_batch
mode = sys.argv[1]
spec, arena, eng = build(True, "None", "None")
cpu_batch = make_batch(6, seed=21)
batch = [t.to("cuda:0") for t in cpu_batch]
torch.cuda.synchronize()
eng.zero_grad()
if mode == "persist":
    eng.forward(*batch)
else:
    eng.forward(*(t.to("cuda:0") for t in cpu_batch))
eng.backward()
torch.cuda.synchronize()
print(mode, "ok", eng.losses_host(), flush=True)
_lib.poll_error()
