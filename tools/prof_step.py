"""Profiling target: two warm-up steps then ONE training step at a given batch (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from tests.test_step_gpu import build
from tests.synth import make_batch
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
gm = sys.argv[2] if len(sys.argv) > 2 else "Reg"
dp = sys.argv[3] if len(sys.argv) > 3 else "None"
spec, arena, eng = build(True, gm, dp)
base = make_batch(min(B, 32), seed=1)
batch = [t.repeat((B // t.shape[0],) + (1,) * (t.dim() - 1)).to("cuda:0") for t in base]
for it in range(3):
    eng.zero_grad(); eng.forward(*batch); eng.backward(); eng.adam_step(it + 1, 1e-4)
torch.cuda.synchronize()
print("done", eng.losses_host())
