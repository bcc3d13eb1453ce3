"""One training step on cuda:0 with a sync + error poll after every phase (bring-up aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import _lib
from tests.test_step_gpu import build
from tests.synth import make_batch
gm = sys.argv[1] if len(sys.argv) > 1 else "None"
dp = sys.argv[2] if len(sys.argv) > 2 else "None"
B = int(sys.argv[3]) if len(sys.argv) > 3 else 6
spec, arena, eng = build(True, gm, dp)
batch = [t.to("cuda:0") for t in make_batch(B, seed=21)]
for it in range(3):
    eng.zero_grad()
    eng.forward(*batch)
    torch.cuda.synchronize(); print("fwd ok", eng.losses_host(), flush=True)
    eng.backward()
    torch.cuda.synchronize(); print("bwd ok", flush=True)
    eng.adam_step(it + 1, 1e-4)
    torch.cuda.synchronize(); print("adam ok", flush=True)
_lib.poll_error()
print("DONE")
