"""Per-layer activation error of the CUDA forward against the UN-INJECTED bf16-emulating oracle (where does
the z error enter?).  python tools/diag_layers_emu.py [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import gabril_oracle as O
from tests.synth import make_batch
from tests.test_step_gpu import DEV, build, our_activations, rel
B = int(sys.argv[1]) if len(sys.argv) > 1 else 6
spec, arena, eng = build(True, "Reg", "None")
batch = make_batch(B, seed=21)
eng.zero_grad(); eng.forward(*(t.to(DEV) for t in batch)); torch.cuda.synchronize()
ours = our_activations(eng)
sds = {m: {k: v.cpu() for k, v in sd.items()} for m, sd in arena.state_dicts().items()}
obs, heat, c = O.prepare_for_bc(batch[0], batch[1], 2, True)
for emu in (True, False):
    taps = {}
    z = O.encoder_forward(sds["encoder"], obs, emulate_bf16=emu, taps=taps)
    if not emu:
        # fp32 oracle has no taps for the rectified intermediate tensors in the same places; only z
        print(f"fp32 oracle: z rel {rel(ours['z'], z):.3e}")
        continue
    for k in ["a1", "a2", "a3", "r-1", "h0", "r0", "h1", "r1", "a5", "z"]:
        a, b = ours[k], taps[k]
        d = (a - b).abs()
        nz = (d > 0).float().mean().item()
        print(f"bf16emu {k:4s} rel {rel(a, b):.3e}  frac differing {nz:.4f}  max|d| {d.max().item():.3e}  max|ref| {b.abs().max().item():.3e}")
