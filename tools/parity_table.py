"""End-to-end parity table of the CUDA training step against the CPU oracle, UN-INJECTED (the oracle runs
its own forward): losses, z, per-parameter gradients for BASELINE configs 1-3 and the RGB stack.

  python tools/parity_table.py [--batches 6 64] [--out profiles/parity_r2.txt]

Two oracles: fp32 (the reference's arithmetic) and bf16-emulating (weights / stored activations rounded
to bf16 like the tensor-core path, fp32 accumulation, fp32 gradient of that forward).  north_star's
tolerances are rel 1e-3 on losses and rel 1e-2 on gradients.  tests/test_step_gpu.py asserts the bounds
this table supports."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.synth import make_batch  # noqa: E402
from tests.test_step_gpu import DEV, build, oracle_step, rel  # noqa: E402

CONFIGS = [(True, "None", "None", "cfg1 gaze.method=None"), (True, "Reg", "None", "cfg2 Reg"),
           (True, "Reg", "GMD", "cfg3 Reg+GMD"), (False, "Reg", "GMD", "RGB Reg+GMD")]


def measure(gray, gm, dp, B, seed=21):
    spec, arena, eng = build(gray, gm, dp)
    batch = make_batch(B, seed=seed)
    A = None
    if dp == "GMD":
        A = torch.rand(B, 1, 20, 38, generator=torch.Generator().manual_seed(3))
        eng.injected_A = A.to(DEV)
    eng.zero_grad()
    eng.forward(*(t.to(DEV) for t in batch))
    eng.backward()
    actor, reg, total = eng.losses_host()
    grads = arena.grad_dicts()
    z = eng.buf["z"].float().cpu().permute(0, 3, 1, 2)
    rows = {}
    for tag, emu in (("fp32", False), ("bf16emu", True)):
        r_total, r_actor, r_reg, r_grads, ex = oracle_step(arena, batch, gray, gm, dp, A, emu, None)
        e = {"actor": abs(actor - r_actor) / abs(r_actor), "total": abs(total - r_total) / abs(r_total),
             "reg": abs(reg - r_reg) / abs(r_reg) if gm != "None" else 0.0, "z": rel(z, ex["z"].detach())}
        worst_w, worst_b, cos_min, num, den, per = 0.0, 0.0, 1.0, 0.0, 0.0, []
        for mod in grads:
            for k, gv in grads[mod].items():
                a, b = gv.cpu().flatten().double(), r_grads[mod][k].flatten().double()
                r = ((a - b).norm() / (b.norm() + 1e-300)).item()
                c = (a @ b / (a.norm() * b.norm() + 1e-300)).item()
                per.append((f"{mod}.{k}", r, c))
                if k.endswith("bias"):
                    worst_b = max(worst_b, r)
                else:
                    worst_w = max(worst_w, r)
                cos_min = min(cos_min, c)
                num += ((a - b) ** 2).sum().item(); den += (b ** 2).sum().item()
        e.update(grad_w=worst_w, grad_b=worst_b, cos=cos_min, grad_all=(num / den) ** 0.5, per=per)
        rows[tag] = e
    return rows


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", type=int, nargs="+", default=[6, 64])
    ap.add_argument("--out", default="")
    ap.add_argument("--per-param", action="store_true")
    a = ap.parse_args()
    lines = [f"# parity_table: CUDA step vs un-injected CPU oracle, GPU {torch.cuda.get_device_name(0)}",
             "# columns: relative errors of the three losses and of z; worst per-parameter gradient rel-l2 over "
             "weights / over biases; rel-l2 of the whole gradient arena; min per-parameter cosine",
             f"{'config':16s} {'B':>3s} {'oracle':8s} {'actor':>9s} {'reg':>9s} {'total':>9s} {'z':>9s} "
             f"{'grad_w':>9s} {'grad_b':>9s} {'grad_all':>9s} {'cos_min':>8s}"]
    for gray, gm, dp, name in CONFIGS:
        for B in a.batches:
            rows = measure(gray, gm, dp, B)
            for tag, e in rows.items():
                lines.append(f"{name:16s} {B:3d} {tag:8s} {e['actor']:9.2e} {e['reg']:9.2e} {e['total']:9.2e} "
                             f"{e['z']:9.2e} {e['grad_w']:9.2e} {e['grad_b']:9.2e} {e['grad_all']:9.2e} {e['cos']:8.5f}")
                if a.per_param:
                    for k, r, c in e["per"]:
                        lines.append(f"      {k:52s} rel {r:9.2e} cos {c:8.5f}")
            print("\n".join(lines[-2:]), flush=True)
    text = "\n".join(lines) + "\n"
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        open(a.out, "w").write(text)
    print(text)
