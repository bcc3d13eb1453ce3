"""On-GPU diagnostic: every dense contraction of the encoder (fwd / dgrad / wgrad) through the
tcgen05 path vs torch (fp32 math on the same bf16-rounded operands).  Prints one line per case and
a JSON summary to gpurun_out/diag_conv.json.  Usage: python tools/diag_conv.py [--naive] [--batch N]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

ap = argparse.ArgumentParser()
ap.add_argument("--naive", action="store_true")
ap.add_argument("--batch", type=int, default=3)
ap.add_argument("--only", type=str, default="")
args = ap.parse_args()
if args.naive:
    os.environ["GABRIL_NAIVE_CONV"] = "1"

import torch
import torch.nn.functional as F

import gabril_carla_b200  # noqa: F401
from gabril_carla_b200 import _lib, ops

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
_lib.init()

# (name, H, W, Cin, Cout, R, S, stride, pad)
LAYERS = [
    ("conv2", 90, 160, 32, 64, 4, 4, 2, 1),
    ("conv3", 45, 80, 64, 128, 4, 4, 2, 1),
    ("conv4", 22, 40, 128, 128, 3, 3, 1, 0),
    ("res3x3", 20, 38, 128, 32, 3, 3, 1, 1),
    ("res1x1", 20, 38, 32, 128, 1, 1, 1, 0),
    ("conv5", 20, 38, 128, 128, 5, 5, 1, 2),
    ("conv6", 20, 38, 128, 64, 5, 5, 1, 2),
    ("linear_small", 1, 1, 256, 256, 1, 1, 1, 0),
]
results = []


def rel_err(a, b):
    a = a.float(); b = b.float()
    return float((a - b).norm() / (b.norm() + 1e-30)), float((a - b).abs().max())


def record(name, what, got, ref, tol=2e-2):
    r, m = rel_err(got, ref)
    flag = _lib.load().gabril_poll_error()
    ok = (r < tol) and flag == 0 and bool(torch.isfinite(got.float()).all())
    print(f"{name:14s} {what:8s} rel={r:.3e} maxabs={m:.3e} flag={flag} {'OK' if ok else 'FAIL'}",
          flush=True)
    results.append(dict(layer=name, op=what, rel=r, maxabs=m, flag=flag, ok=ok))


def run_layer(name, H, W, Cin, Cout, R, S, stride, pad, N):
    g = torch.Generator(device="cpu").manual_seed(hash(name) % 1000)
    OH, OW = ops.conv_out_hw(H, W, R, S, stride, pad)
    x = (torch.randn(N, H, W, Cin, generator=g) * 0.5).to(dev).to(torch.bfloat16)
    w_oihw = (torch.randn(Cout, Cin, R, S, generator=g) / (Cin * R * S) ** 0.5).to(dev).to(torch.bfloat16)
    bias = torch.randn(Cout, generator=g).to(dev)
    w_int = ops.w_oihw_to_internal(w_oihw)
    x_nchw = x.float().permute(0, 3, 1, 2)
    # ---- fwd with bias + relu
    y = torch.full((N, OH, OW, Cout), float("nan"), device=dev, dtype=torch.bfloat16)
    ops.conv_fwd(x, w_int, bias, y, R, S, stride, pad, relu=True)
    torch.cuda.synchronize()
    ref = F.relu(F.conv2d(x_nchw, w_oihw.float(), bias, stride=stride, padding=pad)).permute(0, 2, 3, 1)
    record(name, "fwd", y, ref)
    # ---- fwd with resid + rowmask, no relu
    resid = (torch.randn(N, OH, OW, Cout, generator=g)).to(dev).to(torch.bfloat16)
    rowmask = (torch.rand(N * OH * OW, generator=g) > 0.3).float().to(dev)
    y2 = torch.full((N, OH, OW, Cout), float("nan"), device=dev, dtype=torch.bfloat16)
    ops.conv_fwd(x, w_int, None, y2, R, S, stride, pad, relu=False, resid=resid, rowmask=rowmask)
    torch.cuda.synchronize()
    ref2 = (F.conv2d(x_nchw, w_oihw.float(), None, stride=stride, padding=pad).permute(0, 2, 3, 1)
            + resid.float()) * rowmask.view(N, OH, OW, 1)
    record(name, "fwd_rm", y2, ref2)
    # ---- dgrad with addend + gate
    dy = (torch.randn(N, OH, OW, Cout, generator=g) * 0.5).to(dev).to(torch.bfloat16)
    addend = torch.randn(N, H, W, Cin, generator=g).to(dev).to(torch.bfloat16)
    gate = torch.randn(N, H, W, Cin, generator=g).to(dev).to(torch.bfloat16)
    dx = torch.full((N, H, W, Cin), float("nan"), device=dev, dtype=torch.bfloat16)
    ops.conv_dgrad(dy, w_int, dx, R, S, stride, pad, addend=addend, gate=gate)
    torch.cuda.synchronize()
    dy_nchw = dy.float().permute(0, 3, 1, 2)
    dref = torch.nn.grad.conv2d_input((N, Cin, H, W), w_oihw.float(), dy_nchw, stride=stride,
                                      padding=pad).permute(0, 2, 3, 1)
    dref = (dref + addend.float()) * (gate.float() > 0)
    record(name, "dgrad", dx, dref)
    # ---- wgrad (atomic accumulate on zeros)
    dw = torch.zeros(Cout, R * S, Cin, device=dev, dtype=torch.float32)
    ops.conv_wgrad(dy, x, dw, R, S, stride, pad)
    torch.cuda.synchronize()
    wref = torch.nn.grad.conv2d_weight(x_nchw, (Cout, Cin, R, S), dy_nchw, stride=stride, padding=pad)
    wref = ops.w_oihw_to_internal(wref)
    record(name, "wgrad", dw, wref)


t0 = time.time()
for L in LAYERS:
    if args.only and L[0] not in args.only.split(","):
        continue
    try:
        run_layer(*L, N=args.batch)
    except Exception as e:  # keep going: one call must report every layer
        print(f"{L[0]:14s} EXCEPTION {type(e).__name__}: {e}", flush=True)
        results.append(dict(layer=L[0], op="exception", ok=False, err=str(e)))

# pre_actor-shaped linear: fwd split-K atomic fp32, dgrad (MN-major weights), wgrad direct store
try:
    if not args.only or "pre_actor" in args.only:
        g = torch.Generator(device="cpu").manual_seed(7)
        Bn, K, Nf = 96, 48640, 256
        z = (torch.randn(Bn, K, generator=g) * 0.5).to(dev).to(torch.bfloat16)
        wp = (torch.randn(Nf, K, generator=g) / K ** 0.5).to(dev).to(torch.bfloat16)
        out = torch.zeros(Bn, Nf, device=dev, dtype=torch.float32)
        ops.conv_fwd(z.view(Bn, 1, 1, K), wp.view(Nf, 1, K), None, out, 1, 1, 1, 0,
                     out_f32_atomic=True, splits=19)
        torch.cuda.synchronize()
        record("pre_actor", "fwd", out, z.float() @ wp.float().t(), tol=1e-3)
        dzf = (torch.randn(Bn, Nf, generator=g)).to(dev).to(torch.bfloat16)
        dz = torch.full((Bn, 1, 1, K), float("nan"), device=dev, dtype=torch.bfloat16)
        ops.conv_dgrad(dzf.view(Bn, 1, 1, Nf), wp.view(Nf, 1, K), dz, 1, 1, 1, 0)
        torch.cuda.synchronize()
        record("pre_actor", "dgrad", dz.view(Bn, K), dzf.float() @ wp.float())
        dwp = torch.full((Nf, 1, K), float("nan"), device=dev, dtype=torch.float32)
        ops.conv_wgrad(dzf.view(Bn, 1, 1, Nf), z.view(Bn, 1, 1, K), dwp, 1, 1, 1, 0, direct_store=True)
        torch.cuda.synchronize()
        record("pre_actor", "wgrad", dwp.view(Nf, K), dzf.float().t() @ z.float(), tol=1e-3)
except Exception as e:
    print(f"pre_actor EXCEPTION {type(e).__name__}: {e}", flush=True)
    results.append(dict(layer="pre_actor", op="exception", ok=False, err=str(e)))

os.makedirs("gpurun_out", exist_ok=True)
tag = "naive" if args.naive else "tc"
with open(f"gpurun_out/diag_conv_{tag}.json", "w") as f:
    json.dump(results, f, indent=1)
n_ok = sum(1 for r in results if r["ok"])
print(f"SUMMARY {tag}: {n_ok}/{len(results)} ok in {time.time() - t0:.1f}s")
