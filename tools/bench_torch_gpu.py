"""BASELINE.md §2 "B2 — the kernel bar": the reference model as stock PyTorch (cuDNN / cuBLAS) on the
SAME B200, so that the tcgen05 path has a same-box number to beat.  A TOOL: nothing in the package or
in bench.py imports it.

The model below is a plain torch.nn restatement of the reference architecture (Encoder
vlm_gaze/models/linear_models.py:124-207 with the in-place-ReLU Residual of :78-121, heads
train_bc.py:79-82, get_gaze_mask gaze_utils.py:7-30, heat maps data_loader_robomimic.py:85-139,
204-278, Reg-MSE loss train_bc.py:145-169,290-297) on the reference's execution path: torch.autocast
+ cuDNN convolutions + torch.optim.Adam (train_bc.py:100-115,241; base_trainer.py:133-143), eager and
(if inductor works offline) torch.compile.  /root/reference is absent on the GPU box, hence the
restatement; tests/test_b2_model_cpu.py checks it against the oracle on CPU.

  python tools/bench_torch_gpu.py [--batch 1024 128] [--steps 10] [--out gpurun_out/baseline_torch_gpu.json]
"""
import argparse
import json
import math
import os
import sys
import time

import torch
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


class Residual(nn.Module):
    def __init__(self, c, rh):
        super().__init__()
        self._block = nn.Sequential(nn.ReLU(True), nn.Conv2d(c, rh, 3, 1, 1, bias=False), nn.ReLU(True),
                                    nn.Conv2d(rh, c, 1, 1, bias=False))

    def forward(self, x):
        return x + self._block(x)     # the in-place ReLU rectifies x first (quirk F7)


class Encoder(nn.Module):
    def __init__(self, cin, emb=64, nh=128, nres=2, rh=32):
        super().__init__()
        self._conv_1 = nn.Conv2d(cin, nh // 4, 4, 2, 1)
        self._conv_2 = nn.Conv2d(nh // 4, nh // 2, 4, 2, 1)
        self._conv_3 = nn.Conv2d(nh // 2, nh, 4, 2, 1)
        self._conv_4 = nn.Conv2d(nh, nh, 3, 1, 0)
        self._res = nn.ModuleList([Residual(nh, rh) for _ in range(nres)])
        self._conv_5 = nn.Conv2d(nh, nh, 5, 1, 2)
        self._conv_6 = nn.Conv2d(nh, emb, 5, 1, 2)

    def forward(self, x):
        x = F.relu(self._conv_1(x))
        x = F.relu(self._conv_2(x))
        x = F.relu(self._conv_3(x))
        x = self._conv_4(x)
        for r in self._res:
            x = r(x)
        x = F.relu(x)
        x = F.relu(self._conv_5(x))
        return self._conv_6(x)


def gaussian_kernel_1d(sigma, device):
    size = int(4 * sigma + 1)
    size += 1 - size % 2
    x = torch.arange(size, dtype=torch.float32, device=device) - size // 2
    k = torch.exp(-x ** 2 / (2 * sigma ** 2))
    return k / k.sum()


def heat_maps(gaze, H, W, sigma=30.0, alpha=0.7, P=5):
    """[B,S,P*2] -> [B,S,H,W]: scatter + banded separable blur + min-max, alpha-decay aggregate, min-max."""
    B, S = gaze.shape[:2]
    g = gaze.reshape(B, S, P, 2)
    gx, gy = g[..., 0], g[..., 1]
    valid = (gx >= 0) & (gy >= 0)
    xi = (gx.clamp(0, 1) * (W - 1)).long().clamp(0, W - 1)
    yi = (gy.clamp(0, 1) * (H - 1)).long().clamp(0, H - 1)
    flat = torch.zeros(B * S, H * W, device=gaze.device)
    flat.scatter_add_(1, (yi * W + xi).reshape(B * S, P), valid.float().reshape(B * S, P))
    k = gaussian_kernel_1d(sigma, gaze.device)
    R = k.numel() // 2
    d = flat.reshape(B * S, 1, H, W)
    raw = F.conv2d(F.conv2d(d, k.view(1, 1, -1, 1), padding=(0, R)), k.view(1, 1, 1, -1), padding=(R, 0))

    def mm(x):
        lo, hi = x.amin(dim=(-2, -1), keepdim=True), x.amax(dim=(-2, -1), keepdim=True)
        return (x - lo) / (hi - lo + 1e-8)
    base = mm(raw).reshape(B, S, H, W)
    rows = [base[:, 0]]
    for s in range(1, S):
        w = torch.tensor([alpha ** (s - j) for j in range(s + 1)], device=gaze.device)
        rows.append((base[:, :s + 1] * w.view(1, s + 1, 1, 1)).sum(1))
    return mm(torch.stack(rows, 1))


def format_obs(frames, gray):
    x = frames.float() / 255.0
    x = x.permute(0, 1, 4, 2, 3)
    if gray:
        x = 0.299 * x[:, :, 0:1] + 0.587 * x[:, :, 1:2] + 0.114 * x[:, :, 2:3]
    B, S, C, H, W = x.shape
    return x.reshape(B, S * C, H, W)


def gaze_mask(z, beta, size):
    B, C, h, w = z.shape
    p = F.softmax(z.abs().sum(1).reshape(B, h * w) / beta, dim=-1).reshape(B, 1, h, w)
    up = F.interpolate(p, size=size, mode="bicubic")
    flat = up.reshape(B, 1, -1)
    hi = flat.max(-1).values.reshape(B, 1, 1, 1)
    lo = flat.min(-1).values.reshape(B, 1, 1, 1)
    return (up - lo) / (hi - lo)


class Policy(nn.Module):
    def __init__(self, cin):
        super().__init__()
        self.encoder = Encoder(cin)
        self.pre_actor = nn.Sequential(nn.Flatten(), nn.Linear(48640, 256))
        self.actor = nn.Sequential(nn.Linear(256, 256), nn.ReLU(), nn.Linear(256, 7))

    def losses(self, obs, heat, actions, use_reg=True, beta=50.0, lam=10.0):
        z = self.encoder(obs)
        logits = self.actor(self.pre_actor(z))
        actor = F.mse_loss(logits.float(), actions)
        reg = torch.zeros((), device=obs.device)
        if use_reg:
            g2 = gaze_mask(z, beta, obs.shape[-2:])      # fp32 islands under autocast: softmax, sum
            reg = F.mse_loss(heat[:, -1:].float(), g2.float())
        return lam * reg + actor, actor, reg


def synth(B, device, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    frames = torch.randint(0, 256, (min(B, 64), 2, 180, 320, 3), dtype=torch.uint8, generator=g)
    gaze = torch.rand(min(B, 64), 2, 10, generator=g)
    gaze[torch.rand(min(B, 64), 2, 10, generator=g) < 0.2] = -1.0
    act = torch.rand(min(B, 64), 2, 7, generator=g)
    reps = (B + 63) // 64
    return tuple(t.repeat((reps,) + (1,) * (t.dim() - 1))[:B].contiguous().to(device) for t in (frames, gaze, act))


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def bench_step(B, steps, mode, channels_last, include_prep, compile_=False, use_reg=True):
    dev = "cuda:0"
    torch.manual_seed(42)
    model = Policy(2).to(dev)
    if channels_last:
        model = model.to(memory_format=torch.channels_last)
    opt = torch.optim.Adam(model.parameters(), lr=5e-4)
    frames, gaze, act = synth(B, dev)
    loss_fn = model.losses
    if compile_:
        loss_fn = torch.compile(model.losses)
    obs0, heat0 = format_obs(frames, True), heat_maps(gaze, 180, 320)
    if channels_last:
        obs0 = obs0.contiguous(memory_format=torch.channels_last)

    def step():
        if include_prep:
            obs, heat = format_obs(frames, True), heat_maps(gaze, 180, 320)
            if channels_last:
                obs = obs.contiguous(memory_format=torch.channels_last)
        else:
            obs, heat = obs0, heat0
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode == "bf16")):
            total, actor, reg = loss_fn(obs, heat, act[:, -1], use_reg)
        opt.zero_grad(set_to_none=True)
        total.backward()
        opt.step()
    ms = timed(step, steps)
    return {"batch": B, "mode": mode, "channels_last": channels_last, "with_preprocessing": include_prep,
            "compile": compile_, "gaze_method": "Reg" if use_reg else "None", "ms_per_step": ms,
            "samples_per_s": B / ms * 1e3}


LAYERS = [  # name, Cin, Cout, k, stride, pad, H, W (input)
    ("conv_1", 2, 32, 4, 2, 1, 180, 320), ("conv_2", 32, 64, 4, 2, 1, 90, 160), ("conv_3", 64, 128, 4, 2, 1, 45, 80),
    ("conv_4", 128, 128, 3, 1, 0, 22, 40), ("res3x3", 128, 32, 3, 1, 1, 20, 38), ("res1x1", 32, 128, 1, 1, 0, 20, 38),
    ("conv_5", 128, 128, 5, 1, 2, 20, 38), ("conv_6", 128, 64, 5, 1, 2, 20, 38)]


def bench_layers(B, steps):
    dev, out = "cuda:0", []
    for name, ci, co, k, st, pd, H, W in LAYERS:
        x = torch.randn(B, ci, H, W, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = torch.randn(co, ci, k, k, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
        y = F.conv2d(x, w, None, st, pd)
        dy = torch.randn_like(y)
        flop = 2.0 * y.numel() * ci * k * k
        args = (dy, x, w, None, [st, st], [pd, pd], [1, 1], False, [0, 0], 1)
        t_f = timed(lambda: F.conv2d(x, w, None, st, pd), steps)
        t_d = timed(lambda: torch.ops.aten.convolution_backward(*args, [True, False, False]), steps)
        t_w = timed(lambda: torch.ops.aten.convolution_backward(*args, [False, True, False]), steps)
        out.append({"layer": name, "batch": B, "fwd_ms": t_f, "dgrad_ms": t_d, "wgrad_ms": t_w,
                    "fwd_tflops": flop / t_f / 1e9, "dgrad_tflops": flop / t_d / 1e9, "wgrad_tflops": flop / t_w / 1e9})
        print(json.dumps(out[-1]), flush=True)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, nargs="+", default=[1024, 128])
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--out", default="gpurun_out/baseline_torch_gpu.json")
    ap.add_argument("--no-compile", action="store_true")
    a = ap.parse_args()
    torch.backends.cudnn.benchmark = True
    res = {"torch": torch.__version__, "cudnn": torch.backends.cudnn.version(), "gpu": torch.cuda.get_device_name(0),
           "steps": [], "layers": []}
    for B in a.batch:
        for mode, cl, prep in (("bf16", True, True), ("bf16", True, False), ("bf16", False, True), ("fp32", False, True)):
            if mode == "fp32" and B > 512:
                pass
            try:
                r = bench_step(B, a.steps, mode, cl, prep)
            except Exception as e:  # noqa: BLE001
                r = {"batch": B, "mode": mode, "channels_last": cl, "with_preprocessing": prep, "error": repr(e)[:200]}
            res["steps"].append(r)
            print(json.dumps(r), flush=True)
            torch.cuda.empty_cache()
        if not a.no_compile:
            t0 = time.time()
            try:
                r = bench_step(B, a.steps, "bf16", True, False, compile_=True)
                r["compile_wall_s"] = time.time() - t0
            except Exception as e:  # noqa: BLE001
                r = {"batch": B, "mode": "bf16", "compile": True, "error": repr(e)[:300]}
            res["steps"].append(r)
            print(json.dumps(r), flush=True)
            torch.cuda.empty_cache()
    for B in a.batch:
        res["layers"] += bench_layers(B, a.steps)
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(res, f, indent=1)
