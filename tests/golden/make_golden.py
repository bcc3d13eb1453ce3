#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference
through oracle/ref_import.py, CPU fp32, use_amp=false, use_compile=false) on the seeded synthetic
inputs of tests/synth.py.  The reference's own tests pin no numbers for this path, so
these vectors are what pins the oracle (tests/test_oracle_cpu.py) and the CUDA path (tests -m gpu).

    python tests/golden/make_golden.py        # needs /root/reference; rewrites the fixtures

Inputs are not stored (they are regenerated from the seeds; a sha256 of their bytes is stored and
checked), big outputs are stored as the non-zero band / a strided subsample + a sha256 of the whole.
"""
import hashlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402
from tests.synth import make_batch, make_weights  # noqa: E402


def sha(t: torch.Tensor) -> str:
    return hashlib.sha256(t.contiguous().numpy().tobytes()).hexdigest()


def load_weights(trainer, w):
    trainer.encoder.load_state_dict(w["encoder"])
    trainer.pre_actor.load_state_dict(w["pre_actor"])
    trainer.actor.load_state_dict(w["actor"])


def grad_summary(module, prefix, out):
    for k, p in module.named_parameters():
        g = p.grad.detach().float()
        out[f"{prefix}/{k}/norm"] = np.float64(g.double().norm().item())
        out[f"{prefix}/{k}/sum"] = np.float64(g.double().sum().item())
        out[f"{prefix}/{k}/head"] = g.flatten()[:256].numpy().copy()


def main():
    ref = ref_import.load()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    torch.manual_seed(0)
    B = 4
    img, gaze, act = make_batch(B, seed=11)
    out = {"in/img_sha": sha(img), "in/gaze": gaze.numpy(), "in/act": act.numpy()}

    # ---- preprocessing: prepare_for_bc gray / RGB, aggregate on / off; forward() ----
    gp = ref.GazePreprocessor(img_height=180, img_width=320, gaze_sigma=30.0, gaze_coeff=0.8,
                              maxpoints=5, device="cpu", temporal_alpha=0.7)
    for gray in (True, False):
        obs, heat, center = gp.prepare_for_bc(img, gaze, frame_stack=2, grayscale=gray, aggregate_stack=True)
        tag = "gray" if gray else "rgb"
        out[f"prep/{tag}/obs_sha"] = sha(obs)
        out[f"prep/{tag}/obs_sub"] = obs[:, :, ::9, ::8].numpy().copy()
        out[f"prep/{tag}/center"] = np.int64(center)
    out["prep/heat_band"] = heat[:, :, 60:120].numpy().copy()
    out["prep/heat_outside_absmax"] = np.float32(max(heat[:, :, :60].abs().max().item(),
                                                     heat[:, :, 120:].abs().max().item()))
    _, heat_noagg, _ = gp.prepare_for_bc(img, gaze, frame_stack=2, grayscale=True, aggregate_stack=False)
    out["prep/heat_noagg_band"] = heat_noagg[:, :, 60:120].numpy().copy()
    base = gp.forward(gaze.view(B, 2, 5, 2))          # per-frame maps [B,T,1,H,W]
    out["prep/base_band"] = base[:, :, 0, 60:120].numpy().copy()
    out["prep/kernel_1d"] = gp.kernel_1d.flatten().numpy().copy()
    # longer sequence than the stack (L=5 > S=2) and shorter (L=1): integer window gather
    img5, gaze5, _ = make_batch(2, L=5, seed=12, edge_cases=False)
    obs5, heat5, c5 = gp.prepare_for_bc(img5, gaze5, frame_stack=2, grayscale=True)
    out["prep/L5/img_sha"] = sha(img5)
    out["prep/L5/gaze"] = gaze5.numpy()
    out["prep/L5/obs_sha"] = sha(obs5)
    out["prep/L5/heat_band"] = heat5[:, :, 60:120].numpy().copy()
    out["prep/L5/center"] = np.int64(c5)
    img1, gaze1, _ = make_batch(2, L=1, seed=13, edge_cases=False)
    obs1, heat1, c1 = gp.prepare_for_bc(img1, gaze1, frame_stack=2, grayscale=True)
    out["prep/L1/img_sha"] = sha(img1)
    out["prep/L1/gaze"] = gaze1.numpy()
    out["prep/L1/obs_sha"] = sha(obs1)
    out["prep/L1/heat_band"] = heat1[:, :, 60:120].numpy().copy()
    # sigma = 20 (train_bc_confounded.yaml:19): band rows 40..139
    gp20 = ref.GazePreprocessor(img_height=180, img_width=320, gaze_sigma=20.0, gaze_coeff=0.8,
                                maxpoints=5, device="cpu", temporal_alpha=0.7)
    _, heat20, _ = gp20.prepare_for_bc(img, gaze, frame_stack=2, grayscale=True)
    out["prep/sigma20/heat_band"] = heat20[:, :, 40:140, ::2].numpy().copy()

    # ---- get_gaze_mask fwd + grad, apply_gmd_dropout ----
    g = torch.Generator().manual_seed(21)
    z = torch.randn(B, 64, 20, 38, generator=g).requires_grad_(True)
    m = ref.get_gaze_mask(z, 50.0, (180, 320))
    out["mask/z"] = z.detach().numpy().copy()
    out["mask/out_sub"] = m.detach()[:, 0, ::3, ::4].numpy().copy()
    out["mask/out_sum"] = np.float64(m.detach().double().sum().item())
    tgt = heat[:, -1:]
    loss = torch.nn.functional.mse_loss(tgt, m)
    loss.backward()
    out["mask/mse"] = np.float32(loss.item())
    out["mask/dz"] = z.grad.numpy().copy()
    A = torch.rand(B, 1, 20, 38, generator=g)
    out["gmd/A"] = A.numpy().copy()
    real_rand = torch.rand
    try:
        torch.rand = lambda *a, **k: A.clone()
        zz = z.detach()
        out["gmd/train"] = ref.apply_gmd_dropout(zz, heat, test_mode=False)[:, :4].numpy().copy()
        out["gmd/test"] = ref.apply_gmd_dropout(zz, heat, test_mode=True)[:, :4].numpy().copy()
    finally:
        torch.rand = real_rand

    # ---- BCTrainer.compute_loss + backward: losses and gradient summaries ----
    for gray, gm, dp in ((True, "None", "None"), (True, "Reg", "None"), (True, "Reg", "GMD"),
                         (False, "Reg", "GMD")):
        cfg = ref_import.default_cfg(**{"gaze.method": gm, "dropout.method": dp, "model.grayscale": gray})
        tr = ref_import.make_trainer(cfg, seed=42)
        load_weights(tr, make_weights(2 if gray else 6, seed=7))
        tr.set_train_mode()
        batch = {"obs": {"image": img, cfg.data.gaze_key if hasattr(cfg.data, "gaze_key") else "gaze_coords": gaze},
                 "actions": act}
        try:
            torch.rand = lambda *a, **k: A.clone()
            total, bs, metrics = tr.compute_loss(batch)
        finally:
            torch.rand = real_rand
        total.backward()
        tag = f"step/{'gray' if gray else 'rgb'}_{gm}_{dp}"
        out[f"{tag}/total"] = np.float64(total.item())
        out[f"{tag}/actor"] = np.float64(metrics["Loss/actor"])
        out[f"{tag}/reg"] = np.float64(metrics["Loss/reg"])
        out[f"{tag}/bs"] = np.int64(bs)
        grad_summary(tr.encoder, f"{tag}/encoder", out)
        grad_summary(tr.pre_actor, f"{tag}/pre_actor", out)
        grad_summary(tr.actor, f"{tag}/actor_head", out)
        print(tag, total.item(), metrics)

    # ---- Adam + cosine_warmup schedule trajectory (optim.py:11-28,60-84; base_trainer.py:138-149) ----
    g2 = torch.Generator().manual_seed(31)
    w = torch.nn.Parameter(torch.randn(64, generator=g2))
    cfg = ref_import.default_cfg()
    cfg.training.epochs = 4
    opt = ref.build_optimizer([w], cfg.optimizer)
    sched, per_step = ref.build_scheduler(opt, 300, cfg.training, cfg.scheduler, 1)   # T = 1200 steps
    grads = torch.randn(1200, 64, generator=g2) * 0.1
    out["sched/w0"] = w.detach().numpy().copy()
    out["sched/grads_sha"] = sha(grads)
    lrs, snaps = [], {}
    for i in range(1200):
        lrs.append(opt.param_groups[0]["lr"])
        w.grad = grads[i].clone()
        opt.step()
        if per_step:
            sched.step()
        if i + 1 in (1, 2, 10, 600, 1200):
            snaps[i + 1] = w.detach().numpy().copy()
    out["sched/lr"] = np.asarray(lrs, dtype=np.float64)
    for k, v in snaps.items():
        out[f"sched/w_after_{k}"] = v

    path = os.path.join(HERE, "reference_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
