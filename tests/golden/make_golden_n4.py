#!/usr/bin/env python
"""Golden vectors for SURVEY.md §8f row N4 — the other gaze methods of train_bc.py — from the UNMODIFIED
reference imported through oracle/ref_import.py (CPU fp32, use_amp=false, use_compile=false):

  Mask     enc_in = xx * gg_for_mul                                   train_bc.py:260-261
  ViSaRL   enc_in = cat([xx, gg_for_cat]), 2x input channels           train_bc.py:60-61,262-264
  AGIL     second encoder on xx * gg_for_mul, z = 0.5 (z + z_agil)    train_bc.py:70-77,268-272
  GRIL     gaze-coordinate head on z_flat, clamped MSE                train_bc.py:84-86,182-192,293-294
each at gaze.ratio 1.0 and 0.5 (tests/synth.py::make_batch_n1: sample 2 does not use gaze at 0.5).

It also RECORDS that the two remaining methods cannot run in the reference itself:
  Contrastive (train_bc.py:173-180) slices image channels out of `gg`, which in this fork is the
              [B, S, H, W] heat-map stack: gg[:, S:] is empty and the encoder raises;
  Oreo        (train_bc.py:275-289) rearranges the quantizer's [B, h*w] index tensor with a 3-D pattern
              ('b h w -> (b h w)'): einops raises.
The exception types are stored so that tests/test_n4_cpu.py can state the fact without the reference.

    python tests/golden/make_golden_n4.py     # needs /root/reference; rewrites reference_vectors_n4.npz
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402
from tests.golden.make_golden import grad_summary, load_weights, sha  # noqa: E402
from tests.synth import make_batch_n1, make_weights, make_weights_gril  # noqa: E402


def main():
    ref_import.load()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    torch.manual_seed(0)
    img, gaze, act = make_batch_n1()
    out = {"in/img_sha": sha(img), "in/gaze": gaze.numpy(), "in/act": act.numpy()}
    batch = {"obs": {"image": img, "gaze_coords": gaze}, "actions": act}
    for method in ("Mask", "ViSaRL", "AGIL", "GRIL"):
        for ratio in (1.0, 0.5):
            cfg = ref_import.default_cfg(**{"gaze.method": method, "dropout.method": "None", "model.grayscale": True,
                                            "gaze.ratio": ratio})
            tr = ref_import.make_trainer(cfg, seed=42)
            tr.gaze_ratio = float(ratio)
            load_weights(tr, make_weights(4 if method == "ViSaRL" else 2, seed=7))
            if method == "AGIL":
                tr.encoder_agil.load_state_dict(make_weights(2, seed=8)["encoder"])
            if method == "GRIL":
                tr.gril_gaze_coord_predictor.load_state_dict(make_weights_gril(seed=9))
            tr.set_train_mode()
            total, bs, metrics = tr.compute_loss(batch)
            total.backward()
            tag = f"step/{method}_r{ratio}"
            out[f"{tag}/total"] = np.float64(total.item())
            out[f"{tag}/actor"] = np.float64(metrics["Loss/actor"])
            out[f"{tag}/reg"] = np.float64(metrics["Loss/reg"])
            grad_summary(tr.encoder, f"{tag}/encoder", out)
            grad_summary(tr.pre_actor, f"{tag}/pre_actor", out)
            grad_summary(tr.actor, f"{tag}/actor_head", out)
            if method == "AGIL":
                grad_summary(tr.encoder_agil, f"{tag}/encoder_agil", out)
            if method == "GRIL":
                grad_summary(tr.gril_gaze_coord_predictor, f"{tag}/gril", out)
            print(tag, total.item(), metrics)
    # ---- the two methods the reference itself cannot execute ----
    for method, dp in (("Contrastive", "None"), ("Reg", "Oreo")):
        cfg = ref_import.default_cfg(**{"gaze.method": method, "dropout.method": dp, "model.grayscale": True,
                                        "dropout.vqvae_path": "/nonexistent/vqvae.pth"})   # randomly initialised quantizer
        tr = ref_import.make_trainer(cfg, seed=42)
        load_weights(tr, make_weights(2, seed=7))
        tr.set_train_mode()
        try:
            tr.compute_loss(batch)
            err = "no error"
        except Exception as e:  # noqa: BLE001
            err = f"{type(e).__name__}: {str(e).splitlines()[0][:160]}"
        out[f"dead/{method}_{dp}"] = np.array(err)
        print("dead", method, dp, "->", err)
        assert err != "no error"
    path = os.path.join(HERE, "reference_vectors_n4.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, len(out), "arrays", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
