"""Generates tests/golden/reference_vectors_gp.npz by running the UNMODIFIED reference
(/root/reference, imported through oracle/ref_import.py) for BASELINE config 4: the gaze-predictor
AutoEncoder(Encoder, Decoder(1)) of train_gaze_predictor.py:50-64 with the loss of compute_loss
(:83-101) on a fixed synthetic batch.  Run in the build container only:
    python tests/golden/make_golden_gp.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402
from tests.synth import make_batch, make_weights_gp  # noqa: E402


def main():
    ref = ref_import.load()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    B = 2
    img, gaze, _ = make_batch(B, seed=31)
    out = {"in/gaze": gaze.numpy()}
    gp = ref.GazePreprocessor(img_height=180, img_width=320, gaze_sigma=30.0, gaze_coeff=0.8, maxpoints=5,
                              device="cpu", temporal_alpha=0.7)
    for gray in (True, False):
        tag = "gray" if gray else "rgb"
        c_in = 2 if gray else 6
        w = make_weights_gp(c_in, seed=9)
        model = ref.AutoEncoder(ref.Encoder(c_in, 64, 128, 2, 32), ref.Decoder(1, 64, 128, 2, 32))
        model.encoder.load_state_dict(w["encoder"])
        model.decoder.load_state_dict(w["decoder"])
        model.train()
        obs, target, _ = gp.prepare_for_gaze_predictor(obs_image_seq=img, gaze_seq=gaze, frame_stack=2, grayscale=gray)
        pred = model(obs)
        loss = torch.nn.MSELoss()(pred, target)
        loss.backward()
        out[f"{tag}/target_shape"] = np.array(target.shape)
        out[f"{tag}/target_band"] = target[:, 0, 60:120, ::2].numpy().copy()
        out[f"{tag}/loss"] = np.float32(loss.item())
        out[f"{tag}/pred_sub"] = pred.detach()[:, 0, ::6, ::8].numpy().copy()
        for mod_name, mod in (("encoder", model.encoder), ("decoder", model.decoder)):
            for k, p in mod.named_parameters():
                g = p.grad
                out[f"{tag}/grad_norm/{mod_name}.{k}"] = np.float32(g.norm().item())
                out[f"{tag}/grad_head/{mod_name}.{k}"] = g.flatten()[:16].numpy().copy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_vectors_gp.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, len(out), "arrays")


if __name__ == "__main__":
    main()
