"""Golden vectors for the inference path from the UNMODIFIED reference modules
(imported through oracle/ref_import.py): the statements of BCAgent._process_observation and
_predict_control (vlm_gaze/eval/my_agents/bc_agent.py:228-305) executed on the reference's own
Encoder (in eval()), apply_gmd_dropout(test_mode=True), pre_actor and actor for dp_method None /
GMD / IGMD, gray and RGB, with a synthetic predicted gaze map g (values outside [0,1] included, to
exercise the clipping of :278-279).  bc_agent.py itself imports carla and cannot be imported here;
its model-side statements are reproduced verbatim below.  Run in the build container only:
    python tests/golden/make_golden_agent.py
"""
import os
import sys

import numpy as np
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402
from tests.synth import make_batch, make_weights  # noqa: E402


def synthetic_g(B, seed):
    g = torch.Generator().manual_seed(seed)
    low = torch.rand(B, 1, 5, 8, generator=g) * 1.4 - 0.2          # some values < 0 and > 1
    return torch.nn.functional.interpolate(low, size=(180, 320), mode="bilinear", align_corners=False)


def main():
    ref = ref_import.load()
    B = 2
    img, _, _ = make_batch(B, seed=41)
    out = {}
    for gray in (True, False):
        tag = "gray" if gray else "rgb"
        c_in = 2 if gray else 6
        w = make_weights(c_in, seed=13)
        encoder = ref.Encoder(c_in, 64, 128, 2, 32)
        pre_actor = nn.Sequential(nn.Flatten(start_dim=1), nn.Linear(20 * 38 * 64, 256))
        actor = nn.Sequential(nn.Linear(256, 256), nn.ReLU(), nn.Linear(256, 7))
        encoder.load_state_dict(w["encoder"]); pre_actor.load_state_dict(w["pre_actor"]); actor.load_state_dict(w["actor"])
        encoder.eval()
        # ---- _process_observation (:250-269) on an already stacked uint8 [B,S,H,W,3] ----
        x = img.float() / 255.0
        x = x.permute(0, 1, 4, 2, 3)
        if gray:
            x = (0.299 * x[:, :, 0:1] + 0.587 * x[:, :, 1:2] + 0.114 * x[:, :, 2:3])
        Bb, S, C, H, W = x.shape
        model_input = x.reshape(Bb, S * C, H, W)
        for dp in ("None", "GMD", "IGMD"):
            with torch.no_grad():
                g = None
                if dp in ("GMD", "IGMD"):
                    g = synthetic_g(B, 5).clone()
                    g[g < 0] = 0
                    g[g > 1] = 1
                dropout_mask = g[:, -1:] if dp == "IGMD" else None
                z = encoder(model_input, dropout_mask=dropout_mask)
                if dp == "GMD":
                    z = ref.apply_gmd_dropout(z, g[:, -1:], test_mode=True)
                action = actor(pre_actor(z))
            out[f"{tag}/{dp}/action"] = action.numpy().copy()
            out[f"{tag}/{dp}/z_sub"] = z[:, ::8, ::3, ::5].numpy().copy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_vectors_agent.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items() if k.endswith("action")})


if __name__ == "__main__":
    main()
