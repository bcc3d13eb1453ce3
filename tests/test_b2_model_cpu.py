"""CPU: the stock-PyTorch restatement timed by tools/bench_torch_gpu.py (BASELINE.md §2 "B2") computes
the reference step — checked against the oracle (itself pinned to the unmodified reference)."""
import importlib.util
import os

import torch

from oracle import gabril_oracle as O
from tests.synth import make_batch, make_weights

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_b2_model_matches_oracle():
    spec = importlib.util.spec_from_file_location("bench_torch_gpu", os.path.join(ROOT, "tools", "bench_torch_gpu.py"))
    T = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(T)
    w = make_weights(2, seed=5)
    model = T.Policy(2)
    enc = {k.replace("_residual_stack._layers", "_res"): v for k, v in w["encoder"].items()}
    model.encoder.load_state_dict(enc)
    model.pre_actor.load_state_dict(w["pre_actor"])
    model.actor.load_state_dict(w["actor"])
    img, gaze, act = make_batch(2, seed=3)
    obs, heat = T.format_obs(img, True), T.heat_maps(gaze, 180, 320)
    total, actor, reg = model.losses(obs, heat, act[:, -1])
    r_total, r_actor, r_reg, ex = O.bc_losses(w["encoder"], w["pre_actor"], w["actor"], img, gaze, act)
    assert torch.equal(obs, ex["obs"]) and (heat - ex["heat"]).abs().max() <= 1e-6
    for a, b in ((total, r_total), (actor, r_actor), (reg, r_reg)):
        assert abs(a.item() - b.item()) <= 1e-5 * abs(b.item())
