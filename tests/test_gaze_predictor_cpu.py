"""CPU: the oracle's gaze-predictor restatement (oracle.decoder_forward / gaze_predictor_loss) against
golden vectors produced by the unmodified reference AutoEncoder (tests/golden/make_golden_gp.py), and
the host-side parameter plumbing of GazePredictorSpec (reference state_dict round trip)."""
import os

import numpy as np
import pytest
import torch

from oracle import gabril_oracle as O
from tests.synth import make_batch, make_weights_gp

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors_gp.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.mark.parametrize("gray", [True, False], ids=["gray", "rgb"])
def test_oracle_matches_reference_autoencoder(gold, gray):
    tag = "gray" if gray else "rgb"
    img, gaze, _ = make_batch(2, seed=31)
    assert np.array_equal(gaze.numpy(), gold["in/gaze"])
    w = make_weights_gp(2 if gray else 6, seed=9)
    sds = {m: {k: v.clone().requires_grad_(True) for k, v in sd.items()} for m, sd in w.items()}
    loss, pred, target = O.gaze_predictor_loss(sds["encoder"], sds["decoder"], img, gaze, grayscale=gray)
    assert tuple(target.shape) == tuple(gold[f"{tag}/target_shape"])            # [B,1,H,W]
    assert np.abs(target[:, 0, 60:120, ::2].numpy() - gold[f"{tag}/target_band"]).max() <= 1e-6
    assert abs(loss.item() - float(gold[f"{tag}/loss"])) <= 1e-5 * abs(float(gold[f"{tag}/loss"]))
    ref_pred = gold[f"{tag}/pred_sub"]
    assert np.abs(pred.detach()[:, 0, ::6, ::8].numpy() - ref_pred).max() <= 1e-4 * np.abs(ref_pred).max()
    loss.backward()
    for mod in ("encoder", "decoder"):
        for k, p in sds[mod].items():
            gn = float(gold[f"{tag}/grad_norm/{mod}.{k}"])
            assert abs(p.grad.norm().item() - gn) <= 1e-4 * gn + 1e-12, (mod, k)
            head = gold[f"{tag}/grad_head/{mod}.{k}"]
            assert np.abs(p.grad.flatten()[:16].numpy() - head).max() <= 1e-4 * (np.abs(head).max() + gn * 1e-3), (mod, k)


def test_gaze_predictor_spec_round_trip():
    import gabril_carla_b200  # noqa: F401
    from gabril_carla_b200.gaze_predictor import autoencoder_state_dict, load_autoencoder_state_dict
    from gabril_carla_b200.model import GazePredictorSpec, ParamArena
    spec = GazePredictorSpec(2)
    arena = ParamArena(spec, "cpu")
    w = make_weights_gp(2, seed=9)
    arena.load_reference(w)
    sd = autoencoder_state_dict(arena)
    # AutoEncoder.state_dict() keys / shapes, fp32
    assert set(sd) == {f"{m}.{k}" for m in w for k in w[m]}
    for m in w:
        for k, v in w[m].items():
            assert sd[f"{m}.{k}"].dtype == torch.float32 and torch.equal(sd[f"{m}.{k}"], v), (m, k)
    other = ParamArena(spec, "cpu")
    load_autoencoder_state_dict(other, {f"module.{k}": v for k, v in sd.items()})  # DDP-prefixed keys
    assert torch.equal(other.p, arena.p)
    # ConvTranspose2d(32, 1, 4, 2, 1) weight as the patch-GEMM matrix [32][1][32] (taps in columns 0..15)
    wp = arena.P("dec.ct4.w")
    assert wp.shape == (32, 1, 32) and torch.equal(wp[:, 0, :16], w["decoder"]["_conv_trans_4.weight"].reshape(32, 16))
    assert (wp[:, 0, 16:] == 0).all()
