"""CPU (-m "not gpu"): pins oracle/gabril_oracle.py against the golden vectors produced by the
unmodified reference (tests/golden/reference_vectors.npz, generator tests/golden/make_golden.py)
and against the known-answer facts of SURVEY.md §8c.  Integer / byte work is compared bit-exactly,
fp32 work within the tolerance written at each assert."""
import hashlib
import math
import os

import numpy as np
import pytest
import torch

from oracle import gabril_oracle as O
from tests.synth import make_batch, make_weights

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD, allow_pickle=False)


def sha(t):
    return hashlib.sha256(t.contiguous().numpy().tobytes()).hexdigest()


def T(a):
    return torch.from_numpy(np.asarray(a))


@pytest.fixture(scope="module")
def batch(gold):
    img, gaze, act = make_batch(4, seed=11)
    # the seeded generator must reproduce the inputs the golden vectors were made from
    assert sha(img) == str(gold["in/img_sha"])
    assert torch.equal(gaze, T(gold["in/gaze"])) and torch.equal(act, T(gold["in/act"]))
    return img, gaze, act


# ---- image formatting: bit-exact (u8 -> fp32 /255 -> gray) --------------------------------------
@pytest.mark.parametrize("gray", [True, False])
def test_format_obs_bit_exact(gold, batch, gray):
    img, gaze, _ = batch
    obs, heat, center = O.prepare_for_bc(img, gaze, 2, gray)
    tag = "gray" if gray else "rgb"
    assert center == int(gold[f"prep/{tag}/center"]) == 1
    assert sha(obs) == str(gold[f"prep/{tag}/obs_sha"])
    assert torch.equal(obs[:, :, ::9, ::8], T(gold[f"prep/{tag}/obs_sub"]))


@pytest.mark.parametrize("L,seed", [(5, 12), (1, 13)])
def test_stack_window_gather(gold, L, seed):
    """L != S: integer window indices clamp(i,0,L-1), front padding (data_loader_robomimic.py:144-157)."""
    img, gaze, _ = make_batch(2, L=L, seed=seed, edge_cases=False)
    assert sha(img) == str(gold[f"prep/L{L}/img_sha"])
    obs, heat, center = O.prepare_for_bc(img, gaze, 2, True)
    assert sha(obs) == str(gold[f"prep/L{L}/obs_sha"])
    assert (heat[:, :, 60:120] - T(gold[f"prep/L{L}/heat_band"])).abs().max() <= 1e-6
    assert O.stack_indices(5, 4, 2) == [3, 4] and O.stack_indices(1, 0, 2) == [0, 0]
    assert O.stack_indices(3, 0, 4) == [0, 0, 0, 0] and O.stack_indices(6, 2, 4) == [0, 0, 1, 2]


# ---- gaze heat maps ------------------------------------------------------------------------------
def test_heatmaps_match_reference(gold, batch):
    img, gaze, _ = batch
    _, heat, _ = O.prepare_for_bc(img, gaze, 2, True)
    ref_band = T(gold["prep/heat_band"])
    assert (heat[:, :, 60:120] - ref_band).abs().max() <= 1e-6          # fp32, summation order only
    assert heat[:, :, :60].abs().max() == 0 and heat[:, :, 120:].abs().max() == 0   # F6 row band
    assert float(gold["prep/heat_outside_absmax"]) == 0.0
    _, heat_na, _ = O.prepare_for_bc(img, gaze, 2, True, aggregate=False)
    assert (heat_na[:, :, 60:120] - T(gold["prep/heat_noagg_band"])).abs().max() <= 1e-6
    base = O.frame_heatmaps(gaze.view(4, 2, 5, 2), 180, 320, 30.0)
    assert (base.reshape(4, 2, 180, 320)[:, :, 60:120] - T(gold["prep/base_band"])).abs().max() <= 1e-6
    # empty frame (sample 0, frame 0) -> all zeros; its aggregate channel 0 too
    assert heat[0, 0].abs().max() == 0


def test_heatmaps_sigma20(gold, batch):
    img, gaze, _ = batch
    _, heat, _ = O.prepare_for_bc(img, gaze, 2, True, sigma=20.0)
    assert (heat[:, :, 40:140, ::2] - T(gold["prep/sigma20/heat_band"])).abs().max() <= 1e-6
    assert heat[:, :, :40].abs().max() == 0 and heat[:, :, 140:].abs().max() == 0


def test_known_answers_kernel_and_peak(gold, batch):
    k = O.gaussian_kernel_1d(30.0)
    assert k.numel() == 121 and abs(k.sum().item() - 1) < 1e-6
    assert abs(k[60].item() - 0.0139060570) < 1e-9 and abs(k[0].item() - 0.00188198) < 1e-8
    assert torch.equal(k, T(gold["prep/kernel_1d"]))                     # same fp32 op order
    assert O.gaussian_kernel_1d(20.0).numel() == 81
    img, gaze, _ = batch
    base = O.frame_heatmaps(gaze.view(4, 2, 5, 2), 180, 320, 30.0)
    pk = base[3, 0].max().item()              # eps on raw amplitudes (F8): peak = 1 - O(1e-8 / k[R]^2)
    assert 1 - 2e-4 < pk < 1.0


def test_pixel_indices_bit_exact():
    c = torch.tensor([[[[0.0, 0.0], [1.0, 1.0], [-1.0, 0.4], [0.4, -1.0], [0.5, 0.5], [0.999999, 0.999999],
                        [2.0, 3.0], [1e-9, 1e-9]]]])
    valid, xi, yi = O.pixel_indices(c, 180, 320)
    assert valid.flatten().tolist() == [True, True, False, False, True, True, True, True]
    assert xi.flatten().tolist()[:2] == [0, 319] and yi.flatten().tolist()[:2] == [0, 179]
    assert xi.flatten().tolist()[4:] == [159, 318, 319, 0] and yi.flatten().tolist()[4:] == [89, 178, 179, 0]
    g = torch.Generator().manual_seed(3)
    r = torch.rand(1, 1, 4096, 2, generator=g)
    _, xi, yi = O.pixel_indices(r, 180, 320)
    ex = (r[..., 0].clamp(0, 1) * 319).long().clamp(0, 319)
    ey = (r[..., 1].clamp(0, 1) * 179).long().clamp(0, 179)
    assert torch.equal(xi, ex) and torch.equal(yi, ey)


def test_closed_form_splat_equals_blur(batch):
    """The closed form the CUDA kernel evaluates vs the conv formulation."""
    _, gaze, _ = batch
    coords = gaze.view(4, 2, 5, 2)
    raw = O.banded_splat_closed_form(coords, 180, 320, 30.0)
    mn = raw.amin(dim=(-1, -2), keepdim=True)
    mx = raw.amax(dim=(-1, -2), keepdim=True)
    norm = (raw - mn) / (mx - mn + 1e-8)
    base = O.frame_heatmaps(coords, 180, 320, 30.0).reshape(4, 2, 180, 320)
    assert (norm - base).abs().max() <= 5e-7


# ---- get_gaze_mask / GMD -------------------------------------------------------------------------
def test_gaze_mask_and_grad(gold, batch):
    img, gaze, _ = batch
    z = T(gold["mask/z"]).clone().requires_grad_(True)
    m = O.gaze_mask(z, 50.0, (180, 320))
    assert (m.detach()[:, 0, ::3, ::4] - T(gold["mask/out_sub"])).abs().max() <= 2e-6
    assert abs(m.detach().double().sum().item() - float(gold["mask/out_sum"])) <= 1e-6 * float(gold["mask/out_sum"])
    _, heat, _ = O.prepare_for_bc(img, gaze, 2, True)
    loss = torch.nn.functional.mse_loss(heat[:, -1:], m)
    loss.backward()
    assert abs(loss.item() - float(gold["mask/mse"])) <= 1e-6 * float(gold["mask/mse"])
    ref = T(gold["mask/dz"])
    assert ((z.grad - ref).norm() / ref.norm()).item() <= 1e-4


def test_bicubic_taps_match_interpolate():
    for n_in, n_out in ((20, 180), (38, 320), (180, 20), (320, 38)):
        idx, w = O.bicubic_taps(n_in, n_out)
        x = torch.randn(1, 1, 1, n_in, generator=torch.Generator().manual_seed(n_in))
        ref = torch.nn.functional.interpolate(x, size=(1, n_out), mode="bicubic")[0, 0, 0]
        got = (x[0, 0, 0][idx] * w).sum(-1)
        assert (got - ref).abs().max() <= 1e-5      # randn-scale inputs; 1.2e-6 at softmax scale


def test_gmd_exact_mask(gold, batch):
    img, gaze, _ = batch
    _, heat, _ = O.prepare_for_bc(img, gaze, 2, True)
    z = T(gold["mask/z"])
    A = T(gold["gmd/A"])
    tr = O.gmd_dropout(z, heat, A, test_mode=False)
    te = O.gmd_dropout(z, heat, A, test_mode=True)
    # 0/1 mask: the compare A < K must agree except where |A-K| is below fp32 noise (none here)
    assert torch.equal(tr[:, :4] != 0, T(gold["gmd/train"]) != 0)
    assert (tr[:, :4] - T(gold["gmd/train"])).abs().max() == 0
    assert (te[:, :4] - T(gold["gmd/test"])).abs().max() <= 1e-5
    K = O.gmd_keep_prob(heat, (20, 38))
    assert 0.3 - 1e-6 <= K.min().item() and K.max().item() <= 1.0 + 1e-6


# ---- whole step: losses and gradient summaries ---------------------------------------------------
@pytest.mark.parametrize("gray,gm,dp", [(True, "None", "None"), (True, "Reg", "None"),
                                        (True, "Reg", "GMD"), (False, "Reg", "GMD")])
def test_step_losses_and_grads(gold, batch, gray, gm, dp):
    img, gaze, act = batch
    w = make_weights(2 if gray else 6, seed=7)
    w = {m: {k: v.clone().requires_grad_(True) for k, v in sd.items()} for m, sd in w.items()}
    total, actor, reg, ex = O.bc_losses(w["encoder"], w["pre_actor"], w["actor"], img, gaze, act,
                                        gaze_method=gm, dp_method=dp, grayscale=gray, gmd_A=T(gold["gmd/A"]))
    total.backward()
    tag = f"step/{'gray' if gray else 'rgb'}_{gm}_{dp}"
    for name, got in (("total", total), ("actor", actor), ("reg", reg)):
        ref = float(gold[f"{tag}/{name}"])
        assert abs(got.item() - ref) <= 1e-5 * max(abs(ref), 1e-3), (name, got.item(), ref)
    for mod, pref in (("encoder", "encoder"), ("pre_actor", "pre_actor"), ("actor", "actor_head")):
        for k, v in w[mod].items():
            rn = float(gold[f"{tag}/{pref}/{k}/norm"])
            gn = v.grad.double().norm().item()
            assert abs(gn - rn) <= 1e-4 * rn + 1e-9, (mod, k, gn, rn)
            head = T(gold[f"{tag}/{pref}/{k}/head"])
            got = v.grad.flatten()[:256]
            assert (got - head).norm() <= 2e-4 * head.norm() + 1e-9, (mod, k)


def test_residual_inplace_relu_quirk():
    """Residual(x) = relu(x) + W2 relu(W1 * relu(x))."""
    w = make_weights(2, seed=7)["encoder"]
    x = torch.randn(1, 2, 180, 320, generator=torch.Generator().manual_seed(1))
    z = O.encoder_forward(w, x)
    assert tuple(z.shape) == (1, 64, 20, 38) and O.encoder_out_hw(180, 320) == (20, 38)
    assert sum(v.numel() for v in w.values()) == 1_009_184


# ---- optimiser / schedule ------------------------------------------------------------------------
def test_adam_cosine_warmup_trajectory(gold):
    lr_ref = gold["sched/lr"]
    base_lr, T_total = 5e-4, 1200
    w = T(gold["sched/w0"]).clone()
    m, v = torch.zeros_like(w), torch.zeros_like(w)
    g2 = torch.Generator().manual_seed(31)
    torch.randn(64, generator=g2)
    grads = torch.randn(1200, 64, generator=g2) * 0.1
    assert sha(grads) == str(gold["sched/grads_sha"])
    assert lr_ref[0] == 0.0                                      # first optimizer step has lr 0 (A11)
    for i in range(1200):
        lr = base_lr * O.cosine_warmup_factor(i, 500, T_total, 1e-6, base_lr)
        assert math.isclose(lr, lr_ref[i], rel_tol=1e-12, abs_tol=1e-18)
        O.adam_update(w, grads[i], m, v, i + 1, lr)
        if i + 1 in (1, 2, 10, 600, 1200):
            ref = T(gold[f"sched/w_after_{i + 1}"])
            assert (w - ref).abs().max() <= 1e-6, i


# ---- live cross-check against the imported reference (only where /root/reference exists) --------
def test_oracle_vs_live_reference():
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("/root/reference not mounted (GPU box); golden vectors cover this")
    ref = ref_import.load()
    img, gaze, _ = make_batch(3, seed=99, edge_cases=False)
    gp = ref.GazePreprocessor(img_height=180, img_width=320, gaze_sigma=30.0, gaze_coeff=0.8,
                              maxpoints=5, device="cpu", temporal_alpha=0.7)
    obs_r, heat_r, _ = gp.prepare_for_bc(img, gaze, frame_stack=2, grayscale=True)
    obs_o, heat_o, _ = O.prepare_for_bc(img, gaze, 2, True)
    assert torch.equal(obs_r, obs_o)
    assert (heat_r - heat_o).abs().max() <= 1e-6
    enc = ref.Encoder(2, 64, 128, 2, 32)
    x = obs_r
    assert (enc(x) - O.encoder_forward(dict(enc.state_dict()), x)).abs().max() <= 1e-4
