"""-m gpu parity tests of the memory-bound kernels, through the C ABI (ctypes), against the CPU
oracle (oracle/gabril_oracle.py) on identical seeded inputs.  Integer / byte work is bit-exact;
fp32 tolerances are written next to each assert."""
import pytest
import torch
import torch.nn.functional as F

from oracle import gabril_oracle as O
from tests.synth import make_batch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def G():
    import gabril_carla_b200  # noqa: F401
    from gabril_carla_b200 import _lib, gaze, ops
    _lib.init()

    class NS:
        pass
    ns = NS()
    ns.ops, ns.gaze, ns.lib = ops, gaze, _lib
    return ns


@pytest.mark.parametrize("gray", [True, False])
@pytest.mark.parametrize("L,S", [(2, 2), (5, 2), (1, 2), (3, 4)])
def test_obs_format_bit_exact(G, gray, L, S):
    img, _, _ = make_batch(3, L=L, seed=1, smooth=False)
    got = G.ops.obs_format(img.to(DEV), S, L - 1, gray).cpu()
    ref = O.format_obs(O.gather_stack(img, L - 1, S), gray)
    assert got.shape == ref.shape
    assert torch.equal(got, ref)  # bit-exact fp32 (u8/255, then rounded mul/add chain)


@pytest.mark.parametrize("gray,S", [(True, 2), (False, 2), (True, 1), (True, 4), (False, 1)])
def test_obs_patches_exact(G, gray, S):
    img, _, _ = make_batch(2, L=3, seed=2, smooth=False)
    got = G.ops.obs_patches(img.to(DEV), S, 2, gray).float().cpu()
    obs = O.format_obs(O.gather_stack(img, 2, S), gray)  # [B,C0,H,W]
    B, C0, H, W = obs.shape
    cols = F.unfold(obs, kernel_size=4, stride=2, padding=1)  # [B, C0*16, OH*OW], (c, r, s) major
    OHW = cols.shape[-1]
    cols = cols.reshape(B, C0, 16, OHW).permute(0, 3, 2, 1).reshape(B * OHW, 16 * C0)
    ref = cols.to(torch.bfloat16).float()
    assert torch.equal(got[:, :16 * C0], ref)
    assert torch.count_nonzero(got[:, 16 * C0:]) == 0


def _pre(G, **kw):
    return G.gaze.GazePreprocessor(device=DEV, **kw)


def test_gaze_splat_matches_oracle(G):
    _, gaze, _ = make_batch(16, L=2, seed=3)
    pre = _pre(G)
    got = pre.build_stack_heatmaps(gaze.to(DEV), 2, 1).cpu()
    ref = O.stack_heatmaps(gaze, 2, 1, 180, 320, 30.0)
    assert got.shape == ref.shape == (16, 2, 180, 320)
    # closed form vs the reference's two-pass blur: summation order only (1.8e-7)
    assert (got - ref).abs().max().item() <= 1e-6
    # row-band quirk (F6): rows outside [60, 120) are exactly zero
    assert torch.count_nonzero(got[:, :, :60]) == 0 and torch.count_nonzero(got[:, :, 120:]) == 0
    # empty frame -> all-zero map
    assert torch.count_nonzero(got[0, 0]) == 0
    # eps effect on a single-point per-frame map (F8): peak = 1 - 5.2e-5, not 1
    one = torch.full((1, 1, 10), -1.0); one[0, 0, :2] = torch.tensor([0.5, 0.5])
    pk = pre.forward(one.to(DEV)).max().item()
    assert 0.9999 < pk < 1.0 and abs(pk - O.frame_heatmaps(one, 180, 320, 30.0).max().item()) <= 1e-6


def test_gaze_splat_peak_pixels_bit_exact(G):
    """Integer pixel indexing: with one valid point per frame, the arg-max of the map is exactly
    the (clamped, truncated) pixel of the reference (row clipped into the band)."""
    g = torch.Generator().manual_seed(5)
    B = 64
    pts = torch.full((B, 1, 5, 2), -1.0)
    pts[:, 0, 0] = torch.rand(B, 2, generator=g)
    pts[0, 0, 0] = torch.tensor([1.0, 1.0]); pts[1, 0, 0] = torch.tensor([0.0, 0.0])
    pts[2, 0, 0] = torch.tensor([0.999999, 0.5]); pts[3, 0, 0] = torch.tensor([1.5, 0.25])
    pre = _pre(G)
    maps = pre.forward(pts.reshape(B, 1, 10).to(DEV)).cpu()[:, 0, 0]
    valid, xi, yi = O.pixel_indices(pts, 180, 320)
    am = maps.reshape(B, -1).argmax(1)
    yy, xx = am // 320, am % 320
    assert torch.equal(xx, xi[:, 0, 0])
    assert torch.equal(yy, yi[:, 0, 0].clamp(60, 119))


def test_gaze_splat_per_frame_and_multiscale(G):
    _, gaze, _ = make_batch(8, L=4, seed=4)
    pre = _pre(G)
    got = pre.forward(gaze.to(DEV)).cpu()
    ref = O.frame_heatmaps(gaze, 180, 320, 30.0)
    assert got.shape == ref.shape == (8, 4, 1, 180, 320)
    assert (got - ref).abs().max().item() <= 1e-6
    # stack 4 alpha-decay, sequence longer than the stack, sigma 20 (train_bc_confounded.yaml)
    pre20 = _pre(G, gaze_sigma=20.0)
    got = pre20.build_stack_heatmaps(gaze.to(DEV), 3, 3).cpu()
    ref = O.stack_heatmaps(gaze, 3, 3, 180, 320, 20.0)
    assert (got - ref).abs().max().item() <= 2e-6
    # multiscale mode
    prem = _pre(G, temporal_mode="multiscale", temporal_sigmas=[30.0, 24.0, 18.0],
                temporal_coeffs=[1.0, 0.8, 0.6])
    got = prem.build_stack_heatmaps(gaze.to(DEV), 3, 3).cpu()
    ref = O.stack_heatmaps(gaze, 3, 3, 180, 320, 30.0, mode="multiscale", sigmas=[30.0, 24.0, 18.0],
                           coeffs=[1.0, 0.8, 0.6])
    assert (got - ref).abs().max().item() <= 2e-6


def test_prepare_for_bc_api(G):
    img, gaze, _ = make_batch(4, L=2, seed=6)
    pre = _pre(G)
    obs, heat, c = pre.prepare_for_bc(img.to(DEV), gaze.to(DEV), 2, grayscale=True)
    robs, rheat, rc = O.prepare_for_bc(img, gaze, 2, True, 180, 320, 30.0)
    assert c == rc == 1 and obs.dtype == torch.float32 and heat.dtype == torch.float32
    assert torch.equal(obs.cpu(), robs)
    assert (heat.cpu() - rheat).abs().max().item() <= 1e-6
    obs2, last, _ = pre.prepare_for_gaze_predictor(img.to(DEV), gaze.to(DEV), 2, grayscale=False)
    assert obs2.shape == (4, 6, 180, 320) and last.shape == (4, 1, 180, 320)
    assert (last.cpu() - rheat[:, -1:]).abs().max().item() <= 1e-6


def test_gmd_mask(G):
    _, gaze, _ = make_batch(8, seed=7)
    heat = O.stack_heatmaps(gaze, 2, 1, 180, 320, 30.0)
    g = torch.Generator().manual_seed(0)
    A = torch.rand(8, 1, 20, 38, generator=g)
    keep = torch.empty(8, 760, device=DEV)
    mask = G.ops.gmd_mask(heat.to(DEV), A.reshape(8, 760).to(DEV), 20, 38, 0.7, False, keep_out=keep).cpu()
    K = O.gmd_keep_prob(heat, (20, 38)).reshape(8, 760)
    assert (keep.cpu() - K).abs().max().item() <= 2e-6  # fp32 bicubic summation order
    ref_mask = (A.reshape(8, 760) < K).float()
    decided = (A.reshape(8, 760) - K).abs() > 1e-5  # away from the comparison boundary: exact
    assert torch.equal(mask[decided], ref_mask[decided])
    assert set(mask.unique().tolist()) <= {0.0, 1.0}
    # drop-in function, train and test mode
    z = torch.randn(8, 64, 20, 38, generator=g)
    out = G.gaze.apply_gmd_dropout(z.to(DEV), heat.to(DEV), False, A=A.to(DEV)).cpu()
    ref = O.gmd_dropout(z, heat, A)
    same = decided.reshape(8, 1, 20, 38).expand_as(z)
    assert torch.equal(out[same], ref[same])
    out_t = G.gaze.apply_gmd_dropout(z.to(DEV), heat.to(DEV), True, A=A.to(DEV)).cpu()
    assert (out_t - O.gmd_dropout(z, heat, A, test_mode=True)).abs().max().item() <= 1e-5


def test_gmd_in_kernel_philox_statistics(G):
    """A ~ U[0,1) drawn inside the mask kernel (Philox4x32-10 keyed by a device-resident (seed, step)): the
    mask is a Bernoulli(K~) field — empirical keep rate within 4 sigma of mean(K~) overall and per K~ bin —
    reproducible for a given state, different after gabril_rng_advance and for another draw id or seed."""
    B = 64
    _, gaze, _ = make_batch(B, seed=9)
    heat = O.stack_heatmaps(gaze, 2, 1, 180, 320, 30.0).to(DEV)
    keep = torch.empty(B, 760, device=DEV)
    st = torch.tensor([1234, 0], dtype=torch.int64, device=DEV)
    m0 = G.ops.gmd_mask(heat, None, 20, 38, 0.7, False, keep_out=keep, rng_state=st, draw_id=0).clone()
    m0b = G.ops.gmd_mask(heat, None, 20, 38, 0.7, False, rng_state=st, draw_id=0).clone()
    assert torch.equal(m0, m0b) and set(m0.unique().tolist()) <= {0.0, 1.0}
    m_other_draw = G.ops.gmd_mask(heat, None, 20, 38, 0.7, False, rng_state=st, draw_id=1).clone()
    G.ops.rng_advance(st)
    assert st.tolist() == [1234, 1]
    m1 = G.ops.gmd_mask(heat, None, 20, 38, 0.7, False, rng_state=st, draw_id=0).clone()
    st2 = torch.tensor([99, 0], dtype=torch.int64, device=DEV)
    m_seed = G.ops.gmd_mask(heat, None, 20, 38, 0.7, False, rng_state=st2, draw_id=0).clone()
    K = keep.cpu().double().flatten()
    n = K.numel()
    for m in (m0, m1, m_other_draw, m_seed):
        x = m.cpu().double().flatten()
        sigma = (K * (1 - K)).sum().sqrt() / n
        assert abs(x.mean() - K.mean()) <= 4 * sigma, (x.mean().item(), K.mean().item(), sigma.item())
        for lo, hi in ((0.3, 0.35), (0.35, 0.6), (0.6, 1.01)):        # keep rate follows K~ bin by bin
            sel = (K >= lo) & (K < hi)
            if sel.sum() > 2000:
                s_bin = (K[sel] * (1 - K[sel])).sum().sqrt() / sel.sum()
                assert abs(x[sel].mean() - K[sel].mean()) <= 4.5 * s_bin
    # independent streams: agreement rate between two masks = E[K^2 + (1-K)^2], not 1
    expect = (K * K + (1 - K) * (1 - K)).mean().item()
    for other in (m1, m_other_draw, m_seed):
        agree = (m0 == other).double().mean().item()
        assert abs(agree - expect) <= 0.02, (agree, expect)
    # neighbouring elements are uncorrelated (lag-1 autocorrelation of the uniform-ised residual)
    r = (m0.cpu().double().flatten() - K)
    ac = (r[1:] * r[:-1]).mean() / (r * r).mean()
    assert abs(ac.item()) <= 0.02


def test_get_gaze_mask_fwd_bwd(G):
    g = torch.Generator().manual_seed(8)
    z = torch.randn(6, 64, 20, 38, generator=g) * 3.0
    w = torch.rand(6, 1, 180, 320, generator=g)
    zr = z.clone().requires_grad_(True)
    ref = O.gaze_mask(zr, 50.0, (180, 320))
    (ref * w).sum().backward()
    zd = z.to(DEV).requires_grad_(True)
    got = G.gaze.get_gaze_mask(zd, 50.0, (180, 320))
    (got * w.to(DEV)).sum().backward()
    assert got.shape == (6, 1, 180, 320)
    assert (got.detach().cpu() - ref.detach()).abs().max().item() <= 5e-6
    gerr = (zd.grad.cpu() - zr.grad).abs().max().item() / zr.grad.abs().max().item()
    assert gerr <= 1e-4, gerr  # fp32 adjoint of the same chain, relative to the largest entry


def test_reg_mse_fused(G):
    g = torch.Generator().manual_seed(9)
    B = 5
    z = (torch.randn(B, 64, 20, 38, generator=g) * 2.0).to(torch.bfloat16).float()
    _, gaze, _ = make_batch(B, seed=10)
    heat = O.stack_heatmaps(gaze, 2, 1, 180, 320, 30.0)
    zr = z.clone().requires_grad_(True)
    loss = O.reg_loss(zr, heat, torch.ones(B), 50.0, (180, 320))
    (10.0 * loss).backward()
    z_nhwc = z.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).to(DEV)
    heat_d = heat.to(DEV)
    dz = torch.empty_like(z_nhwc)
    part = torch.empty(B, device=DEV)
    scale = 1.0 / (B * 180 * 320)
    G.ops.reg_mse(z_nhwc, heat_d[:, 1], heat_d.stride(0), 180, 320, 50.0, 10.0 * scale, dz, part)
    got_loss = part.sum().item() * scale
    assert abs(got_loss - loss.item()) <= 1e-4 * abs(loss.item())  # loss tolerance (north_star: 1e-3)
    dref = zr.grad.permute(0, 2, 3, 1)
    derr = (dz.float().cpu() - dref).norm().item() / dref.norm().item()
    assert derr <= 5e-3, derr  # bf16 output rounding (2^-9 per element)


def test_heads_and_adam(G):
    g = torch.Generator().manual_seed(11)
    B, A = 37, 7
    raw = torch.randn(B, 32, generator=g)
    bias = torch.randn(32, generator=g)
    act = torch.rand(B, 2, A, generator=g)
    part = torch.empty(B, device=DEV); dlog = torch.empty(B, 32, device=DEV, dtype=torch.bfloat16)
    db = torch.zeros(32, device=DEV); logits = torch.empty(B, A, device=DEV)
    actd = act.to(DEV)
    G.ops.mse_head(raw.to(DEV), bias.to(DEV), actd[:, 1, :], A, part, 1.0 / (B * A), dlogits=dlog,
                   dbias=db, logits_out=logits)
    l = (raw[:, :A] + bias[:A]).requires_grad_(True)
    loss = F.mse_loss(l, act[:, 1]); loss.backward()
    assert abs(part.sum().item() / (B * A) - loss.item()) <= 1e-6 * max(1.0, loss.item())
    assert (logits.cpu() - l.detach()).abs().max().item() <= 1e-6
    assert (dlog[:, :A].float().cpu() - l.grad).abs().max().item() <= 2 ** -8 * l.grad.abs().max().item()
    assert torch.count_nonzero(dlog[:, A:]) == 0
    assert (db[:A].cpu() - l.grad.sum(0)).abs().max().item() <= 1e-6
    # colsum
    dy = torch.randn(1000, 64, generator=g).to(torch.bfloat16)
    out = torch.zeros(64, device=DEV)
    G.ops.colsum_bf16(dy.to(DEV), out)
    assert (out.cpu() - dy.float().sum(0)).abs().max().item() <= 1e-3
    # Adam vs torch.optim.Adam, three steps
    n = 4096
    p0 = torch.randn(n, generator=g); p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([p], lr=5e-4)
    pd, m, v = p0.to(DEV), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    sh = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    for step in range(1, 4):
        gr = torch.randn(n, generator=g)
        p.grad = gr.clone(); opt.step()
        G.ops.adam(pd, gr.to(DEV), m, v, sh, step, 5e-4)
    assert (pd.cpu() - p.detach()).abs().max().item() <= 1e-6
    assert torch.equal(sh.cpu(), pd.cpu().to(torch.bfloat16))
