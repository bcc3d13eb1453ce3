"""-m gpu: every dense contraction of the encoder / heads (fwd, dgrad, wgrad) through the tcgen05
path (C ABI via ctypes) against fp64 CPU convolution of the same bf16-rounded operands."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

# (name, H, W, Cin, Cout, R, S, stride, pad)
LAYERS = [
    ("conv1_gemm_gray", 1, 1, 32, 32, 1, 1, 1, 0),
    ("conv1_gemm_rgb", 1, 1, 128, 32, 1, 1, 1, 0),
    ("conv2", 90, 160, 32, 64, 4, 4, 2, 1),
    ("conv3", 45, 80, 64, 128, 4, 4, 2, 1),
    ("conv3_p0", 47, 82, 64, 128, 4, 4, 2, 0),   # the engine's form (zero-bordered input): parity classes stacked along N
    ("conv4", 22, 40, 128, 128, 3, 3, 1, 0),
    ("res3x3", 20, 38, 128, 32, 3, 3, 1, 1),
    ("res1x1", 20, 38, 32, 128, 1, 1, 1, 0),
    ("conv5", 20, 38, 128, 128, 5, 5, 1, 2),
    ("conv6", 20, 38, 128, 64, 5, 5, 1, 2),
    ("actor0", 1, 1, 256, 256, 1, 1, 1, 0),
    ("actor2", 1, 1, 256, 32, 1, 1, 1, 0),
]


@pytest.fixture(scope="module")
def ops():
    import gabril_carla_b200  # noqa: F401
    from gabril_carla_b200 import _lib, ops
    _lib.init()
    return ops


def rel(a, b):
    return ((a.double() - b).norm() / (b.norm() + 1e-30)).item()


def pack_signs(t):
    """Host restatement of the sign words (include/gabril_b200.h): word k covers elements 32k .. 32k+31,
    bit j (j < 16) = (element 2j > 0), bit 16 + j = (element 2j + 1 > 0) -- the low / high halves of the
    sixteen packed bf16x2 words the epilogue holds."""
    pos = (t.float().reshape(-1, 32) > 0).to(torch.int64)
    e = torch.arange(32, dtype=torch.int64)
    words = (pos << ((e % 2) * 16 + e // 2)).sum(1)
    return torch.where(words >= 2 ** 31, words - 2 ** 32, words).to(torch.int32)


@pytest.mark.parametrize("N", [3, 40])
@pytest.mark.parametrize("layer", LAYERS, ids=[l[0] for l in LAYERS])
def test_conv_fwd_dgrad_wgrad(ops, layer, N):
    name, H, W, Cin, Cout, R, S, stride, pad = layer
    if H == 1:
        N = N * 700 + 13  # linear layers: rows
    from gabril_carla_b200 import _lib
    g = torch.Generator().manual_seed(len(name) * 7 + N)
    OH, OW = ops.conv_out_hw(H, W, R, S, stride, pad)
    x = (torch.randn(N, H, W, Cin, generator=g) * 0.5).to(torch.bfloat16)
    w = (torch.randn(Cout, Cin, R, S, generator=g) / (Cin * R * S) ** 0.5).to(torch.bfloat16)
    bias = torch.randn(Cout, generator=g)
    resid = torch.randn(N, OH, OW, Cout, generator=g).to(torch.bfloat16)
    rowmask = (torch.rand(N * OH * OW, generator=g) > 0.3).float()
    dy = (torch.randn(N, OH, OW, Cout, generator=g) * 0.5).to(torch.bfloat16)
    addend = torch.randn(N, H, W, Cin, generator=g).to(torch.bfloat16)
    gate = torch.randn(N, H, W, Cin, generator=g).to(torch.bfloat16)
    xd, wd = x.to(DEV), ops.w_oihw_to_internal(w).to(DEV)
    x64, w64, dy64 = x.double().permute(0, 3, 1, 2), w.double(), dy.double().permute(0, 3, 1, 2)
    # forward: bias + residual + ReLU + row mask epilogue
    y = torch.full((N, OH, OW, Cout), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.conv_fwd(xd, wd, bias.to(DEV), y, R, S, stride, pad, relu=True, resid=resid.to(DEV),
                 rowmask=rowmask.to(DEV))
    ref = F.relu(F.conv2d(x64, w64, bias.double(), stride=stride, padding=pad).permute(0, 2, 3, 1)
                 + resid.double()) * rowmask.double().view(N, OH, OW, 1)
    assert rel(y.cpu(), ref) <= 4e-3  # bf16 output rounding (2^-9)
    # the same launch with packed ReLU sign words: identical output, words = (stored value > 0)
    y2 = torch.full_like(y, float("nan"))
    sb = torch.full((y.numel() // 32,), -1, device=DEV, dtype=torch.int32)
    ops.conv_fwd(xd, wd, bias.to(DEV), y2, R, S, stride, pad, relu=True, resid=resid.to(DEV),
                 rowmask=rowmask.to(DEV), bits_out=sb)
    assert torch.equal(y2, y) and torch.equal(sb.cpu(), pack_signs(y.cpu()))
    # dgrad: + addend, gated
    dx = torch.full((N, H, W, Cin), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.conv_dgrad(dy.to(DEV), wd, dx, R, S, stride, pad, addend=addend.to(DEV), gate=gate.to(DEV))
    dref = torch.nn.grad.conv2d_input((N, Cin, H, W), w64, dy64, stride=stride, padding=pad).permute(0, 2, 3, 1)
    dref = (dref + addend.double()) * (gate.double() > 0)
    assert rel(dx.cpu(), dref) <= 4e-3
    # gate read as packed sign words instead of the tensor: bit-identical result
    dx2 = torch.full_like(dx, float("nan"))
    ops.conv_dgrad(dy.to(DEV), wd, dx2, R, S, stride, pad, addend=addend.to(DEV), gate_bits=pack_signs(gate).to(DEV))
    assert torch.equal(dx2, dx)
    # wgrad: fp32 accumulate (split over CTAs with fp32 atomics), and direct store for linears
    dw = torch.zeros(Cout, R * S, Cin, device=DEV, dtype=torch.float32)
    ops.conv_wgrad(dy.to(DEV), xd, dw, R, S, stride, pad)
    wref = ops.w_oihw_to_internal(torch.nn.grad.conv2d_weight(x64, (Cout, Cin, R, S), dy64, stride=stride, padding=pad))
    assert rel(dw.cpu(), wref) <= 2e-5
    # the same launch with the bias gradient fused in (gabril_conv_wgrad_dbias_next): db += column sums of dy
    # (fp32 sums of the bf16 values, like gabril_colsum_bf16), dW unchanged up to the atomics' order
    dw2 = torch.zeros_like(dw)
    db = torch.full((Cout,), 0.5, device=DEV)
    ops.conv_wgrad(dy.to(DEV), xd, dw2, R, S, stride, pad, dbias=db)
    assert rel(dw2.cpu(), wref) <= 2e-5
    assert rel(db.cpu() - 0.5, dy.double().sum((0, 1, 2))) <= 2e-5
    _lib.poll_error()


@pytest.mark.parametrize("N", [1, 3, 40])
@pytest.mark.parametrize("layer", [("conv6", 20, 38, 128, 64, 5, 2), ("res3x3", 20, 38, 128, 32, 3, 1)], ids=["conv6", "res3x3"])
def test_conv_fwd_pairs_is_bit_identical(ops, layer, N):
    """gabril_conv_fwd_pairs (two adjacent output pixels stacked along N, pair-packed weights) vs gabril_conv_fwd."""
    from gabril_carla_b200 import _lib
    name, H, W, Cin, Cout, R, pad = layer
    g = torch.Generator().manual_seed(77 + N + Cout)
    x = (torch.randn(N, H, W, Cin, generator=g) * 0.5).to(torch.bfloat16).to(DEV)
    w = ops.w_oihw_to_internal((torch.randn(Cout, Cin, R, R, generator=g) / (Cin * R * R) ** 0.5).to(torch.bfloat16)).to(DEV)
    bias = torch.randn(Cout, generator=g).to(DEV)
    y_ref = torch.full((N, H, W, Cout), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.conv_fwd(x, w, bias, y_ref, R, R, 1, pad, relu=False)
    wp = torch.full((2 * Cout, R * (R + 1), Cin), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.conv_pack_wpairs(w, wp, R, R)
    wpc, wc = wp.cpu().view(2, Cout, R, R + 1, Cin), w.cpu().view(Cout, R, R, Cin)
    assert torch.equal(wpc[0, :, :, :R], wc) and torch.equal(wpc[1, :, :, 1:], wc)
    assert (wpc[0, :, :, R] == 0).all() and (wpc[1, :, :, 0] == 0).all()
    y = torch.full_like(y_ref, float("nan"))
    for _ in range(2):
        ops.conv_fwd_pairs(x, wp, bias, y, R, R, pad, relu=False)
    torch.cuda.synchronize()
    _lib.poll_error()
    assert torch.equal(y, y_ref)
    # with the per-pixel multiplier of the in-encoder gaze dropout
    rowmask = (torch.rand(N * H * W, generator=g) > 0.3).float().to(DEV)
    ops.conv_fwd(x, w, bias, y_ref, R, R, 1, pad, relu=False, rowmask=rowmask)
    y.fill_(float("nan"))
    ops.conv_fwd_pairs(x, wp, bias, y, R, R, pad, relu=False, rowmask=rowmask)
    torch.cuda.synchronize()
    _lib.poll_error()
    assert torch.equal(y, y_ref)


def test_pre_actor_shapes(ops):
    from gabril_carla_b200 import _lib
    g = torch.Generator().manual_seed(7)
    Bn, K, Nf = 96, 48640, 256
    z = (torch.randn(Bn, K, generator=g) * 0.5).to(torch.bfloat16)
    wp = (torch.randn(Nf, K, generator=g) / K ** 0.5).to(torch.bfloat16)
    dzf = torch.randn(Bn, Nf, generator=g).to(torch.bfloat16)
    zd, wd, dd = z.to(DEV), wp.to(DEV), dzf.to(DEV)
    out = torch.zeros(Bn, Nf, device=DEV, dtype=torch.float32)
    ops.conv_fwd(zd.view(Bn, 1, 1, K), wd.view(Nf, 1, K), None, out, 1, 1, 1, 0, out_f32_atomic=True, splits=19)
    assert rel(out.cpu(), z.double() @ wp.double().t()) <= 1e-5
    dz = torch.full((Bn, 1, 1, K), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.conv_dgrad(dd.view(Bn, 1, 1, Nf), wd.view(Nf, 1, K), dz, 1, 1, 1, 0)
    assert rel(dz.view(Bn, K).cpu(), dzf.double() @ wp.double()) <= 4e-3
    dwp = torch.full((Nf, 1, K), float("nan"), device=DEV, dtype=torch.float32)
    ops.conv_wgrad(dd.view(Bn, 1, 1, Nf), zd.view(Bn, 1, 1, K), dwp, 1, 1, 1, 0, direct_store=True)
    assert rel(dwp.view(Nf, K).cpu(), dzf.double().t() @ z.double()) <= 1e-5
    _lib.poll_error()


def test_unsupported_shapes_fail_loudly(ops):
    x = torch.zeros(2, 4, 4, 96, device=DEV, dtype=torch.bfloat16)
    dy = torch.zeros(2, 4, 4, 32, device=DEV, dtype=torch.bfloat16)
    dw = torch.zeros(32, 1, 96, device=DEV)
    with pytest.raises(RuntimeError):
        ops.conv_wgrad(dy, x, dw, 1, 1, 1, 0)  # Cin = 96 is not 32 / 64 / k*128


@pytest.mark.parametrize("N", [1, 3, 40])
def test_res_block_fwd_fused_is_bit_identical(ops, N):
    """gabril_res_block_fwd (one launch) vs the two convolutions it replaces: same bf16 outputs and sign words,
    and both vs fp64 convolutions of the same bf16 operands."""
    from gabril_carla_b200 import _lib
    H, W, C, Ch = 20, 38, 128, 32
    g = torch.Generator().manual_seed(100 + N)
    x = F.relu(torch.randn(N, H, W, C, generator=g) * 0.5).to(torch.bfloat16)
    w3 = (torch.randn(Ch, C, 3, 3, generator=g) / (C * 9) ** 0.5).to(torch.bfloat16)
    w1 = (torch.randn(C, Ch, 1, 1, generator=g) / Ch ** 0.5).to(torch.bfloat16)
    xd, w3d, w1d = x.to(DEV), ops.w_oihw_to_internal(w3).to(DEV), ops.w_oihw_to_internal(w1).to(DEV)
    h_ref = torch.full((N, H, W, Ch), float("nan"), device=DEV, dtype=torch.bfloat16)
    o_ref = torch.full((N, H, W, C), float("nan"), device=DEV, dtype=torch.bfloat16)
    sbh_ref, sbo_ref = ops.sign_bits_like(h_ref), ops.sign_bits_like(o_ref)
    ops.conv_fwd(xd, w3d, None, h_ref, 3, 3, 1, 1, relu=True, bits_out=sbh_ref)
    ops.conv_fwd(h_ref, w1d, None, o_ref, 1, 1, 1, 0, relu=True, resid=xd, bits_out=sbo_ref)
    h = torch.full_like(h_ref, float("nan"))
    o = torch.full_like(o_ref, float("nan"))
    sbh, sbo = torch.full_like(sbh_ref, -1), torch.full_like(sbo_ref, -1)
    for _ in range(2):   # twice: the second launch starts from used barriers / TMEM of nobody (fresh CTAs), same result
        ops.res_block_fwd(xd, w3d, w1d, h, sbh, o, sbo)
    torch.cuda.synchronize()
    _lib.poll_error()
    assert torch.equal(h, h_ref) and torch.equal(sbh, sbh_ref)
    assert torch.equal(o, o_ref) and torch.equal(sbo, sbo_ref)
    h64 = F.relu(F.conv2d(x.double().permute(0, 3, 1, 2), w3.double(), padding=1))
    assert rel(h.cpu(), h64.permute(0, 2, 3, 1)) <= 4e-3
    o64 = F.relu(x.double().permute(0, 3, 1, 2) + F.conv2d(h.cpu().double().permute(0, 3, 1, 2), w1.double()))
    assert rel(o.cpu(), o64.permute(0, 2, 3, 1)) <= 4e-3


# ---- flat-slab convolutions (one TMA slab per filter row, taps as shifted descriptor views) ----
FLAT = [
    # name, H, W, Cin, Cout, R, pad  (stride-1 'same'/valid convs on a zero-bordered grid)
    ("res3x3", 20, 38, 128, 32, 3, 1),
    ("conv4", 22, 40, 128, 128, 3, 0),
    ("conv6", 20, 38, 128, 64, 5, 2),
    ("dec1", 20, 38, 64, 128, 3, 1),
]


@pytest.mark.parametrize("N", [2, 37])
@pytest.mark.parametrize("layer", FLAT, ids=[l[0] for l in FLAT])
def test_conv_flat_fwd_dgrad_wgrad(ops, layer, N):
    from gabril_carla_b200 import _lib
    name, H, W, Cin, Cout, R, pad = layer
    g = torch.Generator().manual_seed(len(name) * 11 + N)
    Hg, Wg = H + 2 * pad, W + 2 * pad
    OH, OW = Hg - R + 1, Wg - R + 1
    x = (torch.randn(N, H, W, Cin, generator=g) * 0.5).to(torch.bfloat16)
    w = (torch.randn(Cout, Cin, R, R, generator=g) / (Cin * R * R) ** 0.5).to(torch.bfloat16)
    bias = torch.randn(Cout, generator=g)
    dy = (torch.randn(N, OH, OW, Cout, generator=g) * 0.5).to(torch.bfloat16)
    xp = torch.zeros(N, Hg, Wg, Cin, dtype=torch.bfloat16)
    xp[:, pad:pad + H, pad:pad + W] = x
    xd, wd = xp.to(DEV), ops.w_oihw_to_internal(w).to(DEV)
    x64, w64, dy64 = x.double().permute(0, 3, 1, 2), w.double(), dy.double().permute(0, 3, 1, 2)
    # forward into a tensor with a 1-pixel border (borders must stay untouched)
    y = torch.full((N, OH + 2, OW + 2, Cout), 7.0, device=DEV, dtype=torch.bfloat16)
    ops.conv_flat(xd, wd, y, R, R, bias=bias.to(DEV), valid=(OH, OW), out_geom=(OH + 2, OW + 2, 1, 1), relu=True)
    ref = F.relu(F.conv2d(x64, w64, bias.double(), padding=pad).permute(0, 2, 3, 1))
    yc = y.cpu()
    assert rel(yc[:, 1:-1, 1:-1], ref) <= 4e-3
    assert (yc[:, 0] == 7).all() and (yc[:, -1] == 7).all() and (yc[:, :, 0] == 7).all() and (yc[:, :, -1] == 7).all()
    # sign words of the stored pixels only: the words of the (never stored) border stay zero
    y2 = torch.zeros_like(y)
    sb = torch.zeros(y.numel() // 32, device=DEV, dtype=torch.int32)
    ops.conv_flat(xd, wd, y2, R, R, bias=bias.to(DEV), valid=(OH, OW), out_geom=(OH + 2, OW + 2, 1, 1), relu=True,
                  bits_out=sb)
    assert torch.equal(y2[:, 1:-1, 1:-1], y[:, 1:-1, 1:-1]) and torch.equal(sb.cpu(), pack_signs(y2.cpu()))
    # dgrad: dy zero-bordered by R-1 on the x grid; every pixel of the x grid is produced, gated
    dyp = torch.zeros(N, Hg + R - 1, Wg + R - 1, Cout, dtype=torch.bfloat16)
    dyp[:, R - 1:R - 1 + OH, R - 1:R - 1 + OW] = dy
    gate = torch.randn(N, Hg, Wg, Cin, generator=g).to(torch.bfloat16)
    if Cout % 64 == 0:  # the slab contraction runs over 64-channel chunks
        dx = torch.full((N, Hg, Wg, Cin), float("nan"), device=DEV, dtype=torch.bfloat16)
        ops.conv_flat(dyp.to(DEV), wd, dx, R, R, gate=gate.to(DEV), valid=(Hg, Wg), out_geom=(Hg, Wg, 0, 0),
                      transpose_w=True)
        dref = torch.nn.grad.conv2d_input((N, Cin, Hg, Wg), w64, dy64).permute(0, 2, 3, 1) * (gate.double() > 0)
        assert rel(dx.cpu(), dref) <= 4e-3
        dx2 = torch.full_like(dx, float("nan"))
        ops.conv_flat(dyp.to(DEV), wd, dx2, R, R, gate_bits=pack_signs(gate).to(DEV), valid=(Hg, Wg),
                      out_geom=(Hg, Wg, 0, 0), transpose_w=True)
        assert torch.equal(dx2, dx)
    # wgrad on the x grid: dy stored at the top-left of a zero tensor with x's geometry
    dyg = torch.zeros(N, Hg, Wg, Cout, dtype=torch.bfloat16)
    dyg[:, :OH, :OW] = dy
    dw = torch.zeros(Cout, R * R, Cin, device=DEV, dtype=torch.float32)
    ops.conv_wgrad_flat(dyg.to(DEV), xd, dw, R, R, 0)
    wref = ops.w_oihw_to_internal(torch.nn.grad.conv2d_weight(x64, (Cout, Cin, R, R), dy64, padding=pad))
    assert rel(dw.cpu(), wref) <= 2e-5
    dw2 = torch.zeros_like(dw)
    db = torch.zeros(Cout, device=DEV)
    ops.conv_wgrad_flat(dyg.to(DEV), xd, dw2, R, R, 0, dbias=db)   # fused bias gradient
    assert rel(dw2.cpu(), wref) <= 2e-5
    assert rel(db.cpu(), dy.double().sum((0, 1, 2))) <= 2e-5
    _lib.poll_error()


@pytest.mark.parametrize("N", [2, 19])
def test_conv2_s2d_wgrad(ops, N):
    """conv_2's weight gradient on the space-to-depth grid (2x2 taps, dY one pixel down-right, zero-bordered dY): the
    engine's call, served by the weight gradient with the shift on dY; vs the fp64 flat sum it stands for."""
    from gabril_carla_b200 import _lib
    SH, SW, Cx, Cy = 47, 82, 128, 64
    g = torch.Generator().manual_seed(31 + N)
    x = (torch.randn(N, SH, SW, Cx, generator=g) * 0.5).to(torch.bfloat16)
    dy = torch.zeros(N, SH, SW, Cy, dtype=torch.bfloat16)
    dy[:, 1:46, 1:81] = (torch.randn(N, 45, 80, Cy, generator=g) * 0.5).to(torch.bfloat16)
    P = N * SH * SW
    X, D = x.reshape(P, Cx).double(), dy.reshape(P, Cy).double()
    ref = torch.zeros(Cy, 4, Cx, dtype=torch.float64)
    for r in range(2):
        for s_ in range(2):
            sh = SW + 1 - (r * SW + s_)          # dw[cy][r,s][cx] = sum_q dy[q + SW + 1][cy] x[q + r*SW + s][cx]
            ref[:, r * 2 + s_] = D[sh:].t() @ X[:P - sh]
    dw = torch.zeros(Cy, 4, Cx, device=DEV)
    db = torch.zeros(Cy, device=DEV)
    ops.conv_wgrad_flat(dy.to(DEV), x.to(DEV), dw, 2, 2, SW + 1, dbias=db)
    assert rel(dw.cpu(), ref) <= 2e-5
    assert rel(db.cpu(), dy.double().sum((0, 1, 2))) <= 2e-5
    _lib.poll_error()


# ---- conv_1 over the 4x4 space-to-depth frame image (csrc/conv1_s4d.cu) ---------------------------
@pytest.mark.parametrize("gray,S,L", [(True, 2, 2), (False, 2, 3), (True, 1, 2), (True, 3, 5)])
def test_obs_s4d_bit_exact(ops, gray, S, L):
    """u8 frames -> bf16 cells: window gather (integer, bit-exact), /255, gray, rounding, cell layout."""
    from oracle import gabril_oracle as O
    from tests.synth import make_batch
    img, _, _ = make_batch(3, L=L, seed=2, smooth=False)
    B, _, H, W, _ = img.shape
    SH, SW = H // 4 + 2, W // 4 + 2
    C0 = S * (1 if gray else 3)
    Cp = ops.conv1_kp(S, gray)
    x4 = torch.full((B * SH * SW + 1, Cp), float("nan"), device=DEV, dtype=torch.bfloat16)
    x4[-1] = 0
    ops.obs_s4d(img.to(DEV), S, L - 1, gray, SH, SW, x4)
    obs = O.format_obs(O.gather_stack(img, L - 1, S), gray)            # [B, C0, H, W] fp32
    canvas = torch.zeros(B, C0, 4 * SH, 4 * SW)
    canvas[:, :, 4:4 + H, 4:4 + W] = obs                               # cell (u, v) = rows 4(u-1) .. of the image
    ref = canvas.reshape(B, C0, SH, 4, SW, 4).permute(0, 2, 4, 3, 5, 1).reshape(B * SH * SW, 16 * C0)
    got = x4[:-1].float().cpu()
    assert torch.equal(got[:, :16 * C0], ref.to(torch.bfloat16).float())
    assert torch.count_nonzero(got[:, 16 * C0:]) == 0                  # channel padding (RGB: 96 -> 128)


@pytest.mark.parametrize("gray,N", [(True, 3), (False, 2)])
def test_conv1_s4d_fwd_wgrad(ops, gray, N):
    """conv_1 + ReLU as a 2x2 flat convolution over the cells vs fp64 F.conv2d of the same bf16 operands:
    forward in space-to-depth storage (zero border quadrants, sign words), per-pixel mask, and the
    weight gradient folded back to conv_1's own layout."""
    S, H, W, C1 = 2, 180, 320, 32
    C0 = S * (1 if gray else 3)
    Cp = ops.conv1_kp(S, gray)
    SH, SW, H1, W1 = H // 4 + 2, W // 4 + 2, H // 2, W // 2
    g = torch.Generator().manual_seed(11)
    x = torch.rand(N, C0, H, W, generator=g).to(torch.bfloat16)
    w = (torch.randn(C1, C0, 4, 4, generator=g) * 0.3).to(torch.bfloat16)
    bias = torch.randn(C1, generator=g) * 0.2
    canvas = torch.zeros(N, C0, 4 * SH, 4 * SW, dtype=torch.bfloat16)
    canvas[:, :, 4:4 + H, 4:4 + W] = x
    x4 = torch.zeros(N * SH * SW + 1, Cp, dtype=torch.bfloat16)
    x4[:-1, :16 * C0] = canvas.reshape(N, C0, SH, 4, SW, 4).permute(0, 2, 4, 3, 5, 1).reshape(N * SH * SW, 16 * C0)
    x4 = x4.to(DEV)
    w1 = torch.zeros(C1, 1, Cp)
    w1[:, 0, :16 * C0] = w.float().permute(0, 2, 3, 1).reshape(C1, 16 * C0)      # master layout (kh, kw, ci)
    w4 = torch.empty(4 * C1, 2, 2 * Cp, device=DEV, dtype=torch.bfloat16)
    ops.conv1_s4d_expand(w1.to(DEV), w4, C0)
    rowmask = (torch.rand(N * SH * SW * 4, generator=g) > 0.3).float() * 0.5 + 0.25
    y = torch.full((N, SH, SW, 4 * C1), float("nan"), device=DEV, dtype=torch.bfloat16)
    sb = torch.full((y.numel() // 32,), -1, device=DEV, dtype=torch.int32)
    ops.conv1_s4d_fwd(x4, w4, bias.to(DEV), y, N, SH, SW, H1, W1, relu=True, rowmask=rowmask.to(DEV), bits_out=sb)
    torch.cuda.synchronize()
    ref = F.relu(F.conv2d(x.double(), w.double(), bias.double(), stride=2, padding=1))          # [N, C1, H1, W1]
    refp = torch.zeros(N, C1, 2 * SH, 2 * SW, dtype=torch.float64)
    refp[:, :, 1:1 + H1, 1:1 + W1] = ref
    ref_s2d = refp.reshape(N, C1, SH, 2, SW, 2).permute(0, 2, 4, 3, 5, 1).reshape(N, SH, SW, 4 * C1)
    ref_s2d = ref_s2d * rowmask.double().reshape(N, SH, SW, 4, 1).expand(N, SH, SW, 4, C1).reshape(N, SH, SW, 4 * C1)
    got = y.float().cpu()
    assert torch.isfinite(got).all()
    assert rel(got, ref_s2d) <= 4e-3
    border = refp.reshape(N, C1, SH, 2, SW, 2).permute(0, 2, 4, 3, 5, 1).reshape(N, SH, SW, 4 * C1) == 0
    inside = torch.zeros(N, 1, 2 * SH, 2 * SW, dtype=torch.bool)
    inside[:, :, 1:1 + H1, 1:1 + W1] = True
    inside = inside.reshape(N, 1, SH, 2, SW, 2).permute(0, 2, 4, 3, 5, 1).expand(N, SH, SW, 2, 2, C1).reshape(N, SH, SW, 4 * C1)
    assert torch.count_nonzero(got[~inside]) == 0                      # padding pixels store exact zeros
    assert torch.equal(sb.cpu(), pack_signs(y.cpu()))
    # weight gradient: dy in the same storage (zero at padding pixels, like the gated gradient)
    dy = (torch.randn(N, SH, SW, 4 * C1, generator=g) * 0.1).to(torch.bfloat16) * inside
    dw4 = torch.zeros(4 * C1, 2, 2 * Cp, device=DEV)
    ops.conv1_s4d_wgrad(dy.to(DEV), x4, dw4, N, SH, SW, H1, W1)
    dw1 = torch.zeros(C1, 1, Cp, device=DEV)
    dw1[:, 0, 0] = 1.0                                                  # fold accumulates
    ops.conv1_s4d_fold(dw4, dw1, C0)
    dyi = dy.double().reshape(N, SH, SW, 2, 2, C1).permute(0, 5, 1, 3, 2, 4).reshape(N, C1, 2 * SH, 2 * SW)
    dyi = dyi[:, :, 1:1 + H1, 1:1 + W1]
    wd = w.double().requires_grad_(True)
    (F.conv2d(x.double(), wd, None, stride=2, padding=1) * dyi).sum().backward()
    ref_dw = wd.grad.permute(0, 2, 3, 1).reshape(C1, 16 * C0)
    got_dw = dw1.cpu()[:, 0].double()
    got_dw[:, 0] -= 1.0
    assert rel(got_dw[:, :16 * C0], ref_dw) <= 2e-3
    assert torch.count_nonzero(got_dw[:, 16 * C0:]) == 0
    # fused bias gradient: the 4*C1 columns fold onto conv_1's C1 channels
    dw4b = torch.zeros_like(dw4)
    db = torch.zeros(C1, device=DEV)
    ops.conv1_s4d_wgrad(dy.to(DEV), x4, dw4b, N, SH, SW, H1, W1, dbias=db)
    assert rel(dw4b.cpu(), dw4.double().cpu()) <= 2e-5
    assert rel(db.cpu(), dy.double().reshape(-1, 4, C1).sum((0, 1))) <= 2e-5
