"""Seeded synthetic Bench2Drive-shaped batches shared by tests, smoke and bench."""
import torch


def make_batch(B, L=2, H=180, W=320, P=5, A=7, seed=0, smooth=True, edge_cases=True):
    g = torch.Generator().manual_seed(seed)
    if smooth:
        # low-frequency structure + noise so that bicubic / softmax see something non-trivial
        low = torch.rand(B, L, 6, 10, 3, generator=g)
        img = torch.nn.functional.interpolate(low.permute(0, 1, 4, 2, 3).reshape(B * L, 3, 6, 10),
                                              size=(H, W), mode="bilinear", align_corners=False)
        img = img.reshape(B, L, 3, H, W).permute(0, 1, 3, 4, 2)
        img = (img * 200 + torch.rand(B, L, H, W, 3, generator=g) * 55).clamp(0, 255).to(torch.uint8)
    else:
        img = torch.randint(0, 256, (B, L, H, W, 3), generator=g, dtype=torch.uint8)
    gaze = torch.rand(B, L, P, 2, generator=g)
    nvalid = torch.randint(0, P + 1, (B, L), generator=g)
    for b in range(B):
        for t in range(L):
            gaze[b, t, nvalid[b, t]:] = -1.0
    if edge_cases and B >= 4:
        gaze[0, 0] = -1.0                       # empty frame
        gaze[1, -1, 0] = torch.tensor([0.0, 0.0])
        gaze[1, -1, 1] = torch.tensor([1.0, 1.0])  # exactly on the border
        gaze[2, -1, 1] = gaze[2, -1, 0]         # duplicate point (valid or padding alike)
        gaze[2, -1, 0] = torch.tensor([0.5, 0.5]); gaze[2, -1, 1] = torch.tensor([0.5, 0.5])
        gaze[3, -1, 0] = torch.tensor([-1.0, 0.4])  # negative on one axis only => invalid
    actions = torch.rand(B, L, A, generator=g)
    return img.contiguous(), gaze.reshape(B, L, P * 2).contiguous(), actions.contiguous()


def make_batch_n1(seed=17):
    """Batch of 4 for the `gaze.ratio < 1` selection (train_bc.py:236-240): samples 0/1 are ordinary
    frames (content hash r = 0 => always selected), samples 2/3 are black frames with one or two lit
    pixels whose hash is r = 0.7656 (dropped at ratio 0.5) and r = 0.1699 (kept)."""
    img, gaze, act = make_batch(4, seed=seed)
    img[2:] = 0
    img[2, 1, 100, 200] = torch.tensor([255, 255, 255], dtype=torch.uint8)
    img[2, 0, 17, 5] = torch.tensor([33, 77, 120], dtype=torch.uint8)
    img[3, 1, 60, 300] = torch.tensor([128, 0, 0], dtype=torch.uint8)
    return img.contiguous(), gaze, act


# reference state_dict layout: key -> shape, c = encoder input channels
def reference_shapes(c_in, action_dim=7):
    enc = {"_conv_1.weight": (32, c_in, 4, 4), "_conv_1.bias": (32,),
           "_conv_2.weight": (64, 32, 4, 4), "_conv_2.bias": (64,),
           "_conv_3.weight": (128, 64, 4, 4), "_conv_3.bias": (128,),
           "_conv_4.weight": (128, 128, 3, 3), "_conv_4.bias": (128,),
           "_residual_stack._layers.0._block.1.weight": (32, 128, 3, 3),
           "_residual_stack._layers.0._block.3.weight": (128, 32, 1, 1),
           "_residual_stack._layers.1._block.1.weight": (32, 128, 3, 3),
           "_residual_stack._layers.1._block.3.weight": (128, 32, 1, 1),
           "_conv_5.weight": (128, 128, 5, 5), "_conv_5.bias": (128,),
           "_conv_6.weight": (64, 128, 5, 5), "_conv_6.bias": (64,)}
    pre = {"1.weight": (256, 48640), "1.bias": (256,)}
    act = {"0.weight": (256, 256), "0.bias": (256,), "2.weight": (action_dim, 256), "2.bias": (action_dim,)}
    return {"encoder": enc, "pre_actor": pre, "actor": act}


def make_weights(c_in, seed=7, action_dim=7):
    """Deterministic (torch CPU generator) He-scaled weights and small non-zero biases in the
    reference's state_dict layout — independent of LAPACK (orthogonal_) so that the committed
    golden vectors can be regenerated bit-identically anywhere."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for mod, shapes in reference_shapes(c_in, action_dim).items():
        sd = {}
        for k, shp in shapes.items():
            if k.endswith("bias"):
                sd[k] = (torch.rand(shp, generator=g) - 0.5) * 0.1
            else:
                fan_in = 1
                for d in shp[1:]:
                    fan_in *= d
                gain = 2.0 ** 0.5 if len(shp) == 4 else 1.0
                sd[k] = torch.randn(shp, generator=g) * (gain / fan_in ** 0.5)
        out[mod] = sd
    return out


def decoder_shapes():
    """Decoder(1, 64, 128, 2, 32).state_dict() shapes (linear_models.py:210-268); nn.ConvTranspose2d
    weights are [in, out, k, k]."""
    return {
        "_conv_1.weight": (128, 64, 3, 3), "_conv_1.bias": (128,),
        "_residual_stack._layers.0._block.1.weight": (32, 128, 3, 3),
        "_residual_stack._layers.0._block.3.weight": (128, 32, 1, 1),
        "_residual_stack._layers.1._block.1.weight": (32, 128, 3, 3),
        "_residual_stack._layers.1._block.3.weight": (128, 32, 1, 1),
        "_conv_trans_1.weight": (128, 128, 3, 3), "_conv_trans_1.bias": (128,),
        "_conv_trans_2.weight": (128, 64, 4, 4), "_conv_trans_2.bias": (64,),
        "_conv_trans_3.weight": (64, 32, 4, 4), "_conv_trans_3.bias": (32,),
        "_conv_trans_4.weight": (32, 1, 4, 4), "_conv_trans_4.bias": (1,),
    }


def make_weights_gp(c_in, seed=9):
    """Deterministic weights of the gaze-predictor AutoEncoder: {"encoder": sd, "decoder": sd}."""
    enc = make_weights(c_in, seed=seed)["encoder"]
    g = torch.Generator().manual_seed(seed + 1000)
    dec = {}
    for k, shp in decoder_shapes().items():
        if k.endswith("bias"):
            dec[k] = (torch.rand(shp, generator=g) - 0.5) * 0.1
        else:
            # fan-in of a transposed convolution: in_channels * k*k / stride^2 taps reach one output
            fan_in = shp[0] * shp[2] * shp[3] if "trans" in k else shp[1] * shp[2] * shp[3]
            if "trans" in k and shp[2] == 4:
                fan_in //= 4
            dec[k] = torch.randn(shp, generator=g) * ((2.0 / fan_in) ** 0.5)
    return {"encoder": enc, "decoder": dec}


def make_weights_gril(seed=9, z_dim=256, max_points=5):
    """Deterministic weights of the GRIL gaze-coordinate head nn.Sequential(Linear(z, z), ReLU, Linear(z, 2P))
    (train_bc.py:84-86), reference state_dict keys."""
    g = torch.Generator().manual_seed(seed)
    return {"0.weight": torch.randn(z_dim, z_dim, generator=g) / z_dim ** 0.5,
            "0.bias": (torch.rand(z_dim, generator=g) - 0.5) * 0.1,
            "2.weight": torch.randn(2 * max_points, z_dim, generator=g) / z_dim ** 0.5,
            "2.bias": (torch.rand(2 * max_points, generator=g) - 0.5) * 0.1}
