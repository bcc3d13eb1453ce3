"""Callable, autograd-enabled drop-ins for the reference's network modules over libgabril_b200's kernels
(SURVEY.md §8b "Python surface to keep"; INTEGRATION.md level 2):

  Encoder(input_channels, embedding_dim, num_hiddens, num_residual_layers, num_residual_hiddens)
      .forward(inputs, dropout_mask=None)        vlm_gaze/models/linear_models.py:124-207
  make_pre_actor() = nn.Sequential(nn.Flatten(), Linear(48640, 256))        train_bc.py:79
  make_actor()     = nn.Sequential(Linear(256, 256), nn.ReLU(), Linear(256, action_dim))   train_bc.py:81

Parameters are ordinary fp32 nn.Parameters in the REFERENCE layout (OIHW, [out, in]) under the reference's
state_dict keys, so torch.optim / torch.save / load_state_dict(strict=True) work unchanged and a checkpoint
written by the reference loads as is.  The arithmetic is the tcgen05 / TMA kernels: every forward re-derives
the kernels' internal weight layout (bf16, [Cout][taps][Cin]) from the parameters when they changed, and
backward returns the kernels' gradients converted back to the reference layout.  No CPU path.

One forward may be in flight per module instance (the kernels keep the activations of the last forward in
the module's buffers, like the fused StepEngine does); a second forward before backward raises in backward.
"""
from typing import Optional

import torch
import torch.nn as nn

from . import ops
from .model import ParamArena, PolicySpec, to_reference


class _Params(nn.Module):
    """weight (+ bias) holder that gives the reference's state_dict key names."""

    def __init__(self, w_shape, bias: bool):
        super().__init__()
        self.weight = nn.Parameter(torch.zeros(w_shape))
        if bias:
            self.bias = nn.Parameter(torch.zeros(w_shape[0]))


def _weight_init(m: _Params, conv: bool):
    """weight_init (linear_models.py:6-16): orthogonal (gain sqrt(2) for convs), zero bias."""
    nn.init.orthogonal_(m.weight.data, nn.init.calculate_gain("relu") if conv else 1.0)
    if hasattr(m, "bias"):
        m.bias.data.fill_(0.0)


class _EncoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, inputs, dropout_mask, *params):
        ctx.mod = mod
        ctx.ticket = mod._run_forward(inputs, dropout_mask)
        eng = mod._engine
        return eng.buf["z"].permute(0, 3, 1, 2).to(torch.float32).contiguous()   # reference layout [B,E,h,w]

    @staticmethod
    def backward(ctx, dz):
        grads = ctx.mod._run_backward(dz, ctx.ticket)
        return (None, None, None) + tuple(grads)


class Encoder(nn.Module):
    """Encoder(input_channels, embedding_dim, num_hiddens, num_residual_layers, num_residual_hiddens)
    (linear_models.py:124-207): forward(inputs fp32 [B,C,H,W], dropout_mask=None) -> fp32 [B,E,h,w].
    dropout_mask: gaze heat maps [B,S,H,W] for the in-encoder gaze dropout after conv_1 / conv_2
    (linear_models.py:194-199; train mode draws, eval mode scales by the keep probability)."""

    def __init__(self, input_channels: int, embedding_dim: int = 64, num_hiddens: int = 128,
                 num_residual_layers: int = 2, num_residual_hiddens: int = 32):
        super().__init__()
        nh, c = num_hiddens, input_channels
        self.cfg = (c, embedding_dim, nh, num_residual_layers, num_residual_hiddens)
        self._conv_1 = _Params((nh // 4, c, 4, 4), True)
        self._conv_2 = _Params((nh // 2, nh // 4, 4, 4), True)
        self._conv_3 = _Params((nh, nh // 2, 4, 4), True)
        self._conv_4 = _Params((nh, nh, 3, 3), True)
        stack = nn.Module()
        stack._layers = nn.ModuleList()
        for _ in range(num_residual_layers):
            layer = nn.Module()
            layer._block = nn.ModuleDict({"1": _Params((num_residual_hiddens, nh, 3, 3), False),
                                          "3": _Params((nh, num_residual_hiddens, 1, 1), False)})
            stack._layers.append(layer)
        self._residual_stack = stack
        self._conv_5 = _Params((nh, nh, 5, 5), True)
        self._conv_6 = _Params((embedding_dim, nh, 5, 5), True)
        for m in self.modules():
            if isinstance(m, _Params):
                _weight_init(m, conv=True)
        self._engine = None
        self._versions = None
        self._ticket = 0

    # ---- engine plumbing -------------------------------------------------------------------------
    def _ensure_engine(self, device, H, W):
        from .engine import StepEngine
        from .gaze import GazePreprocessor
        c, e, nh, nr, rh = self.cfg
        if self._engine is None or (self._engine.spec.H, self._engine.spec.W) != (H, W) or \
                self._engine.device != device:
            spec = PolicySpec(c, H, W, e, nh, nr, rh)
            arena = ParamArena(spec, device)
            self._engine = StepEngine(spec, arena, GazePreprocessor(img_height=H, img_width=W, device=str(device)),
                                      gaze_method="None", dp_method="None", grayscale=True, frame_stack=c)
            self._versions = None
        return self._engine

    def _sync_params(self):
        """Parameters (reference layout) -> the kernels' internal fp32 / bf16 arenas, when they changed."""
        sd = {k: v for k, v in self.state_dict().items()}
        versions = tuple((p.data_ptr(), p._version) for p in self.parameters())
        if versions != self._versions:
            self._engine.arena.load_reference({"encoder": sd})
            self._versions = versions

    def _run_forward(self, inputs, dropout_mask):
        if not inputs.is_cuda:
            raise RuntimeError("Encoder: CUDA tensor required (no CPU fallback)")
        B, C, H, W = inputs.shape
        if C != self.cfg[0]:
            raise RuntimeError(f"Encoder: expected {self.cfg[0]} input channels, got {C}")
        eng = self._ensure_engine(inputs.device, H, W)
        self._sync_params()
        if B != eng.B:
            eng._alloc(B)
        s, b = eng.spec, eng.buf
        # fp32 NCHW -> 4x4 space-to-depth bf16 cells (host-side layout plumbing; csrc/conv1_s4d.cu)
        SH, SW = eng.SH, eng.SW
        canvas = inputs.new_zeros(B, C, 4 * SH, 4 * SW)
        canvas[:, :, 4:4 + H, 4:4 + W] = inputs
        cells = canvas.reshape(B, C, SH, 4, SW, 4).permute(0, 2, 4, 3, 5, 1).reshape(B * SH * SW, 16 * C)
        b["x4"][:-1, :16 * C] = cells.to(torch.bfloat16)
        igmd = None
        if dropout_mask is not None:
            g = dropout_mask.to(torch.float32)
            g = (g.unsqueeze(1) if g.dim() == 3 else g).contiguous()
            if "rm1" not in b:
                dev, f32 = inputs.device, torch.float32
                b["rm1"] = torch.zeros(B * SH * SW * 4, device=dev, dtype=f32)
                b["rm2"] = torch.zeros(B * SH * SW, device=dev, dtype=f32)
                b["K1"] = torch.empty(B, s.H1 * s.W1, device=dev, dtype=f32)
                b["K2"] = torch.empty(B, s.H2 * s.W2, device=dev, dtype=f32)
            test = not self.training
            ops.gmd_mask(g, None, s.H1, s.W1, 0.7, test, rowmask=b["rm1"], ws=(b["K1"], b["mnmx"]), layout="s2d",
                         grid=(SH, SW), rng_state=eng.rng_state, draw_id=1)
            ops.gmd_mask(g, None, s.H2, s.W2, 0.7, test, rowmask=b["rm2"], ws=(b["K2"], b["mnmx"]), layout="bordered",
                         grid=(SH, SW), rng_state=eng.rng_state, draw_id=2)
            ops.rng_advance(eng.rng_state)
            igmd = (b["rm1"], b["rm2"])
        eng._encoder_fwd(None, igmd)
        self._ticket += 1
        return self._ticket

    def _run_backward(self, dz, ticket):
        if ticket != self._ticket:
            raise RuntimeError("Encoder: backward of a forward that is no longer the module's last one "
                               "(one forward in flight per module instance)")
        eng = self._engine
        eng.buf["dz"].copy_(dz.permute(0, 2, 3, 1))          # fp32 NCHW -> bf16 NHWC
        eng.arena.g.zero_()
        eng._encoder_bwd()
        gd = eng.arena.grad_dicts()["encoder"]
        return [gd[k].to(p.dtype) if p.requires_grad else None for k, p in self.state_dict(keep_vars=True).items()]

    def forward(self, inputs, dropout_mask=None):
        return _EncoderFn.apply(self, inputs, dropout_mask, *self.parameters())


class _LinearFn(torch.autograd.Function):
    """y = x W^T + b on the tcgen05 kernel (bf16 operands, fp32 accumulate), fp32 in / out."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        M, K = x.shape
        N = weight.shape[0]
        Np = (N + 31) // 32 * 32 if N <= 128 else (N + 127) // 128 * 128
        xb = x.to(torch.bfloat16).contiguous()
        wb = torch.zeros(Np, 1, K, device=x.device, dtype=torch.bfloat16)
        wb[:N, 0] = weight.to(torch.bfloat16)
        # split-K over enough CTAs to cover the machine; every split writes its own plane (deterministic)
        tiles = ((M + 127) // 128) * max(1, Np // 128)
        splits = max(1, min(K // 64, 296 // tiles))
        planes = ops.conv_fwd_planes(K, 1, 1, splits)
        yp = torch.empty(planes, M, Np, device=x.device, dtype=torch.float32)
        ops.conv_fwd(xb.view(M, 1, 1, K), wb, None, yp, 1, 1, 1, 0, out_f32_atomic=2, splits=splits)
        y = yp.sum(0) if planes > 1 else yp[0]
        ctx.save_for_backward(xb, wb)
        ctx.dims = (M, K, N, Np, bias is not None)
        return y[:, :N] + bias if bias is not None else y[:, :N].contiguous()

    @staticmethod
    def backward(ctx, dy):
        xb, wb = ctx.saved_tensors
        M, K, N, Np, has_bias = ctx.dims
        dyb = torch.zeros(M, Np, device=dy.device, dtype=torch.bfloat16)
        dyb[:, :N] = dy
        dx = torch.empty(M, K, device=dy.device, dtype=torch.bfloat16)
        ops.conv_dgrad(dyb.view(M, 1, 1, Np), wb, dx.view(M, 1, 1, K), 1, 1, 1, 0)
        dw = torch.zeros(Np, 1, K, device=dy.device, dtype=torch.float32)
        ops.conv_wgrad(dyb.view(M, 1, 1, Np), xb.view(M, 1, 1, K), dw, 1, 1, 1, 0)
        db = dy.sum(0) if has_bias else None   # [N] reduction of the incoming fp32 gradient (plumbing)
        return dx.to(torch.float32), dw[:N, 0], db


class Linear(nn.Module):
    """nn.Linear's parameters and state_dict keys; forward / backward on the tcgen05 GEMM.  in_features must
    be a multiple of 64 (48640, 256 in the reference's heads)."""

    def __init__(self, in_features: int, out_features: int):
        super().__init__()
        if in_features % 64:
            raise NotImplementedError("Linear: in_features must be a multiple of 64")
        self.in_features, self.out_features = in_features, out_features
        self.weight = nn.Parameter(torch.zeros(out_features, in_features))
        self.bias = nn.Parameter(torch.zeros(out_features))
        nn.init.orthogonal_(self.weight.data)      # weight_init (linear_models.py:6-10, train_bc.py:80,82)

    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError("Linear: CUDA tensor required (no CPU fallback)")
        return _LinearFn.apply(x.to(torch.float32), self.weight, self.bias)


def make_pre_actor(flat: int = 48640, z_dim: int = 256) -> nn.Sequential:
    """pre_actor = nn.Sequential(nn.Flatten(), nn.Linear(flat, z_dim)) (train_bc.py:78-80): keys 1.weight / 1.bias."""
    return nn.Sequential(nn.Flatten(), Linear(flat, z_dim))


def make_actor(z_dim: int = 256, action_dim: int = 7) -> nn.Sequential:
    """actor = nn.Sequential(Linear, ReLU, Linear) (train_bc.py:81-82): keys 0.* / 2.*."""
    return nn.Sequential(Linear(z_dim, z_dim), nn.ReLU(), Linear(z_dim, action_dim))
