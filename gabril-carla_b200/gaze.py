"""Drop-in mirrors of the reference's gaze operators over libgabril_b200's kernels.

  GazePreprocessor   vlm_gaze/data_utils/data_loader_robomimic.py:19-379
  get_gaze_mask      vlm_gaze/data_utils/gaze_utils.py:7-30
  apply_gmd_dropout  vlm_gaze/data_utils/gaze_utils.py:33-52

Same names, argument meaning, return shapes / dtypes and error behaviour (Python exceptions) as
the reference.  No CPU path: tensors must live on a CUDA (sm_100a) device.
"""
from typing import List, Optional

import torch

from . import ops


def _gauss_taps(sigma: float, eps: float = 0.0) -> torch.Tensor:
    """The reference's 1-D kernel (data_loader_robomimic.py:71-79; :280-297 with eps=1e-8),
    computed with the same torch fp32 ops on the host so the taps are bit-identical."""
    size = int(4 * sigma + 1)
    if size % 2 == 0:
        size += 1
    x = torch.arange(size).float() - size // 2
    k = torch.exp(-x ** 2 / (2 * sigma ** 2))
    return k / (k.sum() + eps) if eps else k / k.sum()


class GazePreprocessor(torch.nn.Module):
    """GPU gaze heat-map generation, frame-stack gather and image formatting."""

    def __init__(self, img_height: int = 180, img_width: int = 320, gaze_sigma: float = 30.0,
                 gaze_coeff: float = 0.8, maxpoints: int = 5, device: str = "cuda",
                 temporal_k: int = 0, temporal_alpha: float = 0.7, temporal_beta: float = 0.8,
                 temporal_gamma: float = 1.0, temporal_use_future: bool = True,
                 temporal_mode: str = "alpha_decay", temporal_sigmas: Optional[list] = None,
                 temporal_coeffs: Optional[list] = None, temporal_offset_start: int = 0):
        super().__init__()
        self.img_height, self.img_width = img_height, img_width
        self.maxpoints = maxpoints
        self.gaze_sigma, self.gaze_coeff = gaze_sigma, gaze_coeff
        self.device = device
        self.temporal_k = int(max(0, temporal_k))
        self.temporal_alpha = float(temporal_alpha)
        self.temporal_beta, self.temporal_gamma = float(temporal_beta), float(temporal_gamma)
        self.temporal_use_future = bool(temporal_use_future)
        self.temporal_mode = str(temporal_mode)
        self.temporal_sigmas = [float(x) for x in temporal_sigmas] if temporal_sigmas is not None else None
        self.temporal_coeffs = [float(x) for x in temporal_coeffs] if temporal_coeffs is not None else None
        self.temporal_offset_start = int(max(0, temporal_offset_start))
        k = _gauss_taps(gaze_sigma)
        self.kernel_size = k.numel()
        self.register_buffer("kernel_1d", k.to(device))
        self._plan_cache = {}

    # ---- plans: per-(S, mode) tap tables and aggregation coefficients ------------------------
    def _plan(self, S: int, multiscale: bool, dev):
        key = (S, multiscale, str(dev))
        if key in self._plan_cache:
            return self._plan_cache[key]
        if multiscale:
            taps: List[torch.Tensor] = []
            coeffs = [1.0] * S
            for s in range(S):
                sg = float(self.temporal_sigmas[min(self.temporal_offset_start + s, len(self.temporal_sigmas) - 1)])
                taps.append(_gauss_taps(sg, 1e-8))
                if self.temporal_coeffs:
                    coeffs[s] = float(self.temporal_coeffs[min(self.temporal_offset_start + s,
                                                                len(self.temporal_coeffs) - 1)])
            agg = [1.0 if j <= s else 0.0 for s in range(S) for j in range(S)]
            plan = dict(taps=taps, agg=agg, step=coeffs, norm_base=False)
        else:
            taps = [self.kernel_1d.detach().cpu()] * S
            a = self.temporal_alpha
            # coefficient alpha**(s-j) evaluated in Python double, cast to fp32 (:268)
            agg = [float(torch.tensor(a ** (s - j), dtype=torch.float32)) if j <= s else 0.0
                   for s in range(S) for j in range(S)]
            plan = dict(taps=taps, agg=agg, step=None, norm_base=True)
        stride = max(t.numel() for t in taps)
        tab = torch.zeros(S, stride, dtype=torch.float32)
        for s, t in enumerate(taps):
            tab[s, :t.numel()] = t
        plan["ktab"] = tab.to(dev)
        plan["Lk"] = [t.numel() for t in taps]
        self._plan_cache[key] = plan
        return plan

    def _coords(self, gaze: torch.Tensor) -> torch.Tensor:
        if gaze.dim() == 4:
            gaze = gaze.reshape(gaze.shape[0], gaze.shape[1], -1)
        if gaze.dim() == 2:
            gaze = gaze.unsqueeze(0)
        if gaze.dim() != 3 or gaze.shape[-1] != self.maxpoints * 2:
            raise ValueError(f"gaze coords must be [B,T,{self.maxpoints * 2}] or [B,T,{self.maxpoints},2], "
                             f"got {tuple(gaze.shape)}")
        return gaze.to(torch.float32).contiguous()

    def splat_into(self, gaze_seq: torch.Tensor, S: int, center: int, out: torch.Tensor,
                   aggregate: bool = True):
        """Hot-path entry: [B,L,P*2] -> out [B,S,H,W] (build_stack_heatmaps or, aggregate=False,
        the per-frame maps of forward())."""
        g = self._coords(gaze_seq)
        multiscale = aggregate and self.temporal_mode == "multiscale" and bool(self.temporal_sigmas)
        plan = self._plan(S, multiscale, g.device)
        if max(plan["Lk"]) > self.img_height:
            raise RuntimeError("Calculated padded input size per channel is smaller than the kernel size "
                               f"(kernel {max(plan['Lk'])} > image height {self.img_height})")
        ops.gaze_splat(g, out, plan["ktab"], plan["Lk"], plan["agg"] if aggregate else None, plan["step"],
                       plan["norm_base"], aggregate, S, center, self.maxpoints, self.img_height,
                       self.img_width)
        return out

    # ---- reference API --------------------------------------------------------------------
    def forward(self, gaze_coords: torch.Tensor) -> torch.Tensor:
        """[B,T,P*2] | [B,T,P,2] | [T,P*2] -> per-frame heat maps [B,T,1,H,W] (:85-139)."""
        g = self._coords(gaze_coords)
        B, T = g.shape[0], g.shape[1]
        if T > 8:
            # the splat kernel handles up to 8 frames per call: chunk the time axis
            outs = [self.forward(g[:, t0:t0 + 8]) for t0 in range(0, T, 8)]
            return torch.cat(outs, dim=1)
        out = torch.empty(B, T, self.img_height, self.img_width, device=g.device, dtype=torch.float32)
        self.splat_into(g, T, T - 1, out, aggregate=False)
        return out.unsqueeze(2)

    @staticmethod
    def _gather_last_s_frames(seq: torch.Tensor, center_idx: int, stack_len: int) -> torch.Tensor:
        assert seq.dim() >= 2, f"Expected [B, L, ...], got {seq.shape}"
        L = seq.shape[1]
        idxs = [min(max(i, 0), L - 1) for i in range(center_idx - (stack_len - 1), center_idx + 1)]
        return seq.index_select(1, torch.tensor(idxs, device=seq.device, dtype=torch.long))

    @staticmethod
    def extract_image_stack_around_center(images_seq, center_idx, frame_stack):
        if images_seq.dim() != 5:
            return images_seq
        return GazePreprocessor._gather_last_s_frames(images_seq, center_idx, frame_stack)

    @staticmethod
    def extract_gaze_stack_around_center(gaze_seq, center_idx, frame_stack):
        if gaze_seq.dim() < 3:
            return gaze_seq
        return GazePreprocessor._gather_last_s_frames(gaze_seq, center_idx, frame_stack)

    @staticmethod
    def _format_obs_image(images: torch.Tensor, frame_stack: int, grayscale: bool) -> torch.Tensor:
        """[B,S,H,W,3] uint8 -> [B,S*C',H,W] fp32 (:179-202)."""
        if images.dtype != torch.uint8 or images.dim() != 5 or images.shape[1] != frame_stack:
            raise NotImplementedError("gabril-carla_b200 formats uint8 [B,S,H,W,3] frame stacks only")
        return ops.obs_format(images.contiguous(), frame_stack, frame_stack - 1, grayscale)

    def build_stack_heatmaps(self, gaze_seq: torch.Tensor, frame_stack: int, center_idx: int) -> torch.Tensor:
        g = self._coords(gaze_seq)
        out = torch.empty(g.shape[0], frame_stack, self.img_height, self.img_width, device=g.device,
                          dtype=torch.float32)
        return self.splat_into(g, frame_stack, center_idx, out, aggregate=True)

    def prepare_for_bc(self, obs_image_seq: torch.Tensor, gaze_seq: torch.Tensor, frame_stack: int,
                       grayscale: bool = False, aggregate_stack: bool = True):
        """-> (obs_image [B,S*C',H,W] fp32, gaze_heatmaps [B,S,H,W] fp32, center_idx) (:318-360).
        The frame gather is fused into the formatting kernel (no index_select copy)."""
        center_idx = (obs_image_seq.shape[1] - 1) if obs_image_seq.dim() > 1 else 0
        obs = ops.obs_format(obs_image_seq.contiguous(), frame_stack, center_idx, grayscale)
        g = self._coords(gaze_seq)
        heat = torch.empty(g.shape[0], frame_stack, self.img_height, self.img_width, device=g.device,
                           dtype=torch.float32)
        self.splat_into(g, frame_stack, center_idx, heat, aggregate=aggregate_stack)
        return obs, heat, center_idx

    def prepare_for_gaze_predictor(self, obs_image_seq, gaze_seq, frame_stack, grayscale=False):
        """-> (obs_image, last aggregated heat map [B,1,H,W], center_idx) (:362-379)."""
        obs, heat, center_idx = self.prepare_for_bc(obs_image_seq, gaze_seq, frame_stack, grayscale, True)
        return obs, heat[:, -1].unsqueeze(1), center_idx


class _GazeMaskFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, beta, H, W):
        zc = z.contiguous()
        ctx.save_for_backward(zc)
        ctx.cfg = (float(beta), int(H), int(W))
        return ops.gaze_mask_fwd(zc, H, W, float(beta))

    @staticmethod
    def backward(ctx, grad_out):
        (zc,) = ctx.saved_tensors
        beta, H, W = ctx.cfg
        dz = ops.gaze_mask_bwd(zc, grad_out.contiguous().to(torch.float32), H, W, beta)
        return dz, None, None, None


def get_gaze_mask(z_oreo: torch.Tensor, beta: float, target_size):
    """|z| channel sum -> softmax(./beta) -> bicubic -> per-sample min-max (gaze_utils.py:7-30).
    z_oreo: fp32 [B,C,h,w] (reference layout); differentiable; returns fp32 [B,1,H,W]."""
    if not z_oreo.is_cuda:
        raise RuntimeError("get_gaze_mask: CUDA tensor required (no CPU fallback)")
    H, W = int(target_size[0]), int(target_size[1])
    return _GazeMaskFn.apply(z_oreo.to(torch.float32), beta, H, W)


class _RowMaskFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, rowmask):
        ctx.save_for_backward(rowmask)
        return ops.rowmask_apply_nchw(z.contiguous(), rowmask)

    @staticmethod
    def backward(ctx, grad_out):
        (rowmask,) = ctx.saved_tensors
        return ops.rowmask_apply_nchw(grad_out.contiguous(), rowmask), None


def apply_gmd_dropout(z: torch.Tensor, g: torch.Tensor, test_mode: bool = False, A: torch.Tensor = None):
    """Gaze-mask dropout (gaze_utils.py:33-52).  z fp32 [B,C,h,w]; g [B,S,H,W] or [B,H,W].
    `A` optionally injects the uniform draw (parity tests); by default it is drawn with
    torch.rand([B,1,h,w]) before anything else, like the reference."""
    if not z.is_cuda:
        raise RuntimeError("apply_gmd_dropout: CUDA tensor required (no CPU fallback)")
    B, C, h, w = z.shape
    if A is None:
        A = torch.rand([B, 1, h, w], device=z.device)
    if g.dim() == 3:
        g = g.unsqueeze(1)
    rowmask = ops.gmd_mask(g.to(torch.float32).contiguous(), A.reshape(B, h * w).contiguous(), h, w, 0.7,
                           test_mode)
    return _RowMaskFn.apply(z.to(torch.float32), rowmask)
