"""Gaze-mask predictor pre-training on one B200 (BASELINE config 4).

Mirrors GazePredictorTrainer.compute_loss (vlm_gaze/train/train_gaze_predictor.py:83-101) +
loss.backward() + optimizer.step(): AutoEncoder(Encoder, Decoder(out_channels=1))
(vlm_gaze/models/linear_models.py:124-207, 210-282, 356-367) regressing the last aggregated gaze
heat map of GazePreprocessor.prepare_for_gaze_predictor (data_loader_robomimic.py:362-379) with
nn.MSELoss.  The encoder half is StepEngine's; the decoder runs on the same tcgen05 kernels
(nn.ConvTranspose2d = data gradient of a convolution, see include/gabril_b200.h) plus the fused
_conv_trans_4 + MSE head (csrc/gp_head.cu).  No PyTorch math, no CPU fallback.
"""
import json
import os
from typing import Dict

import torch

from . import ops
from .engine import StepEngine
from .gaze import GazePreprocessor
from .model import GazePredictorSpec, ParamArena


class GazePredictorEngine(StepEngine):
    def __init__(self, spec: GazePredictorSpec, arena: ParamArena, gaze_pre: GazePreprocessor, *,
                 grayscale: bool = True, frame_stack: int = 2, grad_accum: int = 1):
        super().__init__(spec, arena, gaze_pre, gaze_method="None", dp_method="None", grayscale=grayscale,
                         frame_stack=frame_stack, grad_accum=grad_accum)

    def _alloc(self, B):
        super()._alloc(B)
        s, dev, b = self.spec, self.device, self.buf
        bf = torch.bfloat16

        def act(h, w, c):
            return torch.empty(B, h, w, c, device=dev, dtype=bf)
        b["q-1"] = act(s.h, s.w, s.NH)
        for i in range(s.NR):
            b[f"g{i}"] = act(s.h, s.w, s.RH)
            b[f"q{i}"] = act(s.h, s.w, s.NH)
        b["t1"], b["t2"], b["t3"] = act(s.H3, s.W3, s.NH), act(s.H2, s.W2, s.C2), act(s.H1, s.W1, s.C1)
        b["dt1"], b["dt2"], b["dt3"] = act(s.H3, s.W3, s.NH), act(s.H2, s.W2, s.C2), act(s.H1, s.W1, s.C1)
        b["dq_a"], b["dq_b"], b["dg"] = act(s.h, s.w, s.NH), act(s.h, s.w, s.NH), act(s.h, s.w, s.RH)
        b["gp_out"] = torch.empty(B, s.H, s.W, device=dev, dtype=torch.float32)
        b["gp_dout"] = torch.empty(B, s.H, s.W, device=dev, dtype=torch.float32)
        b["gp_patches"] = torch.empty(B * s.H1 * s.W1, 32, device=dev, dtype=bf)
        b["gp_part"] = torch.zeros(B, device=dev, dtype=torch.float32)

    # ------------------------------------------------------------------------------------------
    def forward(self, frames: torch.Tensor, gaze: torch.Tensor, actions=None):
        """frames u8 [B,L,H,W,3], gaze f32 [B,L,P*2] -> device tensor losses = [mse, 0, mse]."""
        s, a = self.spec, self.arena
        B, L = frames.shape[0], frames.shape[1]
        if B != self.B:
            self._alloc(B)
        b = self.buf
        center = L - 1
        # prepare_for_gaze_predictor: image stack + causally aggregated heat maps, last one is the target
        ops.obs_patches_s2d(frames, self.S, center, self.grayscale, self.SH, self.SW, b["patches"])
        self.gaze_pre.splat_into(gaze, self.S, center, b["heat"], aggregate=True)
        self._encoder_fwd(None)
        # ---- Decoder.forward (linear_models.py:271-282) ----
        ops.conv_fwd(b["z"], a.WB("dec.conv1.w"), a.P("dec.conv1.b"), b["q-1"], 3, 3, 1, 1, relu=True)  # F7
        for i in range(s.NR):
            prev = b[f"q{i - 1}"]
            ops.conv_fwd(prev, a.WB(f"dec.res{i}.c3.w"), None, b[f"g{i}"], 3, 3, 1, 1, relu=True)
            ops.conv_fwd(b[f"g{i}"], a.WB(f"dec.res{i}.c1.w"), None, b[f"q{i}"], 1, 1, 1, 0, relu=True, resid=prev)
        last = b[f"q{s.NR - 1}"]
        ops.convt_fwd(last, a.WB("dec.ct1.w"), a.P("dec.ct1.b"), b["t1"], 3, 3, 1, 0, relu=True)
        ops.convt_fwd(b["t1"], a.WB("dec.ct2.w"), a.P("dec.ct2.b"), b["t2"], 4, 4, 2, 1, relu=True)
        ops.convt_fwd(b["t2"], a.WB("dec.ct3.w"), a.P("dec.ct3.b"), b["t3"], 4, 4, 2, 1, relu=True)
        # ---- _conv_trans_4 + MSE (train_gaze_predictor.py:96-98) ----
        n_el = B * s.H * s.W
        b["gp_part"].zero_()
        target = b["heat"][:, self.S - 1]
        ops.gp_head_fwd(b["t3"], a.WB("dec.ct4.w"), a.P("dec.ct4.b"), target, b["heat"].stride(0), b["gp_dout"],
                        b["gp_part"], a.G("dec.ct4.b"), 1.0 / (n_el * self.grad_accum), out=b["gp_out"])
        ops.finalize_losses(b["gp_part"], None, B, 1.0 / n_el, 0.0, 0.0, b["losses"])
        return b["losses"]

    def backward(self):
        s, a, b, B = self.spec, self.arena, self.buf, self.B
        M1 = B * s.H1 * s.W1
        # _conv_trans_4: two patch GEMMs over dL/dout laid out as the k4/s2/p1 window of every input pixel
        ops.gp_dout_patches(b["gp_dout"], b["gp_patches"])
        ops.conv_wgrad(b["t3"].view(M1, 1, 1, s.C1), b["gp_patches"].view(M1, 1, 1, 32), a.G("dec.ct4.w"), 1, 1, 1, 0)
        ops.convt_dgrad(b["gp_patches"].view(M1, 1, 1, 32), a.WB("dec.ct4.w"), b["dt3"].view(M1, 1, 1, s.C1),
                        1, 1, 1, 0, gate=b["t3"].view(M1, 1, 1, s.C1))
        # _conv_trans_3 .. _conv_trans_1 (weight gradient: conv wgrad with x and dy exchanged)
        ops.colsum_bf16(b["dt3"], a.G("dec.ct3.b"))
        ops.conv_wgrad(b["t2"], b["dt3"], a.G("dec.ct3.w"), 4, 4, 2, 1)
        ops.convt_dgrad(b["dt3"], a.WB("dec.ct3.w"), b["dt2"], 4, 4, 2, 1, gate=b["t2"])
        ops.colsum_bf16(b["dt2"], a.G("dec.ct2.b"))
        ops.conv_wgrad(b["t1"], b["dt2"], a.G("dec.ct2.w"), 4, 4, 2, 1)
        ops.convt_dgrad(b["dt2"], a.WB("dec.ct2.w"), b["dt1"], 4, 4, 2, 1, gate=b["t1"])
        last = b[f"q{s.NR - 1}"]
        ops.colsum_bf16(b["dt1"], a.G("dec.ct1.b"))
        ops.conv_wgrad(last, b["dt1"], a.G("dec.ct1.w"), 3, 3, 1, 0)
        dt, other = b["dq_a"], b["dq_b"]
        ops.convt_dgrad(b["dt1"], a.WB("dec.ct1.w"), dt, 3, 3, 1, 0, gate=last)
        # decoder residual stack, last block first
        for i in reversed(range(s.NR)):
            prev = b[f"q{i - 1}"]
            ops.conv_wgrad(dt, b[f"g{i}"], a.G(f"dec.res{i}.c1.w"), 1, 1, 1, 0)
            ops.conv_dgrad(dt, a.WB(f"dec.res{i}.c1.w"), b["dg"], 1, 1, 1, 0, gate=b[f"g{i}"])
            ops.conv_wgrad(b["dg"], prev, a.G(f"dec.res{i}.c3.w"), 3, 3, 1, 1)
            ops.conv_dgrad(b["dg"], a.WB(f"dec.res{i}.c3.w"), other, 3, 3, 1, 1, addend=dt, gate=prev)
            dt, other = other, dt
        # decoder _conv_1 (z carries no activation: no gate)
        ops.colsum_bf16(dt, a.G("dec.conv1.b"))
        ops.conv_wgrad(dt, b["z"], a.G("dec.conv1.w"), 3, 3, 1, 1)
        ops.conv_dgrad(dt, a.WB("dec.conv1.w"), b["dz"], 3, 3, 1, 1)
        self._encoder_bwd()
        self._accum_count += 1
        if self.on_all_grads_ready is not None:
            self.on_all_grads_ready()

    def output(self) -> torch.Tensor:
        """The predicted map of the last forward, reference layout [B,1,H,W] fp32."""
        return self.buf["gp_out"].unsqueeze(1)

    def loss_host(self) -> float:
        return float(self.buf["losses"][0].item())


# ------------------------------------------------------------------------------------------------
# checkpoint layout of GazePredictorTrainer.save_for_epoch (train_gaze_predictor.py:103-120)
# ------------------------------------------------------------------------------------------------
def autoencoder_state_dict(arena: ParamArena) -> Dict[str, torch.Tensor]:
    """AutoEncoder.state_dict(): keys `encoder.*` / `decoder.*`, fp32, reference shapes."""
    out = {}
    for mod, sd in arena.state_dicts().items():
        for k, v in sd.items():
            out[f"{mod}.{k}"] = v.cpu()
    return out


def load_autoencoder_state_dict(arena: ParamArena, sd: Dict[str, torch.Tensor]):
    split = {"encoder": {}, "decoder": {}}
    for k, v in sd.items():
        k = k.replace("_orig_mod.", "").replace("module.", "")
        mod, rest = k.split(".", 1)
        split[mod][rest] = v
    arena.load_reference(split)


def gaze_predictor_params(cfg_model, cfg_data, cfg_gaze, ckpt_dir: str, epoch: int) -> dict:
    """The 11 keys of params.json (train_gaze_predictor.py:107-119)."""
    return {
        "model_type": "gaze_predictor", "grayscale": cfg_model["grayscale"], "stack": cfg_data["frame_stack"],
        "embedding_dim": cfg_model["embedding_dim"], "num_hiddens": cfg_model["num_hiddens"],
        "num_residual_layers": cfg_model["num_residual_layers"],
        "num_residual_hiddens": cfg_model["num_residual_hiddens"], "gaze_mask_sigma": cfg_gaze["sigma"],
        "gaze_mask_coeff": cfg_gaze["coeff"], "models_path": str(ckpt_dir), "epochs": epoch,
    }


def save_gaze_predictor(arena: ParamArena, ckpt_dir: str, epoch: int, params: dict, final: bool = False):
    """ExperimentLogger.save_checkpoint layout (train/common/logging.py:70-86): model_ep{E}.torch (+
    model.torch on the final epoch) holding the AutoEncoder state_dict -- the file bc_agent.py:83-94
    loads through params['gaze_predictor_path'] -- and params.json beside it."""
    os.makedirs(ckpt_dir, exist_ok=True)
    sd = autoencoder_state_dict(arena)
    torch.save(sd, os.path.join(ckpt_dir, f"model_ep{epoch}.torch"))
    if final:
        torch.save(sd, os.path.join(ckpt_dir, "model.torch"))
    with open(os.path.join(ckpt_dir, "params.json"), "w") as f:
        json.dump(params, f, indent=2)
