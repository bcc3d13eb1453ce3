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

from . import ddp, ops
from .engine import StepEngine
from .gaze import GazePreprocessor
from .model import GazePredictorSpec, ParamArena


class GazePredictorEngine(StepEngine):
    def __init__(self, spec: GazePredictorSpec, arena: ParamArena, gaze_pre: GazePreprocessor, *,
                 grayscale: bool = True, frame_stack: int = 2, grad_accum: int = 1):
        super().__init__(spec, arena, gaze_pre, gaze_method="None", dp_method="None", grayscale=grayscale,
                         frame_stack=frame_stack, grad_accum=grad_accum)
        self.use_graph = False   # StepEngine.train_step's graph split follows the policy's backward

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
        ops.obs_s4d(frames, self.S, center, self.grayscale, self.SH, self.SW, b["x4"])
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
        closing = self._accum_count + 1 >= self.grad_accum    # hooks fire on the window-closing micro-step
        if closing and self.on_heads_grads_ready is not None:   # decoder gradients complete: first all-reduce bucket
            self.on_heads_grads_ready()
        self._encoder_bwd()
        self._accum_count += 1
        if closing and self.on_all_grads_ready is not None:
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


# ------------------------------------------------------------------------------------------------
# GazePredictorTrainer (train_gaze_predictor.py:22-121): the five BaseTrainer hooks over the engine
# ------------------------------------------------------------------------------------------------
def _trainer_classes():
    from .trainer import ArenaModule, BaseTrainer, _StepLoss

    class _AutoEncoderFacade:
        """state_dict()/load_state_dict()/parameters() of the reference's AutoEncoder module."""

        def __init__(self, arena):
            self.arena = arena
            self.encoder, self.decoder = ArenaModule(arena, "encoder"), ArenaModule(arena, "decoder")
            self.training = True

        def state_dict(self):
            return autoencoder_state_dict(self.arena)

        def load_state_dict(self, sd, strict: bool = True):
            load_autoencoder_state_dict(self.arena, sd)

        def parameters(self):
            return self.encoder.parameters() + self.decoder.parameters()

        def train(self, mode: bool = True):
            self.training = mode
            return self

        def eval(self):
            return self.train(False)

    class GazePredictorTrainer(BaseTrainer):
        def build_models(self):
            c = self.cfg
            if self.device.type != "cuda":
                raise RuntimeError("GazePredictorTrainer needs a CUDA device (sm_100a); there is no CPU fallback")
            S = int(c.data.frame_stack)
            spec = GazePredictorSpec(S * (1 if c.model.grayscale else 3), int(c.data.img_height),
                                     int(c.data.img_width), int(c.model.embedding_dim), int(c.model.num_hiddens),
                                     int(c.model.num_residual_layers), int(c.model.num_residual_hiddens))
            self.spec = spec
            self.arena = ParamArena(spec, self.device)
            self.arena.init_reference_style()
            ddp.broadcast_arena(self.arena.p)
            self.arena.refresh_shadow()
            g = c.gaze
            self.gaze_preprocessor = GazePreprocessor(
                img_height=int(c.data.img_height), img_width=int(c.data.img_width),
                gaze_sigma=float(g.get("sigma", g.get("mask_sigma", 30.0))),
                gaze_coeff=float(g.get("coeff", g.get("mask_coeff", 0.8))), maxpoints=int(g.max_points),
                device=str(self.device), temporal_mode=str(g.get("temporal_mode", "alpha_decay")),
                temporal_sigmas=g.get("temporal_sigmas", None), temporal_coeffs=g.get("temporal_coeffs", None),
                temporal_offset_start=int(g.get("temporal_offset_start", 0)))
            self.engine = GazePredictorEngine(spec, self.arena, self.gaze_preprocessor,
                                              grayscale=bool(c.model.grayscale), frame_stack=S,
                                              grad_accum=int(c.training.gradient_accumulation_steps))
            self.reducer = ddp.attach(self.engine, self.arena)
            self.model = _AutoEncoderFacade(self.arena)

        def get_optim_params(self):
            return self.model.parameters()

        def set_train_mode(self):
            self.model.train()

        def compute_loss(self, batch):
            """-> (loss tensor, batch size, {"loss": float}) like train_gaze_predictor.py:83-101."""
            gaze_key = self.cfg.data.get("gaze_key", "gaze_coords")
            frames = batch["obs"]["image"].to(self.device, non_blocking=True)
            gaze = batch["obs"][gaze_key].to(self.device, non_blocking=True)
            if gaze.dim() == 4:
                gaze = gaze.reshape(gaze.shape[0], gaze.shape[1], -1)
            losses = self.engine.forward(frames.contiguous(), gaze.contiguous().float())
            value = self.engine.loss_host()
            anchor = torch.zeros((), device=self.device, requires_grad=True)
            loss = _StepLoss.apply(anchor, self, losses[0])
            return loss, int(frames.shape[0]), {"loss": value, "Loss/total": value}

        def _run_backward(self):
            self.engine.backward()

        def zero_grad(self):
            self.engine.zero_grad()

        def optimizer_step(self):
            self.reducer.finish()
            self.opt_step += 1
            s = self.schedule
            self.engine.adam_step(self.opt_step, s.lr, weight_decay=s.weight_decay, decoupled=s.decoupled,
                                  grad_scale=self.reducer.grad_scale)

        def save_for_epoch(self, epoch: int):
            c = self.cfg
            g = c.gaze
            params = gaze_predictor_params(
                dict(grayscale=bool(c.model.grayscale), embedding_dim=int(c.model.embedding_dim),
                     num_hiddens=int(c.model.num_hiddens), num_residual_layers=int(c.model.num_residual_layers),
                     num_residual_hiddens=int(c.model.num_residual_hiddens)),
                dict(frame_stack=int(c.data.frame_stack)),
                dict(sigma=float(g.get("sigma", g.get("mask_sigma", 30.0))),
                     coeff=float(g.get("coeff", g.get("mask_coeff", 0.8)))),
                str(self.checkpoint_dir), int(epoch))
            save_gaze_predictor(self.arena, str(self.checkpoint_dir), int(epoch), params,
                                final=(int(epoch) == int(c.training.epochs)))
            self._print_rank0(f"Saved checkpoint: {self.checkpoint_dir}/model_ep{epoch}.torch")

    return GazePredictorTrainer


def __getattr__(name):
    if name == "GazePredictorTrainer":
        return _trainer_classes()
    raise AttributeError(name)


def main(argv=None):
    """`python -m gabril_carla_b200.train_gaze_predictor key=value ...` (train_gaze_predictor.py:124-140);
    defaults of train_gaze.yaml that differ from train_bc_base.yaml: seed 1, step scheduler."""
    import sys
    from .config import load_config
    args = ["training.seed=1", "scheduler.type=step"] + list(sys.argv[1:] if argv is None else argv)
    cfg = load_config(args)
    trainer = _trainer_classes()(cfg)
    trainer.train()
    if os.environ.get("RANK", "0") == "0":
        print("Training completed!")
    ddp.destroy_distributed_if_initialized()
    return trainer
