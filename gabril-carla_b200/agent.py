"""Inference side of the BC policy: the model part of
vlm_gaze/eval/my_agents/bc_agent.py — `setup` (:40-135: params.json + ep{E}_encoder/_pre_actor/_actor.pth
[+ the gaze predictor's model.torch]), `_process_observation` (:228-271: frame stack, the first frame
repeated, uint8 -> /255 -> gray) and `_predict_control` (:273-305: [gaze predictor ->] encoder ->
test-mode GMD / IGMD -> pre_actor -> actor) — on the forward kernels of the training path.

The CARLA side of BCAgent (sensors, route planner, vector_to_control, video overlay) needs the
simulator and a graphics stack: unavailable on this box, out of scope.  The batch-1
control step is launch-bound (≈25 small kernels), so it is captured once into a CUDA graph and
replayed per observation: one H2D copy of the new frame, one graph launch, one 28-byte D2H read.
There is no CPU fallback.
"""
import json
import os
from typing import Optional

import torch

from . import _lib
from .engine import StepEngine
from .gaze import GazePreprocessor
from .gaze_predictor import GazePredictorEngine, load_autoencoder_state_dict
from .model import GazePredictorSpec, ParamArena, PolicySpec

_UNSUPPORTED = ("ViSaRL", "Mask", "AGIL", "GRIL", "Contrastive")


def _load_sd(path):
    return torch.load(path, map_location="cpu", weights_only=True)


class BCPolicy:
    """policy = BCPolicy.from_checkpoint(params_path); policy.reset(); v = policy.run_step(rgb_u8)"""

    def __init__(self, spec: PolicySpec, arena: ParamArena, *, dp_method: str = "None", grayscale: bool = True,
                 frame_stack: int = 2, gaze_predictor: Optional[GazePredictorEngine] = None,
                 use_graph: bool = True):
        if arena.device.type != "cuda":
            raise RuntimeError("BCPolicy needs a CUDA device (sm_100a); there is no CPU fallback")
        _lib.init()
        self.spec, self.arena, self.device = spec, arena, arena.device
        self.dp_method, self.grayscale, self.stack = dp_method, bool(grayscale), int(frame_stack)
        self.engine = StepEngine(spec, arena, GazePreprocessor(device=str(self.device)), gaze_method="None",
                                 dp_method=dp_method, grayscale=self.grayscale, frame_stack=self.stack)
        self.gaze_predictor = gaze_predictor
        if dp_method in ("GMD", "IGMD") and gaze_predictor is None:
            raise RuntimeError(f"dropout.method={dp_method} needs the gaze predictor at inference "
                               "(bc_agent.py:83-94): params.json must name gaze_predictor_path")
        self.use_graph = bool(use_graph)
        self._graph = None
        self._frames = torch.zeros(1, self.stack, spec.H, spec.W, 3, device=self.device, dtype=torch.uint8)
        self._host_frame = torch.empty(spec.H, spec.W, 3, dtype=torch.uint8).pin_memory()
        self._host_out = torch.empty(spec.A, dtype=torch.float32).pin_memory()
        self._no_gaze = torch.full((1, self.stack, 10), -1.0, device=self.device)
        self._n_frames = 0

    # ---- bc_agent.py:40-135 --------------------------------------------------------------------
    @classmethod
    def from_checkpoint(cls, params_path: str, device="cuda:0", use_graph: bool = True, H: int = 180,
                        W: int = 320) -> "BCPolicy":
        with open(os.path.join(params_path, "params.json")) as f:
            p = json.load(f)
        if p["gaze_method"] in _UNSUPPORTED or p["dp_method"] == "Oreo":
            raise NotImplementedError(f"gaze_method={p['gaze_method']} dp_method={p['dp_method']} is outside "
                                      "the B200 hot path (SURVEY.md §8f N4)")
        S, gray = int(p["stack"]), bool(p["grayscale"])
        cin = S * (1 if gray else 3)
        spec = PolicySpec(cin, H, W, int(p["embedding_dim"]), int(p["num_hiddens"]),
                          int(p["num_residual_layers"]), int(p["num_residual_hiddens"]), int(p["z_dim"]),
                          int(p["action_dim"]))
        arena = ParamArena(spec, device)
        mp, E = p["models_path"], p["epochs"]
        arena.load_reference({m: _load_sd(os.path.join(mp, f"ep{E}_{m}.pth"))
                              for m in ("encoder", "pre_actor", "actor")})
        gp = None
        if p["dp_method"] in ("GMD", "IGMD"):
            gspec = GazePredictorSpec(cin, H, W, int(p["embedding_dim"]), int(p["num_hiddens"]),
                                      int(p["num_residual_layers"]), int(p["num_residual_hiddens"]))
            garena = ParamArena(gspec, device)
            load_autoencoder_state_dict(garena, _load_sd(p["gaze_predictor_path"]))
            garena.refresh_shadow()
            gp = GazePredictorEngine(gspec, garena, GazePreprocessor(device=str(device)), grayscale=gray,
                                     frame_stack=S)
        return cls(spec, arena, dp_method=p["dp_method"], grayscale=gray, frame_stack=S, gaze_predictor=gp,
                   use_graph=use_graph)

    # ---- bc_agent.py:228-271 -------------------------------------------------------------------
    def reset(self):
        self._n_frames = 0

    def _process_observation(self, image_rgb) -> torch.Tensor:
        """image_rgb: uint8 [H,W,3] (numpy or tensor, already at the model's resolution).  Updates the
        device-side frame stack (first observation repeated S times, then a FIFO) and returns it as
        u8 [1,S,H,W,3]; /255, the gray weights and the stacking happen in the conv_1 patch kernel."""
        t = torch.as_tensor(image_rgb)
        if t.dtype != torch.uint8 or tuple(t.shape) != (self.spec.H, self.spec.W, 3):
            raise ValueError(f"expected uint8 [{self.spec.H},{self.spec.W},3], got {t.dtype} {tuple(t.shape)}")
        if t.is_cuda:
            new = t
        else:
            self._host_frame.copy_(t)
            new = self._host_frame
        if self._n_frames == 0:
            for s in range(self.stack):
                self._frames[0, s].copy_(new, non_blocking=True)
        else:
            if self.stack > 1:
                self._frames[0, :-1].copy_(self._frames[0, 1:].clone())
            self._frames[0, -1].copy_(new, non_blocking=True)
        self._n_frames += 1
        return self._frames

    # ---- bc_agent.py:273-305 -------------------------------------------------------------------
    def _forward(self, frames: torch.Tensor) -> torch.Tensor:
        g = None
        if self.gaze_predictor is not None:
            gaze = self._no_gaze if frames.shape[0] == 1 else torch.full(
                (frames.shape[0], self.stack, 10), -1.0, device=self.device)
            self.gaze_predictor.forward(frames, gaze)
            g = self.gaze_predictor.output()          # clipped into [0,1] inside StepEngine.predict
        return self.engine.predict(frames, g)

    def predict(self, frames: torch.Tensor) -> torch.Tensor:
        """frames u8 [B,S,H,W,3] on the device -> action vectors fp32 [B, action_dim] (no graph)."""
        return self._forward(frames.contiguous())

    def _predict_control(self, frames: Optional[torch.Tensor] = None):
        """The batch-1 control step on the internal frame stack -> numpy [action_dim]."""
        if frames is not None and frames.data_ptr() != self._frames.data_ptr():
            self._frames.copy_(frames)
        if not self.use_graph:
            out = self._forward(self._frames)
        else:
            if self._graph is None:
                side = torch.cuda.Stream(device=self.device)
                side.wait_stream(torch.cuda.current_stream(self.device))
                with torch.cuda.stream(side):       # warm up outside the capture (allocations, attrs)
                    for _ in range(2):
                        self._forward(self._frames)
                torch.cuda.current_stream(self.device).wait_stream(side)
                torch.cuda.synchronize(self.device)
                self._graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._graph):
                    self._graph_out = self._forward(self._frames)
            self._graph.replay()
            out = self._graph_out
        self._host_out.copy_(out[0], non_blocking=False)
        _lib.poll_error()
        return self._host_out.numpy().copy()

    def run_step(self, image_rgb):
        """One control step: observation -> action vector (what BCAgent.run_step feeds to
        vector_to_control, bc_agent.py:303-305)."""
        self._process_observation(image_rgb)
        return self._predict_control()
