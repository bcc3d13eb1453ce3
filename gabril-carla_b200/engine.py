"""The fused GABRIL / GABRIL+GMD training step on one B200.

Mirrors BCTrainer.compute_loss (vlm_gaze/train/train_bc.py:203-299) + loss.backward() +
optimizer.step() (vlm_gaze/train/common/base_trainer.py:129-149) as an explicit kernel schedule:
every arithmetic operation below is one of libgabril_b200's sm_100a kernels; PyTorch only owns the
buffers, the stream and the RNG draw of GMD's uniform tensor.

Activation layout: NHWC bf16.  Saved-for-backward tensors are the post-ReLU activations
themselves (ReLU backward is a `gate > 0` predicate in the dgrad epilogue).  Quirks reproduced
(SURVEY F6-F8): Residual's in-place ReLU (skip branch sees relu(x)), conv_4 activated by the first
Residual, the regulariser reading the post-dropout z under GMD.
"""
import os
from typing import Dict, Optional

import torch

from . import _lib, ops
from .gaze import GazePreprocessor
from .model import ParamArena, PolicySpec


class StepEngine:
    def __init__(self, spec: PolicySpec, arena: ParamArena, gaze_pre: GazePreprocessor, *,
                 gaze_method: str = "Reg", dp_method: str = "None", grayscale: bool = True,
                 frame_stack: int = 2, beta: float = 50.0, lam: float = 10.0,
                 aggregate: bool = True, grad_accum: int = 1, seed: int = 0,
                 prob_dist_type: str = "MSE", gaze_ratio: float = 1.0):
        if gaze_method in ("Contrastive",) or dp_method == "Oreo":
            # recorded from the unmodified reference by tests/golden/make_golden_n4.py
            raise NotImplementedError(
                f"gaze.method={gaze_method} / dropout.method={dp_method}: this path raises inside the reference "
                "itself (Contrastive feeds the encoder an empty channel slice of the heat-map stack, "
                "train_bc.py:173-176; Oreo rearranges a 2-D index tensor with a 3-D einops pattern, train_bc.py:280)")
        if gaze_method not in ("None", "Reg", "Teacher", "Mask", "ViSaRL", "AGIL", "GRIL"):
            raise NotImplementedError(f"gaze.method={gaze_method} is not a method of the reference "
                                      "(None, Teacher, Reg, Mask, ViSaRL, AGIL, GRIL)")
        if gaze_method in ("Mask", "ViSaRL", "AGIL") and not grayscale:
            # xx [B, S*3, H, W] * gg [B, S, H, W] / cat to 2x channels: the reference's own shapes disagree for RGB
            raise NotImplementedError(f"gaze.method={gaze_method} needs model.grayscale=true (train_bc.py:261,264,270 "
                                      "multiply / concatenate the [B,S,H,W] heat maps with the frame stack)")
        if dp_method not in ("None", "GMD", "IGMD"):
            raise NotImplementedError(f"dropout.method={dp_method} is outside the B200 hot path "
                                      "(supported: None, GMD, IGMD)")
        _lib.init()
        self.spec, self.arena, self.gaze_pre = spec, arena, gaze_pre
        self.use_reg = gaze_method in ("Reg", "Teacher")
        self.gaze_method = gaze_method
        # how the heat maps enter conv_1's input (csrc/conv1_s4d.cu): 1 = multiplied in (Mask), 2 = concatenated (ViSaRL)
        self.in_mode = {"Mask": 1, "ViSaRL": 2}.get(gaze_method, 0)
        self.use_agil, self.use_gril = gaze_method == "AGIL", gaze_method == "GRIL"
        want_c0 = frame_stack * (1 if grayscale else 3) * (2 if gaze_method == "ViSaRL" else 1)
        if spec.C0 != want_c0:
            raise ValueError(f"PolicySpec has {spec.C0} input channels, gaze.method={gaze_method} needs {want_c0} "
                             "(train_bc.py:60-61)")
        self.needs_heat = self.use_reg or dp_method in ("GMD", "IGMD") or self.in_mode != 0 or self.use_agil
        self.needs_sel = self.needs_heat or self.use_gril
        self.use_gmd = dp_method == "GMD"
        self.use_igmd = dp_method == "IGMD"   # in-encoder dropout, linear_models.py:194-199
        self.grayscale, self.S = grayscale, frame_stack
        self.beta, self.lam, self.aggregate = float(beta), float(lam), aggregate
        self.grad_accum = max(1, int(grad_accum))
        self.device = arena.device
        self.B = 0
        self.buf: Dict[str, torch.Tensor] = {}
        # GMD / IGMD uniform draws: a Philox stream inside the mask kernel, (seed, step) kept in device
        # memory so that the draw is part of the captured step (gaze_utils.py:36 draws with torch.rand)
        self.rng_state = torch.tensor([int(seed) & 0x7FFFFFFFFFFFFFFF, 0], dtype=torch.int64, device=self.device)
        self.injected_A: Optional[torch.Tensor] = None  # parity tests inject GMD's uniform draw
        self.injected_A_igmd = None                     # ... and IGMD's pair ([B,H1*W1], [B,H2*W2])
        if prob_dist_type not in ops.DIST_KINDS:
            raise ValueError(f"Invalid prob_dist_type: {prob_dist_type}")   # train_bc.py:171
        self.prob_dist_type = prob_dist_type
        self.gaze_ratio = float(gaze_ratio)
        self._accum_count = 0
        # The gaze branch (heat-map splat -> regulariser -> loss scalars) only meets the main chain at
        # GMD's mask and at conv_6's gradient, so it runs on a side stream: the issue-bound splat
        # overlaps the HBM-bound patch kernel / conv_1, the regulariser overlaps the small-grid heads
        # GEMMs (forward and backward), which leave most of the SMs idle.
        self.overlap = os.environ.get("GABRIL_NO_OVERLAP", "0") != "1"
        self._side = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None
        self._ev = {k: torch.cuda.Event() for k in ("in", "heat", "z", "head", "reg")} if self._side else {}
        self._reg_pending = False
        self.use_graph = os.environ.get("GABRIL_GRAPH", "0") == "1"      # train_step() only
        self._graphs, self._warm_B, self._warm_store_B = {}, 0, 0
        self._own_in, self._own_inputs = {}, False
        self._colsum_pending, self._cs_i = False, 0
        self.fused_dbias = os.environ.get("GABRIL_FUSED_DBIAS", "1") != "0"   # A/B switch: 0 = separate colsum passes
        self.conv6_pairs = os.environ.get("GABRIL_CONV6_PAIRS", "1") != "0"   # A/B switch: 0 = plain conv_6 forward
        self.fused_res = os.environ.get("GABRIL_FUSED_RES", "1") != "0"       # A/B switch: 0 = two launches per block
        self._cs_ev = [torch.cuda.Event() for _ in range(8)] if self._side else []
        if self._side:
            self._ev["colsum"] = torch.cuda.Event()
        self.on_heads_grads_ready = None   # hooks for the data-parallel bucket all-reduce
        self.on_all_grads_ready = None

    # ------------------------------------------------------------------------------------------
    def _alloc(self, B):
        s, dev = self.spec, self.device
        bf, f32 = torch.bfloat16, torch.float32

        def act(h, w, c):
            return torch.empty(B, h, w, c, device=dev, dtype=bf)
        # conv_1 / conv_2 live on a space-to-depth grid: a1s[B,SH,SW,(qh,qw,C1)] holds conv_1 output
        # pixel (2*sh+qh-1, 2*sw+qw-1) (a one-pixel zero border included), so that the k4/s2/p1
        # conv_2 is a 2x2 stride-1 flat-slab convolution; a2p / da2p are conv_2's output and its
        # gradient with a one-pixel zero border on the same [SH, SW] grid (conv_3 reads them with pad 0).
        self.SH, self.SW = s.H2 + 2, s.W2 + 2
        assert s.H1 % 2 == 0 and s.W1 % 2 == 0 and (4 * s.C1) % 64 == 0 and s.C2 % 64 == 0
        assert s.H % 4 == 0 and s.W % 4 == 0 and self.SH == s.H // 4 + 2 and self.SW == s.W // 4 + 2
        zact = lambda h, w, c: torch.zeros(B, h, w, c, device=dev, dtype=bf)  # noqa: E731

        def encoder_bufs(k1p):
            """Activations, sign words and gradient tensors of one encoder pass."""
            e = {}
            # conv_1 input: the frames as a 4x4 space-to-depth bf16 image on the [SH, SW] grid (csrc/conv1_s4d.cu),
            # one spare zero cell at the end (rows of the overlapping-row tensor map span two cells)
            e["x4"] = torch.zeros(B * self.SH * self.SW + 1, k1p, device=dev, dtype=bf)
            e["w4"] = torch.zeros(4 * s.C1, 2, 2 * k1p, device=dev, dtype=bf)
            e["dw4"] = torch.zeros(4 * s.C1, 2, 2 * k1p, device=dev, dtype=f32)
            e["a1s"], e["a2p"] = zact(self.SH, self.SW, 4 * s.C1), zact(self.SH, self.SW, s.C2)
            e["a3"] = act(s.H3, s.W3, s.NH)
            e["r-1"] = act(s.h, s.w, s.NH)  # relu(conv_4): the input of the first Residual
            for i in range(s.NR):
                e[f"h{i}"] = act(s.h, s.w, s.RH)
                e[f"r{i}"] = act(s.h, s.w, s.NH)
            e["a5"], e["z"] = act(s.h, s.w, s.NH), act(s.h, s.w, s.E)
            # conv_6's weights packed for the pixel-pair forward (ops.conv_fwd_pairs), rebuilt from the bf16 shadow per step
            e["w6p"] = torch.zeros(2 * s.E, 5 * 6, s.NH, device=dev, dtype=bf)
            # packed ReLU sign words of every rectified activation (1 bit per element, written by the
            # forward epilogues, read by the dgrad epilogues in place of the activation itself); zeroed:
            # words of never-stored border pixels must read as "not positive"
            for name in ["a1s", "a2p", "a3", "a5"] + [f"r{i}" for i in range(-1, s.NR)] + [f"h{i}" for i in range(s.NR)]:
                e["sb_" + name] = ops.sign_bits_like(e[name])
            # gradients (same shapes as the activations they belong to)
            e["dz"] = act(s.h, s.w, s.E)
            e["da5"] = act(s.h, s.w, s.NH)
            e["dt_a"], e["dt_b"] = act(s.h, s.w, s.NH), act(s.h, s.w, s.NH)
            e["dh"] = act(s.h, s.w, s.RH)
            e["da3"] = act(s.H3, s.W3, s.NH)
            e["da2p"], e["da1s"] = zact(self.SH, self.SW, s.C2), zact(self.SH, self.SW, 4 * s.C1)
            return e
        b = encoder_bufs(s.K1p)
        # gaze.method=AGIL: a second encoder pass with its own weights (prefix "agil") and buffers
        self.buf2 = encoder_bufs(s.k1p_of(s.C0a)) if self.use_agil else None
        b["heat"] = torch.empty(B, self.S, s.H, s.W, device=dev, dtype=f32)
        b["rowmask"] = torch.empty(B, s.n, device=dev, dtype=f32)
        b["sel"] = torch.empty(B, device=dev, dtype=f32)
        b["sel_stats"] = torch.empty(2, device=dev, dtype=f32)
        b["K"] = torch.empty(B, s.n, device=dev, dtype=f32)
        if self.use_igmd:
            # masks in the storage order of the activations they multiply (cells of border pixels
            # are never read for a stored value: those rows are zeroed by the conv epilogues)
            b["rm1"] = torch.zeros(B * self.SH * self.SW * 4, device=dev, dtype=f32)
            b["rm2"] = torch.zeros(B * self.SH * self.SW, device=dev, dtype=f32)
            b["K1"] = torch.empty(B, s.H1 * s.W1, device=dev, dtype=f32)
            b["K2"] = torch.empty(B, s.H2 * s.W2, device=dev, dtype=f32)
        b["mnmx"] = torch.empty(2, device=dev, dtype=torch.int32)
        # split-K for the pre_actor GEMM: enough CTAs to cover the machine; every split stores its partial
        # product into its own plane and bias_act_cast sums the planes in order (deterministic forward)
        m_tiles = (B + 127) // 128
        n_tiles = s.Z // 128 if s.Z >= 128 else 1
        self.pre_splits = max(1, min(s.flat // 64, (2 * 148) // max(1, m_tiles * n_tiles)))
        self.pre_planes = ops.conv_fwd_planes(s.flat, 1, 1, self.pre_splits)
        b["zf_raw"] = torch.empty(self.pre_planes, B, s.Z, device=dev, dtype=f32)
        b["zf"] = torch.empty(B, s.Z, device=dev, dtype=bf)
        b["act1"] = torch.empty(B, s.Z, device=dev, dtype=bf)
        b["lg_raw"] = torch.empty(B, s.AP, device=dev, dtype=f32)
        b["logits"] = torch.empty(B, s.A, device=dev, dtype=f32)
        b["dlog"] = torch.empty(B, s.AP, device=dev, dtype=bf)
        b["actor_part"] = torch.empty(B, device=dev, dtype=f32)
        b["reg_part"] = torch.zeros(B, device=dev, dtype=f32)
        b["losses"] = torch.zeros(3, device=dev, dtype=f32)
        b["stats"] = torch.empty(B, 4, device=dev, dtype=f32)
        # gradients (same shapes as the activations they belong to)
        b["dact1"] = torch.empty(B, s.Z, device=dev, dtype=bf)
        b["dzf"] = torch.empty(B, s.Z, device=dev, dtype=bf)
        b["dz_reg"] = act(s.h, s.w, s.E)
        if self.use_gril:
            b["g_h"] = torch.empty(B, s.Z, device=dev, dtype=bf)
            b["g_raw"] = torch.empty(B, s.AP, device=dev, dtype=f32)
            b["g_dpred"] = torch.empty(B, s.AP, device=dev, dtype=bf)
            b["g_dh"] = torch.empty(B, s.Z, device=dev, dtype=bf)
            b["g_dzf"] = torch.empty(B, s.Z, device=dev, dtype=bf)
        self.buf, self.B = b, B

    # ------------------------------------------------------------------------------------------
    def zero_grad(self):
        self.arena.g.zero_()
        self._accum_count = 0

    def forward(self, frames: torch.Tensor, gaze: torch.Tensor, actions: torch.Tensor):
        """frames u8 [B,L,H,W,3], gaze f32 [B,L,P*2], actions f32 [B,L,A] or [B,A] (all on device).
        Runs the forward pass, the losses and the loss-side gradient seeds.  Returns the device
        tensor losses = [actor, reg, total]."""
        B, L = frames.shape[0], frames.shape[1]
        if B != self.B:
            self._alloc(B)
        b = self.buf
        center = L - 1
        # ---- preprocessing (prepare_for_bc, data_loader_robomimic.py:318-360) ----
        # ---- which samples use gaze (train_bc.py:230-240); all of them at gaze.ratio >= 1 ----
        self.sel = self.sel_stats = None
        if self.gaze_ratio < 1.0 and self.needs_sel:
            self.sel, self.sel_stats = ops.ivg_select(frames, self.S, center, self.grayscale,
                                                      self.gaze_ratio, b["sel"], b["sel_stats"])
        self._heat_done = False
        if self.in_mode or self.use_agil:
            # the heat maps are part of conv_1's input (Mask / ViSaRL / AGIL's second encoder): splat first
            self.gaze_pre.splat_into(gaze, self.S, center, b["heat"], aggregate=self.aggregate)
            self._heat_done = True
        if self.in_mode:
            ops.obs_s4d_gaze(frames, None, b["heat"], self.sel, self.in_mode, self.S, center, self.SH, self.SW, b["x4"])
        else:
            ops.obs_s4d(frames, self.S, center, self.grayscale, self.SH, self.SW, b["x4"])
        if self.use_agil:
            ops.obs_s4d_gaze(frames, None, b["heat"], self.sel, 1, self.S, center, self.SH, self.SW, self.buf2["x4"])
        return self._forward_rest(gaze, actions, center)

    def forward_store(self, store, step_idx: torch.Tensor):
        """Device-resident feed: `store` is a DemoStore living in HBM, step_idx int64
        [B] the sampled steps.  Replaces SequenceDataset.get_item + DataLoader + H2D
        (robomimic/utils/dataset.py:503-633, train_bc.py:204-207): the frame window table is built on
        the device and conv_1's patch kernel, the content hash and the gaze / action gathers read the
        store through it, so the uint8 batch [B,S,H,W,3] never exists.  Same results, bit for bit, as
        forward(*store.batch(step_idx))."""
        if store.images.device != self.device or store.S != self.S:
            raise RuntimeError("forward_store needs a DemoStore on the engine's device with the same "
                               f"frame_stack (store on {store.images.device}, S={store.S})")
        B = step_idx.numel()
        if B != self.B:
            self._alloc(B)
        b = self.buf
        if "widx" not in b:
            b["widx"] = torch.empty(B, self.S, device=self.device, dtype=torch.int64)
        center = self.S - 1
        widx = ops.store_windows(step_idx.to(self.device, torch.int64).contiguous(), store.first, self.S,
                                 out=b["widx"])
        self.sel = self.sel_stats = None
        if self.gaze_ratio < 1.0 and self.needs_sel:
            self.sel, self.sel_stats = ops.ivg_select_store(store.images, widx, self.grayscale,
                                                            self.gaze_ratio, b["sel"], b["sel_stats"])
        gaze = ops.store_gather_rows(store.gaze, widx)
        actions = ops.store_gather_rows(store.actions, widx)
        self._heat_done = False
        if self.in_mode or self.use_agil:
            self.gaze_pre.splat_into(gaze, self.S, center, b["heat"], aggregate=self.aggregate)
            self._heat_done = True
        if self.in_mode:
            ops.obs_s4d_gaze(store.images, widx, b["heat"], self.sel, self.in_mode, self.S, center, self.SH, self.SW, b["x4"])
        else:
            ops.obs_s4d_store(store.images, widx, self.grayscale, self.SH, self.SW, b["x4"])
        if self.use_agil:
            ops.obs_s4d_gaze(store.images, widx, b["heat"], self.sel, 1, self.S, center, self.SH, self.SW, self.buf2["x4"])
        return self._forward_rest(gaze, actions, center)

    def predict(self, frames: torch.Tensor, g: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Inference: BCAgent._process_observation's formatting + _predict_control
        (eval/my_agents/bc_agent.py:228-305).  frames u8 [B,S,H,W,3]; g fp32 [B,C,H,W] = the gaze
        predictor's map (only its last channel is used, clipped into [0,1] like :278-279), needed for
        dropout.method GMD / IGMD, which run in test mode: z * K instead of z * (A < K)
        (gaze_utils.py:49-51, linear_models.py:194-199 with the encoder in eval()).  Forward kernels
        only; gradients, losses and Adam state are not touched.  Returns the action vector
        fp32 [B, action_dim] (a view of an engine buffer, overwritten by the next call)."""
        s, a = self.spec, self.arena
        B, L = frames.shape[0], frames.shape[1]
        if B != self.B:
            self._alloc(B)
        b = self.buf
        if self.in_mode or self.use_agil:
            # bc_agent.py:274-289: the gaze predictor's clipped map takes the heat maps' place (Mask / AGIL: broadcast
            # over the stack; ViSaRL concatenates it, which only fits the trained 2S-channel conv_1 if g has S channels)
            if g is None:
                raise RuntimeError(f"gaze.method={self.gaze_method} needs the predicted gaze map g at inference "
                                   "(bc_agent.py:274-277)")
            gh = g.float().clamp(0.0, 1.0)
            if self.in_mode == 2 and gh.shape[1] != self.S:
                raise RuntimeError(f"ViSaRL: cat([obs, g]) has {self.S + gh.shape[1]} channels, the encoder {s.C0} "
                                   "(bc_agent.py:281 has the same mismatch)")
            gh = gh.expand(B, self.S, s.H, s.W).contiguous()
            if self.in_mode:
                ops.obs_s4d_gaze(frames, None, gh, None, self.in_mode, self.S, L - 1, self.SH, self.SW, b["x4"])
            else:
                ops.obs_s4d(frames, self.S, L - 1, self.grayscale, self.SH, self.SW, b["x4"])
            if self.use_agil:
                ops.obs_s4d_gaze(frames, None, gh, None, 1, self.S, L - 1, self.SH, self.SW, self.buf2["x4"])
        else:
            ops.obs_s4d(frames, self.S, L - 1, self.grayscale, self.SH, self.SW, b["x4"])
        rowmask = igmd = None
        if self.use_gmd or self.use_igmd:
            if g is None:
                raise RuntimeError(f"dropout.method={'GMD' if self.use_gmd else 'IGMD'} needs the "
                                   "predicted gaze map g at inference (bc_agent.py:276-279)")
            gl = g[:, -1:].float().clamp(0.0, 1.0).contiguous()
            dummy = b["K"]   # A is not read in test mode
            if self.use_gmd:
                rowmask = ops.gmd_mask(gl, dummy, s.h, s.w, 0.7, True, rowmask=b["rowmask"],
                                       ws=(b["K"], b["mnmx"]))
            else:
                ops.gmd_mask(gl, b["K1"], s.H1, s.W1, 0.7, True, rowmask=b["rm1"], ws=(b["K1"], b["mnmx"]),
                             layout="s2d", grid=(self.SH, self.SW))
                ops.gmd_mask(gl, b["K2"], s.H2, s.W2, 0.7, True, rowmask=b["rm2"], ws=(b["K2"], b["mnmx"]),
                             layout="bordered", grid=(self.SH, self.SW))
                igmd = (b["rm1"], b["rm2"])
        if self.use_agil:        # z = (z + encoder_agil(obs * g)) / 2, then GMD's test-mode scaling (bc_agent.py:291-295)
            self._encoder_fwd(None, igmd)
            self._encoder_fwd(None, None, pfx="agil", b=self.buf2, c0=s.C0a)
            ops.agil_combine(b["z"], self.buf2["z"], None, rowmask.view(-1) if rowmask is not None else None)
        else:
            self._encoder_fwd(rowmask, igmd)
        self._heads_logits()
        ops.mse_head(b["lg_raw"], a.P("act.2.b"), b["logits"], s.A, b["actor_part"], 0.0,
                     logits_out=b["logits"])
        return b["logits"]

    def _forward_rest(self, gaze, actions, center):
        s, a, b, B = self.spec, self.arena, self.buf, self.B
        if self.needs_heat and not self._heat_done:
            if self.overlap:
                main = torch.cuda.current_stream(self.device)
                self._ev["in"].record(main)
                if not torch.cuda.is_current_stream_capturing():
                    gaze.record_stream(self._side)
                with torch.cuda.stream(self._side):
                    self._side.wait_event(self._ev["in"])
                    self.gaze_pre.splat_into(gaze, self.S, center, b["heat"], aggregate=self.aggregate)
                    self._ev["heat"].record(self._side)
                if self.use_gmd or self.use_igmd:
                    main.wait_event(self._ev["heat"])      # the keep maps read the heat maps
            else:
                self.gaze_pre.splat_into(gaze, self.S, center, b["heat"], aggregate=self.aggregate)
        rowmask = None
        if self.use_gmd:
            # A ~ U[0,1) (gaze_utils.py:36): injected by the parity tests, else drawn inside the kernel
            A = self.injected_A.reshape(B, s.n) if self.injected_A is not None else None
            ops.gmd_mask(b["heat"], A, s.h, s.w, 0.7, False, rowmask=b["rowmask"],
                         ws=(b["K"], b["mnmx"]), sel=self.sel, rng_state=self.rng_state, draw_id=0)
            rowmask = b["rowmask"]
        igmd = None
        if self.use_igmd:
            # apply_gmd_dropout after conv_1 + ReLU (90x160) and conv_2 + ReLU (45x80): the same keep
            # map as GMD, interpolated to each size, its own uniform draw per layer
            A1 = A2 = None
            if self.injected_A_igmd is not None:
                A1, A2 = (t.reshape(B, -1) for t in self.injected_A_igmd)
            ops.gmd_mask(b["heat"], A1, s.H1, s.W1, 0.7, False, rowmask=b["rm1"], ws=(b["K1"], b["mnmx"]),
                         sel=self.sel, layout="s2d", grid=(self.SH, self.SW), rng_state=self.rng_state, draw_id=1)
            ops.gmd_mask(b["heat"], A2, s.H2, s.W2, 0.7, False, rowmask=b["rm2"], ws=(b["K2"], b["mnmx"]),
                         sel=self.sel, layout="bordered", grid=(self.SH, self.SW), rng_state=self.rng_state,
                         draw_id=2)
            igmd = (b["rm1"], b["rm2"])
        if self.use_gmd or self.use_igmd:
            ops.rng_advance(self.rng_state)     # next step draws a fresh stream
        if self.use_agil:
            # second encoder on obs * gg_for_mul; z = 0.5 (z + z_agil) for the samples that use gaze; GMD's
            # mask (applied to the combined z in the reference, train_bc.py:273-274) rides on the combine
            self._encoder_fwd(None, igmd)
            self._encoder_fwd(None, None, pfx="agil", b=self.buf2, c0=s.C0a)
            ops.agil_combine(b["z"], self.buf2["z"], self.sel, rowmask.view(-1) if rowmask is not None else None)
        else:
            self._encoder_fwd(rowmask, igmd)
        self._gaze_seq, self._center = gaze, center      # GRIL's targets: the raw coordinates at the centre step
        self._heads_fwd(actions, center)
        return b["losses"]

    def _encoder_fwd(self, rowmask=None, igmd=None, pfx="enc", b=None, c0=None):
        """Encoder.forward (linear_models.py:191-207) from the patch matrix to z.  rowmask: GMD mask
        fused into conv_6's epilogue; igmd = (mask after conv_1, mask after conv_2): the stored a1 / a2
        are the dropped-out activations, so their `> 0` gates carry the mask through the backward."""
        s, a, B = self.spec, self.arena, self.B
        b = self.buf if b is None else b
        c0 = s.C0 if c0 is None else c0
        # conv_1 = a 2x2 flat convolution over the 4x4 space-to-depth frame image; its zero-inflated
        # weights are re-derived from the fp32 master every step (16 K elements)
        ops.conv1_s4d_expand(a.P(f"{pfx}.conv1.w"), b["w4"], c0)
        ops.conv1_s4d_fwd(b["x4"], b["w4"], a.P(f"{pfx}.conv1.b"), b["a1s"], B, self.SH, self.SW, s.H1, s.W1,
                          relu=True, rowmask=igmd[0] if igmd else None, bits_out=b["sb_a1s"])
        ops.conv_flat(b["a1s"], a.WB(f"{pfx}.conv2.w"), b["a2p"], 2, 2, bias=a.P(f"{pfx}.conv2.b"),
                      valid=(s.H2, s.W2), out_geom=(self.SH, self.SW, 1, 1), relu=True,
                      rowmask=igmd[1] if igmd else None, bits_out=b["sb_a2p"])
        ops.conv_fwd(b["a2p"], a.WB(f"{pfx}.conv3.w"), a.P(f"{pfx}.conv3.b"), b["a3"], 4, 4, 2, 0, relu=True,
                     bits_out=b["sb_a3"])
        # conv_4 has no activation of its own; the first Residual's in-place ReLU rectifies it (F7)
        ops.conv_fwd(b["a3"], a.WB(f"{pfx}.conv4.w"), a.P(f"{pfx}.conv4.b"), b["r-1"], 3, 3, 1, 0, relu=True,
                     bits_out=b["sb_r-1"])
        fused_res = self.fused_res and s.NH == 128 and s.RH == 32
        for i in range(s.NR):
            prev = b[f"r{i - 1}"]
            if fused_res:   # both convolutions of the block in one launch: h never round-trips through HBM
                ops.res_block_fwd(prev, a.WB(f"{pfx}.res{i}.c3.w"), a.WB(f"{pfx}.res{i}.c1.w"), b[f"h{i}"],
                                  b[f"sb_h{i}"], b[f"r{i}"], b[f"sb_r{i}"])
                continue
            ops.conv_fwd(prev, a.WB(f"{pfx}.res{i}.c3.w"), None, b[f"h{i}"], 3, 3, 1, 1, relu=True,
                         bits_out=b[f"sb_h{i}"])
            # relu(x) + block(x); the next consumer (next Residual / the stack's final ReLU)
            # rectifies it again, so the rectified value is what gets stored
            ops.conv_fwd(b[f"h{i}"], a.WB(f"{pfx}.res{i}.c1.w"), None, b[f"r{i}"], 1, 1, 1, 0, relu=True,
                         resid=prev, bits_out=b[f"sb_r{i}"])
        last = b[f"r{s.NR - 1}"]
        ops.conv_fwd(last, a.WB(f"{pfx}.conv5.w"), a.P(f"{pfx}.conv5.b"), b["a5"], 5, 5, 1, 2, relu=True,
                     bits_out=b["sb_a5"])
        if s.E in (32, 64) and s.NH % 64 == 0 and s.w % 2 == 0 and self.conv6_pairs:
            # narrow layer (N = 64): two adjacent output pixels stacked along N, bit-identical to conv_fwd
            ops.conv_pack_wpairs(a.WB(f"{pfx}.conv6.w"), b["w6p"], 5, 5)
            ops.conv_fwd_pairs(b["a5"], b["w6p"], a.P(f"{pfx}.conv6.b"), b["z"], 5, 5, 2, relu=False,
                               rowmask=rowmask.view(-1) if rowmask is not None else None)
        else:
            ops.conv_fwd(b["a5"], a.WB(f"{pfx}.conv6.w"), a.P(f"{pfx}.conv6.b"), b["z"], 5, 5, 1, 2, relu=False,
                         rowmask=rowmask.view(-1) if rowmask is not None else None)

    def _heads_fwd(self, actions, center):
        s, a, b, B = self.spec, self.arena, self.buf, self.B
        side_reg = self.overlap and self.use_reg
        scale = 1.0 / self.grad_accum
        if side_reg:
            main = torch.cuda.current_stream(self.device)
            self._ev["z"].record(main)                      # z (and sel) are complete on the main stream
            with torch.cuda.stream(self._side):
                self._side.wait_event(self._ev["z"])
                reg_scale = self._reg(scale)
        self._heads_logits()
        if actions.dim() == 3:
            act_view = actions[:, center if center < actions.shape[1] else -1, :]
        else:
            act_view = actions
        assert act_view.stride(-1) == 1 and act_view.dtype == torch.float32
        ops.mse_head(b["lg_raw"], a.P("act.2.b"), act_view, s.A, b["actor_part"],
                     scale / (B * s.A), dlogits=b["dlog"], dbias=a.G("act.2.b"), logits_out=b["logits"])
        if side_reg:
            self._ev["head"].record(main)
            with torch.cuda.stream(self._side):
                self._side.wait_event(self._ev["head"])     # actor_part
                self._finalize(reg_scale)
                self._ev["reg"].record(self._side)
            self._reg_pending = True                        # joined in backward() / losses_host()
            return b["losses"]
        return self._finalize(self._reg(scale))

    def _join_reg(self):
        if self._reg_pending:
            torch.cuda.current_stream(self.device).wait_event(self._ev["reg"])
            self._reg_pending = False

    def _heads_logits(self):
        """pre_actor + actor up to the raw (bias-free) logits (train_bc.py:79-82,290-291)."""
        s, a, b, B = self.spec, self.arena, self.buf, self.B
        ops.conv_fwd(b["z"].view(B, 1, 1, s.flat), a.WB("pre.w"), None, b["zf_raw"], 1, 1, 1, 0,
                     out_f32_atomic=2, splits=self.pre_splits)
        ops.bias_act_cast(b["zf_raw"], a.P("pre.b"), b["zf"], relu=False)
        ops.conv_fwd(b["zf"].view(B, 1, 1, s.Z), a.WB("act.0.w"), a.P("act.0.b"), b["act1"], 1, 1, 1, 0,
                     relu=True)
        b["lg_raw"].zero_()
        ops.conv_fwd(b["act1"].view(B, 1, 1, s.Z), a.WB("act.2.w"), None, b["lg_raw"], 1, 1, 1, 0,
                     out_f32_atomic=True, splits=1)

    def _reg(self, scale):
        s, a, b, B = self.spec, self.arena, self.buf, self.B
        # ---- regulariser (train_bc.py:145-150,168-169) on the (post-dropout) z ----
        reg_scale = 0.0
        if self.use_gril:
            # gaze-coordinate head on z_flat (train_bc.py:84-86,182-192): Linear -> ReLU -> Linear, clamped MSE
            # against the raw gaze coordinates of the centre step over the samples that use gaze
            ops.conv_fwd(b["zf"].view(B, 1, 1, s.Z), a.WB("gril.0.w"), a.P("gril.0.b"), b["g_h"], 1, 1, 1, 0, relu=True)
            b["g_raw"].zero_()
            ops.conv_fwd(b["g_h"].view(B, 1, 1, s.Z), a.WB("gril.2.w"), None, b["g_raw"], 1, 1, 1, 0,
                         out_f32_atomic=True, splits=1)
            gz = self._gaze_seq
            tgt = gz[:, self._center] if gz.dim() >= 3 else gz
            assert tgt.stride(-1) == 1 and tgt.dtype == torch.float32 and tgt.shape[-1] >= 2 * s.GP
            ops.gril_head(b["g_raw"], a.P("gril.2.b"), tgt, 2 * s.GP, self.sel, b["g_dpred"], a.G("gril.2.b"),
                          b["reg_part"], self.lam * scale)
            return 1.0
        if self.use_reg:
            # mean over the selected samples (and, for MSE, their pixels); with gaze.ratio < 1 the
            # number selected is a device scalar: sel_stats[1] = B / n_selected rescales both uses
            reg_scale = 1.0 / (B * s.H * s.W) if self.prob_dist_type == "MSE" else 1.0 / B
            g1 = b["heat"][:, self.S - 1]
            sdev = self.sel_stats[1:] if self.sel_stats is not None else None
            ops.reg_dist(b["z"], g1, b["heat"].stride(0), s.H, s.W, self.beta,
                         self.lam * scale * reg_scale, b["dz_reg"], b["reg_part"],
                         kind=self.prob_dist_type, sel=self.sel, scale_dev=sdev, stats=b["stats"])
        return reg_scale

    def _finalize(self, reg_scale):
        s, b, B = self.spec, self.buf, self.B
        ops.finalize_losses(b["actor_part"], b["reg_part"] if (self.use_reg or self.use_gril) else None, B,
                            1.0 / (B * s.A), reg_scale, self.lam, b["losses"],
                            reg_scale_dev=(self.sel_stats[1:] if self.use_reg and self.sel_stats is not None
                                           else None))
        return b["losses"]

    # ------------------------------------------------------------------------------------------
    def backward(self):
        """The rest of loss.backward(): every weight / bias gradient accumulated into arena.g."""
        # the bucket hooks fire only on the micro-step that closes the accumulation window (DDP's
        # no_sync semantics); the engine decides, the hooks themselves are unconditional
        closing = self._accum_count + 1 >= self.grad_accum
        self._backward_heads()
        if closing and self.on_heads_grads_ready is not None:
            self.on_heads_grads_ready()
        self._backward_encoders()
        self._accum_count += 1
        if closing and self.on_all_grads_ready is not None:
            self.on_all_grads_ready()

    def _backward_heads(self):
        """actor / pre_actor backward down to dz (+ the regulariser's share and GMD's mask)."""
        s, a, b = self.spec, self.arena, self.buf
        B = self.B
        direct = self.grad_accum == 1  # pre_actor dW: plain stores when nothing accumulates
        rowmask = b["rowmask"].view(-1) if self.use_gmd else None
        # actor.2
        ops.conv_wgrad(b["dlog"].view(B, 1, 1, s.AP), b["act1"].view(B, 1, 1, s.Z), a.G("act.2.w"), 1, 1, 1, 0)
        ops.conv_dgrad(b["dlog"].view(B, 1, 1, s.AP), a.WB("act.2.w"), b["dact1"].view(B, 1, 1, s.Z),
                       1, 1, 1, 0, gate=b["act1"].view(B, 1, 1, s.Z))
        # actor.0
        fb = self.fused_dbias
        if not fb:
            ops.colsum_bf16(b["dact1"], a.G("act.0.b"))
        ops.conv_wgrad(b["dact1"].view(B, 1, 1, s.Z), b["zf"].view(B, 1, 1, s.Z), a.G("act.0.w"), 1, 1, 1, 0,
                       dbias=a.G("act.0.b") if fb else None)
        zf_addend = None
        if self.use_gril:   # the coordinate head's share of d z_flat
            ops.conv_wgrad(b["g_dpred"].view(B, 1, 1, s.AP), b["g_h"].view(B, 1, 1, s.Z), a.G("gril.2.w"), 1, 1, 1, 0)
            ops.conv_dgrad(b["g_dpred"].view(B, 1, 1, s.AP), a.WB("gril.2.w"), b["g_dh"].view(B, 1, 1, s.Z),
                           1, 1, 1, 0, gate=b["g_h"].view(B, 1, 1, s.Z))
            ops.colsum_bf16(b["g_dh"], a.G("gril.0.b"))
            ops.conv_wgrad(b["g_dh"].view(B, 1, 1, s.Z), b["zf"].view(B, 1, 1, s.Z), a.G("gril.0.w"), 1, 1, 1, 0)
            ops.conv_dgrad(b["g_dh"].view(B, 1, 1, s.Z), a.WB("gril.0.w"), b["g_dzf"].view(B, 1, 1, s.Z), 1, 1, 1, 0)
            zf_addend = b["g_dzf"].view(B, 1, 1, s.Z)
        ops.conv_dgrad(b["dact1"].view(B, 1, 1, s.Z), a.WB("act.0.w"), b["dzf"].view(B, 1, 1, s.Z), 1, 1, 1, 0,
                       addend=zf_addend)
        # pre_actor
        self._join_reg()                                    # dz_reg, losses (side stream)
        if not fb:
            ops.colsum_bf16(b["dzf"], a.G("pre.b"))
        ops.conv_wgrad(b["dzf"].view(B, 1, 1, s.Z), b["z"].view(B, 1, 1, s.flat), a.G("pre.w"), 1, 1, 1, 0,
                       direct_store=direct, dbias=a.G("pre.b") if fb else None)
        ops.conv_dgrad(b["dzf"].view(B, 1, 1, s.Z), a.WB("pre.w"), b["dz"].view(B, 1, 1, s.flat), 1, 1, 1, 0,
                       addend=b["dz_reg"].view(B, 1, 1, s.flat) if self.use_reg else None)
        if rowmask is not None:
            # z' = z * M  =>  dz = dz' * M (the regulariser part is already zero where z' == 0)
            self._mask_rows(b["dz"], rowmask)

    # ------------------------------------------------------------------------------------------
    GRAPH_KEYS = 4   # pointer-keyed graphs kept before the engine switches to its own input buffers

    def train_step(self, frames: torch.Tensor, gaze: torch.Tensor, actions: torch.Tensor):
        """zero_grad + forward + backward.  With use_graph (and no gradient accumulation over
        micro-batches) the step is two CUDA-graph replays, split at the point where the heads'
        gradients are complete so that the data-parallel bucket hooks still fire between them: ~60
        kernel launches become 2, which is what a 128-sample shard (1.6 ms of device time) needs when
        eight ranks share the host's cores.  GMD / IGMD steps are captured too (the uniform draw is a
        Philox stream inside the mask kernel).  Graphs are keyed by the input buffers; a caller that
        hands fresh tensors every step (more than GRAPH_KEYS distinct buffer sets) is served from
        engine-owned input buffers instead (one device-to-device copy, one graph), so the cache is
        bounded.  The first step at a new batch size runs eagerly (allocations), tracing always does."""
        graphable = self.use_graph and self.grad_accum == 1 and ops._TRACE is None
        B = frames.shape[0]
        if not graphable or self._warm_B != B:
            self.zero_grad()
            self.forward(frames, gaze, actions)
            self.backward()
            self._warm_B = B if B == self.B else 0
            return self.buf["losses"]
        key = (frames.data_ptr(), gaze.data_ptr(), actions.data_ptr(), tuple(frames.shape), tuple(actions.shape))
        g = self._graphs.get(key)
        if g is None:
            n_ptr_keys = sum(1 for k in self._graphs if k[0] != "own")
            if n_ptr_keys >= self.GRAPH_KEYS or self._own_inputs:
                # unstable input pointers: stage into our own buffers, one graph per shape
                self._own_inputs = True
                key = ("own", tuple(frames.shape), tuple(gaze.shape), tuple(actions.shape))
                st = self._own_in.get(key)
                if st is None:
                    st = self._own_in[key] = tuple(torch.empty_like(t, memory_format=torch.contiguous_format)
                                                   for t in (frames, gaze, actions))
                for dst, src in zip(st, (frames, gaze, actions)):
                    dst.copy_(src, non_blocking=True)
                frames, gaze, actions = st
                g = self._graphs.get(key)
            if g is None:
                g = self._graphs[key] = self._capture(lambda: self.forward(frames, gaze, actions),
                                                      keep=(frames, gaze, actions))
        return self._replay(g)

    def train_step_store(self, store, step_idx: torch.Tensor):
        """train_step for the device-resident feed (forward_store): the sampled step indices are copied
        into a static buffer, so one graph per (store, batch size) serves every step."""
        graphable = self.use_graph and self.grad_accum == 1 and ops._TRACE is None
        B = step_idx.numel()
        if not graphable or self._warm_store_B != B:
            self.zero_grad()
            self.forward_store(store, step_idx)
            self.backward()
            self._warm_store_B = B if B == self.B else 0
            return self.buf["losses"]
        key = ("store", id(store), B)
        st = self._own_in.get(key)
        if st is None:
            st = self._own_in[key] = torch.empty(B, device=self.device, dtype=torch.int64)
        st.copy_(step_idx.reshape(-1), non_blocking=True)
        g = self._graphs.get(key)
        if g is None:
            g = self._graphs[key] = self._capture(lambda: self.forward_store(store, st), keep=(store, st))
        return self._replay(g)

    def _replay(self, g):
        # graphs exist only for grad_accum == 1: every replayed step closes its window, so both
        # bucket hooks fire on every step (the captured zero_grad never runs on the host)
        g[0].replay()
        if self.on_heads_grads_ready is not None:
            self.on_heads_grads_ready()
        g[1].replay()
        self._accum_count = 1
        if self.on_all_grads_ready is not None:
            self.on_all_grads_ready()
        return self.buf["losses"]

    def _capture(self, forward_fn, keep=()):
        torch.cuda.synchronize(self.device)
        ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(ga, capture_error_mode="thread_local"):
            self.zero_grad()
            forward_fn()
            self._backward_heads()
        with torch.cuda.graph(gb, pool=ga.pool(), capture_error_mode="thread_local"):
            self._backward_encoders()
        return ga, gb, keep   # the graphs hold raw pointers: `keep` pins the captured input tensors

    def _backward_encoders(self):
        """Encoder backward from b["dz"]; gaze.method=AGIL: both encoders, dz split by the where() of the combine."""
        if self.use_agil:
            ops.agil_split(self.buf["dz"], self.buf2["dz"], self.sel)
            self._encoder_bwd()
            self._encoder_bwd(pfx="agil", b=self.buf2, c0=self.spec.C0a)
        else:
            self._encoder_bwd()

    def _encoder_bwd(self, pfx="enc", b=None, c0=None):
        """Backward of the encoder from b["dz"] (gradient w.r.t. z) to every encoder weight gradient."""
        s, a, B = self.spec, self.arena, self.B
        b = self.buf if b is None else b
        c0 = s.C0 if c0 is None else c0
        # conv_6 / conv_5
        last, sb_last = b[f"r{s.NR - 1}"], b[f"sb_r{s.NR - 1}"]
        fb = self.fused_dbias   # bias gradients inside the weight-gradient launches (no separate pass over dY)
        if not fb:
            self._colsum(b["dz"], a.G(f"{pfx}.conv6.b"))
        ops.conv_wgrad(b["dz"], b["a5"], a.G(f"{pfx}.conv6.w"), 5, 5, 1, 2, dbias=a.G(f"{pfx}.conv6.b") if fb else None)
        ops.conv_dgrad(b["dz"], a.WB(f"{pfx}.conv6.w"), b["da5"], 5, 5, 1, 2, gate_bits=b["sb_a5"])
        if not fb:
            self._colsum(b["da5"], a.G(f"{pfx}.conv5.b"))
        ops.conv_wgrad(b["da5"], last, a.G(f"{pfx}.conv5.w"), 5, 5, 1, 2, dbias=a.G(f"{pfx}.conv5.b") if fb else None)
        dt, other = b["dt_a"], b["dt_b"]
        ops.conv_dgrad(b["da5"], a.WB(f"{pfx}.conv5.w"), dt, 5, 5, 1, 2, gate_bits=sb_last)
        # residual stack, last block first
        for i in reversed(range(s.NR)):
            prev = b[f"r{i - 1}"]
            ops.conv_wgrad(dt, b[f"h{i}"], a.G(f"{pfx}.res{i}.c1.w"), 1, 1, 1, 0)
            ops.conv_dgrad(dt, a.WB(f"{pfx}.res{i}.c1.w"), b["dh"], 1, 1, 1, 0, gate_bits=b[f"sb_h{i}"])
            ops.conv_wgrad(b["dh"], prev, a.G(f"{pfx}.res{i}.c3.w"), 3, 3, 1, 1)
            ops.conv_dgrad(b["dh"], a.WB(f"{pfx}.res{i}.c3.w"), other, 3, 3, 1, 1, addend=dt,
                           gate_bits=b[f"sb_r{i - 1}"])
            dt, other = other, dt
        # conv_4 .. conv_1
        if not fb:
            self._colsum(dt, a.G(f"{pfx}.conv4.b"))
        ops.conv_wgrad(dt, b["a3"], a.G(f"{pfx}.conv4.w"), 3, 3, 1, 0, dbias=a.G(f"{pfx}.conv4.b") if fb else None)
        ops.conv_dgrad(dt, a.WB(f"{pfx}.conv4.w"), b["da3"], 3, 3, 1, 0, gate_bits=b["sb_a3"])
        if not fb:
            self._colsum(b["da3"], a.G(f"{pfx}.conv3.b"))
        # conv_3 over the zero-bordered a2p (pad 0): the border gradients are gated to zero by a2p itself
        ops.conv_wgrad(b["da3"], b["a2p"], a.G(f"{pfx}.conv3.w"), 4, 4, 2, 0, dbias=a.G(f"{pfx}.conv3.b") if fb else None)
        ops.conv_dgrad(b["da3"], a.WB(f"{pfx}.conv3.w"), b["da2p"], 4, 4, 2, 0, gate_bits=b["sb_a2p"])
        if not fb:
            self._colsum(b["da2p"], a.G(f"{pfx}.conv2.b"))
        # conv_2 on the space-to-depth grid: dY pixel (i, j) sits at flat row q + SW + 1 (the rows before it are
        # border cells: zero, so the fused column sums miss nothing)
        ops.conv_wgrad_flat(b["da2p"], b["a1s"], a.G(f"{pfx}.conv2.w"), 2, 2, self.SW + 1,
                            flop_pixels=B * s.H2 * s.W2, dbias=a.G(f"{pfx}.conv2.b") if fb else None)
        ops.conv_flat(b["da2p"], a.WB(f"{pfx}.conv2.w"), b["da1s"], 2, 2, gate_bits=b["sb_a1s"], transpose_w=True,
                      flop_pixels=B * s.H2 * s.W2)  # algorithmic FLOPs of the k4/s2 dgrad
        M1 = B * self.SH * self.SW * 4
        if not fb:
            self._colsum(b["da1s"].view(M1, s.C1), a.G(f"{pfx}.conv1.b"))
        b["dw4"].zero_()
        ops.conv1_s4d_wgrad(b["da1s"], b["x4"], b["dw4"], B, self.SH, self.SW, s.H1, s.W1,
                            dbias=a.G(f"{pfx}.conv1.b") if fb else None)
        ops.conv1_s4d_fold(b["dw4"], a.G(f"{pfx}.conv1.w"), c0)
        if self._colsum_pending:
            main = torch.cuda.current_stream(self.device)
            self._ev["colsum"].record(self._side)
            main.wait_event(self._ev["colsum"])             # bias gradients complete before Adam / all-reduce
            self._colsum_pending = False

    def _colsum(self, dy, db):
        """Bias gradient db += column sums of dy.  An HBM-bound reader with no consumer before the
        optimiser: issued on the side stream so that it runs beside the tensor-bound wgrad / dgrad
        GEMMs of the same layer instead of between them."""
        if not self.overlap:
            return ops.colsum_bf16(dy, db)
        ev = self._cs_ev[self._cs_i % len(self._cs_ev)]
        self._cs_i += 1
        ev.record(torch.cuda.current_stream(self.device))   # dy is complete on the main stream
        with torch.cuda.stream(self._side):
            self._side.wait_event(ev)
            ops.colsum_bf16(dy, db)
        self._colsum_pending = True

    def activation(self, name: str) -> torch.Tensor:
        """Stored activation `name` as a plain NHWC tensor (tests / debugging): undoes the
        space-to-depth / zero-border storage of conv_1 and conv_2 outputs."""
        s, b = self.spec, self.buf
        if name == "a1":
            t = b["a1s"].view(self.B, self.SH, self.SW, 2, 2, s.C1).permute(0, 1, 3, 2, 4, 5)
            return t.reshape(self.B, 2 * self.SH, 2 * self.SW, s.C1)[:, 1:1 + s.H1, 1:1 + s.W1].contiguous()
        if name == "a2":
            return b["a2p"][:, 1:1 + s.H2, 1:1 + s.W2].contiguous()
        return b[name]

    def _mask_rows(self, t, rowmask):
        # dz * M as a 1x1 "identity-free" epilogue: reuse the dgrad rowmask path would need a second
        # GEMM; a dedicated vectorised kernel is cheaper
        ops.rowscale_bf16(t, rowmask)

    # ------------------------------------------------------------------------------------------
    def adam_step(self, step: int, lr: float, *, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0,
                  decoupled=False, grad_scale=1.0):
        a = self.arena
        ops.adam(a.p, a.g, a.m, a.v, a.wb, step, lr, beta1, beta2, eps, weight_decay, decoupled, grad_scale)

    def losses_host(self):
        """One device->host read: (actor, reg, total)."""
        self._join_reg()
        v = self.buf["losses"].tolist()
        return v[0], v[1], v[2]


class HostFeeder:
    """Pinned-host -> device staging of one batch on a side stream (the DataLoader(pin_memory) +
    `.to(device, non_blocking=True)` boundary of train_bc.py:204-207), double buffered so the copy
    of batch i+1 overlaps the kernels of batch i."""

    def __init__(self, device, depth: int = 2):
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self.depth = depth
        self.slots = [None] * depth
        self.events = [torch.cuda.Event() for _ in range(depth)]
        self.free_events = [None] * depth
        self.i = 0

    def stage(self, frames: torch.Tensor, gaze: torch.Tensor, actions: torch.Tensor):
        """Start the H2D copies of one host batch; returns a ticket for `wait`."""
        k = self.i % self.depth
        self.i += 1
        if self.slots[k] is None or self.slots[k][0].shape != frames.shape:
            self.slots[k] = tuple(torch.empty(t.shape, dtype=t.dtype, device=self.device)
                                  for t in (frames, gaze, actions))
        with torch.cuda.stream(self.stream):
            if self.free_events[k] is not None:
                self.stream.wait_event(self.free_events[k])  # previous consumer of this slot done
            for dst, src in zip(self.slots[k], (frames, gaze, actions)):
                dst.copy_(src, non_blocking=True)
            self.events[k].record(self.stream)
        return k

    def wait(self, k):
        """Make the compute stream wait for slot k; returns the device tensors."""
        torch.cuda.current_stream(self.device).wait_event(self.events[k])
        return self.slots[k]

    def release(self, k):
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.free_events[k] = ev
