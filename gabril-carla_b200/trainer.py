"""Drop-in trainer surface for the train_bc.py hot path.

`BaseTrainer` mirrors vlm_gaze/train/common/base_trainer.py:18-192 (constructor order, the five
hooks build_models / get_optim_params / set_train_mode / compute_loss / save_for_epoch, the
epoch / gradient-accumulation / scheduler / save-interval loop and the TensorBoard scalar names);
`BCTrainer` mirrors vlm_gaze/train/train_bc.py:20-335 for gaze.method in {None, Reg, Teacher} and
dropout.method in {None, GMD}.  The arithmetic is the fused sm_100a kernel schedule of
engine.StepEngine; torch owns buffers, streams, RNG and NCCL only.  Differences by design:
bf16 tensor-core math instead of fp16 autocast + GradScaler (training.use_amp is accepted and
ignored), no torch.compile (training.use_compile ignored), one flat parameter arena with a fused
Adam instead of torch.optim, and our bucketed all-reduce instead of three DDP wrappers.
"""
import datetime
import json
import os
import random
from pathlib import Path
from typing import Dict, Iterable, Optional, Tuple

import torch

from . import ddp
from .config import Cfg, to_plain
from .data import ShardedBatches, load_store


class ExperimentLogger:
    """Directory naming and params.json writer of vlm_gaze/train/common/logging.py:14-86."""

    def __init__(self, cfg, task: str, rank: int):
        self.cfg, self.rank = cfg, rank
        stamp = datetime.datetime.now().strftime("%Y_%m_%d_%H_%M_%S")
        self.save_dir = f"{stamp}_{self._build_save_tag()}"
        self.log_dir = Path(cfg.logging.log_dir) / task / self.save_dir
        self.ckpt_dir = Path(cfg.logging.checkpoint_dir) / task / self.save_dir
        if rank == 0:
            self.ckpt_dir.mkdir(parents=True, exist_ok=True)
        self.writer = None
        if rank == 0:
            try:
                from torch.utils.tensorboard import SummaryWriter
                self.writer = SummaryWriter(str(self.log_dir))
            except Exception:
                self.writer = None  # tensorboard absent: scalars are still printed

    def _build_save_tag(self) -> str:
        c = self.cfg
        s = (f"s{c.training.seed}_n{c.data.num_episodes}_stack{c.data.frame_stack}"
             f"_gray{c.model.grayscale}_bs{c.data.batch_size}_lr{c.optimizer.lr}")
        sch = c.scheduler.type
        s += {"step": f"_step{c.scheduler.step_size}", "cosine": f"_cosine_eta{c.scheduler.eta_min}",
              "cosine_warm_restarts": f"_coswr_T0{c.scheduler.T_0}",
              "onecycle": f"_onecycle_pct{c.scheduler.pct_start}"}.get(sch, "")
        if c.get("tag"):
            s += f"_{c.tag}"
        if c.data.get("gaze_key"):
            s += f"_{c.data.gaze_key}"
        if c.gaze.method:
            s += f"_{c.gaze.method}"
        if c.dropout.method:
            s += f"_{c.dropout.method}"
        return s

    def rank0_print(self, msg: str):
        if self.rank == 0:
            print(msg, flush=True)

    def add_scalars(self, metrics: Dict[str, float], step: int):
        if self.writer is None:
            return
        for k, v in metrics.items():
            self.writer.add_scalar(k, v, step)
        self.writer.flush()

    def save_params_json(self, params: Dict):
        if not self.cfg.logging.get("save_params", True):
            return
        with open(self.ckpt_dir / "params.json", "w") as f:
            json.dump(params, f, indent=2)


class LRSchedule:
    """build_optimizer + build_scheduler of vlm_gaze/train/common/optim.py:11-108, evaluated on a
    one-element host-side optimizer so that every scheduler type keeps torch's exact semantics
    (including the lr = 0 first step of cosine_warmup); the lr it yields drives the fused Adam."""

    def __init__(self, cfg, loader_len: int):
        oc, sc, tc = cfg.optimizer, cfg.scheduler, cfg.training
        if oc.type not in ("adam", "adamw"):
            raise ValueError(f"Unknown optimizer type: {oc.type}")
        self.decoupled = oc.type == "adamw"
        self.weight_decay = float(oc.weight_decay)
        self._p = torch.nn.Parameter(torch.zeros(1))
        self._opt = torch.optim.SGD([self._p], lr=float(oc.lr))
        accum = max(1, int(tc.gradient_accumulation_steps))
        L = torch.optim.lr_scheduler
        t = sc.type
        self.per_batch = True
        if t == "none":
            self._s, self.per_batch = None, False
        elif t == "step":
            self._s, self.per_batch = L.StepLR(self._opt, step_size=sc.step_size, gamma=sc.gamma), False
        elif t == "cosine":
            self._s, self.per_batch = L.CosineAnnealingLR(self._opt, T_max=tc.epochs, eta_min=float(sc.eta_min)), False
        elif t == "cosine_warm_restarts":
            self._s = L.CosineAnnealingWarmRestarts(self._opt, T_0=sc.T_0, T_mult=sc.T_mult, eta_min=float(sc.eta_min))
        elif t == "cosine_warmup":
            import math
            total = max(1, (loader_len * tc.epochs) // accum)
            warm, base = sc.warmup_steps, float(oc.lr)
            ratio = float(sc.eta_min) / max(1e-12, base)

            def factor(step):
                if step < warm:
                    return step / float(max(1, warm))
                prog = (step - warm) / float(max(1, total - warm))
                return ratio + (1 - ratio) * 0.5 * (1.0 + math.cos(math.pi * prog))
            self._s = L.LambdaLR(self._opt, factor)
        elif t == "onecycle":
            self._s = L.OneCycleLR(self._opt, max_lr=float(oc.lr), total_steps=tc.epochs * loader_len,
                                   pct_start=sc.pct_start, anneal_strategy=sc.anneal_strategy,
                                   div_factor=sc.div_factor, final_div_factor=sc.final_div_factor)
        else:
            raise ValueError(f"Unknown scheduler type: {t}")

    @property
    def lr(self) -> float:
        return self._opt.param_groups[0]["lr"]

    def step(self):
        if self._s is not None:
            self._opt.step()
            self._s.step()


class _StepLoss(torch.autograd.Function):
    """Makes the engine's total loss a tensor whose .backward() runs the kernel backward, so the
    reference loop `loss, bs, m = trainer.compute_loss(batch); (loss / accum).backward()`
    (base_trainer.py:130-136) drives this trainer unchanged.  The 1/accum factor is baked into the
    engine's gradient seeds, hence the incoming grad_output value is not applied again."""

    @staticmethod
    def forward(ctx, anchor, trainer, total):
        ctx.trainer = trainer
        return total.detach().clone()

    @staticmethod
    def backward(ctx, grad_out):
        ctx.trainer._run_backward()
        return torch.zeros((), device=grad_out.device), None, None


class ArenaModule:
    """state_dict()/load_state_dict()/parameters() facade of one reference module (encoder,
    pre_actor, actor) over the flat parameter arena; keys and shapes are the reference's
    (SURVEY.md §8a row 14), tensors are fp32 OIHW / [out,in] regardless of the internal layout."""

    def __init__(self, arena, name: str):
        self.arena, self.name, self.training = arena, name, True

    def state_dict(self):
        return self.arena.state_dicts()[self.name]

    def load_state_dict(self, sd, strict: bool = True):
        self.arena.load_reference({self.name: sd})

    def parameters(self):
        return [self.arena.P(e.name) for e in self.arena.entries if e.module == self.name]

    def train(self, mode: bool = True):
        self.training = mode
        return self

    def eval(self):
        return self.train(False)


class BaseTrainer:
    """Trainer skeleton (base_trainer.py:18-192)."""

    def __init__(self, cfg):
        self.cfg = cfg
        (self.is_distributed, self.rank, self.world_size, self.local_rank,
         self.device) = ddp.init_distributed_if_enabled(cfg.training)
        self._set_seed(int(cfg.training.seed))
        self._setup_data()
        self._setup_model()
        self._setup_optim_scheduler()
        self.experiment = ExperimentLogger(cfg, cfg.data.task, self.rank)
        self.writer = self.experiment.writer
        self.checkpoint_dir = self.experiment.ckpt_dir
        self.log_dir = self.experiment.log_dir

    # hooks --------------------------------------------------------------------------------
    def build_models(self):
        raise NotImplementedError

    def get_optim_params(self) -> Iterable:
        raise NotImplementedError

    def set_train_mode(self):
        raise NotImplementedError

    def compute_loss(self, batch) -> Tuple[torch.Tensor, int, Dict[str, float]]:
        raise NotImplementedError

    def save_for_epoch(self, epoch: int):
        raise NotImplementedError

    def optimizer_step(self):
        raise NotImplementedError

    def zero_grad(self):
        raise NotImplementedError

    def flush_grads(self):
        """All-reduce gradients of an unfinished accumulation window (epoch tail)."""
        red = getattr(self, "reducer", None)
        if red is not None:
            red.flush()

    def poll_device_errors(self):
        """Once per epoch: raise if a tcgen05 kernel reported an mbarrier-wait timeout (the kernels
        only set a device flag), instead of checkpointing silently corrupt weights."""
        if self.device.type == "cuda":
            from . import _lib
            torch.cuda.synchronize(self.device)
            _lib.poll_error()

    # setup ----------------------------------------------------------------------------------
    def _set_seed(self, seed: int):
        import numpy as np
        torch.manual_seed(seed)
        if torch.cuda.is_available():
            torch.cuda.manual_seed_all(seed)
        np.random.seed(seed)
        random.seed(seed)

    def _print_rank0(self, msg: str):
        if self.rank == 0:
            print(msg, flush=True)

    def _setup_data(self):
        c = self.cfg
        store = load_store(str(c.data.hdf5_path), int(c.data.frame_stack), c.data, seed=int(c.training.seed))
        if self.device.type == "cuda":
            free, _ = torch.cuda.mem_get_info(self.device)
            nbytes = store.images.numel()
            store = store.to(self.device) if nbytes < 0.5 * free else store.pin()
        self.train_dataset = store
        # micro-batch = data.batch_size // gradient_accumulation_steps (train/common/data.py:64)
        accum = max(1, int(c.training.gradient_accumulation_steps))
        self.train_loader = ShardedBatches(store, max(1, int(c.data.batch_size) // accum), self.rank, self.world_size,
                                           seed=int(c.training.seed), shuffle=True,
                                           fused=getattr(self, "fused_feed", False))
        self.train_sampler = self.train_loader

    def _setup_model(self):
        self.build_models()

    def _setup_optim_scheduler(self):
        self.schedule = LRSchedule(self.cfg, len(self.train_loader))
        self.scheduler, self.batch_scheduler_update = self.schedule, self.schedule.per_batch
        self.opt_step = 0

    # loop -----------------------------------------------------------------------------------
    def train(self):
        c = self.cfg
        accum = max(1, int(c.training.gradient_accumulation_steps))
        for epoch in range(int(c.training.epochs)):
            # the reference reshuffles every epoch on one GPU too (DataLoader(shuffle=sampler is None),
            # train/common/data.py:62-66), so the permutation seed advances unconditionally
            self.train_sampler.set_epoch(epoch)
            self.set_train_mode()
            epoch_total, epoch_count, sums = 0.0, 0, {}
            self.zero_grad()
            i = -1
            for i, batch in enumerate(self.train_loader):
                loss, bs, metrics = self.compute_loss(batch)
                (loss / accum).backward()
                if (i + 1) % accum == 0:
                    self.optimizer_step()
                    self.zero_grad()
                    if self.batch_scheduler_update:
                        self.schedule.step()
                epoch_total += float(metrics.get("Loss/total", 0.0)) * bs if "Loss/total" in metrics \
                    else float(loss.item()) * bs
                epoch_count += bs
                for k, v in metrics.items():
                    if k != "Loss/total":
                        sums[k] = sums.get(k, 0.0) + float(v) * bs
            # epoch tail (base_trainer.py:161-167): the reference steps when the SAMPLE count is not a
            # multiple of accum; with no pending micro-step its gradients are None and the step is a
            # no-op.  Pending micro-steps were not all-reduced yet (the bucket hooks fire only when a
            # window closes), so flush the reducer first: DDP reduces on every backward.
            if epoch_count % accum != 0 and (i + 1) % accum != 0:
                self.flush_grads()
                self.optimizer_step()
            self.zero_grad()
            self.poll_device_errors()
            avg = epoch_total / max(1, epoch_count)
            if not self.batch_scheduler_update:
                self.schedule.step()
            lr = self.schedule.lr
            self._print_rank0(f"Epoch {epoch + 1}: Loss={avg:.4f}, LR={lr:.6f}")
            if self.writer is not None:
                self.experiment.add_scalars({"Loss/epoch": avg, "LR": lr,
                                             **{k: v / max(1, epoch_count) for k, v in sums.items()}}, epoch)
            self.last_epoch_loss = avg
            if (epoch + 1) % int(c.training.save_interval) == 0 or (epoch + 1) == int(c.training.epochs):
                if self.rank == 0:
                    self.save_for_epoch(epoch + 1)


class BCTrainer(BaseTrainer):
    """GABRIL / GABRIL+GMD behaviour-cloning trainer (train_bc.py:20-335) over the fused step."""
    # HBM-resident stores feed StepEngine.forward_store directly; a host-resident
    # store (too large for HBM) still goes through batch dicts + H2D like the reference
    fused_feed = True

    def __init__(self, cfg):
        super().__init__(cfg)
        self.gaze_ratio = float(cfg.gaze.ratio)

    # ---- hooks ---------------------------------------------------------------------------------
    def build_models(self):
        from .engine import StepEngine
        from .gaze import GazePreprocessor
        from .model import ParamArena, PolicySpec
        c = self.cfg
        if self.device.type != "cuda":
            raise RuntimeError("BCTrainer needs a CUDA device (sm_100a); there is no CPU fallback")
        S = int(c.model.frame_stack)
        base_c = S * (1 if c.model.grayscale else 3)
        cin = base_c * (2 if c.gaze.method == "ViSaRL" else 1)          # train_bc.py:60-61
        spec = PolicySpec(cin, int(c.data.img_height), int(c.data.img_width), int(c.model.embedding_dim),
                          int(c.model.num_hiddens), int(c.model.num_residual_layers),
                          int(c.model.num_residual_hiddens), int(c.model.z_dim), int(c.data.action_dim),
                          agil_channels=base_c if c.gaze.method == "AGIL" else 0,           # train_bc.py:70-77
                          gril_points=int(c.gaze.max_points) if c.gaze.method == "GRIL" else 0)   # train_bc.py:84-86
        self.spec = spec
        self.arena = ParamArena(spec, self.device)
        self.arena.init_reference_style()          # RNG state set by _set_seed, like the reference
        ddp.broadcast_arena(self.arena.p)          # DDP constructor semantics: rank 0's parameters
        self.arena.refresh_shadow()
        self.gaze_preprocessor = GazePreprocessor(
            img_height=int(c.data.img_height), img_width=int(c.data.img_width),
            gaze_sigma=float(c.gaze.mask_sigma), gaze_coeff=float(c.gaze.mask_coeff),
            maxpoints=int(c.gaze.max_points), device=str(self.device),
            temporal_alpha=float(c.gaze.temporal_alpha), temporal_mode=str(c.gaze.temporal_mode),
            temporal_sigmas=c.gaze.temporal_sigmas, temporal_coeffs=c.gaze.temporal_coeffs,
            temporal_offset_start=int(c.gaze.temporal_offset_start))
        self.engine = StepEngine(spec, self.arena, self.gaze_preprocessor, gaze_method=c.gaze.method,
                                 dp_method=c.dropout.method, grayscale=bool(c.model.grayscale),
                                 frame_stack=S, beta=float(c.gaze.beta), lam=float(c.gaze.lambda_weight),
                                 aggregate=bool(c.gaze.get("temporal_flag", True)),
                                 grad_accum=int(c.training.gradient_accumulation_steps),
                                 seed=int(c.training.seed) + 7919 * self.rank,
                                 prob_dist_type=str(c.gaze.prob_dist_type),
                                 gaze_ratio=float(c.gaze.ratio))
        self.reducer = ddp.attach(self.engine, self.arena)
        self.encoder = ArenaModule(self.arena, "encoder")
        self.pre_actor = ArenaModule(self.arena, "pre_actor")
        self.actor = ArenaModule(self.arena, "actor")
        self.gril_gaze_coord_predictor = (ArenaModule(self.arena, "gril_gaze_coord_predictor")
                                          if c.gaze.method == "GRIL" else None)
        self.encoder_agil = ArenaModule(self.arena, "encoder_agil") if c.gaze.method == "AGIL" else None

    def _modules(self):
        return [m for m in (self.encoder, self.pre_actor, self.actor, self.encoder_agil,
                            self.gril_gaze_coord_predictor) if m is not None]

    def get_optim_params(self):
        return [p for m in self._modules() for p in m.parameters()]

    def set_train_mode(self):
        for m in self._modules():
            m.train()

    def compute_loss(self, batch):
        """batch dict -> (total loss tensor, batch size, {"Loss/actor", "Loss/reg"}) like
        train_bc.py:203-299.  The returned tensor's .backward() runs the kernel backward."""
        gaze_key = self.cfg.data.get("gaze_key", "gaze_coords")
        dev = self.device
        use_graph = bool(self.cfg.training.get("cuda_graph", True)) and self.engine.grad_accum == 1
        if "store" in batch:   # device-resident feed (data.ShardedBatches(fused=True))
            if use_graph:
                self.engine.use_graph = True
                losses = self.engine.train_step_store(batch["store"], batch["step_idx"])
                self._bwd_done = True
            else:
                losses = self.engine.forward_store(batch["store"], batch["step_idx"])
            actor, reg, total = self.engine.losses_host()
            anchor = torch.zeros((), device=dev, requires_grad=True)
            loss = _StepLoss.apply(anchor, self, losses[2])
            return loss, int(batch["step_idx"].numel()), {"Loss/actor": actor, "Loss/reg": reg,
                                                          "Loss/total": total}
        frames = batch["obs"]["image"].to(dev, non_blocking=True)
        gaze = batch["obs"][gaze_key].to(dev, non_blocking=True)
        actions = batch["actions"].to(dev, non_blocking=True)
        if gaze.dim() == 4:
            gaze = gaze.reshape(gaze.shape[0], gaze.shape[1], -1)
        if use_graph:
            # training.cuda_graph=true: the whole step (zero_grad + forward + backward) is replayed from
            # CUDA graphs captured over static input buffers (StepEngine.train_step); the later
            # loss.backward() then has nothing left to launch
            st = getattr(self, "_static_in", None)
            if st is None or st[0].shape != frames.shape or st[2].shape != actions.shape:
                st = self._static_in = (torch.empty_like(frames, memory_format=torch.contiguous_format),
                                        torch.empty_like(gaze, dtype=torch.float32, memory_format=torch.contiguous_format),
                                        torch.empty_like(actions, dtype=torch.float32, memory_format=torch.contiguous_format))
            st[0].copy_(frames); st[1].copy_(gaze); st[2].copy_(actions)
            self.engine.use_graph = True
            losses = self.engine.train_step(*st)
            self._bwd_done = True
        else:
            losses = self.engine.forward(frames.contiguous(), gaze.contiguous().float(), actions.float())
        actor, reg, total = self.engine.losses_host()        # one D2H read (the reference does three)
        anchor = torch.zeros((), device=dev, requires_grad=True)
        loss = _StepLoss.apply(anchor, self, losses[2])
        return loss, int(frames.shape[0]), {"Loss/actor": actor, "Loss/reg": reg, "Loss/total": total}

    def _run_backward(self):
        if getattr(self, "_bwd_done", False):
            self._bwd_done = False
            return
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
        """ep{E}_encoder.pth / _actor.pth / _pre_actor.pth (fp32 state_dicts with the reference's
        keys) + params.json with the 14 keys bc_agent.py:47-60 reads (train_bc.py:301-335)."""
        c = self.cfg
        d = Path(self.checkpoint_dir)
        d.mkdir(parents=True, exist_ok=True)
        sds = self.arena.state_dicts()
        # + ep{E}_gril_gaze_coord_predictor.pth / ep{E}_encoder_agil.pth when those modules exist (train_bc.py:310-315)
        for mod in ("encoder", "actor", "pre_actor", "gril_gaze_coord_predictor", "encoder_agil"):
            if mod in sds:
                torch.save({k: v.detach().float().cpu().contiguous() for k, v in sds[mod].items()},
                           d / f"ep{epoch}_{mod}.pth")
        if c.logging.get("save_params", True):
            self.experiment.save_params_json({
                "gaze_method": c.gaze.method, "dp_method": c.dropout.method,
                "grayscale": bool(c.model.grayscale), "stack": int(c.model.frame_stack),
                "embedding_dim": int(c.model.embedding_dim), "num_embeddings": int(c.dropout.num_embeddings),
                "num_hiddens": int(c.model.num_hiddens), "num_residual_layers": int(c.model.num_residual_layers),
                "num_residual_hiddens": int(c.model.num_residual_hiddens), "z_dim": int(c.model.z_dim),
                "gaze_predictor_path": str(c.gaze.get("predictor_path", "")),
                "models_path": str(d), "epochs": int(epoch), "action_dim": int(c.data.action_dim)})
        self._print_rank0(f"Saved checkpoint: {d}/ep{epoch}_*.pth")


def main(argv=None):
    """`python -m gabril_carla_b200.train_bc key=value ... [--gaze_method ... --dp_method GMD
    --result_save_dir DIR]` (train_bc.py:337-351)."""
    import sys
    import yaml
    from .config import load_config
    cfg = load_config(sys.argv[1:] if argv is None else argv)
    if os.environ.get("RANK", "0") == "0":
        print(yaml.safe_dump(to_plain(cfg), sort_keys=False))
    trainer = BCTrainer(cfg)
    trainer.train()
    if os.environ.get("RANK", "0") == "0":
        print("Training completed!")
    ddp.destroy_distributed_if_initialized()
    return trainer
