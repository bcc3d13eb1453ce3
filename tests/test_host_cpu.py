"""CPU (-m "not gpu"): host logic of the drop-in surface — config tree / overrides / legacy flags,
the demonstration store and sharded sampler, the lr schedule, the experiment directory tag and the
parameter arena's reference <-> internal layout round trip."""
import math
import os

import numpy as np
import pytest
import torch

import gabril_carla_b200  # noqa: F401
from gabril_carla_b200 import config as C
from gabril_carla_b200.data import ShardedBatches, synthetic_store
from gabril_carla_b200.model import ParamArena, PolicySpec
from gabril_carla_b200.trainer import ExperimentLogger, LRSchedule
from tests.synth import make_weights, reference_shapes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")


# ---- config ----------------------------------------------------------------------------------------
def test_defaults_match_reference_tree():
    cfg = C.load_config([])
    assert cfg.gaze.method == "Reg" and cfg.dropout.method == "None"
    assert cfg.model.frame_stack == cfg.data.frame_stack == 2          # ${data.frame_stack}
    assert cfg.model.grayscale is True and cfg.data.action_dim == 7
    assert cfg.optimizer.lr == 5e-4 and cfg.scheduler.eta_min == 1e-6
    assert cfg.gaze.beta == 50.0 and cfg.gaze.lambda_weight == 10.0 and cfg.gaze.mask_sigma == 30.0
    assert cfg.training.distributed.backend == "nccl"


def test_hydra_style_and_legacy_overrides():
    cfg = C.load_config(["gaze.method=Reg", "dropout.method=GMD", "data.batch_size=128", "optimizer.lr=1e-3",
                         "model.grayscale=false", "data.frame_stack=4", "--config-name=train_bc"])
    assert cfg.dropout.method == "GMD" and cfg.data.batch_size == 128 and cfg.optimizer.lr == 1e-3
    assert cfg.model.grayscale is False and cfg.model.frame_stack == 4
    cfg = C.load_config(["--gaze_method", "None", "--dp_method=GMD", "--result_save_dir", "/tmp/x", "--seed", "3"])
    assert cfg.gaze.method == "None" and cfg.dropout.method == "GMD"
    assert cfg.logging.checkpoint_dir == "/tmp/x" and cfg.training.seed == 3
    with pytest.raises(SystemExit):
        C.load_config(["gaze.method=Bogus"])
    with pytest.raises(SystemExit):
        C.load_config(["--no_such_flag", "1"])


def test_yaml_file_with_string_floats(tmp_path):
    p = tmp_path / "c.yaml"
    p.write_text("defaults:\n  - train_bc_base\n  - _self_\noptimizer:\n  lr: 5e-4\nscheduler:\n  type: step\n  eta_min: 1e-6\n"
                 "gaze:\n  temporal_sigmas: [30.0, 24.0]\n")
    cfg = C.load_config([f"--config={p}", "training.epochs=3"])
    assert isinstance(cfg.optimizer.lr, float) and cfg.optimizer.lr == 5e-4
    assert cfg.scheduler.type == "step" and cfg.training.epochs == 3 and cfg.gaze.temporal_sigmas == [30.0, 24.0]


# ---- data -------------------------------------------------------------------------------------------
def test_store_windows_and_batch_format():
    st = synthetic_store(40, frame_stack=2, H=18, W=32, demo_len=16, seed=1)
    w = st.window_indices(torch.tensor([0, 1, 16, 17, 39]))
    assert w.tolist() == [[0, 0], [0, 1], [16, 16], [16, 17], [38, 39]]   # first frame repeated at a demo start
    b = st.batch(torch.tensor([0, 5, 16]))
    assert b["obs"]["image"].shape == (3, 2, 18, 32, 3) and b["obs"]["image"].dtype == torch.uint8
    assert b["obs"]["gaze_coords"].shape == (3, 2, 10) and b["actions"].shape == (3, 2, 7)
    assert torch.equal(b["obs"]["image"][0, 0], b["obs"]["image"][0, 1])
    st4 = synthetic_store(40, frame_stack=4, H=18, W=32, demo_len=16, seed=1)
    assert st4.window_indices(torch.tensor([17])).tolist() == [[16, 16, 16, 17]]


def test_sharded_batches_match_distributed_sampler():
    from torch.utils.data import DistributedSampler
    st = synthetic_store(37, frame_stack=2, H=18, W=32, seed=2)
    for world in (1, 2, 4):
        seen = []
        for rank in range(world):
            sb = ShardedBatches(st, 8, rank, world, seed=5)
            sb.set_epoch(3)
            ds = DistributedSampler(range(37), num_replicas=world, rank=rank, shuffle=True, seed=5, drop_last=False)
            ds.set_epoch(3)
            assert sb.indices().tolist() == list(iter(ds))
            assert len(sb) == math.ceil(len(list(iter(ds))) / 8)
            seen += sb.indices().tolist()
            n = sum(b["actions"].shape[0] for b in sb)
            assert n == sb.per_rank
        assert set(seen) == set(range(37))


# ---- schedule / logger --------------------------------------------------------------------------------
def test_lr_schedule_matches_reference_trajectory():
    gold = np.load(GOLD)
    cfg = C.load_config(["training.epochs=4"])
    s = LRSchedule(cfg, 300)
    assert s.per_batch
    for i in range(1200):
        assert math.isclose(s.lr, float(gold["sched/lr"][i]), rel_tol=1e-12, abs_tol=1e-18), i
        s.step()
    s2 = LRSchedule(C.load_config(["scheduler.type=step", "scheduler.step_size=2", "scheduler.gamma=0.5"]), 10)
    assert not s2.per_batch
    lrs = []
    for _ in range(5):
        lrs.append(s2.lr)
        s2.step()
    assert lrs == [5e-4, 5e-4, 2.5e-4, 2.5e-4, 1.25e-4]
    assert LRSchedule(C.load_config(["scheduler.type=none"]), 10).per_batch is False


def test_experiment_dirs_and_tag(tmp_path):
    cfg = C.load_config([f"logging.checkpoint_dir={tmp_path}", f"logging.log_dir={tmp_path}", "dropout.method=GMD",
                         "tag=abc", "data.batch_size=64"])
    lg = ExperimentLogger(cfg, cfg.data.task, rank=0)
    tag = lg.save_dir.split("_", 6)[-1]
    assert tag == "s42_n200_stack2_grayTrue_bs64_lr0.0005_abc_gaze_coords_Reg_GMD"
    assert lg.ckpt_dir.parent == tmp_path / "Mixed_" and lg.ckpt_dir.is_dir()
    lg.save_params_json({"epochs": 1})
    assert (lg.ckpt_dir / "params.json").exists()


# ---- parameter arena --------------------------------------------------------------------------------
@pytest.mark.parametrize("cin", [2, 6])
def test_arena_reference_layout_round_trip(cin):
    spec = PolicySpec(cin)
    arena = ParamArena(spec, "cpu")
    w = make_weights(cin, seed=3)
    arena.load_reference(w)
    out = arena.state_dicts()
    shapes = reference_shapes(cin)
    for mod in ("encoder", "pre_actor", "actor"):
        assert list(out[mod].keys()) == list(shapes[mod].keys())         # nn.Module.state_dict() order
        for k, v in out[mod].items():
            assert tuple(v.shape) == shapes[mod][k] and v.dtype == torch.float32
            assert torch.equal(v, w[mod][k]), (mod, k)
    assert spec.total == sum(v.numel() for sd in w.values() for v in sd.values()) or spec.total >= 13_528_871
    # `module.` / `_orig_mod.` prefixes are stripped like bc_agent.py:107-128
    arena.load_reference({"actor": {"module._orig_mod." + k: v for k, v in w["actor"].items()}})
    with pytest.raises(RuntimeError):
        arena.load_reference({"actor": {"0.weight": w["actor"]["0.weight"]}})


# ---- trainer loop (host logic on CPU with recording hooks) ----------------------------------------
class _LoopProbe:
    """BaseTrainer with the five hooks replaced by recorders: exercises the epoch / accumulation /
    tail / scheduler logic of base_trainer.py:122-192 without a GPU."""

    @staticmethod
    def make(tmp_path, *extra):
        from gabril_carla_b200.trainer import BaseTrainer

        class Probe(BaseTrainer):
            def build_models(self):
                self.log = []

                class Red:
                    def __init__(s): s.flushed = 0
                    def flush(s): s.flushed += 1
                self.reducer = Red()

            def set_train_mode(self): pass

            def compute_loss(self, batch):
                idx = batch["actions"].shape[0]
                self.log.append(("loss", idx, batch["actions"][:, -1, 0].clone()))
                t = torch.zeros((), requires_grad=True)
                return t * 1.0, idx, {"Loss/total": 0.0}

            def optimizer_step(self): self.log.append(("step", self.schedule.lr))
            def zero_grad(self): pass
            def save_for_epoch(self, epoch): self.log.append(("save", epoch))
        cfg = C.load_config(["data.hdf5_path=synthetic:40", "data.img_height=18", "data.img_width=32",
                             f"logging.checkpoint_dir={tmp_path}", f"logging.log_dir={tmp_path}",
                             "training.device=cpu", *extra])
        return Probe(cfg)


def test_single_gpu_epochs_reshuffle(tmp_path):
    tr = _LoopProbe.make(tmp_path, "data.batch_size=8", "training.epochs=2", "scheduler.type=none")
    tr.train()
    losses = [e for e in tr.log if e[0] == "loss"]
    assert len(losses) == 10                                           # 5 batches x 2 epochs
    ep0 = torch.cat([e[2] for e in losses[:5]]); ep1 = torch.cat([e[2] for e in losses[5:]])
    assert not torch.equal(ep0, ep1) and torch.equal(ep0.sort().values, ep1.sort().values)
    sb = tr.train_loader
    sb.set_epoch(0); a = sb.indices().clone(); sb.set_epoch(1)
    assert not torch.equal(a, sb.indices())


def test_accumulation_micro_batch_tail_and_schedule(tmp_path):
    # batch 8 with accum 2 -> micro-batch 4 (data.py:64): 40 samples = 10 micro-steps = 5 optimizer steps
    tr = _LoopProbe.make(tmp_path, "data.batch_size=8", "training.gradient_accumulation_steps=2",
                         "training.epochs=1", "scheduler.warmup_steps=2")
    assert tr.train_loader.bs == 4 and len(tr.train_loader) == 10
    tr.train()
    steps = [e for e in tr.log if e[0] == "step"]
    assert len(steps) == 5 and tr.reducer.flushed == 0
    # cosine_warmup over total = len(loader) * epochs // accum = 5 optimizer steps, lr(0) = 0
    ref = [5e-4 * O_factor(i, 2, 5, 1e-6, 5e-4) for i in range(5)]
    assert [s[1] for s in steps] == pytest.approx(ref, rel=1e-12, abs=1e-18)
    # 36 samples, micro-batch 4, accum 2 -> 9 micro-steps: the tail window holds one micro-step, but the
    # reference steps only when the SAMPLE count is not a multiple of accum (36 % 2 == 0: no tail step)
    tr = _LoopProbe.make(tmp_path, "data.hdf5_path=synthetic:36", "data.batch_size=8", "scheduler.type=none",
                         "training.gradient_accumulation_steps=2", "training.epochs=1")
    tr.train()
    assert len([e for e in tr.log if e[0] == "step"]) == 4 and tr.reducer.flushed == 0
    # 35 samples: 9 micro-steps (the last one of 3 samples), 35 % 2 != 0 -> flush the reducer, then step
    tr = _LoopProbe.make(tmp_path, "data.hdf5_path=synthetic:35", "data.batch_size=8", "scheduler.type=none",
                         "training.gradient_accumulation_steps=2", "training.epochs=1")
    tr.train()
    assert len([e for e in tr.log if e[0] == "step"]) == 5 and tr.reducer.flushed == 1


def O_factor(step, warm, total, eta_min, base):
    from oracle.gabril_oracle import cosine_warmup_factor
    return cosine_warmup_factor(step, warm, total, eta_min, base)
