"""-m gpu: the drop-in trainer surface (BCTrainer hooks, reference-style loop, checkpoint layout)."""
import json
import os

import pytest
import torch

from oracle import gabril_oracle as O
from tests.synth import reference_shapes

pytestmark = pytest.mark.gpu


def make_cfg(tmp_path, *extra):
    import gabril_carla_b200  # noqa: F401
    from gabril_carla_b200.config import load_config
    return load_config(["data.hdf5_path=synthetic:96", "data.batch_size=32", "training.epochs=2",
                        "training.save_interval=1", "scheduler.type=none", "optimizer.lr=1e-4",
                        f"logging.checkpoint_dir={tmp_path}", f"logging.log_dir={tmp_path}", *extra])


def test_compute_loss_matches_oracle_and_reference_loop_style(tmp_path):
    from gabril_carla_b200.trainer import BCTrainer
    tr = BCTrainer(make_cfg(tmp_path, "dropout.method=GMD"))
    tr.set_train_mode()
    batch = next(iter(tr.train_loader))
    A = torch.rand(32, 1, 20, 38, generator=torch.Generator().manual_seed(4))
    tr.engine.injected_A = A.to(tr.device)
    sds = {m: {k: v.cpu() for k, v in sd.items()} for m, sd in tr.arena.state_dicts().items()}
    tr.zero_grad()
    loss, bs, metrics = tr.compute_loss(batch)           # the reference's hook signature
    assert bs == 32 and isinstance(metrics["Loss/actor"], float) and loss.dim() == 0
    (loss / 1).backward()                                # the reference's loop statement
    tr.optimizer_step()
    torch.cuda.synchronize()
    cpu = lambda t: t.cpu()
    total, actor, reg, _ = O.bc_losses(sds["encoder"], sds["pre_actor"], sds["actor"],
                                       cpu(batch["obs"]["image"]), cpu(batch["obs"]["gaze_coords"]),
                                       cpu(batch["actions"]), gaze_method="Reg", dp_method="GMD", gmd_A=A)
    # bf16 activations vs the fp32 oracle: losses rel 1e-3 x safety 5 (north_star tolerance family)
    assert abs(metrics["Loss/actor"] - actor.item()) <= 5e-3 * abs(actor.item())
    assert abs(metrics["Loss/reg"] - reg.item()) <= 5e-3 * abs(reg.item())
    assert abs(loss.item() - total.item()) <= 5e-3 * abs(total.item())
    # parameters moved
    new = tr.arena.state_dicts()["actor"]["2.weight"].cpu()
    assert (new - sds["actor"]["2.weight"]).abs().max() > 0


def test_train_two_epochs_and_checkpoint_layout(tmp_path):
    from gabril_carla_b200.trainer import BCTrainer
    tr = BCTrainer(make_cfg(tmp_path))
    first = {}
    tr.train()
    d = tr.checkpoint_dir
    assert d.parent.name == "Mixed_"
    for ep in (1, 2):
        for mod in ("encoder", "pre_actor", "actor"):
            assert (d / f"ep{ep}_{mod}.pth").exists()
    params = json.load(open(d / "params.json"))
    for k in ("gaze_method", "dp_method", "grayscale", "stack", "embedding_dim", "num_embeddings", "num_hiddens",
              "num_residual_layers", "num_residual_hiddens", "z_dim", "gaze_predictor_path", "models_path",
              "epochs", "action_dim"):
        assert k in params, k
    assert params["epochs"] == 2 and params["stack"] == 2 and params["models_path"] == str(d)
    # the reference loader (bc_agent.py:96-128): torch.load(weights_only=True), strict key/shape match
    shapes = reference_shapes(2)
    sds = {}
    for mod in ("encoder", "pre_actor", "actor"):
        sd = torch.load(d / f"ep2_{mod}.pth", weights_only=True)
        assert list(sd.keys()) == list(shapes[mod].keys())
        for k, v in sd.items():
            assert tuple(v.shape) == shapes[mod][k] and v.dtype == torch.float32 and v.device.type == "cpu"
            assert torch.isfinite(v).all()
        sds[mod] = sd
    # the saved weights drive the reference forward (oracle restatement) to finite actions
    batch = next(iter(tr.train_loader))
    obs, heat, c = O.prepare_for_bc(batch["obs"]["image"].cpu()[:2], batch["obs"]["gaze_coords"].cpu()[:2], 2, True)
    z = O.encoder_forward(sds["encoder"], obs)
    _, logits = O.heads_forward(sds["pre_actor"], sds["actor"], z)
    assert logits.shape == (2, 7) and torch.isfinite(logits).all()
    assert tr.last_epoch_loss == tr.last_epoch_loss  # not NaN


def test_unsupported_methods_raise(tmp_path):
    from gabril_carla_b200.trainer import BCTrainer
    # the two paths that raise inside the reference itself (tests/golden/make_golden_n4.py)
    with pytest.raises(NotImplementedError):
        BCTrainer(make_cfg(tmp_path, "gaze.method=Contrastive"))
    with pytest.raises(NotImplementedError):
        BCTrainer(make_cfg(tmp_path, "dropout.method=Oreo"))


@pytest.mark.parametrize("method", ["Mask", "ViSaRL", "AGIL", "GRIL"])
def test_other_gaze_methods_train_and_checkpoint(tmp_path, method):
    """SURVEY §8f N4 through the trainer: one epoch, and the extra checkpoint files of train_bc.py:310-315."""
    from gabril_carla_b200.trainer import BCTrainer
    tr = BCTrainer(make_cfg(tmp_path, f"gaze.method={method}", "training.epochs=1"))
    tr.train()
    d = tr.checkpoint_dir
    enc = torch.load(d / "ep1_encoder.pth", weights_only=True)
    assert tuple(enc["_conv_1.weight"].shape) == (32, 4 if method == "ViSaRL" else 2, 4, 4)
    assert (d / "ep1_encoder_agil.pth").exists() == (method == "AGIL")
    assert (d / "ep1_gril_gaze_coord_predictor.pth").exists() == (method == "GRIL")
    if method == "GRIL":
        g = torch.load(d / "ep1_gril_gaze_coord_predictor.pth", weights_only=True)
        assert {k: tuple(v.shape) for k, v in g.items()} == {"0.weight": (256, 256), "0.bias": (256,),
                                                             "2.weight": (10, 256), "2.bias": (10,)}
    assert tr.last_epoch_loss == tr.last_epoch_loss


def test_cuda_graph_step_matches_eager_loop(tmp_path):
    """training.cuda_graph=true: the reference-style loop driven by graph replays computes the eager
    loop's step.  The two trainers are compared STEP BY STEP FROM IDENTICAL WEIGHTS (the graph trainer's
    arena is overwritten with the eager trainer's before every step): fp32 atomics (split-K wgrad, bias
    column sums, the split-K pre_actor GEMM) make any two runs of the same step differ by ~1e-7 relative
    (tools/determinism.py, profiles/r2_determinism.txt: eager vs eager == eager vs graph), and Adam's
    first steps are sign-like (p -= lr * g / |g|), so a free-running 3-step trajectory amplifies one
    flipped near-zero gradient element chaotically (loss 4 -> 53 -> 2.8 at this lr): that comparison
    (rounds 1's version of this test) is unsound at any fixed bound.  Per step the bounds are tight."""
    from gabril_carla_b200.trainer import BCTrainer
    lr = 1e-4
    trs = {}
    for mode in ("false", "true"):
        tr = BCTrainer(make_cfg(tmp_path / mode, f"training.cuda_graph={mode}", "training.epochs=1"))
        tr.fused_feed = False
        tr.train_loader.fused = False
        tr.set_train_mode()
        trs[mode] = tr
    eager, graph = trs["false"], trs["true"]
    assert torch.equal(eager.arena.p, graph.arena.p)          # same seed -> same initial weights
    n_steps = 0
    for be, bg in zip(eager.train_loader, graph.train_loader):
        assert torch.equal(be["actions"], bg["actions"])      # same sampler stream
        for name in ("p", "m", "v", "wb"):                    # identical weights / Adam state before the step
            getattr(graph.arena, name).copy_(getattr(eager.arena, name))
        out = {}
        for tag, tr, batch in (("eager", eager, be), ("graph", graph, bg)):
            tr.zero_grad()
            loss, bs, m = tr.compute_loss(batch)
            loss.backward()
            torch.cuda.synchronize()
            g = tr.arena.g.clone()
            tr.optimizer_step()
            torch.cuda.synchronize()
            out[tag] = (m, g, tr.arena.p.clone())
        (me, ge, pe), (mg, gg, pg) = out["eager"], out["graph"]
        for k in ("Loss/actor", "Loss/reg", "Loss/total"):
            assert abs(me[k] - mg[k]) <= 1e-5 * abs(me[k]), (k, me[k], mg[k])
        assert ((ge - gg).norm() / ge.norm()).item() <= 1e-5
        assert (pe - pg).abs().max().item() <= 2 * lr * 1.01   # one sign-like Adam step apart at most
        assert ((pe - pg).norm() / pe.norm()).item() <= 1e-6
        n_steps += 1
    assert n_steps == 3 and len(graph.engine._graphs) == 1
    assert all(v == v for v in me.values())                   # finite after three steps
