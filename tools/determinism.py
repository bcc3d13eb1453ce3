"""Where does the CUDA-graph vs eager loss divergence of tests/test_trainer_gpu.py come from?

(1) engine level: the same step (same weights, same batch) run R times eagerly and R times by graph
    replay -> per-parameter max |g_i - g_0| and whether the gradient arena is bit-identical run to run.
(2) trainer level: the 3-step reference-style loop of test_cuda_graph_step_matches_eager_loop, twice
    eager and twice with training.cuda_graph=true, same seeds -> loss trajectories and final weight
    spread (eager vs eager tells fp32-atomic ordering apart from a capture race).

python tools/determinism.py [B] [R]     (writes a human-readable report to stdout)
"""
import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gabril_carla_b200  # noqa: E402,F401
from gabril_carla_b200 import _lib  # noqa: E402
from gabril_carla_b200.engine import StepEngine  # noqa: E402
from gabril_carla_b200.gaze import GazePreprocessor  # noqa: E402
from gabril_carla_b200.model import ParamArena, PolicySpec  # noqa: E402
from tests.synth import make_batch  # noqa: E402

DEV = "cuda:0"
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
R = int(sys.argv[2]) if len(sys.argv) > 2 else 4


def engine_level():
    spec = PolicySpec(2)
    arena = ParamArena(spec, DEV)
    arena.init_reference_style(seed=0)
    eng = StepEngine(spec, arena, GazePreprocessor(device=DEV), gaze_method="Reg", dp_method="None")
    batch = tuple(t.to(DEV) for t in make_batch(B, seed=21))
    out = {}
    for mode in ("eager", "graph"):
        eng.use_graph = mode == "graph"
        grads, losses = [], []
        for r in range(R + 2):
            eng.train_step(*batch)
            torch.cuda.synchronize()
            if r >= 2:   # the first graph-mode call runs eagerly (allocations), the second captures
                grads.append(arena.g.clone())
                losses.append(eng.losses_host())
        out[mode] = (grads, losses)
        ident = all(torch.equal(g, grads[0]) for g in grads)
        print(f"[engine {mode}] B={B}: gradient arena bit-identical over {R} runs: {ident}")
        print(f"   losses: {[tuple(round(v, 7) for v in l) for l in losses]}")
        for e in arena.entries:
            sl = slice(e.offset, e.offset + e.numel)
            d = max((g[sl] - grads[0][sl]).abs().max().item() for g in grads)
            n = grads[0][sl].abs().max().item()
            if d > 0:
                print(f"   {e.name:18s} max|dg| {d:.3e}  (max|g| {n:.3e}, rel {d / (n + 1e-30):.2e})")
    ge, gg = out["eager"][0][0], out["graph"][0][0]
    print(f"[engine eager vs graph] max|dg| {(ge - gg).abs().max().item():.3e}, "
          f"rel-norm {((ge - gg).norm() / ge.norm()).item():.3e}")
    _lib.poll_error()


def trainer_level():
    from gabril_carla_b200.config import load_config
    from gabril_carla_b200.trainer import BCTrainer
    runs = {}
    for mode in ("false", "false", "true", "true"):
        with tempfile.TemporaryDirectory() as tmp:
            cfg = load_config(["data.hdf5_path=synthetic:96", "data.batch_size=32", "training.epochs=1",
                               "training.save_interval=1", "scheduler.type=none", "optimizer.lr=1e-4",
                               f"logging.checkpoint_dir={tmp}", f"logging.log_dir={tmp}",
                               f"training.cuda_graph={mode}"])
            tr = BCTrainer(cfg)
            tr.fused_feed = False
            tr.train_loader.fused = False
            tr.set_train_mode()
            losses = []
            for batch in tr.train_loader:
                tr.zero_grad()
                loss, bs, m = tr.compute_loss(batch)
                loss.backward()
                tr.optimizer_step()
                losses.append(m["Loss/total"])
            torch.cuda.synchronize()
            runs.setdefault(mode, []).append((losses, tr.arena.p.clone()))
    for a, b, tag in ((runs["false"][0], runs["false"][1], "eager vs eager"),
                      (runs["true"][0], runs["true"][1], "graph vs graph"),
                      (runs["false"][0], runs["true"][0], "eager vs graph")):
        rel = [abs(x - y) / abs(x) for x, y in zip(a[0], b[0])]
        print(f"[trainer {tag}] losses {a[0]} | {b[0]}")
        print(f"   rel diff per step {['%.2e' % r for r in rel]}; final weights max|dp| {(a[1] - b[1]).abs().max().item():.3e}")


if __name__ == "__main__":
    engine_level()
    trainer_level()
