"""On-hardware multi-GPU parity of the CUDA path (reference
semantics vlm_gaze/train/common/distributed.py:55-68, train_bc.py:116-123; the reference's own check
is test/test_ddp_train_one_epoch.py:23-63).

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/ddp_parity.py \
      [--global-batch 64] [--steps 3] [--out profiles/r2_ddp_parity_nN.txt]

For graphs off and on, with the SAME global batch and the SAME initial weights on every rank:
  (a) non-GMD (gaze.method=Reg): N ranks on contiguous shards, bucketed all-reduce, 1/N folded into
      Adam  vs  ONE rank running the concatenated batch (computed locally on every rank, no reducer):
      reduced gradient arena * 1/N, post-Adam weights after `steps` steps, all-reduced mean losses.
  (b) GMD: batch-global min/max stays rank-local, so the N-rank gradient is the mean of
      the per-rank single-shard gradients: local arena.g before the all-reduce, averaged with a plain
      dist.all_reduce, vs the reducer's result.
  (c) every step: weights bit-identical across ranks, bytes_reduced == 4 * arena.n per step (both
      buckets fired, also under CUDA-graph replay).
Exit code 0 only if every check passes; rank 0 prints / writes the table.
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gabril_carla_b200  # noqa: E402,F401
from gabril_carla_b200 import _lib, ddp  # noqa: E402
from gabril_carla_b200.engine import StepEngine  # noqa: E402
from gabril_carla_b200.gaze import GazePreprocessor  # noqa: E402
from gabril_carla_b200.model import ParamArena, PolicySpec  # noqa: E402
from tests.synth import make_batch  # noqa: E402


def make_engine(device, dp_method, attach):
    spec = PolicySpec(2)
    arena = ParamArena(spec, device)
    arena.init_reference_style(seed=42)     # same seed on every rank ...
    ddp.broadcast_arena(arena.p)            # ... and rank 0's values anyway (DDP constructor semantics)
    arena.refresh_shadow()
    eng = StepEngine(spec, arena, GazePreprocessor(device=str(device)), gaze_method="Reg", dp_method=dp_method)
    red = ddp.attach(eng, arena) if attach else None
    return spec, arena, eng, red


def rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def run(args):
    is_dist, rank, world, local_rank, device = ddp.init_distributed_if_enabled()
    assert is_dist and world >= 2, "launch under torchrun with >= 2 ranks"
    torch.cuda.set_device(device)
    GB = args.global_batch
    lo, hi = ddp.shard_range(GB, rank, world)
    full = tuple(t.to(device) for t in make_batch(GB, seed=77))
    shard = tuple(t[lo:hi].contiguous() for t in full)
    lines, ok = [], True

    def report(tag, value, bound):
        nonlocal ok
        good = value <= bound
        ok = ok and good
        lines.append(f"{tag:78s} {value:.3e}  (bound {bound:.0e}) {'ok' if good else 'FAIL'}")

    for graph in (False, True):
        g_tag = "graph" if graph else "eager"
        # ---------------- (a) Reg, no dropout: N ranks vs one rank on the concatenated batch ----------------
        spec, arena, eng, red = make_engine(device, "None", attach=True)
        spec1, arena1, eng1, _ = make_engine(device, "None", attach=False)
        eng.use_graph = eng1.use_graph = graph
        # step 1 starts from identical weights: only the fp32 summation order differs (per-sample
        # bf16 roundings are scale-invariant under the power-of-two 1/N).  Later steps start from
        # weights that differ by Adam's sign-like updates of near-zero gradient elements, and the loss
        # at lr 1e-4 from an orthogonal init is not yet contracting: informational bound only.
        gb = lambda st: 1e-4 if st == 1 else 5e-2  # noqa: E731
        for step in range(1, args.steps + 1):
            before = red.bytes_reduced
            eng.train_step(*shard)
            red.finish()
            eng1.train_step(*full)
            torch.cuda.synchronize()
            if red.bytes_reduced - before != 4 * arena.n:
                ok = False
                lines.append(f"[{g_tag}] step {step}: bytes_reduced {red.bytes_reduced - before} != {4 * arena.n} FAIL")
            g_n = arena.g * red.grad_scale
            split = arena.offset_of("enc.conv6.w")
            report(f"[{g_tag}] step {step} heads-bucket gradient, N ranks vs 1 rank (rel l2)", rel(g_n[:split], arena1.g[:split]), gb(step))
            report(f"[{g_tag}] step {step} encoder-bucket gradient, N ranks vs 1 rank (rel l2)", rel(g_n[split:], arena1.g[split:]), gb(step))
            losses = eng.buf["losses"].clone()
            dist.all_reduce(losses)
            losses /= world
            l1 = eng1.buf["losses"]
            report(f"[{g_tag}] step {step} mean-over-ranks losses vs 1-rank losses (max rel)",
                   ((losses - l1).abs() / l1.abs().clamp_min(1e-12)).max().item(), 1e-5 if step == 1 else 2e-2)
            eng.adam_step(step, 1e-4, grad_scale=red.grad_scale)
            eng1.adam_step(step, 1e-4)
            ref = arena.p.clone()
            dist.broadcast(ref, src=0)
            same = bool(torch.equal(ref, arena.p))
            flag = torch.tensor([0 if same else 1], device=device)
            dist.all_reduce(flag)
            if flag.item():
                ok = False
            lines.append(f"[{g_tag}] step {step} weights bit-identical across {world} ranks: {flag.item() == 0}")
        # Adam's first steps are sign-like (p -= lr * g / |g|): a gradient element whose sign differs
        # between the two summation orders moves by 2 * lr per step; bound = steps * 2 * lr
        report(f"[{g_tag}] post-Adam weights after {args.steps} steps, N ranks vs 1 rank (max abs)",
               (arena.p - arena1.p).abs().max().item(), args.steps * 2e-4 * 1.01)
        report(f"[{g_tag}] post-Adam weights after {args.steps} steps, N ranks vs 1 rank (rel l2)", rel(arena.p, arena1.p), 1e-4)
        del eng, eng1, arena, arena1
        # ---------------- (b) Reg + GMD: reducer result == mean of per-rank shard gradients ----------------
        spec, arena, eng, red = make_engine(device, "GMD", attach=True)
        spec2, arena2, eng2, _ = make_engine(device, "GMD", attach=False)
        eng.use_graph = eng2.use_graph = graph
        B = hi - lo
        A = torch.rand(B, spec.n, generator=torch.Generator().manual_seed(1000 + rank)).to(device)
        eng.injected_A = eng2.injected_A = A
        for step in range(1, 3):
            eng.train_step(*shard)
            red.finish()
            eng2.train_step(*shard)
            torch.cuda.synchronize()
        mean_local = arena2.g.clone()
        dist.all_reduce(mean_local)
        mean_local /= world
        report(f"[{g_tag}] GMD: reduced gradient vs mean of per-rank shard gradients (rel l2)",
               rel(arena.g * red.grad_scale, mean_local), 1e-5)
        l_loc, l_red = eng2.buf["losses"], eng.buf["losses"]
        report(f"[{g_tag}] GMD: losses of the reduced run vs the local run on the same shard (max rel)",
               ((l_loc - l_red).abs() / l_loc.abs().clamp_min(1e-12)).max().item(), 1e-6)
        del eng, eng2, arena, arena2
    _lib.poll_error()
    bad = torch.tensor([0 if ok else 1], device=device)
    dist.all_reduce(bad)
    if rank == 0:
        head = (f"# ddp_parity: world {world}, global batch {GB}, {args.steps} steps, GPU {torch.cuda.get_device_name(0)}, "
                f"NCCL {'.'.join(map(str, torch.cuda.nccl.version()))}\n")
        text = head + "\n".join(lines) + f"\nRESULT: {'PASS' if bad.item() == 0 else 'FAIL'}\n"
        print(text, flush=True)
        if args.out:
            os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
            with open(args.out, "w") as f:
                f.write(text)
    ddp.destroy_distributed_if_initialized()
    return int(bad.item() != 0)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--global-batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--out", default="")
    sys.exit(run(ap.parse_args()))
