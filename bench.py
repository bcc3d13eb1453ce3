#!/usr/bin/env python
"""bench.py — GABRIL train samples/sec (320x180 stack, fwd+bwd+gaze reg + Adam) on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one pass of the train_bc.py hot path (BCTrainer.compute_loss + backward + Adam,
vlm_gaze/train/train_bc.py:203-299, train/common/base_trainer.py:129-149) over one synthetic
Bench2Drive-shaped batch.  Workload (BASELINE.json configs[1] / configs[4]): gaze.method=Reg,
grayscale stack 2, 180x320, action_dim 7, GLOBAL batch 1024 sharded over the N ranks (strong
scaling, the configuration BASELINE.json names); --per-gpu-batch switches to the reference's
DDP semantics (data.batch_size per rank, weak scaling).

`value`  : device-timed, inputs resident in HBM (CUDA events, max over ranks).
`e2e`    : the same K steps through the public API with the batch in pinned HOST memory: the H2D
           copy of every step's inputs and the D2H read of its losses are inside the timed region.
`roofline`: the tcgen05 implicit-GEMM kernel (all conv / linear launches of one step), algorithmic
           FLOPs / summed CUDA-event launch time vs the measured bf16 peak.
`cpu_baseline` / --impl reference: the CPU restatement of the reference step (oracle/, torch CPU
           fp32, all host threads) on a bounded sample (batch 32).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GLOBAL_BATCH = 1024
SPINUP_S = 1.5   # untimed steps before the first measurement (GPU clock ramp of a fresh box), see main_ours
FLOP_PER_SAMPLE = {True: 5.3816e9, False: 5.4996e9}  # fwd+bwd, gray / RGB


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--global-batch", type=int, default=GLOBAL_BATCH)
    ap.add_argument("--per-gpu-batch", type=int, default=0, help="weak scaling: fixed batch per rank")
    ap.add_argument("--gaze-method", default="Reg")
    ap.add_argument("--dp-method", default="None")
    ap.add_argument("--rgb", action="store_true")
    ap.add_argument("--cpu-batch", type=int, default=32)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--graph", action="store_true", help="CUDA-graph replays of the step (StepEngine.train_step)")
    ap.add_argument("--no-graph", action="store_true", help="eager launches")
    ap.add_argument("--no-store-feed", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the short runs of the other configurations")
    ap.add_argument("--store-steps", type=int, default=8192, help="steps held by the HBM-resident store leg")
    ap.add_argument("--trace-out", default="")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# CPU leg: the oracle restatement of the reference step, timed on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_reference_run(args, steps, warmup):
    import torch
    from oracle import gabril_oracle as O
    import gabril_carla_b200  # noqa: F401
    from gabril_carla_b200.model import ParamArena, PolicySpec
    from tests.synth import make_batch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    gray = not args.rgb
    spec = PolicySpec(2 if gray else 6)
    arena = ParamArena(spec, "cpu")
    arena.init_reference_style(seed=42)
    sds = arena.state_dicts()
    params = [v.requires_grad_(True) for sd in sds.values() for v in sd.values()]
    opt = torch.optim.Adam(params, lr=5e-4)
    B = args.cpu_batch
    batch = make_batch(B, seed=0)
    A = torch.rand(B, 1, spec.h, spec.w) if args.dp_method == "GMD" else None
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        total, _, _, _ = O.bc_losses(sds["encoder"], sds["pre_actor"], sds["actor"], *batch,
                                     gaze_method=args.gaze_method, dp_method=args.dp_method,
                                     grayscale=gray, gmd_A=A)
        opt.zero_grad(set_to_none=True)
        total.backward()
        opt.step()
        float(total.item())
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    ms = 1e3 * sum(times) / len(times)
    return dict(value=B / (ms / 1e3), unit="samples/s", cores=cores, kind="port",
                sample=f"batch {B} x {steps} steps after {warmup} warm-up, fp32 torch-CPU restatement "
                       f"(oracle/gabril_oracle.py) of compute_loss+backward+Adam, {cores} threads",
                ms_per_step=ms)


def workload_config(args, B, world):
    gray = not args.rgb
    return {"workload": f"train_bc.py GABRIL step gaze.method={args.gaze_method} dropout.method={args.dp_method} "
                        f"({'grayscale' if gray else 'RGB'} stack 2, 180x320, action_dim 7, Adam), "
                        "synthetic Bench2Drive-shaped frames + gaze coords",
            "global_batch": B * world, "per_gpu_batch": B, "parallelism": f"dp{world}",
            "l2": "per-step working set (>6 MB/sample activations, >6 GB per step) exceeds the 126 MB L2; no flush needed",
            "spinup": f"{SPINUP_S} s of untimed steps before the first measurement (clock ramp of a fresh box), then W warm-up steps"}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference_run(args, args.steps, args.warmup)
    line = {"impl": "reference", "metric": "GABRIL train samples/sec (320x180 stack, fwd+bwd+gaze reg)",
            "value": r["value"], "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, args.cpu_batch, 1),
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU leg
# ------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock / power / throttle reasons sampled DURING the timed regions: NVML every 5 ms (a short
    timed region still gets tens of samples), nvidia-smi as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.source = index, [], False, "nvml"
        self.sm_max = None

    def _nvml_loop(self):
        import pynvml as N
        N.nvmlInit()
        h = N.nvmlDeviceGetHandleByIndex(self.index)
        self.sm_max = float(N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM))
        bits = [getattr(N, n, 0) for n in ("nvmlClocksEventReasonHwSlowdown", "nvmlClocksEventReasonHwThermalSlowdown",
                                           "nvmlClocksEventReasonSwThermalSlowdown", "nvmlClocksEventReasonSwPowerCap")]
        if not all(bits):
            bits = [getattr(N, n, 0) for n in ("nvmlClocksThrottleReasonHwSlowdown", "nvmlClocksThrottleReasonHwThermalSlowdown",
                                               "nvmlClocksThrottleReasonSwThermalSlowdown", "nvmlClocksThrottleReasonSwPowerCap")]
        get_reasons = getattr(N, "nvmlDeviceGetCurrentClocksEventReasons", None) or N.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            r = get_reasons(h)
            self.rows.append((float(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)), N.nvmlDeviceGetPowerUsage(h) / 1e3,
                              [bool(r & b) for b in bits]))
            time.sleep(0.005)

    def _smi_loop(self):
        self.source = "nvidia-smi"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                p = [x.strip() for x in out.strip().split(",")]
                if len(p) >= 7:
                    self.sm_max = float(p[1])
                    self.rows.append((float(p[0]), float(p[2]), [x.lower().startswith("active") for x in p[3:7]]))
            except Exception:
                pass
            time.sleep(0.1)

    def run(self):
        try:
            self._nvml_loop()
        except Exception:
            self._smi_loop()

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock samples"]}
        sm = sorted(r[0] for r in self.rows)
        reasons = [n for i, n in enumerate(self.NAMES) if any(r[2][i] for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.sm_max, "reasons": reasons, "samples": len(self.rows),
                "power_w_max": max(r[1] for r in self.rows), "source": self.source}


def synth_batch(B, seed):
    """B DISTINCT synthetic Bench2Drive-shaped samples (tests.synth.make_batch in chunks of 64, one seed per
    chunk), pinned on the host."""
    import torch
    from tests.synth import make_batch
    parts = [make_batch(min(64, B - lo), seed=seed + 7919 * (lo // 64)) for lo in range(0, B, 64)]
    return tuple(torch.cat([p[i] for p in parts]).contiguous().pin_memory() for i in range(3))


def quick_run(device, gaze_method, dp_method, rgb, B, steps=10, warmup=3, seed=0):
    """Short device-timed run of another configuration (CUDA-graph replays, inputs resident in HBM)."""
    import torch
    from gabril_carla_b200.engine import StepEngine
    from gabril_carla_b200.gaze import GazePreprocessor
    from gabril_carla_b200.model import ParamArena, PolicySpec
    spec = PolicySpec(6 if rgb else 2)
    arena = ParamArena(spec, device)
    arena.init_reference_style(seed=42)
    eng = StepEngine(spec, arena, GazePreprocessor(device=str(device)), gaze_method=gaze_method,
                     dp_method=dp_method, grayscale=not rgb, seed=seed)
    eng.use_graph = True
    base = synth_batch(min(B, 256), seed=300)
    reps = (B + base[0].shape[0] - 1) // base[0].shape[0]
    batch = tuple(t.repeat((reps,) + (1,) * (t.dim() - 1))[:B].contiguous().to(device) for t in base)
    n = [0]

    def step():
        n[0] += 1
        eng.train_step(*batch)
        eng.adam_step(n[0], 5e-4)
    for _ in range(max(3, warmup)):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    losses = eng.losses_host()
    del eng, arena, batch
    torch.cuda.empty_cache()
    return {"value": B / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms, "per_gpu_batch": B,
            "gaze_method": gaze_method, "dp_method": dp_method, "rgb": bool(rgb), "steps": steps,
            "cuda_graph": True, "loss_total": losses[2]}


def load_tc_traffic(args, B):
    """dram__bytes_read.sum + dram__bytes_write.sum of every tc_gemm_kernel launch of ONE step, from the
    committed ncu pass (profiles/r2c_tc_traffic.json, else the older ones); only quoted for the configuration it was taken on."""
    try:
        name = next(n for n in ("r2c_tc_traffic.json", "r2b_tc_traffic.json", "r2_tc_traffic.json", "r1_tc_traffic.json")
                    if os.path.exists(os.path.join(ROOT, "profiles", n)))
        with open(os.path.join(ROOT, "profiles", name)) as f:
            t = json.load(f)
        if B == 1024 and not args.rgb and args.gaze_method == "Reg" and args.dp_method == "None":
            return t["dram_bytes_per_step"], t["source"]
    except Exception:
        pass
    return None, None


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p["hbm_gbs"], p["bf16_tflops"], p["bf16_tflops_sustained"], "measured"
    except Exception:
        return 6650.0, 1590.0, 1400.0, "fallback"


def main_ours(args):
    import torch
    import torch.distributed as dist
    import gabril_carla_b200  # noqa: F401
    from gabril_carla_b200 import _lib, ddp, ops
    from gabril_carla_b200.engine import HostFeeder, StepEngine
    from gabril_carla_b200.gaze import GazePreprocessor
    from gabril_carla_b200.model import ParamArena, PolicySpec
    from tests.synth import make_batch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl ours) needs a CUDA device: there is no CPU fallback")
    is_dist, rank, world, local_rank, device = ddp.init_distributed_if_enabled()
    torch.cuda.set_device(device)
    if args.per_gpu_batch:
        B, scaling = args.per_gpu_batch, "weak"
    else:
        lo, hi = ddp.shard_range(args.global_batch, rank, world)
        B, scaling = hi - lo, "strong"
    gray = not args.rgb
    spec = PolicySpec(2 if gray else 6)
    arena = ParamArena(spec, device)
    arena.init_reference_style(seed=42)
    ddp.broadcast_arena(arena.p)
    arena.refresh_shadow()
    pre = GazePreprocessor(device=str(device))
    eng = StepEngine(spec, arena, pre, gaze_method=args.gaze_method, dp_method=args.dp_method,
                     grayscale=gray, seed=1234 + rank)
    red = ddp.attach(eng, arena)
    # CUDA-graph replays of the step: on by default (GABRIL_GRAPH=0 / --no-graph for eager launches)
    eng.use_graph = os.environ.get("GABRIL_GRAPH", "1") != "0"
    if args.graph:
        eng.use_graph = True
    if args.no_graph:
        eng.use_graph = False

    # synthetic batch of this rank's shard: B distinct samples, pinned on the host
    host = synth_batch(B, seed=100 + 1000 * rank)
    dev_batch = tuple(t.to(device) for t in host)
    h2d = sum(t.numel() * t.element_size() for t in host)
    lr, step_no = 5e-4, [0]
    launches = [0]

    def step(batch):
        step_no[0] += 1
        eng.train_step(*batch)      # zero_grad + forward + backward (CUDA-graph replays when enabled)
        red.finish()
        eng.adam_step(step_no[0], lr, grad_scale=red.grad_scale)

    def barrier():
        if is_dist:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- count our launches in one step (every traced call is one kernel launch of ours, the
    #      GMD mask call is three) and collect the roofline evidence ----
    for _ in range(max(3, args.warmup)):
        step(dev_batch)
    barrier()
    # Clock spin-up: the first process on a fresh box starts on a GPU that has not reached its steady clocks -
    # measured on one box: per-launch event times 5.7 % longer in the first process (traced ~50 ms after its
    # first kernel) than in the next process of the same build.  ~1.5 s of untimed steps before any measurement;
    # every rank runs the same count (the steps contain collectives).
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(5):
        step(dev_batch)
    ev1.record()
    barrier()
    spin_ms = torch.tensor([ev0.elapsed_time(ev1) / 5.0], device=device)
    if is_dist:
        dist.all_reduce(spin_ms, op=dist.ReduceOp.MAX)
    n_spin = int(min(2000, max(0.0, SPINUP_S * 1e3 / max(float(spin_ms.item()), 0.05))))
    for _ in range(n_spin):
        step(dev_batch)
    barrier()
    # the traced step runs eagerly and serially (no side stream), so that every launch is timed alone
    overlap, eng.overlap = eng.overlap, False
    # three traced steps, per-launch median: a single step's event times wander by a few per cent with the
    # clocks (sw_power_cap), and the roofline fraction below is the ratio of two of them
    passes = []
    for _ in range(3):
        with ops.trace() as tr:
            step(dev_batch)
        passes.append(tr.rows())
    rows = passes[0]
    if all(len(q) == len(rows) for q in passes):
        rows = [r[:4] + (sorted(q[i][4] for q in passes)[1],) for i, r in enumerate(rows)]
    eng.overlap = overlap
    launches_per_step = len(rows) + (2 if args.dp_method == "GMD" else 0)
    TC_KERNELS = ("tc_gemm_kernel", "res_block_fwd_kernel")   # the tcgen05 / TMEM / TMA GEMM kernels
    tc_ms = sum(r[4] for r in rows if r[0] in TC_KERNELS)
    tc_flop = sum(r[2] for r in rows if r[0] in TC_KERNELS)
    traced_ms = sum(r[4] for r in rows)

    # ---- timed region 1: device-resident inputs ----
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t_host = time.perf_counter()
    for _ in range(args.steps):
        step(dev_batch)
    host_enqueue_ms = (time.perf_counter() - t_host) * 1e3 / args.steps   # no sync inside: pure enqueue
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    losses = eng.losses_host()

    # ---- timed region 2: end to end from pinned host memory, losses read back every step ----
    e2e_ms = None
    d2h = 12
    if not args.no_e2e:
        feeder = HostFeeder(device)
        out_host = torch.empty(3, dtype=torch.float32).pin_memory()
        for _ in range(3):  # warm the staging buffers
            k = feeder.stage(*host); step(feeder.wait(k)); feeder.release(k)
        barrier()
        e0.record()
        nxt = feeder.stage(*host)
        for i in range(args.steps):
            k = nxt
            if i + 1 < args.steps:
                nxt = feeder.stage(*host)       # H2D of step i+1 overlaps the kernels of step i
            step(feeder.wait(k))
            feeder.release(k)
            out_host.copy_(eng.buf["losses"], non_blocking=False)  # D2H read of the step's losses
        e1.record()
        barrier()
        e2e_ms = e0.elapsed_time(e1) / args.steps
    # ---- timed region 3: HBM-resident demonstration store, a fresh random batch of
    #      step indices every step, frames read through the window table (no batch tensor, no H2D) ----
    store_ms = None
    if not args.no_store_feed:
        from gabril_carla_b200.data import DemoStore
        n_store = args.store_steps
        g = torch.Generator(device=device).manual_seed(7 + rank)
        imgs = torch.randint(0, 256, (n_store, spec.H, spec.W, 3), device=device, dtype=torch.uint8, generator=g)
        gz = torch.rand(n_store, 10, device=device, generator=g)
        gz[torch.rand(n_store, 10, device=device, generator=g) < 0.2] = -1.0
        store = DemoStore(imgs, gz, torch.rand(n_store, spec.A, device=device, generator=g),
                          torch.arange(0, n_store, 250), eng.S).to(device)
        idxs = [torch.randint(0, n_store, (B,), device=device, generator=g) for _ in range(args.steps)]

        def store_step(idx):
            step_no[0] += 1
            eng.train_step_store(store, idx)   # graph replays: the step indices are staged into a static buffer
            red.finish()
            eng.adam_step(step_no[0], lr, grad_scale=red.grad_scale)

        for i in range(3):
            store_step(idxs[i % len(idxs)])
        barrier()
        e0.record()
        for i in range(args.steps):
            store_step(idxs[i])
        e1.record()
        barrier()
        store_ms = e0.elapsed_time(e1) / args.steps
        del store, imgs
    sampler.stop_flag = True
    sampler.join(timeout=2)
    _lib.poll_error()

    t = torch.tensor([ms, e2e_ms or 0.0, store_ms or 0.0], device=device, dtype=torch.float64)
    if is_dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = t[0].item(), (t[1].item() if e2e_ms is not None else None)
    store_ms = t[2].item() if store_ms is not None else None
    if is_dist:
        dist.barrier()   # every rank leaves the measured part together; only rank 0 reports
    if rank != 0:
        return
    hbm, tf_burst, tf_sust, how = load_peaks()
    total = B * world
    achieved = tc_flop / (tc_ms * 1e-3) / 1e12 if tc_ms > 0 else 0.0
    traffic, traffic_src = load_tc_traffic(args, B)
    line = {
        "metric": "GABRIL train samples/sec (320x180 stack, fwd+bwd+gaze reg)",
        "value": total / (ms * 1e-3), "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(args, B, world),
        "clocks": sampler.summary(),
        "gpu_launches": launches_per_step * args.steps,
        "cuda_graph": bool(eng.use_graph),
        "host_enqueue_ms_per_step": host_enqueue_ms,   # rank 0; close to ms_per_step = launch-bound
        "e2e": {"value": (total / (e2e_ms * 1e-3)) if e2e_ms else None, "unit": "samples/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "note": "pinned host batch -> double-buffered H2D on a copy stream -> step -> losses D2H, per rank"},
        "roofline": {"bound": "tensor", "kernel": "tcgen05 GEMM kernels: tc_gemm_kernel (every conv / linear fwd, dgrad, wgrad launch of one step, "
                                                       "bias gradients included) + res_block_fwd_kernel (the two Residual blocks' forward)",
                     "achieved": achieved, "peak": tf_sust, "unit": "TFLOP/s", "frac": achieved / tf_sust,
                     "peak_source": f"bf16_tflops_sustained, {how}", "traffic": traffic,
                     "traffic_unit": "bytes per step (all launches of those kernels), like achieved",
                     "traffic_source": (f"CONSTANT from a committed ncu pass, not measured in this run: {traffic_src}"
                                        if traffic_src else None),
                     "algorithmic_flop_per_step": tc_flop,
                     "tc_ms_per_step": tc_ms, "all_traced_ms_per_step": traced_ms,
                     "step_flop_frac_of_peak": FLOP_PER_SAMPLE[gray] * B / (ms * 1e-3) / 1e12 / tf_sust},
        "losses_last_step": {"actor": losses[0], "reg": losses[1], "total": losses[2]},
    }
    if store_ms:
        line["store_feed"] = {"value": total / (store_ms * 1e-3), "unit": "samples/s", "ms_per_step": store_ms,
                              "store_steps": args.store_steps, "h2d_bytes_per_step": 0,
                              "note": "HBM-resident uint8 frame store, random step indices per step, "
                                      "StepEngine.train_step_store (window table + frame kernel reading the store)"}
    # ---- the other BASELINE.json configurations on one GPU (short runs; the headline stays cfg2 / 1024) ----
    if world == 1 and not args.no_extra:
        extra = {}
        for key, (gm, dpm, rgb, bb) in {"cfg3_gmd": ("Reg", "GMD", False, B), "cfg1_bc": ("None", "None", False, B),
                                       "rgb": (args.gaze_method, args.dp_method, True, B),
                                       "b128": (args.gaze_method, args.dp_method, False, 128),
                                       "b512": (args.gaze_method, args.dp_method, False, 512),
                                       "b2000": (args.gaze_method, args.dp_method, False, 2000)}.items():
            try:
                extra[key] = quick_run(device, gm, dpm, rgb, bb)
            except Exception as e:  # noqa: BLE001
                extra[key] = {"error": repr(e)[:200]}
        line["extra_configs"] = extra
    if args.trace_out:
        agg = {}
        for k, lab, work, unit, t_ms in rows:
            a = agg.setdefault((k, lab, unit), [0, 0.0, 0.0])
            a[0] += 1; a[1] += work; a[2] += t_ms
        with open(args.trace_out, "w") as f:
            f.write("kernel,label,launches,work,unit,ms,rate(TFLOP/s|GB/s),share_of_traced\n")
            for (k, lab, unit), (n, work, t_ms) in sorted(agg.items(), key=lambda kv: -kv[1][2]):
                rate = work / (t_ms * 1e-3) / (1e12 if unit == "flop" else 1e9) if t_ms > 0 else 0
                f.write(f"{k},{lab},{n},{work:.4g},{unit},{t_ms:.4f},{rate:.1f},{t_ms / traced_ms:.4f}\n")
    if not args.no_cpu_baseline and world == 1:
        r = cpu_reference_run(args, 5, 2)
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        main_reference(a)
    else:
        main_ours(a)
    try:
        from gabril_carla_b200 import ddp as _d
        _d.destroy_distributed_if_initialized()
    except Exception:
        pass
