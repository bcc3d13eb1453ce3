"""Data-parallel plumbing: one process per GPU, gradients summed over ranks with bucketed
all-reduces of the flat gradient arena, overlapped with the rest of backward.

Replaces the reference's three DistributedDataParallel reducers (vlm_gaze/train/common/
distributed.py:55-68, applied at train_bc.py:116-123) and its process-group bootstrap
(distributed.py:11-47).  Because the arena is laid out in reverse execution order, the first
bucket [actor, pre_actor] (50 MB, ready right after the heads' backward) travels over NVLink
while the encoder backward is still computing; the encoder bucket follows at the end.  The
1/world_size averaging is folded into the fused Adam kernel (grad_scale).
"""
import os
from typing import List, Tuple

import torch
import torch.distributed as dist


def init_distributed_if_enabled(cfg_training=None) -> Tuple[bool, int, int, int, torch.device]:
    """(is_distributed, rank, world_size, local_rank, device) — torchrun env wins over the config
    flag, exactly like distributed.py:24-43."""
    env_has_dist = all(k in os.environ for k in ("LOCAL_RANK", "RANK", "WORLD_SIZE"))
    if env_has_dist:
        backend = "nccl"
        if cfg_training is not None and hasattr(cfg_training, "distributed"):
            backend = getattr(cfg_training.distributed, "backend", "nccl")
        if not torch.cuda.is_available():
            backend = "gloo"
        local_rank = int(os.environ["LOCAL_RANK"])
        if torch.cuda.is_available():
            torch.cuda.set_device(local_rank)
            device = torch.device(f"cuda:{local_rank}")
        else:
            device = torch.device("cpu")
        if not dist.is_initialized():
            kw = {"device_id": device} if device.type == "cuda" else {}
            dist.init_process_group(backend=backend, **kw)
        return True, dist.get_rank(), dist.get_world_size(), local_rank, device
    dev = "cuda:0"
    if cfg_training is not None:
        dev = getattr(cfg_training, "device", dev)
    return False, 0, 1, 0, torch.device(dev if torch.cuda.is_available() else "cpu")


def destroy_distributed_if_initialized():
    try:
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of a global batch for this rank (equal sizes required so that
    the mean of per-rank means equals the global mean)."""
    if n_items % world:
        raise ValueError(f"global batch {n_items} is not divisible by world size {world}")
    per = n_items // world
    return rank * per, (rank + 1) * per


def broadcast_arena(p: torch.Tensor, src: int = 0):
    """Initial parameter broadcast (DDP constructor semantics)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(p, src=src)


class GradReducer:
    """Bucketed SUM all-reduce of a flat gradient arena on a side stream."""

    def __init__(self, flat_grad: torch.Tensor, boundaries: List[int]):
        """boundaries: element offsets [0, b1, ..., n] delimiting the buckets in readiness order."""
        self.g = flat_grad
        self.bounds = boundaries
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.cuda = flat_grad.is_cuda
        self.comm = torch.cuda.Stream(device=flat_grad.device) if self.cuda else None
        self.pending = []
        self.bytes_reduced = 0

    def reduce_bucket(self, i: int):
        """Call when every gradient in bucket i has been enqueued on the current stream."""
        if self.world == 1:
            return
        lo, hi = self.bounds[i], self.bounds[i + 1]
        view = self.g[lo:hi]
        self.bytes_reduced += view.numel() * view.element_size()
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.g.device))
            with torch.cuda.stream(self.comm):
                self.comm.wait_event(ev)
                self.pending.append(dist.all_reduce(view, op=dist.ReduceOp.SUM, async_op=True))
        else:
            self.pending.append(dist.all_reduce(view, op=dist.ReduceOp.SUM, async_op=True))

    def flush(self):
        """Reduce every bucket now (gradients of an accumulation window that did not close)."""
        for i in range(len(self.bounds) - 1):
            self.reduce_bucket(i)

    def finish(self):
        """Order the optimizer after every outstanding bucket (stream-level wait, no host sync on CUDA)."""
        for w in self.pending:
            w.wait()
        self.pending = []

    @property
    def grad_scale(self) -> float:
        return 1.0 / self.world


def attach(engine, arena) -> GradReducer:
    """Wire a GradReducer into a StepEngine's gradient-ready hooks: bucket 0 = actor + pre_actor,
    bucket 1 = encoder."""
    split = arena.offset_of("enc.conv6.w")
    red = GradReducer(arena.g, [0, split, arena.n])

    # the engine calls the hooks only on the micro-step that closes an accumulation window (and on
    # every CUDA-graph replayed step), so they carry no condition of their own
    engine.on_heads_grads_ready = lambda: red.reduce_bucket(0)
    engine.on_all_grads_ready = lambda: red.reduce_bucket(1)
    return red
