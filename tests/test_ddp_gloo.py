"""CPU (-m "not gpu"): the data-parallel host logic at world_size 2 over gloo — shard ranges, the
initial parameter broadcast, the bucketed SUM all-reduce of the flat gradient arena with the
1/world scale applied afterwards (DDP's averaged gradients, vlm_gaze/train/common/distributed.py:55-68),
and the shard semantics of SURVEY.md §8e: mean of per-rank gradients == gradient on the
concatenated batch for the non-GMD losses."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import gabril_carla_b200  # noqa: F401
from gabril_carla_b200 import ddp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      LOCAL_RANK=str(rank), WORLD_SIZE=str(world))
    try:
        torch.set_num_threads(2)
        is_dist, r, w, lr, dev = ddp.init_distributed_if_enabled()
        assert is_dist and r == rank and w == world and dev.type == "cpu"
        assert dist.get_backend() == "gloo"
        # --- initial broadcast: every rank ends with rank 0's parameters
        p = torch.full((1000,), float(rank + 1))
        ddp.broadcast_arena(p)
        assert torch.all(p == 1.0)
        # --- bucketed all-reduce of a flat arena, buckets in readiness order
        n = 1003
        g = torch.arange(n, dtype=torch.float32) * (rank + 1)
        red = ddp.GradReducer(g, [0, 400, n])
        red.reduce_bucket(0)
        red.reduce_bucket(1)
        red.finish()
        expect = torch.arange(n, dtype=torch.float32) * sum(range(1, world + 1))
        assert torch.equal(g, expect)
        assert red.grad_scale == 1.0 / world and red.bytes_reduced == 4 * n
        # --- flush(): every bucket at once (epoch-tail micro-steps of an unfinished accumulation window)
        g3 = torch.ones(10) * (rank + 1)
        red3 = ddp.GradReducer(g3, [0, 4, 10])
        red3.flush()
        red3.finish()
        assert torch.all(g3 == float(sum(range(1, world + 1)))) and red3.bytes_reduced == 40
        # --- attach(): the hooks are unconditional; the engine decides when a window closes
        class _Eng:
            on_heads_grads_ready = on_all_grads_ready = None
        class _Arena:
            g = torch.ones(8) * (rank + 1); n = 8
            @staticmethod
            def offset_of(name): return 3
        e = _Eng()
        red4 = ddp.attach(e, _Arena)
        e.on_heads_grads_ready(); e.on_all_grads_ready(); red4.finish()
        assert torch.all(_Arena.g == float(sum(range(1, world + 1)))) and red4.bytes_reduced == 32
        # --- shard semantics with the oracle's losses (checker only)
        from oracle import gabril_oracle as O
        from tests.synth import make_batch, make_weights
        B = 4
        img, gaze, act = make_batch(B, seed=3)
        wts = make_weights(2, seed=5)
        lo, hi = ddp.shard_range(B, rank, world)

        def grads(sl):
            ww = {m: {k: v.clone().requires_grad_(True) for k, v in sd.items()} for m, sd in wts.items()}
            total, *_ = O.bc_losses(ww["encoder"], ww["pre_actor"], ww["actor"], img[sl], gaze[sl], act[sl],
                                    gaze_method="Reg", dp_method="None")
            total.backward()
            return torch.cat([v.grad.flatten() for k, v in ww["actor"].items()] +
                             [ww["encoder"]["_conv_6.weight"].grad.flatten()])
        local = grads(slice(lo, hi))
        red2 = ddp.GradReducer(local, [0, local.numel() // 2, local.numel()])
        red2.reduce_bucket(0)
        red2.reduce_bucket(1)
        red2.finish()
        local.mul_(red2.grad_scale)
        full = grads(slice(0, B))
        err = ((local - full).norm() / full.norm()).item()
        assert err <= 1e-5, err
        ddp.destroy_distributed_if_initialized()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))


@pytest.mark.timeout(300)
def test_world2_gloo_reducer_and_shard_semantics():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=280) for _ in range(world)]
    for p in procs:
        p.join(timeout=30)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_shard_range():
    assert ddp.shard_range(1024, 0, 8) == (0, 128) and ddp.shard_range(1024, 7, 8) == (896, 1024)
    assert [ddp.shard_range(1024, r, 2) for r in range(2)] == [(0, 512), (512, 1024)]
    with pytest.raises(ValueError):
        ddp.shard_range(1000, 0, 3)


def test_single_process_is_not_distributed(monkeypatch):
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    is_dist, rank, world, lr, dev = ddp.init_distributed_if_enabled()
    assert (is_dist, rank, world, lr) == (False, 0, 1, 0)
