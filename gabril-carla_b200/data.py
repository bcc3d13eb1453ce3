"""Batch source for the trainer: Bench2Drive-shaped demonstrations in the batch-dict format
BCTrainer.compute_loss consumes (vlm_gaze/train/train_bc.py:204-207):

    {'obs': {'image': u8 [B,L,H,W,3], <gaze_key>: f32 [B,L,P*2]}, 'actions': f32 [B,L,A]}

with L = frame_stack (robomimic SequenceDataset with frame_stack=S, seq_length=1 returns the S
frames ending at the sampled step, the first frame of a demo repeated as front padding:
vlm_gaze/robomimic/utils/dataset.py:589-633).  The reference's HDF5 reader is out of scope
(SURVEY.md §8f N2); here the store is a set of flat tensors — synthetic ("synthetic:N"), or loaded
from an .npz/.pt file with arrays `images` u8 [N,H,W,3], `gaze` f32 [N,P*2], `actions` f32 [N,A] and
`demo_start` int64 [D] — kept resident in HBM when it fits so that a batch is one index gather on
the device instead of a CPU DataLoader + H2D copy.
"""
import math
import os
from typing import Dict, Iterator, List, Optional

import torch


class DemoStore:
    def __init__(self, images: torch.Tensor, gaze: torch.Tensor, actions: torch.Tensor,
                 demo_start: torch.Tensor, frame_stack: int, gaze_key: str = "gaze_coords"):
        assert images.dtype == torch.uint8 and images.dim() == 4 and images.shape[-1] == 3
        self.images, self.gaze, self.actions = images, gaze.float(), actions.float()
        self.S, self.gaze_key = int(frame_stack), gaze_key
        n = images.shape[0]
        # first step index of the demo every step belongs to (for the front padding)
        starts = torch.zeros(n, dtype=torch.long)
        ds = demo_start.long().clamp(0, max(0, n - 1))
        starts[ds] = ds
        self.first = torch.cummax(starts, 0).values
        self.n = n

    def __len__(self):
        return self.n

    def to(self, device):
        for k in ("images", "gaze", "actions", "first"):
            setattr(self, k, getattr(self, k).to(device))
        return self

    def pin(self):
        if torch.cuda.is_available():
            for k in ("images", "gaze", "actions"):
                setattr(self, k, getattr(self, k).pin_memory())
        return self

    @property
    def device(self):
        return self.images.device

    def window_indices(self, idx: torch.Tensor) -> torch.Tensor:
        """[B] step indices -> [B,S] frame indices: steps idx-S+1..idx clamped to the demo start."""
        idx = idx.to(self.first.device)
        offs = torch.arange(-(self.S - 1), 1, device=idx.device)
        return torch.maximum(idx[:, None] + offs[None, :], self.first[idx][:, None])

    def batch(self, idx: torch.Tensor) -> Dict:
        w = self.window_indices(idx)
        B, S = w.shape
        flat = w.reshape(-1)
        return {"obs": {"image": self.images[flat].view(B, S, *self.images.shape[1:]),
                        self.gaze_key: self.gaze[flat].view(B, S, -1)},
                "actions": self.actions[flat].view(B, S, -1)}


def synthetic_store(n: int, frame_stack: int, H: int = 180, W: int = 320, P: int = 5, A: int = 7,
                    seed: int = 0, demo_len: int = 256, gaze_key: str = "gaze_coords") -> DemoStore:
    """Seeded Bench2Drive-shaped demonstrations: low-frequency RGB structure + noise, up to P gaze
    points per frame (-1 padded), actions in [0,1).  Built in chunks to bound host memory."""
    g = torch.Generator().manual_seed(seed)
    imgs = torch.empty(n, H, W, 3, dtype=torch.uint8)
    for lo in range(0, n, 256):
        hi = min(n, lo + 256)
        low = torch.rand(hi - lo, 3, 6, 10, generator=g)
        up = torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=False)
        noise = torch.randint(0, 56, (hi - lo, H, W, 3), generator=g, dtype=torch.uint8)
        imgs[lo:hi] = (up.permute(0, 2, 3, 1) * 200).to(torch.uint8) + noise
    gaze = torch.rand(n, P, 2, generator=g)
    nvalid = torch.randint(0, P + 1, (n,), generator=g)
    gaze[torch.arange(P)[None, :] >= nvalid[:, None]] = -1.0
    actions = torch.rand(n, A, generator=g)
    demo_start = torch.arange(0, n, demo_len)
    return DemoStore(imgs, gaze.reshape(n, P * 2), actions, demo_start, frame_stack, gaze_key)


def load_store(path: str, frame_stack: int, cfg_data=None, seed: int = 0) -> DemoStore:
    gaze_key = getattr(cfg_data, "gaze_key", "gaze_coords") if cfg_data is not None else "gaze_coords"
    H = getattr(cfg_data, "img_height", 180) if cfg_data is not None else 180
    W = getattr(cfg_data, "img_width", 320) if cfg_data is not None else 320
    A = getattr(cfg_data, "action_dim", 7) if cfg_data is not None else 7
    if path.startswith("synthetic"):
        n = int(path.split(":")[1]) if ":" in path else 4096
        return synthetic_store(n, frame_stack, H, W, 5, A, seed=seed, gaze_key=gaze_key)
    if path.endswith(".npz"):
        import numpy as np
        d = np.load(path)
        t = {k: torch.from_numpy(d[k]) for k in d.files}
    elif path.endswith((".pt", ".pth")):
        t = torch.load(path, weights_only=True)
    elif path.endswith((".hdf5", ".h5")):
        raise NotImplementedError(
            "robomimic HDF5 input is outside the B200 hot path (SURVEY.md §8f N2): convert the "
            "episodes to an .npz/.pt store (images, gaze, actions, demo_start)")
    else:
        raise ValueError(f"unknown dataset format: {path}")
    gz = t.get(gaze_key, t.get("gaze"))
    return DemoStore(t["images"], gz.reshape(gz.shape[0], -1), t["actions"],
                     t.get("demo_start", torch.zeros(1, dtype=torch.long)), frame_stack, gaze_key)


class StoreBatch(dict):
    """Batch of an HBM-resident store: {"store", "step_idx"}.  The reference's batch-dict keys
    ("obs", "actions") are materialised on first access, for callers that look inside a batch."""

    def __init__(self, store: DemoStore, step_idx: torch.Tensor):
        super().__init__(store=store, step_idx=step_idx)

    def __missing__(self, key):
        if key in ("obs", "actions"):
            self.update(self["store"].batch(self["step_idx"]))
            return dict.__getitem__(self, key)
        raise KeyError(key)


class ShardedBatches:
    """DistributedSampler(shuffle=True, drop_last=False) + DataLoader(batch_size=per-rank batch)
    semantics (base_trainer.py:88-92, data.py:62-71): a seeded permutation per epoch, padded by
    wrapping to a multiple of the world size, rank r takes indices r::world."""

    def __init__(self, store: DemoStore, batch_size: int, rank: int = 0, world: int = 1, seed: int = 0,
                 shuffle: bool = True, fused: bool = False):
        self.fused = bool(fused)
        self.store, self.bs, self.rank, self.world, self.seed, self.shuffle = store, int(batch_size), rank, world, seed, shuffle
        self.epoch = 0
        self.per_rank = math.ceil(len(store) / world)

    def set_epoch(self, epoch: int):
        self.epoch = epoch

    def __len__(self):
        return math.ceil(self.per_rank / self.bs)

    def indices(self) -> torch.Tensor:
        n = len(self.store)
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            perm = torch.randperm(n, generator=g)
        else:
            perm = torch.arange(n)
        total = self.per_rank * self.world
        if total > n:
            reps = math.ceil((total - n) / n)
            perm = torch.cat([perm] + [perm] * reps)[:total]
        return perm[self.rank:total:self.world]

    def __iter__(self) -> Iterator[Dict]:
        idx = self.indices().to(self.store.device)
        for lo in range(0, idx.numel(), self.bs):
            if self.fused and self.store.device.type == "cuda":
                # device-resident feed: the trainer hands (store, step indices) to
                # StepEngine.forward_store, no batch tensors are built
                yield StoreBatch(self.store, idx[lo:lo + self.bs].contiguous())
            else:
                yield self.store.batch(idx[lo:lo + self.bs])
