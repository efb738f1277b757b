"""Multi-GPU plumbing: one process per GPU, frames (trajectories) sharded by rank with no data-path collective;
the only exchange is the final gather of per-frame outputs (the reference's train.py:79-123 uses the
same one-process-per-GPU layout with torch.distributed)."""
import os

import torch
import torch.distributed as dist


def init_from_env(n_gpus=1, backend=None):
    """Reads RANK / WORLD_SIZE / LOCAL_RANK (torchrun) and initialises the process group when WORLD_SIZE > 1.
    Returns (rank, world, local_rank)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world, local


def shard_range(n_items, rank, world):
    """Contiguous shard of trajectories for this rank (never split inside a trajectory: prev_output / prev_lc
    recur per trajectory, default_trainer.py:247-249, 283-290)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_frame_outputs(out, world, keys=("depth", "depth_lc")):
    """all_gather of the per-frame result maps [B,H,W] -> [world*B,H,W] on every rank (NCCL over NVLink)."""
    if world <= 1:
        return out
    res = {}
    for k in keys:
        if k in out and out[k] is not None:
            t = out[k].contiguous()
            full = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), device=t.device, dtype=t.dtype)
            dist.all_gather_into_tensor(full, t)
            res[k] = full
    return res


class FrameOutputCollector:
    """Keeps the per-frame result maps of the steps a rank runs and exchanges them once at the end
    (north_star: "a single NCCL gather at the end"), so no collective sits between the steps of the path."""

    def __init__(self, n_steps, world, keys=("depth", "depth_lc")):
        self.n_steps, self.world, self.keys = n_steps, world, keys
        self.buf = {}
        self.count = 0

    def add(self, out):
        """Copies this step's maps into the rank-local buffer (on the current stream)."""
        i = self.count % self.n_steps
        for k in self.keys:
            t = out.get(k)
            if t is None:
                continue
            if k not in self.buf:
                self.buf[k] = torch.empty((self.n_steps,) + tuple(t.shape), device=t.device, dtype=t.dtype)
            self.buf[k][i].copy_(t, non_blocking=True)
        self.count += 1

    def finish(self):
        """-> {key: [world, n_steps, B, H, W]} on every rank (all_gather over NVLink); rank-local view if world = 1."""
        res = {}
        for k, t in self.buf.items():
            if self.world <= 1:
                res[k] = t.unsqueeze(0)
                continue
            full = torch.empty((self.world * t.shape[0],) + tuple(t.shape[1:]), device=t.device, dtype=t.dtype)
            dist.all_gather_into_tensor(full, t)                       # concatenated along dim 0 (gloo needs this form)
            res[k] = full.view((self.world,) + tuple(t.shape))
        return res
