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
