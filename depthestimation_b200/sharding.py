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
    """Keeps the per-frame result maps of the steps a rank runs and brings every rank's maps to every rank (the
    path's only exchange).  No collective sits between the steps: each step's maps are copied into a
    rank-local buffer on the step's own stream, and -- on CUDA -- an all_gather of just that step's slice is issued
    on a side stream behind an event, so it runs while the following steps compute instead of as
    one 2 GB exchange after the last step.  finish() waits for the side stream.  On CPU tensors (gloo tests) the
    exchange is a single all_gather in finish().  Every rank must add() the same steps in the same order."""

    def __init__(self, n_steps, world, keys=("depth", "depth_lc"), overlap=True):
        self.n_steps, self.world, self.keys = n_steps, world, keys
        self.buf, self.full = {}, {}
        self.count = 0
        self.overlap = overlap
        self.comm = None                # side stream of the per-step gathers

    def _alloc(self, k, t):
        self.buf[k] = torch.empty((self.n_steps,) + tuple(t.shape), device=t.device, dtype=t.dtype)
        if self.world > 1 and t.is_cuda and self.overlap:
            # [step, rank, ...]: one step's gather writes one contiguous block
            self.full[k] = torch.empty((self.n_steps, self.world) + tuple(t.shape), device=t.device, dtype=t.dtype)
            if self.comm is None:
                self.comm = torch.cuda.Stream(t.device)

    def reserve(self, out):
        """Allocates every buffer for outputs shaped like `out` now (so that no cudaMalloc -- which synchronises the
        whole device -- happens inside a timed region) and returns self."""
        for k in self.keys:
            t = out.get(k)
            if t is not None and k not in self.buf:
                self._alloc(k, t)
        return self

    def reset(self):
        """Forget the recorded steps, keep the buffers."""
        self.count = 0
        return self

    def add(self, out):
        """Copies this step's maps into the rank-local buffer (on the current stream) and queues their gather."""
        i = self.count % self.n_steps
        self.count += 1
        got = []
        for k in self.keys:
            t = out.get(k)
            if t is None:
                continue
            if k not in self.buf:
                self._alloc(k, t)
            self.buf[k][i].copy_(t, non_blocking=True)
            got.append(k)
        if self.comm is not None and got:
            ready = torch.cuda.Event()
            ready.record()                                             # on the step's stream, after the copies
            with torch.cuda.stream(self.comm):
                self.comm.wait_event(ready)
                for k in got:
                    dist.all_gather_into_tensor(self.full[k][i], self.buf[k][i])

    def finish(self):
        """-> {key: [world, n_steps, B, H, W]} on every rank; rank-local view if world = 1."""
        res = {}
        if self.comm is not None:
            torch.cuda.current_stream(self.comm.device).wait_stream(self.comm)
        for k, t in self.buf.items():
            if self.world <= 1:
                res[k] = t.unsqueeze(0)
            elif k in self.full:
                res[k] = self.full[k].transpose(0, 1)                  # [world, n_steps, ...] (a view)
            else:
                full = torch.empty((self.world * t.shape[0],) + tuple(t.shape[1:]), device=t.device, dtype=t.dtype)
                dist.all_gather_into_tensor(full, t)                   # concatenated along dim 0 (gloo needs this form)
                res[k] = full.view((self.world,) + tuple(t.shape))
        return res


def pin_rank_to_cores(local_rank, ranks_on_host):
    """Gives each rank of a host its own slice of the cores the process may run on (launch threads of eight ranks
    otherwise migrate over the same cores).  Returns the cores it now runs on (None where unsupported)."""
    try:
        cores = sorted(os.sched_getaffinity(0))
    except AttributeError:
        return None
    per = len(cores) // max(1, ranks_on_host)
    if per < 1:
        return cores
    mine = cores[local_rank * per:(local_rank + 1) * per]
    try:
        os.sched_setaffinity(0, mine)
    except OSError:
        return cores
    return mine
