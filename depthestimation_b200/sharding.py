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
            # NCCL's own banner / debug lines ("NCCL version ...", printed when the environment sets NCCL_DEBUG) go to
            # stderr: a caller's stdout (bench.py: ONE JSON line) stays its own
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
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
    rank-local buffer on the step's own stream.  On CUDA the exchange is cut into rounds of `chunk` steps: once all
    steps of a round have been added, ONE all_gather of that round's slice is issued on a side stream behind their
    events, so it runs while the next rounds compute and only the last round's exchange remains after the last step.
    Why rounds and not single steps (measured on 8 x B200, C2): a collective kernel holds its SMs and spins until every
    peer has launched the matching one; with several batches in flight steps complete out of order and differently on
    every rank, and 40 per-step gathers cost 8 % (efficiency 0.92), one gather after the last step 4 % (0.958), none at
    all 0.8 % (0.992).  A round is issued when every stream has finished its step, i.e. when the ranks are nearly in
    phase.  chunk = 0: a single all_gather in finish().  On CPU tensors (gloo tests) always a single all_gather in
    finish().  Every rank must add() the same steps in the same order."""

    def __init__(self, n_steps, world, keys=("depth", "depth_lc"), chunk=0):
        self.n_steps, self.world, self.keys = n_steps, world, keys
        self.buf, self.full = {}, {}
        self.count = 0
        self.chunk = int(chunk) if chunk and chunk < n_steps else 0
        self.comm = None                # side stream of the per-round gathers
        self.pending = []               # events of the steps added since the last gather
        self.done = 0                   # steps already handed to a gather

    def _alloc(self, k, t):
        self.buf[k] = torch.empty((self.n_steps,) + tuple(t.shape), device=t.device, dtype=t.dtype)
        if self.world > 1 and t.is_cuda:
            # [rank, step, ...] as the result wants it; a round's gather lands in a staging block [rank, chunk, ...]
            self.full[k] = torch.empty((self.world, self.n_steps) + tuple(t.shape), device=t.device, dtype=t.dtype)
            if self.chunk:
                self.stage = getattr(self, "stage", {})
                self.stage[k] = torch.empty((self.world, self.chunk) + tuple(t.shape), device=t.device, dtype=t.dtype)
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
        self.count, self.done, self.pending = 0, 0, []
        return self

    def _gather_round(self, lo, hi):
        """all_gather of steps [lo, hi) on the side stream, behind the events of those steps."""
        with torch.cuda.stream(self.comm):
            for ev in self.pending:
                self.comm.wait_event(ev)
            for k in self.buf:
                if k not in self.full:
                    continue
                n = hi - lo
                stage = self.stage[k][:, :n] if n == self.chunk else torch.empty(
                    (self.world, n) + tuple(self.buf[k].shape[1:]), device=self.buf[k].device, dtype=self.buf[k].dtype)
                if n == self.chunk:
                    dist.all_gather_into_tensor(self.stage[k], self.buf[k][lo:hi])
                else:
                    dist.all_gather_into_tensor(stage, self.buf[k][lo:hi])
                self.full[k][:, lo:hi].copy_(stage, non_blocking=True)      # [rank, round] -> [rank, step]
        self.pending = []
        self.done = hi

    def add(self, out):
        """Copies this step's maps into the rank-local buffer (on the current stream); closes a round when it is full."""
        i = self.count % self.n_steps
        self.count += 1
        got = False
        for k in self.keys:
            t = out.get(k)
            if t is None:
                continue
            if k not in self.buf:
                self._alloc(k, t)
            self.buf[k][i].copy_(t, non_blocking=True)
            got = True
        if self.comm is not None and got:
            ev = torch.cuda.Event()
            ev.record()                                                # on the step's stream, after the copies
            self.pending.append(ev)
            if self.count - self.done >= self.chunk:
                self._gather_round(self.done, self.count)

    def finish(self):
        """-> {key: [world, n_steps, B, H, W]} on every rank; rank-local view if world = 1."""
        res = {}
        if self.comm is not None:
            if self.count > self.done:
                self._gather_round(self.done, self.count)             # a last, partial round
            torch.cuda.current_stream(self.comm.device).wait_stream(self.comm)
        for k, t in self.buf.items():
            if self.world <= 1:
                res[k] = t.unsqueeze(0)
            elif self.comm is not None and k in self.full:
                res[k] = self.full[k]
            else:
                full = self.full[k].view((self.world * t.shape[0],) + tuple(t.shape[1:])) if k in self.full else \
                    torch.empty((self.world * t.shape[0],) + tuple(t.shape[1:]), device=t.device, dtype=t.dtype)
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
