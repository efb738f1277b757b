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
    finish().  Every rank must add() the same steps in the same order.

    Result layout: finish() -> {key: [world, rounds, steps_per_round, B, H, W]}; step s of rank r is
    res[key][r, s // steps_per_round, s % steps_per_round] (at(res[key], s) gives [world, B, H, W]).  It is a VIEW of
    round-major storage [rounds, world, steps_per_round, ...]: a round's all_gather writes its final place directly
    (a rank-major [world, n_steps, ...] result would need a second pass over everything received -- at N = 8 that
    copy was as much HBM traffic as the gather itself).  Entries past the last step of a partial last round are
    unspecified."""

    def __init__(self, n_steps, world, keys=("depth", "depth_lc"), chunk=0):
        self.n_steps, self.world, self.keys = n_steps, world, keys
        self.buf, self.full = {}, {}
        self.count = 0
        self.chunk = int(chunk) if chunk and chunk < n_steps else 0
        self.per_round = self.chunk if self.chunk else n_steps
        self.rounds = (n_steps + self.per_round - 1) // self.per_round
        self.comm = None                # side stream of the per-round gathers
        self.pending = []               # events of the steps added since the last gather
        self.done = 0                   # steps already handed to a gather

    def at(self, gathered, step):
        """[world, rounds, steps_per_round, ...] -> the maps of one step on every rank, [world, ...]."""
        return gathered[:, step // self.per_round, step % self.per_round]

    def _alloc(self, k, t):
        # rank-local maps, padded to whole rounds so that every round's slice is a full, contiguous block
        self.buf[k] = torch.empty((self.rounds * self.per_round,) + tuple(t.shape), device=t.device, dtype=t.dtype)
        if self.world > 1 and t.is_cuda:
            self.full[k] = torch.empty((self.rounds, self.world, self.per_round) + tuple(t.shape), device=t.device,
                                       dtype=t.dtype)
            if self.chunk and self.comm is None:
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

    def _gather_round(self, rnd):
        """all_gather of round `rnd` (its whole slice, also when the last round is only partly filled) on the side
        stream, behind the events of its steps, straight into its final place."""
        lo = rnd * self.per_round
        with torch.cuda.stream(self.comm):
            for ev in self.pending:
                self.comm.wait_event(ev)
            for k in self.buf:
                if k in self.full:
                    dist.all_gather_into_tensor(self.full[k][rnd], self.buf[k][lo:lo + self.per_round])
        self.pending = []
        self.done = min(lo + self.per_round, self.n_steps)

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
            if self.count - self.done >= self.per_round:
                self._gather_round(self.done // self.per_round)

    def finish(self):
        """-> {key: [world, rounds, steps_per_round, B, H, W]} on every rank (world = 1: the rank-local maps)."""
        res = {}
        if self.comm is not None:
            if self.count > self.done:
                self._gather_round(self.done // self.per_round)       # a last, partial round
            torch.cuda.current_stream(self.comm.device).wait_stream(self.comm)
        for k, t in self.buf.items():
            shape = (self.rounds, self.per_round) + tuple(t.shape[1:])
            if self.world <= 1:
                res[k] = t.view((1,) + shape)
            elif self.comm is not None and k in self.full:
                res[k] = self.full[k].permute(1, 0, *range(2, self.full[k].dim()))
            else:
                if k in self.full:                                     # CUDA, chunk = 0: one round
                    dist.all_gather_into_tensor(self.full[k][0], t)
                    res[k] = self.full[k].permute(1, 0, *range(2, self.full[k].dim()))
                else:                                                  # CPU tensors (gloo): concatenated along dim 0
                    flat = torch.empty((self.world * t.shape[0],) + tuple(t.shape[1:]), device=t.device, dtype=t.dtype)
                    dist.all_gather_into_tensor(flat, t)
                    res[k] = flat.view((self.world,) + shape)
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
