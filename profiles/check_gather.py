"""torchrun check of FrameOutputCollector on real GPUs / NCCL: for several round sizes (0 = one gather at the end, 1 =
per step, 3 = leaves a partial last round, 5) every rank ends up with every rank's map of every step.
Usage: python -m torch.distributed.run --nproc-per-node N profiles/check_gather.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from depthestimation_b200 import sharding  # noqa: E402

rank, world, local = sharding.init_from_env()
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
n_steps, S = 10, 5
streams = [torch.cuda.Stream(dev) for _ in range(S)]
ok = True
for chunk in (0, 1, 3, 5):
    coll = sharding.FrameOutputCollector(n_steps, world, chunk=chunk)
    for rep in range(2):                                    # second pass reuses the buffers after reset()
        coll.reset()
        for i in range(n_steps):
            with torch.cuda.stream(streams[i % S]):
                m = torch.full((2, 4, 6), float(1000 * rank + 10 * i + rep), device=dev)
                torch.cuda._sleep(200000 * ((i * 7 + rank * 3) % 4))          # steps finish out of order
                coll.add({"depth": m, "depth_lc": m + 0.5})
        for st in streams:
            torch.cuda.current_stream(dev).wait_stream(st)
        res = coll.finish()
        torch.cuda.synchronize(dev)
        for k, add in (("depth", 0.0), ("depth_lc", 0.5)):
            got = res[k]                                    # [world, rounds, steps per round, ...] (a view)
            assert tuple(got.shape) == (world, coll.rounds, coll.per_round, 2, 4, 6), got.shape
            assert torch.equal(coll.at(got, 7), got[:, 7 // coll.per_round, 7 % coll.per_round])
            got = got.reshape(world, coll.rounds * coll.per_round, 2, 4, 6)[:, :n_steps]
            exp = torch.tensor([[1000.0 * r + 10 * i + rep + add for i in range(n_steps)] for r in range(world)], device=dev)
            good = bool((got[:, :, 0, 0, 0] == exp).all()) and bool((got == got[:, :, :1, :1, :1]).all())
            ok = ok and good
            if not good and rank == 0:
                print("MISMATCH chunk", chunk, k, got[:, :, 0, 0, 0].tolist())
flag = torch.tensor([1.0 if ok else 0.0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("check_gather:", "ok" if flag.item() == 1.0 else "FAILED", "world", world)
dist.destroy_process_group()
