"""Multi-GPU host logic on CPU: world_size-2 gloo run of the shard / gather plumbing (sharding.py)."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_items, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    from depthestimation_b200 import sharding
    r, w, _ = sharding.init_from_env(world, backend="gloo")
    lo, hi = sharding.shard_range(n_items, r, w)
    frames = torch.arange(lo, hi, dtype=torch.float32).reshape(-1, 1, 1).repeat(1, 2, 3)   # "depth maps" of my shard
    # gloo has no all_gather_into_tensor for ragged shards: pad to the largest shard like the GPU path's equal B
    per = (n_items + w - 1) // w
    pad = torch.full((per, 2, 3), -1.0)
    pad[:hi - lo] = frames
    out = sharding.gather_frame_outputs({"depth": pad, "depth_lc": None}, w, keys=("depth", "depth_lc"))
    # the end-of-run exchange used by bench.py: 3 steps of per-rank maps, one all_gather
    coll = sharding.FrameOutputCollector(3, w, keys=("depth", "depth_lc"))
    for step in range(3):
        coll.add({"depth": pad + 100.0 * step + 1000.0 * r, "depth_lc": None})
    full = coll.finish()
    q.put((r, lo, hi, out["depth"][:, 0, 0].tolist(), sorted(out.keys()),
           tuple(full["depth"].shape), torch.stack([coll.at(full["depth"], s)[:, 0, 0, 0] for s in range(3)], 1).tolist(),
           sorted(full.keys())))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_and_gather_world2():
    world, n_items = 2, 7
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, world, port, n_items, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    # shards are contiguous, disjoint and cover everything
    assert [(r[1], r[2]) for r in res] == [(0, 4), (4, 7)]
    for r in res:
        got = [v for v in r[3] if v >= 0]
        assert got == [float(i) for i in range(n_items)]        # every rank holds every frame, in shard order
        assert r[4] == ["depth"]
        assert r[5] == (world, 1, 3, 4, 2, 3) and r[7] == ["depth"]       # [rank, round, step in round, frames, H, W]
        # [rank, step] -> first frame of that rank's shard + 100*step + 1000*rank
        assert r[6] == [[0.0, 100.0, 200.0], [1004.0, 1104.0, 1204.0]]


def test_shard_range_properties():
    sys.path.insert(0, ROOT)
    from depthestimation_b200 import sharding
    for n in (0, 1, 7, 64):
        for w in (1, 2, 3, 8):
            spans = [sharding.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1
