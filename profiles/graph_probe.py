"""Is bench.py's multi-stream step host-bound?  Times eager launches vs CUDA-graph replays of DpvLcPipeline.step with S
batches in flight.  Usage: python profiles/graph_probe.py [S] [steps]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from depthestimation_b200.pipeline import DpvLcPipeline, synthetic_batch  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 6
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 24
dev = torch.device("cuda", 0)
host, meta = synthetic_batch("C2_mono_lc_b8", seed=0)
x = {k: v.to(dev) for k, v in host.items()}
pipes = [DpvLcPipeline(meta, dev) for _ in range(S)]
streams = [torch.cuda.Stream(dev) for _ in range(S)]
main = torch.cuda.current_stream(dev)


def region(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    for st in streams:
        st.wait_stream(main)
    t0 = time.perf_counter()
    for i in range(steps):
        with torch.cuda.stream(streams[i % S]):
            fn(i % S)
    t_host = time.perf_counter() - t0
    for st in streams:
        main.wait_stream(st)
    e1.record(main)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, t_host * 1e3 / steps


for i in range(S):
    with torch.cuda.stream(streams[i]):
        pipes[i].step(x)
torch.cuda.synchronize()
print("eager : %.3f ms/step on the GPU, host issue %.3f ms/step" % region(lambda i: pipes[i].step(x)))
graphs, outs = [], []
for i in range(S):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=streams[i]):
        outs.append(pipes[i].step(x))
    graphs.append(g)
torch.cuda.synchronize()
ref = pipes[0].step(x)["logp_lc"].clone()
graphs[0].replay()
torch.cuda.synchronize()
print("graph replay == eager:", bool(torch.equal(outs[0]["logp_lc"], ref)))
print("graphs: %.3f ms/step on the GPU, host issue %.3f ms/step" % region(lambda i: graphs[i].replay()))
print("graphs: %.3f ms/step on the GPU, host issue %.3f ms/step" % region(lambda i: graphs[i].replay()))
