#!/usr/bin/env python
"""bench.py — DPV + light-curtain-fusion hot path throughput (frames/sec) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]

One "step" = one pass of the whole hot path (plane sweep -> log-softmax -> warp_feature -> log-softmax(+resi) ->
full-res log-softmax+stats+carry -> 5 x [UF collapse, plan, DP, curtain planes, sense+fuse]) over one batch of
synthetic KITTI-shaped frames per GPU (BASELINE.json configs[1]: default_exp7_lc --lc, batch 8 per GPU).
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

_REAL_STDOUT = None


def emit(line):
    data = (line + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line + "\n")
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


METRIC = "DPV+LC-fusion frames/sec"
UNIT = "frames/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="C2_mono_lc_b8")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--launch", default="auto", choices=["auto", "graph", "eager"],
                    help="value loop: replay one captured CUDA graph per batch slot, or launch kernel by kernel; auto = "
                         "graphs (a step is ~80 launches for ~3.4 ms of GPU work: replay measured 3.35 ms per step at "
                         "N = 1 against 3.41-3.53 ms kernel by kernel, and several ranks on one host need it)")
    ap.add_argument("--hit-noise", type=float, default=0.01,
                    help="sigma of the simulator's intensity noise (the reference hard-codes 0.01, lc/light_curtain.py:270-285)")
    ap.add_argument("--gather", default="round", choices=["round", "step", "end", "none"],
                    help="exchange of the per-frame depth maps between ranks: one all_gather per round of S steps on a "
                         "side stream (default), one per step, one after the last step, or none (diagnostic only: the "
                         "path requires the exchange)")
    ap.add_argument("--planner-pack", type=int, default=1, choices=[1, 2],
                    help="maps per CTA pair of the DP planner in the throughput region (diagnostic: 2 halves the SMs a "
                         "solve holds and measured 1-5 %% lower throughput, DESIGN.md 6); the single-batch pass uses 1")
    ap.add_argument("--streams", type=int, default=6,
                    help="independent batches in flight (steps are issued round-robin on this many CUDA streams)")
    return ap.parse_args()


# --------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the same step (reference Python restated in oracle/, lcsim restated in C++)
# --------------------------------------------------------------------------------------------------
def oracle_step(batch, meta, lc_cache={}):
    """One frame-batch through the CPU oracle; returns the fused log-DPV.  Same inputs as DpvLcPipeline.step."""
    from oracle import dpv_oracle as O
    from oracle import lc_oracle as LO
    d = meta["d_candi"]
    B = batch["feat_all"].shape[0]
    outs = []
    for b in range(B):                                    # the reference is serial over the batch (models.py:617,984)
        K, rays = batch["K"][b], batch["rays"][0]
        cost = O.plane_sweep_cost(batch["feat_all"][b:b + 1, -1], batch["feat_all"][b:b + 1, :-1], d,
                                  batch["R"][b, :-1], batch["t"][b, :-1], K, rays, 10.0, "L2")
        bv = O.logsoftmax_stats(cost, d)[0]
        if meta.get("warp_feature", True):
            O.warp_feature_diag(batch["feat_raw"][b:b + 1], d, batch["R"][b], batch["t"][b], K, rays)
        O.logsoftmax_stats(bv, d, batch["bv_resi"][b:b + 1])
        logp, mean, var, ent = O.logsoftmax_stats(batch["dec_logits"][b:b + 1], d)
        O.carry_quarter(logp)
        if meta["iterations"] > 0:
            key = (meta["H"], meta["W"], meta["D"], meta["R"])
            if key not in lc_cache:
                lc_cache[key] = LO.LightCurtainOracle(meta["K_up"], meta["W"], meta["H"], d, meta["d_candi_up"])
                lc_cache[key].flow()
            depth = np.repeat(np.repeat(batch["dmaps"][b].numpy(), 4, 0), 4, 1).astype(np.float32)
            logp = LO.lc_loop(logp, depth, lc_cache[key], meta["K_up"], O.uf_params(meta["dataset"]),
                              meta["iterations"], meta["step"], meta["spread_iter"], meta["spread_n"])
        outs.append(logp)
    return torch.cat(outs)


def time_cpu(workload, steps, warmup, frames_per_step=1):
    from depthestimation_b200.pipeline import synthetic_batch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    batch, meta = synthetic_batch(workload, seed=0, B=frames_per_step)
    for _ in range(max(warmup, 0)):
        oracle_step(batch, meta)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle_step(batch, meta)
    dt = (time.perf_counter() - t0) / steps
    return frames_per_step / dt, dt, cores, meta


class ClockSampler:
    """SM clock and throttle reasons while the GPU runs the step (B200_PROFILING.md's clocks line: `nvidia-smi -lms 200`
    started BEFORE the timed region and killed after it, so that its start-up — it attaches to every GPU of the box —
    stays out of the region).  start() launches it, wait_ready() returns once the first row is in, rows are
    time-stamped on arrival; summary(t0, t1) uses the rows that arrived in [t0, t1] (the caller keeps the GPU under
    the same load until at least one has).  DPV_BENCH_CLOCKS=off disables it."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.p, self.t = index, [], None, None

    def start(self):
        if os.environ.get("DPV_BENCH_CLOCKS", "smi") == "off":
            return self
        try:
            self.p = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.p = None
        return self

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def wait_ready(self, timeout=5.0):
        t_end = time.perf_counter() + timeout
        while self.p and not self.rows and time.perf_counter() < t_end:
            time.sleep(0.01)

    def count(self, t0, t1=None):
        return sum(1 for t, _ in self.rows if t >= t0 and (t1 is None or t <= t1))

    def stop(self):
        if self.p:
            self.p.terminate()
            self.t.join(timeout=2)

    def summary(self, t0=None, t1=None):
        rows = [r for t, r in self.rows if (t0 is None or t >= t0) and (t1 is None or t <= t1)]
        sm = [float(r[1]) for r in rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
def pick_slots(requested, steps):
    """Batches in flight: the largest count <= requested that divides the timed steps, so that every slot runs the
    same number of steps and the region has no tail in which the last batches run alone (>= 4 where possible)."""
    requested = max(1, min(requested, steps))
    for s_ in range(requested, 0, -1):
        if steps % s_ == 0 and (s_ >= min(4, requested) or s_ == 1):
            return s_
    return requested


def run_b200(args):
    import torch.distributed as dist
    from depthestimation_b200 import _lib, sharding
    from depthestimation_b200.pipeline import STREAMED_KEYS, DpvLcPipeline, synthetic_batch
    rank, world, local = sharding.init_from_env(args.gpus)
    cores = sharding.pin_rank_to_cores(local, world)    # before any helper thread / the clock sampler starts
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    host, meta = synthetic_batch(args.workload, seed=rank)
    B = meta["B"]
    S = pick_slots(args.streams, args.steps)
    hit_noise = args.hit_noise if meta["iterations"] > 0 else 0.0
    pack = args.planner_pack
    pipes = [DpvLcPipeline(meta, dev, hit_noise=hit_noise, noise_seed=1000 * rank + sl, planner_pack=pack) for sl in range(S)]
    streams = [torch.cuda.Stream(dev) for _ in range(S)]
    pipe = pipes[0]
    x = {k: v.to(dev) for k, v in host.items()}
    # every batch slot has its own resident copy of the per-step inputs: `value` leaves them where they are, `e2e`
    # refills them from pinned host memory before each step (the captured graphs read them in both cases)
    xs = [dict(x, **{k: x[k].clone() for k in STREAMED_KEYS}) for _ in range(S)]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak_gbs, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    main = torch.cuda.current_stream(dev)

    def sync():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    graphs = None                                       # per batch slot: (CUDAGraph, its static output dict)
    launches_per_step = None

    def capture_graphs():
        """A step is ~80 kernel launches for a few ms of GPU work; with several ranks on one host the launching
        threads fall behind the GPUs.  One CUDA graph per batch slot (captured after an eager step has sized every
        buffer) makes a step one graph launch.  The graph holds exactly the kernels the eager step launches."""
        got = []
        n_before = _lib.launch_count()
        for sl in range(S):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=streams[sl]):
                out = pipes[sl].step(xs[sl])
            got.append((g, out))
        return got, (_lib.launch_count() - n_before) // S

    collectors = {}

    def collector_for(n, sample_out):
        """One collector per step count, its buffers allocated on first use (always outside the timed region: the
        same step count is run untimed first) and reused after that."""
        if n not in collectors:
            collectors[n] = sharding.FrameOutputCollector(n, world if args.gather != "none" else 1,
                                                          chunk={"round": S, "step": 1}.get(args.gather, 0)).reserve(sample_out)
        return collectors[n].reset()

    def run_steps(n, step_events=None):
        """n steps, step i on stream i % S: successive (independent) batches overlap -- the latency-bound planner of
        one batch runs beside the streaming kernels of the next.  Each step's depth maps go to every rank through the
        collector's side-stream all_gather while later steps compute."""
        start = torch.cuda.Event()
        start.record(main)
        for st in streams:
            st.wait_event(start)
        coll = collector_for(n, graphs[0][1] if graphs is not None else last_out[0])
        for i in range(n):
            with torch.cuda.stream(streams[i % S]):
                if step_events is not None:
                    a = torch.cuda.Event(enable_timing=True)
                    a.record()
                if graphs is not None:
                    graphs[i % S][0].replay()
                    coll.add(graphs[i % S][1])
                else:
                    coll.add(pipes[i % S].step(xs[i % S]))
                if step_events is not None:
                    b = torch.cuda.Event(enable_timing=True)
                    b.record()
                    step_events.append((a, b))
        for st in streams:
            done = torch.cuda.Event()
            done.record(st)
            main.wait_event(done)
        coll.finish()                                   # waits for the last per-step gathers (NCCL, side stream)

    # ---- value: inputs resident in HBM --------------------------------------------------------
    sampler = ClockSampler(local).start()               # before the region: its start-up must not land inside it
    last_out = [pipes[0].step(xs[0])]                   # sizes every buffer of slot 0; the shapes the collector needs
    torch.cuda.synchronize(dev)
    run_steps(max(args.warmup, S))
    sync()
    launch_mode = "eager"
    if args.launch in ("graph", "auto"):
        try:
            graphs, launches_per_step = capture_graphs()
            launch_mode = "cuda-graph replay, one graph per batch slot"
            sync()
            run_steps(max(args.warmup, S))              # warm-up of the replays themselves
            sync()
            run_steps(args.steps)                       # the timed call's own shape once untimed: its collector buffers
            sync()                                      # are allocated here, not inside the region
        except Exception as e:                          # stay measurable: kernel-by-kernel launches
            graphs = None
            launch_mode = "eager (graph capture failed: %s)" % str(e).splitlines()[0][:120]
            torch.cuda.synchronize(dev)

    def load_pass():
        """S steps without the collector (no collective, no barrier): keeps the GPU under the step's load."""
        for sl in range(S):
            with torch.cuda.stream(streams[sl]):
                if graphs is not None:
                    graphs[sl][0].replay()
                else:
                    pipes[sl].step(xs[sl])
        torch.cuda.synchronize(dev)

    def set_kernel_events(on):
        for p_ in pipes:
            p_.fuse_events = [] if on else None
            p_.planner_events = [] if on else None
            p_.lsm_events = None

    if graphs is None:
        run_steps(args.steps)                           # as above for eager launches
        sync()
    sampler.wait_ready()
    load_pass()                                         # the GPU is under the step's load when the region starts
    set_kernel_events(graphs is None)                   # events inside a captured graph cannot be timed
    step_ev = []
    n0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    t_load0 = time.perf_counter()
    e0.record(main)
    run_steps(args.steps, step_ev)
    e1.record(main)
    sync()
    n1 = _lib.launch_count()
    fuse_in = [ev for p_ in pipes for ev in (p_.fuse_events or [])]
    set_kernel_events(False)
    extra = 0
    while sampler.p and sampler.count(t_load0) == 0 and time.perf_counter() - t_load0 < 3.0:
        load_pass()                                     # a short region can fall between two 200 ms samples: keep the
        extra += 1                                      # same load (untimed, rank-local) until one has been taken
    t_load1 = time.perf_counter()
    sampler.stop()
    clk_summary = sampler.summary(t_load0, t_load1)
    clk_summary["window"] = "timed region" if extra == 0 else "timed region + %d untimed passes of the same steps" % extra
    ms = e0.elapsed_time(e1)
    launches = n1 - n0 if graphs is None else launches_per_step * args.steps
    # how long each step took from its first to its last kernel (S of them overlap): a stretched run names its step
    per_step = sorted(a.elapsed_time(b) for a, b in step_ev)
    step_spread = {"min": round(per_step[0], 3), "median": round(per_step[len(per_step) // 2], 3),
                   "max": round(per_step[-1], 3), "note": "ms from a step's first to its last kernel, %d in flight" % S}
    ms_t = torch.tensor([ms], device=dev)
    ms_ranks = [round(ms / args.steps, 4)]
    if world > 1:
        every = torch.empty(world, device=dev)
        dist.all_gather_into_tensor(every, ms_t)                 # each rank's own time: a slow GPU or host shows here
        ms_ranks = [round(float(v) / args.steps, 4) for v in every.tolist()]
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms = float(ms_t.item())
    ms_per_step = ms / args.steps
    value = world * B * args.steps / (ms / 1e3)

    # ---- one batch in flight (eager, CUDA events around single launches): the latency of a batch and the kernels'
    # own durations -- the condition the ncu launch list measures too.  Right after the timed region, same data.
    replay, graphs = graphs, None
    set_kernel_events(False)
    pipe.set_planner_pack(1)                            # one batch alone: the latency setting (the graphs keep theirs)
    with torch.cuda.stream(main):
        pipe.step(xs[0])                                # warm: eager launches after graph replays
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        pipe.fuse_events, pipe.planner_events, pipe.lsm_events = [], [], []
        s0.record(main)
        for _ in range(3):
            pipe.step(xs[0])
        s1.record(main)
    torch.cuda.synchronize(dev)
    single_batch_ms = s0.elapsed_time(s1) / 3
    fuse_solo = [a.elapsed_time(b) for a, b in pipe.fuse_events]
    plan_solo = [a.elapsed_time(b) for a, b in pipe.planner_events]
    lsm_solo = [a.elapsed_time(b) for a, b in pipe.lsm_events]
    set_kernel_events(False)
    pipe.set_planner_pack(pack)
    graphs = replay
    # the same, replaying slot 0's captured graph alone (what a caller with one batch and a captured step gets)
    single_batch_graph_ms = None
    if graphs is not None and pack == 1:
        with torch.cuda.stream(streams[0]):
            graphs[0][0].replay()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record()
            for _ in range(3):
                graphs[0][0].replay()
            g1.record()
        torch.cuda.synchronize(dev)
        single_batch_graph_ms = g0.elapsed_time(g1) / 3

    # ---- e2e: host buffers in, host results out, copies inside the timed region ------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, pipes, streams, host, xs, graphs, dev, world, B, STREAMED_KEYS, sync, sharding)

    # ---- roofline of the dominant HBM kernel (sense+fuse: one read + one write of the volume) and the kernel that
    # dominates by TIME (the DP planner: latency-bound, SURVEY 8d reports it in microseconds and SMs, not GB/s)
    alg = pipe.algorithmic_bytes()
    roof = None
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    traffic = json.load(open(traffic_file)) if os.path.exists(traffic_file) else {}
    t_lsm = statistics.mean(lsm_solo) / 1e3
    lsm = {"kernel": "lsm_stats_tma_kernel" if meta["D"] in (64, 128) else "lsm_stats_kernel", "bound": "hbm",
           "achieved": round(alg["lsm_full"] / t_lsm / 1e9, 1), "peak": peak_gbs, "peak_source": peak_src, "unit": "GB/s",
           "frac": round(alg["lsm_full"] / t_lsm / 1e9 / peak_gbs, 4), "traffic": traffic.get("lsm_stats_tma_kernel"),
           "launch_ms": round(t_lsm * 1e3, 4), "launches_timed": len(lsm_solo),
           "algorithmic_bytes_per_launch": alg["lsm_full"],
           "timed_with": "CUDA events around each launch, 1 batch in flight (pass after the timed region)",
           "share_of_single_batch_step": round(sum(lsm_solo) / 3 / single_batch_ms, 4),
           "role": "full-resolution log-softmax + statistics + carry (TMA tiles): reads the logits, writes the DPV"}
    if not fuse_solo:
        roof = lsm                       # a workload without LC iterations: this is the kernel that moves the most bytes
    if fuse_solo:
        t = statistics.mean(fuse_solo) / 1e3
        ach = alg["sense_fuse"] / t / 1e9
        nz_bytes = 0 if pipe.noise_pool is None else pipe.noise_pool[0].numel() * 4      # read once per launch
        roof = {"kernel": "sense_fuse_kernel", "bound": "hbm", "achieved": round(ach, 1), "peak": peak_gbs,
                "peak_source": peak_src, "unit": "GB/s", "frac": round(ach / peak_gbs, 4), "traffic": None,
                "launch_ms": round(t * 1e3, 4), "launches_timed": len(fuse_solo),
                "algorithmic_bytes_per_launch": alg["sense_fuse"], "noise_bytes_per_launch_not_counted": nz_bytes,
                "timed_with": "CUDA events around each launch, 1 batch in flight (pass after the timed region)",
                "share_of_single_batch_step": round(sum(fuse_solo) / 3 / single_batch_ms, 4),
                "role": "dominant HBM kernel (named by hand: the kernel that moves the most bytes per step)"}
        if fuse_in:
            tin = [a.elapsed_time(b) for a, b in fuse_in]
            roof["in_flight"] = {"launch_ms": round(statistics.mean(tin), 4), "batches": S,
                                 "note": "event-to-event inside the timed region: includes the slices other batches' "
                                         "kernels get on the same SMs, so it is not the kernel's own duration"}
        roof["traffic"] = traffic.get("sense_fuse_kernel")
        roof["also"] = {"lsm": lsm}
        if plan_solo:
            nmaps = B * (1 + 2 * len(meta["step"]))
            roof["dominant_by_time"] = {
                "kernel": "planner_solve_kernel", "class": "latency-bound (W sequential rays per map)",
                "launch_us": round(statistics.mean(plan_solo) * 1e3, 1), "launches_per_step": meta["iterations"],
                "maps_per_launch": nmaps, "sms_occupied": 2 * nmaps, "sms_total": 148,
                "share_of_single_batch_step": round(sum(plan_solo) / 3 / single_batch_ms, 4)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, dt, ncores, _ = time_cpu(args.workload, 1, 1)            # one warm-up like the reference arm (flow field,
        cpu = {"value": round(v, 5), "unit": UNIT, "cores": ncores, "kind": "port",     # handles, first-touch)
               "sample": "1 frame of the same workload through oracle/ after 1 warm-up frame (torch-CPU restatement + "
                         "restated lcsim C++; leaner than the reference's own Python: no D x C warp, no repeat()), "
                         "%.1f s" % dt}
    if rank == 0:
        line = {"metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": args.workload, "frames_per_gpu": B, "H": meta["H"], "W": meta["W"],
                           "D": meta["D"], "C": meta["C"], "T": meta["T"], "lc_iterations": meta["iterations"],
                           "curtains_per_iteration": 1 + 2 * len(meta["step"]), "hit_noise": hit_noise,
                           "noise": "pre-generated pool per batch slot" if hit_noise > 0 else "none",
                           "parallelism": "batch-shard x%d" % world, "streams_in_flight": S, "launch": launch_mode,
                           "planner_pack": {"throughput_region": pack, "single_batch_pass": 1},
                           "gather": ({"round": "one all_gather per round of %d steps on a side stream" % S,
                                       "step": "per-step all_gather on a side stream", "end": "one all_gather after "
                                       "the last step", "none": "DISABLED (diagnostic run)"}[args.gather]
                                      if world > 1 else "none (1 rank)"),
                           "rank_cores": None if cores is None else len(cores),
                           "l2_policy": "inputs larger than L2 (0.9 GB streamed per frame)"},
                "clocks": clk_summary, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu,
                "ms_per_frame": round(ms_per_step / B, 4), "single_batch_ms": round(single_batch_ms, 4),
                "single_batch_note": "latency of ONE batch of %d frame(s) through the whole step, nothing else in "
                                     "flight (eager launches)" % B,
                "single_batch_graph_ms": None if single_batch_graph_ms is None else round(single_batch_graph_ms, 4),
                "step_spread_ms": step_spread, "ms_per_step_by_rank": ms_ranks}
        emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_e2e(args, pipes, streams, host, xs, graphs, dev, world, B, keys, sync, sharding):
    """Same step through the public API with HOST buffers: every step copies its inputs from pinned host memory into
    its batch slot's device buffers (H2D) and reads depth / variance / LC depth back (D2H), on the step's own stream,
    so the copies of one batch overlap the kernels of the others.  Launch mode as in the `value` loop."""
    S = len(pipes)
    m = pipes[0].m
    pin = {k: host[k].pin_memory() for k in keys}
    res_host = [{k: torch.empty(B, m["H"], m["W"]).pin_memory() for k in ("depth", "var", "depth_lc")} for _ in range(S)]
    main = torch.cuda.current_stream(dev)
    h2d = sum(pin[k].numel() * 4 for k in keys)
    d2h = sum(v.numel() * 4 for v in res_host[0].values())
    e2e_coll = {}

    def loop(n):
        start = torch.cuda.Event()
        start.record(main)
        for st in streams:
            st.wait_event(start)
        if n not in e2e_coll:
            e2e_coll[n] = sharding.FrameOutputCollector(n, world, chunk=S).reserve(graphs[0][1] if graphs is not None else pipes[0].step(xs[0]))
        coll = e2e_coll[n].reset()
        for i in range(n):
            sl = i % S
            with torch.cuda.stream(streams[sl]):
                for k in keys:
                    xs[sl][k].copy_(pin[k], non_blocking=True)
                if graphs is not None:
                    graphs[sl][0].replay()
                    out = graphs[sl][1]
                else:
                    out = pipes[sl].step(xs[sl])
                coll.add(out)
                for k in res_host[sl]:
                    if out.get(k) is not None:
                        res_host[sl][k].copy_(out[k], non_blocking=True)
        for st in streams:
            done = torch.cuda.Event()
            done.record(st)
            main.wait_event(done)
        coll.finish()
        torch.cuda.synchronize(dev)

    loop(max(args.warmup, S))
    sync()
    loop(args.steps)                                    # the timed call's shape once untimed (buffers, pinned pages)
    sync()
    t0 = time.perf_counter()
    loop(args.steps)
    sync()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    per_rank_gbs = h2d * args.steps / dt / 1e9
    return {"value": round(world * B * args.steps / dt, 2), "unit": UNIT, "h2d_bytes_per_step": h2d,
            "d2h_bytes_per_step": d2h, "ms_per_step": round(dt / args.steps * 1e3, 4),
            "h2d_gbs_per_rank": round(per_rank_gbs, 1), "h2d_gbs_all_ranks": round(per_rank_gbs * world, 1),
            "limiter": "host link: every step uploads its inputs (%.0f MB, of which the decoder logits are %.0f MB) from "
                       "pinned host memory; one rank is PCIe-bound, several ranks share the host's memory / PCIe "
                       "bandwidth" % (h2d / 1e6, pin["dec_logits"].numel() * 4 / 1e6)}


def run_reference(args):
    """Reference arm: the reference's own CPU algorithm for this path (oracle port; lcsim cannot be compiled here,
    DESIGN.md) on all host cores.  Each step = 1 frame of the workload.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from depthestimation_b200.pipeline import WORKLOADS
    v, dt, cores, meta = time_cpu(args.workload, args.steps, min(args.warmup, 1))
    w = WORKLOADS[args.workload]
    line = {"impl": "reference", "metric": METRIC, "value": round(v, 5), "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 2), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "frames_per_gpu": w["B"], "H": w["H"], "W": w["W"], "D": w["D"],
                       "C": w["C"], "T": w["T"], "lc_iterations": w["iterations"],
                       "curtains_per_iteration": 1 + 2 * len(w["step"])},
            "cpu_baseline": {"value": round(v, 5), "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "each step = 1 frame of the workload through oracle/ (serial over batch like the "
                                       "reference, models.py:984)"},
            "e2e": {"value": round(v, 5), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(json.dumps(line))


if __name__ == "__main__":
    a = parse()
    # stdout carries ONE JSON line: whatever libraries print there (NCCL's version banner does, on rank 0, when the
    # environment sets NCCL_DEBUG) goes to stderr instead; emit() writes the line to the real stdout
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
