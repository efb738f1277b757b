"""Generates tests/golden/planner_ref.npz from the REFERENCE's own planner: external/lcsim/lcsim/src/planning.cpp
compiled verbatim (oracle/ref_planner.py) and run on

  * the reference's fixture (external/lcsim/python/example/dpv.npy: intrinsics, 128 ranges, the 128 x 400 field), at
    the two laser time steps the path uses (3.5e-5 s: lc/light_curtain.py:283, and 1e-4 s),
  * the benchmark geometry (768 rays, 128 ranges 5-40 m, KITTI-shaped intrinsics of pipeline.synthetic_batch).

Planner inputs: the per-column camera angles of the device model (oracle/lcsim LCDevice.valid_angles(), which
reproduces the fixture's own `angles` to 2e-5 deg and is what the GPU library computes bit for bit), the ranges,
laser_to_cam = inverse of lTc (translate x by 0.2 m -> x - 0.2, exactly).  Maps: the fixture field, a uniform random
map, maps quantised to a few levels (ties everywhere: the strict-">" / lowest-index / smaller-laser-penalty rules of
planning.cpp:83-104, :198-217 decide), a map with one non-zero entry.  An all-zero map is NOT included: the reference
dereferences a null trajectory on it (planning.cpp:101-104,:122).

Run in the build container (needs /root/reference):  python tests/golden/make_planner_ref.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import lcsim as olcsim  # noqa: E402
from oracle import ref_planner as RP  # noqa: E402


def lTc():
    T = np.eye(4, dtype=np.float32)
    T[0, 3] = 0.2
    return T


def device(K, W, H, timestep):
    cp = dict(width=W, height=H, matrix=np.asarray(K, np.float32), distortion=[0.] * 5, hit_mode=1, hit_noise=0.)
    lp = dict(lTc=lTc(), fov=80., laser_timestep=timestep)
    return olcsim.LCDevice(CAMERA_PARAMS=cp, LASER_PARAMS=lp)


def maps_for(R, W, seed, first=None):
    gen = np.random.default_rng(seed)
    one = np.zeros((R, W), np.float32)
    one[R // 3, W // 2] = 1.0
    m = [gen.random((R, W), dtype=np.float32), np.round(gen.random((R, W)) * 4).astype(np.float32) / 4,
         (gen.random((R, W)) > 0.9).astype(np.float32), one]
    return ([first] if first is not None else []) + m


def main():
    fx = np.load(os.path.join(HERE, "lcsim_dpv_fixture.npz"))
    W, H = int(fx["size"][0]), int(fx["size"][1])
    laser_to_cam = np.linalg.inv(lTc().astype(np.float64)).astype(np.float32)
    assert laser_to_cam[0, 3] == np.float32(-0.2)
    out = dict(laser_to_cam=laser_to_cam)
    cases = [("fixture_t35", fx["intrinsics"], W, H, fx["r_candi"], 3.5e-5, fx["field"]),
             ("fixture_t100", fx["intrinsics"], W, H, fx["r_candi"], 1.0e-4, fx["field"])]
    from depthestimation_b200.pipeline import kitti_intrinsics
    K, _ = kitti_intrinsics(192, 64)
    Kup = K * 4
    Kup[2, 2] = 1
    cases.append(("bench_768", Kup, 768, 256, np.linspace(5.0, 40.0, 128).astype(np.float32), 3.5e-5, None))
    for name, Kc, w, h, ranges, ts, first in cases:
        dev = device(Kc, w, h, ts)
        angles = dev.valid_angles()
        ref = RP.RefPlanner(angles, ranges, laser_to_cam, 14000.0, ts)
        maps = maps_for(len(ranges), w, seed=len(name) + w, first=first)
        paths = [ref.get_design_points(m) for m in maps]      # a path ends early at a node without a successor
        counts = np.array([len(p) for p in paths], np.int32)
        xz = np.full((len(maps), w, 2), np.nan, np.float32)
        for i, p in enumerate(paths):
            xz[i, :len(p)] = p
        out.update({name + "_K": np.asarray(Kc, np.float32), name + "_size": np.array([w, h]), name + "_ranges": ranges,
                    name + "_timestep": np.float32(ts), name + "_angles": angles,
                    name + "_maps": np.stack(maps).astype(np.float32), name + "_xz": xz,
                    name + "_counts": counts,
                    name + "_edges": np.int64(ref.edge_count())})
        print(name, "edges", ref.edge_count(), "path lengths", counts.tolist())
    np.savez_compressed(os.path.join(HERE, "planner_ref.npz"), **out)
    print("wrote planner_ref.npz, %.1f KB" % (os.path.getsize(os.path.join(HERE, "planner_ref.npz")) / 1e3))


if __name__ == "__main__":
    main()
