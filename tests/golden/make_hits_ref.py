"""Generates tests/golden/hits_ref.npz from the REFERENCE's own per-pixel curtain return: Util::intersect
(external/lcsim/lcsim/src/util.cpp:47-53) and DatumProcessor::computeDepthHits (src/dprocessor.cpp:942-993, hit modes
0 and 1, noise off), compiled from where they lie by oracle/ref_hits.py, with calculateSurface's two per-pixel rules
around them restated in numpy (the NaN rule :824-829, the range uncertainty :884-887).

Inputs (stored in the file): the plane coefficients of three curtains, the pixel rays, the z-to-range map and the hit
look-up table of the 64 x 32 golden camera -- these come from the restated device model (their Eigen arithmetic cannot
be compiled here) -- and a depth image with NaN pixels.  Outputs: surface points, thickness and quantised intensity
per pixel as the reference's functions produce them.

Run in the build container (needs /root/reference):  python tests/golden/make_hits_ref.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import lcsim as olcsim  # noqa: E402
from oracle import ref_hits  # noqa: E402


def reference_return(planes, nx, ny, depth, zt, lut, mode):
    main = ref_hits.intersect(planes[:, 0:4], nx, ny)
    bot = ref_hits.intersect(planes[:, 4:8], nx, ny)
    top = ref_hits.intersect(planes[:, 8:12], nx, ny)
    surf = main.copy()
    surf[(surf[..., 2] < 0) | (np.abs(surf[..., 1]) > 3.0)] = np.nan            # dprocessor.cpp:824-829
    d = bot[..., :3].astype(np.float64) - top[..., :3].astype(np.float64)      # :884-886: pow(float, 2) is double
    unc = np.sqrt((d ** 2).sum(-1)).astype(np.float32)
    return ref_hits.compute_depth_hits(surf, unc, depth, zt, lut, mode), unc


def main():
    assert ref_hits.build() is not None
    g = np.load(os.path.join(HERE, "light_curtain.npz"))
    H, W = g["logp"].shape[2:]
    lTc = np.eye(4, dtype=np.float32)
    lTc[0, 3] = 0.2
    cp = dict(width=W, height=H, matrix=np.asarray(g["Kup"], np.float32), distortion=[0.] * 5, hit_mode=1, hit_noise=0.)
    dev = olcsim.LCDevice(CAMERA_PARAMS=cp, LASER_PARAMS=dict(lTc=lTc, fov=80., laser_timestep=3.5e-5))
    depth = np.repeat(np.repeat(g["dmaps"], 4, 0), 4, 1).astype(np.float32)
    depth[3:6, 10:14] = np.nan
    nx, ny = dev.rays()
    out = dict(Kup=g["Kup"], depth=depth, nn_x=nx, nn_y=ny, ztoramap=dev.ztoramap(), lut=dev.hit_lut())
    for i, pts in enumerate((g["path0"], g["path1"], g["path2"][5:40])):
        _, _, _, planes = dev.get_return(depth, pts, True, debug=True)
        out["pts%d" % i], out["planes%d" % i] = np.asarray(pts, np.float32), planes
        for mode in (0, 1):
            img, unc = reference_return(planes, nx, ny, depth, out["ztoramap"], out["lut"], mode)
            out["image%d_mode%d" % (i, mode)] = img
        out["thick%d" % i] = unc
    np.savez_compressed(os.path.join(HERE, "hits_ref.npz"), **out)
    print("wrote hits_ref", {k: np.shape(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
