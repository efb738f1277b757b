import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_lc as T
import torch
from depthestimation_b200.lc.light_curtain import LightCurtain
from oracle import lc_oracle as LO
g = T.load("plan_m1")
D, W = g["uf"].shape; H = 32
eye = np.eye(4, dtype=np.float32)
P = {"intr_rgb": g["Kup"], "dist_rgb": [0.] * 5, "size_rgb": [W, H], "intr_lc": g["Kup"], "dist_lc": [0.] * 5,
     "size_lc": [W, H], "rTc": eye, "lTc": T._lTc(), "laser_fov": 80., "d_candi": g["d_candi"], "r_candi": g["d_candi"],
     "d_candi_up": g["d_candi_up"], "r_candi_up": g["d_candi_up"], "name": "m1", "laser_timestep": 3.5e-5,
     "device": torch.device("cuda", 0), "expand_A": D, "expand_B": len(g["d_candi_up"])}
lc = LightCurtain(); lc.init(P)
cfg = {"step": int(g["step"]), "interval": int(g["interval"])}
for seed in (int(g["seed"]), 1, 2, 3, 4):
    np.random.seed(seed)
    pts, vis = lc.plan_m1_high(T.cu(g["uf"]), cfg)
    olc = LO.LightCurtainOracle(g["Kup"], W, H, g["d_candi"], g["d_candi_up"])
    np.random.seed(seed)
    ref, _ = olc.plan_m1(torch.as_tensor(g["uf"]).clone(), cfg)
    print("seed", seed, [round(float(np.all(p.cpu().numpy() == r, axis=1).mean()), 3) if p.shape == r.shape else "shape" for p, r in zip(pts, ref)])
gl = T.load("light_curtain")
lc2 = T._golden_lc(gl)
Hh, Ww = gl["logp"].shape[2:]
uf2 = torch.exp(T.cu(gl["logp"]))[0].mean(1); uf2 = uf2 / uf2.sum(0, keepdim=True)
olc = LO.LightCurtainOracle(gl["Kup"], Ww, Hh, gl["d_candi"], gl["d_candi_up"])
for seed in (5, 6, 7, 8):
    np.random.seed(seed); got, _ = lc2.plan_m1_high(uf2, {"step": 2, "interval": 8})
    np.random.seed(seed); ref, _ = olc.plan_m1(uf2.cpu().clone(), {"step": 2, "interval": 8})
    print("field2 seed", seed, [round(float(np.all(p.cpu().numpy() == r, axis=1).mean()), 3) if p.shape == r.shape else "shape" for p, r in zip(got, ref)])
