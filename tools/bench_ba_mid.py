"""GPU: optimize(10) on a mid-size BA problem (cameras, points from argv); VSLAM_BA_PCG_BLOCKS picks the PCG partition."""
import sys, time, os
import numpy as np, torch
sys.path.insert(0, ".")
from visual_slam_b200.ba import BAEngine
from visual_slam_b200.ba_problem import synthetic_ba
nc = int(sys.argv[1]); npts = int(sys.argv[2])
p = synthetic_ba(n_cams=nc, n_pts=npts, obs_per_pt=8.0, seed=3)
eng = BAEngine(p)
print(os.environ.get("VSLAM_BA_PCG_BLOCKS"), eng.dims())
best = 1e9
for rep in range(6):
    eng.set_state(p.cams, p.pts, p.obs_uv, p.scale_meas, p.K)
    torch.cuda.synchronize(); t = time.perf_counter()
    st = eng.optimize(10)
    torch.cuda.synchronize(); best = min(best, time.perf_counter() - t)
print("  optimize(10) %.3f ms  %.0f it/s  pcg %d cost %.9g" % (best * 1e3, st["iterations"] / best, st["pcg_iterations"], st["final_cost"]))
