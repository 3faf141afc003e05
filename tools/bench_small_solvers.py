"""GPU: C3 (10 KF x 5k pts) LM iterations/s with each linear solver of the reduced camera system
(block-Jacobi PCG, register-resident Gauss-Jordan, shared-memory LDL^T), plus a correctness spot check."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from visual_slam_b200.ba import BAEngine, direct_solve  # noqa: E402
from visual_slam_b200.ba_problem import synthetic_ba  # noqa: E402

rng = np.random.default_rng(1)
for n in (6, 12, 30, 42, 54, 60):
    A = rng.normal(size=(n, n)); S = A @ A.T + n * np.eye(n); b = rng.normal(size=n)
    for kern in ("auto", "ldlt"):
        x, status = direct_solve(torch.from_numpy(S).cuda(), torch.from_numpy(b).cuda(), kern)
        err = np.abs(x.cpu().numpy() - np.linalg.solve(S, b)).max() / np.abs(b).max()
        print("n %2d %-5s status %d err %.2e" % (n, kern, status, err))

p = synthetic_ba(n_cams=10, n_pts=5000, obs_per_pt=6.0, seed=0)
for kind in ("pcg", "direct", "ldlt", "auto"):
    eng = BAEngine(p)
    eng.set_linear_solver(kind)
    best = 1e9
    for rep in range(6):
        eng.set_state(p.cams, p.pts, p.obs_uv, p.scale_meas, p.K)
        torch.cuda.synchronize()
        t = time.perf_counter()
        st = eng.optimize(10)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t)
    eng.set_state(p.cams, p.pts, p.obs_uv, p.scale_meas, p.K)
    lin = eng.linearize()
    eng.trial(1e-5 * lin["max_diag"])
    tr = np.median(eng.time_stage("trial", 8, None)[1:]) * 1e3
    sc = np.median(eng.time_stage("schur", 8, None)[1:]) * 1e3
    print("%-6s optimize(10) %.3f ms -> %.0f LM it/s  (iters %d trials %d pcg %d cost %.9g)  trial %.1f us  schur(+solve if fused) %.1f us" % (
        kind, best * 1e3, st["iterations"] / best, st["iterations"], st["trials"], st["pcg_iterations"], st["final_cost"], tr, sc))
    eng.close()
