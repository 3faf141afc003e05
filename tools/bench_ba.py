"""GPU: times BA LM iterations at C3 (10 KF x 5k pts) and C4 (200 KF x 100k pts)."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from visual_slam_b200.ba import BAEngine  # noqa: E402
from visual_slam_b200.ba_problem import synthetic_ba  # noqa: E402

variant = "units"
if "--chunks" in sys.argv:
    sys.argv.remove("--chunks"); variant = "chunks"
which = sys.argv[1:] or ["c3", "c4"]
cfg = {"c3": dict(n_cams=10, n_pts=5000, obs_per_pt=6.0, seed=0),
       "c4": dict(n_cams=200, n_pts=100000, obs_per_pt=10.0, seed=1)}
for name in which:
    t0 = time.time()
    p = synthetic_ba(**cfg[name])
    t1 = time.time()
    eng = BAEngine(p)
    eng.set_schur_variant(variant)
    torch.cuda.synchronize()
    t2 = time.time()
    print(name, "cams %d pts %d obs %d | gen %.2fs create %.2fs" % (p.n_cams, p.n_pts, p.n_obs, t1 - t0, t2 - t1), eng.dims())
    for rep in range(3):
        eng.set_state(p.cams, p.pts, p.obs_uv, p.scale_meas, p.K)
        torch.cuda.synchronize()
        t = time.time()
        st = eng.optimize(10)
        torch.cuda.synchronize()
        dt = time.time() - t
        print("  optimize(10): %.2f ms  -> %.1f LM iters/s (trials %d, pcg iters %d) cost %.6g -> %.6g" % (
            dt * 1e3, st["iterations"] / dt, st["trials"], st["pcg_iterations"], st["initial_cost"], st["final_cost"]))
    # stage timing
    eng.set_state(p.cams, p.pts, p.obs_uv, p.scale_meas, p.K)
    torch.cuda.synchronize()
    t = time.time(); lin = eng.linearize(); torch.cuda.synchronize(); tl = time.time() - t
    t = time.time(); lin = eng.linearize(); torch.cuda.synchronize(); tl = time.time() - t
    t = time.time(); tr = eng.trial(1e-5 * lin["max_diag"]); torch.cuda.synchronize(); tt = time.time() - t
    print("  linearize %.3f ms (%.1f GB/s algorithmic), trial %.3f ms, pcg iters %d status %d" % (
        tl * 1e3, p.linearize_bytes() / tl / 1e9, tt * 1e3, tr["pcg_iters"], tr["pcg_status"]))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    eng.trial(1e-5 * lin["max_diag"])
    for stage in ("linearize", "schur", "trial"):
        ms = eng.time_stage(stage, 6, flush)[1:]
        msw = eng.time_stage(stage, 6, None)[1:]
        extra = " (%.1f GB/s algorithmic cold)" % (p.linearize_bytes() / (np.median(ms) * 1e-3) / 1e9) if stage == "linearize" else ""
        print("  %-9s cold %.1f us  warm %.1f us%s" % (stage, np.median(ms) * 1e3, np.median(msw) * 1e3, extra))
    b2b = eng.time_stage("linearize_back_to_back", 20)[0]
    print("  linearize back-to-back x20: %.1f us per launch (%.1f GB/s algorithmic)" % (b2b * 1e3, p.linearize_bytes() / (b2b * 1e-3) / 1e9))
    del flush
    eng.close()
