"""Short BA run for ncu: C4 (or C3) linearise + one LM trial after one warm-up pass."""
import sys
import torch
sys.path.insert(0, ".")
from visual_slam_b200.ba import BAEngine  # noqa: E402
from visual_slam_b200.ba_problem import synthetic_ba  # noqa: E402
variant = "units"
if "--chunks" in sys.argv:
    sys.argv.remove("--chunks"); variant = "chunks"
name = sys.argv[1] if len(sys.argv) > 1 else "c4"
cfg = {"c3": dict(n_cams=10, n_pts=5000, obs_per_pt=6.0, seed=0),
       "c4": dict(n_cams=200, n_pts=100000, obs_per_pt=10.0, seed=1),
       "w17": dict(n_cams=17, n_pts=4000, obs_per_pt=8.0, seed=3),
       "w45": dict(n_cams=45, n_pts=4000, obs_per_pt=8.0, seed=3)}[name]
p = synthetic_ba(**cfg)
eng = BAEngine(p)
eng.set_schur_variant(variant)
for _ in range(2):
    lin = eng.linearize()
    tr = eng.trial(1e-5 * lin["max_diag"])
torch.cuda.synchronize()
print(lin, tr)
