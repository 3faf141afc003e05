"""GPU timeline of BA LM iterations (kernel start / duration / gap to the previous kernel) through torch.profiler
(CUPTI activity records cover the library's own launches too).  Usage: python tools/timeline_ba.py [c3|c4|w17|w45] [solver]"""
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
from visual_slam_b200.ba import BAEngine  # noqa: E402
from visual_slam_b200.ba_problem import synthetic_ba  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
cfg = {"c3": dict(n_cams=10, n_pts=5000, obs_per_pt=6.0, seed=0),
       "c4": dict(n_cams=200, n_pts=100000, obs_per_pt=10.0, seed=1),
       # mid-size windows: the single-CTA four-rows-per-lane PCG, the cluster PCG
       "w17": dict(n_cams=17, n_pts=4000, obs_per_pt=8.0, seed=3),
       "w45": dict(n_cams=45, n_pts=4000, obs_per_pt=8.0, seed=3)}[name]
p = synthetic_ba(**cfg)
eng = BAEngine(p)
if len(sys.argv) > 2:
    eng.set_linear_solver(sys.argv[2])
for _ in range(3):
    eng.set_state(p.cams, p.pts, p.obs_uv, p.scale_meas, p.K)
    eng.optimize(10)
torch.cuda.synchronize()
eng.set_state(p.cams, p.pts, p.obs_uv, p.scale_meas, p.K)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    st = eng.optimize(10)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start
prev_end = t0
print("%-44s %9s %8s %8s" % ("kernel", "start us", "dur us", "gap us"))
rows = []
for e in ev:
    s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
    rows.append((e.name[:44], s, d, e.time_range.start - prev_end))
    prev_end = max(prev_end, e.time_range.end)
n_per = max(1, len(rows) // 10)
for r in rows[: 2 * n_per + 2]:
    print("%-44s %9.1f %8.1f %8.1f" % r)
print("... total %d kernels, span %.1f us, sum of durations %.1f us, sum of positive gaps %.1f us" % (
    len(rows), rows[-1][1] + rows[-1][2], sum(r[2] for r in rows), sum(max(r[3], 0) for r in rows)))
