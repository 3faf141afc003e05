"""GPU: per-CTA timeline of the fused BA linearisation kernel (start/end by kind)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from visual_slam_b200.ba import BAEngine  # noqa: E402
from visual_slam_b200.ba_problem import synthetic_ba  # noqa: E402
name = sys.argv[1] if len(sys.argv) > 1 else "c4"
cfg = {"c3": dict(n_cams=10, n_pts=5000, obs_per_pt=6.0, seed=0),
       "c4": dict(n_cams=200, n_pts=100000, obs_per_pt=10.0, seed=1)}[name]
p = synthetic_ba(**cfg)
eng = BAEngine(p)
eng.linearize()
for rep in range(2):
    t = eng.trace_linearize()
print("marks us (stamp before, controller end, stamp after):", [round(x / 1e3, 1) for x in eng.last_trace_marks])
print("ctas", len(t), "span %.1f us" % (t[:, 2].max() / 1e3))
for kind, nm in ((0, "camera"), (1, "point"), (2, "scale")):
    m = t[:, 0] == kind
    if not m.any():
        continue
    d = (t[m, 2] - t[m, 1]) / 1e3
    print("%-7s n=%5d  dur us: min %.1f med %.1f p90 %.1f max %.1f | start max %.1f end max %.1f" % (
        nm, m.sum(), d.min(), np.median(d), np.percentile(d, 90), d.max(), t[m, 1].max() / 1e3, t[m, 2].max() / 1e3))
ends = np.sort(t[:, 2]) / 1e3
print("end-time percentiles us:", [round(float(np.percentile(ends, q)), 1) for q in (10, 50, 75, 90, 95, 99, 100)])
late = np.argsort(-t[:, 2])[:8]
for i in late:
    print("  cta %5d kind %d start %.1f end %.1f" % (i, t[i, 0], t[i, 1] / 1e3, t[i, 2] / 1e3))
# duration against start time, and the number of resident CTAs over time
st, en = t[:, 1] / 1e3, t[:, 2] / 1e3
edges = np.arange(0, st.max() + 5, 5.0)
for kind, nm in ((0, "camera"), (1, "point")):
    m = t[:, 0] == kind
    row = []
    for a, b in zip(edges[:-1], edges[1:]):
        s = m & (st >= a) & (st < b)
        row.append("%4.1f-%4.1f: n=%4d med %.1f" % (a, b, s.sum(), np.median((en - st)[s]) if s.any() else 0.0))
    print(nm, "duration by start time |", " | ".join(row))
grid = np.arange(0, en.max(), 2.5)
print("resident CTAs at t:", [(float(x), int(((st <= x) & (en > x)).sum())) for x in grid])
