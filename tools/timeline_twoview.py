"""GPU: per-kernel durations of the batched two-view path (undistort + 5-point RANSAC + recoverPose) at the bench.py
configuration, through torch.profiler (CUPTI kernel records)."""
import sys
from collections import defaultdict

import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
from visual_slam_b200 import synth, twoview  # noqa: E402

pairs, n, hyp = 1024, 512, 256
dev = torch.device("cuda")
data = [synth.two_view_pair(s, n) for s in range(32)]
p1 = torch.from_numpy(np.stack([d[0] for d in data])).to(dev).repeat(pairs // 32, 1, 1).contiguous()
p2 = torch.from_numpy(np.stack([d[1] for d in data])).to(dev).repeat(pairs // 32, 1, 1).contiguous()
n_pts = torch.full((pairs,), n, dtype=torch.int32, device=dev)
words = torch.from_numpy(np.random.default_rng(3).integers(0, 2 ** 32, size=(pairs, hyp, 5), dtype=np.uint64)
                         .astype(np.uint32).view(np.int32)).to(dev)
K = synth.ICL_K


def one():
    x1, x2 = twoview.undistort_points(p1, K), twoview.undistort_points(p2, K)
    o = twoview.essential_ransac_batch(x1, x2, n_pts, words, 3.0 / K[0])
    return o, twoview.recover_pose_batch(o["E"], p1, p2, n_pts, K, 50.0)


for _ in range(2):
    one()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); one(); e1.record(); torch.cuda.synchronize()
print("one batch: %.3f ms" % e0.elapsed_time(e1))
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    one()
    torch.cuda.synchronize()
tot = defaultdict(float)
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        tot[e.name[:70]] += e.time_range.end - e.time_range.start
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print("%-72s %9.1f us" % (k, v))
