"""Short matching run for ncu: C2 batch (default 592 pairs), knn2 (+rev) + filter, 3 passes."""
import sys
import torch
sys.path.insert(0, ".")
from visual_slam_b200 import matching, synth  # noqa: E402
B = int(sys.argv[1]) if len(sys.argv) > 1 else 592
qn, tn = synth.descriptors_batch(8, seed=0)
q = torch.from_numpy(qn).cuda().repeat(B // 8, 1, 1)
t = torch.from_numpy(tn).cuda().repeat(B // 8, 1, 1)
for _ in range(3):
    out = matching.match_batch(q, t, 0.8, True)
torch.cuda.synchronize()
print(int(out["n_keep"].sum()))
