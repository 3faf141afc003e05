"""GPU: times the kNN-2 kernel variants (vslam_hamming_set_variant) at C2 and the popc probe."""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from visual_slam_b200 import _cabi, matching, synth  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 592
lib = _cabi.lib()
v = ctypes.c_double()
_cabi.check(lib.vslam_microbench_popc(4096, ctypes.byref(v), None))
print("popc/s probe: %.4e  (per SM per clk @1.965GHz: %.2f)" % (v.value, v.value / 148 / 1.965e9))
qn, tn = synth.descriptors_batch(8, seed=0)
q = torch.from_numpy(qn).cuda().repeat(B // 8, 1, 1)
t = torch.from_numpy(tn).cuda().repeat(B // 8, 1, 1)
for rev in (False, True):
    for var in range(0, 7):
        lib.vslam_hamming_set_variant(var)
        for _ in range(3):
            matching.knn2_batch(q, t, want_rev=rev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 5
        for _ in range(n):
            matching.knn2_batch(q, t, want_rev=rev)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        pd = B * 2048 * 2048 / (ms * 1e-3)
        print("rev=%d variant=%d  %.3f ms/launch  %.1f Gpair-dist/s  %.2f Gpopc-equiv/s (8/pair)  %.3g q/s"
              % (rev, var, ms, pd / 1e9, 8 * pd / 1e9, B * 2048 / (ms * 1e-3)))
lib.vslam_hamming_set_variant(0)
